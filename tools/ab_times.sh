#!/bin/bash
# development aid: A/B timing of kernel variants built with `make dev DEVNAME=... DEVFLAGS=...`
for lib in "$@"; do
    echo "== $lib ${NBODY_SYM_VARIANT:+(NBODY_SYM_VARIANT=$NBODY_SYM_VARIANT)}"
    NBODY_GPU_LIB=$PWD/parallel-nbody_b200/$lib python tools/quick_bench.py ${CASES:-65536:1:20 262144:1:6}
done
