#!/bin/bash
# development aid: time the fast force kernel variants (NBODY_FAST_VARIANT) at N=65536
for v in ${VARIANTS:-0 1 2 3 4 5 6 7 8 9}; do
    echo "variant $v: $(NBODY_FAST_VARIANT=$v python tools/quick_bench.py ${CASE:-65536:1:10} 2>&1 | grep 'n=')"
done
