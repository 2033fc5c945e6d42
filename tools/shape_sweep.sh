#!/bin/bash
# development aid (NBODY_DEV build): 128x4 (variant 1) against 64x8 (variant 2) over N, default chunks
DEV=$PWD/parallel-nbody_b200/libnbody_gpu_dev.so
for n in "$@"; do
  for v in 1 2; do
    NBODY_GPU_LIB=$DEV NBODY_SYM_VARIANT=$v python tools/c3_shapes.py one $n 0
  done
done
