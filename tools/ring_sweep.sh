#!/bin/bash
# development aid: step time and DRAM bytes of the force kernel for several lengths of the row-sum slice ring
# (make dev DEVNAME=_r2048 DEVFLAGS="-DNBODY_SYM_RING=2048" ...)
mkdir -p gpurun_out
for r in _r1024 _r2048 "" _r8192; do
  lib=$PWD/parallel-nbody_b200/libnbody_gpu_dev$r.so
  echo "== ring ${r:-_r4096}"
  NBODY_GPU_LIB=$lib python tools/quick_bench.py 65536:1:100 262144:1:10 | tail -n 2
  NBODY_GPU_LIB=$lib NBODY_DEV_VRANK=3 python tools/vrank_times.py one 262144
  NBODY_GPU_LIB=$lib ncu --metrics dram__bytes_write.sum,dram__bytes_read.sum,gpu__time_duration.sum --clock-control none -k regex:forces_sym -s 4 -c 1 python tools/quick_bench.py 262144:1:3 2>&1 | grep -E "dram__|gpu__time"
done
