#!/bin/bash
# development aid, one gpurun call: bits (tools/state_hash.py) and times of this build against an earlier build kept
# under tools/_ab/ (libnbody_gpu_base.so, libnbody_gpu_dev_base.so; git-ignored, they travel with gpurun)
# (build them from the commit to compare with: git worktree add /tmp/basewt <commit>; make -C /tmp/basewt/parallel-nbody_b200/csrc all dev;
#  cp the two .so files to tools/_ab/libnbody_gpu_base.so and tools/_ab/libnbody_gpu_dev_base.so)
set -u
mkdir -p gpurun_out
BASE=$PWD/tools/_ab/libnbody_gpu_base.so
DEVBASE=$PWD/tools/_ab/libnbody_gpu_dev_base.so
DEVNEW=$PWD/parallel-nbody_b200/libnbody_gpu_dev.so
timeout 300 python tools/state_hash.py > gpurun_out/hash_new.log 2>&1; echo "hash new rc=$?"
NBODY_GPU_LIB=$BASE timeout 300 python tools/state_hash.py > gpurun_out/hash_base.log 2>&1; echo "hash base rc=$?"
cat gpurun_out/hash_new.log
diff gpurun_out/hash_new.log gpurun_out/hash_base.log && echo BITS_SAME_AS_BASE
CASES="65536:1:100 131072:1:30 262144:1:10"
for rep in 1 2; do
timeout 300 python tools/quick_bench.py $CASES > gpurun_out/quick_new$rep.log 2>&1; echo "quick new rc=$?"
NBODY_GPU_LIB=$BASE timeout 300 python tools/quick_bench.py $CASES > gpurun_out/quick_base$rep.log 2>&1; echo "quick base rc=$?"
tail -n 3 gpurun_out/quick_new$rep.log gpurun_out/quick_base$rep.log
done
for v in 0 7; do
  for lib in $DEVNEW $DEVBASE; do
    echo "8-way emulation vrank $v $(basename $lib): $(NBODY_GPU_LIB=$lib NBODY_DEV_VRANK=$v timeout 120 python tools/vrank_times.py one 262144 2>&1 | tail -n 1)"
  done
done 2>&1 | tee gpurun_out/vrank_ab.log
