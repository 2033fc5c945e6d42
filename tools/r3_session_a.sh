#!/bin/bash
# One gpurun call (round 3): GPU tests of the build, the 1-GPU bench line, then ncu of the exact-mode kernels
# (forces_exact_wide_kernel at N = 4096, small_world_kernel at 32 and 128 bodies) -- plain runs first.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -s > gpurun_out/r3_pytest_gpu_3.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r3_pytest_gpu_3.log
grep -E "radius x|passed|failed|rc=" gpurun_out/r3_pytest_gpu_3.log | tail -8
python bench.py > gpurun_out/r3_bench_1gpu_b.json 2> gpurun_out/r3_bench_1gpu_b.err || { echo "bench failed"; exit 1; }
python -c "import json;d=json.loads(open('gpurun_out/r3_bench_1gpu_b.json').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'],d['e2e']['value'],d['roofline']['frac'])"
EX="python tools/quick_bench.py 4096:0:40"
S32="python tools/quick_bench.py 32:0:1000"
S128="python tools/quick_bench.py 128:0:1000"
$EX > gpurun_out/r3_plain_exact4096.log 2>&1 || { echo "plain exact failed"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:forces_exact_wide -s 4 -c 2 -f -o gpurun_out/r3_prof_exact4096 $EX > gpurun_out/r3_ncu_full_exact4096.log 2>&1
$S32 > gpurun_out/r3_plain_small32.log 2>&1 || { echo "plain small32 failed"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:small_world -s 1 -c 1 -f -o gpurun_out/r3_prof_small32 $S32 > gpurun_out/r3_ncu_full_small32.log 2>&1
$S128 > gpurun_out/r3_plain_small128.log 2>&1 || { echo "plain small128 failed"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:small_world -s 1 -c 1 -f -o gpurun_out/r3_prof_small128 $S128 > gpurun_out/r3_ncu_full_small128.log 2>&1
ls -la gpurun_out/*.ncu-rep
cat gpurun_out/r3_plain_exact4096.log gpurun_out/r3_plain_small32.log gpurun_out/r3_plain_small128.log | grep "n="
tail -n 2 gpurun_out/r3_ncu_full_exact4096.log gpurun_out/r3_ncu_full_small32.log gpurun_out/r3_ncu_full_small128.log
