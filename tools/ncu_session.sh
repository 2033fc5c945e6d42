#!/bin/bash
# One gpurun call: plain runs first (each must exit 0), then the ncu passes of profiles/README.md (round 2).
set -u
mkdir -p gpurun_out
C4="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-configs"
C3="python bench.py --bodies 65536 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-configs"
DEVLIB=$PWD/parallel-nbody_b200/libnbody_gpu_dev.so
$C4 > gpurun_out/r2_plain_c4.json 2> gpurun_out/r2_plain_c4.err || { echo "plain C4 failed"; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_c4.csv $C4 > gpurun_out/r2_ncu_launches_c4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:forces_sym -s 4 -c 2 -f -o gpurun_out/r2_prof_c4 $C4 > gpurun_out/r2_ncu_full_c4.log 2>&1
$C3 > gpurun_out/r2_plain_c3.json 2> gpurun_out/r2_plain_c3.err || { echo "plain C3 failed"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:forces_sym -s 4 -c 2 -f -o gpurun_out/r2_prof_c3 $C3 > gpurun_out/r2_ncu_full_c3.log 2>&1
NBODY_GPU_LIB=$DEVLIB NBODY_DEV_VRANK=3 python tools/vrank_times.py one 262144 > gpurun_out/r2_plain_8way.json 2> gpurun_out/r2_plain_8way.err || { echo "plain 8way failed"; exit 1; }
NBODY_GPU_LIB=$DEVLIB NBODY_DEV_VRANK=3 ncu --set full --clock-control none --import-source on -k regex:forces_sym -s 3 -c 1 -f -o gpurun_out/r2_prof_8way python tools/vrank_times.py one 262144 > gpurun_out/r2_ncu_full_8way.log 2>&1
ls -la gpurun_out/*.ncu-rep
tail -n 3 gpurun_out/r2_ncu_full_c4.log gpurun_out/r2_ncu_full_c3.log gpurun_out/r2_ncu_full_8way.log
