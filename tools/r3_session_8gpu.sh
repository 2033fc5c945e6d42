#!/bin/bash
# One gpurun --gpus 8 call (round 3): the multi-GPU / multi-process / host-driver GPU tests of the committed build,
# then the 8-GPU bench line (with the C5 block and the three parity entries).
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest -m gpu tests/test_multi_gpu.py tests/test_multi_process.py tests/test_cell_order.py tests/test_host_driver.py -q -rs > gpurun_out/r3_pytest_multi_gpu_8gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r3_pytest_multi_gpu_8gpu.log
tail -5 gpurun_out/r3_pytest_multi_gpu_8gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r3_bench_8gpu.json 2> gpurun_out/r3_bench_8gpu.err
echo "bench rc=$?"
tail -c 3000 gpurun_out/r3_bench_8gpu.json
tail -5 gpurun_out/r3_bench_8gpu.err
