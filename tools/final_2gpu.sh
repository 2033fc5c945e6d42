#!/bin/bash
# one gpurun --gpus 2 call: the multi-GPU parity tests (one process driving 2 devices, and one process per GPU with
# CUDA IPC / NCCL), then the 2-GPU bench line
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multi_gpu.py tests/test_multi_process.py tests/test_cell_order.py -m gpu -q --durations=3 > gpurun_out/pytest_2gpu.log 2>&1; echo "pytest rc=$?"; tail -n 6 gpurun_out/pytest_2gpu.log
GPUS=2 STEPS=10 bash tools/scale_run.sh
