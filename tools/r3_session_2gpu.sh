#!/bin/bash
# One gpurun --gpus 2 call (round 3): the whole GPU suite on 2 GPUs (multi-GPU, multi-process and host-driver tests
# included: COMMTIME > 0, nbody-par --gpus 2), then the 2-GPU bench line with the C5 block forced on (development
# override) and the three parity entries, then the exact-mode timing of the 16-row kernel.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest -m gpu tests -x -q -rs > gpurun_out/r3_pytest_gpu_2gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r3_pytest_gpu_2gpu.log
tail -6 gpurun_out/r3_pytest_gpu_2gpu.log
NBODY_BENCH_C5=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r3_bench_2gpu.json 2> gpurun_out/r3_bench_2gpu.err
echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r3_bench_2gpu.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'])
print(json.dumps(d['parity'])[:1500])
print(json.dumps(d.get('configs'))[:1200])
PY
tail -3 gpurun_out/r3_bench_2gpu.err
python tools/quick_bench.py 4096:0:200 2>&1 | grep "n=" | tee gpurun_out/r3_exact4096_rows16.log
