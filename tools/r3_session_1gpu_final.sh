#!/bin/bash
# One gpurun call (round 3, final build): the GPU suite, the 1-GPU bench line, the exact-mode timing at N = 4096.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest -m gpu tests -x -q -rs > gpurun_out/r3_pytest_gpu_final.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r3_pytest_gpu_final.log
tail -4 gpurun_out/r3_pytest_gpu_final.log
python bench.py > gpurun_out/r3_bench_1gpu_final.json 2> gpurun_out/r3_bench_1gpu_final.err; echo "bench rc=$?"
python -c "import json;d=json.loads(open('gpurun_out/r3_bench_1gpu_final.json').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'],d['e2e']['value'],d['roofline']['frac'],d['configs']['c2_4096x1000_exact'])"
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')"
