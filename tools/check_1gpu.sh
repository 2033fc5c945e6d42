#!/bin/bash
# one gpurun call on one B200: GPU test-suite, smoke, the bench line (no ncu)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --durations=5 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 10 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/bench_1gpu.json").read().strip().splitlines()[-1])
print("bench value %.4e ms/step %.3f e2e %.4e traffic %s frac %.3f pipe %.3f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["traffic"], d["roofline"]["frac"], d["roofline"]["pipe_frac"]))
print(d["config"]["l2"])
PY
