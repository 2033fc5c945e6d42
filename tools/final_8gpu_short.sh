#!/bin/bash
# one gpurun --gpus 8 call, kept short: the 8-GPU bench line of BASELINE config 4 and config 5 (N = 1048576)
set -u
mkdir -p gpurun_out
GPUS="8" STEPS=10 bash tools/scale_run.sh
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus 8 --bodies 1048576 --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/scale_8_N1048576.json 2> gpurun_out/scale_8_N1048576.err
python -c "import json,sys; d=json.load(open(sys.argv[1])); print(d['n_gpus'], d['config']['n'], '%.4g' % d['value'], '%.3f ms/step' % d['ms_per_step'], d['phase_ms_per_step'], 'e2e %.4g' % d['e2e']['value'], d.get('parity'))" gpurun_out/scale_8_N1048576.json || tail -n 5 gpurun_out/scale_8_N1048576.err
