#!/bin/bash
# development aid: the driver's 1->8 scaling run (bench.py at N = 1, 2, 4, 8 GPUs of one box)
mkdir -p gpurun_out
for n in ${GPUS:-1 2 4 8}; do
    if [ "$n" = 1 ]; then
        python bench.py --gpus 1 --steps ${STEPS:-10} --warmup 3 --no-cpu-baseline > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
    else
        python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n \
            bench.py --gpus $n --steps ${STEPS:-10} --warmup 3 > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
    fi
    python -c "import json,sys; d=json.load(open(sys.argv[1])); print(d['n_gpus'], '%.4g' % d['value'], '%.3f ms/step' % d['ms_per_step'], d['phase_ms_per_step'], 'e2e %.4g' % d['e2e']['value'])" gpurun_out/scale_$n.json || tail -5 gpurun_out/scale_$n.err
done
