#!/bin/bash
# development aid: per-rank phase times of the sharded step at 2 and 4 GPUs (peer exchange and NCCL exchange)
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.log 2>&1
for cfg in "2 0" "2 -1" "4 0" "4 -1"; do
    set -- $cfg
    NBODY_BENCH_P2P=$2 NBODY_PROFILE_SPLIT=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 \
        --master-port 2952$1 bench.py --gpus $1 --steps 4 --warmup 3 --no-e2e > gpurun_out/diag_$1_$2.json 2> gpurun_out/diag_$1_$2.err
    echo "== gpus $1 p2p $2"
    python -c "import json,sys; d=json.load(open(sys.argv[1])); print('%.4g' % d['value'], d['phase_ms_per_step'])" gpurun_out/diag_$1_$2.json
    grep -a "nbody_gpu\]" gpurun_out/diag_$1_$2.err | sed 's/.*\[nbody_gpu\]/[nbody_gpu]/' | grep "step [23]:" | sort -k3,3n -k5,5n
done
