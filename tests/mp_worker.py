"""Worker of tests/test_multi_process.py and of bench.py's `parity` block: ONE RANK of a one-process-per-GPU job
(the shape nbody-par has under MPI and the driver's scaling run has under torchrun), launched by
`python -m torch.distributed.run --nproc-per-node P tests/mp_worker.py ...`.

Every rank creates the sharded handle (cfg.rank / nranks / nccl_id; exchange over CUDA-IPC peer memory or NCCL),
runs  step(k1) -> fetch -> set_state(fetched) -> step(k2) -> fetch  (which covers the per-step exchange, the
cell-order re-sort with its completion of the sharded velocities, the own-rows upload and the fetch), and every
rank's result is compared bit for bit with the same sequence on a plain one-GPU handle.  Rank 0 prints one JSON
line."""
import argparse
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")


def state_hash(arrays):
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def run_sequence(g, k1, k2):
    g.step(k1)
    first = g.fetch()
    g.set_state(first[0], first[1], first[4], first[5])
    g.step(k2)
    return first, g.fetch()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bodies", type=int, default=20000)
    ap.add_argument("--mode", type=int, default=1)
    ap.add_argument("--p2p", type=int, default=0)
    ap.add_argument("--k1", type=int, default=3)
    ap.add_argument("--k2", type=int, default=2)
    ap.add_argument("--sort-interval", type=int, default=0)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    import pnb200
    nb = pnb200.load()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    idt = torch.zeros(nb.NCCL_ID_BYTES, dtype=torch.uint8, device="cuda")
    if rank == 0:
        idt.copy_(torch.frombuffer(bytearray(nb.nccl_unique_id()), dtype=torch.uint8))
    dist.broadcast(idt, 0)
    nccl_id = bytes(idt.cpu().numpy().tobytes())

    n = args.bodies
    x, y, xv, yv, mass, radius = nb.reference_init(n)
    x[n - 1], y[n - 1] = x[1], y[1]          # a coincident pair across row blocks
    if args.sort_interval:
        xv *= 3.0e4                          # bodies really move between re-sorts (DELTA_T is 5e-6)
        yv *= 3.0e4
    kw = dict(mode=args.mode, sort_interval=args.sort_interval)
    g = nb.NBodyGPU(x, y, xv, yv, mass, radius, device=local, rank=rank, nranks=world, nccl_id=nccl_id, p2p=args.p2p, **kw)
    desc = g.describe()
    first, last = run_sequence(g, args.k1, args.k2)
    g.close()
    # the same sequence on ONE GPU (every rank does it on its own device: every rank's copy is checked)
    g1 = nb.NBodyGPU(x, y, xv, yv, mass, radius, device=local, **kw)
    first1, last1 = run_sequence(g1, args.k1, args.k2)
    g1.close()
    same = all(np.array_equal(a.view(np.uint64), b.view(np.uint64)) for a, b in zip(first + last, first1 + last1))
    flag = torch.tensor([1 if same else 0], dtype=torch.int64, device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(json.dumps({"world": world, "n": n, "mode": args.mode, "p2p": args.p2p, "exchange": desc.get("exchange"),
                          "graph_steps": desc.get("graph_steps"), "kernel": desc.get("kernel"),
                          "bits_equal_to_1gpu_on_every_rank": bool(flag.item()),
                          "sha256_sharded": state_hash(last), "sha256_1gpu": state_hash(last1)}), flush=True)
    dist.destroy_process_group()
    return 0 if flag.item() else 1


if __name__ == "__main__":
    sys.exit(main())
