"""world_size-2 gloo test of the multi-rank host logic (no GPU): the NCCL-id style bootstrap broadcast,
the row partition, and the all-gather-per-step structure -- each rank advances only its own rows with the
oracle's row form, gathers the new positions, and the result must equal the single-rank oracle bit for
bit (the property that makes exact mode independent of the GPU count)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, steps, out_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle_lib
    import pnb200
    nb = pnb200.load()
    # bootstrap: rank 0 creates an opaque 128-byte id, everyone receives the same bytes
    idt = torch.zeros(nb.NCCL_ID_BYTES, dtype=torch.uint8)
    if rank == 0:
        idt.copy_(torch.arange(nb.NCCL_ID_BYTES, dtype=torch.uint8))
    dist.broadcast(idt, 0)
    assert idt.tolist() == list(range(nb.NCCL_ID_BYTES))

    L = oracle_lib.lib()
    b = oracle_lib.init_bodies(n)               # every rank has the full initial state (MPI_Bcast)
    lo, hi = nb.partition(n, world, rank)
    rows = -(-n // world)
    for _ in range(steps):
        xf, yf = np.zeros(n), np.zeros(n)
        L.oracle_forces_rows(n, 1, lo, hi, 1, b.x, b.y, b.mass, b.radius, xf, yf)
        b.xf[lo:hi], b.yf[lo:hi] = xf[lo:hi], yf[lo:hi]
        cnt = hi - lo
        xv, yv = b.xv[lo:hi].copy(), b.yv[lo:hi].copy()
        L.oracle_velocities(cnt, 1, np.ascontiguousarray(xf[lo:hi]), np.ascontiguousarray(yf[lo:hi]),
                            np.ascontiguousarray(b.mass[lo:hi]), xv, yv)
        xn, yn = np.zeros(cnt), np.zeros(cnt)
        L.oracle_positions(cnt, 1024, 768, np.ascontiguousarray(b.x[lo:hi]), np.ascontiguousarray(b.y[lo:hi]),
                           xv, yv, xn, yn)
        b.xv[lo:hi], b.yv[lo:hi] = xv, yv
        # padded equal-count all-gather of the new positions (what ncclAllGather does on the GPUs)
        send = torch.zeros(rows, 2, dtype=torch.float64)
        send[:cnt, 0] = torch.from_numpy(xn)
        send[:cnt, 1] = torch.from_numpy(yn)
        recv = [torch.zeros(rows, 2, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(recv, send)
        full = torch.cat(recv)[:n].numpy()
        b.x[:], b.y[:] = full[:, 0], full[:, 1]
    # gather velocities and forces at the end (nbody_gpu_fetch in multi-rank mode)
    for arr in (b.xv, b.yv, b.xf, b.yf):
        send = torch.zeros(rows, dtype=torch.float64)
        send[:hi - lo] = torch.from_numpy(arr[lo:hi].copy())
        recv = [torch.zeros(rows, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(recv, send)
        arr[:] = torch.cat(recv)[:n].numpy()
    if rank == 0:
        np.savez(out_path, x=b.x, y=b.y, xf=b.xf, yf=b.yf, xv=b.xv, yv=b.yv)
    dist.destroy_process_group()


@pytest.mark.parametrize("n", [96, 101])
def test_two_rank_row_sharding_matches_single_rank_oracle(tmp_path, oracle, n):
    steps = 6
    out = str(tmp_path / "r0.npz")
    mp.spawn(_worker, args=(2, _free_port(), n, steps, out), nprocs=2, join=True)
    got = np.load(out)
    ref = oracle.init_bodies(n)
    oracle.run(ref, steps, trig=1)
    for f in ("x", "y", "xf", "yf", "xv", "yv"):
        assert np.array_equal(got[f].view(np.uint64), getattr(ref, f).view(np.uint64)), f
