"""world_size-2 (and 8) gloo tests of the multi-rank HOST logic of the library itself (no GPU): every rank asks
libnbody_gpu.so how IT shares the bodies (nbody_gpu_shard_plan: the very partition / virtual-rank / row-block code
the device path uses), the ranks exchange their answers with torch.distributed (gloo), and together the answers
must tile the problem exactly once -- rows, virtual ranks, row blocks and pair evaluations -- with the pair
work balanced.  The reference's counterpart is get_recvcounts / get_displacements / get_bounds
(nbody-par.c:225-263), whose rule the library also exports (nbody_gpu_partition_reference).  The device side of
the same path is tests/test_multi_process.py (needs >= 2 GPUs) and bench.py's `parity` block."""
import os
import socket
import sys

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


def _worker(rank, world, port, cases, out_path):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import pnb200
    nb = pnb200.load()
    # bootstrap shape of the multi-process job: rank 0 owns an opaque 128-byte id, everyone receives the same bytes
    idt = torch.zeros(nb.NCCL_ID_BYTES, dtype=torch.uint8)
    if rank == 0:
        idt.copy_(torch.arange(nb.NCCL_ID_BYTES, dtype=torch.uint8))
    dist.broadcast(idt, 0)
    assert idt.tolist() == list(range(nb.NCCL_ID_BYTES))
    ok = True
    for n, mode in cases:
        p = nb.shard_plan(n, mode, world, rank)
        mine = torch.tensor([p["lo"], p["hi"], p["v0"], p["nv"], p["row_blocks"], p["chunks"], p["long_chunks"],
                             int(p["symmetric"]), p["pair_evaluations"]], dtype=torch.int64)
        every = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(every, mine)
        plans = [t.tolist() for t in every]
        # rows: consecutive, disjoint, covering [0, n)
        assert plans[0][0] == 0 and plans[-1][1] == n
        for a, b in zip(plans, plans[1:]):
            assert a[1] == b[0]
        assert plans[rank][:2] == list(nb.partition(n, world, rank))
        # every rank took the same decisions
        assert len({(q[5], q[6], q[7]) for q in plans}) == 1
        total = sum(q[8] for q in plans)
        if plans[0][7]:   # symmetric kernel
            assert [q[2] for q in plans] == [r * (8 // world) for r in range(world)] and all(q[3] == 8 // world for q in plans)
            nblocks = -(-n // 512)
            assert sum(q[4] for q in plans) == nblocks
            # each unordered pair once, except inside a row block, where both directions are evaluated
            tail = n - (nblocks - 1) * 512
            inside = (nblocks - 1) * 512 * 512 + tail * tail
            assert total == (n * n - inside) // 2 + inside
            if n >= 65536:   # many row blocks per virtual rank: the boustrophedon assignment balances the triangle
                worst = max(q[8] for q in plans)
                assert worst <= 1.02 * total / world, (n, world, [q[8] for q in plans])
        else:
            assert total == n * n
        single = nb.shard_plan(n, mode, 1, 0)
        if plans[0][7]:
            assert single["chunks"] == plans[0][5] and single["long_chunks"] == plans[0][6]   # a function of n only
    if rank == 0:
        with open(out_path, "w") as f:
            f.write("ok" if ok else "bad")
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 8])
def test_ranks_tile_the_problem_exactly_once(tmp_path, nb, world):
    out = str(tmp_path / "result.txt")
    cases = [(262144, 1), (65536, 1), (1048576, 1), (20000, 1), (8192 + 77, 1), (4096, 0), (4097, 0), (128, 0), (5, 0), (3000, 1)]
    mp.spawn(_worker, args=(world, _free_port(), cases, out), nprocs=world, join=True)
    assert open(out).read() == "ok"


def test_partition_reference_rule_and_ours_cover_the_same_bodies(nb):
    for n, p in ((128, 8), (10, 4), (1000, 7), (262144, 8)):
        counts, displs = nb.partition_reference(n, p)
        assert sum(counts) == n and displs[0] == 0
        assert sum(nb.partition(n, p, r)[1] - nb.partition(n, p, r)[0] for r in range(p)) == n
