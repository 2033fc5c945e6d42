"""Host logic of the symmetric force kernel's dispatch order (no GPU): nbody_gpu_sym_work_items returns the very
table the device path uploads (one CTA per (row block, j-chunk) item).  The kernel adds the reaction sums of one
(virtual rank, 1024-body segment) to `vacc` in ascending row-block order, every CTA waiting for the previous block
of its chain (forces_sym.cuh, turn_wait) -- the hand-over a serial split-K reduction uses.  That is free of
deadlock only if the CTA waited for is dispatched EARLIER (lower CTA index), and it fixes the order of the sums --
the reference's order of touching XF(c) in the triangular loop (nbody-seq.c:87-107) is a fixed one too -- only if
every chain is complete.  Checked here for every rank of 1, 2, 4, 8 GPUs."""
import numpy as np
import pytest

import pnb200

RB, SEG, V = 512, 1024, 8


def vrank(B):
    r = B % (2 * V)
    return r if r < V else 2 * V - 1 - r


def vseq(B):
    return 2 * (B // (2 * V)) + (0 if B % (2 * V) < V else 1)


@pytest.mark.parametrize("n", [8192, 20000, 65536, 100001, 262144, 300000, 1048576])
@pytest.mark.parametrize("nranks", [1, 2, 4, 8])
def test_dispatch_order_feeds_every_chain_in_block_order(n, nranks):
    nb = pnb200.load()
    if n == 1048576 and nranks not in (1, 8):
        pytest.skip("the large case on 1 and 8 ranks only")
    nblocks = -(-n // RB)
    seen_blocks = set()
    for rank in range(nranks):
        w = nb.sym_work_items(n, nranks, rank)
        blocks, chunks, start = w["blocks"], w["chunks"], w["chunk_start"]
        nchunks = len(start) - 1
        assert start[0] == 0 and start[-1] == n and np.all(np.diff(start) > 0)
        nv = V // nranks
        owned = [B for B in range(nblocks) if rank * nv <= vrank(B) < (rank + 1) * nv]
        seen_blocks.update(owned)
        # exactly the live items, each once
        want = {(B, s) for B in owned for s in range(nchunks) if start[s + 1] > B * RB}
        got = list(zip(blocks.tolist(), chunks.tolist()))
        assert len(got) == len(set(got)) and set(got) == want
        index = {it: k for k, it in enumerate(got)}
        # long chunks first, the one-tile chunks last (they fill the tail of the launch)
        lens = np.diff(start)[chunks]
        first_short = np.argmax(lens < lens[0]) if np.any(lens < lens[0]) else len(lens)
        assert np.all(lens[:first_short] == lens[0]) and np.all(lens[first_short:] <= 256)
        # chains: for every item and every segment of its chunk with bodies above the block, the previous block of the
        # same virtual rank has an item for the same chunk, earlier in dispatch order
        prev_of = {}
        last = {}
        for B in owned:
            prev_of[B] = last.get(vrank(B))
            last[vrank(B)] = B
        min_gap = None
        for k, (B, s) in enumerate(got):
            c0, c1 = int(start[s]), int(start[s + 1])
            for sstart in range(c0, c1, SEG):
                send = min(c1, sstart + SEG)
                lo = max(sstart, (B + 1) * RB)
                if lo >= send:
                    continue
                Bp = prev_of[B]
                if Bp is None:
                    assert vseq(B) == 0
                    continue
                assert vseq(Bp) == vseq(B) - 1
                # the previous block contributes to this segment too (its rows end lower), through an earlier CTA
                assert max(sstart, (Bp + 1) * RB) < send
                kp = index[(Bp, s)]
                assert kp < k, (n, nranks, rank, B, s, Bp)
                gap = k - kp
                min_gap = gap if min_gap is None else min(min_gap, gap)
        assert min_gap is None or min_gap >= 1
        # row-sum groups: a row block's long-chunk items and its short-chunk items, each a run of consecutive items
        # with the chunks ascending (the fold adds the slices in that order)
        groups, guards, ring = w["groups"], w["guards"], w["ring"]
        ngroups = int(groups.max()) + 1
        first = np.full(ngroups, -1)
        last = np.full(ngroups, -1)
        for k, g in enumerate(groups.tolist()):
            if first[g] < 0:
                first[g] = k
            assert last[g] in (-1, k - 1), "group %d is not a run of consecutive items" % g
            last[g] = k
            if k > first[g]:
                assert blocks[k] == blocks[k - 1] and chunks[k] == chunks[k - 1] + 1
            assert (lens[k] < lens[0]) == (lens[first[g]] < lens[0])      # long and short chunks do not mix
        assert np.all(np.diff(first) > 0)                                   # group ids in dispatch order
        # ring of slices: every item but its group's last (which folds) stores into slot k % ring; the guard is the
        # group of the slot's previous tenant, and that group's fold (its last item) comes earlier in dispatch order
        tenant = {}
        for k, g in enumerate(groups.tolist()):
            if k == last[g]:
                assert guards[k] == -1
                continue
            want = tenant.get(k % ring, -1)
            assert guards[k] == want
            if want >= 0:
                assert last[want] < k and want != g
            tenant[k % ring] = g
    assert seen_blocks == set(range(nblocks))


def test_sym_work_items_rejects_bad_arguments():
    nb = pnb200.load()
    with pytest.raises(nb.NBodyGPUError):
        nb.sym_work_items(65536, 3, 0)      # 3 does not divide the 8 virtual ranks
    with pytest.raises(nb.NBodyGPUError):
        nb.sym_work_items(65536, 2, 2)


def test_groups_longer_than_the_slice_ring_are_refused():
    """A row block with more items than the ring has slots would wait for its own fold (item k + ring shares item
    k's slot and group): the library must not build such a table.  With the default split that is n above ~16.7
    million; the plan then names the directed kernel, and the table is refused."""
    nb = pnb200.load()
    ring = 2048
    n_ok, n_big = ring * 8192, (ring + 2) * 8192
    assert nb.shard_plan(n_ok, nb.MODE_FAST, 8, 0)["symmetric"]
    assert nb.shard_plan(n_ok, nb.MODE_FAST, 8, 0)["chunks"] - nb.shard_plan(n_ok, nb.MODE_FAST, 8, 0)["long_chunks"] <= ring
    assert not nb.shard_plan(n_big, nb.MODE_FAST, 8, 0)["symmetric"]
    with pytest.raises(nb.NBodyGPUError, match="slice ring"):
        nb.sym_work_items(n_big, 8, 0)


@pytest.mark.gpu
def test_init_refuses_a_split_that_wraps_the_slice_ring():
    """cfg.jsplit beyond the ring size with the symmetric kernel insisted on: EINVAL at init instead of a step that
    spins until its timeout; without the insistence the handle takes the directed kernel."""
    nb = pnb200.load()
    n = 4096 * 512
    x, y, xv, yv, m, r = nb.reference_init(n)
    with pytest.raises(nb.NBodyGPUError, match="ring"):
        nb.NBodyGPU(x, y, xv, yv, m, r, mode=nb.MODE_FAST, device=0, symmetric=1, jsplit=4096)
