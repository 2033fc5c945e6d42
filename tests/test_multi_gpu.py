"""Row sharding over several GPUs of one process (cfg.ngpus = K, ncclCommInitAll + ncclAllGather):
results must be bit-identical to the single-GPU run in BOTH modes (each body's sum is formed whole on
one GPU in the single-GPU order; fast-mode chunking depends on N only).  Skipped with < 2 GPUs; the
one-process-per-GPU shape is exercised by bench.py under torchrun and by tests/test_multirank_cpu.py."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(nb, b, mode, ngpus, steps):
    g = nb.NBodyGPU(b.x, b.y, b.xv, b.yv, b.mass, b.radius, mode=mode, ngpus=ngpus)
    g.step(steps)
    out = g.fetch()
    g.close()
    return out


@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("n,steps", [(128, 50), (1001, 8), (4096, 3), (5, 20)])
def test_sharded_equals_single_gpu_bits(nb, oracle, mode, n, steps):
    k = nb.device_count()
    if k < 2:
        pytest.skip("needs >= 2 GPUs")
    b = oracle.init_bodies(n)
    ref = _run(nb, b, mode, 1, steps)
    for p in sorted({2, min(k, 3), k}):
        got = _run(nb, b, mode, p, steps)
        for a, c in zip(ref, got):
            assert np.array_equal(a.view(np.uint64), c.view(np.uint64)), (mode, n, p)


@pytest.mark.parametrize("n,steps", [(3000, 5), (20000, 2), (777, 6)])
def test_symmetric_kernel_same_bits_on_1_2_4_8_gpus(nb, oracle, n, steps):
    """The symmetric fast kernel sums its partial results through 8 fixed virtual ranks, so the bits do
    not depend on how many GPUs (1, 2, 4 or 8) share them; other GPU counts use the directed kernel."""
    k = nb.device_count()
    if k < 2:
        pytest.skip("needs >= 2 GPUs")
    b = oracle.init_bodies(n)
    b.x[n - 1], b.y[n - 1] = b.x[1], b.y[1]          # a coincident pair across row blocks
    def run(p, p2p=0):
        g = nb.NBodyGPU(b.x, b.y, b.xv, b.yv, b.mass, b.radius, mode=1, ngpus=p, symmetric=1, p2p=p2p)
        g.step(steps)
        out = g.fetch()
        if p > 1:
            want = "ncclAllGather" if p2p < 0 else "peer-memory"
            assert want in g.describe()["exchange"], g.describe()
        g.close()
        return out
    ref = run(1)
    chk = b.copy()
    oracle.run(chk, steps, trig=0, threads=8)
    np.testing.assert_allclose(ref[0], chk.x, rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(ref[4], chk.xv, rtol=1e-10, atol=1e-12)
    for p in (2, 4, 8):
        if p > k:
            continue
        for p2p in (1, -1):   # fused peer-memory exchange, and the NCCL all-gather path
            got = run(p, p2p)
            for a, c in zip(ref, got):
                assert np.array_equal(a.view(np.uint64), c.view(np.uint64)), (n, p, p2p)


def test_sharded_exact_matches_oracle_and_golden(nb, oracle):
    k = nb.device_count()
    if k < 2:
        pytest.skip("needs >= 2 GPUs")
    b = oracle.init_bodies(128)
    got = _run(nb, b, 0, k, 1000)
    assert oracle.format_state(*got) == oracle.golden_text("128_MY_REF_OUTPUT_128x1000.txt")
    oracle.run(b, 1000, trig=1)
    for a, f in zip(got, ("x", "y", "xf", "yf", "xv", "yv")):
        assert np.array_equal(a.view(np.uint64), getattr(b, f).view(np.uint64)), f


def test_forces_only_sharded(nb, oracle):
    k = nb.device_count()
    if k < 2:
        pytest.skip("needs >= 2 GPUs")
    b = oracle.init_bodies(3000)
    g1 = nb.NBodyGPU(b.x, b.y, b.xv, b.yv, b.mass, b.radius, mode=1, ngpus=1)
    g2 = nb.NBodyGPU(b.x, b.y, b.xv, b.yv, b.mass, b.radius, mode=1, ngpus=k)
    a, c = g1.forces(), g2.forces()
    assert np.array_equal(a[0], c[0]) and np.array_equal(a[1], c[1])
