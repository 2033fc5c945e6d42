"""GPU parity tests: the CUDA path, called through the C ABI, against the oracle.

EXACT mode bar: bit-identical (uint64 view) to the oracle with detmath trig -- which
tests/test_oracle.py pins bit-for-bit to the reference source compiled against the same
header -- and byte-identical text against the reference's two golden outputs.
FAST mode bar (north_star: "within a stated max relative error on final positions"):
positions 1e-12, velocities 1e-10 relative; forces 1e-8 * max|F| absolute, against the
oracle with the C library's trig (what the reference links).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

FIELDS = ("x", "y", "xf", "yf", "xv", "yv")
POS_RTOL = 1e-12
VEL_RTOL = 1e-10
FRC_RTOL_OF_MAX = 1e-8


def make(nb, b, **kw):
    return nb.NBodyGPU(b.x, b.y, b.xv, b.yv, b.mass, b.radius, **kw)


def assert_bits(got, b, what=""):
    for g, f in zip(got, FIELDS):
        want = getattr(b, f)
        same = g.view(np.uint64) == want.view(np.uint64)
        assert same.all(), "%s %s: %d of %d differ, first at %d: got %r want %r" % (
            what, f, (~same).sum(), same.size, int(np.argmin(same)), g[np.argmin(same)], want[np.argmin(same)])


def assert_fast_close(got, b, what=""):
    x, y, xf, yf, xv, yv = got
    for g, w in ((x, b.x), (y, b.y)):
        np.testing.assert_allclose(g, w, rtol=POS_RTOL, atol=1e-12, err_msg=what + " position")
    for g, w in ((xv, b.xv), (yv, b.yv)):
        np.testing.assert_allclose(g, w, rtol=VEL_RTOL, atol=1e-12, err_msg=what + " velocity")
    fmax = max(np.abs(b.xf).max(), np.abs(b.yf).max(), 1.0)
    for g, w in ((xf, b.xf), (yf, b.yf)):
        assert np.abs(g - w).max() <= FRC_RTOL_OF_MAX * fmax, what + " force"


@pytest.mark.parametrize("n,steps", [(2, 50), (32, 500), (127, 40), (128, 1000), (129, 40), (257, 30),
                                     (777, 20), (1000, 10), (2048, 4)])
def test_exact_bit_identical_to_oracle(nb, oracle, n, steps):
    b = oracle.init_bodies(n)
    g = make(nb, b, mode=nb.MODE_EXACT)
    g.step(steps)
    got = g.fetch()
    oracle.run(b, steps, trig=1)
    assert_bits(got, b, "n=%d steps=%d" % (n, steps))


def test_exact_golden_128x1000_text(nb, oracle):
    """BASELINE config 1 / bin/tests.sh:34-41: 128 bodies x 1000 steps vs tests/128_MY_REF_OUTPUT."""
    b = oracle.init_bodies(128)
    g = make(nb, b, mode=nb.MODE_EXACT)
    g.step(1000)
    assert oracle.format_state(*g.fetch()) == oracle.golden_text("128_MY_REF_OUTPUT_128x1000.txt")


def test_exact_golden_32x100000_text(nb, oracle):
    """bin/nbody-sanity-check:16,28: 32 bodies x 100000 steps vs bin/REF_OUTPUT."""
    b = oracle.init_bodies(32)
    g = make(nb, b, mode=nb.MODE_EXACT)
    g.step(100000)
    assert oracle.format_state(*g.fetch()) == oracle.golden_text("REF_OUTPUT_32x100000.txt")


def test_fast_goldens_text(nb, oracle):
    for n, steps, name in ((128, 1000, "128_MY_REF_OUTPUT_128x1000.txt"), (32, 100000, "REF_OUTPUT_32x100000.txt")):
        b = oracle.init_bodies(n)
        g = make(nb, b, mode=nb.MODE_FAST)
        g.step(steps)
        assert oracle.format_state(*g.fetch()) == oracle.golden_text(name), name


def test_exact_4096_config2_prefix_and_graph_independence(nb, oracle):
    """BASELINE config 2 (N=4096, exact). The CPU oracle needs ~0.5 s per step at this size, so the
    bit comparison covers the first 12 steps; the full 1000 steps are checked for independence
    of how the step loop is issued (graph vs plain launches, chunked calls)."""
    n = 4096
    b = oracle.init_bodies(n)
    g = make(nb, b, mode=nb.MODE_EXACT)
    g.step(12)
    got = g.fetch()
    ref = b.copy()
    oracle.run(ref, 12, trig=1, threads=8)
    assert_bits(got, ref, "n=4096 steps=12")
    g.step(988)
    full = g.fetch()
    g2 = make(nb, b, mode=nb.MODE_EXACT, use_graph=-1)
    for k in (1, 2, 3, 494, 500):
        g2.step(k)
    full2 = g2.fetch()
    for a, c in zip(full, full2):
        assert np.array_equal(a.view(np.uint64), c.view(np.uint64))


@pytest.mark.parametrize("symmetric", [-1, 1])
@pytest.mark.parametrize("n,steps", [(2, 50), (128, 1000), (777, 20), (2048, 4), (4096, 6)])
def test_fast_within_tolerance_of_reference_math(nb, oracle, n, steps, symmetric):
    """symmetric = -1: directed kernel; 1: each unordered pair evaluated once (+f to i, -f to j)."""
    b = oracle.init_bodies(n)
    g = make(nb, b, mode=nb.MODE_FAST, symmetric=symmetric)
    g.step(steps)
    got = g.fetch()
    oracle.run(b, steps, trig=0, threads=8)
    assert_fast_close(got, b, "n=%d" % n)


def test_fast_jsplit_variants_agree(nb, oracle):
    b = oracle.init_bodies(3000)
    ref = b.copy()
    oracle.run(ref, 3, trig=0, threads=8)
    for sym in (-1, 1):
        for js in (1, 2, 3, 7, 11):
            g = make(nb, b, mode=nb.MODE_FAST, jsplit=js, symmetric=sym)
            g.step(3)
            assert_fast_close(g.fetch(), ref, "jsplit=%d sym=%d" % (js, sym))


def _kat_bodies(oracle):
    """Hand-made cases the reference has no fixture for: coincident bodies (atan2(0,0) convention),
    clamped distance (dsqr < mindsqr), dy = 0 with dx < 0 (sin(pi) != 0), far pairs."""
    xs = [100.0, 100.0, 100.0, 50.0, 150.0, 100.0, 900.5, 900.25, 3.0, 1000.0]
    ys = [100.0, 100.0, 100.0, 100.0, 100.0, 700.0, 300.125, 300.5, 3.0, 760.0]
    n = len(xs)
    b = oracle.Bodies(n)
    b.x[:] = xs
    b.y[:] = ys
    b.radius[:] = np.linspace(1.0, 40.0, n)
    b.mass[:] = b.radius ** 3
    b.xv[:] = np.linspace(-4, 4, n)
    b.yv[:] = np.linspace(3, -3, n)
    return b


def test_forces_only_known_cases(nb, oracle):
    b = _kat_bodies(oracle)
    L = oracle.lib()
    xf, yf = np.zeros(b.n), np.zeros(b.n)
    L.oracle_forces_symmetric(b.n, 1, b.x, b.y, b.mass, b.radius, xf, yf)
    gx, gy = make(nb, b, mode=nb.MODE_EXACT).forces()
    assert np.array_equal(gx.view(np.uint64), xf.view(np.uint64))
    assert np.array_equal(gy.view(np.uint64), yf.view(np.uint64))
    # coincident triple 0,1,2: lower index is pushed +x by each higher coincident body
    L.oracle_forces_symmetric(b.n, 0, b.x, b.y, b.mass, b.radius, xf, yf)
    scale = max(np.abs(xf).max(), np.abs(yf).max())
    for sym in (-1, 1):
        fx, fy = make(nb, b, mode=nb.MODE_FAST, symmetric=sym).forces()
        assert np.abs(fx - xf).max() <= 1e-12 * scale, sym
        assert np.abs(fy - yf).max() <= 1e-12 * scale, sym


@pytest.mark.parametrize("symmetric", [-1, 1])
def test_fast_coincident_bodies_across_blocks(nb, oracle, symmetric):
    """Exactly coincident bodies far apart in index (different row blocks and tiles of the symmetric
    kernel), including a triple, and a coincident pair inside one block: the reference's
    atan2(0,0) = 0 convention gives the lower index (+force, 0), the higher (-force, 0)."""
    n = 3000
    b = oracle.init_bodies(n)
    for i, j in ((3, 2900), (3, 1500), (700, 701), (511, 512), (1023, 2047), (0, n - 1)):
        b.x[j], b.y[j] = b.x[i], b.y[i]
    xf, yf = np.zeros(n), np.zeros(n)
    oracle.lib().oracle_forces_symmetric(n, 0, b.x, b.y, b.mass, b.radius, xf, yf)
    fx, fy = make(nb, b, mode=nb.MODE_FAST, symmetric=symmetric).forces()
    scale = max(np.abs(xf).max(), np.abs(yf).max())
    assert np.abs(fx - xf).max() <= 1e-12 * scale
    assert np.abs(fy - yf).max() <= 1e-12 * scale


@pytest.mark.parametrize("mode", [0, 1])
def test_wall_bounce_all_edges(nb, oracle, mode):
    """compute_positions edge cases (nbody-seq.c:145-159): each wall, both corners, exact hits."""
    xdim, ydim = 64, 48
    pts = [(0.0, 10.0, -5.0, 0.0), (63.99999, 10.0, 5.0, 0.0), (10.0, 0.0, 0.0, -5.0), (10.0, 47.99999, 0.0, 5.0),
           (0.0, 0.0, -4.0, -4.0), (63.99999, 47.99999, 4.0, 4.0), (32.0, 24.0, 1.0, -1.0), (63.0, 47.0, 0.0, 0.0)]
    n = len(pts)
    b = oracle.Bodies(n)
    for k, (x, y, xv, yv) in enumerate(pts):
        b.x[k], b.y[k], b.xv[k], b.yv[k] = x, y, xv, yv
    b.radius[:] = 1.0
    b.mass[:] = 1.0
    g = nb.NBodyGPU(b.x, b.y, b.xv, b.yv, b.mass, b.radius, xdim=xdim, ydim=ydim, mode=mode)
    g.step(3)
    got = g.fetch()
    oracle.run(b, 3, trig=1 if mode == 0 else 0, xdim=xdim, ydim=ydim)
    if mode == 0:
        assert_bits(got, b, "bounce")
    else:
        assert_fast_close(got, b, "bounce")
    assert got[0].min() >= 0 and got[0].max() < xdim


def test_set_state_and_relaunch(nb, oracle):
    b = oracle.init_bodies(300)
    g = make(nb, b, mode=nb.MODE_EXACT)
    g.step(7)
    g.set_state(b.x, b.y, b.xv, b.yv)
    g.step(5)
    got = g.fetch()
    oracle.run(b, 5, trig=1)
    assert_bits(got, b, "after set_state")
    total, per = g.launch_count()
    assert per[0] == 12 and per[1] == 12 and per[3] == 0     # 12 force + 12 integrate launches, no NCCL


def test_large_n_properties_fast(nb, oracle):
    """BASELINE config 3 size (N=65536): too large for the CPU oracle in a test, so check
    size-independent properties: momentum balance of the pair sums (Newton's third law:
    sum of forces ~ 0 up to rounding), bodies stay inside the world, run-to-run determinism,
    and a sampled-row comparison against the oracle's row form."""
    n = 65536
    b = oracle.init_bodies(n)
    g = make(nb, b, mode=nb.MODE_FAST)          # default at this size: the symmetric kernel
    gd = make(nb, b, mode=nb.MODE_FAST, symmetric=-1)
    fxd, fyd = gd.forces()
    fx, fy = g.forces()
    fs = max(np.abs(fxd).max(), np.abs(fyd).max())
    assert np.abs(fx - fxd).max() <= 1e-11 * fs and np.abs(fy - fyd).max() <= 1e-11 * fs
    fx2, fy2 = g.forces()
    assert np.array_equal(fx, fx2) and np.array_equal(fy, fy2)
    scale = np.abs(fx).sum()
    assert abs(fx.sum()) <= 1e-9 * scale and abs(fy.sum()) <= 1e-9 * np.abs(fy).sum()
    rows = 24
    xf, yf = np.zeros(n), np.zeros(n)
    oracle.lib().oracle_forces_rows(n, 0, 1000, 1000 + rows, 8, b.x, b.y, b.mass, b.radius, xf, yf)
    sl = slice(1000, 1000 + rows)
    np.testing.assert_allclose(fx[sl], xf[sl], rtol=1e-10, atol=1e-10 * np.abs(xf[sl]).max())
    np.testing.assert_allclose(fy[sl], yf[sl], rtol=1e-10, atol=1e-10 * np.abs(yf[sl]).max())
    g.step(2)
    x, y, _, _, _, _ = g.fetch()
    assert x.min() >= 0 and x.max() < 1024 and y.min() >= 0 and y.max() < 768


def test_large_n_exact_sampled_rows(nb, oracle):
    n = 20000
    b = oracle.init_bodies(n)
    gx, gy = make(nb, b, mode=nb.MODE_EXACT).forces()
    xf, yf = np.zeros(n), np.zeros(n)
    for lo in (0, 9999, n - 16):
        oracle.lib().oracle_forces_rows(n, 1, lo, lo + 16, 8, b.x, b.y, b.mass, b.radius, xf, yf)
        sl = slice(lo, lo + 16)
        assert np.array_equal(gx[sl].view(np.uint64), xf[sl].view(np.uint64))
        assert np.array_equal(gy[sl].view(np.uint64), yf[sl].view(np.uint64))


@pytest.mark.parametrize("n,steps", [(128, 200), (32, 0), (3000, 3)])
def test_render_matches_reference_display(nb, oracle, n, steps):
    """display() (nbody-seq.c:326-354) as a CUDA rasteriser: byte-identical to the CPU restatement."""
    xdim, ydim = (1024, 768) if n < 1000 else (320, 200)
    b = oracle.init_bodies(n, xdim, ydim)
    g = nb.NBodyGPU(b.x, b.y, b.xv, b.yv, b.mass, b.radius, xdim=xdim, ydim=ydim, mode=nb.MODE_EXACT)
    g.step(steps)
    img = g.render(xdim, ydim)
    oracle.run(b, steps, trig=1, xdim=xdim, ydim=ydim)
    ref = oracle.display(b, xdim, ydim)
    assert img.shape == ref.shape and np.array_equal(img, ref)
    assert ref.any()     # something was painted


def _rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.mark.parametrize("n,steps", [(65536, 200), (262144, 50)])
def test_baseline_configs_3_and_4_full_size_properties(nb, oracle, n, steps):
    """BASELINE configs 3 (N=65536 x 200 steps) and 4 (N=262144 x 50 steps), fast mode, at FULL size.  The CPU
    oracle would need hours here, so the checks are size-independent: (1) the symmetric kernel (each pair
    once, reactions reduced across lanes/warps/blocks) and the directed kernel (every body sums its own row)
    are two independent summation structures of the same physics and must agree after the whole run to far
    better than the stated tolerance; (2) run-to-run bit determinism; (3) bodies stay inside the world;
    (4) the last step's pair forces obey Newton's third law (sum ~ 0 against sum |F|); (5) a sample of rows of
    the FINAL state's forces is recomputed by the CPU oracle from the final positions."""
    b = oracle.init_bodies(n)
    gs = make(nb, b, mode=nb.MODE_FAST, symmetric=1)
    gd = make(nb, b, mode=nb.MODE_FAST, symmetric=-1)
    gs.step(steps)
    gd.step(steps)
    s = gs.fetch()
    d = gd.fetch()
    assert _rel(s[0], d[0]) <= 1e-12 and _rel(s[1], d[1]) <= 1e-12          # positions
    assert _rel(s[4], d[4]) <= 1e-10 and _rel(s[5], d[5]) <= 1e-10          # velocities
    fscale = max(np.abs(d[2]).max(), np.abs(d[3]).max())
    assert np.abs(s[2] - d[2]).max() <= 1e-8 * fscale and np.abs(s[3] - d[3]).max() <= 1e-8 * fscale
    g2 = make(nb, b, mode=nb.MODE_FAST, symmetric=1)
    g2.step(steps)
    for a, c in zip(s, g2.fetch()):
        assert np.array_equal(a.view(np.uint64), c.view(np.uint64))
    assert s[0].min() >= 0 and s[0].max() < 1024 and s[1].min() >= 0 and s[1].max() < 768
    assert abs(s[2].sum()) <= 1e-9 * np.abs(s[2]).sum() and abs(s[3].sum()) <= 1e-9 * np.abs(s[3]).sum()
    # forces at the final positions (the state after the last step) against the CPU oracle, for the bodies that sit
    # at the edges of EVERY row block of the symmetric kernel (first and last slot of each 512-slot block of the
    # cell order: block-boundary rows, last lane of a rotation group, tiles next to the diagonal, the partial last
    # tile) plus one random slot per block: 3 x N/512 rows x N columns on the CPU
    order = oracle.cell_order(nb, b.x, b.y)          # the order of the only sort of this run (steps < 256)
    rows = oracle.block_sample(order)
    assert len(rows) >= 2 * (n // 512)
    fx, fy = gs.forces()
    xf, yf = oracle.forces_for(rows, s[0], s[1], b.mass, b.radius)
    sc = max(np.abs(xf).max(), np.abs(yf).max())
    assert np.abs(fx[rows] - xf).max() <= 1e-10 * sc and np.abs(fy[rows] - yf).max() <= 1e-10 * sc
    print("n=%d: %d sampled rows, max |dF| / max|F| = %.2e" % (n, len(rows), max(np.abs(fx[rows] - xf).max(), np.abs(fy[rows] - yf).max()) / sc))


@pytest.mark.parametrize("n,steps", [(65536, 200), (262144, 50)])
def test_baseline_configs_3_and_4_sampled_trajectory_against_oracle(nb, oracle, n, steps):
    """Every step of BASELINE configs 3 and 4 at FULL size against the reference's math without O(N^2) CPU work per
    step: each step the oracle computes the pair-force sums of a sample of bodies (block-boundary slots of the
    cell order + random ones) from the GPU's current positions, advances those bodies with the oracle's
    compute_velocities / compute_positions, and the GPU's next state must agree on them: forces 1e-8 max|F|,
    velocities 1e-10, positions 1e-12 (the stated fast-mode tolerance)."""
    b = oracle.init_bodies(n)
    g = make(nb, b, mode=nb.MODE_FAST)
    assert g.describe()["cell_order"] is True
    order = oracle.cell_order(nb, b.x, b.y)
    rng = np.random.default_rng(7)
    edge = oracle.block_sample(order, per_block_extra=0)
    rows = np.unique(np.concatenate([rng.choice(edge, 40, replace=False), rng.integers(0, n, 24),
                                     order[[0, 511, 512, n - 1]]])).astype(np.int32)
    x, y, xv, yv = b.x.copy(), b.y.copy(), b.xv.copy(), b.yv.copy()
    worst = dict(pos=0.0, vel=0.0, frc=0.0)
    for step in range(steps):
        xf, yf = oracle.forces_for(rows, x, y, b.mass, b.radius)
        xn, yn, xvn, yvn = oracle.step_bodies(rows, x, y, xv, yv, xf, yf, b.mass)
        g.step(1)
        x, y, gxf, gyf, xv, yv = g.fetch()
        fmax = max(np.abs(gxf).max(), np.abs(gyf).max())
        worst["frc"] = max(worst["frc"], np.abs(gxf[rows] - xf).max() / fmax, np.abs(gyf[rows] - yf).max() / fmax)
        worst["pos"] = max(worst["pos"], (np.abs(x[rows] - xn) / np.maximum(np.abs(xn), 1.0)).max(),
                           (np.abs(y[rows] - yn) / np.maximum(np.abs(yn), 1.0)).max())
        worst["vel"] = max(worst["vel"], (np.abs(xv[rows] - xvn) / np.maximum(np.abs(xvn), 1e-2)).max(),
                           (np.abs(yv[rows] - yvn) / np.maximum(np.abs(yvn), 1e-2)).max())
        assert worst["frc"] <= FRC_RTOL_OF_MAX and worst["pos"] <= POS_RTOL and worst["vel"] <= VEL_RTOL, (step, worst)
    print("n=%d steps=%d worst one-step deviation from the oracle: %r" % (n, steps, worst))


@pytest.mark.parametrize("n,mode,kw", [(300, 0, {}), (3000, 1, dict(symmetric=1)), (17000, 1, dict(symmetric=1, sort=1))])
def test_fetch_range_and_forces_leave_the_state_alone(nb, oracle, n, mode, kw):
    """nbody_gpu_fetch_range returns the same numbers as the slice of a full fetch (cell order included), rejects
    ranges outside [0, n), and nbody_gpu_forces does not disturb what fetch returns (the LAST STEP's forces)."""
    b = oracle.init_bodies(n)
    g = make(nb, b, mode=mode, **kw)
    g.step(3)
    full = g.fetch()
    for first, count in ((0, n), (0, 1), (n - 1, 1), (n // 3, n // 2), (5, 0)):
        part = g.fetch_range(first, count)
        for a, c in zip(full, part):
            assert np.array_equal(a[first:first + count].view(np.uint64), c.view(np.uint64)), (first, count)
    with pytest.raises(nb.NBodyGPUError):
        g.fetch_range(n - 1, 2)
    fx, fy = g.forces()                       # forces at the CURRENT positions: not the last step's
    assert not np.array_equal(fx, full[2])
    again = g.fetch()
    for a, c in zip(full, again):
        assert np.array_equal(a.view(np.uint64), c.view(np.uint64))
    g.step(2)                                 # and the run goes on as if forces() had not been called
    h = make(nb, b, mode=mode, **kw)
    h.step(5)
    for a, c in zip(g.fetch(), h.fetch()):
        assert np.array_equal(a.view(np.uint64), c.view(np.uint64))


def test_pinned_host_buffers_round_trip(nb, oracle):
    """nbody_gpu_pin_host: set_state / fetch through page-locked caller arrays give the same bits as pageable ones."""
    n = 20000
    b = oracle.init_bodies(n)
    g = make(nb, b, mode=nb.MODE_FAST)
    g.step(2)
    plain = g.fetch()
    host = [np.empty(n) for _ in range(6)]
    for a in host:
        nb.pin_host(a)
    try:
        g.fetch(out=host)
        for a, c in zip(plain, host):
            assert np.array_equal(a.view(np.uint64), c.view(np.uint64))
        g.set_state(host[0], host[1], host[4], host[5])
        g.step(2)
        pinned = g.fetch(out=host)
        h = make(nb, b, mode=nb.MODE_FAST)
        h.step(2)
        mid = h.fetch()
        h.set_state(mid[0], mid[1], mid[4], mid[5])
        h.step(2)
        for a, c in zip(pinned, h.fetch()):
            assert np.array_equal(a.view(np.uint64), c.view(np.uint64))
    finally:
        g.sync()
        for a in host:
            nb.unpin_host(a)


@pytest.mark.parametrize("jsplit", [0, 1, 5])
def test_reaction_sums_handed_from_cta_to_cta(nb, oracle, jsplit):
    """The symmetric kernel adds a row block's reaction sums to vacc[v][j] when it is that block's turn in the chain
    of its (virtual rank, 1024-body segment), and the last CTA of a row block folds the row sums its siblings left in
    the slice ring -- forces_sym.cuh.  Ragged N (last segment, tile and row block are
    partial), chunk shapes with one segment per CTA (jsplit = 0: 1024-body chunks + one-tile chunks) and with many
    (jsplit = 1: one chunk of 41 segments; 5: 9 segments each): forces of block-boundary and random bodies against
    the oracle's row form (nbody-seq.c:87-107), and -- the order of the additions being fixed -- the same bits from
    a second launch on the same handle (the counts and sums are left clean), after steps, and from a second handle."""
    n = 40000 + 37
    b = oracle.init_bodies(n)
    b.x[n - 1], b.y[n - 1] = b.x[5], b.y[5]           # a coincident pair across the whole index range
    g = make(nb, b, mode=nb.MODE_FAST, jsplit=jsplit)
    d = g.describe()
    assert d["symmetric"] and d["cell_order"], d
    # the partial sums are O(N): 8 reaction vectors + 2 row-sum vectors + the fixed ring of slices
    assert d["partial_sum_bytes"] <= 16 * 10 * (n + 8) + 16 * 512 * 4096, d
    order = oracle.cell_order(nb, b.x, b.y)
    rows = np.unique(np.concatenate([oracle.block_sample(order, per_block_extra=1), [0, 5, n - 1]])).astype(np.int32)
    wxf, wyf = oracle.forces_for(rows, b.x, b.y, b.mass, b.radius)
    fx, fy = g.forces()
    full_x, full_y = oracle.forces_for(np.arange(0, n, 53, dtype=np.int32), b.x, b.y, b.mass, b.radius)
    fmax = max(np.abs(wxf).max(), np.abs(wyf).max(), np.abs(full_x).max(), np.abs(full_y).max())
    assert np.abs(fx[rows] - wxf).max() <= FRC_RTOL_OF_MAX * fmax and np.abs(fy[rows] - wyf).max() <= FRC_RTOL_OF_MAX * fmax
    assert np.abs(fx[::53] - full_x).max() <= FRC_RTOL_OF_MAX * fmax and np.abs(fy[::53] - full_y).max() <= FRC_RTOL_OF_MAX * fmax
    fx2, fy2 = g.forces()
    assert np.array_equal(fx.view(np.uint64), fx2.view(np.uint64)) and np.array_equal(fy.view(np.uint64), fy2.view(np.uint64))
    g.step(3)
    h = make(nb, b, mode=nb.MODE_FAST, jsplit=jsplit)
    h.step(1)
    h.step(2)
    for a, c in zip(g.fetch(), h.fetch()):
        assert np.array_equal(a.view(np.uint64), c.view(np.uint64))
    assert abs(fx.sum()) <= 1e-6 * np.abs(fx).sum() and abs(fy.sum()) <= 1e-6 * np.abs(fy).sum()   # Newton's third law


def test_hand_overs_are_race_free_over_many_launches(nb, oracle):
    """A race in the CTA-to-CTA hand-overs of the symmetric kernel (turn counters, slice ring reuse: the ring wraps
    a dozen times per launch at this size, ragged N) would show as run-to-run differences in the last bits.  40
    force launches on one handle, and the same 40 steps on a second handle split differently: identical bits."""
    n = 150000 + 11
    b = oracle.init_bodies(n)
    g = make(nb, b, mode=nb.MODE_FAST)
    h = make(nb, b, mode=nb.MODE_FAST)
    g.step(40)
    for k in (1, 7, 2, 30):
        h.step(k)
    for a, c in zip(g.fetch(), h.fetch()):
        assert np.array_equal(a.view(np.uint64), c.view(np.uint64))
    f1 = g.forces()
    for _ in range(5):
        f2 = g.forces()
        assert np.array_equal(f1[0].view(np.uint64), f2[0].view(np.uint64)) and np.array_equal(f1[1].view(np.uint64), f2[1].view(np.uint64))
