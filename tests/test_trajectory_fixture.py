"""Accumulated-trajectory parity of fast mode against the REFERENCE PROGRAM's own final state.

tests/golden/trajectory_<N>x<STEPS>.npz hold what the reference's nbody-seq.c (compiled where it lies, stock libm;
oracle/_ref/nbody-seq-hex, made by tests/golden/make_trajectory_fixtures.py) prints after a whole run from its own
initialiser: nothing is re-seeded on the way, so this bounds the drift of the fast kernels over the step counts of
BASELINE configs 3 and 4 (200 and 50 steps), not the error of one step (north_star: "within a stated max relative
error on final positions"; loop nbody-seq.c:457-464, surface nbody-seq.c:356-365).
"""
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = [(8192, 200), (16384, 50)]

# Stated tolerances on the FINAL state (the repository's fast-mode gates):
POS_TOL = 1e-12     # |x - x_ref| <= POS_TOL * max(1, |x_ref|)
VEL_TOL = 1e-10     # same form on the velocities
FRC_TOL = 1e-8      # |F - F_ref| <= FRC_TOL * max|F_ref|


def load(n, steps):
    z = np.load(os.path.join(GOLDEN, "trajectory_%dx%d.npz" % (n, steps)))
    assert int(z["n"]) == n and int(z["steps"]) == steps
    return [np.ascontiguousarray(z[k]) for k in ("x", "y", "xf", "yf", "xv", "yv")]


@pytest.mark.parametrize("n,steps", CASES)
def test_fixture_is_the_reference_initialiser_s_world(nb, n, steps):
    """CPU: shape and sanity of the committed fixtures (they came from the reference binary, not from this repo)."""
    x, y, xf, yf, xv, yv = load(n, steps)
    for a in (x, y, xf, yf, xv, yv):
        assert a.shape == (n,) and np.isfinite(a).all()
    assert x.min() >= 0 and x.max() < 1024 and y.min() >= 0 and y.max() < 768     # compute_positions' walls (nbody-seq.c:141-160)


def errors(got, ref):
    x, y, xf, yf, xv, yv = got
    rx, ry, rxf, ryf, rxv, ryv = ref
    pos = max(np.max(np.abs(x - rx) / np.maximum(1.0, np.abs(rx))), np.max(np.abs(y - ry) / np.maximum(1.0, np.abs(ry))))
    vel = max(np.max(np.abs(xv - rxv) / np.maximum(1.0, np.abs(rxv))), np.max(np.abs(yv - ryv) / np.maximum(1.0, np.abs(ryv))))
    fmax = max(np.abs(rxf).max(), np.abs(ryf).max())
    frc = max(np.abs(xf - rxf).max(), np.abs(yf - ryf).max()) / fmax
    return pos, vel, frc


@pytest.mark.gpu
@pytest.mark.parametrize("n,steps", CASES)
@pytest.mark.parametrize("variant", ["symmetric_default", "symmetric_resort16", "symmetric_unsorted", "directed"])
def test_fast_mode_final_state_against_the_reference_program(nb, n, steps, variant):
    ref = load(n, steps)
    b = nb.reference_init(n)
    kw = dict(symmetric_default=dict(symmetric=1, sort=1),
              symmetric_resort16=dict(symmetric=1, sort=1, sort_interval=16),
              symmetric_unsorted=dict(symmetric=1, sort=-1),
              directed=dict(symmetric=-1))[variant]
    g = nb.NBodyGPU(*b, mode=nb.MODE_FAST, device=0, **kw)
    try:
        g.step(steps)
        got = g.fetch()
    finally:
        g.close()
    pos, vel, frc = errors(got, ref)
    print("\ntrajectory %dx%d %-20s final-state error: positions %.3e velocities %.3e forces %.3e (of max|F|)"
          % (n, steps, variant, pos, vel, frc))
    assert pos <= POS_TOL and vel <= VEL_TOL and frc <= FRC_TOL, (pos, vel, frc)


@pytest.mark.gpu
@pytest.mark.parametrize("n,steps", CASES)
def test_exact_mode_final_state_against_the_reference_program(nb, n, steps):
    """Exact mode swaps libm's trig for detmath (same formula, same order): it must land on the reference program's
    final state within the same gates, too."""
    ref = load(n, steps)
    b = nb.reference_init(n)
    g = nb.NBodyGPU(*b, mode=nb.MODE_EXACT, device=0)
    try:
        g.step(steps)
        got = g.fetch()
    finally:
        g.close()
    pos, vel, frc = errors(got, ref)
    print("\ntrajectory %dx%d exact final-state error: positions %.3e velocities %.3e forces %.3e" % (n, steps, pos, vel, frc))
    assert pos <= POS_TOL and vel <= VEL_TOL and frc <= FRC_TOL, (pos, vel, frc)
