"""Development aid: error of the fast-mode force kernel against the oracle (libm trig, the reference's own formula)
on the bodies at the row-block boundaries of the cell order + random ones, at BASELINE config 3 and 4 sizes:
    max |dF| / max|F|          (the quantity the stated tolerance 1e-8 bounds)
    max per-body |dF| / |F|    (relative to the body's own force)
    worst one-step deviation of position / velocity over the first steps of the sampled trajectory.
Run once per library build (NBODY_GPU_LIB=...): profiles/r2_error_table.log compares the second- and the
third-order rsqrt correction of the symmetric pair (forces_sym.cuh)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import pnb200, oracle_lib

nb = pnb200.load()
print("library:", nb.LIB_PATH)
for n, steps in ((65536, 6), (262144, 4)):
    b = oracle_lib.init_bodies(n)
    g = nb.NBodyGPU(b.x, b.y, b.xv, b.yv, b.mass, b.radius, mode=nb.MODE_FAST)
    order = oracle_lib.cell_order(nb, b.x, b.y)
    rows = oracle_lib.block_sample(order, per_block_extra=1)
    rows = rows[:: max(1, len(rows) // 400)].astype(np.int32)
    fx, fy = g.forces()
    xf, yf = oracle_lib.forces_for(rows, b.x, b.y, b.mass, b.radius)
    fmax = max(np.abs(xf).max(), np.abs(yf).max())
    dabs = np.hypot(fx[rows] - xf, fy[rows] - yf)
    rel_body = dabs / np.maximum(np.hypot(xf, yf), 1e-300)
    signed = ((fx[rows] - xf) * xf + (fy[rows] - yf) * yf) / np.maximum(xf * xf + yf * yf, 1e-300)   # along the force
    x, y, xv, yv = b.x.copy(), b.y.copy(), b.xv.copy(), b.yv.copy()
    wpos = wvel = wtol = 0.0
    for _ in range(steps):
        wxf, wyf = oracle_lib.forces_for(rows, x, y, b.mass, b.radius)
        wx, wy, wxv, wyv = oracle_lib.step_bodies(rows, x, y, xv, yv, wxf, wyf, b.mass)
        g.step(1)
        x, y, _, _, xv, yv = g.fetch()
        wpos = max(wpos, (np.abs(x[rows] - wx) / np.maximum(np.abs(wx), 1.0)).max(), (np.abs(y[rows] - wy) / np.maximum(np.abs(wy), 1.0)).max())
        wvel = max(wvel, (np.abs(xv[rows] - wxv) / np.maximum(np.abs(wxv), 1e-2)).max(), (np.abs(yv[rows] - wyv) / np.maximum(np.abs(wyv), 1e-2)).max())
        # share of the tests' velocity tolerance used: |dv| <= 1e-12 + 1e-10 |v|
        wtol = max(wtol, (np.abs(xv[rows] - wxv) / (1e-12 + 1e-10 * np.abs(wxv))).max(), (np.abs(yv[rows] - wyv) / (1e-12 + 1e-10 * np.abs(wyv))).max())
    print("N=%d: signed relative error along the force: mean %.3e median %.3e" % (n, signed.mean(), np.median(signed)))
    print("N=%d, %d sampled bodies: max|dF|/max|F| = %.2e (tolerance 1e-8), max per-body |dF|/|F| = %.2e, median %.2e; "
          "one-step deviation over %d steps: position %.2e (tolerance 1e-12), velocity %.2e (tolerance 1e-10; %.3f of the tests' |dv| <= 1e-12 + 1e-10 |v|)"
          % (n, len(rows), dabs.max() / fmax, rel_body.max(), np.median(rel_body), steps, wpos, wvel, wtol), flush=True)
    g.close()
