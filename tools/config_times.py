"""Development aid: wall-clock + device time of the BASELINE configs 1-3 through the C ABI (bench.py's `configs`
block is the contract; this prints the same figures with more detail)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import pnb200, oracle_lib

nb = pnb200.load()
cases = [(32, 0, 100000), (128, 0, 1000), (32, 1, 100000), (128, 1, 1000), (256, 0, 1000), (512, 0, 500), (1024, 0, 300),
         (4096, 0, 1000), (65536, 1, 200)]
if len(sys.argv) > 1:
    cases = [tuple(int(v) for v in a.split(":")) for a in sys.argv[1:]]
for n, mode, steps in cases:
    b = oracle_lib.init_bodies(n)
    g = nb.NBodyGPU(b.x, b.y, b.xv, b.yv, b.mass, b.radius, mode=mode)
    g.step(2)
    g.sync()
    g.set_state(b.x, b.y, b.xv, b.yv)
    t0 = time.perf_counter()
    g.step(steps)
    g.sync()
    wall = time.perf_counter() - t0
    g.set_state(b.x, b.y, b.xv, b.yv)
    ms = g.step_timed(steps)
    print("n=%d mode=%s steps=%d: wall %.4f s, device %.4f s, %.3f us/step, %.3e pair-interactions/s  %s" %
          (n, "fast" if mode else "exact", steps, wall, ms * 1e-3, ms * 1e3 / steps, steps * n * (n - 1.0) / (ms * 1e-3),
           g.describe().get("kernel")), flush=True)
    g.close()
