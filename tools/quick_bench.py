"""Quick device-side timing of the step loop (development aid; bench.py is the contract)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import pnb200, oracle_lib

nb = pnb200.load()
tf, mhz = nb.fp64_peak(0, 300)
print("fp64 peak %.2f TFLOP/s  (clock ~%.0f MHz)" % (tf, mhz), flush=True)
cases = [(int(a.split(":")[0]), int(a.split(":")[1]), int(a.split(":")[2])) for a in sys.argv[1:]] or \
    [(4096, 0, 200), (65536, 1, 10), (262144, 1, 3)]
for n, mode, steps in cases:
    b = oracle_lib.init_bodies(n)
    g = nb.NBodyGPU(b.x, b.y, b.xv, b.yv, b.mass, b.radius, mode=mode)
    g.step_timed(2)
    ms = g.step_timed(steps)
    pairs = steps * n * (n - 1.0)
    print("n=%d mode=%d steps=%d: %.3f ms/step  %.3e pairs/s  algorithmic %.2f TFLOP/s" %
          (n, mode, steps, ms / steps, pairs / (ms * 1e-3), 10 * pairs / (ms * 1e-3) / 1e12), flush=True)
    g.close()
