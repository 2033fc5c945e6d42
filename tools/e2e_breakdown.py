"""Development aid: wall-clock split of the end-to-end step (set_state / step+sync / fetch) through the C ABI."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import pnb200, oracle_lib

nb = pnb200.load()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
b = oracle_lib.init_bodies(n)
g = nb.NBodyGPU(b.x, b.y, b.xv, b.yv, b.mass, b.radius, mode=1)
g.step(2)
host = g.fetch()
x, y, xf, yf, xv, yv = host
acc = [0.0, 0.0, 0.0]
reps = 5
for _ in range(reps):
    t0 = time.perf_counter(); g.set_state(x, y, xv, yv)
    t1 = time.perf_counter(); g.step(1); g.sync()
    t2 = time.perf_counter(); g.fetch(out=host)
    t3 = time.perf_counter()
    acc[0] += t1 - t0; acc[1] += t2 - t1; acc[2] += t3 - t2
print("n=%d  set_state %.3f ms  step+sync %.3f ms  fetch %.3f ms" % (n, *[1e3 * a / reps for a in acc]))
