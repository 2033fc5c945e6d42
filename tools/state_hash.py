"""Development aid: sha256 of the fetched state after a few steps, for bit-level A/B comparison of two builds
(NBODY_GPU_LIB=/path/to/other/libnbody_gpu.so python tools/state_hash.py)."""
import hashlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import pnb200, oracle_lib

nb = pnb200.load()
for n, steps, kw in ((3000, 3, dict(symmetric=1)), (20000, 3, dict(symmetric=1, sort=1)), (65536 + 100, 2, {}), (262144, 1, {})):
    b = oracle_lib.init_bodies(n)
    b.x[n - 1], b.y[n - 1] = b.x[1], b.y[1]
    g = nb.NBodyGPU(b.x, b.y, b.xv, b.yv, b.mass, b.radius, mode=1, **kw)
    g.step(steps)
    h = hashlib.sha256()
    for a in g.fetch():
        h.update(np.ascontiguousarray(a).tobytes())
    print(n, steps, kw, h.hexdigest()[:16], flush=True)
    g.close()
