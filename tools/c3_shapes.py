"""Development aid (NBODY_DEV build): N = 65536 step time for the two CTA shapes of the symmetric kernel and a few
j-chunk lengths.  NBODY_GPU_LIB must point at libnbody_gpu_dev.so; NBODY_SYM_VARIANT=1|2 pins 128x4 | 64x8."""
import os, sys, subprocess, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
if len(sys.argv) > 1 and sys.argv[1] == "one":
    import pnb200
    nb = pnb200.load()
    n, jsplit = int(sys.argv[2]), int(sys.argv[3])
    x, y, xv, yv, m, r = nb.reference_init(n)
    g = nb.NBodyGPU(x, y, xv, yv, m, r, mode=1, jsplit=jsplit)
    g.step_timed(4)
    ms = g.step_timed(40) / 40
    print(json.dumps({"n": n, "variant": os.environ.get("NBODY_SYM_VARIANT", "default"), "jsplit": jsplit, "ms_per_step": ms,
                      "pairs_per_s": n * (n - 1.0) / ms * 1e3, "chunk": g.describe()["chunk"]}))
    sys.exit(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
dev = os.path.join(ROOT, "parallel-nbody_b200", "libnbody_gpu_dev.so")
for variant in ("1", "2"):
    for jsplit in (0, n // 2048, n // 1024, n // 512):
        env = dict(os.environ, NBODY_GPU_LIB=dev, NBODY_SYM_VARIANT=variant)
        r = subprocess.run([sys.executable, __file__, "one", str(n), str(jsplit)], env=env, capture_output=True, text=True)
        print(r.stdout.strip() or r.stderr[-400:], flush=True)
