"""Development aid (NBODY_DEV build only): time the force kernel of ONE virtual rank -- what one GPU of an 8-way
shard runs -- on a single device, for every virtual rank.  NBODY_GPU_LIB must point at libnbody_gpu_dev.so.
Shows how evenly the eight GPUs of a strong-scaling run are loaded without paying for eight."""
import os, sys, subprocess, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
if len(sys.argv) > 1 and sys.argv[1] == "one":
    import pnb200
    nb = pnb200.load()
    n = int(sys.argv[2])
    x, y, xv, yv, m, r = nb.reference_init(n)
    g = nb.NBodyGPU(x, y, xv, yv, m, r, mode=1)
    g.step_profiled(2)
    p = g.step_profiled(6)
    print(json.dumps({"vrank": os.environ.get("NBODY_DEV_VRANK", "all"), "force_ms": p["force"] / 6}))
    sys.exit(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
dev = os.path.join(ROOT, "parallel-nbody_b200", "libnbody_gpu_dev.so")
for v in ["all"] + [str(k) for k in range(8)]:
    env = dict(os.environ, NBODY_GPU_LIB=dev)
    if v != "all":
        env["NBODY_DEV_VRANK"] = v
    r = subprocess.run([sys.executable, __file__, "one", str(n)], env=env, capture_output=True, text=True)
    print(r.stdout.strip() or r.stderr[-500:], flush=True)
