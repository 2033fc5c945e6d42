#!/usr/bin/env python
"""bench.py -- FP64 pair-interactions/s of the all-pairs N-body step on B200.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun)
    python bench.py --impl reference --gpus N --steps K --warmup W

Workload (BASELINE.json config 4, the one `metric` is quoted on): N = 262144 bodies from the
reference initialiser (srand(27102015), world 1024x768; reference nbody/nbody-par.c:208-222),
fast mode, STRONG scaling: the same 262144 bodies are sharded over the GPUs (row blocks of the symmetric
kernel); the per-step exchange that replaces MPI_Allgatherv (nbody-par.c:913) is fused into the kernels over
NVLink peer memory (or one ncclAllGather when the GPUs cannot map each other).
A step = one force + integrate (+ exchange) pass.  value = K * N * (N-1) / t, t = device time of
the K-step loop on the production path (nbody_gpu_step_bench: CUDA-graph launches, CUDA events on the library's
stream, a 192 MiB L2 flush before every graph launch inside the span), max over ranks.

e2e: the same metric through the C ABI with HOST buffers every step: nbody_gpu_set_state (H2D) ->
nbody_gpu_step(1) -> nbody_gpu_fetch (D2H), wall clock; the caller's six arrays are page-locked once
(nbody_gpu_pin_host), like a long-lived `struct world`.

roofline: the force kernel is FP64-pipe bound (neither HBM nor tensor): achieved = algorithmic
flops by the reference's own model (nbody-seq.c:367-380: 20 flop per unordered pair = 10 per
directed pair-interaction) / the kernel's mean duration from per-launch CUDA events (a second pass of K steps
right after the timed one: nbody_gpu_step_profiled, plain launches with an event between the phases);
peak = the DFMA rate measured live by nbody_gpu_fp64_peak (MEASURED_PEAKS.json has
no FP64 figure).  pipe_frac = FP64-pipe instructions actually issued per pair (SASS count,
confirmed by ncu sm__inst_executed_pipe_fp64) x pairs/s / (peak/2): the utilisation figure the
north star's "% of FP64 pipe peak" means (the reciprocal square root makes one unordered pair 17 pipe
instructions = 8.5 per interaction -- 15 = 7.5 in far-field tiles -- so the algorithmic-flop fraction cannot approach 1).
traffic = DRAM bytes per force-kernel launch from the committed ncu captures (profiles/r2_traffic.json,
written by tools/ncu_traffic.py from the .ncu-rep), keyed by kernel, N and GPU count.

parity (N > 1): after the timed region every rank re-runs 3 steps from the initial state on the sharded handle and
on a one-GPU handle on its own device; `bits_equal_to_1gpu` is true only if all six arrays agree bit for bit on
every rank.

configs (N = 1): the other BASELINE configs through the same C ABI: C1 32 x 100000 and 128 x 1000 (exact, the
one-launch step loop; beside the reference's own nbody-seq timed on this host), C2 4096 x 1000 exact, C3 65536 x 200 fast.

--impl reference: the reference's own parallel program (unmodified nbody-par.c built by oracle/Makefile against
the fork/shared-memory MPI stand-in, one rank per host core) for ONE full step of the same N = 262144 workload
(kind "reference"), cross-checked by the oracle port's rows sample; the port sample alone (kind "port") when the
binary is missing or a full step would take more than ~7 minutes on this host.  The oracle is only ever
executed in that arm and in cpu_baseline.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

# stdout carries exactly one JSON line: NCCL's banner (NCCL_DEBUG=VERSION prints to stdout and ignores
# NCCL_DEBUG_FILE) is raised to WARN, which prints the same banner to the debug file
if os.environ.get("NCCL_DEBUG", "").upper().startswith("V"):
    os.environ["NCCL_DEBUG"] = "WARN"
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))   # oracle_lib (cpu_baseline leg and --impl reference only)

METRIC = "fp64_pair_interactions_per_s"
UNIT = "pair-interactions/s"
N_DEFAULT = 262144
XDIM, YDIM = 1024, 768


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--jsplit", type=int, default=0, help="development: number of j-chunks (0 = library default)")
    ap.add_argument("--n", "--bodies", dest="n", type=int, default=N_DEFAULT)   # under torchrun use --bodies (--n is ambiguous there)
    ap.add_argument("--mode", default="fast", choices=["fast", "exact"])
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the BASELINE config 1-3 block of the N = 1 line")
    ap.add_argument("--no-parity", action="store_true", help="skip the bits-equal-to-one-GPU block of the N > 1 line")
    ap.add_argument("--no-reference-full-step", action="store_true",
                    help="--impl reference: port sample only (skip the reference binary's full step)")
    ap.add_argument("--reference-full-step-limit", type=float, default=420.0,
                    help="--impl reference: longest predicted full step (s) the reference binary is run for")
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks + throttle reasons while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index):
        self.idx = device_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "25"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append((time.perf_counter(), ln.strip()))

    def stop(self, t0=None, t1=None):
        """Summary of the samples taken between t0 and t1 (perf_counter; all samples if none fall inside)."""
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        inside = [ln for (ts, ln) in self.lines if t0 is not None and t0 <= ts <= t1 + 0.15]
        for ln in (inside or [ln for (_, ln) in self.lines]):
            f = [t.strip() for t in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for nm, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------------------------
def cpu_sample(oracle_lib, b, n, threads, budget_s):
    """Bounded CPU sample of the same workload: rows [0,R) x all N columns, oracle row form with
    libm trig on `threads` host threads.  Returns (pairs/s, description)."""
    L = oracle_lib.lib()
    xf, yf = np.zeros(n), np.zeros(n)
    rows = max(threads, 8)
    t0 = time.perf_counter()
    L.oracle_forces_rows(n, 0, 0, rows, threads, b.x, b.y, b.mass, b.radius, xf, yf)
    dt = time.perf_counter() - t0
    rate = rows * (n - 1.0) / dt
    rows2 = int(max(rows, min(n, budget_s * rate / (n - 1.0))))
    rows2 = max(threads, rows2 // threads * threads)
    t0 = time.perf_counter()
    L.oracle_forces_rows(n, 0, 0, rows2, threads, b.x, b.y, b.mass, b.radius, xf, yf)
    dt = time.perf_counter() - t0
    return rows2 * (n - 1.0) / dt, rows2, dt


def cpu_seq_sample(oracle_lib, b, n, budget_s):
    """nbody-seq's own form: triangular walk, one core; 2 directed interactions per evaluation."""
    L = oracle_lib.lib()
    xf, yf = np.zeros(n), np.zeros(n)
    rows = 4
    t0 = time.perf_counter()
    ev = L.oracle_forces_symmetric_sample(n, 0, 0, rows, b.x, b.y, b.mass, b.radius, xf, yf)
    dt = time.perf_counter() - t0
    rows2 = max(rows, int(budget_s * (ev / dt) / n))
    t0 = time.perf_counter()
    ev = L.oracle_forces_symmetric_sample(n, 0, 0, rows2, b.x, b.y, b.mass, b.radius, xf, yf)
    dt = time.perf_counter() - t0
    return 2.0 * ev / dt, rows2, dt


def run_reference_binary(n, steps, timeout=900, threads_per_rank=1):
    """The reference's OWN parallel program (unmodified nbody-par.c built by oracle/Makefile against the
    fork/shared-memory MPI stand-in; MAXBODIES raised) on all host cores, pure MPI (one rank per core).
    Returns None if oracle/_ref/nbody-par-big is not built or the run fails."""
    exe = os.path.join(ROOT, "oracle", "_ref", "nbody-par-big")
    if not os.path.exists(exe):
        return None
    import tempfile
    import oracle_lib
    cores = max(1, (os.cpu_count() or 1) // threads_per_rank)   # ranks
    hybrid = ["--openmp"] if threads_per_rank > 1 else []       # nbody-par.c:302,366,395: the OpenMP loops are opt-in
    with tempfile.TemporaryDirectory() as d:
        ppm = os.path.join(d, "nbody.ppm")
        oracle_lib.write_ppm(ppm, XDIM, YDIM)
        env = dict(os.environ, MPISHIM_NP=str(cores), OMP_NUM_THREADS=str(threads_per_rank))
        try:
            r = subprocess.run([exe, str(n), "0", ppm, str(steps)] + hybrid, env=env, cwd=d, stdout=subprocess.DEVNULL,
                               stderr=subprocess.PIPE, text=True, timeout=timeout)
        except (OSError, subprocess.TimeoutExpired):
            return None
    took = [ln for ln in r.stderr.splitlines() if "N-body took" in ln]
    if r.returncode != 0 or not took:
        return None
    secs = float(took[0].split(":")[1].split()[0])
    return {"kind": "reference", "binary": "oracle/_ref/nbody-par-big (reference nbody-par.c + MPI stand-in)",
            "ranks": cores, "threads_per_rank": threads_per_rank, "openmp": bool(hybrid), "n": n, "steps": steps, "seconds": secs,
            "value": steps * n * (n - 1.0) / max(secs, 1e-9), "unit": UNIT}


def reference_binary_check(n_small=16384, steps=2):
    """The reference binary at a size it finishes in seconds: per-pair cost does not depend on N, so this pins the
    port's rate to the real thing (and predicts how long a full step of the headline N takes)."""
    return run_reference_binary(n_small, steps, timeout=300)


def reference_seq_seconds(n, steps):
    """The reference's sequential program (oracle/_ref/nbody-seq, its Makefile's flags) on one host core: seconds of
    its own "N-body took" line, or None if it is not built."""
    exe = os.path.join(ROOT, "oracle", "_ref", "nbody-seq")
    if not os.path.exists(exe):
        return None
    import tempfile
    import oracle_lib
    with tempfile.TemporaryDirectory() as d:
        ppm = os.path.join(d, "nbody.ppm")
        oracle_lib.write_ppm(ppm, XDIM, YDIM)
        try:
            r = subprocess.run([exe, str(n), "0", ppm, str(steps)], cwd=d, stdout=subprocess.DEVNULL,
                               stderr=subprocess.PIPE, text=True, timeout=120)
        except (OSError, subprocess.TimeoutExpired):
            return None
    took = [ln for ln in r.stderr.splitlines() if "N-body took" in ln]
    return float(took[0].split(":")[1].split()[0]) if (r.returncode == 0 and took) else None


def run_reference(args, rank, world):
    """The reference arm: CPU only, rank 0 only."""
    if rank != 0:
        return
    import oracle_lib
    n = args.n
    threads = os.cpu_count() or 1
    b = oracle_lib.init_bodies(n)
    L = oracle_lib.lib()
    xf, yf = np.zeros(n), np.zeros(n)
    # port sample: one "step" = a rows sample worth ~1 s, so that steps+warmup stay well within a minute
    t0 = time.perf_counter()
    L.oracle_forces_rows(n, 0, 0, 4 * threads, threads, b.x, b.y, b.mass, b.radius, xf, yf)
    probe = time.perf_counter() - t0
    rows = max(threads, int(1.0 / probe * 4 * threads) // threads * threads)
    rows = min(rows, n)
    xn, yn = np.zeros(n), np.zeros(n)

    def one_step():
        L.oracle_forces_rows(n, 0, 0, rows, threads, b.x, b.y, b.mass, b.radius, xf, yf)
        L.oracle_velocities(rows, 0, xf, yf, b.mass, b.xv, b.yv)
        L.oracle_positions(rows, XDIM, YDIM, b.x, b.y, b.xv, b.yv, xn, yn)

    for _ in range(args.warmup):
        one_step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        one_step()
    dt = time.perf_counter() - t0
    port_value = args.steps * rows * (n - 1.0) / dt
    port = {"value": port_value, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": "rows [0,%d) x all %d columns per step (%.3g of a full step), oracle row form, libm trig, "
                      "%d + %d such steps" % (rows, n, rows / float(n), args.warmup, args.steps)}

    # the reference program itself: first at a size it finishes in seconds (predicts a full step), then ONE full step
    chk = reference_binary_check()
    # north_star's "MPI+OpenMP" shape of the reference: cores/4 ranks x 4 threads with --openmp, at the small size
    hybrid = run_reference_binary(16384, 2, timeout=300, threads_per_rank=4) if (chk and threads >= 4) else None
    full = None
    if chk and not args.no_reference_full_step:
        predicted = n * (n - 1.0) / chk["value"]
        if predicted <= args.reference_full_step_limit:
            full = run_reference_binary(n, 1, timeout=int(2.5 * predicted) + 120)
    workload = "BASELINE config 4: N=%d bodies, reference initialiser (srand 27102015), world %dx%d" % (n, XDIM, YDIM)
    if full:
        value, ms_per_step = full["value"], full["seconds"] * 1e3
        cpu = {"value": value, "unit": UNIT, "cores": threads, "kind": "reference",
               "sample": "ONE full step of the same N=%d workload by the reference's own nbody-par.c (unmodified; MAXBODIES "
                         "raised, int overflows of the init fixed; built against a fork/shared-memory MPI stand-in), %d ranks x 1 "
                         "thread, its own 'N-body took' timer: %.1f s; --steps/--warmup apply to the port cross-check only"
                         % (n, threads, full["seconds"]),
               "port_cross_check": dict(port, steps=args.steps, warmup=args.warmup), "reference_check_small": chk}
        note = ("one real full step of the reference program; ms_per_step is that step; steps/warmup describe what was "
                "timed (the requested --steps/--warmup drive the port cross-check only, see requested_steps)")
        steps_timed, warmup_timed = 1, 0
    else:
        value, ms_per_step = port_value, dt / args.steps * 1e3 * (n / float(rows))
        cpu = dict(port)
        if chk:
            cpu["reference_check"] = chk
        note = "port sample (the reference binary is missing or a full step would exceed the limit); ms_per_step is the sample time scaled to a full step (x N/rows)"
        steps_timed, warmup_timed = args.steps, args.warmup
    if hybrid:
        cpu["reference_check_hybrid"] = hybrid
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps_timed, "warmup": warmup_timed, "requested_steps": args.steps, "requested_warmup": args.warmup,
            "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": {"workload": workload + "; CPU: nbody-par MPI ranks on the host cores", "n": n,
                                            "mode": "reference", "parallelism": "%d host cores" % threads},
            "cpu_baseline": cpu,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "note": note}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------
def state_sha256(arrays):
    import hashlib
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def load_traffic(kernel, n, gpus):
    """DRAM bytes per force-kernel launch from the committed ncu captures (profiles/r2_traffic.json)."""
    path = os.path.join(ROOT, "profiles", "r2_traffic.json")
    try:
        with open(path) as f:
            table = json.load(f)
    except (OSError, ValueError):
        return None, None
    for e in table.get("captures", []):
        if e.get("kernel") == kernel and e.get("n") == n and e.get("gpus") == gpus:
            return e.get("dram_bytes_read", 0.0) + e.get("dram_bytes_write", 0.0), \
                "profiles/r2_traffic.json <- %s (dram__bytes_read.sum + dram__bytes_write.sum per launch)" % e.get("source")
    return None, None


def other_configs(nb, local_rank, peak_tf):
    """BASELINE configs 1-3 through the same C ABI on one GPU (config 4 is the line itself, config 5 needs 8 GPUs)."""
    out = {}

    def run(n, mode, steps, reps=1):
        x, y, xv, yv, mass, radius = nb.reference_init(n, XDIM, YDIM)
        g = nb.NBodyGPU(x, y, xv, yv, mass, radius, xdim=XDIM, ydim=YDIM, mode=mode, device=local_rank)
        g.step(2)
        g.sync()
        best = None
        for _ in range(reps):
            g.set_state(x, y, xv, yv)
            g.sync()
            t0 = time.perf_counter()
            g.step(steps)
            g.sync()
            wall = time.perf_counter() - t0
            best = wall if best is None else min(best, wall)
        g.set_state(x, y, xv, yv)
        ms = g.step_bench(steps, flush_l2=False)
        d = g.describe()
        far = g.far_fraction() if d.get("cell_order") else 0.0
        g.close()
        return best, ms * 1e-3, d, far

    for n, steps in ((32, 100000), (128, 1000)):
        wall, dev, d, _ = run(n, nb.MODE_EXACT, steps, reps=2)
        cpu_s = reference_seq_seconds(n, steps)
        out["c1_%dx%d_exact" % (n, steps)] = {
            "seconds": wall, "device_seconds": dev, "steps_per_s": steps / wall, "kernel": d["kernel"],
            "launches_for_the_run": 1 if d.get("single_launch_loop") else None,
            "cpu_seconds_reference_nbody_seq_1core": cpu_s, "speedup_vs_nbody_seq": (cpu_s / wall) if cpu_s else None,
            "checked_by": "tests/test_gpu_parity.py::test_exact_golden_* (byte-identical text, bit-identical state)"}
    wall, dev, d, _ = run(4096, nb.MODE_EXACT, 1000)
    out["c2_4096x1000_exact"] = {"seconds": wall, "ms_per_step": dev * 1e3 / 1000, "steps_per_s": 1000 / wall, "kernel": d["kernel"],
                                 "pair_interactions_per_s": 1000 * 4096 * 4095.0 / dev,
                                 "checked_by": "tests/test_gpu_parity.py::test_exact_4096_config2_prefix_and_graph_independence"}
    wall, dev, d, far = run(65536, nb.MODE_FAST, 200)
    rate = 200 * 65536 * 65535.0 / dev
    ipp = 8.5 * (1.0 - far) + 7.5 * far
    out["c3_65536x200_fast"] = {"seconds": wall, "ms_per_step": dev * 1e3 / 200, "steps_per_s": 200 / wall, "kernel": d["kernel"],
                                "pair_interactions_per_s": rate, "far_field_fraction": far,
                                "pipe_frac": ipp * rate / (peak_tf * 1e12 / 2.0),
                                "ncu": "profiles/r2_ncu_forces_sym_N65536.txt",
                                "checked_by": "tests/test_gpu_parity.py::test_baseline_configs_3_and_4_*"}
    return out


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return 0

    import torch
    import pnb200
    nb = pnb200.load()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    nccl_id = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        idt = torch.zeros(nb.NCCL_ID_BYTES, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(nb.nccl_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        nccl_id = bytes(idt.cpu().numpy().tobytes())

    n = args.n
    mode = nb.MODE_FAST if args.mode == "fast" else nb.MODE_EXACT
    # the reference's synthetic initial state (nbody-par.c:208-222) from the library's own host helper
    bx, by, bxv, byv, bmass, bradius = nb.reference_init(n, XDIM, YDIM)
    p2p_cfg = int(os.environ.get("NBODY_BENCH_P2P", "0"))   # development: -1 = NCCL exchange
    g = nb.NBodyGPU(bx, by, bxv, byv, bmass, bradius, xdim=XDIM, ydim=YDIM, mode=mode,
                    device=local_rank, rank=rank, nranks=world, nccl_id=nccl_id, jsplit=args.jsplit, p2p=p2p_cfg)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def maxreduce(v):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    peak_tf, _ = nb.fp64_peak(local_rank, 200)

    # ---- device-resident timed region: the production path (CUDA graphs) -------------------
    g.step(args.warmup)
    g.step_bench(2, flush_l2=True)      # untimed: the flush scratch and the graphs exist afterwards
    g.step_profiled(1, flush_l2=True)   # untimed: the event pool of the per-phase pass exists afterwards
    g.sync()
    l0 = sum(g.launch_count()[1][:3])   # our kernels only (force, integrate, others), not NCCL's
    # the sampler (a fork of nvidia-smi: tens of ms on an 8-GPU box) starts BEFORE the barrier, so that rank 0 does
    # not enter the timed region late and leave the other ranks waiting for it in their first exchange
    sampler = ClockSampler(local_rank)   # every rank records its own GPU
    sampler.start()
    time.sleep(0.3)
    barrier()
    t_region0 = time.perf_counter()
    t_ms = g.step_bench(args.steps, flush_l2=True)
    launches = sum(g.launch_count()[1][:3]) - l0
    barrier()
    # ---- the same K steps once more with an event between the phases (plain launches): kernel durations -----
    prof = g.step_profiled(args.steps, flush_l2=True)
    t_region1 = time.perf_counter()
    barrier()
    clocks = sampler.stop(t_region0, t_region1)
    if dist is not None:
        # per-rank clock record: [median SM MHz, max W, samples, throttled?] of every GPU, gathered on all ranks
        mine = torch.tensor([float(clocks.get("sm_mhz") or -1.0), float(clocks.get("power_w_max") or -1.0),
                             float(clocks.get("samples") or 0), float(len(clocks.get("reasons") or []))],
                            dtype=torch.float64, device="cuda")
        allc = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allc, mine)
        per_rank = [[float(v) for v in t.cpu().tolist()] for t in allc]
        clocks["per_rank_sm_mhz"] = [r[0] for r in per_rank]
        clocks["per_rank_power_w_max"] = [r[1] for r in per_rank]
        clocks["per_rank_samples"] = [int(r[2]) for r in per_rank]
        if any(r[3] > 0 for r in per_rank) and not clocks.get("reasons"):
            clocks["reasons"] = ["see other ranks"]
    per_rank_phase = None
    if dist is not None:
        mine_ph = torch.tensor([prof["force"], prof["integrate"], prof["exchange"], t_ms], dtype=torch.float64, device="cuda")
        all_ph = [torch.zeros_like(mine_ph) for _ in range(world)]
        dist.all_gather(all_ph, mine_ph)
        per_rank_phase = {"force": [float(t[0]) / args.steps for t in all_ph], "integrate": [float(t[1]) / args.steps for t in all_ph],
                          "exchange": [float(t[2]) / args.steps for t in all_ph], "graph_pass": [float(t[3]) / args.steps for t in all_ph]}
    t_ms = maxreduce(t_ms)
    prof_ms = maxreduce(prof["total"])
    force_ms = maxreduce(prof["force"])
    integ_ms = maxreduce(prof["integrate"])
    exch_ms = maxreduce(prof["exchange"])
    pairs_per_step = n * (n - 1.0)
    value = args.steps * pairs_per_step / (t_ms * 1e-3)

    # ---- end to end through the C ABI with host buffers every step --------------------------
    e2e = None
    if not args.no_e2e:
        host = g.fetch()                       # the caller's six arrays, reused every step like the reference's world struct
        for a in host:
            nb.pin_host(a)                     # allocated once, page-locked once (nbody_gpu_pin_host)
        x, y, xf, yf, xv, yv = host
        e_steps = max(2, min(args.steps, 5))
        # Multi-rank with the peer-memory exchange: every rank keeps ITS rows on the host, like the reference's ranks
        # between two MPI_Allgatherv calls (nbody-par.c:225-263): set_state uploads the rank's own rows (the library
        # stores them into every GPU), fetch_range brings the rank's own rows back.  The whole state still crosses
        # PCIe both ways every step, an N-th per rank.  Otherwise (one GPU, NCCL exchange) every rank moves all of it.
        own_rows = world > 1 and "peer-memory" in g.describe().get("exchange", "")
        lo, hi = nb.partition(n, world, rank) if own_rows else (0, n)
        mine = [a[lo:hi] for a in host]

        def fetch_mine():
            if own_rows:
                g.fetch_range(lo, hi - lo, out=mine)
            else:
                g.fetch(out=host)

        g.set_state(x, y, xv, yv)              # untimed round: first-touch of the pinned pages
        g.step(1)
        fetch_mine()
        barrier()
        t0 = time.perf_counter()
        split = [0.0, 0.0, 0.0]
        for _ in range(e_steps):
            ta = time.perf_counter()
            g.set_state(x, y, xv, yv)          # H2D: positions + velocities
            tb = time.perf_counter()
            g.step(1)
            tc = time.perf_counter()
            fetch_mine()                       # D2H: the six columns the reference prints (waits for the step)
            td = time.perf_counter()
            split[0] += tb - ta
            split[1] += tc - tb
            split[2] += td - tc
        barrier()
        e_dt = maxreduce(time.perf_counter() - t0)
        e2e = {"value": e_steps * pairs_per_step / e_dt, "unit": UNIT,
               "h2d_bytes_per_step": 4 * 8 * n, "d2h_bytes_per_step": 6 * 8 * n, "steps": e_steps,
               "ms_per_step": 1e3 * e_dt / e_steps,
               "call": "nbody_gpu_set_state + nbody_gpu_step(1) + nbody_gpu_fetch per step, wall clock, host arrays "
                       "page-locked once with nbody_gpu_pin_host",
               "rank0_ms_per_step": {"set_state": 1e3 * split[0] / e_steps, "step_enqueue": 1e3 * split[1] / e_steps,
                                     "fetch_incl_wait_for_step": 1e3 * split[2] / e_steps},
               "h2d_bytes_per_step_per_rank": 4 * 8 * (hi - lo), "d2h_bytes_per_step_per_rank": 6 * 8 * (hi - lo),
               "rows_per_rank": "own rows [lo, hi) of nbody_gpu_partition on every rank (set_state + fetch_range)" if own_rows
               else "all rows on every rank (set_state + fetch)"}
        if own_rows:
            # the pieces the ranks fetched are one consistent state: rank 0 also fetches everything and the hashes of its
            # own slice must agree (cheap, untimed)
            full = g.fetch()
            e2e["own_rows_consistent_with_full_fetch"] = bool(all(np.array_equal(f[lo:hi], m) for f, m in zip(full, mine)))

    # ---- multi-rank parity: the sharded handle against one GPU, bit for bit, on every rank ---
    parity = None
    if world > 1 and not args.no_parity:
        k = 3
        g.set_state(bx, by, bxv, byv)
        g.step(k)
        sharded = g.fetch()
        g1 = nb.NBodyGPU(bx, by, bxv, byv, bmass, bradius, xdim=XDIM, ydim=YDIM, mode=mode, device=local_rank, jsplit=args.jsplit)
        g1.step(k)
        single = g1.fetch()
        g1.close()
        same = all(np.array_equal(a.view(np.uint64), c.view(np.uint64)) for a, c in zip(sharded, single))
        flag = torch.tensor([1 if same else 0], dtype=torch.int64, device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        parity = {"bits_equal_to_1gpu": bool(flag.item()), "steps": k, "ranks_checked": world,
                  "sha256": state_sha256(sharded), "sha256_1gpu": state_sha256(single),
                  "what": "set_state(initial) + step(%d) + fetch on the %d-rank handle vs a one-GPU handle on the same device, "
                          "all six arrays, every rank (MIN-reduced)" % (k, world)}

    # ---- more multi-rank parity (each against one GPU, bit for bit, MIN over ranks) and BASELINE config 5 -----------
    def fresh_nccl_id():
        idt = torch.zeros(nb.NCCL_ID_BYTES, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.frombuffer(bytearray(nb.nccl_unique_id()), dtype=torch.uint8))
        dist.broadcast(idt, 0)
        return bytes(idt.cpu().numpy().tobytes())

    def sharded_against_one_gpu(arrays, k, timed_steps=0, **kw):
        gs = nb.NBodyGPU(*arrays, xdim=XDIM, ydim=YDIM, mode=mode, device=local_rank, rank=rank, nranks=world,
                         nccl_id=fresh_nccl_id(), **kw)
        out = {}
        if timed_steps:
            gs.step(2)
            gs.sync()
            barrier()
            ms = maxreduce(gs.step_bench(timed_steps, flush_l2=True))
            nn = len(arrays[0])
            out.update({"steps": timed_steps, "ms_per_step": ms / timed_steps,
                        "pair_interactions_per_s": timed_steps * nn * (nn - 1.0) / (ms * 1e-3)})
            gs.set_state(*arrays[:4])
        gs.step(k)
        sharded = gs.fetch()
        d = gs.describe()
        gs.close()
        g1 = nb.NBodyGPU(*arrays, xdim=XDIM, ydim=YDIM, mode=mode, device=local_rank,
                         **{a: v for a, v in kw.items() if a != "p2p"})
        g1.step(k)
        single = g1.fetch()
        g1.close()
        same = all(np.array_equal(a.view(np.uint64), c.view(np.uint64)) for a, c in zip(sharded, single))
        flag = torch.tensor([1 if same else 0], dtype=torch.int64, device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        out["parity"] = {"bits_equal_to_1gpu": bool(flag.item()), "steps": k, "ranks_checked": world,
                         "sha256": state_sha256(sharded), "exchange": d.get("exchange"), "sorts": d.get("sorts")}
        return out

    desc = g.describe()
    far = g.far_fraction() if desc.get("cell_order") else 0.0
    g.close()   # one sharded handle at a time: the blocks below create their own, one after the other

    c5 = None
    if world > 1 and not args.no_parity:
        init = (bx, by, bxv, byv, bmass, bradius)
        # a run that crosses re-sorts of the cell order (sort_interval = 2: sorts before steps 2 and 4)
        try:
            parity["across_resorts"] = dict(sharded_against_one_gpu(init, 5, sort_interval=2)["parity"], sort_interval=2)
            # the same step over the ncclAllGather exchange instead of the peer-memory stores
            parity["nccl_exchange"] = dict(sharded_against_one_gpu(init, 3, p2p=-1)["parity"], p2p=-1)
        except nb.NBodyGPUError as e:   # the headline line must not depend on these blocks
            parity["extra_blocks_error"] = str(e)
        # BASELINE config 5 is quoted on 8 GPUs; NBODY_BENCH_C5=1 (development) runs the block on any N > 1
        if (world == 8 or os.environ.get("NBODY_BENCH_C5") == "1") and n == N_DEFAULT and not args.no_configs:
            n5 = 1048576   # BASELINE config 5: N = 1 048 576 x 10 steps on 8 GPUs (nbody-par.c:901-921 is the loop)
            try:
                c5 = sharded_against_one_gpu(nb.reference_init(n5, XDIM, YDIM), 3, timed_steps=10)
                c5["kernel"] = "forces_sym_kernel<64,8>"
                c5["timing"] = ("nbody_gpu_step_bench(10) on the %d-rank handle, CUDA-graph path, L2 flushed per launch, "
                                "max over ranks" % world)
            except nb.NBodyGPUError as e:   # the headline line must not depend on this block
                c5 = {"error": str(e)}

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return 0

    # ---- roofline of the dominant kernel (force kernel) -------------------------------------
    rows = (n + world - 1) // world
    launch_pairs = rows * (n - 1.0)                       # pair-interactions one force launch evaluates
    launch_s = force_ms * 1e-3 / args.steps
    algo_tflops = 10.0 * launch_pairs / launch_s / 1e12   # reference flop model: 10 per directed pair
    ipp = desc.get("fp64_instr_per_pair") or None
    if ipp and far > 0.0:
        # far-field tiles issue 15 FP64-pipe instructions per unordered pair instead of 17 (forces_sym.cuh)
        ipp = ipp * (1.0 - far) + 7.5 * far
    traffic, traffic_source = load_traffic(desc["kernel"], n, world)
    roofline = {"bound": "fp64", "achieved": algo_tflops, "peak": peak_tf, "unit": "TFLOP/s",
                "frac": algo_tflops / peak_tf, "traffic": traffic, "traffic_source": traffic_source,
                "peak_source": "builder-measured live: nbody_gpu_fp64_peak DFMA kernel on this GPU "
                               "(MEASURED_PEAKS.json has no FP64 entry; theory 148 SM x 64 lanes x 2 x 1.965 GHz = 37.2)",
                "kernel": desc["kernel"], "evaluation": "symmetric (each unordered pair once)" if desc["symmetric"]
                else "directed", "kernel_ms_per_launch": launch_s * 1e3,
                "kernel_share_of_step": force_ms / prof_ms,
                "timing": "mean of per-launch CUDA events over a second pass of the same K steps (plain launches, "
                          "%.3f ms per step) right after the graph-launched timed pass (%.3f ms per step)"
                          % (prof_ms / args.steps, t_ms / args.steps),
                "algorithmic_flops_per_pair": 10, "pairs_per_launch": launch_pairs}
    if ipp:
        roofline["far_field_fraction"] = far
        roofline["fp64_pipe_instr_per_pair"] = ipp
        roofline["pipe_frac"] = ipp * (launch_pairs / launch_s) / (peak_tf * 1e12 / 2.0)

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        import oracle_lib   # the checker, timed as the CPU baseline: the only place this arm touches oracle/
        b = oracle_lib.init_bodies(n, XDIM, YDIM)
        threads = os.cpu_count() or 1
        v, rws, dt = cpu_sample(oracle_lib, b, n, threads, args.cpu_seconds)
        v1, rws1, dt1 = cpu_seq_sample(oracle_lib, b, n, min(6.0, args.cpu_seconds / 2))
        cpu = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": "rows [0,%d) x all %d columns of the same N=%d workload, oracle row form (nbody-par pass B), "
                         "libm trig, %.1f s" % (rws, n, n, dt),
               "seq_1core": {"value": v1, "cores": 1,
                             "sample": "triangular walk rows [0,%d) (nbody-seq form, 2 interactions per evaluation), "
                                       "%.1f s" % (rws1, dt1)}}

        chk = reference_binary_check()
        if chk:
            cpu["reference_check"] = chk

    configs = None
    if world == 1 and not args.no_configs and n == N_DEFAULT:
        configs = other_configs(nb, local_rank, peak_tf)

    mb = lambda v: "%.0f MB" % (v / 1e6)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t_ms / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "BASELINE config 4: N=%d bodies, reference initialiser (srand 27102015), world %dx%d; "
                                   "GPU: %s mode, row blocks sharded over %d GPU(s), per-step exchange: %s"
                                   % (n, XDIM, YDIM, args.mode, world, desc.get("exchange", "none")),
                       "n": n, "mode": args.mode, "parallelism": "row-shard x%d" % world,
                       "step_path": "CUDA graph of %d steps per launch" % desc["graph_steps"] if desc.get("graph_steps") else "plain launches",
                       "l2": "192 MiB scratch overwritten before every graph launch, inside the timed span (L2 flush); "
                             "the force kernel's partial sums (%s of device memory per GPU) are meant to live in the L2, "
                             "its DRAM traffic per launch is roofline.traffic (%s)"
                             % (mb(desc.get("partial_sum_bytes", 0)), mb(traffic) if traffic else "no capture for this configuration")},
            "steps_per_s": args.steps / (t_ms * 1e-3),
            "phase_ms_per_step": {"force": force_ms / args.steps, "integrate": integ_ms / args.steps,
                                  "exchange": exch_ms / args.steps, "whole_step_plain_launches": prof_ms / args.steps},
            "clocks": clocks, "gpu_launches": launches, "roofline": roofline}
    if per_rank_phase:
        line["phase_ms_per_step_per_rank"] = per_rank_phase
    if e2e:
        line["e2e"] = e2e
    if parity:
        line["parity"] = parity
    if cpu:
        line["cpu_baseline"] = cpu
    if configs:
        line["configs"] = configs
    if c5:
        line.setdefault("configs", {})["c5_1048576x10"] = c5
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
