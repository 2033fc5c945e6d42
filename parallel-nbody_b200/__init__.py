"""parallel-nbody_b200 -- host-side Python mirror of the C ABI in include/nbody_gpu.h.

The product is libnbody_gpu.so (hand-written sm_100a CUDA behind a C ABI); this module is a
thin ctypes binding used by tests/, bench.py and __graft_entry__.py.  It adds no arithmetic
of its own and has no fallback: if the library is missing or no CUDA device is usable, the
calls raise.

Reference mapping (jfdev001/parallel-nbody): NBodyGPU(...) ~ calloc(world) + initialize +
MPI_Bcast (nbody/nbody-par.c:810,870-874); .step(k) ~ the step loop (nbody-par.c:901-921);
.fetch() ~ reading world->bodies for print() (nbody-par.c:740-750).
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NBODY_GPU_LIB") or os.path.join(_HERE, "libnbody_gpu.so")   # override: A/B runs of two builds

MODE_EXACT = 0
MODE_FAST = 1
NCCL_ID_BYTES = 128
STATIC_MAX_BODIES = 32768          # NBODY_GPU_STATIC_MAX_BODIES
STATIC_SCRATCH_BYTES = 16 << 20     # NBODY_GPU_STATIC_SCRATCH_BYTES

_dp = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")


class NBodyGPUError(RuntimeError):
    pass


class _Cfg(ctypes.Structure):
    _fields_ = [("n", ctypes.c_int), ("xdim", ctypes.c_int), ("ydim", ctypes.c_int),
                ("mode", ctypes.c_int), ("ngpus", ctypes.c_int), ("device", ctypes.c_int),
                ("rank", ctypes.c_int), ("nranks", ctypes.c_int), ("nccl_id", ctypes.c_void_p),
                ("jsplit", ctypes.c_int), ("use_graph", ctypes.c_int), ("symmetric", ctypes.c_int), ("p2p", ctypes.c_int),
                ("sort", ctypes.c_int), ("sort_interval", ctypes.c_int)]


_lib = None


def build():
    """Compile libnbody_gpu.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    import subprocess
    subprocess.run(["make", "-C", os.path.join(_HERE, "csrc"), "--no-print-directory"], check=True)


def lib():
    """Load libnbody_gpu.so; raises NBodyGPUError if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise NBodyGPUError("libnbody_gpu.so is not built (run `make -C parallel-nbody_b200/csrc` "
                            "or __graft_entry__.build()); there is no CPU fallback")
    L = ctypes.CDLL(LIB_PATH)
    vp, ci = ctypes.c_void_p, ctypes.c_int
    odp = ctypes.c_void_p  # optional double* (may be NULL)
    L.nbody_gpu_init.argtypes = [ctypes.POINTER(vp), ctypes.POINTER(_Cfg)] + [_dp] * 6
    L.nbody_gpu_set_state.argtypes = [vp] + [_dp] * 4
    L.nbody_gpu_step.argtypes = [vp, ci]
    L.nbody_gpu_step_timed.argtypes = [vp, ci, ctypes.POINTER(ctypes.c_float)]
    L.nbody_gpu_step_bench.argtypes = [vp, ci, ci, ctypes.POINTER(ctypes.c_float)]
    L.nbody_gpu_step_profiled.argtypes = [vp, ci, ci, ctypes.POINTER(ctypes.c_float)]
    L.nbody_gpu_forces.argtypes = [vp, _dp, _dp]
    L.nbody_gpu_fetch.argtypes = [vp] + [odp] * 6
    L.nbody_gpu_fetch_range.argtypes = [vp, ci, ci] + [odp] * 6
    L.nbody_gpu_render.argtypes = [vp, np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")]
    L.nbody_gpu_far_fraction.argtypes = [vp, ctypes.POINTER(ctypes.c_double)]
    L.nbody_gpu_sync.argtypes = [vp]
    L.nbody_gpu_destroy.argtypes = [vp]
    L.nbody_gpu_destroy.restype = None
    L.nbody_gpu_last_error.restype = ctypes.c_char_p
    L.nbody_gpu_describe.argtypes = [vp, ctypes.c_char_p, ci]
    L.nbody_gpu_launch_count.argtypes = [vp, ctypes.POINTER(ctypes.c_longlong)]
    L.nbody_gpu_launch_count.restype = ctypes.c_longlong
    L.nbody_gpu_partition.argtypes = [ci, ci, ci, ctypes.POINTER(ci), ctypes.POINTER(ci)]
    L.nbody_gpu_partition.restype = None
    L.nbody_gpu_partition_reference.argtypes = [ci, ci, ctypes.POINTER(ci), ctypes.POINTER(ci)]
    L.nbody_gpu_partition_reference.restype = None
    L.nbody_gpu_cell_key.argtypes = [ctypes.c_double, ctypes.c_double, ci, ci]
    L.nbody_gpu_cell_key.restype = ctypes.c_uint
    L.nbody_gpu_reference_init.argtypes = [ci, ci, ci] + [_dp] * 6
    L.nbody_gpu_reference_init.restype = None
    L.nbody_gpu_nccl_unique_id.argtypes = [ctypes.c_char_p]
    L.nbody_gpu_device_count.restype = ci
    L.nbody_gpu_fp64_peak.argtypes = [ci, ci, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]
    L.nbody_gpu_abi_version.restype = ci
    L.nbody_gpu_shard_plan.argtypes = [ci, ci, ci, ci, ctypes.POINTER(ci), ctypes.POINTER(ctypes.c_longlong)]
    ip = ctypes.POINTER(ci)
    if not os.environ.get("NBODY_GPU_LIB") or hasattr(L, "nbody_gpu_sym_work_items"):   # an older build under A/B test may lack it
        L.nbody_gpu_sym_work_items.argtypes = [ci, ci, ci, ip, ip, ip, ip, ci, ip, ip, ci, ip, ip]
    L.nbody_gpu_pin_host.argtypes = [vp, ctypes.c_size_t]
    L.nbody_gpu_forces_static.argtypes = [ci, vp, vp, vp, vp]
    L.nbody_gpu_unpin_host.argtypes = [vp]
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise NBodyGPUError("nbody_gpu error %d: %s" % (rc, lib().nbody_gpu_last_error().decode()))


def device_count():
    return lib().nbody_gpu_device_count()


def nccl_unique_id():
    buf = ctypes.create_string_buffer(NCCL_ID_BYTES)
    _check(lib().nbody_gpu_nccl_unique_id(buf))
    return buf.raw


def partition(n, nranks, rank):
    lo, hi = ctypes.c_int(), ctypes.c_int()
    lib().nbody_gpu_partition(n, nranks, rank, ctypes.byref(lo), ctypes.byref(hi))
    return lo.value, hi.value


def partition_reference(n, nranks):
    counts = (ctypes.c_int * nranks)()
    displs = (ctypes.c_int * nranks)()
    lib().nbody_gpu_partition_reference(n, nranks, counts, displs)
    return list(counts), list(displs)


def forces_static(n, pos_xy_ptr, mass_radius_ptr, force_xy_ptr, stream=None):
    """nbody_gpu_forces_static: the fast-mode force pass on DEVICE arrays the caller owns (integer addresses:
    interleaved x/y, mass/radius in, xf/yf out; 2n doubles each, 16-byte aligned, on the current device), enqueued
    on `stream` (a cudaStream_t as an integer; None = the default stream).  Allocates nothing -- the partial sums
    sit in a fixed array inside the library -- so n may not exceed STATIC_MAX_BODIES.  For callers whose device
    memory belongs to someone else's pool, or who cannot afford a cudaMalloc stall inside their loop."""
    _check(lib().nbody_gpu_forces_static(int(n), pos_xy_ptr, mass_radius_ptr, force_xy_ptr, stream))


def shard_plan(n, mode, nranks, rank):
    """How rank `rank` of `nranks` GPUs shares n bodies (host arithmetic of the library's sharding)."""
    out = (ctypes.c_int * 8)()
    ev = ctypes.c_longlong()
    _check(lib().nbody_gpu_shard_plan(n, mode, nranks, rank, out, ctypes.byref(ev)))
    return dict(lo=out[0], hi=out[1], v0=out[2], nv=out[3], row_blocks=out[4], chunks=out[5], long_chunks=out[6],
                symmetric=bool(out[7]), pair_evaluations=ev.value)


def sym_work_items(n, nranks=1, rank=0):
    """The symmetric kernel's work items of one rank in dispatch order: dict of int32 arrays `blocks`, `chunks`, `groups`,
    `guards` (one entry per item = CTA), `chunk_start` (chunk s covers bodies chunk_start[s]..chunk_start[s+1]) and
    the length `ring` of the row-sum slice ring."""
    count, nchunks, ring = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    _check(lib().nbody_gpu_sym_work_items(n, nranks, rank, None, None, None, None, 0, ctypes.byref(count), None, 0,
                                          ctypes.byref(nchunks), ctypes.byref(ring)))
    cols = {k: np.empty(count.value, dtype=np.int32) for k in ("blocks", "chunks", "groups", "guards")}
    start = np.empty(nchunks.value + 1, dtype=np.int32)
    ip = ctypes.POINTER(ctypes.c_int)
    _check(lib().nbody_gpu_sym_work_items(n, nranks, rank, cols["blocks"].ctypes.data_as(ip), cols["chunks"].ctypes.data_as(ip),
                                          cols["groups"].ctypes.data_as(ip), cols["guards"].ctypes.data_as(ip), count.value,
                                          ctypes.byref(count), start.ctypes.data_as(ip), nchunks.value + 1,
                                          ctypes.byref(nchunks), ctypes.byref(ring)))
    cols["chunk_start"] = start
    cols["ring"] = ring.value
    return cols


def reference_init(n, xdim=1024, ydim=768):
    """The reference's synthetic initial state (nbody-par.c:208-222): (x, y, xv, yv, mass, radius)."""
    out = [np.empty(n) for _ in range(6)]
    lib().nbody_gpu_reference_init(n, xdim, ydim, *out)
    return tuple(out)


def cell_key(x, y, xdim=1024, ydim=768):
    """Hilbert cell index the device sorts positions by (host mirror)."""
    return int(lib().nbody_gpu_cell_key(float(x), float(y), xdim, ydim))


def fp64_peak(device=-1, ms_budget=200):
    """Measured FP64 DFMA peak of the device: (TFLOP/s, implied SM MHz)."""
    tf, mhz = ctypes.c_double(), ctypes.c_double()
    _check(lib().nbody_gpu_fp64_peak(device, ms_budget, ctypes.byref(tf), ctypes.byref(mhz)))
    return tf.value, mhz.value


def pin_host(a):
    """Page-lock a numpy array the caller keeps for the whole run (set_state / fetch then copy at full PCIe speed)."""
    _check(lib().nbody_gpu_pin_host(a.ctypes.data_as(ctypes.c_void_p), a.nbytes))


def unpin_host(a):
    _check(lib().nbody_gpu_unpin_host(a.ctypes.data_as(ctypes.c_void_p)))


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class NBodyGPU:
    """One N-body world on the GPU(s). Arrays are host numpy float64 of length n."""

    def __init__(self, x, y, xv, yv, mass, radius, xdim=1024, ydim=768, mode=MODE_EXACT,
                 ngpus=1, device=-1, rank=0, nranks=1, nccl_id=None, jsplit=0, use_graph=0, symmetric=0, p2p=0, sort=0,
                 sort_interval=0):
        x, y, xv, yv, mass, radius = map(_f64, (x, y, xv, yv, mass, radius))
        self.n = int(x.shape[0])
        for a in (y, xv, yv, mass, radius):
            if a.shape != (self.n,):
                raise ValueError("all body arrays must have the same length")
        self._id = ctypes.create_string_buffer(nccl_id, NCCL_ID_BYTES) if nccl_id else None
        cfg = _Cfg(self.n, xdim, ydim, mode, ngpus, device, rank, nranks,
                   ctypes.cast(self._id, ctypes.c_void_p) if self._id else None, jsplit, use_graph, symmetric, p2p, sort, sort_interval)
        self._h = ctypes.c_void_p()
        _check(lib().nbody_gpu_init(ctypes.byref(self._h), ctypes.byref(cfg), x, y, xv, yv, mass, radius))

    def set_state(self, x, y, xv, yv):
        _check(lib().nbody_gpu_set_state(self._h, _f64(x), _f64(y), _f64(xv), _f64(yv)))

    def step(self, steps=1):
        _check(lib().nbody_gpu_step(self._h, int(steps)))

    def step_timed(self, steps=1):
        """Run `steps` steps bracketed by CUDA events on the library's stream; returns milliseconds."""
        ms = ctypes.c_float()
        _check(lib().nbody_gpu_step_timed(self._h, int(steps), ctypes.byref(ms)))
        return ms.value

    def step_bench(self, steps=1, flush_l2=True):
        """The production step path (CUDA graphs / one-launch loop) between two CUDA events, with an L2 flush before
        every launch unit; returns milliseconds."""
        ms = ctypes.c_float()
        _check(lib().nbody_gpu_step_bench(self._h, int(steps), int(bool(flush_l2)), ctypes.byref(ms)))
        return ms.value

    def step_profiled(self, steps=1, flush_l2=False):
        """Plain launches with events between phases: dict(total, force, integrate, exchange) in ms."""
        out = (ctypes.c_float * 4)()
        _check(lib().nbody_gpu_step_profiled(self._h, int(steps), int(bool(flush_l2)), out))
        return dict(total=out[0], force=out[1], integrate=out[2], exchange=out[3])

    def forces(self):
        xf, yf = np.empty(self.n), np.empty(self.n)
        _check(lib().nbody_gpu_forces(self._h, xf, yf))
        return xf, yf

    def fetch(self, out=None):
        """(x, y, xf, yf, xv, yv): the six columns the reference's print() shows.  `out`: six float64 arrays of
        length n to fill instead of allocating new ones."""
        if out is None:
            out = [np.empty(self.n) for _ in range(6)]
        else:
            out = list(out)
            for a in out:
                if a.dtype != np.float64 or a.shape != (self.n,) or not a.flags.c_contiguous:
                    raise ValueError("fetch(out=...) needs six contiguous float64 arrays of length n")
        _check(lib().nbody_gpu_fetch(self._h, *[a.ctypes.data_as(ctypes.c_void_p) for a in out]))
        return tuple(out)

    def fetch_range(self, first, count, out=None):
        """fetch() for the bodies [first, first + count) only: six arrays of length count."""
        if out is None:
            out = [np.empty(count) for _ in range(6)]
        else:
            out = list(out)
            for a in out:
                if a.dtype != np.float64 or a.shape != (count,) or not a.flags.c_contiguous:
                    raise ValueError("fetch_range(out=...) needs six contiguous float64 arrays of length count")
        _check(lib().nbody_gpu_fetch_range(self._h, int(first), int(count), *[a.ctypes.data_as(ctypes.c_void_p) for a in out]))
        return tuple(out)

    def render(self, xdim, ydim):
        """The reference's display() of the current positions: uint8 image (ydim, xdim, 3)."""
        img = np.zeros((ydim, xdim, 3), dtype=np.uint8)
        _check(lib().nbody_gpu_render(self._h, img))
        return img

    def far_fraction(self):
        """Share of the symmetric kernel's pair evaluations currently in far-field tiles (0 if not in cell order)."""
        f = ctypes.c_double()
        _check(lib().nbody_gpu_far_fraction(self._h, ctypes.byref(f)))
        return f.value

    def sync(self):
        _check(lib().nbody_gpu_sync(self._h))

    def describe(self):
        import json
        buf = ctypes.create_string_buffer(1024)
        lib().nbody_gpu_describe(self._h, buf, 1024)
        return json.loads(buf.value.decode())

    def launch_count(self):
        out = (ctypes.c_longlong * 4)()
        total = lib().nbody_gpu_launch_count(self._h, out)
        return total, list(out)

    def close(self):
        if self._h:
            lib().nbody_gpu_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
