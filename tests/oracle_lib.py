"""ctypes loader for oracle/liboracle.so (the CPU checker). Test infrastructure only."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

_dp = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_lib = None


def build():
    subprocess.run(["make", "-C", ORACLE_DIR, "--no-print-directory"], check=True,
                   stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is not None:
        return _lib
    path = os.path.join(ORACLE_DIR, "liboracle.so")
    if not os.path.exists(path):
        build()
    L = ctypes.CDLL(path)
    i = ctypes.c_int
    L.oracle_init_bodies.argtypes = [i, i, i] + [_dp] * 6
    L.oracle_forces_symmetric.argtypes = [i, i] + [_dp] * 6
    L.oracle_forces_rows.argtypes = [i, i, i, i, i] + [_dp] * 6
    L.oracle_forces_row_list.argtypes = [i, i, i, np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS"), i] + [_dp] * 6
    L.oracle_forces_symmetric_sample.argtypes = [i, i, i, i] + [_dp] * 6
    L.oracle_forces_symmetric_sample.restype = ctypes.c_longlong
    L.oracle_velocities.argtypes = [i, i] + [_dp] * 5
    L.oracle_positions.argtypes = [i, i, i] + [_dp] * 6
    L.oracle_run.argtypes = [i, i, i, i, i, i] + [_dp] * 8
    L.oracle_run.restype = i
    L.oracle_display.argtypes = [i, i, i, _dp, _dp, _dp, np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")]
    L.oracle_display.restype = None
    L.oracle_nr_flops.argtypes = [ctypes.c_longlong, ctypes.c_longlong]
    L.oracle_nr_flops.restype = ctypes.c_longlong
    for f in ("oracle_det_atan2_v", "oracle_libm_atan2_v", "oracle_det_sincos_v", "oracle_libm_sincos_v"):
        getattr(L, f).argtypes = [i, _dp, _dp, _dp]
    _lib = L
    return L


class Bodies:
    """Structure-of-arrays state: x, y, xf, yf, xv, yv, mass, radius (float64, length n)."""

    FIELDS = ("x", "y", "xf", "yf", "xv", "yv", "mass", "radius")

    def __init__(self, n):
        self.n = n
        for f in self.FIELDS:
            setattr(self, f, np.zeros(n, dtype=np.float64))

    def copy(self):
        o = Bodies(self.n)
        for f in self.FIELDS:
            getattr(o, f)[:] = getattr(self, f)
        return o


def init_bodies(n, xdim=1024, ydim=768):
    b = Bodies(n)
    lib().oracle_init_bodies(n, xdim, ydim, b.x, b.y, b.xv, b.yv, b.mass, b.radius)
    return b


def run(b, steps, trig, xdim=1024, ydim=768, threads=1):
    """Advance b in place by `steps` steps. trig: 0 = libm, 1 = detmath."""
    rc = lib().oracle_run(b.n, xdim, ydim, steps, trig, threads,
                          b.x, b.y, b.xf, b.yf, b.xv, b.yv, b.mass, b.radius)
    assert rc == 0
    return b


def forces_for(rows, x, y, mass, radius, trig=0, threads=0):
    """Pair-force sums of the listed bodies only (oracle row form): (xf[rows], yf[rows])."""
    n = len(x)
    rows = np.ascontiguousarray(rows, dtype=np.int32)
    xf, yf = np.zeros(n), np.zeros(n)
    lib().oracle_forces_row_list(n, trig, len(rows), rows, threads or (os.cpu_count() or 1),
                                 np.ascontiguousarray(x), np.ascontiguousarray(y), mass, radius, xf, yf)
    return xf[rows], yf[rows]


def step_bodies(rows, x, y, xv, yv, xf_rows, yf_rows, mass, trig=0, xdim=1024, ydim=768):
    """compute_velocities + compute_positions (oracle) for the listed bodies, given their force sums:
    returns (x', y', xv', yv') of those bodies."""
    rows = np.asarray(rows)
    m = len(rows)
    sxv, syv = np.ascontiguousarray(xv[rows]), np.ascontiguousarray(yv[rows])
    lib().oracle_velocities(m, trig, np.ascontiguousarray(xf_rows), np.ascontiguousarray(yf_rows),
                            np.ascontiguousarray(mass[rows]), sxv, syv)
    xn, yn = np.zeros(m), np.zeros(m)
    lib().oracle_positions(m, xdim, ydim, np.ascontiguousarray(x[rows]), np.ascontiguousarray(y[rows]), sxv, syv, xn, yn)
    return xn, yn, sxv, syv


def cell_order(nb, x, y, xdim=1024, ydim=768):
    """Slot order of the library's cell order for these positions: bodies sorted by (Hilbert cell key, index),
    from the library's host mirror of the key.  order[s] = body in slot s."""
    keys = np.fromiter((nb.cell_key(float(a), float(b), xdim, ydim) for a, b in zip(x, y)), dtype=np.int64, count=len(x))
    return np.lexsort((np.arange(len(x)), keys))


def block_sample(order, rb=512, per_block_extra=1, seed=1):
    """Bodies that sit at row-block boundaries of the symmetric kernel (first and last slot of every block of
    `rb` slots) plus `per_block_extra` random slots per block."""
    n = len(order)
    rng = np.random.default_rng(seed)
    slots = []
    for b0 in range(0, n, rb):
        b1 = min(n, b0 + rb)
        slots += [b0, b1 - 1]
        slots += list(rng.integers(b0, b1, per_block_extra))
    return np.unique(order[np.array(slots)])


def display(b, xdim=1024, ydim=768):
    """The reference's display() on the CPU: uint8 image (ydim, xdim, 3)."""
    img = np.zeros((ydim, xdim, 3), dtype=np.uint8)
    lib().oracle_display(b.n, xdim, ydim, b.x, b.y, b.radius, img)
    return img


def format_state(x, y, xf, yf, xv, yv):
    """The reference's print(): nbody-seq.c:356-365."""
    return "".join("%10.3f %10.3f %10.3f %10.3f %10.3f %10.3f\n" % t
                   for t in zip(x, y, xf, yf, xv, yv))


def golden_text(name):
    with open(os.path.join(GOLDEN_DIR, name)) as f:
        return f.read()


def write_ppm(path, xdim=1024, ydim=768):
    """A P6 file of the shipped nbody.ppm's dimensions; only the header is read when secs_per_update = 0."""
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (xdim, ydim))
        f.write(bytes(xdim * ydim * 3))


def ref_binary(name):
    p = os.path.join(REF_DIR, name)
    return p if os.path.exists(p) else None


def run_ref_hex(binary, n, steps, ppm):
    """Run an oracle/_ref '-hex' build and parse its %a state dump into six float64 arrays."""
    out = subprocess.run([binary, str(n), "0", ppm, str(steps)], check=True,
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL).stdout.decode()
    rows = [[float.fromhex(t) for t in line.split()] for line in out.splitlines() if line.strip()]
    a = np.array(rows, dtype=np.float64)
    assert a.shape == (n, 6)
    return [np.ascontiguousarray(a[:, k]) for k in range(6)]
