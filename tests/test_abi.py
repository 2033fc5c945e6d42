"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/nbody_gpu.h declares,
its pure host helpers match the reference's rules, and without a GPU it fails loudly (no fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "nbody_gpu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(nbody_gpu_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(nb):
    L = ctypes.CDLL(nb.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 15
    for nm in names:
        assert hasattr(L, nm), nm
    assert L.nbody_gpu_abi_version() == 5


def test_header_compiles_as_c(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "nbody_gpu.h"\nint main(void){ nbody_gpu_cfg c = {0}; return c.n; }\n')
    import subprocess
    subprocess.run(["/usr/bin/gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", str(src),
                    "-o", str(tmp_path / "t.o")], check=True)


def ref_partition(n, p):
    """Python restatement of nbody-par.c:225-250."""
    counts = [n // p] * p
    r = n % p
    while r > 0:
        counts[r] += 1
        r -= 1
    displs = [0] * p
    for k in range(1, p):
        displs[k] = displs[k - 1] + counts[k - 1]
    return counts, displs


@pytest.mark.parametrize("n,p", [(128, 1), (128, 8), (10, 4), (32, 2), (10000, 6), (7, 8), (262144, 8), (1000, 7)])
def test_partitions(nb, n, p):
    assert nb.partition_reference(n, p) == ref_partition(n, p)
    rows = -(-n // p)
    covered = []
    for r in range(p):
        lo, hi = nb.partition(n, p, r)
        assert lo == min(n, r * rows) and hi == min(n, (r + 1) * rows)
        covered += list(range(lo, hi))
    assert covered == list(range(n))   # every body owned exactly once, in order


def test_argument_errors(nb):
    z = np.zeros(4)
    with pytest.raises(nb.NBodyGPUError, match="mode"):
        nb.NBodyGPU(z, z, z, z, z + 1, z + 1, mode=7)
    with pytest.raises(nb.NBodyGPUError, match="xdim"):
        nb.NBodyGPU(z, z, z, z, z + 1, z + 1, xdim=0)
    with pytest.raises(ValueError):
        nb.NBodyGPU(z, z[:3], z, z, z, z)


def test_no_silent_cpu_fallback(nb):
    """Without a usable CUDA device the product path must raise, never compute on the CPU."""
    if nb.device_count() > 0:
        pytest.skip("a GPU is present")
    z = np.ones(8)
    with pytest.raises(nb.NBodyGPUError, match="CUDA|cuda|device"):
        nb.NBodyGPU(z, z, z, z, z, z)
    with pytest.raises(nb.NBodyGPUError):
        nb.fp64_peak(0, 10)


def test_product_sources_do_not_reference_the_oracle():
    """oracle/ is test infrastructure: nothing under the package, include/ or nbody/ may name it."""
    bad = []
    for top in ("parallel-nbody_b200", "include", "nbody"):
        for dp, _, files in os.walk(os.path.join(ROOT, top)):
            for f in files:
                if f.endswith((".cu", ".cuh", ".h", ".c", ".py", "Makefile")):
                    txt = open(os.path.join(dp, f), errors="ignore").read()
                    if re.search(r"#\s*include[^\n]*oracle|liboracle|oracle_[a-z_]+\s*\(|import\s+oracle|-loracle", txt):
                        bad.append(os.path.join(dp, f))
    assert not bad, bad


@pytest.mark.parametrize("n", [1, 32, 128, 4096, 50000])
def test_library_initialiser_is_the_reference_one(nb, oracle, n):
    """nbody_gpu_reference_init (host helper used by bench.py) against the oracle's restatement of
    nbody-par.c:208-222, which tests/test_oracle.py pins to the reference program's own output."""
    got = nb.reference_init(n)
    want = oracle.init_bodies(n)
    for a, f in zip(got, ("x", "y", "xv", "yv", "mass", "radius")):
        assert np.array_equal(a.view(np.uint64), getattr(want, f).view(np.uint64)), f


def test_cell_key_is_a_hilbert_curve(nb):
    """The sort key of the cell order: a bijection of the 65536 x 65536 cells onto 32 bits whose consecutive
    indices are edge-neighbouring cells (checked on the cells of a 64 x 64 pixel world, which the scale maps to a
    1024-cell stride, and on unit cells of a 65535-pixel world), monotone in nothing else; out-of-range and NaN
    positions clamp instead of wrapping."""
    side = 65535
    cells = {}
    for cx in range(40):
        for cy in range(40):
            k = nb.cell_key(cx, cy, side, side)          # scale 1: pixel = cell
            assert k not in cells
            cells[k] = (cx, cy)
    keys = sorted(cells)
    steps = [abs(cells[a][0] - cells[b][0]) + abs(cells[a][1] - cells[b][1]) for a, b in zip(keys, keys[1:]) if b == a + 1]
    assert steps and all(s == 1 for s in steps)          # consecutive indices = neighbouring cells
    # an aligned 2^k x 2^k square of cells occupies one contiguous index range (locality of the curve)
    sq = sorted(k for k, (cx, cy) in cells.items() if cx < 32 and cy < 32)
    assert sq[-1] - sq[0] == 32 * 32 - 1
    assert nb.cell_key(-5.0, 3.0, side, side) == nb.cell_key(0.0, 3.0, side, side)
    assert nb.cell_key(1e9, 3.0, side, side) == nb.cell_key(65535.0, 3.0, side, side)
    assert nb.cell_key(float("nan"), 0.0, side, side) == nb.cell_key(0.0, 0.0, side, side)
    # the reference world: 1024 x 768 pixels map to 64 cells per pixel; a sub-cell move keeps the key
    assert nb.cell_key(100.0, 100.0) == nb.cell_key(100.001, 100.001)
    assert nb.cell_key(100.0, 100.0) != nb.cell_key(101.0, 100.0)
