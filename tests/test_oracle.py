"""Pins the oracle (oracle/nbody_oracle.c) before anything trusts it.

1. text level: the reference's own golden outputs: tests/128_MY_REF_OUTPUT
   (128 bodies x 1000 steps, bin/tests.sh:34-41) and bin/REF_OUTPUT (32 x 100000,
   bin/nbody-sanity-check:16,28), byte for byte, with libm trig AND with detmath trig.
2. bit level: the reference source itself, compiled where it lies by oracle/Makefile into
   oracle/_ref (stock libm, and trig interposed by include/nbody_detmath.h), state dumped
   with %a.  Skipped only if oracle/_ref has not been built (no /root/reference and no
   prebuilt binaries).
"""
import os
import subprocess

import numpy as np
import pytest

FIELDS = ("x", "y", "xf", "yf", "xv", "yv")


def bits_equal(a, b):
    return np.array_equal(np.asarray(a).view(np.uint64), np.asarray(b).view(np.uint64))


@pytest.mark.parametrize("trig", [0, 1])
def test_golden_128x1000(oracle, trig):
    b = oracle.init_bodies(128)
    oracle.run(b, 1000, trig)
    assert oracle.format_state(b.x, b.y, b.xf, b.yf, b.xv, b.yv) == oracle.golden_text("128_MY_REF_OUTPUT_128x1000.txt")


@pytest.mark.parametrize("trig", [0, 1])
def test_golden_32x100000(oracle, trig):
    b = oracle.init_bodies(32)
    oracle.run(b, 100000, trig)
    assert oracle.format_state(b.x, b.y, b.xf, b.yf, b.xv, b.yv) == oracle.golden_text("REF_OUTPUT_32x100000.txt")


@pytest.mark.parametrize("binary,trig", [("nbody-seq-det-hex", 1), ("nbody-seq-hex", 0)])
@pytest.mark.parametrize("n,steps", [(2, 100), (128, 1000), (777, 20), (2048, 3)])
def test_bit_identical_to_reference_source(oracle, ppm, binary, trig, n, steps):
    exe = oracle.ref_binary(binary)
    if exe is None:
        pytest.skip("oracle/_ref/%s not built" % binary)
    ref = oracle.run_ref_hex(exe, n, steps, ppm)
    b = oracle.init_bodies(n)
    oracle.run(b, steps, trig)
    for r, f in zip(ref, FIELDS):
        assert bits_equal(r, getattr(b, f)), f


@pytest.mark.parametrize("binary", ["nbody-seq", "nbody-seq-det"])
def test_reference_builds_reproduce_goldens(oracle, ppm, binary):
    """The unpatched reference source, with libm and with detmath interposed, prints both goldens."""
    exe = oracle.ref_binary(binary)
    if exe is None:
        pytest.skip("oracle/_ref/%s not built" % binary)
    for n, steps, name in ((128, 1000, "128_MY_REF_OUTPUT_128x1000.txt"), (32, 100000, "REF_OUTPUT_32x100000.txt")):
        out = subprocess.run([exe, str(n), "0", ppm, str(steps)], check=True, stdout=subprocess.PIPE,
                             stderr=subprocess.PIPE)
        assert out.stdout.decode() == oracle.golden_text(name)
        assert b"took" in out.stderr   # bin/nbody-sanity-check:18 greps for it


@pytest.mark.parametrize("trig", [0, 1])
def test_row_form_equals_triangular_walk(oracle, trig):
    """The per-body order (-f(j,i) for j<i, then +f(i,j) for j>i, j ascending) reproduces the
    reference's symmetric triangular loop bit for bit: the property the GPU kernel relies on."""
    for n in (2, 3, 64, 501):
        b = oracle.init_bodies(n)
        b.x[n // 2] = b.x[0]
        b.y[n // 2] = b.y[0]          # a coincident pair: atan2(0,0) orientation convention
        L = oracle.lib()
        xs, ys = np.zeros(n), np.zeros(n)
        xr, yr = np.zeros(n), np.zeros(n)
        L.oracle_forces_symmetric(n, trig, b.x, b.y, b.mass, b.radius, xs, ys)
        L.oracle_forces_rows(n, trig, 0, n, 3, b.x, b.y, b.mass, b.radius, xr, yr)
        assert bits_equal(xs, xr) and bits_equal(ys, yr)


def test_threads_do_not_change_bits(oracle):
    a = oracle.init_bodies(300)
    b = a.copy()
    oracle.run(a, 5, 1, threads=1)
    oracle.run(b, 5, 1, threads=4)
    for f in FIELDS:
        assert bits_equal(getattr(a, f), getattr(b, f))


def test_init_matches_reference_formula(oracle):
    """nbody-seq.c:440-449: integer grid positions, radius/mass formula, velocity lattice."""
    n = 1000
    b = oracle.init_bodies(n)
    assert np.all(b.x == np.floor(b.x)) and b.x.min() >= 0 and b.x.max() < 1024
    assert np.all(b.y == np.floor(b.y)) and b.y.max() < 768
    idx = np.arange(n, dtype=np.float64)
    r = 1 + ((idx * idx + 1.0) * 1280.0) / (25.0 * (float(n) * n + 1.0))
    assert bits_equal(b.radius, r)
    assert bits_equal(b.mass, r * r * r)
    assert bits_equal(b.xv, np.round(b.xv * 2000) / 2000.0) and b.xv.min() >= -5 and b.xv.max() < 5
    # large n: formed in double, finite and monotone (the reference's int arithmetic overflows here)
    big = oracle.init_bodies(70000)
    assert np.all(np.diff(big.radius) > 0) and big.radius[-1] < 53


def test_flop_model(oracle):
    # nbody-seq.c:367-380
    assert oracle.lib().oracle_nr_flops(128, 1000) == 1000 * (20 * (128 * 127 // 2) + 22 * 128)
    assert oracle.lib().oracle_nr_flops(262144, 50) == 50 * (20 * (262144 * 262143 // 2) + 22 * 262144)


@pytest.mark.parametrize("ranks,threads,flag", [(1, 1, None), (2, 1, None), (3, 2, "--openmp"), (8, 1, None)])
def test_reference_nbody_par_reproduces_goldens(oracle, ppm, ranks, threads, flag):
    """The reference's parallel program (unmodified nbody-par.c + the MPI stand-in of oracle/mpi_shim) is a
    text-level oracle: any rank count, with and without --openmp, prints the golden state (its bits differ
    from nbody-seq's because the summation order depends on the partition, SURVEY section 0 point 5)."""
    exe = oracle.ref_binary("nbody-par")
    if exe is None:
        pytest.skip("oracle/_ref/nbody-par not built")
    env = dict(os.environ, MPISHIM_NP=str(ranks), OMP_NUM_THREADS=str(threads))
    for n, steps, name in ((128, 1000, "128_MY_REF_OUTPUT_128x1000.txt"), (32, 20000, None)):
        cmd = [exe, str(n), "0", ppm, str(steps)] + ([flag] if flag else [])
        out = subprocess.run(cmd, check=True, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert b"took" in out.stderr
        if name:
            assert out.stdout.decode() == oracle.golden_text(name)
        else:
            b = oracle.init_bodies(n)
            oracle.run(b, steps, trig=0)
            assert out.stdout.decode() == oracle.format_state(b.x, b.y, b.xf, b.yf, b.xv, b.yv)


@pytest.mark.parametrize("n,dims", [(32, (1024, 768)), (128, (1024, 768)), (3000, (320, 200)), (2, (64, 48))])
def test_oracle_display_is_pinned_to_the_reference_display(oracle, tmp_path, n, dims):
    """display() (nbody-seq.c:326-354) has no fixture in the reference, but the reference program itself makes
    one: with secs_per_update = 1 and one step it paints exactly once, after step 1, into the mmapped P6 file
    (`lastup` starts at 0: nbody-seq.c:389,467-471).  The pixels it wrote must equal oracle_display() of the
    oracle's state after one step (libm trig, like the reference binary) -- which pins the restatement the CUDA
    rasteriser is tested against (tests/test_gpu_parity.py::test_render_matches_reference_display)."""
    exe = oracle.ref_binary("nbody-seq")
    if exe is None:
        pytest.skip("oracle/_ref/nbody-seq not built")
    xdim, ydim = dims
    ppm = str(tmp_path / "paint.ppm")
    oracle.write_ppm(ppm, xdim, ydim)
    hdr = b"P6\n%d %d\n255\n" % (xdim, ydim)
    subprocess.run([exe, str(n), "1", ppm, "1"], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    data = open(ppm, "rb").read()
    assert data.startswith(hdr) and len(data) == len(hdr) + xdim * ydim * 3
    painted = np.frombuffer(data[len(hdr):], dtype=np.uint8).reshape(ydim, xdim, 3)
    b = oracle.init_bodies(n, xdim, ydim)
    oracle.run(b, 1, trig=0, xdim=xdim, ydim=ydim)
    ref = oracle.display(b, xdim, ydim)
    assert painted.any()
    assert np.array_equal(painted, ref)
