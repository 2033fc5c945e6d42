"""The C host driver (nbody/nbody-par, nbody/nbody-seq) behind the reference's command line, and the
reference's own acceptance scripts run UNCHANGED against it through the bin/prun shim.
The scripts are staged, unmodified and git-ignored, under oracle/_ref/bin by oracle/Makefile."""
import os
import shutil
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PAR = os.path.join(ROOT, "nbody", "nbody-par")
SEQ = os.path.join(ROOT, "nbody", "nbody-seq")


def _build():
    subprocess.run(["make", "-C", os.path.join(ROOT, "nbody"), "--no-print-directory"], check=True,
                   stdout=subprocess.DEVNULL)


@pytest.fixture(scope="module")
def workdir(tmp_path_factory, oracle, nb):
    """A directory laid out like the reference's repo root: nbody/nbody-par, nbody.ppm, bin/, tests/."""
    _build()
    d = tmp_path_factory.mktemp("refroot")
    os.makedirs(d / "nbody")
    os.makedirs(d / "bin")
    os.makedirs(d / "tests")
    os.symlink(PAR, d / "nbody" / "nbody-par")
    os.symlink(SEQ, d / "nbody" / "nbody-seq")
    oracle.write_ppm(str(d / "nbody.ppm"))
    shutil.copy(os.path.join(oracle.GOLDEN_DIR, "REF_OUTPUT_32x100000.txt"), d / "bin" / "REF_OUTPUT")
    shutil.copy(os.path.join(oracle.GOLDEN_DIR, "128_MY_REF_OUTPUT_128x1000.txt"), d / "tests" / "128_MY_REF_OUTPUT")
    return d


def _env():
    e = dict(os.environ)
    e["PATH"] = os.path.join(ROOT, "bin") + os.pathsep + e["PATH"]
    e["PRUN_ETC"] = "/nonexistent"
    return e


def test_usage_and_errors_without_gpu_work(tmp_path):
    _build()
    r = subprocess.run([PAR], capture_output=True, text=True)
    assert r.returncode == 1 and "Usage:" in r.stderr and "num_bodies secs_per_update ppm_output_file steps" in r.stderr
    r = subprocess.run([PAR, "8", "0", str(tmp_path / "missing.ppm"), "1"], capture_output=True, text=True)
    assert r.returncode == 1 and "Cannot read" in r.stderr
    bad = tmp_path / "bad.ppm"
    bad.write_bytes(b"P5\n4 4\n255\n")
    r = subprocess.run([PAR, "8", "0", str(bad), "1"], capture_output=True, text=True)
    assert r.returncode == 1 and "Cannot read" in r.stderr


@pytest.mark.gpu
def test_driver_prints_reference_format_and_goldens(workdir, oracle):
    r = subprocess.run(["nbody/nbody-par", "128", "0", "nbody.ppm", "1000"], cwd=workdir, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert r.stdout == oracle.golden_text("128_MY_REF_OUTPUT_128x1000.txt")
    assert "Running N-body with 128 bodies and 1000 steps" in r.stderr
    assert "N-body took:" in r.stderr and "Performance N-body:" in r.stderr and "GFLOPS" in r.stderr
    r = subprocess.run(["nbody/nbody-seq", "32", "0", "nbody.ppm", "100000"], cwd=workdir, capture_output=True, text=True)
    assert r.stdout == oracle.golden_text("REF_OUTPUT_32x100000.txt")
    # fast mode prints the same text
    r = subprocess.run(["nbody/nbody-par", "128", "0", "nbody.ppm", "1000", "--mode", "fast"], cwd=workdir,
                       capture_output=True, text=True)
    assert r.stdout == oracle.golden_text("128_MY_REF_OUTPUT_128x1000.txt")


@pytest.mark.gpu
def test_driver_hex_state_bit_identical_to_reference_source(workdir, oracle, ppm):
    """nbody-par --hex against oracle/_ref/nbody-seq-det-hex (the reference source + detmath)."""
    exe = oracle.ref_binary("nbody-seq-det-hex")
    if exe is None:
        pytest.skip("oracle/_ref not built")
    for n, steps in ((2, 10), (300, 25), (1, 3)):
        r = subprocess.run(["nbody/nbody-par", str(n), "0", "nbody.ppm", str(steps), "--hex"], cwd=workdir,
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        nn = max(n, 2)    # both programs clamp to two bodies
        got = np.array([[float.fromhex(t) for t in ln.split()] for ln in r.stdout.splitlines()])
        ref = oracle.run_ref_hex(exe, nn, steps, ppm)
        for k in range(6):
            assert np.array_equal(got[:, k].view(np.uint64), ref[k].view(np.uint64)), (n, k)


@pytest.mark.gpu
def test_run_xps_csv_shape(workdir):
    r = subprocess.run(["nbody/nbody-par", "64", "0", "nbody.ppm", "10", "--run-xps", "--measure-comm"], cwd=workdir,
                       capture_output=True, text=True)
    f = r.stdout.strip().split(",")
    assert len(f) == 8 and f[3] == "64"      # SIZE,NODES,CPUS_PER_NODE,NBODIES,RTIME,COMM,COMM/RTIME,GFLOPS
    r = subprocess.run(["nbody/nbody-seq", "64", "0", "nbody.ppm", "10", "--run-xps"], cwd=workdir,
                       capture_output=True, text=True)
    assert len(r.stdout.strip().split(",")) == 3


@pytest.mark.gpu
def test_reference_experiment_sweep_script_unchanged(workdir, oracle, nb):
    """bin/experiments.sh run UNCHANGED through the prun shim against the GPU-backed nbody-par: it
    passes `--openmp --measure-comm --run-xps` (experiments.sh:42-46) and appends one CSV line per run
    (nbody-par.c:954-957: SIZE,NODES,CPUS_PER_NODE,NBODIES,RTIME,COMMTIME,COMMTIME:RTIME,GFLOPS).  The shim's
    PRUN_SHIM_MAX_RUNS stops the 144-run sweep after the first runs; NODES / CPUS_PER_NODE must come back from the
    PRUN_ENVIRONMENT file the launcher writes, as in the reference (nbody-par.c:441-517)."""
    script = os.path.join(oracle.REF_DIR, "bin", "experiments.sh")
    if not os.path.exists(script):
        pytest.skip("reference scripts not staged (oracle/_ref/bin)")
    shutil.copy(script, workdir / "bin" / "experiments.sh")
    out = workdir / "xps.csv"
    for f in (out, workdir / ".prun_shim_count"):
        if f.exists():
            f.unlink()
    env = _env()
    env["PRUN_SHIM_MAX_RUNS"] = "7"
    r = subprocess.run(["bash", "bin/experiments.sh", "1", "xps.csv", "1"], cwd=workdir, env=env, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.count("N_BODIES=") == 144          # the script walked its whole grid
    lines = open(out).read().splitlines()
    assert len(lines) == 7
    want = [(512, 2, 1), (512, 2, 4), (512, 2, 16), (512, 4, 1), (512, 4, 4), (512, 4, 16), (512, 6, 1)]
    gpus = nb.device_count()
    for ln, (n, nodes, cpus) in zip(lines, want):
        f = ln.split(",")
        assert len(f) == 8, ln
        size, got_nodes, got_cpus, got_n = int(f[0]), int(f[1]), int(f[2]), int(f[3])
        rtime, comm, ratio, gflops = float(f[4]), float(f[5]), float(f[6]), float(f[7])
        assert (got_nodes, got_cpus, got_n) == (nodes, cpus, n), ln
        assert size == min(nodes * cpus, gpus, n), ln       # the GPUs that shared the rows
        assert rtime > 0 and gflops > 0 and 0 <= ratio <= 1.0, ln
        if size >= 2:
            assert comm > 0, ln                             # device-timed exchange of the sharded step
        else:
            assert comm == 0, ln
    # the sequential sweep's CSV (nbody-seq.c:488): N,rtime,gflops
    script = os.path.join(oracle.REF_DIR, "bin", "sequential-experiments.sh")
    shutil.copy(script, workdir / "bin" / "sequential-experiments.sh")
    (workdir / ".prun_shim_count").unlink()
    env["PRUN_SHIM_MAX_RUNS"] = "2"
    r = subprocess.run(["bash", "bin/sequential-experiments.sh", "seq.csv", "1"], cwd=workdir, env=env, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = open(workdir / "seq.csv").read().splitlines()
    assert [ln.split(",")[0] for ln in lines] == ["512", "1024"] and all(len(ln.split(",")) == 3 for ln in lines)
    (workdir / ".prun_shim_count").unlink()


@pytest.mark.gpu
def test_reference_sanity_check_script_unchanged(workdir, oracle):
    script = os.path.join(oracle.REF_DIR, "bin", "nbody-sanity-check")
    if not os.path.exists(script):
        pytest.skip("reference scripts not staged (oracle/_ref/bin)")
    shutil.copy(script, workdir / "bin" / "nbody-sanity-check")
    r = subprocess.run(["sh", "bin/nbody-sanity-check"], cwd=workdir, env=_env(), capture_output=True, text=True)
    assert "Output ok" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
    assert "wrong output" not in r.stdout and "did not print run time" not in r.stdout


@pytest.mark.gpu
def test_reference_grid_test_script_unchanged(workdir, oracle):
    script = os.path.join(oracle.REF_DIR, "bin", "tests.sh")
    if not os.path.exists(script):
        pytest.skip("reference scripts not staged (oracle/_ref/bin)")
    shutil.copy(script, workdir / "bin" / "tests.sh")
    golden = oracle.golden_text("128_MY_REF_OUTPUT_128x1000.txt")
    for np_, cpus, extra in ((1, 1, []), (2, 4, ["--openmp"]), (8, 16, [])):
        r = subprocess.run(["bash", "bin/tests.sh", str(np_), str(cpus), "128"] + extra, cwd=workdir, env=_env(),
                           capture_output=True, text=True)
        out = open(workdir / "tests" / ("%d_%d_128_mynbody.test.out" % (np_, cpus))).read()
        assert out == golden + "STATUS=PASSED\n", r.stdout[-500:]
        # the script prints `diff golden output`: it must be empty between the banner and the verdict
        body = r.stdout.replace(golden, "")
        assert "<" not in body and ">" not in body


@pytest.mark.gpu
def test_strict_runner(workdir, oracle):
    r = subprocess.run(["bash", os.path.join(ROOT, "bin", "strict-check.sh"), "bin/REF_OUTPUT", "tests/128_MY_REF_OUTPUT"],
                       cwd=workdir, env=_env(), capture_output=True, text=True)
    assert r.returncode == 0, r.stdout
    assert r.stdout.count("STATUS=PASSED") == 25 and "FAILED" not in r.stdout


@pytest.mark.gpu
def test_driver_display_final_writes_ppm(workdir, oracle):
    """secs_per_update / --display-final: the CUDA rasteriser's image lands behind the P6 header."""
    ppm2 = str(workdir / "img.ppm")
    oracle.write_ppm(ppm2, 256, 192)
    r = subprocess.run(["nbody/nbody-par", "64", "0", "img.ppm", "50", "--display-final"], cwd=workdir,
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    data = open(ppm2, "rb").read()
    hdr = b"P6\n256 192\n255\n"
    assert data.startswith(hdr) and len(data) == len(hdr) + 256 * 192 * 3
    img = np.frombuffer(data[len(hdr):], dtype=np.uint8).reshape(192, 256, 3)
    b = oracle.init_bodies(64, 256, 192)
    oracle.run(b, 50, trig=1, xdim=256, ydim=192)
    assert np.array_equal(img, oracle.display(b, 256, 192))
    r = subprocess.run(["nbody/nbody-seq", "64", "1", "img.ppm", "40"], cwd=workdir, capture_output=True, text=True)
    assert r.returncode == 0 and open(ppm2, "rb").read().startswith(hdr)
