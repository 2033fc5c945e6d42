#!/usr/bin/env python
"""Generates tests/golden/trajectory_<N>x<STEPS>.npz: the FINAL state of the reference program itself after a whole
run, for the accumulated-trajectory parity tests of fast mode (north_star: "within a stated max relative error on
final positions"; the loop is nbody-seq.c:457-464, the surface nbody-seq.c:356-365).

The generator is oracle/_ref/nbody-seq-hex: the reference's nbody-seq.c compiled where it lies under
/root/reference by oracle/Makefile (patched only for MAXBODIES and a "%a" print format), stock libm.  It exists in
the build container only, hence the committed fixtures.  One core, about 60 ns per pair:
    8192 x 200 steps ~ 7 min,  16384 x 50 steps ~ 7 min.

    python tests/golden/make_trajectory_fixtures.py [N STEPS]...
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib  # noqa: E402

DEFAULT = [(8192, 200), (16384, 50)]


def main(argv):
    cases = [(int(argv[i]), int(argv[i + 1])) for i in range(0, len(argv) - 1, 2)] or DEFAULT
    exe = oracle_lib.ref_binary("nbody-seq-hex")
    if exe is None:
        sys.exit("oracle/_ref/nbody-seq-hex is missing: run `make -C oracle` where /root/reference exists")
    with tempfile.TemporaryDirectory() as tmp:
        ppm = os.path.join(tmp, "nbody.ppm")
        oracle_lib.write_ppm(ppm)
        for n, steps in cases:
            x, y, xf, yf, xv, yv = oracle_lib.run_ref_hex(exe, n, steps, ppm)
            out = os.path.join(HERE, "trajectory_%dx%d.npz" % (n, steps))
            np.savez_compressed(out, n=n, steps=steps, x=x, y=y, xf=xf, yf=yf, xv=xv, yv=yv,
                                generator="oracle/_ref/nbody-seq-hex (reference nbody-seq.c, stock libm)")
            print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main(sys.argv[1:])
