"""What DESIGN.md says about the shipped force kernel, read back from the SASS of libnbody_gpu.so with cuobjdump
(no GPU needed): the tiles arrive by TMA bulk copies behind mbarriers, nothing spills, the reaction sums are added
by L2 reductions under acquire/release hand-overs, and the far-field loop of the 64 x 8 shape spends 15 FP64-pipe
instructions and one MUFU per unordered pair (the reference's per-pair formula, nbody-seq.c:89-106, restated
without trig).  Skipped when cuobjdump is not installed."""
import collections
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "parallel-nbody_b200", "libnbody_gpu.so")


@pytest.fixture(scope="module")
def sass():
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not installed")
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    funcs, cur = {}, None
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
        if m and cur:
            funcs[cur].append((int(m.group(1), 16), re.sub(r"^@!?U?P\d+\s+", "", m.group(2).strip())))
    return funcs


def _kernel(funcs, pat):
    names = [n for n in funcs if pat in n]
    assert len(names) == 1, names
    return funcs[names[0]]


def _innermost_loops(ins):
    index = {a: k for k, (a, _) in enumerate(ins)}
    loops = []
    for k, (a, t) in enumerate(ins):
        m = re.search(r"BRA(?:\.U)?(?:\.ANY)?\s+(?:!?U?P\d,\s*)?(0x[0-9a-f]+)", t)
        if m and int(m.group(1), 16) <= a and int(m.group(1), 16) in index:
            loops.append((index[int(m.group(1), 16)], k))
    return [ins[lo:hi + 1] for lo, hi in loops
            if not any((l2, h2) != (lo, hi) and lo <= l2 and h2 <= hi for l2, h2 in loops)]


@pytest.mark.parametrize("pat", ["forces_sym_kernelILi64ELi8", "forces_sym_kernelILi128ELi4"])
def test_symmetric_kernel_uses_tma_reductions_and_no_local_memory(sass, pat):
    ops = collections.Counter(t.split()[0] for _, t in _kernel(sass, pat))
    assert any(o.startswith("UBLKCP") for o in ops), "no TMA bulk copy (cp.async.bulk) in the kernel"
    assert any(o.startswith("SYNCS") for o in ops), "no mbarrier instructions"
    assert not any(o.startswith(("STL", "LDL")) for o in ops), "the kernel spills to local memory"
    assert any(o.startswith(("REDG.E.ADD.F64", "ATOMG.E.ADD.F64")) for o in ops), "reaction sums are not L2 reductions"
    assert any(o.startswith("LDG.E.STRONG.GPU") for o in ops), "no acquire load for the hand-overs"
    assert not any(o.startswith(("HMMA", "DMMA", "UTCHMMA", "UTCQMMA")) for o in ops)   # north star: no tensor cores


def test_far_field_loop_is_15_fp64_instructions_per_pair(sass):
    best = None
    for body in _innermost_loops(_kernel(sass, "forces_sym_kernelILi64ELi8")):
        h = collections.Counter(t.split()[0].split(".")[0] for _, t in body)
        if h["MUFU"] < 32 or h["SHFL"] == 0:
            continue                      # not a 32-pair rotation step of the symmetric walk
        fp64 = h["DFMA"] + h["DMUL"] + h["DADD"] + h["DSETP"]
        per_pair = fp64 / h["MUFU"]
        other = (len(body) - fp64 - h["MUFU"]) / h["MUFU"]
        if best is None or per_pair < best[0]:
            best = (per_pair, other)
    assert best is not None
    assert best[0] == 15.0, best          # dx dy | dsqr x2 | f2 q | t w y1 | zi ax ay | zj rx ry
    assert best[1] <= 2.0, best           # LDS / SHFL / address updates amortised over 8 bodies per thread
