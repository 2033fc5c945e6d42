"""Development aid: opcode histogram of the hottest loop (the one with most DFMA) of a kernel in an object file.
usage: python tools/sass_loop_count.py kernels_fast.o 'forces_sym_kernelILi128ELi4ELi1ELi4ELi1EE'"""
import collections, re, subprocess, sys

obj, pat = sys.argv[1], sys.argv[2]
want_shfl = len(sys.argv) > 3 and sys.argv[3] in ("shfl", "allshfl")
want_all = len(sys.argv) > 3 and sys.argv[3] in ("all", "allshfl")


def report(name, body):
    hist = collections.Counter()
    for _, x in body:
        x = re.sub(r"^@!?U?P\d+\s+", "", x)
        hist[x.split()[0].split(".")[0]] += 1
    fp64 = sum(v for k2, v in hist.items() if k2 in ("DFMA", "DMUL", "DADD", "DSETP"))
    mufu = hist.get("MUFU", 0)
    other = sum(hist.values()) - fp64 - mufu
    pairs = max(mufu, 1)
    print(name)
    print("  loop %d instr, %d pairs: FP64 %.2f/pair, MUFU %.2f, other %.2f  -> model %.1f issue cycles/pair" %
          (len(body), pairs, fp64 / pairs, mufu / pairs, other / pairs, (2 * fp64 + 2 * mufu + other) / pairs))
    print("  " + " ".join("%s:%d" % kv for kv in hist.most_common()))

txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True, check=True).stdout
cur, funcs = None, {}
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        funcs[cur] = []
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
    if m and cur:
        funcs[cur].append((int(m.group(1), 16), m.group(2).strip()))
for name, ins in funcs.items():
    if pat not in name:
        continue
    addr_index = {a: k for k, (a, _) in enumerate(ins)}
    best = None
    loops = []
    for k, (a, t) in enumerate(ins):
        m = re.search(r"BRA(?:\.U)?(?:\.ANY)?\s+(?:!?U?P\d,\s*)?(0x[0-9a-f]+)", t)
        if not m:
            continue
        tgt = int(m.group(1), 16)
        if tgt <= a and tgt in addr_index:
            loops.append((addr_index[tgt], k))
    for lo, hi in loops:
        if any((l2, h2) != (lo, hi) and lo <= l2 and h2 <= hi for l2, h2 in loops):
            continue   # not innermost
        body = ins[lo:hi + 1]
        nd = sum(1 for _, x in body if re.match(r"(@!?U?P\d+\s+)?DFMA", x))
        if want_shfl and not any("SHFL" in x for _, x in body):
            continue
        if want_all and nd > 16:
            report(name, body)
        if best is None or nd > best[0]:
            best = (nd, body)
    if want_all:
        continue
    report(name, best[1])
