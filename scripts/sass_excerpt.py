#!/usr/bin/env python
"""profiles/r02_sass_excerpt.txt: what the SASS of the hot kernels says, readable without running a tool.
usage: python scripts/sass_excerpt.py > profiles/r02_sass_excerpt.txt   (needs cuobjdump; no GPU)"""
import collections, os, re, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(root, "harfang3d_b200", "libhgb200.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
funcs, cur = collections.OrderedDict(), None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1); funcs[cur] = []; continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(.*?);", line)
    if m and cur:
        funcs[cur].append(m.group(1).strip())
def op(ins):
    t = ins.split()
    t = t[1] if t[0].startswith("@") else t[0]
    return t.split(".")[0]
tot = collections.Counter()
per = {}
for f, ins in funcs.items():
    c = collections.Counter(op(i) for i in ins)
    per[f] = c
    tot.update(c)
ffma_where = {f: c["FFMA"] for f, c in per.items() if c["FFMA"]}
print("SASS evidence (cuobjdump -sass harfang3d_b200/libhgb200.so, sm_100a, nvcc 12.9, -fmad=false). Regenerate: python scripts/sass_excerpt.py")
print("FFMA2 (contracted packed multiply-add) in the whole library: %d. Scalar FFMA: %d, in %s — the expansion of IEEE-correct float" % (tot["FFMA2"], tot["FFMA"], ", ".join("%s (%d)" % (re.sub(r"^_ZN3hgb\d+", "", f)[:24], n) for f, n in ffma_where.items())))
print("division / square root (div.rn.f32, sqrt.rn.f32), not contractions; tests/test_capi_symbols.py guards both facts.\n")
want = [("k_update_levelILb1ELb0ELi6", ["UBLKCP", "SYNCS", "PREEXIT", "ACQBULK", "CCTL", "DFMA", "FMUL2", "FADD2", "FSETP.GT", "VIMNMX3", "STL", "LDL"]),
        ("k_update_topILb1ELb0E", ["PREEXIT", "ACQBULK", "DFMA", "FMUL2", "STL", "LDL"]),
        ("k_update_level_tmaILb1ELb0E", ["UBLKCP", "SYNCS", "UTMA", "DFMA"]),
        ("k_sample_tracks", ["LDGSTS", "LDGDEPBAR", "DEPBAR", "PREEXIT", "FMUL2"]),
        ("k_compactENS", ["ACQBULK", "PREEXIT", "ATOMG", "MEMBAR", "VOTE", "POPC"])]
for key, pats in want:
    for f, ins in funcs.items():
        if key not in f:
            continue
        c = per[f]
        print("== %s : %d instructions" % (f, len(ins)))
        print("   opcode histogram: " + " ".join("%s:%d" % kv for kv in c.most_common(28)))
        print("   FFMA/FFMA2 (contracted float mul+add, would break bit-exactness): %d ; FMUL2 (packed IEEE products): %d ; DFMA (explicit fma() of the libm restatements): %d" % (c["FFMA"] + c["FFMA2"], c["FMUL2"], c["DFMA"]))
        shown = collections.Counter()
        for i in ins:
            for p in pats:
                if p in i and shown[p] < 2:
                    shown[p] += 1
                    print("     " + i)
                    break
        print()
