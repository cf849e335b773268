#!/usr/bin/env python
"""markdown tables of the committed bench lines (profiles/<prefix>_bench_*.json) — pasted into DESIGN.md / BASELINE.md / profiles/README.md"""
import glob, json, os, re, sys
prefix = sys.argv[1] if len(sys.argv) > 1 else "r02"
root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles")
rows = []
for f in sorted(glob.glob(os.path.join(root, prefix + "_bench_*.json"))):
    name = os.path.basename(f)[len(prefix) + 7:-5]
    lines = [l for l in open(f) if l.startswith("{")]
    if not lines:
        continue
    d = json.loads(lines[-1])
    r = d.get("roofline") or {}
    k = r.get("kernel_ms_per_step") or {}
    rows.append((name, d))
    fac = d.get("e2e_facade") or {}
    print("| %s | %d | %.4f | %.1f | %.1f | %s | %s | %s | %s |" % (
        name, d["n_gpus"], d["ms_per_step"], d["value"] / 1e3, d["e2e"]["value"] / 1e3,
        ("%.1f" % (fac["value"] / 1e3)) if fac.get("value") else "-",
        ("%.3f" % r["frac"]) if r.get("frac") else "-",
        " / ".join("%s %.4f" % (a.replace("k_", ""), b) for a, b in sorted(k.items())),
        (d.get("exchange_check") or {}).get("ok", "-")))
