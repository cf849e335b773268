#!/usr/bin/env python
"""profiles/traffic.json from two `ncu --set full` reports of the bench command (run HERE, after gpurun brought the .ncu-rep files back).

usage: traffic_from_ncu.py <levels.ncu-rep> <others.ncu-rep> <capture label> > profiles/traffic.json

One frame of the level path = a k_update_top launch (the top of the hierarchy) and the k_update_level launches up to the next
k_update_top; its DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum, summed over those launches) are what bench.py quotes as
roofline.traffic beside the algorithmic bytes of the same frame."""
import csv
import json
import re
import subprocess
import sys


def rows(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--print-units", "base"], capture_output=True, text=True, check=True).stdout
    r = list(csv.reader(raw.splitlines()))
    H = r[0]
    col = {k: H.index(k) for k in ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "Grid Size", "Block Size")}
    units = r[1]
    assert units[col["gpu__time_duration.sum"]] in ("ns", "nsecond"), units[col["gpu__time_duration.sum"]]
    assert units[col["dram__bytes_read.sum"]] in ("byte", "B"), units[col["dram__bytes_read.sum"]]
    out = []
    for x in r[2:]:
        f = lambda k: float(x[col[k]].replace(",", ""))
        m = re.search(r"(k_[a-z0-9_]+)", x[col["Kernel Name"]])
        out.append({"kernel": m.group(1) if m else x[col["Kernel Name"]], "time_s": f("gpu__time_duration.sum") * 1e-9,
                    "dram_read": f("dram__bytes_read.sum"), "dram_write": f("dram__bytes_write.sum"), "grid": x[col["Grid Size"]], "block": x[col["Block Size"]]})
    return out


def main():
    lv, others, label = rows(sys.argv[1]), rows(sys.argv[2]), sys.argv[3]
    tops = [i for i, r in enumerate(lv) if r["kernel"] == "k_update_top"]
    assert len(tops) >= 2, "need one whole frame (two k_update_top launches) in the capture"
    frame = lv[tops[0]:tops[1]]
    levels = [r for r in frame if r["kernel"] == "k_update_level"]
    last = max(levels, key=lambda r: r["time_s"])
    doc = {
        "capture": label,
        "source": "ncu --set full --clock-control none of `python bench.py --steps 3 --warmup 3 --no-cpu-baseline` (c4 workload); one frame = 1 k_update_top + %d k_update_level launches" % len(levels),
        # what bench.py quotes as roofline.traffic: every launch of the level path of one frame (k_update_top + k_update_level)
        "k_update_level_bytes_per_launch_sum_per_step": sum(r["dram_read"] + r["dram_write"] for r in frame),
        "k_update_top_bytes_per_step": sum(r["dram_read"] + r["dram_write"] for r in frame if r["kernel"] == "k_update_top"),
        "frame_launches": [{k: r[k] for k in ("kernel", "grid", "block", "time_s", "dram_read", "dram_write")} for r in frame],
        "k_update_level_last_level": {k: last[k] for k in ("dram_read", "dram_write", "time_s")},
        "others": [{"kernel": r["kernel"], "dram_bytes": r["dram_read"] + r["dram_write"], "time_s": r["time_s"]} for r in others],
    }
    json.dump(doc, sys.stdout, indent=1)
    print()


main()
