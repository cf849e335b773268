#!/bin/bash
# parity tests, then bench variants given in $VARIANTS (separated by ';'), 8 steps each, no CPU baseline
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
IFS=';' read -ra V <<< "$VARIANTS"
i=0
for v in "${V[@]}"; do
  # "VAR=val|args" sets an environment variable for that variant
  e=""; case "$v" in *"|"*) e="${v%%|*}"; v="${v#*|}";; esac
  timeout 600 env $e python bench.py --no-cpu-baseline $v > gpurun_out/ab_$i.json 2> gpurun_out/ab_$i.err; echo "variant [$e $v] exit $?"
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/ab_$i.json"))
    print("  ms/step %.4f value %.0f e2e %.0f resident %.0f frac %.3f"%(d["ms_per_step"],d["value"],d["e2e"]["value"],d.get("e2e_facade",{}).get("value",0),d["roofline"]["frac"]), d["roofline"]["kernel_ms_per_step"])
except Exception as e:
    print("  failed", e); print(open("gpurun_out/ab_$i.err").read()[-800:])
PY
  i=$((i+1))
done
