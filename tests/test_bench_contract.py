"""bench.py contract on a machine without a GPU: the reference arm runs (the reference's own CPU implementation, oracle/_ref, or the port)
and prints ONE JSON line with the keys the driver reads; the own arm refuses to run — the path has no CPU fallback."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, timeout=600, cwd=ROOT)


def test_reference_arm_line():
    r = _run("--impl", "reference", "--steps", "2", "--warmup", "1", "--scale", "0.002")
    assert r.returncode == 0, r.stderr[-800:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "scene nodes updated+culled/sec" and d["unit"] == "Mnodes/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 2 and d["vs_baseline"] is None and d["gpu_launches"] == 0
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["config"]["workload"].startswith("c4")
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


@pytest.mark.skipif(torch.cuda.is_available(), reason="needs a machine WITHOUT a CUDA device")
def test_own_arm_refuses_without_cuda():
    r = _run("--steps", "1", "--warmup", "1", "--scale", "0.002", "--no-cpu-baseline")
    assert r.returncode != 0 and "no CPU fallback" in r.stderr and not [l for l in r.stdout.splitlines() if l.startswith("{")]
