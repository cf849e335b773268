"""Multi-GPU parity of the peer-memory visibility gather: needs at least 2 GPUs on the box (skipped otherwise — the round-end
`pytest -m gpu` box has one). The check itself is tests/mgpu_peer_check.py, started under torchrun, one process per GPU."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu


def test_peer_gather_equals_single_context():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs (found %d)" % n)
    n = min(n, 4)
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1", "--master-port", "29533",
           os.path.join(here, "mgpu_peer_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "identical to one context" in r.stdout and "identical to the per-rank lists" in r.stdout
