"""Two GPUs (skipped on a one-GPU box): the peer-memory exchange delivers exactly what the NCCL all-to-all delivers, buffer reuse
included, and a page handed out by one exchange survives the next one (tools/exchange_check.py under torchrun)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_peer_exchange_equals_nccl_and_held_pages_survive():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29533", os.path.join(ROOT, "tools", "exchange_check.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "exchange_check: OK" in r.stdout, (r.stdout[-2000:], r.stderr[-2000:])
