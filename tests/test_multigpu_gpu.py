"""Needs >= 2 GPUs on the box (skipped otherwise): the peer-memory exchange (embbag_fwd_scatter / a2a_push_grad over
NVLink) against the NCCL all-to-all path, bit for bit, and against the reference's golden vectors.  The checker is
tools/check_exchange.py, launched with torch.distributed.run as the driver launches bench.py."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("name", ["c4_terabyte_small_rws", "c3_multihot_sgd"])
def test_peer_exchange_matches_nccl_and_reference(name):
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tools", "check_exchange.py"), name]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("exchange parity OK") == 2
