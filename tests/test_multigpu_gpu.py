"""Needs >= 2 GPUs on the box (skipped otherwise): the model-parallel path -- NCCL all-to-all, peer-memory exchange
(embbag_fwd_scatter / a2a_push_grad over NVLink), replicated small tables, CUDA-graph replay -- against each other bit
for bit and against the fixtures the unmodified reference produced as a 2-process gloo run (tests/golden/dist_w2_*.npz).
The checker is tools/check_exchange.py, launched with torch.distributed.run as the driver launches bench.py.  On a
1-GPU box the same properties are checked by bench.py itself before its timed region whenever it runs with N > 1
(the "parity" key of its JSON line)."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("name,sharders", [("c4_terabyte_small_rws", "naive_chunk,greedy,traffic"),
                                           ("c3_multihot_sgd", "naive")])
def test_exchange_paths_match_each_other_and_the_reference(name, sharders):
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if n < 2:
        pytest.skip("needs 2 GPUs")
    W = 4 if (n >= 4 and name.startswith("c4")) else 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(W), "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "tools", "check_exchange.py"),
           "--case", name, "--sharders", sharders]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("exchange parity OK") == W


def test_peer_allreduce_matches_the_rank_order_sum():
    """ext_dist.PeerAllReduce (own two-shot kernel over NVLink peer memory) on every GPU of the box: bit-equal to the
    rank-order fp32 sum of the operands, eager and as a replayed graph, NCCL's result within rounding (tools/check_allreduce.py)."""
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if n < 2:
        pytest.skip("needs 2 GPUs")
    W = 8 if n >= 8 else (4 if n >= 4 else 2)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(W), "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "tools", "check_allreduce.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("allreduce parity OK") == W
