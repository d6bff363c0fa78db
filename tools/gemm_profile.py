"""A handful of tower-GEMM launches for an ncu capture (no timing here):
    ncu --set full --clock-control none --import-source on -k regex:'gemm_pair|wgrad_streamk' -o gpurun_out/gemm python tools/gemm_profile.py
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dlrm_yx_b200 import ops  # noqa: E402

dev = torch.device("cuda")
bf = torch.bfloat16
M = 32768
torch.manual_seed(0)
for K, N in [(1024, 1024), (16, 512)]:
    x = (torch.randn(M, K, device=dev) * 0.5).to(bf)
    w = (torch.randn(N, K, device=dev) / K ** 0.5).to(bf)
    b = torch.randn(N, device=dev)
    bits = ops.relu_bits_empty(M, N, dev)
    for _ in range(2):
        y = ops.gemm_bf16(x, w, epilogue=ops.EPI_BIAS_RELU, bias=b, relu_bits_out=bits)
    dz = (torch.randn(M, N, device=dev) * 0.1).to(bf)
    gb = ops.relu_bits_empty(M, K, dev).fill_(0x55555555)
    for _ in range(2):
        dx = ops.gemm_bf16(dz, w, b_is_kn=True, epilogue=ops.EPI_RELU_BITS, relu_mask=gb)
    dw = torch.empty(N, K, device=dev)
    db = torch.empty(N, device=dev)
    ws = ops.gemm_wgrad_workspace([N], [K], M, dev)
    for _ in range(2):
        ops.gemm_wgrad_bf16([dz], [x], [dw], [db], ws)
torch.cuda.synchronize()
print("done")
