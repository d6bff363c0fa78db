"""A handful of tower-GEMM launches for an ncu capture (no timing here):
    ncu --set full --clock-control none --import-source on -k regex:'gemm_pair|wgrad' -o gpurun_out/gemm python tools/gemm_profile.py
Launch order: forward (16->512), (512->256), (1024->1024); dgrad 1024x1024; wgrad of the Terabyte top tower; wgrad of the
bottom tower (each followed by its reduce launch)."""
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


def ops_of(K, N):
    x = (torch.randn(M, K, device=dev) * 0.5).to(bf)
    w = (torch.randn(N, K, device=dev) / K ** 0.5).to(bf)
    b = torch.randn(N, device=dev)
    dz = (torch.randn(M, N, device=dev) * 0.1).to(bf)
    return x, w, b, dz


for K, N in [(16, 512), (512, 256), (1024, 1024)]:
    x, w, b, dz = ops_of(K, N)
    bits = ops.relu_bits_empty(M, N, dev)
    ops.gemm_bf16(x, w, epilogue=ops.EPI_BIAS_RELU, bias=b, relu_bits_out=bits)
gb = ops.relu_bits_empty(M, 1024, dev).fill_(0x55555555)
ops.gemm_bf16(dz, w, b_is_kn=True, epilogue=ops.EPI_RELU_BITS, relu_mask=gb)
for layers in ([(480, 1024), (1024, 1024), (1024, 512), (512, 256)], [(16, 512), (512, 256), (256, 128)]):
    dzs = [(torch.randn(M, N, device=dev) * 0.1).to(bf) for K, N in layers]
    xs = [(torch.randn(M, K, device=dev) * 0.5).to(bf) for K, N in layers]
    dws = [torch.empty(N, K, device=dev) for K, N in layers]
    dbs = [torch.empty(N, device=dev) for K, N in layers]
    ws = ops.gemm_wgrad_workspace([N for K, N in layers], [K for K, N in layers], M, dev)
    ops.gemm_wgrad_bf16(dzs, xs, dws, dbs, ws)
torch.cuda.synchronize()
print("done")
