"""tcgen05 tower GEMMs against torch on every shape of the Terabyte / benchmark / Kaggle towers: correctness (fp32 matmul of
the same bf16-rounded operands) and time beside the cuBLASLt kernel torch would run (the measured comparator, never a
fallback).  One JSON line per (kind, shape); exit code 1 if any shape is wrong.

    python tools/gemm_check.py [--batch 32768] [--iters 20] [--only fwd,dgrad,wgrad] [--quick]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dlrm_yx_b200 import ops  # noqa: E402

# (K, N) of the fused layers: Terabyte bottom 13(16)-512-256-128, top 479(480)-1024-1024-512-256; benchmark-script
# 512-512-64 / 100(104)-1024-1024-1024; Kaggle 13(16)-512-256-64-16 / 367(368)-512-256
SHAPES = {"terabyte_bot": [(16, 512), (512, 256), (256, 128)],
          "terabyte_top": [(480, 1024), (1024, 1024), (1024, 512), (512, 256)],
          "bench": [(512, 512), (512, 64), (104, 1024), (1024, 1024)],
          "kaggle": [(16, 512), (256, 64), (64, 16), (368, 512)]}


def timed(fn, iters, reps=10):
    """Median ms per call: `reps` back-to-back calls captured in one CUDA graph (the whole training step runs as a graph,
    so host launch cost must not be part of a kernel's time), replayed `iters` times between CUDA events."""
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            for _ in range(reps):
                fn()
    torch.cuda.synchronize()
    g.replay()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for s, e in ev:
        s.record()
        g.replay()
        e.record()
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(b) for a, b in ev)
    return t[len(t) // 2] / reps


def pack_bits(gate):
    """bool [M, N] -> int32 [M, ceil(N / 32)] gate words as the forward epilogue writes them."""
    M, N = gate.shape
    padded = torch.zeros(M, (N + 31) // 32 * 32, dtype=torch.int64, device=gate.device)
    padded[:, :N] = gate
    words = (padded.view(M, -1, 32) << torch.arange(32, device=gate.device, dtype=torch.int64)).sum(-1)
    return torch.where(words >= 2 ** 31, words - 2 ** 32, words).to(torch.int32)


def rel_err(got, ref):
    return float((got.float() - ref).norm() / ref.norm().clamp_min(1e-30))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32768)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--only", default="fwd,dgrad,wgrad")
    ap.add_argument("--quick", action="store_true", help="one small and one large shape only (first bring-up)")
    args = ap.parse_args()
    dev = torch.device("cuda")
    torch.manual_seed(0)
    M = args.batch
    only = set(args.only.split(","))
    bad = 0
    groups = dict(SHAPES)
    if args.quick:
        groups = {"quick": [(64, 128), (512, 256)]}
    bf = torch.bfloat16
    for gname, shapes in groups.items():
        xs, dzs, ws = [], [], []
        for K, N in shapes:
            x = (torch.randn(M, K, device=dev) * 0.5).to(bf)
            w = (torch.randn(N, K, device=dev) * (1.0 / K ** 0.5)).to(bf)
            b = torch.randn(N, device=dev)
            dz = (torch.randn(M, N, device=dev) * 0.1).to(bf)
            yprev = torch.randn(M, K, device=dev).to(bf)          # the producing layer's output (ReLU mask source)
            xs.append(x)
            dzs.append(dz)
            ws.append(w)
            flops = 2.0 * M * N * K
            if "fwd" in only:
                ref = torch.relu(x.float() @ w.float().t() + b)
                got = ops.gemm_bf16(x, w, epilogue=ops.EPI_BIAS_RELU, bias=b)
                err = rel_err(got, ref)
                y_out, bb = torch.empty_like(got), b.to(bf)
                gbits = ops.relu_bits_empty(M, N, dev)
                ms = timed(lambda: ops.gemm_bf16(x, w, epilogue=ops.EPI_BIAS_RELU, bias=b, out=y_out, relu_bits_out=gbits), args.iters)
                ms_lib = timed(lambda: torch._addmm_activation(bb, x, w.t(), out=y_out), args.iters)
                ok = err < 6e-3
                bad += not ok
                print(json.dumps(dict(kind="fwd", group=gname, M=M, K=K, N=N, rel_err=err, ok=ok, ms=round(ms, 4),
                                      tflops=round(flops / ms / 1e9, 1), cublas_ms=round(ms_lib, 4),
                                      cublas_tflops=round(flops / ms_lib / 1e9, 1))), flush=True)
            if "dgrad" in only:
                ref = (dz.float() @ w.float()) * (yprev.float() > 0)
                bits = pack_bits(yprev.float() > 0)
                got = ops.gemm_bf16(dz, w, b_is_kn=True, epilogue=ops.EPI_RELU_BITS, relu_mask=bits)
                err = rel_err(got, ref)
                got2 = ops.gemm_bf16(dz, w, b_is_kn=True)
                err2 = rel_err(got2, dz.float() @ w.float())
                dx_out = torch.empty_like(got)
                ms = timed(lambda: ops.gemm_bf16(dz, w, b_is_kn=True, epilogue=ops.EPI_RELU_BITS, relu_mask=bits, out=dx_out), args.iters)
                ms_nomask = timed(lambda: ops.gemm_bf16(dz, w, b_is_kn=True, out=dx_out), args.iters)
                ms_lib = timed(lambda: torch.mm(dz, w, out=dx_out), args.iters)
                ok = err < 6e-3 and err2 < 6e-3
                bad += not ok
                print(json.dumps(dict(kind="dgrad", group=gname, M=M, K=N, N=K, rel_err=err, rel_err_nomask=err2, ok=ok,
                                      ms=round(ms, 4), ms_nomask=round(ms_nomask, 4), tflops=round(flops / ms / 1e9, 1),
                                      cublas_ms=round(ms_lib, 4), cublas_tflops=round(flops / ms_lib / 1e9, 1))), flush=True)
        if "wgrad" in only:
            dws = [torch.full((w.shape[0], w.shape[1]), float("nan"), device=dev) for w in ws]
            dbs = [torch.full((w.shape[0],), float("nan"), device=dev) for w in ws]
            wsb = ops.gemm_wgrad_workspace([w.shape[0] for w in ws], [w.shape[1] for w in ws], M, dev)
            ops.gemm_wgrad_bf16(dzs, xs, dws, dbs, wsb)
            ops.gemm_wgrad_bf16(dzs, xs, dws, dbs, wsb)           # second call: the counters were left zero
            errs, errb = [], []
            for dz, x, dw, db in zip(dzs, xs, dws, dbs):
                errs.append(rel_err(dw, dz.float().t() @ x.float()))
                errb.append(rel_err(db, dz.float().sum(dim=0)))
            again = [d.clone() for d in dws]
            ops.gemm_wgrad_bf16(dzs, xs, dws, dbs, wsb)
            det = all(torch.equal(a, b) for a, b in zip(again, dws))
            ms = timed(lambda: ops.gemm_wgrad_bf16(dzs, xs, dws, dbs, wsb), args.iters)
            ms_lib = timed(lambda: [torch.mm(dz.t(), x, out_dtype=torch.float32) for dz, x in zip(dzs, xs)], args.iters)
            per_layer = []
            for dz, x, dw, db in zip(dzs, xs, dws, dbs):
                ws1 = ops.gemm_wgrad_workspace([dz.shape[1]], [x.shape[1]], M, dev)
                t1 = timed(lambda: ops.gemm_wgrad_bf16([dz], [x], [dw], [db], ws1), args.iters)
                t0 = timed(lambda: ops.gemm_wgrad_bf16([dz], [x], [dw], None, ws1), args.iters)
                tl = timed(lambda: torch.mm(dz.t(), x, out_dtype=torch.float32), args.iters)
                per_layer.append(dict(N=dz.shape[1], K=x.shape[1], ms=round(t1, 4), ms_nobias=round(t0, 4), cublas_ms=round(tl, 4)))
            flops = sum(2.0 * M * w.shape[0] * w.shape[1] for w in ws)
            ok = max(errs) < 3e-3 and max(errb) < 3e-3 and det
            bad += not ok
            print(json.dumps(dict(kind="wgrad", group=gname, M=M, layers=shapes, rel_err=errs, rel_err_bias=errb,
                                  deterministic=det, ok=ok, ms=round(ms, 4), tflops=round(flops / ms / 1e9, 1),
                                  cublas_ms=round(ms_lib, 4), cublas_tflops=round(flops / ms_lib / 1e9, 1),
                                  per_layer=per_layer)), flush=True)
    print("gemm_check:", "FAILED (%d)" % bad if bad else "all shapes OK", flush=True)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
