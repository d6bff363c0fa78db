"""Per-kernel timing of the embedding path at BASELINE shapes (CUDA events, L2-exceeding inputs).

    python tools/bench_kernels.py --shape c4 [--rows-cap N] [--iters 20]

Prints one JSON line per (kernel, variant): achieved algorithmic GB/s and fraction of measured HBM peak.
Development tool -- bench.py is the judged entry point.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dlrm_yx_b200 import _lib, ops  # noqa: E402

TERABYTE = [39884406, 39043, 17289, 7420, 20263, 3, 7120, 1543, 63, 38532951, 2953546, 403346, 10,
            2208, 11938, 155, 4, 976, 14, 39979771, 25641295, 39664984, 585935, 12972, 108, 36]
KAGGLE = [1460, 583, 10131227, 2202608, 305, 24, 12517, 633, 3, 93145, 5683, 8351593, 3194, 27,
          14992, 5461306, 10, 5652, 2173, 4, 7046547, 18, 15, 286181, 105, 142572]
SHAPES = {
    "c2": dict(rows=KAGGLE, D=16, B=128, L=1),
    "c3": dict(rows=[1000000] * 8, D=64, B=2048, L=100),
    "c4": dict(rows=TERABYTE, D=128, B=32768, L=1),
    "c4d64": dict(rows=TERABYTE, D=64, B=65536, L=1),          # probe: random 256-byte rows, one per bag
    "c3big": dict(rows=[1000000] * 8, D=64, B=16384, L=100),   # probe: C3 with 8x the bags (no wave-quantisation tail)
    "c3d128": dict(rows=[1000000] * 8, D=128, B=2048, L=100),  # probe: C3 with 512-byte rows
}


def peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p))["hbm_gbs"], "measured"
    return 6650.0, "fallback"


PRE = [None]


def timed(fn, iters, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for s, e in evs:
        if PRE[0] is not None:
            PRE[0]()
        s.record()
        fn()
        e.record()
    torch.cuda.synchronize()
    ms = sorted(s.elapsed_time(e) for s, e in evs)
    return ms[len(ms) // 2], ms[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="c4")
    ap.add_argument("--rows-cap", type=int, default=0)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--idx64", action="store_true")
    ap.add_argument("--flush", action="store_true", help="evict L2 (512 MB memset) before every timed call")
    ap.add_argument("--gstride", type=int, default=0, help="gradient laid out [B, gstride, D] (features 1..T used)")
    ap.add_argument("--only", default="")
    ap.add_argument("--gscale", type=float, default=1.0)
    ap.add_argument("--sorted", action="store_true", help="ids sorted within each bag (the reference's generator)")
    ap.add_argument("--pad", type=int, default=0, help="floats of padding after every table row (accumulator in the row)")
    args = ap.parse_args()
    cfg = SHAPES[args.shape]
    rows = [min(r, args.rows_cap) if args.rows_cap else r for r in cfg["rows"]]
    D, B, L = cfg["D"], cfg["B"], cfg["L"]
    T = len(rows)
    dev = torch.device("cuda:0")
    torch.manual_seed(1)
    t0 = time.time()
    if args.pad:
        full = [torch.zeros((r, D + args.pad), device=dev) for r in rows]
        tw = [f[:, :D] for f in full]
        for w, r in zip(tw, rows):
            w.uniform_(-(1.0 / r) ** 0.5, (1.0 / r) ** 0.5)
        ts = ops.TableSet(tw, acc_in_row=True)
    else:
        tw = [torch.empty((r, D), device=dev).uniform_(-(1.0 / r) ** 0.5, (1.0 / r) ** 0.5) for r in rows]
        tm = [torch.zeros(r, device=dev) for r in rows]
        ts = ops.TableSet(tw, moms=tm)
    idt = torch.int64 if args.idx64 else torch.int32
    # several distinct batches so consecutive iterations touch different rows (inputs exceed L2 anyway)
    nbatch = 4
    batches = []
    for i in range(nbatch):
        idx = torch.cat([torch.randint(0, r, (B * L,), device=dev) for r in rows]).to(idt)
        if args.sorted:
            idx = torch.sort(idx.view(T * B, L), dim=1)[0].reshape(-1).contiguous()
        off = (torch.arange(T * B + 1, device=dev) * L).to(idt)
        batches.append((idx, off))
    N = T * B * L
    torch.cuda.synchronize()
    print("# setup %.1fs, tables %.1f GB, N=%d" % (time.time() - t0, sum(rows) * D * 4 / 1e9, N), flush=True)
    peak, src = peak_gbs()
    ib = 8 if args.idx64 else 4
    fwd_bytes = N * D * 4 + N * ib + T * B * ib + T * B * D * 4
    out = torch.empty((B, T, D), device=dev)
    GS = args.gstride or T
    grad_full = torch.randn((B, GS, D), device=dev) * args.gscale
    grad = grad_full[:, GS - T:, :]
    if args.flush:
        junk = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
        PRE[0] = lambda: junk.zero_()
    it = [0]

    def report(name, ms, mn, nbytes, extra=None):
        d = dict(kernel=name, shape=args.shape, pad=args.pad, ms=round(ms, 4), ms_min=round(mn, 4),
                 gbs=round(nbytes / ms / 1e6, 1), frac=round(nbytes / ms / 1e6 / peak, 3), peak=src)
        if extra:
            d.update(extra)
        print(json.dumps(d), flush=True)

    only = set(args.only.split(",")) if args.only else None
    if not only or "fwd" in only:
        def f():
            idx, off = batches[it[0] % nbatch]
            it[0] += 1
            ops.embbag_forward(ts, idx, off, B, out=out, stride_b=T * D, stride_t=D, validate=False)
        ms, mn = timed(f, args.iters)
        report("embbag_fwd", ms, mn, fwd_bytes)

        def fv():
            idx, off = batches[it[0] % nbatch]
            it[0] += 1
            ops.embbag_forward(ts, idx, off, B, out=out, stride_b=T * D, stride_t=D, validate=True)
        ms, mn = timed(fv, args.iters)
        report("embbag_fwd", ms, mn, fwd_bytes, dict(validate=1))

    def s():
        idx, off = batches[it[0] % nbatch]
        it[0] += 1
        return ops.bwd_sort(ts, idx, off, B)
    if not only or "sort" in only:
        ms, mn = timed(s, args.iters)
        report("bwd_sort", ms, mn, N * 8 * 2, None)

    sorted_batches = [ops.bwd_sort(ts, idx, off, B) for idx, off in batches]
    uniq = [int(ops.bwd_segments(sk)[2].item()) for sk, _ in sorted_batches]
    U = sum(uniq) / len(uniq)
    for opt, name in ((_lib.OPT_SGD, "sgd"), (_lib.OPT_RWSADAGRAD, "rwsadagrad")):
        if only and "bwd" not in only:
            break
        for seg in (0, 32, 64, 128, 256):        # sorted positions per group (test hook 3; 0 = the default choice)
            _lib.set_tunable(3, seg)

            def u():
                sk, sb = sorted_batches[it[0] % nbatch]
                it[0] += 1
                ops.bwd_update(ts, opt, grad, GS * D, D, B, sk, sb, 0.01, 1e-10)
            ms, mn = timed(u, args.iters)
            nbytes = T * B * D * 4 + N * 8 + 2 * U * D * 4 + (2 * U * 4 if opt else 0)
            report("bwd_update_" + name, ms, mn, nbytes, dict(seg=seg, U=int(U)))
    _lib.set_tunable(3, 0)
    if only and "dense" not in only and "ix" not in only:
        return

    if args.shape == "c4" and (not only or "dense" in only):
        for N_ in (128, 256, 512, 1024):
            y = torch.relu(torch.randn(B, N_, device=dev)).to(torch.bfloat16)
            dy = torch.randn(B, N_, device=dev).to(torch.bfloat16)
            db = torch.empty(N_, device=dev)
            ms, mn = timed(lambda: ops.relu_bwd_bias_grad(dy, y, relu=True, db=db), args.iters)
            report("relu_bwd_bias_grad", ms, mn, 3 * B * N_ * 2, dict(N=N_))
            ms, mn = timed(lambda: ops.relu_bwd_bias_grad(dy, y, relu=True), args.iters)
            report("relu_bwd_bias_grad_nodb", ms, mn, 3 * B * N_ * 2, dict(N=N_))
    F = T + 1
    feat = torch.randn((B, F, D), device=dev)
    ms, mn = timed(lambda: ops.interact_dot_fwd(feat), args.iters)
    nz = F * (F - 1) // 2
    report("interact_fwd", ms, mn, B * F * D * 4 + B * (D + nz) * 4)
    if ops.interact_mixed_supported(F, D):
        cols = (D + nz + 7) // 8 * 8
        xb = torch.randn((B, D), device=dev).to(torch.bfloat16)
        doutb = torch.randn((B, cols), device=dev).to(torch.bfloat16)
        featp = torch.empty((B, F, D + 8), device=dev)[:, :, :D]        # pitched input: one TMA bulk copy per sample
        featp.copy_(feat)
        for padded, grid in ((0, 0), (1, 0), (1, 6), (1, 8), (1, 18), (1, 24)):   # CTAs per SM of the persistent grid
            _lib.set_tunable(11, grid)
            fin = featp if padded else feat
            ms, mn = timed(lambda: ops.interact_dot_fwd_mixed(fin, xb, cols), args.iters)
            report("interact_mixed_fwd", ms, mn, B * ((F - 1) * D * 4 + D * 2) + B * cols * 2, dict(bulk_padded=padded, ctas_per_sm=grid))
            ms, mn = timed(lambda: ops.interact_dot_bwd_mixed(fin, xb, doutb), args.iters)
            report("interact_mixed_bwd", ms, mn, B * (2 * (F - 1) * D * 4 + 2 * D * 2 + cols * 2), dict(bulk_padded=padded, ctas_per_sm=grid))
        _lib.set_tunable(10, 0)
        _lib.set_tunable(11, 0)
    dout = torch.randn((B, D + nz), device=dev)
    ms, mn = timed(lambda: ops.interact_dot_bwd(feat, dout), args.iters)
    report("interact_bwd", ms, mn, 2 * B * F * D * 4 + B * (D + nz) * 4)


if __name__ == "__main__":
    main()
