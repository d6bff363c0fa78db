"""bench.py -- DLRM train samples/s on the Terabyte/MLPerf shape (BASELINE.json configs[3]/[4]) through the
B200 sparse hot path, plus the roofline of the embedding-bag kernels and the CPU baseline.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference            # the reference's CPU path (oracle port) on the host cores

A "step" = one full training iteration on one synthetic mini-batch: bottom MLP, table-batched embedding-bag
lookup, (all-to-all,) dot interaction, top MLP, BCE loss, backward, fused row-wise-Adagrad update of the touched
table rows, Adagrad step of the MLPs.  One JSON line on stdout (rank 0).
"""
import argparse
import json
import os
import sys
import threading
import time

import warnings

import numpy as np
import torch

warnings.filterwarnings("ignore", message="Sparse invariant checks")

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="terabyte", choices=["terabyte", "bench", "kaggle"])
    ap.add_argument("--batch", type=int, default=0, help="per-GPU mini-batch (default: the config's)")
    ap.add_argument("--global-batch", type=int, default=0, help="fix the GLOBAL batch (strong scaling)")
    ap.add_argument("--rows-cap", type=int, default=0, help="cap rows per table (debug only; marks the run)")
    ap.add_argument("--allreduce", default="peer", choices=["peer", "nccl"],
                    help="data-parallel gradient sums (towers, replicated tables) at N > 1: the repo's kernel over NVLink peer "
                         "memory (csrc/peer_allreduce.cu) or NCCL (comparator)")
    ap.add_argument("--watchdog-s", type=float, default=900.0,
                    help="N > 1: a rank that has not finished after this many seconds says so on stderr and exits 124 instead of "
                         "spinning in a device barrier until an outer limit kills the box (0 = off)")
    ap.add_argument("--mgpu-overlap", type=int, default=0,
                    help="N > 1 schedule, bit mask (dlrm_net.MGPU_OVERLAP): 0 = the one measured on 2 / 4 / 8 GPUs; 1 separate "
                         "high-priority all-reduce streams, 2 short-lived push CTAs, 4 replicated sort on the side stream, 8 "
                         "replicated lookup on a third stream; 15 = all (measured on 2 GPUs only)")
    ap.add_argument("--grad-wire", default="fp32", choices=["fp32", "bf16"],
                    help="N > 1, bf16 towers: pooled-row gradients of the sharded tables cross NVLink as fp32 or rounded to bf16")
    ap.add_argument("--tower-gemm", default="tc", choices=["tc", "cublas"],
                    help="fused-bf16 towers: tc = the repo's tcgen05/TMEM/TMA GEMM kernels; cublas = cuBLASLt through torch, "
                         "the measured comparator (never a fallback)")
    ap.add_argument("--defer-wgrad", type=int, default=1,
                    help="1: the top tower's weight-gradient launch moves to the end of the backward pass, beside the sparse update; "
                         "2: right behind the interaction backward instead (measured alternative)")
    ap.add_argument("--trace", default="", help="after the timed regions: CUPTI kernel timeline (torch.profiler) of 3 more steps, "
                    "written as text (start / duration / stream / kernel) for the LAST of them; every rank writes <trace>.rank<r>")
    ap.add_argument("--table-dtype", default="fp32", choices=["fp32", "bf16"],
                    help="row storage of the embedding tables (bf16: widened on load, fp32 sums / updates, rounded on store); "
                         "the headline stays fp32")
    ap.add_argument("--mlp", default="fused-bf16", choices=["fused-bf16", "tf32", "fp32", "bf16"],
                    help="fused-bf16: FusedMLP (bf16 tensor-core GEMMs, fused epilogues/optimizer); others: torch towers")
    ap.add_argument("--sharder", default="traffic")
    ap.add_argument("--no-graph", action="store_true",
                    help="launch the step kernel by kernel instead of replaying the captured CUDA graph")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="N>1: 'peer' = exchange fused into the kernels over NVLink peer memory; 'nccl' = all_to_all_single")
    ap.add_argument("--replicate-below", type=int, default=100000,
                    help="N>1: tables with fewer rows are kept on every rank (looked up for the local batch, dense gradient "
                         "all-reduce) instead of being sharded and exchanged; 0 = shard everything as the reference does")
    ap.add_argument("--index-dist", default="uniform", choices=["uniform", "zipf"],
                    help="uniform: the reference's generator (dlrm_data_pytorch.py:1195-1208); zipf: P(row k) ~ k^-1.05")
    ap.add_argument("--no-parity-check", action="store_true",
                    help="N>1: skip the pre-timing correctness checks (peer vs NCCL step, received rows vs direct gather)")
    ap.add_argument("--no-gpu-reference", action="store_true", help="skip the ATen-on-the-same-GPU reference leg (N=1)")
    ap.add_argument("--optimizer", default="rwsadagrad", choices=["rwsadagrad", "sgd"])
    ap.add_argument("--nbatches", type=int, default=8, help="distinct synthetic batches cycled through")
    ap.add_argument("--cpu-batch", type=int, default=0, help="CPU arm batch (default: the config's own batch)")
    ap.add_argument("--cpu-rows-cap", type=int, default=-1,
                    help="CPU arm: cap on rows per table; -1 = none when the host has >= 160 GB available, else 10 M "
                         "(BASELINE.md section 3), else 1 M below 48 GB")
    ap.add_argument("--cpu-steps", type=int, default=20)
    ap.add_argument("--cpu-warmup", type=int, default=5)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-c3-probe", action="store_true", help="skip the C3-shape embedding-bag kernel lines")
    ap.add_argument("--tunable", action="append", default=[], metavar="K=V",
                    help="experiments: dlrmx_set_tunable(K, V) before anything runs (include/dlrmx.h); recorded in config")
    return ap.parse_args()


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy read+write)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""
    BITS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
            0x80: "hw_power_brake_slowdown"}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self.ok = index, [], set(), None, False
        self._stop_ev = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception as e:  # NVML missing: clocks are reported as unknown, never invented
            self.err = repr(e)

    def sample(self):
        if not self.ok:
            return
        try:
            self.samples.append(int(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
            r = int(self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
            for bit, name in self.BITS.items():
                if r & bit:
                    self.reasons.add(name)
        except Exception:
            pass

    def run(self):
        while not self._stop_ev.is_set():
            self.sample()
            self._stop_ev.wait(0.02)

    def stop(self):
        self._stop_ev.set()
        self.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "no NVML samples"}
        return {"sm_mhz": int(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------------
# synthetic workload
# ------------------------------------------------------------------------------------------------------
def config_of(args):
    from dlrm_yx_b200 import data
    cfg = dict(data.CONFIGS[args.config])
    cfg["rows"] = [min(r, args.rows_cap) if args.rows_cap else r for r in cfg["rows"]]
    if args.batch:
        cfg["B"] = args.batch
    return cfg


def draw_ids(rng, B, L, E, dist):
    """Row ids [B, L] of one table: the reference's uniform generator round(U * (E - 1)) (dlrm_data_pytorch.py:1195-1208),
    or a bounded Zipf(1.05) over the rows (inverse CDF of the continuous power law; row 0 is the hottest)."""
    u = rng.random((B, L))
    if dist == "uniform" or E < 2:
        return np.round(u * (E - 1)).astype(np.int32)
    a = 1.0 - 1.05
    k = np.floor(((float(E) ** a - 1.0) * u + 1.0) ** (1.0 / a))
    return (np.clip(k, 1, E) - 1).astype(np.int32)


def make_batches(cfg, B_global, local_tables, rep_tables, lb_slice, n, seed, dist="uniform"):
    """n pinned host batches for one rank: dense X / targets for the LOCAL batch slice, packed int32 lookups of the LOCAL
    sharded tables for the FULL global batch (dlrm_wrap's split, dlrm_s_pytorch.py:145-148) and -- with replicated
    tables -- packed lookups of those tables for the local slice: [X, idx, off, tgt(, idx_rep, off_rep)]."""
    rng = np.random.default_rng(seed)          # same seed on every rank -> every rank draws the same global batch
    T_all, L = len(cfg["rows"]), cfg["L"]
    lb = lb_slice.stop - lb_slice.start
    out = []
    for _ in range(n):
        X = rng.random((B_global, cfg["bot"][0]), dtype=np.float32)
        tgt = np.round(rng.random((B_global, 1))).astype(np.float32)
        parts = {}
        for t in range(T_all):
            ids = draw_ids(rng, B_global, L, cfg["rows"][t], dist)
            if L > 1:
                ids.sort(axis=1)
            if t in local_tables:
                parts[t] = ids.reshape(-1)
            elif t in rep_tables:
                parts[t] = np.ascontiguousarray(ids[lb_slice]).reshape(-1)
        idx = np.concatenate([parts[t] for t in local_tables]) if local_tables else np.zeros(0, np.int32)
        off = (np.arange(len(local_tables) * B_global + 1, dtype=np.int64) * L).astype(np.int32)
        host = [torch.from_numpy(np.ascontiguousarray(X[lb_slice])), torch.from_numpy(idx), torch.from_numpy(off),
                torch.from_numpy(np.ascontiguousarray(tgt[lb_slice]))]
        if rep_tables:
            host.append(torch.from_numpy(np.concatenate([parts[t] for t in rep_tables])))
            host.append(torch.from_numpy((np.arange(len(rep_tables) * lb + 1, dtype=np.int64) * L).astype(np.int32)))
        out.append([t.pin_memory() for t in host])
    return out


def algorithmic_bytes(N, TB, D, U, ib=4, tb=4):
    """forward = rows + indices + offsets + pooled rows; backward+row-wise Adagrad = pooled-row
    gradients + (key, bag) + read/write of the U distinct rows and their accumulators.  tb: bytes per table element."""
    fwd = N * D * tb + N * ib + TB * ib + TB * D * 4
    bwd = TB * D * 4 + N * 8 + 2 * U * D * tb + 2 * U * 4
    return fwd, bwd


# ------------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------------
def build_model(args, cfg, W, exchange=None):
    from dlrm_yx_b200 import data, dlrm_net, embedding, rwsadagrad
    exchange = args.exchange if exchange is None else exchange
    D, T = cfg["D"], len(cfg["rows"])
    ln_bot = np.array(cfg["bot"])
    ln_top = np.array([data.num_interactions(T, D)] + list(cfg["top"][1:]))
    kind = (embedding.Optimizer.EXACT_ROWWISE_ADAGRAD if args.optimizer == "rwsadagrad" else embedding.Optimizer.SGD)
    lr = 0.01
    np.random.seed(123)
    torch.manual_seed(123)
    m = dlrm_net.DLRM_Net(D, np.array(cfg["rows"]), ln_bot, ln_top, arch_interaction_op="dot", sigmoid_bot=-1,
                          sigmoid_top=ln_top.size - 2, batched_emb=True, fuse_sparse_arch=True, sharder=args.sharder,
                          emb_optimizer=kind, emb_learning_rate=lr, emb_init="device", emb_weights_precision=args.table_dtype,
                          replicate_rows_below=(args.replicate_below if W > 1 else 0)).cuda()
    m.sync_replicas()
    if W > 1 and exchange == "peer":
        lb = cfg["B"] if not args.global_batch else args.global_batch // W
        m.enable_peer_exchange(lb, mixed=(args.mlp == "fused-bf16"), grad_wire_bf16=(args.grad_wire == "bf16"))
    if args.mlp == "fused-bf16":
        m.fuse_mlps(lr, "adagrad" if args.optimizer == "rwsadagrad" else "sgd", gemm=args.tower_gemm, allreduce=args.allreduce,
                    grad_wire_bf16=(args.grad_wire == "bf16"))
        m.top_l.defer_wgrad = bool(m.top_l.defer_wgrad and args.defer_wgrad)
        dlrm_net.FLUSH_EARLY = args.defer_wgrad == 2
        if W == 1 and args.defer_wgrad == 1:
            m.bot_l.defer_wgrad = True
        return m, None
    dense = list(m.bot_l.parameters()) + list(m.top_l.parameters())
    if W > 1:
        from dlrm_yx_b200 import extend_distributed as ext
        m.bot_l = ext.DDP(m.bot_l, device_ids=[torch.cuda.current_device()])
        m.top_l = ext.DDP(m.top_l, device_ids=[torch.cuda.current_device()])
    opt = (rwsadagrad.RWSAdagrad(dense, lr=lr) if args.optimizer == "rwsadagrad" else torch.optim.SGD(dense, lr=lr))
    return m, opt


def make_step(args, m, opt):
    autocast = torch.autocast("cuda", dtype=torch.bfloat16, enabled=(args.mlp == "bf16"))

    def step(X, idx, off, tgt, idx_rep=None, off_rep=None):
        o, i = ((off, off_rep), (idx, idx_rep)) if idx_rep is not None else (off, idx)
        with autocast:
            p = m(X, o, i)
        loss = m.loss_fn(p.float(), tgt)
        if opt is None:
            loss.backward()
            m.step_mlps()
        else:
            opt.zero_grad(set_to_none=True)
            loss.backward()
            opt.step()
        return loss
    return step


def _sparse_state(m):
    ts = [("tables", m._batched_module().embedding_weights.data), ("accumulators", m._batched_module().row_momentum)]
    if m.emb_rep is not None:
        ts += [("replicated tables", m.emb_rep.embedding_weights.data), ("replicated accumulators", m.emb_rep.row_momentum)]
    return [(n, t) for n, t in ts if t is not None]


def parity_step_vs_nccl(args, cfg, W, m, step, batch):
    """N > 1, before any timing: ONE training step of the benchmark model (peer-memory exchange) and ONE of a second model
    built from the same seeds with the NCCL all-to-all, on the same batch; loss, every local table row, every
    accumulator and the towers must agree bit for bit (only the transport differs).  Raises on mismatch."""
    m2, opt2 = build_model(args, cfg, W, exchange="nccl")
    step2 = make_step(args, m2, opt2)
    for (n, a), (_, b) in zip(_sparse_state(m), _sparse_state(m2)):
        if not torch.equal(a, b):
            raise SystemExit("bench parity: the two seeded builds differ before the first step (%s)" % n)
    l1 = step(*batch)
    l2 = step2(*batch)
    torch.cuda.synchronize()
    if l1.item() != l2.item():
        raise SystemExit("bench parity: peer exchange loss %r != NCCL exchange loss %r" % (l1.item(), l2.item()))
    checked = 0
    for (n, a), (_, b) in zip(_sparse_state(m), _sparse_state(m2)):
        if not torch.equal(a, b):
            raise SystemExit("bench parity: %s differ between the peer and the NCCL exchange after one step" % n)
        checked += a.numel()
    for ta, tb in ((m.bot_l, m2.bot_l), (m.top_l, m2.top_l)):
        if hasattr(ta, "master") and not torch.equal(ta.master.data, tb.master.data):
            raise SystemExit("bench parity: tower parameters differ between the peer and the NCCL exchange")
    del m2, opt2, step2
    torch.cuda.empty_cache()
    return {"peer_vs_nccl": "bit-exact", "values_compared": int(checked)}


def parity_rows_vs_gather(cfg, W, rank, m, run, batch, nrows=1024):
    """N > 1 with one index per lookup (pooled row == table row): the interaction input that rank 0 receives through the
    benchmarked launch (`run`: the captured graph when there is one) against rows gathered directly from the owning
    ranks' tables BEFORE the step touches them.  Bit-exact or raises.  Returns the number of rows compared."""
    import torch.distributed as dist
    if cfg["L"] != 1 or getattr(m, "_peer", None) is None:
        return 0
    peer = m._peer
    lb, D = peer.lb, peer.D
    n = min(nrows, lb)
    Bg = lb * W
    emb = m._batched_module()
    idx = batch[1]
    expect = {}                                # feature slot -> [n, D] rows of rank 0's first n samples
    slot = 1
    for r in range(W):
        for j in range(peer.tpr[r]):
            rows = torch.empty((n, D), device=idx.device)
            if r == rank:
                w = emb.split_weights()[j]
                rows.copy_(w[idx[j * Bg:j * Bg + n].long()])       # rank 0 owns global samples [0, lb)
            dist.broadcast(rows, r)
            expect[slot] = rows
            slot += 1
    if m.emb_rep is not None and rank == 0:
        ridx = batch[4]
        for j, w in enumerate(m.emb_rep.split_weights()):
            expect[slot + j] = w[ridx[j * lb:j * lb + n].long()].clone()
    run(*batch)
    torch.cuda.synchronize()
    bad = 0
    if rank == 0:
        for sl, rows in expect.items():
            if not torch.equal(peer.feat[:n, sl, :], rows):
                bad += 1
    flag = torch.tensor([bad], device=idx.device)
    dist.all_reduce(flag)
    if int(flag.item()) != 0:
        raise SystemExit("bench parity: %d feature slots of rank 0's interaction input differ from a direct gather"
                         % int(flag.item()))
    return n * len(expect) if rank == 0 else 0


def _ensure_built():
    """The harness (not the product code) compiles libdlrmx.so in-tree when it is missing or stale; with several ranks
    only local rank 0 builds, the others wait for the stamp."""
    import time
    sys.path.insert(0, ROOT)
    from dlrm_yx_b200 import build
    if not build.needs_build():
        return
    if int(os.environ.get("LOCAL_RANK", "0")) == 0:
        build.build()
    else:
        for _ in range(600):
            if not build.needs_build():
                return
            time.sleep(1.0)


def run_ours(args):
    _ensure_built()
    from dlrm_yx_b200 import _lib, ops
    from dlrm_yx_b200 import extend_distributed as ext
    W = int(os.environ.get("WORLD_SIZE", "1"))
    if W != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch N>1 with torch.distributed.run" % (args.gpus, W))
    ext.init_distributed(use_gpu=True, backend="nccl")
    rank = max(ext.my_rank, 0)
    dev = torch.device("cuda", torch.cuda.current_device())
    if W > 1 and args.watchdog_s > 0:
        # cross-rank device barriers spin on the GPU: if one rank never arrives the others would wait for ever
        def _expired():
            sys.stderr.write("bench.py: rank %d not finished after %.0f s (a cross-rank wait that never completed?) -- giving up\n"
                             % (rank, args.watchdog_s))
            sys.stderr.flush()
            os._exit(124)
        wd = threading.Timer(args.watchdog_s, _expired)
        wd.daemon = True
        wd.start()
    if args.mlp == "tf32":
        torch.backends.cuda.matmul.allow_tf32 = True
        torch.backends.cudnn.allow_tf32 = True
    else:
        torch.backends.cuda.matmul.allow_tf32 = False
    cfg = config_of(args)
    lb = cfg["B"] if not args.global_batch else args.global_batch // W
    B_global = lb * W
    T = len(cfg["rows"])
    gpu_ref = None
    if W == 1 and not args.no_gpu_reference:
        try:
            gpu_ref = gpu_reference(args, cfg)      # before our model exists: both hold the full tables in HBM
        except Exception as ex:                     # noqa: BLE001  (the comparator must never cost the headline line)
            gpu_ref = {"error": repr(ex)}
        torch.cuda.empty_cache()
    if W > 1 and args.exchange == "peer":
        # symmetric (peer-mapped) memory needs CUDA VMM + P2P on the box; every rank must take the same branch, so the
        # outcome is agreed on with an all-reduce and the NCCL exchange is the common fallback (same results, slower)
        ok = 1
        try:
            m, opt = build_model(args, cfg, W)
        except Exception as ex:                                  # noqa: BLE001
            sys.stderr.write("bench: peer exchange unavailable on this rank (%r)\n" % (ex,))
            ok, m = 0, None
        flag = torch.tensor([ok], device="cuda")
        torch.distributed.all_reduce(flag, op=torch.distributed.ReduceOp.MIN)
        if int(flag.item()) == 0:
            del m
            torch.cuda.empty_cache()
            args.exchange = "nccl"
            sys.stderr.write("bench: falling back to the NCCL all-to-all exchange\n")
            m, opt = build_model(args, cfg, W)
    else:
        m, opt = build_model(args, cfg, W)
    local_tables = list(range(T)) if W == 1 else list(m.local_emb_indices)
    rep_tables = [] if W == 1 else list(m.replicated_emb_indices)
    lb_slice = slice(rank * lb, (rank + 1) * lb)
    host = make_batches(cfg, B_global, local_tables, rep_tables, lb_slice, args.nbatches, seed=727, dist=args.index_dist)
    devb = [[t.to(dev) for t in b] for b in host]
    step = make_step(args, m, opt)
    parity = None
    if W > 1 and args.exchange == "peer" and not args.no_parity_check:
        parity = parity_step_vs_nccl(args, cfg, W, m, step, devb[0])

    def sync():
        if W > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    def timed(fn, K):
        sync()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        s.record()
        for i in range(K):
            fn(i)
        e.record()
        sync()
        wall = (time.perf_counter() - t0) * 1e3
        ms = s.elapsed_time(e)
        if W > 1:
            t = torch.tensor([ms, wall], device=dev)
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            ms, wall = float(t[0]), float(t[1])
        return ms, wall

    nb = args.nbatches
    for i in range(max(args.warmup, 3)):
        step(*devb[i % nb])
    sync()
    use_graph = not args.no_graph and opt is None
    run = step
    if use_graph:
        # the whole step (forward, backward, sparse update, tower optimizer) as ONE CUDA graph; a batch is copied
        # into the graph's input buffers (device-to-device in region 1, pinned host-to-device in region 2)
        from dlrm_yx_b200.graph import GraphedTrainStep
        l0 = _lib.launch_count()
        run = GraphedTrainStep(step, devb[0], warmup=2)
        launches_per_step = (_lib.launch_count() - l0) // 3          # 2 warm-up steps + the captured one
        for i in range(3):
            run(*devb[i % nb])
        sync()
    if parity is not None:
        parity["rows_checked"] = parity_rows_vs_gather(cfg, W, rank, m, run, devb[1 % nb])
        parity["rows_check"] = ("rank 0's interaction input after the benchmarked launch (%s) == rows gathered directly from "
                                "the owners' tables, bit for bit" % ("CUDA-graph replay" if use_graph else "eager step"))
        ops.check_index_errors()
        sync()

    # ---- timed region 1: inputs resident in HBM --------------------------------------------------
    sampler = ClockSampler(torch.cuda.current_device())
    sampler.start()
    l0 = _lib.launch_count()
    ms_dev, _ = timed(lambda i: run(*devb[i % nb]), args.steps)
    launches = launches_per_step * args.steps if use_graph else _lib.launch_count() - l0

    # ---- timed region 2: end to end, host buffers in, loss out, every step --------------------------
    if use_graph:
        run.prefetch(*host[0])

    def e2e_step(i):
        hb = host[i % nb]
        if use_graph:
            # this step's inputs were put on the wire (pinned H2D, copy stream) while the previous step ran -- the usual
            # one-batch-ahead loader; now: move them in, replay, start the next batch's H2D, read the loss back
            loss = run.run_prefetched()
            run.prefetch(*host[(i + 1) % nb])
            return float(loss.item())
        return float(step(*[t.to(dev, non_blocking=True) for t in hb]).item())      # D2H read of the step's result
    for i in range(2):
        e2e_step(i)
    _, wall_e2e = timed(e2e_step, args.steps)

    if args.trace:
        write_trace(args.trace + (".rank%d" % rank if W > 1 else ""), lambda i: run(*devb[i % nb]), sync)

    # ---- per-kernel durations: CUDA events around every libdlrmx launch.  Graph mode: a second capture of the same
    #      step on ONE stream with external event-record nodes around each kernel; it is replayed K times and the
    #      events are read after each replay, so the kernels are timed inside a dense replayed step, each alone on
    #      the GPU.  Eager mode: events around each eager launch.
    from dlrm_yx_b200 import dlrm_net
    dlrm_net.OVERLAP_STREAMS = False
    prof, ms_prof, prof_mode = None, 0.0, "eager"
    if use_graph:
        try:
            ops.PROFILE = {}
            gprof = GraphedTrainStep(step, devb[0], warmup=1)
            events = ops.PROFILE
            ops.PROFILE = None
            for name in events:                      # keep only the events recorded by the capture itself
                n = len(events[name]) // 2
                events[name] = events[name][n:]
            acc = {k: [0.0] * len(v) for k, v in events.items()}
            t_rep = 0.0
            for i in range(args.steps):
                gprof.load(*devb[i % nb])
                s_, e_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s_.record()
                gprof.graph.replay()
                e_.record()
                torch.cuda.synchronize()
                t_rep += s_.elapsed_time(e_)
                for k, v in events.items():
                    for j, (a, b) in enumerate(v):
                        acc[k][j] += a.elapsed_time(b)
            prof = {k: (len(v) * args.steps, sum(v) / (len(v) * args.steps)) for k, v in acc.items()}
            ms_prof, prof_mode = t_rep, "graph"
            del gprof
        except Exception as ex:                      # external events unsupported: fall back to the eager pass
            import traceback
            sys.stderr.write("bench: graph-event profiling failed (%r), using the eager pass\n%s\n" % (ex, traceback.format_exc()))
            ops.PROFILE = None
            prof = None
    if prof is None:
        ops.PROFILE = {}
        ms_prof, _ = timed(lambda i: step(*devb[i % nb]), args.steps)
        prof = ops.profile_summary()
        ops.PROFILE = None
    dlrm_net.OVERLAP_STREAMS = True
    ms_eager = ms_prof
    clocks = sampler.stop()
    h2d = sum(t.numel() * t.element_size() for t in host[0])
    ops.check_index_errors()
    return dict(W=W, rank=rank, cfg=cfg, lb=lb, B_global=B_global, local_tables=local_tables, rep_tables=rep_tables,
                parity=parity, gpu_ref=gpu_ref, ms_dev=ms_dev,
                wall_e2e=wall_e2e, launches=launches, prof=prof, ms_eager=ms_eager, use_graph=use_graph, prof_mode=prof_mode, clocks=clocks, h2d=h2d, devb=devb, model=m)


def write_trace(path, step_fn, sync):
    """Kernel timeline of one step as CUPTI sees it (never a timed number: the profiler is attached): per kernel the start
    relative to the step's first kernel, the duration, the stream and the name -- what overlaps what."""
    import tempfile
    from torch.profiler import ProfilerActivity, profile
    sync()
    marks = []
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for i in range(3):
            step_fn(i)
            torch.cuda.synchronize()
            marks.append(time.perf_counter())
    tmp = tempfile.mktemp(suffix=".json")
    prof.export_chrome_trace(tmp)
    ev = json.load(open(tmp))["traceEvents"]
    os.remove(tmp)
    ks = sorted((e for e in ev if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "ts" in e), key=lambda e: e["ts"])
    if not ks:
        return
    # the last step = the kernels after the largest gap among the final third of the list is fragile; use three equal groups
    n = len(ks) // 3
    last = ks[len(ks) - n:]
    t0 = last[0]["ts"]
    with open(path, "w") as f:
        f.write("# one step, %d GPU activities; columns: start_us dur_us stream name\n" % len(last))
        for e in last:
            f.write("%9.1f %8.1f %4s %s\n" % (e["ts"] - t0, e.get("dur", 0.0), e.get("args", {}).get("stream", "?"), e["name"][:110]))
        f.write("# step span %.1f us\n" % (last[-1]["ts"] + last[-1].get("dur", 0.0) - t0))


# ------------------------------------------------------------------------------------------------------
# CPU baseline / reference arm: the reference's CPU path (oracle port over the same ATen CPU operators)
# ------------------------------------------------------------------------------------------------------
def host_info():
    model, avail = "unknown", None
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                model = line.split(":", 1)[1].strip()
                break
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable"):
                avail = int(line.split()[1]) / 1048576.0
    except OSError:
        pass
    return model, avail


def cpu_reference(args, steps, warmup):
    """Times oracle/oracle_torch.py (nn.EmbeddingBag sparse + bmm/tril interaction + RWSAdagrad / SGD, all host threads) on
    the benchmark's OWN configuration -- same tables, same batch (32768 at the Terabyte shape), same optimizer -- with the
    protocol of BASELINE.md section 3: 5 warm-up steps, median of >= 20, per-phase wall times.  Only when the host
    cannot hold the 96 GB of tables are rows capped (10 M per table below 160 GB available, 1 M below 48 GB); the cap is
    written into the line's config."""
    from dlrm_yx_b200 import data
    from oracle import oracle_torch as OT
    cfg = config_of(args)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cpu_model, avail = host_info()
    need = sum(cfg["rows"]) * cfg["D"] * 4 / 2 ** 30
    cap = args.cpu_rows_cap
    if cap < 0:
        cap = 0 if (avail is None or avail >= max(160.0, 1.6 * need) or need < 8) else (10000000 if avail >= 48 else 1000000)
        if avail is None and need >= 8:
            cap = 10000000
    rows = [min(r, cap) if cap else r for r in cfg["rows"]]
    D, T, B = cfg["D"], len(rows), (args.cpu_batch or cfg["B"])
    ln_top = [data.num_interactions(T, D)] + list(cfg["top"][1:])
    np.random.seed(123)
    torch.manual_seed(123)
    model = OT.Model(rows, D, cfg["bot"], ln_top, init="torch")
    rng = np.random.default_rng(727)
    batches = [data.random_batch(rows, B, cfg["L"], cfg["bot"][0], fixed=True, rng=rng) for _ in range(2)]
    phases = {"bottom_mlp": [], "embedding_fwd": [], "interaction": [], "top_mlp_loss": [], "backward": [], "optimizer": []}
    times = []

    def one(i, record):
        X, lS_o, lS_i, tgt = batches[i % 2]
        t = [time.perf_counter()]
        for p in model.params():
            p.grad = None
        x = model.bot(X); t.append(time.perf_counter())                               # noqa: E702
        ly = OT.lookup(model.tables, lS_o, lS_i); t.append(time.perf_counter())      # noqa: E702
        z = OT.interact(x, ly, model.op, model.itself); t.append(time.perf_counter())  # noqa: E702
        loss = model.loss_fn(model.top(z), tgt); t.append(time.perf_counter())       # noqa: E702
        loss.backward(); t.append(time.perf_counter())                                # noqa: E702
        model.steps += 1
        model.apply_optimizer(0.01, args.optimizer)
        t.append(time.perf_counter())
        if record:
            for k, a, b in zip(phases, t[:-1], t[1:]):
                phases[k].append((b - a) * 1e3)
            times.append((t[-1] - t[0]) * 1e3)
    for i in range(warmup):
        one(i, False)
    for i in range(steps):
        one(i, True)
    med = float(np.median(times))
    sample = ("median of %d steps after %d warm-ups; batch %d; %s tables%s (sum %.1f M rows, %.1f GB); torch %s CPU ops, "
              "%d threads on %s; host RAM available %s GB"
              % (steps, warmup, B, args.config, (" capped at %d rows" % cap) if cap else " at full size", sum(rows) / 1e6,
                 sum(rows) * D * 4 / 1e9, torch.__version__, cores, cpu_model, "?" if avail is None else "%.0f" % avail))
    base = dict(value=B / (med / 1e3), unit="samples/s", cores=cores, kind="port", sample=sample, cpu_model=cpu_model,
                ms_per_step_median=med, rows_cap=cap or None, batch=B,
                phases_ms_median={k: round(float(np.median(v)), 3) for k, v in phases.items()})
    return base, med, B, rows


def gpu_reference(args, cfg, steps=20, warmup=5):
    """The reference's GPU path on THIS GPU (N = 1): the same ATen CUDA operators the unmodified reference runs with
    --use-gpu -- nn.EmbeddingBag(sparse=True) per table, bmm + tril gather, torch's sparse SGD update (the reference
    refuses rwsadagrad on GPU, dlrm_s_pytorch.py:1636-1637), fp32 cuBLAS towers -- strung together by
    oracle/oracle_torch.py, CUDA-event timed as the reference times itself (dlrm_s_pytorch.py:1844-1846, 1936-1939).
    A reported comparator, never on the product path."""
    from dlrm_yx_b200 import data
    from oracle import oracle_torch as OT
    dev = torch.device("cuda", torch.cuda.current_device())
    D, T, B = cfg["D"], len(cfg["rows"]), cfg["B"]
    ln_top = [data.num_interactions(T, D)] + list(cfg["top"][1:])
    np.random.seed(123)
    torch.manual_seed(123)
    model = OT.Model(cfg["rows"], D, cfg["bot"], ln_top, init="torch", device=dev)
    rng = np.random.default_rng(727)
    batches = []
    for _ in range(4):
        X, lS_o, lS_i, tgt = data.random_batch(cfg["rows"], B, cfg["L"], cfg["bot"][0], fixed=True, rng=rng)
        batches.append((X.to(dev), [o.to(dev) for o in lS_o], [i.to(dev) for i in lS_i], tgt.to(dev)))
    ev = {k: [] for k in ("step", "embedding_fwd", "interaction", "backward", "optimizer")}

    def one(i, record):
        X, lS_o, lS_i, tgt = batches[i % 4]
        marks = [torch.cuda.Event(enable_timing=True) for _ in range(7)]
        for p in model.params():
            p.grad = None
        marks[0].record()
        x = model.bot(X)
        marks[1].record()
        ly = OT.lookup(model.tables, lS_o, lS_i)
        marks[2].record()
        z = OT.interact(x, ly, model.op, model.itself)
        marks[3].record()
        loss = model.loss_fn(model.top(z), tgt)
        marks[4].record()
        loss.backward()
        marks[5].record()
        model.steps += 1
        model.apply_optimizer(0.01, "sgd")
        marks[6].record()
        if record:
            ev["step"].append((marks[0], marks[6]))
            ev["embedding_fwd"].append((marks[1], marks[2]))
            ev["interaction"].append((marks[2], marks[3]))
            ev["backward"].append((marks[4], marks[5]))
            ev["optimizer"].append((marks[5], marks[6]))
    for i in range(warmup):
        one(i, False)
    torch.cuda.synchronize()
    for i in range(steps):
        one(i, True)
    torch.cuda.synchronize()
    med = {k: float(np.median([a.elapsed_time(b) for a, b in v])) for k, v in ev.items()}
    out = {"what": "ATen CUDA nn.EmbeddingBag(sparse) + bmm/tril interaction + sparse SGD + fp32 cuBLAS towers (the "
                   "reference's --use-gpu operators via oracle/oracle_torch.py), same GPU, same shape, batch %d; CUDA events, "
                   "median of %d steps after %d warm-ups" % (B, steps, warmup),
           "optimizer": "sgd (the reference refuses rwsadagrad on GPU)", "ms_per_step": round(med["step"], 4),
           "samples_per_s": B / (med["step"] / 1e3),
           "phases_ms_median": {k: round(v, 4) for k, v in med.items() if k != "step"}}
    del model, batches
    return out


def workload_name(args, cfg, W, B_global):
    base = {"terabyte": "Terabyte/MLPerf shape synthetic: 26 tables (up to 40M rows, 187.8M total), dim 128, "
                        "bot 13-512-256-128, top 479-1024-1024-512-256-1, 1 index/lookup",
            "bench": "dlrm_s_benchmark.sh shape: 8 tables x 1M rows, dim 64, 100 indices/lookup",
            "kaggle": "Criteo Kaggle shape synthetic: 26 tables, dim 16"}[args.config]
    return "%s, global batch %d, %s" % (base, B_global, args.optimizer)


def embbag_c3_probe(peak, iters=20):
    """BASELINE.json configs[2] (bench/dlrm_s_benchmark.sh:36-42: 8 tables x 1M rows, D=64, 100 indices per lookup,
    batch 2048, SGD) as a kernel-level line: the bandwidth-bound embedding-bag of metric (ii), timed with CUDA events,
    4 rotating batches, 2 GB of tables (>> L2).  Algorithmic bytes as in DESIGN.md section 4."""
    from dlrm_yx_b200 import _lib, ops
    dev = torch.device("cuda", torch.cuda.current_device())
    T, R, D, B, L = 8, 1000000, 64, 2048, 100
    sampler = ClockSampler(torch.cuda.current_device())
    sampler.start()
    g = torch.Generator(device=dev)
    g.manual_seed(727)
    tw = [torch.empty((R, D), device=dev).uniform_(-(1.0 / R) ** 0.5, (1.0 / R) ** 0.5, generator=g) for _ in range(T)]
    ts = ops.TableSet(tw)
    N = T * B * L
    batches = []
    for _ in range(4):     # the reference draws L distinct sorted ids per bag (dlrm_data_pytorch.py:1195-1208)
        idx = torch.sort(torch.randint(0, R, (T * B, L), device=dev, generator=g), dim=1)[0].reshape(-1).to(torch.int32)
        batches.append((idx, (torch.arange(T * B + 1, device=dev) * L).to(torch.int32)))
    out = torch.empty((B, T, D), device=dev)
    grad = torch.randn((B, T, D), device=dev, generator=g)
    it = [0]

    def timed(fn):
        for _ in range(10):
            fn()
        torch.cuda.synchronize()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
        for s_, e_ in evs:
            s_.record()
            fn()
            e_.record()
        torch.cuda.synchronize()
        ms = sorted(a.elapsed_time(b) for a, b in evs)
        return ms[len(ms) // 2]

    def fwd():
        idx, off = batches[it[0] % 4]
        it[0] += 1
        ops.embbag_forward(ts, idx, off, B, out=out, stride_b=T * D, stride_t=D, validate=False)
    ms_f = timed(fwd)
    srt = [ops.bwd_sort(ts, idx, off, B) for idx, off in batches]
    U = sum(int(ops.bwd_segments(sk)[2].item()) for sk, _ in srt) / 4.0

    def sort():
        idx, off = batches[it[0] % 4]
        it[0] += 1
        ops.bwd_sort(ts, idx, off, B)
    ms_s = timed(sort)

    def upd():
        sk, sb = srt[it[0] % 4]
        it[0] += 1
        ops.bwd_update(ts, _lib.OPT_SGD, grad, T * D, D, B, sk, sb, 0.01, 0.0)
    ms_u = timed(upd)
    fb = N * D * 4 + N * 4 + T * B * 4 + T * B * D * 4
    ub = T * B * D * 4 + N * 8 + 2 * U * D * 4
    # the same launches over bf16 table rows (weights_precision="bf16"): table bytes halve, everything else stays fp32
    ts16 = ops.TableSet([w.to(torch.bfloat16) for w in tw])

    def fwd16():
        idx, off = batches[it[0] % 4]
        it[0] += 1
        ops.embbag_forward(ts16, idx, off, B, out=out, stride_b=T * D, stride_t=D, validate=False)
    ms_f16 = timed(fwd16)

    def upd16():
        sk, sb = srt[it[0] % 4]
        it[0] += 1
        ops.bwd_update(ts16, _lib.OPT_SGD, grad, T * D, D, B, sk, sb, 0.01, 0.0)
    ms_u16 = timed(upd16)
    fb16 = N * D * 2 + N * 4 + T * B * 4 + T * B * D * 4
    ub16 = T * B * D * 4 + N * 8 + 2 * U * D * 2
    res = {"bf16_table_rows": {
               "embbag_fwd": {"ms": round(ms_f16, 4), "algorithmic_MB": round(fb16 / 1e6, 1), "GBps": round(fb16 / ms_f16 / 1e6, 1),
                              "frac": round(fb16 / ms_f16 / 1e6 / peak, 4)},
               "bwd_update": {"ms": round(ms_u16, 4), "algorithmic_MB": round(ub16 / 1e6, 1), "GBps": round(ub16 / ms_u16 / 1e6, 1),
                              "frac": round(ub16 / ms_u16 / 1e6 / peak, 4)}},
           "workload": "dlrm_s_benchmark.sh shape: 8 tables x 1M rows, dim 64, 100 indices/lookup, batch 2048, SGD",
           "embbag_fwd": {"ms": round(ms_f, 4), "algorithmic_MB": round(fb / 1e6, 1), "GBps": round(fb / ms_f / 1e6, 1),
                          "frac": round(fb / ms_f / 1e6 / peak, 4)},
           "bwd_sort": {"ms": round(ms_s, 4)},
           "bwd_update": {"ms": round(ms_u, 4), "algorithmic_MB": round(ub / 1e6, 1), "GBps": round(ub / ms_u / 1e6, 1),
                          "frac": round(ub / ms_u / 1e6 / peak, 4), "unique_rows": int(U)}}
    res["clocks"] = sampler.stop()
    del tw, ts, ts16, batches, srt, out, grad
    return res


_REAL_STDOUT = None


def _guard_stdout():
    """stdout carries exactly ONE JSON line: everything else that libraries write to file descriptor 1 (NCCL prints its
    version banner there) is sent to stderr; the line itself goes to a duplicate of the original descriptor."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def _emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    _guard_stdout()
    if args.impl == "reference":
        if rank != 0:
            return
        cfg = config_of(args)
        base, ms, B, rows = cpu_reference(args, max(args.steps, 1), max(args.warmup, 1))
        capped = base["rows_cap"] is not None
        line = {"impl": "reference", "metric": "DLRM train samples/sec", "value": base["value"], "unit": "samples/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32",
                "data": "synthetic",
                "config": {"workload": workload_name(args, cfg, 1, B) + (
                               " -- CPU arm: rows capped at %d per table (host RAM), %.1f M rows in all" % (
                                   base["rows_cap"], sum(rows) / 1e6) if capped else ""),
                           "global_batch": B, "per_gpu_batch": B,
                           "device": "host CPU cores, one process (%d threads, %s)" % (base["cores"], base["cpu_model"]),
                           "rows_cap": base["rows_cap"]},
                "cpu_baseline": base,
                "e2e": {"value": base["value"], "unit": "samples/s", "h2d_bytes_per_step": 0,
                        "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        _emit(line)
        return

    if args.mgpu_overlap:
        # one hardware queue per stream: the schedule has three chains of device barriers in flight at once
        os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
        from dlrm_yx_b200 import dlrm_net as _dn
        _dn.set_mgpu_overlap(args.mgpu_overlap)
    if args.tunable:
        from dlrm_yx_b200 import _lib
        for kv in args.tunable:
            k, v = kv.split("=")
            _lib.set_tunable(int(k), int(v))
    r = run_ours(args)
    if r["rank"] != 0:
        return
    cfg, W, lb, Bg = r["cfg"], r["W"], r["lb"], r["B_global"]
    D = cfg["D"]
    K = args.steps
    peak, peak_src = measured_peaks()
    Tl = len(r["local_tables"])
    Nl = Tl * Bg * cfg["L"]
    # distinct rows touched per step on rank 0 (measured on the first resident batch)
    from dlrm_yx_b200 import ops
    emb = r["model"]._batched_module()
    U = 0
    if Tl > 0:
        skey, _ = ops.bwd_sort(emb.table_set(), r["devb"][0][1], r["devb"][0][2], Bg)
        U = int(ops.bwd_segments(skey)[2].item())
    fwd_b, bwd_b = algorithmic_bytes(Nl, Tl * Bg, D, U, tb=2 if args.table_dtype == "bf16" else 4)
    kern = {}
    for name, nbytes in (("embbag_fwd", fwd_b), ("bwd_update", bwd_b)):
        if name in r["prof"]:
            calls, ms = r["prof"][name]
            kern[name] = {"ms": round(ms, 4), "calls_per_step": calls / K}
            if W == 1:      # one launch per step at N = 1: the algorithmic bytes of that launch
                kern[name].update({"algorithmic_MB": round(nbytes / 1e6, 1), "GBps": round(nbytes / ms / 1e6, 1),
                                   "frac": round(nbytes / ms / 1e6 / peak, 4)})
    for name in r["prof"]:
        if name not in kern:
            kern[name] = {"ms": round(r["prof"][name][1], 4), "calls_per_step": r["prof"][name][0] / K}
    f = kern.get("embbag_fwd", {"ms": None})
    traffic = None                       # DRAM bytes per launch of that kernel from the committed ncu --set full capture
    tp = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "traffic.json")
    if os.path.exists(tp) and W == 1 and args.config == "terabyte" and not args.rows_cap:
        tj = json.load(open(tp))
        traffic = tj["embbag_fwd_kernel"]["dram_read_bytes"] + tj["embbag_fwd_kernel"]["dram_write_bytes"]
    gemm_txt = "the repo's tcgen05 GEMM kernels" if args.tower_gemm == "tc" else "cuBLASLt GEMMs (comparator run)"
    mlp_txt = {"fused-bf16": "bf16 in / fp32 accumulate (%s), bf16 activations" % gemm_txt}.get(args.mlp, args.mlp)
    ix_txt = ("single-pass TF32 products, fp32 accumulate, bf16 x / R / dR / dx" if args.mlp == "fused-bf16"
              else "3xTF32 (fp32-accurate), fp32 I/O")
    line = {
        "metric": "DLRM train samples/sec", "value": Bg * K / (r["ms_dev"] / 1e3), "unit": "samples/s",
        "n_gpus": W, "steps": K, "warmup": max(args.warmup, 3), "ms_per_step": r["ms_dev"] / K,
        "higher_is_better": True, "scaling": "strong" if args.global_batch else "weak", "vs_baseline": None,
        "dtype": "%s embedding tables, fp32 pooled rows, gradients and sparse optimizer%s; interaction: %s; MLP GEMMs: %s"
                 % ("bf16-row (widened on load, rounded to nearest on store)" if args.table_dtype == "bf16" else "fp32",
                    " (between GPUs the replicated tables' gradient sums%s cross NVLink rounded to bf16, accumulated in fp32)"
                    % (" and the pooled-row gradients of the sharded tables" if args.grad_wire == "bf16" else "")
                    if (W > 1 and args.mlp == "fused-bf16") else "",
                    ix_txt, mlp_txt),
        "data": "synthetic",
        "config": {"workload": workload_name(args, cfg, W, Bg), "global_batch": Bg, "per_gpu_batch": lb,
                   "index_distribution": args.index_dist,
                   "parallelism": "single GPU" if W == 1 else "table-wise model parallel x%d (%s sharder) for the %d tables "
                   "with >= %d rows, the other %d replicated on every rank (local lookups, dense-gradient all-reduce); "
                   "data-parallel MLPs (flat-gradient all-reduce); all-reduces: %s; %s" % (
                       W, args.sharder, len(cfg["rows"]) - len(r["rep_tables"]), args.replicate_below, len(r["rep_tables"]),
                       ("own two-shot kernel over NVLink peer memory" if (args.allreduce == "peer" and args.mlp == "fused-bf16")
                        else "NCCL"),
                       "exchange fused into the lookup / gradient kernels over NVLink peer memory"
                       if args.exchange == "peer" else "NCCL all-to-all"),
                   "l2": "tables (%.1f GB on rank 0) and %d rotating batches exceed L2; no flush needed"
                         % (sum(cfg["rows"][t] for t in r["local_tables"]) * D * (2 if args.table_dtype == "bf16" else 4) / 1e9,
                            args.nbatches),
                   "rows_cap": args.rows_cap or None, **({"tunables": args.tunable} if args.tunable else {}),
                   **({"mgpu_overlap": args.mgpu_overlap, "grad_wire": args.grad_wire} if W > 1 else {})},
        "e2e": {"value": Bg * K / (r["wall_e2e"] / 1e3), "unit": "samples/s", "h2d_bytes_per_step": r["h2d"],
                "d2h_bytes_per_step": 4, "ms_per_step": r["wall_e2e"] / K},
        "gpu_launches": r["launches"],
        "clocks": r["clocks"],
        "launch_mode": "cuda-graph (whole step)" if r["use_graph"] else "eager",
        "kernels": kern,
    }
    timing = (("CUDA events (external event-record nodes) around each launch inside a single-stream CUDA "
               "graph of the same step, replayed %d times right after the timed regions (%.3f ms/step)"
               if r["prof_mode"] == "graph" else
               "CUDA events around each launch, %d eager single-stream steps right after the timed "
               "regions (%.3f ms/step)") % (K, r["ms_eager"] / K))
    if W == 1 and "GBps" in f:
        line["roofline"] = {"kernel": "embbag_fwd_kernel (table-batched embedding-bag forward)", "bound": "hbm",
                            "achieved": f["GBps"], "peak": peak, "unit": "GB/s", "frac": f["frac"], "traffic": traffic,
                            "peak_source": peak_src, "algorithmic_bytes_per_launch": fwd_b, "ms_per_launch": f["ms"],
                            "timing": timing}
    else:
        # N > 1: the lookup of the sharded tables is NVLink-bound (remote row stores): bytes that leave this rank / time
        remote = (W - 1) / W * Tl * Bg * D * 4
        ms = f["ms"]
        line["roofline"] = {"kernel": "embbag_fwd_kernel<SCATTER> (lookup storing pooled rows into the peers' memory)",
                            "bound": "nvlink", "achieved": round(remote / ms / 1e6, 1) if ms else None, "peak": 770.0,
                            "unit": "GB/s", "frac": round(remote / ms / 1e6 / 770.0, 4) if ms else None, "traffic": None,
                            "peak_source": "B200_PROFILING.md: measured peer copy 770 GB/s per direction (900 nominal)",
                            "algorithmic_bytes_per_launch": remote, "ms_per_launch": ms, "timing": timing}
    if r["parity"] is not None:
        line["parity"] = r["parity"]
    if r["gpu_ref"] is not None:
        line["gpu_reference"] = r["gpu_ref"]
        if "samples_per_s" in r["gpu_ref"]:
            line["gpu_reference"]["ours_over_it"] = round(line["value"] / r["gpu_ref"]["samples_per_s"], 2)
    if W == 1 and args.config == "terabyte" and not args.no_c3_probe:
        try:
            line["embedding_bag_c3"] = embbag_c3_probe(peak)
        except Exception as ex:                       # the probe must never cost the headline line
            line["embedding_bag_c3"] = {"error": repr(ex)}
    if W == 1 and not args.no_cpu_baseline:
        del r["model"], r["devb"]
        torch.cuda.empty_cache()
        line["cpu_baseline"] = cpu_reference(args, args.cpu_steps, args.cpu_warmup)[0]
    _emit(line)


if __name__ == "__main__":
    main()
    if torch.distributed.is_available() and torch.distributed.is_initialized():
        torch.distributed.destroy_process_group()
