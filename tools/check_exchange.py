"""Multi-GPU parity of the model-parallel path.  Run under torchrun with W >= 2 GPUs:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node W --master-addr 127.0.0.1 tools/check_exchange.py \
        [--case c4_terabyte_small_rws] [--sharders naive_chunk,greedy,traffic] [--steps 3] [--out profiles/x.json]

For every sharder in the list and for replication off / on (tables below --replicate-below rows kept on every rank), every
rank builds the SAME model (weights of the reference-generated fixture tests/golden/<case>.npz) several times and trains
`--steps` steps on its slice of the fixture's global batch:

  A  NCCL all-to-all, fp32 towers (torch) -- the reference's structure (dlrm_s_pytorch.py:686-730, extend_distributed.py:405-508)
  B  peer-memory exchange (lookup kernel stores into the peers' interaction input, gradient push kernel), fp32 towers
  C  NCCL all-to-all, fused bf16 towers, bf16-facing interaction (pitched interaction input)
  D  peer-memory exchange, fused bf16 towers  -- the benchmarked configuration
  E  D replayed from ONE CUDA graph (GraphedTrainStep)          -- the benchmarked launch mode

Checks (any failure raises; exit code != 0):
  * B == A and D == C == E bit for bit, every step: losses, every local table, every accumulator, the interaction input --
    only the transport / launch mode differs, and symmetric-memory barriers are reused across >= 3 steps;
  * W == 2 and a reference fixture of the MULTI-PROCESS run exists (tests/golden/dist_w2_<case>_<sharder>.npz, written by
    oracle/gen_golden_dist.py from the unmodified reference under 2-rank gloo): the interaction input equals the
    reference's bit for bit (rank-major feature order), per-rank losses of step 1 at rtol 1e-5, tables and accumulators
    after step 1 at 1e-5 / 1e-6 (+ the stated reassociation slack of hot rows), step 2 losses at 1e-3;
  * replication on: pooled rows identical to the sharded run's (same bits, permuted feature slots), step-1 loss equal to
    the sharded run's at 1e-6 relative (the top tower sees permuted columns), tables after step 1 within 1e-5 / 1e-6 of the
    sharded run (the per-row sum over ranks is associated differently: all-reduce), replicas bit-identical across ranks;
  * bf16 variants track the fp32 ones (loss 2e-2).
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dlrm_yx_b200 import data, dlrm_net, embedding, ops, rwsadagrad  # noqa: E402
from dlrm_yx_b200 import extend_distributed as ext  # noqa: E402
from dlrm_yx_b200.graph import GraphedTrainStep  # noqa: E402


def permute_top_columns(W_ref, D, ref_order, my_order):
    """First top-tower weight [N, D + nz] of a model whose interaction features are ordered x, ref_order... re-expressed
    for the feature order x, my_order... (tril pairs, dlrm_s_pytorch.py:646-656)."""
    F = 1 + len(my_order)
    pos_ref = {t: 1 + k for k, t in enumerate(ref_order)}
    out = W_ref.copy()
    for i in range(1, F):
        for j in range(i):
            ri = pos_ref[my_order[i - 1]]
            rj = 0 if j == 0 else pos_ref[my_order[j - 1]]
            hi, lo = max(ri, rj), min(ri, rj)
            out[:, D + i * (i - 1) // 2 + j] = W_ref[:, D + hi * (hi - 1) // 2 + lo]
    return out


class Case(object):
    def __init__(self, name, dev):
        self.name, self.dev = name, dev
        self.g = g = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
        self.T, self.D, self.B = len(g["meta_ln_emb"]), int(g["meta_misc"][0]), int(g["meta_misc"][1])
        self.lr = float(g["meta_lr"])
        self.sgd = str(g["meta_opt"]) == "sgd"
        self.kind = embedding.Optimizer.SGD if self.sgd else embedding.Optimizer.EXACT_ROWWISE_ADAGRAD
        self.W, self.rank = ext.my_size, ext.my_rank
        assert self.B % self.W == 0, "fixture batch %d not divisible by %d ranks" % (self.B, self.W)
        self.lb = self.B // self.W
        self.tower_gemm = "tc"
        self.allreduce = "peer"
        self.grad_wire = "fp32"

    def build(self, sharder, exchange, bf16, replicate_below):
        g, dev = self.g, self.dev
        np.random.seed(int(g["meta_misc"][4]))
        m = dlrm_net.DLRM_Net(self.D, g["meta_ln_emb"], g["meta_ln_bot"], g["meta_ln_top"], arch_interaction_op="dot",
                              sigmoid_bot=-1, sigmoid_top=len(g["meta_ln_top"]) - 2, batched_emb=True,
                              emb_optimizer=self.kind, emb_learning_rate=self.lr, fuse_sparse_arch=True, sharder=sharder,
                              replicate_rows_below=replicate_below).to(dev)
        ref_order = [i for r in range(self.W) for i in range(self.T)
                     if dlrm_net.sharders.shard(g["meta_ln_emb"], self.W, sharder)[i] == r]
        with torch.no_grad():    # the sharded constructor draws only the local tables: load the fixture's weights
            for w, k in zip(m._batched_module().split_weights(), m.local_emb_indices):
                w.copy_(torch.from_numpy(g["W0_%d" % k]))
            if m.emb_rep is not None:
                for w, k in zip(m.emb_rep.split_weights(), m.replicated_emb_indices):
                    w.copy_(torch.from_numpy(g["W0_%d" % k]))
            for pre, mlp in (("bot", m.bot_l), ("top", m.top_l)):
                lins = [l for l in mlp if isinstance(l, torch.nn.Linear)]
                for j, l in enumerate(lins):
                    Wj = g["%s_W_%d" % (pre, j)]
                    if pre == "top" and j == 0:
                        Wj = permute_top_columns(Wj, self.D, ref_order, m.feature_tables)
                    l.weight.copy_(torch.from_numpy(Wj))
                    l.bias.copy_(torch.from_numpy(g["%s_b_%d" % (pre, j)]))
        m.sync_replicas()
        if exchange == "peer":
            m.enable_peer_exchange(self.lb, mixed=bool(bf16), grad_wire_bf16=bool(bf16) and self.grad_wire == "bf16")
        opt = None
        if bf16:
            m.fuse_mlps(self.lr, "sgd" if self.sgd else "adagrad", gemm=self.tower_gemm, allreduce=self.allreduce,
                        grad_wire_bf16=self.grad_wire == "bf16")
        else:
            dense = list(m.bot_l.parameters()) + list(m.top_l.parameters())
            m.bot_l = ext.DDP(m.bot_l, device_ids=[dev.index])
            m.top_l = ext.DDP(m.top_l, device_ids=[dev.index])
            opt = torch.optim.SGD(dense, lr=self.lr) if self.sgd else rwsadagrad.RWSAdagrad(dense, lr=self.lr)
        return m, opt

    def inputs(self, m):
        g, dev, lb, rank = self.g, self.dev, self.lb, self.rank
        lS_o = [torch.from_numpy(g["off_%d" % k]) for k in range(self.T)]
        lS_i = [torch.from_numpy(g["idx_%d" % k]) for k in range(self.T)]
        idx_all, off_all = data.pack(lS_o, lS_i)
        o, i = m.distribute_batched_emb_data(self.B, off_all, idx_all)
        sl = slice(rank * lb, (rank + 1) * lb)
        X, tgt = torch.from_numpy(g["X"][sl]).to(dev), torch.from_numpy(g["target"][sl]).to(dev)
        mv = lambda t: tuple(u.to(dev) for u in t) if isinstance(t, tuple) else t.to(dev)     # noqa: E731
        return X, mv(o[0]), mv(i[0]), tgt

    def run(self, m, opt, steps, graph=False):
        """-> dict(losses=[per step, this rank], tables=[per step: {table id: tensor}], moms=..., feat=[lb, T, D] in
        GLOBAL table order after step 1)."""
        X, off, idx, tgt = self.inputs(m)
        rep = m.emb_rep is not None
        flat = [X, tgt] + (list(off) + list(idx) if rep else [off, idx])

        def step_fn(X, tgt, *rest):
            o, i = ((rest[0], rest[1]), (rest[2], rest[3])) if rep else (rest[0], rest[1])
            loss = m.loss_fn(m(X, o, i).float(), tgt)
            if opt is not None:
                opt.zero_grad()
            loss.backward()
            if opt is not None:
                opt.step()
            else:
                m.step_mlps()
            return loss
        out = dict(losses=[], tables=[], moms=[], feat=None)
        gs = None
        for s in range(steps):
            if graph and s == 0:
                # the capture's warm-up runs ARE training steps: rewind the model afterwards so that replay s is step s
                state = [t.clone() for t in self.state_tensors(m)]
                gs = GraphedTrainStep(step_fn, flat, warmup=1)
                for dst, src in zip(self.state_tensors(m), state):
                    dst.copy_(src)
            loss = gs(*flat) if gs is not None else step_fn(*flat)
            out["losses"].append(float(loss.item()))
            tabs, moms = {}, {}
            mods = [(m._batched_module(), m.local_emb_indices)] + ([(m.emb_rep, m.replicated_emb_indices)] if rep else [])
            for mod, ids in mods:
                mo = mod.split_momentum()
                for j, k in enumerate(ids):
                    tabs[k] = mod.split_weights()[j].clone()
                    if mo is not None:
                        moms[k] = mo[j].clone()
            out["tables"].append(tabs)
            out["moms"].append(moms)
            if s == 0:
                feat = m._last_feat
                order = m.feature_tables
                glob = torch.empty((self.lb, self.T, self.D), device=self.dev)
                for slot, k in enumerate(order):
                    glob[:, k, :] = feat[:, 1 + slot, :]
                out["feat"] = glob
        ops.check_index_errors()
        return out

    @staticmethod
    def state_tensors(m):
        ts = [m._batched_module().embedding_weights.data]
        if m._batched_module().row_momentum is not None:
            ts.append(m._batched_module().row_momentum)
        if m.emb_rep is not None:
            ts.append(m.emb_rep.embedding_weights.data)
            if m.emb_rep.row_momentum is not None:
                ts.append(m.emb_rep.row_momentum)
        for tower in (m.bot_l, m.top_l):
            if hasattr(tower, "master"):
                ts += [tower.master.data, tower.acc, tower.grad, tower.shadow]
        return ts


def same(a, b, what):
    assert a["losses"] == b["losses"], (what, a["losses"], b["losses"])
    for s in range(len(a["tables"])):
        for k in a["tables"][s]:
            assert torch.equal(a["tables"][s][k], b["tables"][s][k]), "%s: table %d differs at step %d" % (what, k, s + 1)
        for k in a["moms"][s]:
            assert torch.equal(a["moms"][s][k], b["moms"][s][k]), "%s: accumulator %d differs at step %d" % (what, k, s + 1)
    assert torch.equal(a["feat"], b["feat"]), what + ": interaction input differs"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case", default="c4_terabyte_small_rws")
    ap.add_argument("--sharders", default="naive_chunk,greedy,traffic")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--replicate-below", type=int, default=100)
    ap.add_argument("--out", default="")
    ap.add_argument("--tower-gemm", default="tc", choices=["tc", "cublas"], help="GEMMs of the fused bf16 towers (cublas: comparator)")
    ap.add_argument("--allreduce", default="peer", choices=["peer", "nccl"], help="gradient sums of the fused towers / replicated tables")
    ap.add_argument("--mgpu-overlap", type=int, default=0, help="dlrm_net.MGPU_OVERLAP bit mask (see there; 15 = all)")
    ap.add_argument("--grad-wire", default="fp32", choices=["fp32", "bf16"], help="pooled-row gradients between GPUs (bf16 towers)")
    ap.add_argument("--losses-only", action="store_true", help="debug: print the bf16 variants' losses instead of asserting")
    args = ap.parse_args()
    ext.init_distributed(use_gpu=True, backend="nccl")
    W, rank = ext.my_size, ext.my_rank
    dev = torch.device("cuda", torch.cuda.current_device())
    torch.backends.cuda.matmul.allow_tf32 = False
    c = Case(args.case, dev)
    c.tower_gemm = args.tower_gemm
    c.allreduce = args.allreduce
    c.grad_wire = args.grad_wire
    if args.mgpu_overlap:
        dlrm_net.set_mgpu_overlap(args.mgpu_overlap)
    report = {"case": args.case, "world_size": W, "steps": args.steps, "checks": []}

    def note(msg):
        report["checks"].append(msg)
        if rank == 0:
            print("check_exchange:", msg, flush=True)

    hooked = dlrm_net._DistInteract.forward

    def spy(ctx, x, recv, step, itself, out_cols, relu_x=False):     # remember the interaction input of the last forward
        r = hooked(ctx, x, recv, step, itself, out_cols, relu_x)
        step.model._last_feat = step.feat
        return r
    dlrm_net._DistInteract.forward = staticmethod(spy)

    for sharder in args.sharders.split(","):
        base = None
        for rep_below in (0, args.replicate_below):
            tag = "%s%s" % (sharder, ", replicated < %d rows" % rep_below if rep_below else "")
            A = c.run(*c.build(sharder, "nccl", False, rep_below), steps=args.steps)
            B = c.run(*c.build(sharder, "peer", False, rep_below), steps=args.steps)
            same(A, B, tag + ": peer vs NCCL (fp32 towers)")
            note("%s: peer == NCCL bit for bit over %d steps (fp32 towers; losses, tables, accumulators, interaction input)"
                 % (tag, args.steps))
            C = c.run(*c.build(sharder, "nccl", True, rep_below), steps=args.steps)
            Dm = c.build(sharder, "peer", True, rep_below)
            Dr = c.run(*Dm, steps=args.steps)
            if args.losses_only:
                print("rank %d %s gemm=%s  fp32 A %s | bf16 nccl C %s | bf16 peer D %s | feat equal %s" % (
                    rank, tag, args.tower_gemm, A["losses"], C["losses"], Dr["losses"], torch.equal(C["feat"], Dr["feat"])), flush=True)
                continue
            same(C, Dr, tag + ": peer vs NCCL (bf16 towers)")
            E = c.run(*c.build(sharder, "peer", True, rep_below), steps=args.steps, graph=True)
            same(Dr, E, tag + ": CUDA-graph replay vs eager (peer, bf16 towers)")
            note("%s: peer == NCCL == CUDA-graph replay bit for bit over %d steps (bf16 towers, pitched interaction input)"
                 % (tag, args.steps))
            np.testing.assert_allclose(Dr["losses"][0], A["losses"][0], rtol=2e-2)
            # ---- against the reference's own multi-process run (W = 2 fixtures)
            gp = os.path.join(ROOT, "tests", "golden", "dist_w%d_%s_%s.npz" % (W, args.case, sharder))
            if os.path.exists(gp):
                d = np.load(gp)
                ref_order = [i for r in range(W) for i in range(c.T) if d["device_indices"][i] == r]
                ref_feat = np.empty((c.lb, c.T, c.D), np.float32)
                for slot, k in enumerate(ref_order):
                    ref_feat[:, k, :] = d["r%d_feat_ly" % rank][:, slot, :]
                assert np.array_equal(A["feat"].cpu().numpy(), ref_feat), tag + ": pooled rows differ from the reference's"
                np.testing.assert_allclose(A["losses"][0], float(d["r%d_loss1" % rank]), rtol=1e-5)
                for k, w in A["tables"][0].items():
                    ref = d["W1_%d" % k]
                    # 1e-5 relative; the absolute term covers the hot rows of the tiny tables, whose ~100-term gradient sums
                    # are associated differently (segment partials / all-reduce) -- cf. tests/test_kernels_gpu.py
                    np.testing.assert_allclose(w.cpu().numpy(), ref, rtol=1e-5, atol=1e-5)
                for k, mo in A["moms"][0].items():
                    np.testing.assert_allclose(mo.cpu().numpy(), d["mom1_%d" % k], rtol=2e-5, atol=1e-7)
                if args.steps > 1 and not c.sgd:
                    np.testing.assert_allclose(A["losses"][1], float(d["r%d_loss2" % rank]), rtol=1e-3)
                note("%s: interaction input == the reference's 2-process run bit for bit; loss 1e-5; tables / accumulators "
                     "after step 1 at 1e-5" % tag)
            if rep_below == 0:
                base = A
            else:
                assert torch.equal(A["feat"], base["feat"]), tag + ": pooled rows differ from the sharded run's"
                np.testing.assert_allclose(A["losses"][0], base["losses"][0], rtol=1e-6)
                # tables after step 1: gather every table from wherever it lives and compare with the sharded run's owner
                for k in range(c.T):
                    mine = A["tables"][0].get(k)
                    theirs = base["tables"][0].get(k)
                    owner_a = torch.tensor([rank if mine is not None else W], device=dev)
                    owner_b = torch.tensor([rank if theirs is not None else W], device=dev)
                    dist.all_reduce(owner_a, op=dist.ReduceOp.MIN)
                    dist.all_reduce(owner_b, op=dist.ReduceOp.MIN)
                    shape = (int(c.g["meta_ln_emb"][k]), c.D)
                    ta = mine.clone() if mine is not None else torch.empty(shape, device=dev)
                    tb = theirs.clone() if theirs is not None else torch.empty(shape, device=dev)
                    dist.broadcast(ta, int(owner_a.item()))
                    dist.broadcast(tb, int(owner_b.item()))
                    if mine is not None and k in c_rep(A):
                        assert torch.equal(ta, mine), "replicas of table %d diverged across ranks" % k
                    np.testing.assert_allclose(ta.cpu().numpy(), tb.cpu().numpy(), rtol=1e-5, atol=2e-5)
                note("%s: pooled rows bit-identical to the sharded run; step-1 tables within 1e-5 of it; replicas "
                     "bit-identical across ranks" % tag)
    dist.barrier()
    if rank == 0 and args.out:
        with open(os.path.join(ROOT, args.out), "w") as f:
            json.dump(report, f, indent=1)
    print("rank %d: exchange parity OK (%s, %d ranks)" % (rank, args.case, W), flush=True)
    dist.destroy_process_group()


def c_rep(run):
    """Table ids that are present on every rank in this run (replicated): all keys, the caller intersects by broadcast."""
    return set(run["tables"][0].keys())


if __name__ == "__main__":
    main()
