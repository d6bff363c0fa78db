"""TEST INFRASTRUCTURE ONLY -- golden vectors of the reference's MULTI-PROCESS path.

Runs the UNMODIFIED reference (imported from /root/reference through oracle/ref_shim.py) as W = 2 gloo processes on
the CPU: `ext_dist.init_distributed`, `DLRM_Net` built with `ext_dist.my_size > 1` (table-wise sharding block,
dlrm_s_pytorch.py:444-468), towers wrapped in `ext_dist.DDP` as the trainer does (dlrm_s_pytorch.py:1626-1633),
`dlrm_wrap`'s input split (124-167), `distributed_forward` (686-730) with `ext_dist.alltoall` (extend_distributed.py:
405-508, 601-639), loss on the LOCAL batch, backward, optimizer step -- two training steps.

    python -m oracle.gen_golden_dist           (build container only; writes tests/golden/dist_w2_*.npz)

Initial weights and the global batch are those of the single-process fixture of the same name (tests/golden/<case>.npz),
so one batch has a single-process AND a two-process reference answer.  What is recorded per case:
  sharder / device_indices / n_emb_per_rank, per-rank loss of both steps, the interaction input the ranks saw in step 1
  (rank-major feature order), every table and (rwsadagrad) accumulator after steps 1 and 2 (taken from the
  owning rank), the tower parameters after step 2 (identical on both ranks: DDP averages their gradients).
"""
import os
import sys

import numpy as np
import torch
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLDEN = os.path.join(ROOT, "tests", "golden")
# (fixture, sharder, keep the step-2 tables / tower parameters): the second c4 case only differs in the placement, so it
# keeps the small arrays (losses, maps, interaction input, step-1 tables)
CASES = [("c4_terabyte_small_rws", "naive_chunk", True), ("c4_terabyte_small_rws", "greedy", False),
         ("c3_multihot_sgd", "naive", True)]
W = 2


def _worker(rank, port, case, sharder, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(W),
                      LOCAL_RANK=str(rank))
    sys.path.insert(0, ROOT)
    from oracle import ref_shim
    ns = ref_shim.load_reference()
    torch.set_num_threads(2)
    ns.ext_dist.init_distributed(rank=rank, local_rank=rank, size=W, use_gpu=False, backend="gloo")
    g = np.load(os.path.join(GOLDEN, case + ".npz"))
    ln_emb, ln_bot, ln_top = g["meta_ln_emb"], g["meta_ln_bot"], g["meta_ln_top"]
    D, B = int(g["meta_misc"][0]), int(g["meta_misc"][1])
    lr, sgd = float(g["meta_lr"]), str(g["meta_opt"]) == "sgd"
    T = len(ln_emb)
    np.random.seed(int(g["meta_misc"][4]))
    torch.manual_seed(int(g["meta_misc"][4]))
    m = ns.dlrm.DLRM_Net(D, ln_emb, ln_bot, ln_top, arch_interaction_op="dot", sigmoid_bot=-1,
                         sigmoid_top=len(ln_top) - 2, sharder=sharder)
    with torch.no_grad():       # the sharded constructor draws only the local tables: load the fixture's weights
        for E, k in zip(m.emb_l, m.local_emb_indices):
            E.weight.copy_(torch.from_numpy(g["W0_%d" % k]))
        for pre, mlp in (("bot", m.bot_l), ("top", m.top_l)):
            lins = [l for l in mlp if isinstance(l, torch.nn.Linear)]
            for j, l in enumerate(lins):
                l.weight.copy_(torch.from_numpy(g["%s_W_%d" % (pre, j)]))
                l.bias.copy_(torch.from_numpy(g["%s_b_%d" % (pre, j)]))
    m.bot_l = ns.ext_dist.DDP(m.bot_l)
    m.top_l = ns.ext_dist.DDP(m.top_l)
    optim = (torch.optim.SGD(m.parameters(), lr=lr) if sgd else ns.rwsadagrad.RWSAdagrad(m.parameters(), lr=lr))
    X = torch.from_numpy(g["X"])
    tgt = torch.from_numpy(g["target"])
    lS_o = [torch.from_numpy(g["off_%d" % k]) for k in range(T)]
    lS_i = [torch.from_numpy(g["idx_%d" % k]) for k in range(T)]
    sl = ns.ext_dist.get_my_slice(B)
    out = {"device_indices": np.array(m.device_indices), "n_emb_per_rank": np.array(m.n_emb_per_rank),
           "local_emb_indices": np.array(m.local_emb_indices)}
    # record the interaction input of step 1 by wrapping the bound method of THIS instance (the reference code is untouched)
    seen = {}
    orig = m.interact_features

    def spy(x, ly):
        if "ly" not in seen:
            seen["x"] = x.detach().clone()
            seen["ly"] = [t.detach().clone() for t in ly]
        return orig(x, ly)
    m.interact_features = spy
    for step in (1, 2):
        # dlrm_wrap with use_gpu=False passes everything through; its split (dlrm_s_pytorch.py:145-148) is done here
        Z = m(X[sl], [lS_o[i] for i in m.local_emb_indices], [lS_i[i] for i in m.local_emb_indices])
        E = m.loss_fn(Z, tgt[sl])
        optim.zero_grad()
        E.backward()
        optim.step()
        out["loss%d" % step] = np.array(E.item(), dtype=np.float64)
        for E_, k in zip(m.emb_l, m.local_emb_indices):
            out["W%d_%d" % (step, k)] = E_.weight.detach().numpy().copy()
            if not sgd:
                out["mom%d_%d" % (step, k)] = optim.state[E_.weight]["momentum"].numpy().copy()
    out["feat_x"] = seen["x"].numpy()
    # wait() returns W tensors [lb, T_p*D] (extend_distributed.py:482-487); interact_features cats them (dlrm_s_pytorch.py:
    # 629-637): [lb, T, D] with the tables in rank-major order
    out["feat_ly"] = torch.cat(seen["ly"], dim=1).view(seen["x"].shape[0], -1, D).numpy()
    for pre, mlp in (("bot", m.bot_l.module), ("top", m.top_l.module)):
        lins = [l for l in mlp if isinstance(l, torch.nn.Linear)]
        for j, l in enumerate(lins):
            out["%s_W2_%d" % (pre, j)] = l.weight.detach().numpy().copy()
            out["%s_b2_%d" % (pre, j)] = l.bias.detach().numpy().copy()
    ret[rank] = out
    import torch.distributed as dist
    dist.barrier()
    dist.destroy_process_group()


def run_case(case, sharder, full, port):
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(port, case, sharder, ret), nprocs=W, join=True)
    merged = {"meta_case": np.array(case), "meta_sharder": np.array(sharder), "meta_W": np.array(W)}
    for r in range(W):
        for k, v in ret[r].items():
            if not full and (k.startswith(("W2_", "mom2_", "bot_", "top_"))):
                continue
            if k.startswith(("W1_", "W2_", "mom1_", "mom2_")) or k in ("device_indices", "n_emb_per_rank"):
                merged[k] = v                      # tables: owned by exactly one rank; maps: identical on all ranks
            elif k.startswith(("bot_", "top_")):
                if r > 0:
                    assert np.array_equal(merged[k], v), "DDP replicas diverged: " + k
                merged[k] = v
            else:
                merged["r%d_%s" % (r, k)] = v
    path = os.path.join(GOLDEN, "dist_w%d_%s_%s.npz" % (W, case, sharder))
    np.savez_compressed(path, **merged)
    print("wrote", path, "losses", [float(merged["r%d_loss1" % r]) for r in range(W)],
          [float(merged["r%d_loss2" % r]) for r in range(W)])


def main():
    for i, (case, sharder, full) in enumerate(CASES):
        run_case(case, sharder, full, 29610 + i)


if __name__ == "__main__":
    main()
