"""Host-side logic either side of the kernels, on CPU: sharders (bit-exact vs the reference's own output),
packed layouts, batch slicing, and the model-parallel exchange run with 2 gloo processes (pure communication +
autograd plumbing: no kernels are involved when `Request.wait()` is used)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from dlrm_yx_b200 import data, sharders
from oracle import oracle_np as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sharders_match_reference_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "sharders.npz"))
    tables = {"kaggle": g["kaggle"], "terabyte": g["terabyte"], "bench": np.array([1000000] * 8)}
    n = 0
    for key in g.files:
        if key in ("kaggle", "terabyte"):
            continue
        nm, rest = key.split("_", 1)
        alg, W = rest.rsplit("_", 1)
        assert sharders.shard(tables[nm], int(W), alg) == g[key].tolist(), key
        assert O.shard(tables[nm].tolist(), int(W), alg) == g[key].tolist(), key
        n += 1
    assert n == 60
    assert list(data.TERABYTE_ROWS) == g["terabyte"].tolist() and list(data.KAGGLE_ROWS) == g["kaggle"].tolist()


def test_unknown_sharder_raises():
    with pytest.raises(ValueError):
        sharders.shard([1, 2, 3], 2, "nope")


def test_traffic_sharder_balances_table_counts():
    d = sharders.shard(data.TERABYTE_ROWS, 8, "traffic")
    counts = [d.count(r) for r in range(8)]
    assert max(counts) - min(counts) <= 1 and sorted(set(d)) == list(range(8))


def test_get_splits():
    assert sharders.get_splits(26, 8) == [4, 4, 3, 3, 3, 3, 3, 3]
    assert sharders.get_splits(8, 4) == [2, 2, 2, 2]


def test_pack_matches_oracle_and_roundtrips():
    rng = np.random.default_rng(3)
    X, lS_o, lS_i, tgt = data.random_batch([50, 7, 3, 1000], 16, 5, 4, fixed=False, rng=rng)
    idx, off = data.pack(lS_o, lS_i)
    oi, oo = O.pack_batched([o.numpy() for o in lS_o], [i.numpy() for i in lS_i])
    assert np.array_equal(idx.numpy(), oi) and np.array_equal(off.numpy(), oo)
    assert idx.dtype == torch.int32 and off.numel() == 4 * 16 + 1
    bo, bi = O.unpack_batched(idx.numpy(), off.numpy(), 4, 16)
    for k in range(4):
        assert np.array_equal(bo[k], lS_o[k].numpy()) and np.array_equal(bi[k], lS_i[k].numpy())
    li, lo = data.local_packed(lS_o, lS_i, [1, 3])
    ri, ro = O.distribute_batched_local(idx.numpy(), off.numpy(), 4, 16, [1, 3])
    assert np.array_equal(li.numpy(), ri) and np.array_equal(lo.numpy(), ro)


def test_random_batch_distribution_rules():
    rng = np.random.default_rng(1)
    X, lS_o, lS_i, tgt = data.random_batch([1000, 3], 32, 10, 13, fixed=True, rng=rng)
    assert X.shape == (32, 13) and X.dtype == torch.float32 and tgt.shape == (32, 1)
    i0 = lS_i[0].view(32, 10).numpy()
    assert np.all(np.diff(i0, axis=1) > 0)                       # sorted, distinct within a bag
    assert lS_i[1].max() <= 2 and lS_o[1].numel() == 32           # a 3-row table cannot give 10 distinct ids
    assert data.num_interactions(26, 128) == 479 and data.num_interactions(26, 16) == 367
    assert data.num_interactions(8, 64) == 100 and data.num_interactions(3, 2) == 8


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _a2a_worker(rank, W, port, batched, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    from dlrm_yx_b200 import extend_distributed as ext
    ext.init_distributed(rank=rank, local_rank=rank, size=W, use_gpu=False, backend="gloo")
    B, D, tpr = 6, 4, [2, 1]
    rng = np.random.default_rng(5)
    pooled = [rng.normal(size=(B, t, D)).astype(np.float32) for t in tpr]      # every rank knows all (test only)
    coef = [rng.normal(size=(B // W, sum(tpr), D)).astype(np.float32) for _ in range(W)]
    mine = torch.from_numpy(pooled[rank].copy()).requires_grad_(True)
    if batched:
        req = ext.alltoall([mine], tpr, True)
        outs = req.wait()
        got = torch.cat([o.view(B // W, -1, D) for o in outs], dim=1)
    else:
        ins = [mine[:, t, :] for t in range(tpr[rank])]
        req = ext.alltoall(ins, tpr, False)
        outs = req.wait()
        got = torch.cat(outs, dim=1).view(B // W, -1, D)
    (got * torch.from_numpy(coef[rank])).sum().backward()
    np.savez(os.path.join(out_dir, "r%d_%d.npz" % (rank, int(batched))), out=got.detach().numpy(),
             grad=mine.grad.numpy(), slice=np.array([ext.get_my_slice(B).start, ext.get_my_slice(B).stop]))
    ext.barrier()
    torch.distributed.destroy_process_group()


@pytest.mark.parametrize("batched", [False, True])
def test_alltoall_two_ranks_gloo(tmp_path, batched):
    W, B, D, tpr = 2, 6, 4, [2, 1]
    port = _free_port()
    mp.spawn(_a2a_worker, args=(W, port, batched, str(tmp_path)), nprocs=W, join=True)
    rng = np.random.default_rng(5)
    pooled = [rng.normal(size=(B, t, D)).astype(np.float32) for t in tpr]
    coef = [rng.normal(size=(B // W, sum(tpr), D)).astype(np.float32) for _ in range(W)]
    want = O.alltoall_fwd(pooled, W)
    want_grad = O.alltoall_bwd(coef, tpr, W)
    for r in range(W):
        got = np.load(os.path.join(str(tmp_path), "r%d_%d.npz" % (r, int(batched))))
        assert np.array_equal(got["out"], want[r])
        assert np.array_equal(got["grad"], want_grad[r])
        assert got["slice"].tolist() == [r * 3, r * 3 + 3]


# ---------------------------------------------------------------------------------------------------------------------
# multi-process path: fixtures of the reference's own 2-process run (oracle/gen_golden_dist.py) against the single-process
# fixtures, and the host-side input split with replicated tables
# ---------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("case,sharder", [("c4_terabyte_small_rws", "naive_chunk"), ("c4_terabyte_small_rws", "greedy"),
                                           ("c3_multihot_sgd", "naive")])
def test_distributed_reference_fixture_is_rank_major_slices_of_the_single_process_one(golden_dir, case, sharder):
    """What the reference's all-to-all delivers to rank r (extend_distributed.py:405-508): its batch slice of every table's
    pooled rows, tables in RANK-MAJOR order -- bit-identical to slices of the single-process run; the
    oracle's permutation (oracle_np.alltoall_fwd) restates exactly that."""
    g = np.load(os.path.join(golden_dir, case + ".npz"))
    d = np.load(os.path.join(golden_dir, "dist_w2_%s_%s.npz" % (case, sharder)))
    W, T = int(d["meta_W"]), len(g["meta_ln_emb"])
    B = int(g["meta_misc"][1])
    lb = B // W
    dev = d["device_indices"].tolist()
    assert dev == sharders.shard(g["meta_ln_emb"], W, sharder)
    order = [i for r in range(W) for i in range(T) if dev[i] == r]
    pooled = g["pooled"]                                     # [B, T, D] single process
    per_rank = [pooled[:, [i for i in range(T) if dev[i] == r], :] for r in range(W)]
    for r in range(W):
        got = d["r%d_feat_ly" % r]
        assert np.array_equal(got, pooled[r * lb:(r + 1) * lb][:, order, :])
        assert np.array_equal(O.alltoall_fwd(per_rank, W)[r], got)
        assert np.array_equal(d["r%d_feat_x" % r], g["x_bot"][r * lb:(r + 1) * lb])
    # each rank's loss is the mean over ITS slice; with equal slices their mean is the single-process loss -- when the
    # rank-major feature order is the table order (otherwise the top tower's columns meet other features: the reference's
    # multi-process model is a different function of the same weights)
    if order == list(range(T)):
        np.testing.assert_allclose(    np.mean([float(d["r%d_loss1" % r]) for r in range(W)]), float(g["loss"]), rtol=1e-6)


def test_distribute_batched_emb_data_with_replicated_tables():
    """DLRM_Net.distribute_batched_emb_data (dlrm_s_pytorch.py:772-802): local sharded tables x GLOBAL batch, and -- with
    replicated tables -- those tables x this rank's slice, offsets re-based; ragged bags."""
    from dlrm_yx_b200 import dlrm_net
    from dlrm_yx_b200 import extend_distributed as ext
    rng = np.random.default_rng(11)
    rows, B, W = [50, 7, 3, 1000, 9], 8, 2
    X, lS_o, lS_i, tgt = data.random_batch(rows, B, 4, 3, fixed=False, rng=rng)
    idx, off = data.pack(lS_o, lS_i)
    saved = (ext.my_size, ext.my_rank, ext.my_local_rank)
    try:
        for rank in range(W):
            ext.my_size, ext.my_rank, ext.my_local_rank = W, rank, rank
            m = dlrm_net.DLRM_Net()
            m.local_emb_indices = [0, 3] if rank == 0 else [4]
            m.replicated_emb_indices = [1, 2]
            m.emb_rep = object()
            o, i = m.distribute_batched_emb_data(B, off, idx)
            (so, ro), (si, ri) = o[0], i[0]
            ei, eo = data.local_packed(lS_o, lS_i, m.local_emb_indices)
            assert torch.equal(so, eo) and torch.equal(si, ei)
            lb = B // W
            ends = [torch.cat([lS_o[k], torch.tensor([lS_i[k].numel()])]) for k in range(len(rows))]
            exp_o, exp_i = [], []
            for k in m.replicated_emb_indices:
                lo, hi = int(ends[k][rank * lb]), int(ends[k][(rank + 1) * lb])
                exp_o.append(lS_o[k][rank * lb:(rank + 1) * lb] - lo)
                exp_i.append(lS_i[k][lo:hi])
            xi, xo = data.pack(exp_o, exp_i)
            assert torch.equal(ro, xo) and torch.equal(ri, xi)
            m.emb_rep = None
            o, i = m.distribute_batched_emb_data(B, off, idx)
            assert torch.equal(o[0], eo) and torch.equal(i[0], ei)
    finally:
        ext.my_size, ext.my_rank, ext.my_local_rank = saved


def test_installed_module_has_the_surface_the_reference_imports():
    """dlrm_s_pytorch.py:321-334 does `import table_batched_embeddings_ops` and calls
    TableBatchedEmbeddingBags(T, Es, D, optimizer=Optimizer.SGD, learning_rate=..., eps=0.1, stochastic_rounding=False)."""
    import importlib
    import inspect
    from dlrm_yx_b200 import embedding
    embedding.install_table_batched_embeddings_ops()
    mod = importlib.import_module("table_batched_embeddings_ops")
    assert mod.Optimizer.SGD is embedding.Optimizer.SGD
    sig = inspect.signature(mod.TableBatchedEmbeddingBags.__init__)
    names = list(sig.parameters)
    assert names[1:4] == ["num_tables", "num_embeddings", "embedding_dim"]
    for kw in ("optimizer", "learning_rate", "eps", "stochastic_rounding"):
        assert kw in sig.parameters
    assert list(inspect.signature(mod.TableBatchedEmbeddingBags.forward).parameters)[1:] == ["indices", "offsets"]


def test_multi_gpu_defaults_are_the_schedule_measured_on_eight_gpus(monkeypatch):
    """The opt-in multi-GPU schedule (dlrm_net.MGPU_OVERLAP != 0) and the bf16 gradient wire have only been measured on two
    GPUs: nothing may switch them on by default -- not the module, not bench.py's or
    tools/check_exchange.py's arguments, not the model's entry points."""
    import inspect
    import sys
    from dlrm_yx_b200 import dlrm_net, extend_distributed
    assert dlrm_net.MGPU_OVERLAP == 0
    assert inspect.signature(dlrm_net.DLRM_Net.enable_peer_exchange).parameters["grad_wire_bf16"].default is False
    assert inspect.signature(dlrm_net.DLRM_Net.fuse_mlps).parameters["grad_wire_bf16"].default is False
    assert inspect.signature(dlrm_net.DLRM_Net.fuse_mlps).parameters["allreduce"].default == "peer"
    assert inspect.signature(extend_distributed.PeerExchange.__init__).parameters["grad_bf16"].default is False
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import bench
    monkeypatch.setattr(sys, "argv", ["bench.py"])
    args = bench.parse()
    assert args.mgpu_overlap == 0 and args.grad_wire == "fp32" and args.allreduce == "peer"
