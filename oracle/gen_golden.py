"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz by running the UNMODIFIED reference
(imported from /root/reference through oracle/ref_shim.py) on seeded inputs.

Run in the build container:   python -m oracle.gen_golden
The fixtures are small, are committed, and travel to the GPU box where /root/reference does
not exist.  Every array in them is an output of the reference's own code
(DLRM_Net / nn.EmbeddingBag / interact_features / RWSAdagrad / torch.optim.SGD / sharders).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLDEN = os.path.join(ROOT, "tests", "golden")

# (name, seed, ln_emb, D, ln_bot, ln_top, B, L, fixed_L, optimizer)
CASES = [
    # C1: tiny CPU run of BASELINE.json configs[0] (README.md:141; defaults dlrm_s_pytorch.py:1181-1185)
    ("c1_tiny_sgd", 123, [4, 3, 2], 2, [4, 3, 2], [8, 4, 2, 1], 2, 10, False, "sgd"),
    ("c1_tiny_rws", 123, [4, 3, 2], 2, [4, 3, 2], [8, 4, 2, 1], 2, 10, False, "rwsadagrad"),
    # C2-like: Kaggle MLP shape and 26 tables, rows capped so the fixture stays small
    ("c2_kaggle_small_rws", 123,
     [1460, 583, 1000, 1000, 305, 24, 1000, 633, 3, 1000, 1000, 1000, 1000, 27, 1000, 1000, 10,
      1000, 1000, 4, 1000, 18, 15, 1000, 105, 1000],
     16, [13, 512, 256, 64, 16], [367, 512, 256, 1], 128, 1, True, "rwsadagrad"),
    # C3-like: multi-hot pooling, dim 64, variable bag length incl. duplicates across bags
    ("c3_multihot_sgd", 727, [500, 400, 300, 200], 64, [16, 32, 64], [74, 32, 1], 64, 20, False, "sgd"),
    # C4-like: dim 128, L=1, skewed tiny tables next to larger ones (hot rows)
    ("c4_terabyte_small_rws", 123, [3000, 39, 17, 7, 20, 3, 7000, 1500, 63, 4, 10, 2208],
     128, [13, 64, 128], [206, 64, 1], 256, 1, True, "rwsadagrad"),
]


def run_case(ns, name, seed, ln_emb, D, ln_bot, ln_top, B, L, fixed_L, opt):
    np.random.seed(seed)
    torch.manual_seed(seed)
    ln_emb = np.array(ln_emb)
    ln_bot = np.array(ln_bot)
    ln_top = np.array(ln_top)
    m = ns.dlrm.DLRM_Net(D, ln_emb, ln_bot, ln_top, arch_interaction_op="dot",
                         sigmoid_bot=-1, sigmoid_top=ln_top.size - 2)
    out = {}
    for k, E in enumerate(m.emb_l):
        out["W0_%d" % k] = E.weight.detach().numpy().copy()
    for pre, mlp in (("bot", m.bot_l), ("top", m.top_l)):
        j = 0
        for layer in mlp:
            if isinstance(layer, torch.nn.Linear):
                out["%s_W_%d" % (pre, j)] = layer.weight.detach().numpy().copy()
                out["%s_b_%d" % (pre, j)] = layer.bias.detach().numpy().copy()
                j += 1
    # inputs from the reference's own generators (dlrm_data_pytorch.py:1165-1228, 1095-1106)
    X, lS_o, lS_i = ns.dp.generate_dist_input_batch(
        int(ln_bot[0]), ln_emb, B, L, fixed_L, "uniform", 0, 1, -1, 1)
    Tgt = ns.dp.generate_random_output_batch(B, 1, True)
    out["X"] = X.numpy()
    out["target"] = Tgt.numpy()
    for k in range(len(ln_emb)):
        out["off_%d" % k] = lS_o[k].numpy().astype(np.int64)
        out["idx_%d" % k] = lS_i[k].numpy().astype(np.int64)

    if opt == "sgd":
        optim = torch.optim.SGD(m.parameters(), lr=0.1)
    else:
        optim = ns.rwsadagrad.RWSAdagrad(m.parameters(), lr=0.1)

    # forward pieces, exactly the calls sequential_forward makes (dlrm_s_pytorch.py:732-770)
    x = m.apply_mlp(X, m.bot_l)
    ly = m.apply_emb(lS_o, lS_i)
    for t in ly:
        t.retain_grad()
    x.retain_grad()
    z = m.interact_features(x, ly)
    z.retain_grad()
    p = m.apply_mlp(z, m.top_l)
    loss = m.loss_fn(p, Tgt)
    optim.zero_grad()
    loss.backward()
    out["x_bot"] = x.detach().numpy()
    out["pooled"] = np.stack([t.detach().numpy() for t in ly], axis=1)        # [B,T,D]
    out["interact"] = z.detach().numpy()
    out["pred"] = p.detach().numpy()
    out["loss"] = np.array(loss.item(), dtype=np.float32)
    out["d_interact"] = z.grad.numpy().copy()
    out["d_x_bot"] = x.grad.numpy().copy()
    out["d_pooled"] = np.stack([t.grad.numpy() for t in ly], axis=1)          # [B,T,D]
    for k, E in enumerate(m.emb_l):
        g = E.weight.grad
        assert g.is_sparse
        out["graw_idx_%d" % k] = g._indices().numpy().copy()[0]               # uncoalesced
        gc = g.coalesce()
        out["gco_idx_%d" % k] = gc._indices().numpy().copy()[0]
        out["gco_val_%d" % k] = gc._values().numpy().copy()
    optim.step()
    for k, E in enumerate(m.emb_l):
        out["W1_%d" % k] = E.weight.detach().numpy().copy()
        if opt == "rwsadagrad":
            out["mom1_%d" % k] = optim.state[E.weight]["momentum"].numpy().copy()
    # a second step on the same batch exercises non-zero accumulator state
    p2 = m(X, lS_o, lS_i)
    loss2 = m.loss_fn(p2, Tgt)
    optim.zero_grad()
    loss2.backward()
    optim.step()
    out["loss2"] = np.array(loss2.item(), dtype=np.float32)
    for k, E in enumerate(m.emb_l):
        out["W2_%d" % k] = E.weight.detach().numpy().copy()
        if opt == "rwsadagrad":
            out["mom2_%d" % k] = optim.state[E.weight]["momentum"].numpy().copy()
    out["meta_ln_emb"] = ln_emb
    out["meta_ln_bot"] = ln_bot
    out["meta_ln_top"] = ln_top
    out["meta_misc"] = np.array([D, B, L, int(fixed_L), seed])
    out["meta_opt"] = np.array(opt)
    out["meta_lr"] = np.array(0.1)
    return out


def sharder_golden(ns):
    kaggle = [1460, 583, 10131227, 2202608, 305, 24, 12517, 633, 3, 93145, 5683, 8351593, 3194, 27,
              14992, 5461306, 10, 5652, 2173, 4, 7046547, 18, 15, 286181, 105, 142572]
    tera = [39884406, 39043, 17289, 7420, 20263, 3, 7120, 1543, 63, 38532951, 2953546, 403346, 10,
            2208, 11938, 155, 4, 976, 14, 39979771, 25641295, 39664984, 585935, 12972, 108, 36]
    out = {"kaggle": np.array(kaggle), "terabyte": np.array(tera)}
    for nm, Es in (("kaggle", kaggle), ("terabyte", tera), ("bench", [1000000] * 8)):
        for W in (1, 2, 3, 4, 8):
            for alg in ("naive", "naive_chunk", "greedy", "hardcode"):
                out["%s_%s_%d" % (nm, alg, W)] = np.array(ns.sharders.shard(np.array(Es), W, alg))
    return out


def criteo_bin_golden(ns):
    """The reference's CriteoBinDataset (data_loader_terabyte.py:205-252) run on a small synthetic binary file:
    150 records in batches of 64 (the last batch is short), categorical ids up to 2^31-1 plus a few negative ones
    (floor-modulo semantics of `%`), dense counts 0..10^6; per-table and batched layouts, with and without
    max_ind_range."""
    import importlib
    import tempfile
    dlt = importlib.import_module("data_loader_terabyte")
    rng = np.random.default_rng(2024)
    n, bs = 150, 64
    rec = np.zeros((n, 40), dtype=np.int32)
    rec[:, 0] = rng.integers(0, 2, size=n)
    rec[:, 1:14] = (10 ** rng.uniform(0, 6, size=(n, 13))).astype(np.int32) - 1
    rec[:, 14:] = rng.integers(0, 2 ** 31 - 1, size=(n, 26), dtype=np.int64).astype(np.int32)
    rec[::17, 20] = -rec[::17, 20] - 1
    rec[5, 14:] = 0
    out = {"records": rec, "batch_size": np.array(bs)}
    with tempfile.TemporaryDirectory() as d:
        f = os.path.join(d, "day.bin")
        rec.tofile(f)
        cf = os.path.join(d, "counts.npz")
        np.savez(cf, counts=np.arange(26))
        for batched in (False, True):
            for mir in (-1, 1000, 40000000):
                ds = dlt.CriteoBinDataset(f, cf, batch_size=bs, max_ind_range=mir, batched_or_fbgemm_emb=batched)
                out["len"] = np.array(len(ds))
                for i in range(len(ds)):
                    X, lS_o, lS_i, T = ds[i]
                    tag = "%s_%d_%d" % ("batched" if batched else "tables", mir, i)
                    out["X_" + tag] = X.numpy()
                    out["off_" + tag] = lS_o.numpy()
                    out["idx_" + tag] = lS_i.numpy()
                    out["T_" + tag] = T.numpy()
                del ds
    return out


def weighted_pooling_golden(ns):
    """The reference's weighted pooling (weighted_pooling="learned": one learnable weight per table ROW, gathered per
    looked-up index, dlrm_s_pytorch.py:475-478, 544-547) on a small model: pooled rows, their gradients, the gradient
    of the per-row weights and the coalesced sparse table gradients, all from the unmodified DLRM_Net on CPU."""
    np.random.seed(321)
    torch.manual_seed(321)
    ln_emb, D, B = np.array([50, 7, 300]), 16, 32
    ln_bot, ln_top = np.array([5, 8, D]), np.array([D + 6, 8, 1])
    m = ns.dlrm.DLRM_Net(D, ln_emb, ln_bot, ln_top, arch_interaction_op="dot", sigmoid_bot=-1,
                         sigmoid_top=ln_top.size - 2, weighted_pooling="learned")
    with torch.no_grad():
        for v in m.v_W_l:
            v.uniform_(0.5, 1.5)
    X, lS_o, lS_i = ns.dp.generate_dist_input_batch(int(ln_bot[0]), ln_emb, B, 6, False, "uniform", 0, 1, -1, 1)
    Tgt = ns.dp.generate_random_output_batch(B, 1, True)
    out = {"X": X.numpy(), "target": Tgt.numpy(), "D": np.array(D), "B": np.array(B), "ln_emb": ln_emb}
    x = m.apply_mlp(X, m.bot_l)
    ly = m.apply_emb(lS_o, lS_i)
    for t in ly:
        t.retain_grad()
    p = m.apply_mlp(m.interact_features(x, ly), m.top_l)
    loss = m.loss_fn(p, Tgt)
    loss.backward()
    out["loss"] = np.array(loss.item(), dtype=np.float32)
    out["ln_bot"], out["ln_top"] = ln_bot, ln_top
    for pre, mlp in (("bot", m.bot_l), ("top", m.top_l)):
        lins = [l for l in mlp if isinstance(l, torch.nn.Linear)]
        for j, l in enumerate(lins):
            out["%s_W_%d" % (pre, j)] = l.weight.detach().numpy().copy()
            out["%s_b_%d" % (pre, j)] = l.bias.detach().numpy().copy()
    for k in range(len(ln_emb)):
        out["W_%d" % k] = m.emb_l[k].weight.detach().numpy().copy()
        out["vW_%d" % k] = m.v_W_l[k].detach().numpy().copy()
        out["idx_%d" % k] = lS_i[k].numpy().astype(np.int64)
        out["off_%d" % k] = lS_o[k].numpy().astype(np.int64)
        out["pooled_%d" % k] = ly[k].detach().numpy().copy()
        out["d_pooled_%d" % k] = ly[k].grad.numpy().copy()
        out["vW_grad_%d" % k] = m.v_W_l[k].grad.numpy().copy()
        gc = m.emb_l[k].weight.grad.coalesce()
        out["gco_idx_%d" % k] = gc._indices().numpy().copy()[0]
        out["gco_val_%d" % k] = gc._values().numpy().copy()
    return out


def tricks_golden(ns):
    """The reference's QREmbeddingBag (tricks/qr_embedding_bag.py, all three operations) and PrEmbeddingBag
    (tricks/md_embedding_bag.py) on CPU: outputs, coalesced sparse gradients of the tables, projection gradient."""
    import importlib
    qr = importlib.import_module("tricks.qr_embedding_bag")
    md = importlib.import_module("tricks.md_embedding_bag")
    torch.manual_seed(99)
    rng = np.random.default_rng(99)
    n, D, c, B = 1000, 16, 37, 24
    lens = rng.integers(0, 6, size=B)
    idx = torch.from_numpy(rng.integers(0, n, size=int(lens.sum())).astype(np.int64))
    off = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64))
    out = {"idx": idx.numpy(), "off": off.numpy(), "meta": np.array([n, D, c, B])}
    nq = int(np.ceil(n / c))
    Wq = torch.randn(nq, D) * 0.5
    Wr = torch.randn(c, D) * 0.5
    out["Wq"], out["Wr"] = Wq.numpy().copy(), Wr.numpy().copy()
    for op in ("mult", "add", "concat"):
        E = qr.QREmbeddingBag(n, D, c, operation=op, mode="sum", sparse=True, _weight=[Wq.clone(), Wr.clone()])
        y = E(idx, off)
        gy = torch.randn(y.shape)
        y.backward(gy)
        out["qr_%s_out" % op], out["qr_%s_gy" % op] = y.detach().numpy().copy(), gy.numpy().copy()
        for nm, w in (("q", E.weight_q), ("r", E.weight_r)):
            gco = w.grad.coalesce()
            out["qr_%s_g%s_idx" % (op, nm)] = gco._indices().numpy().copy()[0]
            out["qr_%s_g%s_val" % (op, nm)] = gco._values().numpy().copy()
    P = md.PrEmbeddingBag(n, 8, D)
    out["md_W"], out["md_proj"] = P.embs.weight.detach().numpy().copy(), P.proj.weight.detach().numpy().copy()
    y = P(idx, off)
    gy = torch.randn(y.shape)
    y.backward(gy)
    gco = P.embs.weight.grad.coalesce()
    out["md_out"], out["md_gy"] = y.detach().numpy().copy(), gy.numpy().copy()
    out["md_g_idx"], out["md_g_val"] = gco._indices().numpy().copy()[0], gco._values().numpy().copy()
    out["md_gproj"] = P.proj.weight.grad.numpy().copy()
    return out


def main():
    sys.path.insert(0, ROOT)
    from oracle.ref_shim import load_reference
    ns = load_reference()
    os.makedirs(GOLDEN, exist_ok=True)
    cb = criteo_bin_golden(ns)
    np.savez_compressed(os.path.join(GOLDEN, "criteo_bin.npz"), **cb)
    print("wrote criteo_bin.npz")
    tg = tricks_golden(ns)
    np.savez_compressed(os.path.join(GOLDEN, "tricks.npz"), **tg)
    print("wrote tricks.npz")
    wp = weighted_pooling_golden(ns)
    np.savez_compressed(os.path.join(GOLDEN, "weighted_pooling.npz"), **wp)
    print("wrote weighted_pooling.npz")
    if "--only-criteo" in sys.argv or "--only-small" in sys.argv:
        return
    for case in CASES:
        out = run_case(ns, *case)
        path = os.path.join(GOLDEN, case[0] + ".npz")
        np.savez_compressed(path, **out)
        print("wrote", path, os.path.getsize(path), "bytes; loss", out["loss"], out["loss2"])
    sh = sharder_golden(ns)
    np.savez_compressed(os.path.join(GOLDEN, "sharders.npz"), **sh)
    print("wrote sharders.npz")


if __name__ == "__main__":
    main()
