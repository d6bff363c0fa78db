"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the reference's sparse hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module, and only as the CHECKER.  Nothing under dlrm_yx_b200/ imports it.

Each function cites the reference file:line (relative to /root/reference) it restates.
Pinning: the reference's own tests cannot run (no Caffe2; README transcript is stale, SURVEY.md
section 4), so the oracle is pinned against OUTPUTS OF THE REFERENCE ITSELF, imported unmodified
through oracle/ref_shim.py in the build container (tests/test_oracle_vs_reference.py) and
against golden vectors generated from it (tests/golden/, made by oracle/gen_golden.py).

Arithmetic conventions: fp32 everywhere (np.float32), sums of looked-up rows are SEQUENTIAL in
index order (ATen EmbeddingBag sum semantics), integer artefacts (sort order, unique rows, run
offsets, packed offsets, shard maps) are exact.
"""
import numpy as np

f32 = np.float32


# ----------------------------------------------------------------------------------------------
# sharders.py:3-60  (table -> rank placement; integer, bit-exact)
# ----------------------------------------------------------------------------------------------
def get_splits(T, ndevices):
    """sharders.py:3-9."""
    k, m = divmod(T, ndevices)
    return [k] * ndevices if m == 0 else [(k + 1) if i < m else k for i in range(ndevices)]


def shard(Es, ndevices, alg="naive"):
    """sharders.py:23-60: naive = round robin; naive_chunk = contiguous chunks; greedy = each
    table (in order) goes to the rank with the fewest rows so far, ties -> lowest rank;
    hardcode = table 0 on rank 0, the rest on rank 1."""
    T = len(Es)
    if alg == "naive":
        return [t % ndevices for t in range(T)]
    if alg == "naive_chunk":
        out = []
        for r, s in enumerate(get_splits(T, ndevices)):
            out += [r] * s
        return out
    if alg == "greedy":
        load = [0] * ndevices
        out = [0] * T
        for t, E in enumerate(Es):
            r = load.index(min(load))
            load[r] += int(E)
            out[t] = r
        return out
    if alg == "hardcode":
        return [0] + [1] * (T - 1)
    raise ValueError("sharder not found: %s" % alg)


def local_tables(device_indices, rank):
    """dlrm_s_pytorch.py:468 local_emb_indices."""
    return [i for i, r in enumerate(device_indices) if r == rank]


def tables_per_rank(device_indices, W):
    """dlrm_s_pytorch.py:465-467 n_emb_per_rank (uniform m_spa == D so one split per table)."""
    n = [0] * W
    for r in device_indices:
        n[r] += 1
    return n


# ----------------------------------------------------------------------------------------------
# index layouts
# ----------------------------------------------------------------------------------------------
def pack_batched(lS_o, lS_i):
    """dlrm_data_pytorch.py:748-753 (and 839-843, data_loader_terabyte.py:88-93): packed
    table-major layout for the batched op: indices int32 [sum N_t]; offsets int32 [T*B+1] =
    per-table offsets shifted by the running index total, plus the grand total at the end."""
    idx = np.concatenate([np.asarray(x).reshape(-1) for x in lS_i]).astype(np.int32)
    E = [0] + np.cumsum([np.asarray(x).size for x in lS_i]).tolist()
    off = np.concatenate(
        [np.asarray(o).astype(np.int64) + e for o, e in zip(lS_o, E[:-1])] + [np.array([E[-1]])]
    ).astype(np.int32)
    return idx, off


def unpack_batched(indices, offsets, T, B):
    """Inverse of pack_batched: per-table (offsets [B] int64, indices [N_t] int64)."""
    lS_o, lS_i = [], []
    for t in range(T):
        o = offsets[t * B:(t + 1) * B + 1].astype(np.int64)
        lS_i.append(indices[o[0]:o[-1]].astype(np.int64))
        lS_o.append(o[:-1] - o[0])
    return lS_o, lS_i


def distribute_batched_local(indices, offsets, T, B, local_idx):
    """dlrm_s_pytorch.py:772-802, distributed branch: select the local tables of a packed
    batch and re-base the offsets so they index the concatenated local index stream."""
    o_parts, i_parts, run = [], [], 0
    for k in local_idx:
        o = offsets[k * B:(k + 1) * B + 1].astype(np.int64)
        i_parts.append(indices[o[0]:o[-1]])
        o = o - o[0]
        o_parts.append(o + run if not o_parts else (o[1:] + run))
        run += int(o[-1])
    return np.concatenate(i_parts).astype(np.int32), np.concatenate(o_parts).astype(np.int32)


def bag_bounds(off, n_idx, B):
    """nn.EmbeddingBag offsets semantics (call site dlrm_s_pytorch.py:571-577): bag b covers
    [off[b], off[b+1]) and the LAST bag runs to len(indices); empty bags are allowed."""
    off = np.asarray(off).astype(np.int64)
    return off, np.concatenate([off[1:], np.array([n_idx], dtype=np.int64)])


# ----------------------------------------------------------------------------------------------
# a2 / a3: embedding-bag forward (sum pooling)
# ----------------------------------------------------------------------------------------------
def embbag_fwd_table(W, idx, off, per_sample_weights=None):
    """nn.EmbeddingBag(mode='sum') as used by apply_emb (dlrm_s_pytorch.py:571-577):
    out[b,:] = sum_{l in bag b} W[idx[l],:] (* psw[l]); sequential fp32 adds in index order."""
    W = np.asarray(W, dtype=f32)
    idx = np.asarray(idx).astype(np.int64)
    B = len(off)
    s, e = bag_bounds(off, idx.size, B)
    out = np.zeros((B, W.shape[1]), dtype=f32)
    for b in range(B):
        acc = np.zeros(W.shape[1], dtype=f32)
        for l in range(s[b], e[b]):
            row = W[idx[l]]
            if per_sample_weights is not None:
                row = row * f32(per_sample_weights[l])
            acc = (acc + row).astype(f32)
        out[b] = acc
    return out


def apply_emb(Ws, lS_o, lS_i):
    """DLRM_Net.apply_emb, dlrm_s_pytorch.py:526-587 (no weighted pooling, no quantisation):
    list of T pooled tensors [B, D]."""
    return [embbag_fwd_table(W, i, o) for W, o, i in zip(Ws, lS_o, lS_i)]


def embbag_fwd_batched(Ws, indices, offsets, B):
    """Batched module call emb(indices, offsets) -> [B, T, D] (dlrm_s_pytorch.py:589-591,
    321-334; packed layout of pack_batched)."""
    T = len(Ws)
    lS_o, lS_i = unpack_batched(np.asarray(indices), np.asarray(offsets), T, B)
    ly = apply_emb(Ws, lS_o, lS_i)
    return np.stack(ly, axis=1)


# ----------------------------------------------------------------------------------------------
# a4-a6: backward = sort by row + segmented reduce + optimizer
# ----------------------------------------------------------------------------------------------
def sort_artifacts(indices, offsets, T, B, table_row_base):
    """The integer contract of the sorted backward.
    key[p]  = table_row_base[t(p)] + indices[p]     (global row id, table-major)
    bag[p]  = t(p)*B + b(p)                          (which pooled row fed position p)
    sorted by key, STABLE in original position p (== torch.sort(stable=True));
    unique = np.unique(key) ; run_off[u] = first sorted position of unique[u], run_off[U] = N.
    This is what grad.coalesce() (optim/rwsadagrad.py:98) computes implicitly."""
    indices = np.asarray(indices).astype(np.int64)
    offsets = np.asarray(offsets).astype(np.int64)
    N = int(offsets[T * B])
    lens = np.diff(offsets[:T * B + 1])
    bag = np.repeat(np.arange(T * B, dtype=np.int64), lens)
    base = np.asarray(table_row_base, dtype=np.int64)
    key = indices[:N] + base[bag // B]
    order = np.argsort(key, kind="stable")
    skey, sbag = key[order], bag[order]
    if N:
        head = np.concatenate([[True], skey[1:] != skey[:-1]])
    else:
        head = np.zeros(0, dtype=bool)
    run_off = np.concatenate([np.nonzero(head)[0], [N]]).astype(np.int64)
    return dict(key=key, bag=bag, sorted_key=skey, sorted_bag=sbag,
                unique=skey[head], run_off=run_off)


def embbag_bwd_coalesced(grad_out, art, D):
    """a4 + coalesce (optim/rwsadagrad.py:98): per unique row, the sum of grad_out rows of every
    bag that looked it up; accumulated sequentially in stable sorted order, fp32.
    grad_out: [B, T, D] (batched layout) flattened to rows by bag id t*B+b via go_rows()."""
    U = art["unique"].size
    g = np.zeros((U, D), dtype=f32)
    for u in range(U):
        acc = np.zeros(D, dtype=f32)
        for p in range(art["run_off"][u], art["run_off"][u + 1]):
            acc = (acc + grad_out[art["sorted_bag"][p]]).astype(f32)
        g[u] = acc
    return g


def go_rows(grad_btd):
    """[B, T, D] -> [T*B, D] rows addressed by bag id t*B+b."""
    B, T, D = grad_btd.shape
    return np.ascontiguousarray(np.transpose(grad_btd, (1, 0, 2))).reshape(T * B, D)


def sgd_update(Wcat, unique, g, lr):
    """torch.optim.SGD on a sparse grad (dlrm_s_pytorch.py:1640, 1933): W[r] += -lr * g[r]."""
    Wcat[unique] = (Wcat[unique] + f32(-lr) * g).astype(f32)


def rwsadagrad_update(Wcat, mom, unique, g, lr, eps=1e-10, step=1, lr_decay=0.0):
    """RWSAdagrad.step sparse branch, optim/rwsadagrad.py:95-113:
    clr = lr / (1 + (step-1)*lr_decay); mom[r] += mean_d(g^2); std = sqrt(mom[r]) + eps;
    W[r,:] += -clr * g[r,:] / std."""
    clr = lr / (1.0 + (step - 1.0) * lr_decay)
    msq = np.mean((g * g).astype(f32), axis=1, dtype=f32)
    mom[unique] = (mom[unique] + msq).astype(f32)
    std = (np.sqrt(mom[unique]).astype(f32) + f32(eps)).astype(f32)
    Wcat[unique] = (Wcat[unique] + f32(-clr) * (g / std[:, None]).astype(f32)).astype(f32)


# ----------------------------------------------------------------------------------------------
# a7: feature interaction
# ----------------------------------------------------------------------------------------------
def tril_pairs(F, itself=False):
    """torch.tril_indices(F, F, offset) row-major order, dlrm_s_pytorch.py:646-650."""
    off = 0 if itself else -1
    li = [i for i in range(F) for j in range(i + 1 + off)]
    lj = [j for i in range(F) for j in range(i + 1 + off)]
    return np.array(li, dtype=np.int64), np.array(lj, dtype=np.int64)


def interact_dot(x, emb_btd, itself=False):
    """interact_features('dot'), dlrm_s_pytorch.py:627-659: T=[x;E] [B,F,D]; Z=T T^t;
    R = [x | Z[:, li, lj]]."""
    x = np.asarray(x, dtype=f32)
    Tm = np.concatenate([x[:, None, :], np.asarray(emb_btd, dtype=f32)], axis=1)
    Z = np.einsum("bid,bjd->bij", Tm.astype(np.float64), Tm.astype(np.float64))
    li, lj = tril_pairs(Tm.shape[1], itself)
    return np.concatenate([x, Z[:, li, lj].astype(f32)], axis=1)


def interact_dot_bwd(x, emb_btd, dR, itself=False):
    """Autograd of interact_dot: dT = (dZ + dZ^t) T with dZ scattered from dR[:, D:], plus the
    pass-through dR[:, :D] on the dense row.  Returns (dx [B,D], demb [B,T,D])."""
    x = np.asarray(x, dtype=f32)
    B, D = x.shape
    Tm = np.concatenate([x[:, None, :], np.asarray(emb_btd, dtype=f32)], axis=1).astype(np.float64)
    F = Tm.shape[1]
    li, lj = tril_pairs(F, itself)
    dZ = np.zeros((B, F, F), dtype=np.float64)
    dZ[:, li, lj] = dR[:, D:]
    dT = np.einsum("bij,bjd->bid", dZ + np.transpose(dZ, (0, 2, 1)), Tm)
    dx = dT[:, 0, :] + dR[:, :D]
    return dx.astype(f32), dT[:, 1:, :].astype(f32)


def interact_cat(x, emb_btd):
    """interact_features('cat'), dlrm_s_pytorch.py:660-665: pure layout, [B, F, D]."""
    return np.concatenate([np.asarray(x, f32)[:, None, :], np.asarray(emb_btd, f32)], axis=1)


# ----------------------------------------------------------------------------------------------
# a9: all-to-all permutation (extend_distributed.py:405-508)
# ----------------------------------------------------------------------------------------------
def alltoall_fwd(per_rank_pooled, W):
    """per_rank_pooled[r] : [B, T_r, D] (rank r's local tables, full global batch).
    Returns per-rank outputs [B/W, T_global, D] with features in RANK-MAJOR table order
    (All2All_Req.forward 424-434 + All2All_Wait.forward 482-487 + the cat at 629-637)."""
    B = per_rank_pooled[0].shape[0]
    assert B % W == 0
    lb = B // W
    return [np.concatenate([per_rank_pooled[p][r * lb:(r + 1) * lb] for p in range(W)], axis=1)
            for r in range(W)]


def alltoall_bwd(per_rank_grad, tables_per_rank_, W):
    """Transpose of alltoall_fwd (All2All_Wait.backward 490-508, All2All_Req.backward 447-462):
    per_rank_grad[r] : [B/W, T_global, D] -> per-rank [B, T_r, D]."""
    starts = np.concatenate([[0], np.cumsum(tables_per_rank_)])
    return [np.concatenate([per_rank_grad[p][:, starts[r]:starts[r + 1]] for p in range(W)], axis=0)
            for r in range(W)]


# ----------------------------------------------------------------------------------------------
# a8: MLP towers + loss (used only to pin whole-step results on tiny configs)
# ----------------------------------------------------------------------------------------------
def mlp_fwd(x, layers, sigmoid_layer=-1):
    """create_mlp/apply_mlp, dlrm_s_pytorch.py:227-265, 518-524: Linear + ReLU, Sigmoid at
    `sigmoid_layer`.  layers = [(W [m,n], b [m]), ...].  Returns (out, activations)."""
    acts = [np.asarray(x, f32)]
    for i, (Wl, bl) in enumerate(layers):
        z = (acts[-1].astype(np.float64) @ Wl.T.astype(np.float64) + bl).astype(f32)
        if i == sigmoid_layer:
            a = (1.0 / (1.0 + np.exp(-z.astype(np.float64)))).astype(f32)
        else:
            a = np.maximum(z, f32(0))
        acts.append(a)
    return acts[-1], acts


def mlp_bwd(dout, layers, acts, sigmoid_layer=-1):
    """Backward of mlp_fwd.  Returns (dx, [(dW, db), ...])."""
    grads = [None] * len(layers)
    d = dout.astype(np.float64)
    for i in reversed(range(len(layers))):
        a = acts[i + 1].astype(np.float64)
        d = d * a * (1 - a) if i == sigmoid_layer else d * (a > 0)
        grads[i] = ((d.T @ acts[i].astype(np.float64)).astype(f32), d.sum(0).astype(f32))
        d = d @ layers[i][0].astype(np.float64)
    return d.astype(f32), grads


def bce_loss(p, t):
    """torch.nn.BCELoss(reduction='mean'), dlrm_s_pytorch.py:170-178, 506.  Returns (loss, dp)."""
    p64, t64 = p.astype(np.float64), t.astype(np.float64)
    n = p.size
    loss = -np.mean(t64 * np.maximum(np.log(p64), -100) + (1 - t64) * np.maximum(np.log(1 - p64), -100))
    dp = (p64 - t64) / np.maximum(p64 * (1 - p64), 1e-12) / n
    return f32(loss), dp.astype(f32)


def criteo_bin_transform(records, max_ind_range=-1, batched=False):
    """CriteoBinDataset.__getitem__ + _transform_features (data_loader_terabyte.py:236-249, 68-95) on one batch of
    raw records [B, 40] int32 = [label | 13 dense | 26 categorical]:
      x_cat % max_ind_range when max_ind_range > 0 (torch int32 `%`: floor modulo), X = log(float32(dense) + 1),
      lS_i = x_cat.t() as int64 [26, B], lS_o = arange(B) per table, T = float32(label) [B, 1];
      batched: indices = lS_i flattened table-major as int32, offsets = arange(26*B + 1) int32 (88-93)."""
    rec = np.asarray(records, dtype=np.int32).reshape(-1, 40)
    B = rec.shape[0]
    y = rec[:, 0].astype(np.float32).reshape(-1, 1)
    x_int = np.log(rec[:, 1:14].astype(np.float32) + np.float32(1.0)).astype(np.float32)
    x_cat = rec[:, 14:]
    if max_ind_range > 0:
        x_cat = np.mod(x_cat, np.int32(max_ind_range))          # floor modulo, sign of the divisor
    lS_i = np.ascontiguousarray(x_cat.T).astype(np.int64)
    lS_o = np.tile(np.arange(B, dtype=np.int64), (26, 1))
    if batched:
        indices = lS_i.reshape(-1).astype(np.int32)
        offsets = np.arange(26 * B + 1, dtype=np.int32)
        return x_int, offsets, indices, y
    return x_int, lS_o, lS_i, y


def embbag_bwd_weighted(W, indices, offsets, psw, grad_out):
    """Backward of EmbeddingBag(mode="sum") with per_sample_weights for one table, as autograd delivers it to the
    reference (dlrm_s_pytorch.py:544-547, 571-577):
      dpsw[l]  = <grad_out[bag(l)], W[indices[l]]>
      sparse table gradient, coalesced: unique rows (ascending) and, per row, the sum over its lookups in original
      order of  psw[l] * grad_out[bag(l)]  (product rounded to fp32, then added)."""
    indices = np.asarray(indices, dtype=np.int64)
    offsets = np.asarray(offsets, dtype=np.int64)
    N, B = indices.size, offsets.size
    ends = np.concatenate([offsets[1:], [N]])
    bag = np.repeat(np.arange(B), ends - offsets)
    dpsw = np.zeros(N, dtype=f32)
    for l in range(N):
        dpsw[l] = np.dot(grad_out[bag[l]].astype(np.float64), W[indices[l]].astype(np.float64))
    uniq = np.unique(indices)
    pos_of = {int(r): u for u, r in enumerate(uniq)}
    g = np.zeros((uniq.size, W.shape[1]), dtype=f32)
    for l in range(N):
        u = pos_of[int(indices[l])]
        g[u] = (g[u] + (f32(psw[l]) * grad_out[bag[l]]).astype(f32)).astype(f32)
    return dpsw, uniq, g
