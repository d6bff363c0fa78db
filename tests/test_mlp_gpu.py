"""FusedMLP (bf16 tensor-core towers) against the fp32 torch towers that create_mlp builds, and the fused dense
optimizer kernel against the reference formulas.  bf16 operands with fp32 accumulation: outputs / gradients are
compared as stated in the test docstring (the fp32 parity bars apply to the unfused towers, which the
golden-vector tests use).  The optimizer kernel itself is fp32 and held to 1e-6."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _need_gpu():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    torch.backends.cuda.matmul.allow_tf32 = False


def make_tower(ln, sigmoid_layer):
    from dlrm_yx_b200 import dlrm_net
    np.random.seed(3)
    return dlrm_net.DLRM_Net.create_mlp(None, np.array(ln), sigmoid_layer).cuda()


def scale_close(a, b, tol):
    """Relative Frobenius error.  (Max-abs is the wrong yardstick for gradients through ReLU: a pre-activation that
    rounds across zero in bf16 flips its gate and moves single gradient entries by a full |dy*w|.)"""
    a, b = a.float(), b.float()
    err, ref = (a - b).norm().item(), b.norm().item()
    assert err <= tol * ref + 1e-6, (err, ref)


def emulate(fused, x, go):
    """fp32 torch restatement of the tower WITH the bf16 roundings FusedMLP applies (operands and stored activations /
    gradients rounded to bf16, products accumulated in fp32): isolates kernel / plumbing errors from rounding."""
    bf = torch.bfloat16
    saved, h = [], x
    for i, L in enumerate(fused.layers):
        if L.fused:
            hb = h.to(bf).float()
            bias = fused.bias(i) if fused.gemm == "tc" else fused.bias(i).to(bf).float()     # tcgen05 epilogue: fp32 bias
            y = torch.relu(hb @ fused.weight(i).to(bf).float().t() + bias).to(bf).float()
            saved.append((hb, y))
            h = y
        else:
            hl = h.float().detach().requires_grad_()
            W, b = fused.weight(i).clone().requires_grad_(), fused.bias(i).clone().requires_grad_()
            h = torch.sigmoid(torch.nn.functional.linear(hl, W, b))
            saved.append((hl, W, b))
    out, grads, dy = h, [None] * len(fused.layers), go
    for i in reversed(range(len(fused.layers))):
        L = fused.layers[i]
        if L.fused:
            hb, y = saved[i]
            dz = dy.to(bf).float() * (y > 0)
            grads[i] = (dz.t() @ hb, dz.sum(0))
            dy = (dz @ fused.weight(i).to(bf).float()).to(bf).float()
        else:
            hl, W, b = saved[i]
            out.backward(dy)
            grads[i], dy = (W.grad, b.grad), hl.grad
    return out.detach(), dy, grads


@pytest.mark.parametrize("gemm", ["tc", "cublas"])
@pytest.mark.parametrize("ln,sig", [([13, 64, 32, 16], -1), ([479, 128, 64, 1], 2), ([16, 8], -1),
                                    ([13, 512, 256, 128], -1), ([479, 1024, 1024, 512, 256, 1], 4)])
def test_fused_mlp_matches_fp32_tower(ln, sig, gemm):
    """Two bars: (1) against the same-rounding fp32 emulation: 5e-3 relative Frobenius error everywhere (what is left
    is accumulation order and 1-ulp bf16 ties); (2) against the true fp32 tower create_mlp builds: forward 1e-2; the
    gradients 0.15 -- a pre-activation that bf16 rounding moves across zero flips its ReLU gate and changes single
    gradient entries by a full |dy*w|, so the fp32 gradient bar is a sanity bound, not a precision claim.  The
    five-layer 1024-wide Terabyte top tower gets 2e-2 on bar (1) for its gradients: accumulation-order differences
    flip a few gates per layer there too (every GEMM alone is held to 2e-3 by tools/gemm_check.py)."""
    from dlrm_yx_b200 import mlp
    ref = make_tower(ln, sig)
    fused = mlp.FusedMLP(make_tower(ln, sig), lr=0.1, optimizer="sgd", gemm=gemm)
    B = 257
    torch.manual_seed(0)
    x = torch.rand(B, ln[0], device="cuda")
    xr = x.clone().requires_grad_()
    xf = x.clone().requires_grad_()
    yr, yf = ref(xr), fused(xf)
    go = torch.randn_like(yr)
    ye, dxe, ge = emulate(fused, x, go)
    yr.backward(go)
    yf.backward(go)
    gtol = 2e-2 if len(ln) > 4 else 5e-3
    scale_close(yf, ye, 5e-3)
    scale_close(xf.grad, dxe, gtol)
    scale_close(yf, yr, 1e-2)
    scale_close(xf.grad, xr.grad, 0.15)
    lins = [m for m in ref if isinstance(m, torch.nn.Linear)]
    for i, lin in enumerate(lins):
        L = fused.layers[i]
        scale_close(L.gW[:, :L.K], ge[i][0], gtol)
        scale_close(L.gb, ge[i][1], gtol)
        scale_close(L.gW[:, :L.K], lin.weight.grad, 0.15)
        scale_close(L.gb, lin.bias.grad, 0.15)
        if L.K_pad != L.K:
            assert float(L.gW[:, L.K:].abs().max()) == 0.0          # padded input columns never get gradient
    # SGD step: master -= lr * grad, shadow refreshed, grads cleared
    w0 = fused.weight(0).clone()
    g0 = fused.layers[0].gW[:, :ln[0]].clone()
    fused.step()
    torch.testing.assert_close(fused.weight(0), w0 - 0.1 * g0, rtol=1e-6, atol=1e-7)
    assert float(fused.grad.abs().max()) == 0.0
    torch.testing.assert_close(fused.shadow.float(), fused.master.data.to(torch.bfloat16).float())


@pytest.mark.parametrize("B,K", [(4099, 256), (40000, 256), (5000, 520), (3000, 64)])
def test_head_linear_sigmoid_kernels(B, K):
    """Linear(K -> 1) + Sigmoid head against torch on the same bf16 activation (fp32 weights): forward 1e-5 relative,
    backward (dh bf16: 2^-8; gw, gb: 1e-4 of their scale -- different summation order over 4099 rows)."""
    from dlrm_yx_b200 import ops
    torch.manual_seed(4)
    h = torch.randn(B, K, device="cuda").to(torch.bfloat16)
    w = (torch.randn(K, device="cuda") * 0.1).requires_grad_()
    b = torch.randn(1, device="cuda").requires_grad_()
    p = ops.head_fwd(h, w.detach(), b.detach())
    hf = h.float().requires_grad_()
    ref = torch.sigmoid(hf @ w[:, None] + b)
    torch.testing.assert_close(p, ref, rtol=1e-5, atol=1e-6)
    dp = torch.randn(B, 1, device="cuda")
    ref.backward(dp)
    gw, gb = torch.empty(K + 1, device="cuda")[1:], torch.empty(1, device="cuda")     # gw only 4-byte aligned
    for _ in range(2):                                   # counters self-reset
        dh = ops.head_bwd(dp, p, h, w.detach(), gw, gb)
        assert dh.dtype == torch.bfloat16
        assert torch.all((dh.float() - hf.grad).abs() <= 2.0 ** -8 * hf.grad.abs() + 1e-7)
        torch.testing.assert_close(gw, w.grad, rtol=1e-4, atol=1e-4 * float(w.grad.abs().max()))
        torch.testing.assert_close(gb, b.grad, rtol=1e-4, atol=1e-4 * float(b.grad.abs().max()))
    # with the producing Linear + ReLU layer's backward fused in: dh leaves masked by (h > 0), gbp = its column sums
    hr = torch.relu(h.float()).to(torch.bfloat16)
    p2 = ops.head_fwd(hr, w.detach(), b.detach())
    gbp = torch.empty(K + 1, device="cuda")[1:]
    for _ in range(2):
        dz = ops.head_bwd(dp, p2, hr, w.detach(), gw, gb, gb_prev_relu=gbp)
        dzl = (dp * p2 * (1 - p2)) * w.detach()[None, :]
        ref_dz = torch.where(hr.float() > 0, dzl, torch.zeros_like(dzl))
        assert torch.all((dz.float() - ref_dz).abs() <= 2.0 ** -8 * ref_dz.abs() + 1e-7)
        assert bool(((dz.float() == 0) | (hr.float() > 0)).all())
        torch.testing.assert_close(gbp, dz.float().sum(0), rtol=1e-4, atol=1e-4 * float(dz.float().abs().sum(0).max()))


def test_dense_adagrad_kernel_matches_reference_formula():
    """optim/rwsadagrad.py:115-118: sum += g*g ; p -= clr * g / (sqrt(sum) + eps)."""
    from dlrm_yx_b200 import ops
    torch.manual_seed(1)
    n = 100003
    p = torch.randn(n, device="cuda")
    acc = torch.rand(n, device="cuda")
    g = torch.randn(n, device="cuda")
    shadow = torch.empty(n, dtype=torch.bfloat16, device="cuda")
    p0, a0, g0 = p.clone(), acc.clone(), g.clone()
    ops.dense_opt(1, p, acc, g, shadow, lr=0.05, eps=1e-10, grad_scale=0.5)
    gs = g0 * 0.5
    a1 = a0 + gs * gs
    p1 = p0 - 0.05 * gs / (a1.sqrt() + 1e-10)
    torch.testing.assert_close(acc, a1, rtol=1e-6, atol=1e-7)
    torch.testing.assert_close(p, p1, rtol=1e-6, atol=1e-7)
    assert float(g.abs().max()) == 0.0
    torch.testing.assert_close(shadow.float(), p.to(torch.bfloat16).float())


def test_relu_bwd_bias_grad_kernel():
    from dlrm_yx_b200 import ops
    torch.manual_seed(2)
    B, N = 1000, 264
    y = torch.relu(torch.randn(B, N, device="cuda")).to(torch.bfloat16)
    dy = torch.randn(B, N, device="cuda").to(torch.bfloat16)
    dz, part = ops.relu_bwd_bias_grad(dy, y, relu=True, chunks=7)
    want = dy.float() * (y.float() > 0)
    assert torch.equal(dz.float(), want)
    torch.testing.assert_close(part.sum(0), want.sum(0), rtol=1e-5, atol=1e-4)
    dz2, part2 = ops.relu_bwd_bias_grad(dy, y, relu=False, chunks=3)
    assert dz2 is dy
    torch.testing.assert_close(part2.sum(0), dy.float().sum(0), rtol=1e-5, atol=1e-4)
    # in-kernel second stage: deterministic, equal to the sum of the partials, counters self-reset (call it twice)
    for _ in range(2):
        db = torch.full((N,), float("nan"), device="cuda")
        dz3, part3 = ops.relu_bwd_bias_grad(dy, y, relu=True, db=db)
        assert torch.equal(dz3, dz)
        torch.testing.assert_close(db, want.sum(0), rtol=1e-5, atol=1e-4)
    db2 = torch.empty((N,), device="cuda")
    ops.relu_bwd_bias_grad(dy, y, relu=True, db=db2)
    assert torch.equal(db, db2)


@pytest.mark.parametrize("name", ["c4_terabyte_small_rws", "c2_kaggle_small_rws", "c3_multihot_sgd"])
def test_bf16_tower_model_tracks_the_reference_step(golden_dir, name):
    """The bench configuration (fused sparse arch + FusedMLP towers + bf16-facing interaction) against the golden
    vectors of the UNMODIFIED fp32 reference: first-step loss within 1e-2 relative, second-step loss (after one
    update of tables and towers) within 2e-2, and the first update of the embedding tables within 0.2 relative
    Frobenius error of the reference's update (bf16 activations / gradients, ReLU gate flips -- a tracking bound,
    not a parity claim; the fp32 parity bars are held by the unfused towers in test_model_gpu.py)."""
    import os
    from dlrm_yx_b200 import data, dlrm_net, embedding
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    T, D = len(g["meta_ln_emb"]), int(g["meta_misc"][0])
    lr = float(g["meta_lr"])
    sgd = str(g["meta_opt"]) == "sgd"
    np.random.seed(int(g["meta_misc"][4]))
    kind = embedding.Optimizer.SGD if sgd else embedding.Optimizer.EXACT_ROWWISE_ADAGRAD
    m = dlrm_net.DLRM_Net(D, g["meta_ln_emb"], g["meta_ln_bot"], g["meta_ln_top"], arch_interaction_op="dot",
                          sigmoid_bot=-1, sigmoid_top=len(g["meta_ln_top"]) - 2, batched_emb=True,
                          emb_optimizer=kind, emb_learning_rate=lr, fuse_sparse_arch=True).cuda()
    m.fuse_mlps(lr, "sgd" if sgd else "adagrad")
    idx, off = data.pack([torch.from_numpy(g["off_%d" % k]) for k in range(T)],
                         [torch.from_numpy(g["idx_%d" % k]) for k in range(T)])
    idx, off = idx.cuda(), off.cuda()
    X, tgt = torch.from_numpy(g["X"]).cuda(), torch.from_numpy(g["target"]).cuda()
    losses = []
    for step in range(2):
        loss = m.loss_fn(m(X, off, idx), tgt)
        loss.backward()
        m.step_mlps()
        losses.append(loss.item())
        if step == 0:
            num = den = 0.0
            for k, w in enumerate(m._batched_module().split_weights()):
                d_ref = g["W1_%d" % k].astype(np.float64) - g["W0_%d" % k]
                d_got = w.cpu().numpy().astype(np.float64) - g["W0_%d" % k]
                num += ((d_got - d_ref) ** 2).sum()
                den += (d_ref ** 2).sum()
            assert num ** 0.5 <= 0.2 * den ** 0.5, (num ** 0.5, den ** 0.5)
    np.testing.assert_allclose(losses[0], float(g["loss"]), rtol=1e-2)
    np.testing.assert_allclose(losses[1], float(g["loss2"]), rtol=2e-2)


def test_graphed_train_step_equals_eager_steps(golden_dir):
    """GraphedTrainStep (whole step captured as one CUDA graph) against the same steps launched eagerly: same seeded
    model, same batch sequence (the graph's 2 warm-up runs are real steps on the example batch).  Tables and
    accumulators within rtol 1e-4 / atol 1e-6 and losses within 1e-3 (the library GEMMs may pick a different algorithm
    under capture, which moves bf16 roundings; everything in libdlrmx is launch-mode independent)."""
    import os
    from dlrm_yx_b200 import data, dlrm_net, embedding
    from dlrm_yx_b200.graph import GraphedTrainStep
    g = np.load(os.path.join(golden_dir, "c4_terabyte_small_rws.npz"))
    T, D = len(g["meta_ln_emb"]), int(g["meta_misc"][0])
    lr = float(g["meta_lr"])

    def build():
        np.random.seed(int(g["meta_misc"][4]))
        m = dlrm_net.DLRM_Net(D, g["meta_ln_emb"], g["meta_ln_bot"], g["meta_ln_top"], arch_interaction_op="dot",
                              sigmoid_bot=-1, sigmoid_top=len(g["meta_ln_top"]) - 2, batched_emb=True,
                              emb_optimizer=embedding.Optimizer.EXACT_ROWWISE_ADAGRAD, emb_learning_rate=lr,
                              fuse_sparse_arch=True).cuda()
        return m.fuse_mlps(lr, "adagrad")

    idx, off = data.pack([torch.from_numpy(g["off_%d" % k]) for k in range(T)],
                         [torch.from_numpy(g["idx_%d" % k]) for k in range(T)])
    X, tgt = torch.from_numpy(g["X"]).cuda(), torch.from_numpy(g["target"]).cuda()
    b0 = (X, idx.cuda(), off.cuda(), tgt)
    gen = torch.Generator().manual_seed(5)
    rows = torch.from_numpy(np.repeat(g["meta_ln_emb"].astype(np.int64), X.shape[0]))
    b1 = (torch.rand(X.shape, generator=gen).cuda(),
          (torch.rand(idx.shape, generator=gen) * (rows - 1)).round().to(idx.dtype).cuda() if idx.numel() == rows.numel()
          else idx.cuda(), off.cuda(), (torch.rand(tgt.shape, generator=gen) > 0.5).float().cuda())

    def make_step(m):
        def step(X, idx, off, tgt):
            loss = m.loss_fn(m(X, off, idx), tgt)
            loss.backward()
            m.step_mlps()
            return loss
        return step

    me, mg = build(), build()
    se = make_step(me)
    eager_losses = [se(*b0).item(), se(*b0).item()]                 # == the graph's two warm-up steps
    gs = GraphedTrainStep(make_step(mg), b0, warmup=2)
    for b in (b1, b0, b1):
        le = se(*b).item()
        lg = gs(*b).item()
        np.testing.assert_allclose(lg, le, rtol=1e-3)
    for we, wg in zip(me._batched_module().split_weights(), mg._batched_module().split_weights()):
        np.testing.assert_allclose(wg.cpu().numpy(), we.cpu().numpy(), rtol=1e-4, atol=1e-6)
    with pytest.raises(ValueError):
        gs(X[:-1], b0[1], b0[2], tgt)                               # static shapes are enforced


@pytest.mark.parametrize("n", [1, 37, 4096, 32768, 100003])
def test_bce_mean_kernels_match_torch(n):
    """ops.bce_mean against torch.nn.BCELoss(reduction="mean") on the same fp32 inputs (saturated predictions included:
    the -100 clamp of the logs and the 1e-12 clamp of the gradient's denominator): loss 1e-5 relative (summation order), gradient 1e-6."""
    from dlrm_yx_b200 import ops
    torch.manual_seed(n)
    p = torch.rand(n, 1, device="cuda")
    p[::7] = 0.0
    p[3::11] = 1.0
    y = torch.round(torch.rand(n, 1, device="cuda"))
    p1 = p.clone().requires_grad_()
    p2 = p.clone().requires_grad_()
    for _ in range(2):                                   # the counter resets itself
        l1 = ops.bce_mean(p1, y)
    l2 = torch.nn.BCELoss(reduction="mean")(p2, y)
    torch.testing.assert_close(l1, l2, rtol=1e-5, atol=1e-6)
    (l1 * 3.0).backward()
    (l2 * 3.0).backward()
    torch.testing.assert_close(p1.grad, p2.grad, rtol=1e-6, atol=1e-12)
