"""Multi-GPU parity of the model-parallel exchange.  Run under torchrun with N >= 2 GPUs:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/check_exchange.py

Every rank builds the SAME seeded model three times -- NCCL all-to-all (fp32 towers), peer-memory exchange (fp32
towers), peer-memory exchange with the bf16 towers -- trains two steps on its slice of one global batch and checks:
  * peer == NCCL bit for bit (losses, local tables, accumulators): only the transport differs;
  * (contiguous 'naive_chunk' sharding, so the rank-major feature order equals the single-process order) the
    rank-averaged loss and the first update of the local tables against the golden vectors of the unmodified
    single-process reference on the same global batch (first loss rtol 1e-5; with row-wise Adagrad also the
    first table update, 1e-4 -- see the comment at the check);
  * the bf16-tower variant tracks it (loss 1e-2), and its micro-batched form (2 windows) tracks the one-shot form.
Prints 'rank r: exchange parity OK' and exits 0, or raises."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dlrm_yx_b200 import data, dlrm_net, embedding, rwsadagrad  # noqa: E402
from dlrm_yx_b200 import extend_distributed as ext  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c4_terabyte_small_rws"
    ext.init_distributed(use_gpu=True, backend="nccl")
    W, rank = ext.my_size, ext.my_rank
    dev = torch.device("cuda", torch.cuda.current_device())
    torch.backends.cuda.matmul.allow_tf32 = False
    g = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    T, D, B = len(g["meta_ln_emb"]), int(g["meta_misc"][0]), int(g["meta_misc"][1])
    assert B % W == 0, "golden batch %d not divisible by %d ranks" % (B, W)
    lb = B // W
    lr = float(g["meta_lr"])
    sgd = str(g["meta_opt"]) == "sgd"
    kind = embedding.Optimizer.SGD if sgd else embedding.Optimizer.EXACT_ROWWISE_ADAGRAD

    def build(exchange, bf16, micro=1):
        np.random.seed(int(g["meta_misc"][4]))
        m = dlrm_net.DLRM_Net(D, g["meta_ln_emb"], g["meta_ln_bot"], g["meta_ln_top"], arch_interaction_op="dot",
                              sigmoid_bot=-1, sigmoid_top=len(g["meta_ln_top"]) - 2, batched_emb=True,
                              emb_optimizer=kind, emb_learning_rate=lr, fuse_sparse_arch=True, sharder="naive_chunk").to(dev)
        # the sharded constructor draws only the local tables from the RNG (as the reference's does), so the golden
        # initial weights are loaded explicitly
        with torch.no_grad():
            for w, k in zip(m._batched_module().split_weights(), m.local_emb_indices):
                w.copy_(torch.from_numpy(g["W0_%d" % k]))
            for pre, mlp in (("bot", m.bot_l), ("top", m.top_l)):
                lins = [l for l in mlp if isinstance(l, torch.nn.Linear)]
                for j, l in enumerate(lins):
                    l.weight.copy_(torch.from_numpy(g["%s_W_%d" % (pre, j)]))
                    l.bias.copy_(torch.from_numpy(g["%s_b_%d" % (pre, j)]))
        if exchange.startswith("peer"):
            m.enable_peer_exchange(lb, engine="copy" if exchange == "peer-copy" else "sm", mixed=bool(bf16 and micro == 1))
        opt = None
        if bf16:
            m.fuse_mlps(lr, "sgd" if sgd else "adagrad")
            if micro > 1:
                m.set_microbatches(micro)
        else:
            dense = list(m.bot_l.parameters()) + list(m.top_l.parameters())
            m.bot_l = ext.DDP(m.bot_l, device_ids=[dev.index])
            m.top_l = ext.DDP(m.top_l, device_ids=[dev.index])
            opt = torch.optim.SGD(dense, lr=lr) if sgd else rwsadagrad.RWSAdagrad(dense, lr=lr)
        return m, opt

    def run(m, opt):
        local = list(m.local_emb_indices)
        idx, off = data.pack([torch.from_numpy(g["off_%d" % k]) for k in local],
                             [torch.from_numpy(g["idx_%d" % k]) for k in local])
        idx, off = idx.to(dev), off.to(dev)
        sl = slice(rank * lb, (rank + 1) * lb)
        X, tgt = torch.from_numpy(g["X"][sl]).to(dev), torch.from_numpy(g["target"][sl]).to(dev)
        losses, tables = [], []
        for step in range(2):
            loss = m.loss_fn(m(X, off, idx).float(), tgt)
            if opt is not None:
                opt.zero_grad()
            loss.backward()
            if opt is not None:
                opt.step()
            else:
                m.step_mlps()
            lt = loss.detach().clone()
            dist.all_reduce(lt)
            losses.append(lt.item() / W)
            tables.append([w.clone() for w in m._batched_module().split_weights()])
        return losses, tables, local

    ln, tn, local = run(*build("nccl", False))
    lp, tp, _ = run(*build("peer", False))
    lc, tc, _ = run(*build("peer-copy", False))
    assert ln == lp == lc, (ln, lp, lc)
    for step in range(2):
        for a, b, c in zip(tn[step], tp[step], tc[step]):
            assert torch.equal(a, b), "peer exchange differs from NCCL exchange"
            assert torch.equal(a, c), "copy-engine exchange differs from NCCL exchange"
    np.testing.assert_allclose(ln[0], float(g["loss"]), rtol=1e-5)
    if not sgd:
        # Row-wise Adagrad normalises the step by the gradient's own scale, so the W-times larger embedding gradients of
        # the data-parallel run (each rank's loss is the mean over ITS lb samples; only the MLP gradients are averaged
        # by DDP -- the reference behaves the same, dlrm_s_pytorch.py:686-730) still give the single-process update to
        # first order; with plain SGD they do not, so only the first loss is comparable there.
        np.testing.assert_allclose(ln[1], float(g["loss2"]), rtol=1e-3)
        for j, k in enumerate(local):
            np.testing.assert_allclose(tp[0][j].cpu().numpy(), g["W1_%d" % k], rtol=1e-4, atol=1e-5)
    lb16, _, _ = run(*build("peer", True))
    np.testing.assert_allclose(lb16[0], float(g["loss"]), rtol=1e-2)
    np.testing.assert_allclose(lb16[1], ln[1], rtol=2e-2)
    # micro-batched exchange (2 windows of the local batch, transfers overlapping the towers) against the one-shot one:
    # same arithmetic per sample, only the GEMM batch split and the gradient accumulation order differ
    lm2, tm2, _ = run(*build("peer-copy", True, micro=2))
    np.testing.assert_allclose(lm2[0], lb16[0], rtol=2e-3)
    np.testing.assert_allclose(lm2[1], lb16[1], rtol=5e-3)
    dist.barrier()
    print("rank %d: exchange parity OK (%s, %d ranks, local tables %s)" % (rank, name, W, local), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
