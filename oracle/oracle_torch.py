"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's training step with the same PyTorch (ATen)
CPU operators the reference itself calls.  Parity pinned: tests/test_oracle_cpu.py checks this module against
the golden vectors the UNMODIFIED reference produced (tests/golden/*.npz, made by oracle/gen_golden.py).

The arithmetic of this path lives in a third-party dependency of the reference -- PyTorch, un-vendored and
only loosely pinned (requirements.txt:5 `torch`; README.md:371 nightly 11/10/20); this container has torch
2.11.0.  So "the reference's own CPU implementation" of the hot path IS: nn.EmbeddingBag(mode="sum",
sparse=True) forward/backward (dlrm_s_pytorch.py:301, 571-577), torch.bmm + tril_indices gather
(dlrm_s_pytorch.py:637-659), sparse `coalesce` / `add_` in the optimizers (optim/rwsadagrad.py:97-113;
dlrm_s_pytorch.py:1640).  This file strings those same calls together WITHOUT importing /root/reference, so
that it can run on the GPU box's host cores as the timed CPU baseline (bench.py `cpu_baseline` and
`--impl reference`, kind "port") and as the whole-step checker of smoke().

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module.
"""
import numpy as np
import torch
from torch import nn


def build_tables(rows, D, init="numpy", device=None):
    """create_emb (dlrm_s_pytorch.py:267-318): one nn.EmbeddingBag(n, D, mode='sum', sparse=True) per table,
    U(+-sqrt(1/n)) weights; numpy RNG in table order (parity) or torch RNG (large timing runs: the weight is drawn
    once, directly on `device`, instead of nn.EmbeddingBag's own normal_ fill followed by the overwrite)."""
    tables = nn.ModuleList()
    for n in rows:
        n = int(n)
        a = float(np.sqrt(1 / n))
        if init == "numpy":
            E = nn.EmbeddingBag(n, D, mode="sum", sparse=True)
            E.weight.data = torch.tensor(np.random.uniform(low=-a, high=a, size=(n, D)).astype(np.float32),
                                         requires_grad=True)
            if device is not None:
                E = E.to(device)
        else:
            w = torch.empty((n, D), dtype=torch.float32, device=device).uniform_(-a, a)
            E = nn.EmbeddingBag(n, D, mode="sum", sparse=True, _weight=w)
        tables.append(E)
    return tables


def build_mlp(ln, sigmoid_layer):
    """create_mlp (dlrm_s_pytorch.py:227-265)."""
    layers = []
    for i in range(len(ln) - 1):
        n, m = int(ln[i]), int(ln[i + 1])
        L = nn.Linear(n, m, bias=True)
        L.weight.data = torch.tensor(np.random.normal(0.0, np.sqrt(2 / (m + n)), size=(m, n)).astype(np.float32),
                                     requires_grad=True)
        L.bias.data = torch.tensor(np.random.normal(0.0, np.sqrt(1 / m), size=m).astype(np.float32),
                                   requires_grad=True)
        layers += [L, nn.Sigmoid() if i == sigmoid_layer else nn.ReLU()]
    return nn.Sequential(*layers)


def lookup(tables, lS_o, lS_i):
    """apply_emb (dlrm_s_pytorch.py:526-587) without weighted pooling: list of [B, D]."""
    return [E(idx, off) for E, off, idx in zip(tables, lS_o, lS_i)]


def interact(x, ly, op="dot", itself=False):
    """interact_features (dlrm_s_pytorch.py:627-673)."""
    B, d = x.shape
    T = torch.cat([x] + list(ly), dim=1).view(B, -1, d)
    if op == "cat":
        return T
    Z = torch.bmm(T, T.transpose(1, 2))
    li, lj = torch.tril_indices(Z.shape[1], Z.shape[2], offset=0 if itself else -1)
    return torch.cat([x, Z[:, li, lj]], dim=1)


def rwsadagrad_sparse_step(p, grad, mom, lr, eps=1e-10, step=1, lr_decay=0.0):
    """RWSAdagrad.step, sparse branch (optim/rwsadagrad.py:95-113)."""
    clr = lr / (1.0 + (step - 1.0) * lr_decay)
    g = grad.coalesce()
    rows, vals = g._indices(), g._values()
    if vals.numel() == 0:
        return
    mom_up = torch.sparse_coo_tensor(rows, vals.pow(2).mean(dim=1), [p.shape[0]])
    mom.add_(mom_up)
    std = mom.sparse_mask(mom_up.coalesce())._values().sqrt_().add_(eps)
    p.data.add_(torch.sparse_coo_tensor(rows, vals / std.view(-1, 1), p.shape), alpha=-clr)


def adagrad_dense_step(p, grad, acc, lr, eps=1e-10):
    """RWSAdagrad.step, dense branch (optim/rwsadagrad.py:115-118)."""
    acc.addcmul_(grad, grad, value=1.0)
    p.data.addcdiv_(grad, acc.sqrt().add_(eps), value=-lr)


class Model:
    """The pieces of DLRM_Net that lie on the hot path, in the reference's construction order (tables, then
    bottom MLP, then top MLP -- dlrm_s_pytorch.py:469-496) so that numpy-seeded weights match."""

    def __init__(self, rows, D, ln_bot, ln_top, init="numpy", interaction_op="dot", itself=False, device=None):
        self.tables = build_tables(rows, D, init, device)
        self.bot = build_mlp(ln_bot, -1)
        self.top = build_mlp(ln_top, len(ln_top) - 2)
        if device is not None:
            self.bot, self.top = self.bot.to(device), self.top.to(device)
        self.op, self.itself = interaction_op, itself
        self.loss_fn = nn.BCELoss(reduction="mean")
        self.state = {}
        self.steps = 0

    def params(self):
        return list(self.tables.parameters()) + list(self.bot.parameters()) + list(self.top.parameters())

    def forward(self, X, lS_o, lS_i):
        """sequential_forward (dlrm_s_pytorch.py:732-770)."""
        x = self.bot(X)
        ly = lookup(self.tables, lS_o, lS_i)
        z = interact(x, ly, self.op, self.itself)
        return self.top(z)

    def train_step(self, X, lS_o, lS_i, target, lr, optimizer="rwsadagrad", eps=1e-10):
        """One iteration of the train loop (dlrm_s_pytorch.py:1893-1934): forward, loss, backward, step."""
        for p in self.params():
            p.grad = None
        loss = self.loss_fn(self.forward(X, lS_o, lS_i), target)
        loss.backward()
        self.steps += 1
        self.apply_optimizer(lr, optimizer, eps)
        return float(loss.item())

    def apply_optimizer(self, lr, optimizer="rwsadagrad", eps=1e-10):
        """The optimizer half of the iteration: torch.optim.SGD (dlrm_s_pytorch.py:1640) or RWSAdagrad.step
        (optim/rwsadagrad.py:56-122) on the gradients the backward pass left."""
        with torch.no_grad():
            for p in self.params():
                if p.grad is None:
                    continue
                if optimizer == "sgd":                      # torch.optim.SGD: p += -lr * grad (sparse or dense)
                    p.data.add_(p.grad, alpha=-lr)
                elif p.grad.is_sparse:
                    if p not in self.state:
                        self.state[p] = torch.zeros(p.shape[0], device=p.device)
                    rwsadagrad_sparse_step(p, p.grad, self.state[p], lr, eps, self.steps)
                else:
                    if p not in self.state:
                        self.state[p] = torch.zeros_like(p)
                    adagrad_dense_step(p, p.grad, self.state[p], lr, eps)
