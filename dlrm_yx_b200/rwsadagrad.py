"""Row-wise sparse Adagrad and sparse SGD on the GPU, with the reference's optimizer surface.

`RWSAdagrad(params, lr=1e-2, lr_decay=0.0, weight_decay=0.0, initial_accumulator_value=0.0, eps=1e-10)`
mirrors optim/rwsadagrad.py:11-122.  The reference refuses to run it on a GPU
(dlrm_s_pytorch.py:1636-1637, and its accumulator is created on the CPU, rwsadagrad.py:72-78); here the
sparse branch runs in libdlrmx.so:

  * a parameter whose `.grad` is a sparse COO tensor (e.g. from `EmbeddingBagSum`, already coalesced; or
    from torch's own nn.EmbeddingBag, coalesced here) is stepped by `dlrmx_rows_update`:
        mom[r] += mean_d(g[r,d]^2);  W[r,:] += -clr * g[r,:] / (sqrt(mom[r]) + eps)     (rwsadagrad.py:109-113)
  * a module attached with `attach(module)` is updated INSIDE its backward by the fused
    sort -> segmented-reduce -> update kernel (no COO gradient at all); `step()` then only advances the
    step counter of that parameter;
  * dense parameters (the MLPs) take the plain Adagrad branch (rwsadagrad.py:115-118).
`state[p]['momentum']` keeps the reference's name/shape ([rows] fp32) so checkpoints stay exchangeable
(SURVEY.md section 5).
"""
import torch
from torch.optim import Optimizer

from . import _lib, ops


class _FusedHook:
    """What an attached embedding module asks its optimizer for during backward."""

    def __init__(self, opt, kind):
        self.opt, self.kind = opt, kind

    def momentum(self, p):
        return self.opt._momentum(p) if self.kind == _lib.OPT_RWSADAGRAD else None

    def hyper(self, p):
        return self.opt._fused_hyper(p)


class RWSAdagrad(Optimizer):
    def __init__(self, params, lr=1e-2, lr_decay=0.0, weight_decay=0.0, initial_accumulator_value=0.0, eps=1e-10):
        for name, v in (("learning rate", lr), ("lr_decay", lr_decay), ("weight_decay", weight_decay),
                        ("initial_accumulator_value", initial_accumulator_value), ("epsilon", eps)):
            if not 0.0 <= v:
                raise ValueError("Invalid %s value: %s" % (name, v))
        defaults = dict(lr=lr, lr_decay=lr_decay, eps=eps, weight_decay=weight_decay,
                        initial_accumulator_value=initial_accumulator_value)
        super().__init__(params, defaults)
        self._pending = set()      # fused parameters updated in backward since the last step()
        for group in self.param_groups:
            for p in group["params"]:
                self.state[p]["step"] = 0

    # -- fused path --------------------------------------------------------------------------------
    def attach(self, *modules):
        """Fuse this optimizer into the backward of EmbeddingBagSum modules (their weights must already
        be in one of this optimizer's param groups)."""
        for m in modules:
            self._group_of(m.weight)
            m._fused = _FusedHook(self, _lib.OPT_RWSADAGRAD)
            m._ts = None
        return self

    def _group_of(self, p):
        for group in self.param_groups:
            if any(q is p for q in group["params"]):
                return group
        raise ValueError("parameter is not managed by this optimizer")

    def _momentum(self, p):
        st = self.state[p]
        if "momentum" not in st:
            init = self._group_of(p)["initial_accumulator_value"]
            st["momentum"] = torch.full([p.shape[0]], init, dtype=torch.float32, device=p.device)
        elif st["momentum"].device != p.device:
            # the model moved after the optimizer gathered state (or state was restored onto another device): the
            # accumulated history follows the parameter, it is never re-initialised
            st["momentum"] = st["momentum"].to(p.device)
        return st["momentum"]

    def _fused_hyper(self, p):
        group = self._group_of(p)
        if group["weight_decay"] != 0:
            raise RuntimeError("weight_decay option is not compatible with sparse gradients")
        step = self.state[p]["step"] + 1           # step() has not run yet for this iteration
        clr = group["lr"] / (1.0 + (step - 1.0) * group["lr_decay"])
        self._pending.add(p)
        return _lib.OPT_RWSADAGRAD, clr, group["eps"]

    # -- step --------------------------------------------------------------------------------------
    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            for p in group["params"]:
                state = self.state[p]
                if p.grad is None:
                    if p in self._pending:
                        state["step"] += 1
                    continue
                grad = p.grad
                state["step"] += 1
                if group["weight_decay"] != 0:
                    if grad.is_sparse:
                        raise RuntimeError("weight_decay option is not compatible with sparse gradients")
                    grad = grad.add(p, alpha=group["weight_decay"])
                clr = group["lr"] / (1.0 + (state["step"] - 1.0) * group["lr_decay"])
                if grad.is_sparse:
                    grad = grad.coalesce()             # no-op for EmbeddingBagSum's gradients
                    rows = grad._indices()[0].contiguous()
                    vals = grad._values().contiguous()
                    if vals.numel() > 0:
                        ops.rows_update(_lib.OPT_RWSADAGRAD, p.data, self._momentum(p), rows, vals, clr,
                                        group["eps"])
                else:
                    if "sum" not in state:
                        state["sum"] = torch.full_like(p, group["initial_accumulator_value"])
                    state["sum"].addcmul_(grad, grad, value=1.0)
                    std = state["sum"].sqrt().add_(group["eps"])
                    p.addcdiv_(grad, std, value=-clr)
        self._pending.clear()
        return loss


class SparseSGD(torch.optim.SGD):
    """torch.optim.SGD (dlrm_s_pytorch.py:1640) whose sparse-gradient parameters are stepped by libdlrmx and
    to which EmbeddingBagSum modules can be attached for the fused backward.  Dense parameters go through
    torch.optim.SGD unchanged."""

    def attach(self, *modules):
        for m in modules:
            self._group_of(m.weight)
            m._fused = _FusedHook(self, _lib.OPT_SGD)
            m._ts = None
        return self

    _group_of = RWSAdagrad._group_of

    def _fused_hyper(self, p):
        group = self._group_of(p)
        if group["momentum"] != 0 or group["weight_decay"] != 0:
            raise RuntimeError("fused sparse SGD supports neither momentum nor weight_decay")
        return _lib.OPT_SGD, group["lr"], 0.0

    @torch.no_grad()
    def step(self, closure=None):
        sparse = []
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is not None and p.grad.is_sparse:
                    g = p.grad.coalesce()
                    if g._values().numel() > 0:
                        ops.rows_update(_lib.OPT_SGD, p.data, None, g._indices()[0].contiguous(),
                                        g._values().contiguous(), group["lr"], 0.0)
                    sparse.append((p, p.grad))
                    p.grad = None
        try:
            return super().step(closure)
        finally:
            for p, g in sparse:
                p.grad = g
