"""Feature interaction with the reference's semantics (DLRM_Net.interact_features,
dlrm_s_pytorch.py:627-673), computed by libdlrmx.

'dot': the reference runs cat -> bmm(T, T^t) -> tril_indices gather -> cat (5 kernels, 3 intermediates);
here ONE kernel reads the [B, F, D] feature block and writes R = [x | lower-triangle(T T^t)] directly, and
ONE kernel computes the whole backward dT = (dZ + dZ^t) T (+ the pass-through of dR[:, :D] on feature 0).
'cat': pure layout.
"""
import torch

from . import _lib, ops


class DotInteraction(torch.autograd.Function):
    """feat [B, F, D] (feature 0 = bottom-MLP output) -> R [B, D + F(F-1)/2] (or F(F+1)/2 with itself)."""

    @staticmethod
    def forward(ctx, feat, itself):
        ctx.itself = bool(itself)
        ctx.save_for_backward(feat)
        return ops.interact_dot_fwd(feat, ctx.itself)

    @staticmethod
    def backward(ctx, dR):
        (feat,) = ctx.saved_tensors
        return ops.interact_dot_bwd(feat, dR.contiguous(), ctx.itself), None


def _feature_block(x, ly, batched):
    B, d = x.shape
    if batched:      # chunks [B, T_p, D] (or [B, T_p*D]) as returned by the batched op / the all-to-all
        parts = [x.view(B, 1, d)] + [l.view(B, -1, d) for l in ly]
        return torch.cat(parts, dim=1)
    return torch.cat([x] + list(ly), dim=1).view(B, -1, d)


def interact_features(x, ly, arch_interaction_op="dot", arch_interaction_itself=False, batched=False):
    """Same contract as DLRM_Net.interact_features: x [B, D], ly = list of pooled embeddings
    ([B, D] each, or [B, T_p, D] chunks when `batched`)."""
    if not x.is_cuda:
        raise _lib.DlrmxError("interact_features needs CUDA tensors (no CPU fallback)")
    feat = _feature_block(x, ly, batched)
    if arch_interaction_op == "dot":
        return DotInteraction.apply(feat.contiguous(), arch_interaction_itself)
    if arch_interaction_op == "cat":
        return feat
    raise ValueError("ERROR: --arch-interaction-op=" + str(arch_interaction_op) + " is not supported")
