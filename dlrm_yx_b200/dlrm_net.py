"""Host-side mirror of the reference model's hot-path interface (DLRM_Net, dlrm_s_pytorch.py:226-824) with
the sparse path running on libdlrmx: same method names, argument meaning and error behaviour for
`create_mlp / create_emb / create_emb_batched / apply_mlp / apply_emb / apply_emb_batched /
interact_features / forward / sequential_forward / distributed_forward / distribute_batched_emb_data`.

What is deliberately NOT here: `parallel_forward` (single-process multi-GPU, SURVEY.md 2.1 out of scope),
QR/MD embeddings, quantised inference, the fbgemm path.

Two ways through the sparse arch:
  * modular (default, reads like the reference): per-table `EmbeddingBagSum` or one
    `TableBatchedEmbeddingBags`, then `interact_features`;
  * `fuse_sparse_arch=True` with `batched_emb=True` and dot interaction: lookup writes pooled rows straight
    into the [B, 1+T, D] interaction input, one interaction kernel, and in backward one interaction kernel
    whose output gradient feeds the fused sort/reduce/update in place -- no cat / split / contiguous copies.
    In the multi-process form the exchange is completed by one unpack kernel into the same block.
"""
import sys

import numpy as np
import torch
from torch import nn
from torch.autograd.profiler import record_function

from . import _lib, embedding, interaction, ops, sharders
from . import extend_distributed as ext_dist


# ---- second stream for the gradient-independent / tower-independent halves of the sparse path -----------------
# forward : the radix sort of the looked-up rows (needed only by the backward) runs beside lookup, interaction and
#           top MLP;
# backward: the fused reduce + sparse-optimizer update runs beside the bottom-MLP backward and the tower optimizer.
# The main stream re-joins when autograd's backward pass ends (engine callback), so anything the caller does after
# loss.backward() is ordered after the update.  Works eagerly and under CUDA-graph capture (fork / join pattern).
_SIDE = {}
OVERLAP_STREAMS = True        # False: everything on the caller's stream (per-kernel timing wants kernels alone)


def _side_stream(device):
    if not OVERLAP_STREAMS:
        return torch.cuda.current_stream()
    s = _SIDE.get(device)
    if s is None:
        s = _SIDE[device] = torch.cuda.Stream(device=device)
    return s


def _sort_on_side(emb, indices, offsets, B):
    main = torch.cuda.current_stream()
    side = _side_stream(indices.device)
    side.wait_stream(main)
    with torch.cuda.stream(side):
        presorted = emb.sort_lookups(indices, offsets, B)
    return presorted


def _update_on_side(emb, grad, indices, offsets, B, stride_b, stride_t, presorted, before=None, keep=()):
    """Runs `before()` (optional, e.g. the gradient exchange) and the fused update on the side stream, after
    everything enqueued so far on the main stream; the main stream joins at the end of the backward pass.  `keep`:
    tensors the side-stream work reads -- held until the join so their memory cannot be handed out earlier."""
    main = torch.cuda.current_stream()
    side = _side_stream(grad.device)
    side.wait_stream(main)
    with torch.cuda.stream(side):
        if before is not None:
            grad = before()
        if grad is not None:
            emb.fused_backward(grad, indices, offsets, B, stride_b=stride_b, stride_t=stride_t, presorted=presorted)
    held = [grad, presorted, keep]

    if side is main or side == main:
        return

    def join():
        torch.cuda.current_stream().wait_stream(side)
        held.clear()
    torch.autograd.Variable._execution_engine.queue_callback(join)


class _BCELoss(torch.nn.BCELoss):
    """dlrm.loss_fn for --loss-function=bce (dlrm_s_pytorch.py:505-507): torch.nn.BCELoss(reduction="mean"), computed
    by one libdlrmx kernel each way when predictions and targets are fp32 CUDA tensors (ops.bce_mean); anything else
    (CPU tensors in host-side checks, other dtypes) takes torch's implementation."""

    def __init__(self):
        super().__init__(reduction="mean")

    def forward(self, input, target):
        if input.is_cuda and target.is_cuda and input.dtype == torch.float32 and target.dtype == torch.float32 \
                and input.shape == target.shape and input.numel() > 0:
            return ops.bce_mean(input, target)
        return super().forward(input, target)


class _SparseArchLocal(torch.autograd.Function):
    """Single-GPU fused lookup + dot interaction.  forward(x [B,D]) -> R [B, D + F(F-1)/2]."""

    @staticmethod
    def forward(ctx, x, weights, emb, indices, offsets, itself):
        B, D = x.shape
        T = emb.T
        feat = torch.empty((B, 1 + T, D), dtype=torch.float32, device=x.device)
        ops.embbag_forward(emb.table_set(), indices, offsets, B, out=feat[:, 1:, :],
                           stride_b=(1 + T) * D, stride_t=D)
        feat[:, 0, :].copy_(x)
        ctx.emb, ctx.itself = emb, bool(itself)
        ctx.presorted = _sort_on_side(emb, indices, offsets, B) if ctx.needs_input_grad[0] else None
        ctx.save_for_backward(feat, indices, offsets)
        return ops.interact_dot_fwd(feat, ctx.itself)

    @staticmethod
    def backward(ctx, dR):
        feat, indices, offsets = ctx.saved_tensors
        B, F, D = feat.shape
        dfeat = ops.interact_dot_bwd(feat, dR.contiguous(), ctx.itself)
        _update_on_side(ctx.emb, dfeat[:, 1:, :], indices, offsets, B, F * D, D, ctx.presorted, keep=(dfeat,))
        ctx.presorted = None
        return dfeat[:, 0, :], None, None, None, None, None


class _SparseArchLocalMixed(torch.autograd.Function):
    """_SparseArchLocal between bf16 towers: x arrives as bf16, R leaves as bf16 padded to `out_cols` (the top MLP's
    K rounded up to 8); pooled embeddings, their gradients and the sparse update stay fp32."""

    @staticmethod
    def forward(ctx, x, weights, emb, indices, offsets, itself, out_cols):
        B, D = x.shape
        T = emb.T
        # slot 0 stays unused; rows are pitched for the interaction kernels' one-bulk-copy staging (D + 8 when D % 32 == 0)
        feat = ops.alloc_features(B, 1 + T, D, x.device)
        ops.embbag_forward(emb.table_set(), indices, offsets, B, out=feat[:, 1:, :],
                           stride_b=feat.stride(0), stride_t=feat.stride(1))
        x = x.contiguous()
        ctx.emb, ctx.itself = emb, bool(itself)
        ctx.presorted = _sort_on_side(emb, indices, offsets, B) if ctx.needs_input_grad[0] else None
        ctx.save_for_backward(feat, x, indices, offsets)
        return ops.interact_dot_fwd_mixed(feat, x, out_cols, ctx.itself)

    @staticmethod
    def backward(ctx, dR):
        feat, x, indices, offsets = ctx.saved_tensors
        B, F, D = feat.shape
        dfeat, dx = ops.interact_dot_bwd_mixed(feat, x, dR.contiguous(), ctx.itself)
        _update_on_side(ctx.emb, dfeat[:, 1:, :], indices, offsets, B, F * D, D, ctx.presorted, keep=(dfeat,))
        ctx.presorted = None
        return dx, None, None, None, None, None, None


class _ExchangeInteractMixed(torch.autograd.Function):
    """_ExchangeInteract between bf16 towers."""

    @staticmethod
    def forward(ctx, x, recv, itself, out_cols):
        lb, D = x.shape
        T = ext_dist.myreq.a2a_info.global_table_num
        feat = torch.empty((lb, 1 + T, D), dtype=torch.float32, device=x.device)
        ctx.info = ext_dist.complete_into(recv, feat, 1)
        x = x.contiguous()
        ctx.itself = bool(itself)
        ctx.save_for_backward(feat, x)
        return ops.interact_dot_fwd_mixed(feat, x, out_cols, ctx.itself)

    @staticmethod
    def backward(ctx, dR):
        feat, x = ctx.saved_tensors
        dfeat, dx = ops.interact_dot_bwd_mixed(feat, x, dR.contiguous(), ctx.itself)
        send = ext_dist.start_backward_from(ctx.info, dfeat, 1)
        return dx, send, None, None


class _PeerStep(object):
    """Per-step state shared by the micro-batch branches of the peer-memory exchange."""

    def __init__(self, peer, emb, indices, offsets, presorted, nwin):
        self.peer, self.emb, self.indices, self.offsets = peer, emb, indices, offsets
        self.presorted, self.pending, self.first_push = presorted, nwin, True


class _PeerInteract(torch.autograd.Function):
    """Multi-process, exchange fused into the kernels over NVLink peer memory (ext_dist.PeerExchange): the pooled rows
    of window [off, off+ln) of the local batch were stored into this rank's interaction input by the peers' lookup
    kernels; backward pushes the window's gradient slices into the table owners' buffers and, once every window has
    been pushed, runs the fused update of the local tables.  out_cols > 0 selects the bf16-facing interaction kernels
    (x, R, dR, dx in bf16)."""

    @staticmethod
    def forward(ctx, x, step, off, ln, itself, out_cols):
        feat = step.peer.feat[off:off + ln]
        ctx.step, ctx.off, ctx.itself, ctx.mixed = step, off, bool(itself), out_cols > 0
        if ctx.mixed:
            x = x.contiguous()
            ctx.save_for_backward(feat, x)
            return ops.interact_dot_fwd_mixed(feat, x, out_cols, ctx.itself)
        feat[:, 0, :].copy_(x)
        ctx.save_for_backward(feat, x)
        return ops.interact_dot_fwd(feat, ctx.itself)

    @staticmethod
    def backward(ctx, dR):
        feat, x = ctx.saved_tensors
        step = ctx.step
        peer = step.peer
        if ctx.mixed:
            dfeat, dx = ops.interact_dot_bwd_mixed(feat, x, dR.contiguous(), ctx.itself)
        else:
            dfeat = ops.interact_dot_bwd(feat, dR.contiguous(), ctx.itself)
            dx = dfeat[:, 0, :]
        T_r = peer.grad.shape[1]
        first, off = step.first_push, ctx.off
        step.first_push = False
        step.pending -= 1
        last = step.pending == 0

        def exchange():                       # on the side stream: (barrier,) push over NVLink (, barrier)
            peer.push_grad(dfeat, mb_off=off, first=first)
            if not last:
                return None
            grad = peer.grad_ready()
            return grad if T_r > 0 else None
        _update_on_side(step.emb, dfeat, step.indices, step.offsets, peer.W * peer.lb, T_r * peer.D, peer.D,
                        step.presorted if last else None, before=exchange, keep=(dfeat,))
        if last:
            step.presorted = None
        return dx, None, None, None, None, None


class _ExchangeInteract(torch.autograd.Function):
    """Multi-process: completes the all-to-all into the feature block, then dot interaction.
    backward packs the block gradient and starts the transposed exchange (completed by the request's
    own backward, so it overlaps with the bottom-MLP backward exactly as in the reference)."""

    @staticmethod
    def forward(ctx, x, recv, itself):
        lb, D = x.shape
        T = ext_dist.myreq.a2a_info.global_table_num
        feat = torch.empty((lb, 1 + T, D), dtype=torch.float32, device=x.device)
        ctx.info = ext_dist.complete_into(recv, feat, 1)
        feat[:, 0, :].copy_(x)
        ctx.itself = bool(itself)
        ctx.save_for_backward(feat)
        return ops.interact_dot_fwd(feat, ctx.itself)

    @staticmethod
    def backward(ctx, dR):
        (feat,) = ctx.saved_tensors
        dfeat = ops.interact_dot_bwd(feat, dR.contiguous(), ctx.itself)
        send = ext_dist.start_backward_from(ctx.info, dfeat, 1)
        return dfeat[:, 0, :], send, None


class DLRM_Net(nn.Module):
    def create_mlp(self, ln, sigmoid_layer):
        """Linear + ReLU stack, Sigmoid after layer `sigmoid_layer`; weights N(0, sqrt(2/(m+n))), biases
        N(0, sqrt(1/m)) drawn from numpy's global RNG in layer order (dlrm_s_pytorch.py:227-265)."""
        layers = []
        for i in range(ln.size - 1):
            n, m = int(ln[i]), int(ln[i + 1])
            LL = nn.Linear(n, m, bias=True)
            W = np.random.normal(0.0, np.sqrt(2 / (m + n)), size=(m, n)).astype(np.float32)
            bt = np.random.normal(0.0, np.sqrt(1 / m), size=m).astype(np.float32)
            LL.weight.data = torch.tensor(W, requires_grad=True)
            LL.bias.data = torch.tensor(bt, requires_grad=True)
            layers.append(LL)
            layers.append(nn.Sigmoid() if i == sigmoid_layer else nn.ReLU())
        return nn.Sequential(*layers)

    def create_emb(self, m, ln, weighted_pooling=None):
        """One EmbeddingBagSum per (local) table, weights U(+-sqrt(1/n)) from numpy in table order
        (dlrm_s_pytorch.py:267-318)."""
        emb_l = nn.ModuleList()
        v_W_l = []
        for i in range(ln.size):
            if ext_dist.my_size > 1 and i not in self.local_emb_indices:
                continue
            n = int(ln[i])
            a = np.sqrt(1 / n)
            if self.qr_flag and n > self.qr_threshold:        # dlrm_s_pytorch.py:281-289
                from .tricks import QREmbeddingBag
                emb_l.append(QREmbeddingBag(n, m, self.qr_collisions, operation=self.qr_operation, mode="sum", sparse=True))
            elif self.md_flag and n > self.md_threshold:      # 290-299: narrow table + projection to the base dimension
                from .tricks import PrEmbeddingBag
                base = max(m)
                _m = int(m[i])
                EE = PrEmbeddingBag(n, _m, int(base))
                W = np.random.uniform(low=-a, high=a, size=(n, _m)).astype(np.float32)
                EE.embs.weight.data = torch.tensor(W)
                emb_l.append(EE)
            else:
                _m = int(max(m)) if self.md_flag else m
                W = np.random.uniform(low=-a, high=a, size=(n, _m)).astype(np.float32)
                emb_l.append(embedding.EmbeddingBagSum(n, _m, _weight=torch.tensor(W)))
            v_W_l.append(None if weighted_pooling is None else torch.ones(n, dtype=torch.float32))
        return emb_l, v_W_l

    def create_emb_batched(self, D, Es, weighted_pooling=None, learning_rate=0.1):
        """dlrm_s_pytorch.py:321-334: one table-batched module over the local tables, optimizer fused."""
        assert weighted_pooling is None, "Weighted pooling not supported yet!"
        self.Es = Es[self.local_emb_indices] if ext_dist.my_size > 1 else Es
        T = len(self.Es)
        return embedding.TableBatchedEmbeddingBags(
            T, self.Es, D, optimizer=self.emb_optimizer, learning_rate=learning_rate,
            eps=self.emb_eps, stochastic_rounding=False, init=self.emb_init), [None] * T

    def __init__(self, m_spa=None, ln_emb=None, ln_bot=None, ln_top=None, arch_interaction_op=None,
                 arch_interaction_itself=False, sigmoid_bot=-1, sigmoid_top=-1, sync_dense_params=True,
                 loss_threshold=0.0, ndevices=-1, weighted_pooling=None, loss_function="bce",
                 batched_emb=False, sharder="naive", allocation=None,
                 qr_flag=False, qr_operation="mult", qr_collisions=0, qr_threshold=200, md_flag=False, md_threshold=200,
                 # additions (keyword-only in spirit): the fused optimizer of the batched module
                 emb_optimizer=embedding.Optimizer.SGD, emb_learning_rate=0.1, emb_eps=None, emb_init="auto",
                 fuse_sparse_arch=False):
        super().__init__()
        if any(a is None for a in (m_spa, ln_emb, ln_bot, ln_top, arch_interaction_op)):
            return
        if ndevices > 1:
            sys.exit("single-process multi-GPU (parallel_forward) is not supported: launch one process per GPU")
        self.m_spa, self.ln_emb, self.ln_bot, self.ln_top = m_spa, np.asarray(ln_emb), np.asarray(ln_bot), np.asarray(ln_top)
        self.ndevices = ndevices
        self.arch_interaction_op = arch_interaction_op
        self.arch_interaction_itself = arch_interaction_itself
        self.sync_dense_params = sync_dense_params
        self.loss_threshold = loss_threshold
        self.loss_function = loss_function
        self.weighted_pooling = ("learned" if weighted_pooling not in (None, "fixed") else weighted_pooling)
        self.batched_emb = batched_emb
        self.fbgemm_emb = False
        # QR / mixed-dimension embedding tricks (dlrm_s_pytorch.py:425-433), per-table path only
        self.qr_flag, self.qr_collisions, self.qr_operation, self.qr_threshold = qr_flag, qr_collisions, qr_operation, qr_threshold
        self.md_flag, self.md_threshold = md_flag, md_threshold
        if (qr_flag or md_flag) and batched_emb:
            raise _lib.DlrmxError("QR / mixed-dimension embeddings use the per-table path (batched_emb=False), as in the reference")
        self.sharder = sharder
        self.emb_optimizer, self.emb_eps, self.emb_init = emb_optimizer, emb_eps, emb_init
        self.fuse_sparse_arch = bool(fuse_sparse_arch)
        if self.fuse_sparse_arch and not (batched_emb and arch_interaction_op == "dot"):
            raise ValueError("fuse_sparse_arch needs batched_emb=True and arch_interaction_op='dot'")
        self.device_indices = [0]

        if ext_dist.my_size > 1:       # table-wise model parallel (dlrm_s_pytorch.py:444-468)
            n_emb = len(self.ln_emb)
            if n_emb < ext_dist.my_size:
                sys.exit("only (%d) sparse features for (%d) devices, table partitions will fail"
                         % (n_emb, ext_dist.my_size))
            self.n_global_emb = n_emb
            if sharder == "input":
                self.device_indices = list(map(int, allocation.split(",")))
            else:
                self.device_indices = sharders.shard(self.ln_emb, ext_dist.my_size, sharder)
            self.n_emb_per_rank = [0] * ext_dist.my_size
            for i in range(n_emb):
                self.n_emb_per_rank[self.device_indices[i]] += m_spa // int(self.ln_bot[-1])
            self.local_emb_indices = [i for i, j in enumerate(self.device_indices) if j == ext_dist.my_local_rank]
        if batched_emb:
            self.emb_l, w_list = self.create_emb_batched(m_spa, self.ln_emb, weighted_pooling=weighted_pooling,
                                                         learning_rate=emb_learning_rate)
        else:
            self.emb_l, w_list = self.create_emb(m_spa, self.ln_emb, weighted_pooling=weighted_pooling)
        if self.weighted_pooling == "learned":
            self.v_W_l = nn.ParameterList([nn.Parameter(w) for w in w_list])
        else:
            self.v_W_l = w_list
        self.bot_l = self.create_mlp(self.ln_bot, sigmoid_bot)
        self.top_l = self.create_mlp(self.ln_top, sigmoid_top)
        if loss_function == "mse":
            self.loss_fn = torch.nn.MSELoss(reduction="mean")
        elif loss_function == "bce":
            self.loss_fn = _BCELoss()
        else:
            sys.exit("ERROR: --loss-function=" + loss_function + " is not supported")

    def apply_mlp(self, x, layers):
        return layers(x)

    def fuse_mlps(self, lr, optimizer="adagrad", eps=1e-10):
        """Replaces the two nn.Sequential towers by FusedMLP (bf16 tensor-core GEMMs with fused epilogues, flat
        parameter buffers, one-kernel optimizer step).  After this, call `step_mlps()` instead of putting the tower
        parameters into an outer optimizer."""
        from . import mlp
        self.bot_l = mlp.FusedMLP(self.bot_l, lr, optimizer, eps)
        self.top_l = mlp.FusedMLP(self.top_l, lr, optimizer, eps)
        # bf16 hand-over around the interaction (no cast / pad / slice passes): the bottom tower ends in a fused bf16
        # layer, the top tower starts with one, and the tensor-core interaction kernels take the shape
        F, D = 1 + int(self.ln_emb.size), int(self.ln_bot[-1])
        first, last = self.top_l.layers[0], self.bot_l.layers[-1]
        nz = ops.tri_count(F, self.arch_interaction_itself)
        self._mixed = bool(self.fuse_sparse_arch and self.arch_interaction_op == "dot" and first.fused and last.fused
                           and first.K == D + nz and ops.interact_mixed_supported(F, D))
        self.bot_l.out_bf16 = self._mixed
        return self

    def step_mlps(self):
        w = max(ext_dist.my_size, 1)
        self.bot_l.step(w)
        self.top_l.step(w)

    def apply_emb(self, lS_o, lS_i):
        """List of T pooled tensors [B, D] (dlrm_s_pytorch.py:526-587)."""
        ly = []
        for k, sparse_index_group_batch in enumerate(lS_i):
            sparse_offset_group_batch = lS_o[k]
            if self.v_W_l[k] is not None:
                if self.v_W_l[k].device != sparse_index_group_batch.device:     # "fixed" weights are plain tensors
                    self.v_W_l[k] = self.v_W_l[k].to(sparse_index_group_batch.device)
                per_sample_weights = self.v_W_l[k].gather(0, sparse_index_group_batch)
            else:
                per_sample_weights = None
            V = self.emb_l[k](sparse_index_group_batch, sparse_offset_group_batch,
                              per_sample_weights=per_sample_weights)
            ly.append(V)
        d = int(self.ln_bot[-1])
        out = []
        for y in ly:
            if y.shape[1] == d:
                out.append(y)
            else:
                out.extend(y.split(d, dim=1))
        return out

    def apply_emb_batched(self, lS_o, lS_i, device_ids=None):
        """lS_i / lS_o: one-element lists holding the packed indices / offsets (dlrm_s_pytorch.py:589-591;
        one module on one device here, so no parallel_apply threads)."""
        mods = self.emb_l if isinstance(self.emb_l, nn.ModuleList) else [self.emb_l]
        return [m(i, o) for m, i, o in zip(mods, lS_i, lS_o)]

    def interact_features(self, x, ly):
        return interaction.interact_features(x, ly, self.arch_interaction_op, self.arch_interaction_itself,
                                             batched=self.batched_emb)

    def forward(self, dense_x, lS_o, lS_i):
        if ext_dist.my_size > 1:
            return self.distributed_forward(dense_x, lS_o, lS_i)
        return self.sequential_forward(dense_x, lS_o, lS_i)

    def _clamp(self, p):
        if 0.0 < self.loss_threshold < 1.0:
            return torch.clamp(p, min=self.loss_threshold, max=(1.0 - self.loss_threshold))
        return p

    def _batched_module(self):
        return self.emb_l[0] if isinstance(self.emb_l, nn.ModuleList) else self.emb_l

    def sequential_forward(self, dense_x, lS_o, lS_i):
        if not isinstance(self.emb_l, nn.ModuleList):
            self.emb_l = nn.ModuleList([self.emb_l])
        with record_function("module::forward_pass::bottom_mlp"):
            x = self.apply_mlp(dense_x, self.bot_l)
        if self.fuse_sparse_arch:
            with record_function("module::forward_pass::embedding_lookup"):
                emb = self._batched_module()
                if getattr(self, "_mixed", False):
                    z = _SparseArchLocalMixed.apply(x, emb.embedding_weights, emb, lS_i, lS_o,
                                                    self.arch_interaction_itself, self.top_l.layers[0].K_pad)
                else:
                    z = _SparseArchLocal.apply(x, emb.embedding_weights, emb, lS_i, lS_o,
                                               self.arch_interaction_itself)
        else:
            with record_function("module::forward_pass::embedding_lookup"):
                if self.batched_emb:
                    ly = self.apply_emb_batched([lS_o], [lS_i])
                else:
                    ly = self.apply_emb(lS_o, lS_i)
            with record_function("module::forward_pass::interaction"):
                z = self.interact_features(x, ly)
        with record_function("module::forward_pass::top_mlp"):
            p = self.apply_mlp(z, self.top_l)
        return self._clamp(p)

    def enable_peer_exchange(self, local_batch, microbatches=1, engine="sm", mixed=None):
        """Replaces the NCCL all-to-all of distributed_forward by kernels that store straight into the peers'
        memory over NVLink (ext_dist.PeerExchange).  Needs fuse_sparse_arch + batched_emb and a fixed local batch.
        mixed=True (bf16 towers will be fused in: fuse_mlps) lays the interaction input out with the row pitch the
        bf16-facing interaction kernels stage with one bulk copy per sample."""
        if not (self.fuse_sparse_arch and self.batched_emb):
            raise _lib.DlrmxError("enable_peer_exchange needs batched_emb=True and fuse_sparse_arch=True")
        dev = next(self.parameters()).device
        D = int(self.ln_bot[-1])
        pitch = ops.interact_feature_pitch(1 + int(sum(self.n_emb_per_rank)), D) \
            if ((getattr(self, "_mixed", False) if mixed is None else mixed) and engine == "sm") else D
        self._peer = ext_dist.PeerExchange(self.n_emb_per_rank, local_batch, D, dev, engine=engine, feature_pitch=pitch)
        self.set_microbatches(microbatches)
        return self

    def set_microbatches(self, m):
        """Micro-batched exchange (peer exchange + FusedMLP towers only): the local batch is cut into m windows; the
        NVLink transfer of window k+1 (forward) / k-1 (backward) runs beside interaction + top tower of window k."""
        m = max(1, int(m))
        if m > 1 and not hasattr(self.top_l, "accumulate"):
            raise _lib.DlrmxError("micro-batched exchange needs the fused towers (call fuse_mlps first)")
        self._microbatches = m
        if hasattr(self.top_l, "accumulate"):
            self.top_l.accumulate = m > 1
        return self

    def _peer_forward(self, dense_x, lS_o, lS_i):
        peer = self._peer
        emb = self._batched_module()
        M = max(1, int(getattr(self, "_microbatches", 1)))
        lb = peer.lb
        cuts = [lb * m // M for m in range(M + 1)]
        windows = [(cuts[m], cuts[m + 1] - cuts[m]) for m in range(M)]
        main = torch.cuda.current_stream()
        side = _side_stream(dense_x.device)
        side.wait_stream(main)
        with record_function("module::forward_pass::embedding_lookup"):
            with torch.cuda.stream(side):       # lookup + NVLink scatter (and the backward's sort) beside the towers
                arrived = peer.lookup_and_scatter(emb.table_set(), lS_i, lS_o, windows if M > 1 else None)
                presorted = (emb.sort_lookups(lS_i, lS_o, peer.W * lb)
                             if torch.is_grad_enabled() and emb.T > 0 else None)
        with record_function("module::forward_pass::bottom_mlp"):
            x = self.apply_mlp(dense_x, self.bot_l)
        step = _PeerStep(peer, emb, lS_i, lS_o, presorted, M)
        cols = self.top_l.layers[0].K_pad if getattr(self, "_mixed", False) else 0
        outs = []
        for m, (off, ln) in enumerate(windows):
            # window m of the interaction input is complete; window m+1 is still on the wire while m is contracted
            with record_function("module::forward_pass::interaction"):
                main.wait_event(arrived[m])
                z = _PeerInteract.apply(x if M == 1 else x[off:off + ln], step, off, ln, self.arch_interaction_itself,
                                        cols)
            with record_function("module::forward_pass::top_mlp"):
                outs.append(self.apply_mlp(z, self.top_l))
        p = outs[0] if M == 1 else torch.cat(outs, dim=0)
        return self._clamp(p)

    def distributed_forward(self, dense_x, lS_o, lS_i):
        """Local tables x full global batch -> all-to-all (overlapped with the bottom MLP) -> local batch x
        all tables -> interaction -> top MLP (dlrm_s_pytorch.py:686-730)."""
        if not isinstance(self.emb_l, nn.ModuleList):
            self.emb_l = nn.ModuleList([self.emb_l])
        if getattr(self, "_peer", None) is not None:
            return self._peer_forward(dense_x, lS_o, lS_i)
        with record_function("module::forward_pass::embedding_lookup"):
            if self.batched_emb:
                ly = self.apply_emb_batched([lS_o], [lS_i])
            else:
                ly = self.apply_emb(lS_o, lS_i)
            if len(self.emb_l) != len(ly):
                sys.exit("ERROR: corrupted intermediate result in distributed_forward call")
            a2a_req = ext_dist.alltoall(ly, self.n_emb_per_rank, self.batched_emb)
        with record_function("module::forward_pass::bottom_mlp"):
            x = self.apply_mlp(dense_x, self.bot_l)
            if self.fuse_sparse_arch:
                recv = a2a_req.tensor[0]
            else:
                ly = list(a2a_req.wait())
        with record_function("module::forward_pass::interaction"):
            if self.fuse_sparse_arch:
                if getattr(self, "_mixed", False):
                    z = _ExchangeInteractMixed.apply(x, recv, self.arch_interaction_itself, self.top_l.layers[0].K_pad)
                else:
                    z = _ExchangeInteract.apply(x, recv, self.arch_interaction_itself)
            else:
                z = self.interact_features(x, ly)
        with record_function("module::forward_pass::top_mlp"):
            p = self.apply_mlp(z, self.top_l)
        return self._clamp(p)

    def distribute_batched_emb_data(self, batch_size, lS_o, lS_i):
        """Packed global batch (all T tables) -> this rank's packed batch (local tables, offsets re-based),
        dlrm_s_pytorch.py:772-802 (distributed branch).  Unlike the reference it does not assume a fixed
        number of indices per lookup: each table's index range comes from its own offsets."""
        B = batch_size
        if ext_dist.my_size <= 1:
            return [lS_o], [lS_i]
        o_parts, i_parts, run = [], [], 0
        bounds = lS_o[::B].tolist() if lS_o.numel() % B == 1 else None
        for k in self.local_emb_indices:
            o = lS_o[k * B:(k + 1) * B + 1]
            lo, hi = (bounds[k], bounds[k + 1]) if bounds else (int(o[0]), int(o[-1]))
            i_parts.append(lS_i[lo:hi])
            rel = o - lo + run
            o_parts.append(rel if not o_parts else rel[1:])
            run += hi - lo
        return [torch.cat(o_parts, dim=0)], [torch.cat(i_parts, dim=0)]
