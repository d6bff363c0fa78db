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
FLUSH_EARLY = False           # single-GPU path: True launches the top tower's deferred wgrad right behind the interaction backward


def _side_stream(device):
    if not OVERLAP_STREAMS:
        return torch.cuda.current_stream()
    s = _SIDE.get(device)
    if s is None:
        s = _SIDE[device] = torch.cuda.Stream(device=device)
    return s


_COMM = {}

# Scheduling of the multi-process step around the exchange: a bit mask.
#   0 (default): the schedule that has run on 2, 4 and 8 GPUs -- ONE communication stream of default priority for every
#      gradient all-reduce chain, the replicated tables' lookup and sort on the main stream behind the bottom tower, the
#      gradient push as a looping grid of 16 CTAs per SM.
#   1: every all-reduce chain on its own HIGH-priority stream;  2: the gradient push as short-lived CTAs;
#   4: the replicated tables' sort on the side stream;  8: their lookup on a third stream beside the scatter and the
#      bottom tower.  15 = all four, measured on 2 GPUs (1.498 -> 1.418 ms per step with the bf16 gradient wire, parity
#      checks green); the one 8-GPU run of it did NOT finish within its time limit (no error, no output: see DESIGN.md
#      section 5) and there was no GPU time left to find out why -- hence opt-in, and hence separate bits, to bisect.
MGPU_OVERLAP = 0
MGPU_COMM_STREAMS, MGPU_SHORT_PUSH, MGPU_REP_SORT_SIDE, MGPU_REP_LOOKUP_STREAM = 1, 2, 4, 8


def set_mgpu_overlap(mask):
    """Selects the schedule above (before the first step / graph capture) and the matching gradient-push grid."""
    global MGPU_OVERLAP
    MGPU_OVERLAP = int(mask) & 15
    _lib.set_tunable(13, 1 if MGPU_OVERLAP & MGPU_SHORT_PUSH else 0)          # DLRMX_TUNE_PUSH_GRID (csrc/a2a_layout.cu)


def _comm_stream(device, which=0):
    """Stream(s) of the gradient all-reduces that run beside compute.  MGPU_COMM_STREAMS: 0 = the towers', 1 = the replicated
    tables' (each chain is barrier -> sum -> barrier -> consumer, and the two do not depend on each other), both HIGH
    priority: their kernels are small and latency-bound (device barriers, NVLink round trips) and must not queue behind the
    grids of the gradient push / sparse update they run beside."""
    if not OVERLAP_STREAMS:
        return torch.cuda.current_stream()
    split = bool(MGPU_OVERLAP & MGPU_COMM_STREAMS)
    key = (device, which, 1) if split else (device, 0, 0)
    s = _COMM.get(key)
    if s is None:
        s = _COMM[key] = torch.cuda.Stream(device=device, priority=-1 if split else 0)
    return s


def _sort_on_side(emb, indices, offsets, B):
    main = torch.cuda.current_stream()
    side = _side_stream(indices.device)
    side.wait_stream(main)
    with torch.cuda.stream(side):
        presorted = emb.sort_lookups(indices, offsets, B)
    return presorted


def _update_on_side(emb, grad, indices, offsets, B, stride_b, stride_t, presorted, before=None, keep=(), after=None):
    """Runs `before()` (optional, e.g. the gradient exchange) and the fused update on the side stream, after
    everything enqueued so far on the main stream (or after the event `after`, recorded on it earlier); the main stream
    joins at the end of the backward pass.  `keep`: tensors the side-stream work reads -- held until the join so their
    memory cannot be handed out earlier."""
    main = torch.cuda.current_stream()
    side = _side_stream(grad.device)
    if after is not None and side != main:
        side.wait_event(after)
    else:
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


def _flush_deferred_wgrads():
    """The fused top tower's deferred weight-gradient kernel goes out here, behind the interaction backward
    (mlp.flush_deferred).  Returns the event that marks the interaction backward's end on this stream: the sparse update
    forks from THERE, so it runs beside the weight-gradient kernel, not behind it."""
    from . import mlp
    done = torch.cuda.current_stream().record_event()
    mlp.flush_deferred()
    return done


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
        ctx.presorted = _sort_on_side(emb, indices, offsets, B) if ctx.needs_input_grad[0] else None
        ops.embbag_forward(emb.table_set(), indices, offsets, B, out=feat[:, 1:, :],
                           stride_b=(1 + T) * D, stride_t=D)
        feat[:, 0, :].copy_(x)
        ctx.emb, ctx.itself = emb, bool(itself)
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
    def forward(ctx, x, weights, emb, indices, offsets, itself, out_cols, relu_x):
        B, D = x.shape
        T = emb.T
        ctx.relu_x = bool(relu_x)
        # slot 0 stays unused; rows are pitched for the interaction kernels' one-bulk-copy staging (D + 8 when D % 32 == 0)
        feat = ops.alloc_features(B, 1 + T, D, x.device)
        # the sort goes out first: forked here, it runs beside the lookup instead of beside the top tower's GEMMs
        ctx.presorted = _sort_on_side(emb, indices, offsets, B) if ctx.needs_input_grad[0] else None
        ops.embbag_forward(emb.table_set(), indices, offsets, B, out=feat[:, 1:, :],
                           stride_b=feat.stride(0), stride_t=feat.stride(1))
        x = x.contiguous()
        ctx.emb, ctx.itself = emb, bool(itself)
        ctx.save_for_backward(feat, x, indices, offsets)
        return ops.interact_dot_fwd_mixed(feat, x, out_cols, ctx.itself)

    @staticmethod
    def backward(ctx, dR):
        feat, x, indices, offsets = ctx.saved_tensors
        B, F, D = feat.shape
        dfeat, dx = ops.interact_dot_bwd_mixed(feat, x, dR.contiguous(), ctx.itself, relu_x=ctx.relu_x)
        emb, presorted = ctx.emb, ctx.presorted
        if FLUSH_EARLY:
            # the top tower's deferred weight-gradient kernel right behind the interaction backward, the sparse update forked
            # from before it (1.198 ms at the Terabyte shape: the bottom tower's dgrad GEMMs -- one 34 K-register CTA per SM --
            # then wait ~35 us for the update's CTAs to retire and run at half speed beside it)
            done = _flush_deferred_wgrads()
            _update_on_side(emb, dfeat[:, 1:, :], indices, offsets, B, F * D, D, presorted, keep=(dfeat,), after=done)
        else:
            # the sparse update is launched when the backward pass ends: the bottom tower's dgrad GEMMs run alone (they are the
            # only kernels here that cannot share an SM with the update's CTAs), then the weight-gradient kernels of both
            # towers (15 K registers per SM) go out FIRST and the update fills the SMs beside them
            from . import mlp
            mlp._HOLD[0] = True           # the towers' own end-of-pass flush stands back: tail() decides the order

            def tail():
                mlp._HOLD[0] = False
                done = torch.cuda.current_stream().record_event()       # the update forks from here, not from behind the wgrads
                mlp.flush_deferred()
                _update_on_side(emb, dfeat[:, 1:, :], indices, offsets, B, F * D, D, presorted, keep=(dfeat,), after=done)
            torch.autograd.Variable._execution_engine.queue_callback(tail)
        ctx.presorted = None
        return dx, None, None, None, None, None, None, None


class _DistStep(object):
    """What one multi-process training step shares between the forward halves and the autograd nodes."""

    def __init__(self, model, lb):
        self.model, self.lb = model, lb
        self.peer = None            # ext_dist.PeerExchange, or None: NCCL all-to-all
        self.feat = None            # [lb, F, D] interaction input (feature pitch may exceed D)
        self.emb = self.indices = self.offsets = self.presorted = None      # sharded tables (local), global batch
        self.rep = self.rep_indices = self.rep_offsets = self.rep_sorted = None   # replicated tables, local batch
        self.info = None            # NCCL: the exchange descriptor


class _DistInteract(torch.autograd.Function):
    """Multi-process sparse arch behind the exchange: the interaction input [lb, 1 + T_s + T_rep, D] holds the pooled rows
    of the sharded tables (rank-major order, delivered by the peers' lookup kernels over NVLink or unpacked from the NCCL
    all-to-all) followed by those of the replicated tables (looked up locally).  backward: interaction gradient, then --
    on the side stream, beside the bottom-MLP backward -- the transposed exchange + fused update of the local sharded
    tables, and the accumulate / all-reduce / update of the replicated ones.  out_cols > 0 selects the bf16-facing
    interaction kernels (x, R, dR, dx in bf16)."""

    @staticmethod
    def forward(ctx, x, recv, step, itself, out_cols, relu_x=False):
        feat = step.feat
        if step.peer is None:
            step.info = ext_dist.complete_into(recv, feat, 1)
        ctx.step, ctx.itself, ctx.mixed, ctx.relu_x = step, bool(itself), out_cols > 0, bool(relu_x)
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
        if ctx.mixed:
            dfeat, dx = ops.interact_dot_bwd_mixed(feat, x, dR.contiguous(), ctx.itself, relu_x=ctx.relu_x)
        else:
            dfeat = ops.interact_dot_bwd(feat, dR.contiguous(), ctx.itself)
            dx = dfeat[:, 0, :]
        # the interaction backward's end: the sparse updates fork from here; the deferred weight-gradient kernel of the top
        # tower and the all-reduce of its gradients are issued first -- collectives of one process group run in issue order,
        # and this one (10 MB, ready when the replicated tables' sums are) must not wait for that one (the largest)
        done = _flush_deferred_wgrads()
        peer, lb = step.peer, step.lb
        F, D = dfeat.shape[1], dfeat.shape[2]
        T_s = F - 1 - (step.rep.T if step.rep is not None else 0)
        send = None
        if peer is None:      # NCCL: pack + start the transposed all-to-all; _A2AReq.backward completes it and the batched
            send = ext_dist.start_backward_from(step.info, dfeat, 1,     # lookup's own backward runs the fused update
                                                wire_bf16=getattr(step.model, "_grad_wire_bf16", False))

        def sparse_updates():                 # on the side stream
            cur = torch.cuda.current_stream()
            comm = _comm_stream(dfeat.device, 1)
            if step.rep is not None:
                # replicated tables first: their gradient all-reduce (62 MB at the Terabyte shape, ~0.26 ms on 8 GPUs) and
                # row update go to the communication stream and run beside the push / update of the sharded tables
                step.rep.fused_backward(dfeat[:, 1 + T_s:, :], step.rep_indices, step.rep_offsets, lb,
                                        stride_b=dfeat.stride(0), stride_t=dfeat.stride(1), presorted=step.rep_sorted,
                                        comm_stream=comm)
            if peer is not None:
                grad = peer.push_grad(dfeat)
                if step.emb.T > 0:
                    step.emb.fused_backward(grad, step.indices, step.offsets, peer.W * lb, stride_b=grad.stride(0),
                                            stride_t=grad.stride(1), presorted=step.presorted)
            if step.rep is not None and comm != cur:
                cur.wait_stream(comm)         # whoever joins the side stream has joined the communication stream too
            return None
        if peer is not None or step.rep is not None:
            _update_on_side(None, dfeat, None, None, 0, 0, 0, None, before=sparse_updates, keep=(dfeat, step), after=done)
        step.presorted = step.rep_sorted = None
        return dx, send, None, None, None, None


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
        """dlrm_s_pytorch.py:321-334: one table-batched module over the local tables, optimizer fused.  With
        `replicate_rows_below` (multi-process only) the tables below that row count get a second, data-parallel module
        (`self.emb_rep`, embedding.ReplicatedTableBatchedEmbeddingBags) that every rank holds in full."""
        assert weighted_pooling is None, "Weighted pooling not supported yet!"
        self.Es = Es[self.local_emb_indices] if ext_dist.my_size > 1 else Es
        T = len(self.Es)
        kw = dict(optimizer=self.emb_optimizer, learning_rate=learning_rate, eps=self.emb_eps, stochastic_rounding=False,
                  init=self.emb_init, weights_precision=self.emb_weights_precision)
        self.emb_rep = None
        if ext_dist.my_size > 1 and len(self.replicated_emb_indices) > 0:
            rows = Es[self.replicated_emb_indices]
            self.emb_rep = embedding.ReplicatedTableBatchedEmbeddingBags(len(rows), rows, D, **kw)
        return embedding.TableBatchedEmbeddingBags(T, self.Es, D, **kw), [None] * T

    def __init__(self, m_spa=None, ln_emb=None, ln_bot=None, ln_top=None, arch_interaction_op=None,
                 arch_interaction_itself=False, sigmoid_bot=-1, sigmoid_top=-1, sync_dense_params=True,
                 loss_threshold=0.0, ndevices=-1, weighted_pooling=None, loss_function="bce",
                 batched_emb=False, sharder="naive", allocation=None,
                 qr_flag=False, qr_operation="mult", qr_collisions=0, qr_threshold=200, md_flag=False, md_threshold=200,
                 # additions (keyword-only in spirit): the fused optimizer of the batched module
                 emb_optimizer=embedding.Optimizer.SGD, emb_learning_rate=0.1, emb_eps=None, emb_init="auto",
                 fuse_sparse_arch=False, replicate_rows_below=0, emb_weights_precision="fp32"):
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
        self.emb_weights_precision = emb_weights_precision      # "bf16": 16-bit row storage of the batched module
        self.fuse_sparse_arch = bool(fuse_sparse_arch)
        if self.fuse_sparse_arch and not (batched_emb and arch_interaction_op == "dot"):
            raise ValueError("fuse_sparse_arch needs batched_emb=True and arch_interaction_op='dot'")
        self.device_indices = [0]

        self.replicated_emb_indices, self.emb_rep = [], None
        if ext_dist.my_size > 1:       # table-wise model parallel (dlrm_s_pytorch.py:444-468)
            n_emb = len(self.ln_emb)
            # tables kept on every rank instead of being sharded (fused batched path only): ascending table order
            if replicate_rows_below > 0:
                if not (batched_emb and self.fuse_sparse_arch):
                    raise _lib.DlrmxError("replicate_rows_below needs batched_emb=True and fuse_sparse_arch=True")
                self.replicated_emb_indices = [i for i in range(n_emb) if int(self.ln_emb[i]) < replicate_rows_below]
            self.sharded_emb_indices = [i for i in range(n_emb) if i not in set(self.replicated_emb_indices)]
            n_sh = len(self.sharded_emb_indices)
            if n_sh < ext_dist.my_size and not self.replicated_emb_indices:
                sys.exit("only (%d) sparse features for (%d) devices, table partitions will fail"
                         % (n_emb, ext_dist.my_size))
            self.n_global_emb = n_emb
            if sharder == "input":
                owners = list(map(int, allocation.split(",")))
                if len(owners) != n_sh:
                    sys.exit("ERROR: --allocation lists %d devices for %d sharded tables" % (len(owners), n_sh))
            elif n_sh >= ext_dist.my_size:
                owners = sharders.shard(self.ln_emb[self.sharded_emb_indices], ext_dist.my_size, sharder)
            else:                       # fewer sharded tables than ranks (the rest is replicated): one each, some ranks none
                owners = list(range(n_sh))
            # device_indices[i] = owner rank of table i, -1 for a replicated table
            self.device_indices = [-1] * n_emb
            for i, r in zip(self.sharded_emb_indices, owners):
                self.device_indices[i] = int(r)
            self.n_emb_per_rank = [0] * ext_dist.my_size
            for i in self.sharded_emb_indices:
                self.n_emb_per_rank[self.device_indices[i]] += m_spa // int(self.ln_bot[-1])
            self.local_emb_indices = [i for i, j in enumerate(self.device_indices) if j == ext_dist.my_local_rank]
            # feature order of the interaction input: sharded tables in rank-major order (what the reference's cat of the
            # all-to-all outputs produces), then the replicated tables in table order
            self.feature_tables = [i for r in range(ext_dist.my_size) for i in self.sharded_emb_indices
                                   if self.device_indices[i] == r] + list(self.replicated_emb_indices)
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

    def fuse_mlps(self, lr, optimizer="adagrad", eps=1e-10, gemm="tc", allreduce="peer", grad_wire_bf16=False):
        """Replaces the two nn.Sequential towers by FusedMLP (bf16 tensor-core GEMMs with fused epilogues, flat
        parameter buffers, one-kernel optimizer step).  After this, call `step_mlps()` instead of putting the tower
        parameters into an outer optimizer."""
        from . import mlp
        self.bot_l = mlp.FusedMLP(self.bot_l, lr, optimizer, eps, gemm=gemm)
        self.top_l = mlp.FusedMLP(self.top_l, lr, optimizer, eps, gemm=gemm)
        # bf16 hand-over around the interaction (no cast / pad / slice passes): the bottom tower ends in a fused bf16
        # layer, the top tower starts with one, and the tensor-core interaction kernels take the shape
        F, D = 1 + int(self.ln_emb.size), int(self.ln_bot[-1])
        first, last = self.top_l.layers[0], self.bot_l.layers[-1]
        nz = ops.tri_count(F, self.arch_interaction_itself)
        self._mixed = bool(self.fuse_sparse_arch and self.arch_interaction_op == "dot" and first.fused and last.fused
                           and first.K == D + nz and ops.interact_mixed_supported(F, D))
        self.bot_l.out_bf16 = self._mixed
        # the top tower's backward comes first: its weight gradients wait for the end of the pass, where they overlap the
        # sparse update on the side stream instead of delaying the interaction backward
        self.top_l.defer_wgrad = bool(self.fuse_sparse_arch and gemm == "tc")
        if gemm == "tc" and ext_dist.my_size > 1:      # step_mlps() averages over the replicas: start that in the backward
            self.bot_l.dp_world = self.top_l.dp_world = ext_dist.my_size
        if self.emb_rep is not None and self._mixed:
            self.emb_rep.wire_bf16 = True              # mixed-precision mode: the replicated tables' gradients cross NVLink as bf16
        if ext_dist.my_size > 1 and self.fuse_sparse_arch:
            # grad_wire_bf16 (opt-in, mixed-precision mode only): the pooled-row gradients of the sharded tables cross NVLink
            # rounded to bf16, peer-memory and NCCL exchange alike (the same values either way)
            wire = bool(grad_wire_bf16 and self._mixed and D % 8 == 0)
            peer = getattr(self, "_peer", None)
            if peer is not None and peer.grad_bf16 != wire:
                raise _lib.DlrmxError("enable_peer_exchange(grad_wire_bf16=%s) does not match fuse_mlps (grad_wire_bf16=%s)"
                                      % (peer.grad_bf16, wire))
            self._grad_wire_bf16 = wire
        if allreduce not in ("peer", "nccl"):
            raise ValueError("allreduce must be 'peer' (own kernel over NVLink peer memory) or 'nccl'")
        if allreduce == "peer" and ext_dist.my_size > 1:
            # the three data-parallel sums of a step (two towers, replicated tables) leave NCCL: gradient buffers in
            # symmetric memory, ext_dist.PeerAllReduce
            dev = self.bot_l.master.device
            self.bot_l.use_peer_allreduce(ext_dist.PeerAllReduce(self.bot_l.numel, torch.float32, dev))
            self.top_l.use_peer_allreduce(ext_dist.PeerAllReduce(self.top_l.numel, torch.float32, dev))
            if self.emb_rep is not None:
                self.emb_rep.peer_ar = ext_dist.PeerAllReduce(
                    self.emb_rep.grad_sum.numel(), torch.bfloat16 if self.emb_rep.wire_bf16 else torch.float32, dev)
        return self

    def step_mlps(self):
        """Optimizer step of the fused towers (gradient all-reduce across the data-parallel replicas included)."""
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

    def _take_over_bottom_relu(self):
        """Call right after the fused bottom tower's forward: True if the bf16-facing interaction backward may gate dx by
        (x > 0), i.e. apply the ReLU backward of the tower's last layer (x is that layer's output); the tower's backward
        then skips its ReLU-backward pass and takes the bias gradient from its weight-gradient kernel."""
        bot = self.bot_l
        last = getattr(bot, "_last_relu", None)
        if getattr(bot, "gemm", None) != "tc" or last is None or last[0] != len(bot.layers) - 1 or not torch.is_grad_enabled():
            return False
        last[1]["pre_masked"] = "nobias"
        return True

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
                                                    self.arch_interaction_itself, self.top_l.layers[0].K_pad,
                                                    self._take_over_bottom_relu())
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

    def enable_peer_exchange(self, local_batch, mixed=None, grad_wire_bf16=False):
        """Replaces the NCCL all-to-all of distributed_forward by kernels that store straight into the peers'
        memory over NVLink (ext_dist.PeerExchange).  Needs fuse_sparse_arch + batched_emb and a fixed local batch.
        mixed=True (bf16 towers will be fused in: fuse_mlps) lays the interaction input out with the row pitch the
        bf16-facing interaction kernels stage with one bulk copy per sample."""
        if not (self.fuse_sparse_arch and self.batched_emb):
            raise _lib.DlrmxError("enable_peer_exchange needs batched_emb=True and fuse_sparse_arch=True")
        dev = next(self.parameters()).device
        D = int(self.ln_bot[-1])
        n_rep = len(self.replicated_emb_indices)
        F = 1 + int(sum(self.n_emb_per_rank)) + n_rep
        mixed = bool(getattr(self, "_mixed", False) if mixed is None else mixed)
        pitch = ops.interact_feature_pitch(F, D) if mixed else D
        # opt-in with the mixed-precision mode: the pooled-row gradients cross NVLink as bf16 (D % 8 == 0: one 16-byte store
        # per 8 elements); pass the same flag to fuse_mlps
        self._grad_wire_bf16 = bool(grad_wire_bf16) and mixed and D % 8 == 0
        self._peer = ext_dist.PeerExchange(self.n_emb_per_rank, local_batch, D, dev, feature_pitch=pitch,
                                           extra_features=n_rep, grad_bf16=self._grad_wire_bf16)
        return self

    def sync_replicas(self):
        """Replicated tables start (or restart, after loading a checkpoint) from rank 0's rows on every rank."""
        if self.emb_rep is not None:
            self.emb_rep.sync_replicas(0)
        return self

    def _split_inputs(self, lS_o, lS_i):
        """(sharded offsets, indices), (replicated offsets, indices or None): with replicated tables the caller hands
        pairs -- lS_o = (offsets of the local sharded tables x GLOBAL batch, offsets of the replicated tables x LOCAL
        batch), lS_i likewise (what distribute_batched_emb_data returns)."""
        if self.emb_rep is None:
            return (lS_o, lS_i), (None, None)
        if not (isinstance(lS_o, (tuple, list)) and len(lS_o) == 2 and isinstance(lS_i, (tuple, list)) and len(lS_i) == 2):
            raise _lib.DlrmxError("replicated tables: pass lS_o = (sharded_offsets, replicated_offsets) and lS_i likewise")
        return (lS_o[0], lS_i[0]), (lS_o[1], lS_i[1])

    def _replicated_lookup(self, step, feat, T_s, off, idx):
        """Pooled rows of the replicated tables for the LOCAL batch, straight into their feature slots; the sort for
        their backward is issued right behind (MGPU_REP_SORT_SIDE: on the side stream)."""
        rep = self.emb_rep
        step.rep, step.rep_indices, step.rep_offsets = rep, idx, off
        ops.embbag_forward(rep.table_set(), idx, off, step.lb, out=feat[:, 1 + T_s:, :], stride_b=feat.stride(0),
                           stride_t=feat.stride(1))
        if torch.is_grad_enabled():
            if MGPU_OVERLAP & MGPU_REP_SORT_SIDE:
                # on the side stream, where the backward consumes it: on the main stream its eight launches (~80 us at the
                # Terabyte shape) sit between this lookup and the interaction (profiles/r02_timeline_n8_peer_ar_rank0.txt)
                step.rep_sorted = _sort_on_side(rep, idx, off, step.lb)
            else:
                step.rep_sorted = rep.sort_lookups(idx, off, step.lb)

    def _peer_forward(self, dense_x, lS_o, lS_i):
        peer = self._peer
        emb = self._batched_module()
        (off, idx), (roff, ridx) = self._split_inputs(lS_o, lS_i)
        lb, T_s = peer.lb, peer.T
        step = _DistStep(self, lb)
        step.peer, step.feat, step.emb, step.indices, step.offsets = peer, peer.feat, emb, idx, off
        main = torch.cuda.current_stream()
        side = _side_stream(dense_x.device)
        side.wait_stream(main)
        with record_function("module::forward_pass::embedding_lookup"):
            with torch.cuda.stream(side):       # lookup + NVLink scatter (and the backward's sorts) beside the bottom tower
                peer.lookup_and_scatter(emb.table_set() if emb.T > 0 else None, idx, off)   # a rank may hold no sharded table
                peer.wait_rows()
                arrived = side.record_event()
                if torch.is_grad_enabled() and emb.T > 0:
                    step.presorted = emb.sort_lookups(idx, off, peer.W * lb)
        rep_done = None
        if self.emb_rep is not None and (MGPU_OVERLAP & MGPU_REP_LOOKUP_STREAM):
            # local rows of the replicated tables: HBM work on a third stream, beside the NVLink-bound scatter AND the bottom
            # tower (behind the tower on the main stream it was the last thing the interaction waited for)
            rs = _comm_stream(dense_x.device, 1)
            if rs != main:
                rs.wait_stream(main)
            with record_function("module::forward_pass::embedding_lookup"), torch.cuda.stream(rs):
                self._replicated_lookup(step, peer.feat, T_s, roff, ridx)
                rep_done = rs.record_event() if rs != main else None
        with record_function("module::forward_pass::bottom_mlp"):
            x = self.apply_mlp(dense_x, self.bot_l)
        if self.emb_rep is not None and not (MGPU_OVERLAP & MGPU_REP_LOOKUP_STREAM):      # local rows of the replicated tables: HBM work beside the
            with record_function("module::forward_pass::embedding_lookup"):       # NVLink-bound scatter
                self._replicated_lookup(step, peer.feat, T_s, roff, ridx)
        cols = self.top_l.layers[0].K_pad if getattr(self, "_mixed", False) else 0
        with record_function("module::forward_pass::interaction"):
            main.wait_event(arrived)
            if rep_done is not None:
                main.wait_event(rep_done)
            z = _DistInteract.apply(x, None, step, self.arch_interaction_itself, cols,
                                    cols > 0 and self._take_over_bottom_relu())
        with record_function("module::forward_pass::top_mlp"):
            p = self.apply_mlp(z, self.top_l)
        return self._clamp(p)

    def distributed_forward(self, dense_x, lS_o, lS_i):
        """Local tables x full global batch -> all-to-all (overlapped with the bottom MLP) -> local batch x
        all tables -> interaction -> top MLP (dlrm_s_pytorch.py:686-730)."""
        if not isinstance(self.emb_l, nn.ModuleList):
            self.emb_l = nn.ModuleList([self.emb_l])
        if getattr(self, "_peer", None) is not None:
            return self._peer_forward(dense_x, lS_o, lS_i)
        (off, idx), (roff, ridx) = self._split_inputs(lS_o, lS_i)
        with record_function("module::forward_pass::embedding_lookup"):
            if self.batched_emb and self._batched_module().T == 0:
                # a rank without sharded tables (fewer of them than ranks, the rest replicated) still takes part in the
                # exchange, with an empty [B, 0, D] block that requires grad so that its backward half runs here too
                ly = [torch.zeros((dense_x.shape[0] * ext_dist.my_size, 0, int(self.m_spa)), device=dense_x.device,
                                  requires_grad=True)]
            elif self.batched_emb:
                ly = self.apply_emb_batched([off], [idx])
            else:
                ly = self.apply_emb(off, idx)
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
                mixed = getattr(self, "_mixed", False)
                lb, D = x.shape
                T_s = int(sum(self.n_emb_per_rank))
                F = 1 + T_s + len(self.replicated_emb_indices)
                step = _DistStep(self, lb)
                step.feat = ops.alloc_features(lb, F, D, x.device, pitch=None if mixed else D)
                if self.emb_rep is not None:
                    self._replicated_lookup(step, step.feat, T_s, roff, ridx)
                z = _DistInteract.apply(x, recv, step, self.arch_interaction_itself,
                                        self.top_l.layers[0].K_pad if mixed else 0, mixed and self._take_over_bottom_relu())
            else:
                z = self.interact_features(x, ly)
        with record_function("module::forward_pass::top_mlp"):
            p = self.apply_mlp(z, self.top_l)
        return self._clamp(p)

    def distribute_batched_emb_data(self, batch_size, lS_o, lS_i):
        """Packed global batch (all T tables) -> this rank's packed batch (local tables, offsets re-based),
        dlrm_s_pytorch.py:772-802 (distributed branch).  Unlike the reference it does not assume a fixed
        number of indices per lookup: each table's index range comes from its own offsets.  With replicated tables
        the result is ([offsets], [indices]) of PAIRS: (local sharded tables x global batch, replicated tables x this
        rank's slice of the batch)."""
        B = batch_size
        if ext_dist.my_size <= 1:
            return [lS_o], [lS_i]
        bounds = lS_o[::B].tolist() if lS_o.numel() % B == 1 else None

        def gather(tables, b0, b1):
            o_parts, i_parts, run = [], [], 0
            for k in tables:
                o = lS_o[k * B + b0:k * B + b1 + 1]
                if o.numel() == b1 - b0:           # last table without the trailing total: close it with N
                    o = torch.cat([o, o.new_tensor([lS_i.numel()])])
                if b0 == 0 and b1 == B and bounds:
                    lo, hi = bounds[k], bounds[k + 1]
                else:
                    lo, hi = int(o[0]), int(o[-1])
                i_parts.append(lS_i[lo:hi])
                rel = o - lo + run
                o_parts.append(rel if not o_parts else rel[1:])
                run += hi - lo
            if not o_parts:
                return lS_o.new_zeros(1), lS_i[:0]
            return torch.cat(o_parts, dim=0), torch.cat(i_parts, dim=0)

        off, idx = gather(self.local_emb_indices, 0, B)
        if self.emb_rep is None:
            return [off], [idx]
        lb = B // ext_dist.my_size
        roff, ridx = gather(self.replicated_emb_indices, ext_dist.my_rank * lb, (ext_dist.my_rank + 1) * lb)
        return [(off, roff)], [(idx, ridx)]
