"""Embedding modules with the reference's call surfaces, backed by libdlrmx.so (sm_100a CUDA).

* `EmbeddingBagSum`            drop-in for `nn.EmbeddingBag(n, m, mode="sum", sparse=True)` as built by
                               DLRM_Net.create_emb (dlrm_s_pytorch.py:301-308) and called by apply_emb
                               (dlrm_s_pytorch.py:571-577): `.weight`, `E(indices, offsets, per_sample_weights)`.
* `TableBatchedEmbeddingBags`  the `table_batched_embeddings_ops` module dlrm_s_pytorch.py:321-334 imports
  + `Optimizer`                (absent from the reference tree): `emb(indices_i32, offsets_i32[T*B+1]) -> [B,T,D]`
                               with the sparse optimizer fused into backward.
`install_table_batched_embeddings_ops()` publishes this module under the name the reference imports.

Backward never materialises the [N, D] gradient expansion of EmbeddingBag.backward: looked-up rows are
radix-sorted (8 bytes each), and either (fused) the segmented reduce + SGD / row-wise Adagrad update runs
in one kernel, or (parity seam) a COALESCED sparse COO gradient is emitted so that the reference's own
`torch.optim.SGD` / `RWSAdagrad.step` (optim/rwsadagrad.py:97-113) consume it unchanged.

No CPU path: CPU tensors raise (there is no fallback by design).
"""
import enum
import math
import sys

import numpy as np
import torch
from torch import nn

from . import _lib, ops


class Optimizer(enum.Enum):
    """Same member names as table_batched_embeddings_ops.Optimizer (only .SGD is used by the reference,
    dlrm_s_pytorch.py:329)."""
    SGD = 1
    APPROX_ROWWISE_ADAGRAD = 2
    EXACT_ROWWISE_ADAGRAD = 3


_ROWWISE = (Optimizer.APPROX_ROWWISE_ADAGRAD, Optimizer.EXACT_ROWWISE_ADAGRAD)


def _uniform_init_(w, rows, generator=None):
    """create_emb's rule (dlrm_s_pytorch.py:304-306): U(-sqrt(1/n), +sqrt(1/n))."""
    a = math.sqrt(1.0 / max(int(rows), 1))
    w.uniform_(-a, a, generator=generator)


# ------------------------------------------------------------------------------------------------
# batched module (boundary B2)
# ------------------------------------------------------------------------------------------------
class _BatchedLookup(torch.autograd.Function):
    @staticmethod
    def forward(ctx, weights, module, indices, offsets, B):
        out = ops.embbag_forward(module.table_set(), indices, offsets, B)
        ctx.module = module
        ctx.B = B
        ctx.save_for_backward(indices, offsets)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        indices, offsets = ctx.saved_tensors
        ctx.module.fused_backward(grad_out, indices, offsets, ctx.B)
        return None, None, None, None, None


class TableBatchedEmbeddingBags(nn.Module):
    """T embedding tables of a common dimension D looked up by ONE kernel launch.

    Constructor mirrors dlrm_s_pytorch.py:323-334: `(T, Es, D, optimizer=Optimizer.SGD, learning_rate=..,
    eps=.., stochastic_rounding=False)`; `Es` is an int (same rows for all) or a sequence of T row counts.
    `forward(indices, offsets)`: packed table-major int32 (or int64) indices and `offsets[T*B+1]`
    (dlrm_data_pytorch.py:748-753) -> pooled `[B, T, D]` fp32.

    The optimizer is fused into backward (as in the module it stands in for): `embedding_weights.grad`
    stays None, so the outer `optimizer.step()` (dlrm_s_pytorch.py:1650, 1933) skips these parameters.
    Row-wise Adagrad follows optim/rwsadagrad.py:95-113 exactly (mean of squares, sqrt(mom)+eps).
    """

    def __init__(self, num_tables, num_embeddings, embedding_dim, optimizer=Optimizer.SGD, learning_rate=0.01,
                 eps=None, stochastic_rounding=False, lr_decay=0.0, initial_accumulator_value=0.0,
                 init="auto", device=None, weights_precision="fp32"):
        """weights_precision: "fp32" (the reference's batched module) or "bf16" -- 16-bit row storage as the `weights_precision`
        argument of the reference's fbgemm slot offers (dlrm_s_pytorch.py:350-366): rows are widened to fp32 when read, pooled
        and updated in fp32 (fp32 row-wise accumulators), and rounded to nearest even when written (stochastic_rounding=False,
        dlrm_s_pytorch.py:333).  Half the table bytes per lookup and per update."""
        super().__init__()
        if stochastic_rounding:
            raise _lib.DlrmxError("stochastic_rounding is not implemented (bf16 rows are rounded to nearest even)")
        if weights_precision not in ("fp32", "bf16"):
            raise ValueError("weights_precision must be 'fp32' or 'bf16'")
        self.weights_precision = weights_precision
        if isinstance(num_embeddings, (int, np.integer)):
            rows = [int(num_embeddings)] * int(num_tables)
        else:
            rows = [int(e) for e in num_embeddings]
        if len(rows) != int(num_tables):
            raise ValueError("expected %d row counts, got %d" % (num_tables, len(rows)))
        self.T, self.D, self.rows = int(num_tables), int(embedding_dim), rows
        self.optimizer = optimizer
        self.learning_rate = float(learning_rate)
        self.eps = 1e-10 if eps is None else float(eps)
        self.lr_decay = float(lr_decay)
        self.step_count = 0
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        total = sum(rows)
        flat = torch.empty(total * self.D, dtype=torch.float32, device=dev)
        if init == "auto":
            init = "numpy" if total * self.D <= (1 << 26) else "device"
        base = 0
        for r in rows:
            view = flat[base * self.D:(base + r) * self.D].view(r, self.D)
            if init == "numpy":   # same draw order / distribution as create_emb (dlrm_s_pytorch.py:304-308)
                a = np.sqrt(1 / r)
                view.copy_(torch.from_numpy(np.random.uniform(low=-a, high=a, size=(r, self.D)).astype(np.float32)))
            elif init == "device":
                _uniform_init_(view, r)
            elif init != "empty":
                raise ValueError("init must be auto|numpy|device|empty")
            base += r
        if weights_precision == "bf16":
            flat = flat.to(torch.bfloat16)          # the fp32 draw rounded once: the same initial values on a bf16 grid
        self.embedding_weights = nn.Parameter(flat)
        if optimizer in _ROWWISE:
            self.register_buffer("row_momentum", torch.full((total,), float(initial_accumulator_value),
                                                            dtype=torch.float32, device=dev))
        else:
            self.row_momentum = None
        self._ts = None

    # per-table [rows_t, D] views of the flat parameter (state_dict keeps the flat tensor)
    def split_weights(self):
        out, base = [], 0
        for r in self.rows:
            out.append(self.embedding_weights.data[base * self.D:(base + r) * self.D].view(r, self.D))
            base += r
        return out

    def split_momentum(self):
        if self.row_momentum is None:
            return None
        return list(torch.split(self.row_momentum, self.rows))

    def table_set(self):
        key = (self.embedding_weights.data_ptr(), None if self.row_momentum is None else self.row_momentum.data_ptr())
        if self._ts is None or self._ts_key != key:
            self._ts = ops.TableSet(self.split_weights(), self.split_momentum())
            self._ts_key = key
        return self._ts

    def forward(self, indices, offsets):
        total_bags = offsets.numel() - 1
        if total_bags % self.T != 0:
            raise _lib.DlrmxError("offsets must have T*B+1 entries (T=%d), got %d" % (self.T, offsets.numel()))
        return _BatchedLookup.apply(self.embedding_weights, self, indices, offsets, total_bags // self.T)

    def current_lr(self):
        return self.learning_rate / (1.0 + self.step_count * self.lr_decay)

    def sort_lookups(self, indices, offsets, B):
        """(sorted global row ids, bag of each) -- the gradient-independent half of the backward; callers may run it
        early (e.g. on a side stream during the forward pass) and hand the result to `fused_backward`."""
        return ops.bwd_sort(self.table_set(), indices, offsets, B)

    def fused_backward(self, grad_out, indices, offsets, B, stride_b=None, stride_t=None, presorted=None):
        """grad_out: gradient of the pooled rows, [B, T, D] (or any layout given by the strides)."""
        ts = self.table_set()
        if stride_b is None:
            if grad_out.stride(2) != 1 or grad_out.stride(1) % 4 or grad_out.stride(0) % 4:
                grad_out = grad_out.contiguous()
            stride_b, stride_t = grad_out.stride(0), grad_out.stride(1)
        skey, sbag = presorted if presorted is not None else ops.bwd_sort(ts, indices, offsets, B)
        kind = _lib.OPT_RWSADAGRAD if self.optimizer in _ROWWISE else _lib.OPT_SGD
        ops.bwd_update(ts, kind, grad_out, stride_b, stride_t, B, skey, sbag, self.current_lr(), self.eps)
        self.step_count += 1


class ReplicatedTableBatchedEmbeddingBags(TableBatchedEmbeddingBags):
    """Data-parallel twin of the batched module for tables that are cheaper to keep on EVERY rank than to exchange (a
    sharding policy beside the reference's table-wise one, dlrm_s_pytorch.py:444-468: the Terabyte set has 18 tables
    with fewer rows than one mini-batch -- 62 MB in all -- whose pooled rows and gradients would otherwise cross NVLink
    for every sample).  Each rank looks the tables up for its LOCAL batch only; in backward the local pooled-row gradients
    are summed per row into a dense [rows, D] buffer (sorted segmented reduce, deterministic), the buffers are all-reduced,
    and one kernel applies the optimizer to every row with a non-zero summed gradient.  The result equals the sharded
    owner's update up to the order of the floating-point sum over ranks; replicas stay bit-identical to one another."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.register_buffer("grad_sum", torch.zeros_like(self.embedding_weights.data), persistent=False)
        self._gts = None
        self.wire_bf16 = False      # True: the gradient all-reduce moves bf16 (set with the bf16 towers, DLRM_Net.fuse_mlps)
        self.peer_ar = None         # ext_dist.PeerAllReduce over grad_sum.numel() elements (bf16 iff wire_bf16): own kernel, no NCCL

    def _grad_table_set(self):
        if self._gts is None or self._gts_key != self.grad_sum.data_ptr():
            views, base = [], 0
            for r in self.rows:
                views.append(self.grad_sum[base * self.D:(base + r) * self.D].view(r, self.D))
                base += r
            self._gts, self._gts_key = ops.TableSet(views), self.grad_sum.data_ptr()
        return self._gts

    def sync_replicas(self, src=0):
        """Every rank adopts rank `src`'s rows and accumulators (after construction / after loading a checkpoint)."""
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.broadcast(self.embedding_weights.data, src)
            if self.row_momentum is not None:
                dist.broadcast(self.row_momentum, src)

    def fused_backward(self, grad_out, indices, offsets, B, stride_b=None, stride_t=None, presorted=None, comm_stream=None):
        """comm_stream: the all-reduce of the summed gradients and the row update that follows are enqueued THERE (after
        the accumulation on the current stream), so that whatever the caller enqueues next on the current stream -- the
        gradient push and the update of the sharded tables -- runs beside them; the caller joins comm_stream."""
        import torch.distributed as dist
        if stride_b is None:
            if grad_out.stride(2) != 1 or grad_out.stride(1) % 4 or grad_out.stride(0) % 4:
                grad_out = grad_out.contiguous()
            stride_b, stride_t = grad_out.stride(0), grad_out.stride(1)
        gts = self._grad_table_set()
        skey, sbag = presorted if presorted is not None else ops.bwd_sort(self.table_set(), indices, offsets, B)
        # SGD with lr = -1 on the zeroed buffer: grad_sum[row] = sum of the local gradient rows, in sorted (stable) order
        ops.bwd_update(gts, _lib.OPT_SGD, grad_out, stride_b, stride_t, B, skey, sbag, -1.0, 0.0)
        cur = torch.cuda.current_stream()
        if comm_stream is None or comm_stream == cur:
            comm_stream = cur
        else:
            comm_stream.wait_stream(cur)
        with torch.cuda.stream(comm_stream):
            if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
                if self.peer_ar is not None:
                    # the repo's own two-shot sum over NVLink peer memory (csrc/peer_allreduce.cu): the partial sums go into
                    # the symmetric buffer (rounded to bf16 in the mixed-precision mode: half the bytes), are added in fp32
                    # in rank order and rounded once; every replica receives the same bits
                    self.peer_ar.buf.copy_(self.grad_sum)
                    self.peer_ar.allreduce()
                    self.grad_sum.copy_(self.peer_ar.buf)
                elif self.wire_bf16:
                    # bf16 on the wire (the towers' mixed-precision mode only): every rank's partial sums are rounded to
                    # bf16 before they are added, the sum is widened back -- half the NVLink bytes of the one collective
                    # that is as large as the tables themselves; replicas stay bit-identical (NCCL's result is)
                    wire = self.grad_sum.to(torch.bfloat16)
                    dist.all_reduce(wire)
                    self.grad_sum.copy_(wire)
                else:
                    dist.all_reduce(self.grad_sum)
            kind = _lib.OPT_RWSADAGRAD if self.optimizer in _ROWWISE else _lib.OPT_SGD
            total = sum(self.rows)
            ops.dense_rows_update(kind, self.embedding_weights.data.view(total, self.D), self.row_momentum,
                                  self.grad_sum.view(total, self.D), self.current_lr(), self.eps)
        self.step_count += 1


def install_table_batched_embeddings_ops():
    """Publishes this module as `table_batched_embeddings_ops` so that the reference's
    `create_emb_batched` (dlrm_s_pytorch.py:323) imports it unmodified."""
    sys.modules.setdefault("table_batched_embeddings_ops", sys.modules[__name__])


# ------------------------------------------------------------------------------------------------
# per-table module (boundary B3)
# ------------------------------------------------------------------------------------------------
class _BagLookup(torch.autograd.Function):
    @staticmethod
    def forward(ctx, weight, module, indices, offsets, psw):
        ts = module.table_set()
        B = offsets.numel()
        out = ops.embbag_forward(ts, indices, offsets, B, per_sample_weights=psw)
        ctx.module = module
        ctx.has_psw = psw is not None
        if psw is None:
            ctx.save_for_backward(indices, offsets)
        else:
            ctx.save_for_backward(indices, offsets, psw)
        return out.view(B, ts.D)

    @staticmethod
    def _backward_weighted(ctx, grad_out):
        """Weighted pooling (dlrm_s_pytorch.py:544-547): gradient of the per-sample weights, and the table rows'
        gradient with each lookup's weight applied -- lookups sorted by row with their original position as payload
        (every lookup its own bag), runs reduced by dlrmx_bwd_coalesce_weighted; no [N, D] expansion."""
        indices, offsets, psw = ctx.saved_tensors
        m = ctx.module
        B, N = offsets.numel(), indices.numel()
        dev = indices.device
        grad_out = grad_out.contiguous()
        ts = m.table_set()
        # before any row is updated: dpsw[l] = <grad_out[bag(l)], W[idx[l]]>
        dpsw = ops.embbag_psw_grad(m.weight.data, indices, offsets, grad_out) if ctx.needs_input_grad[4] else None
        if N == 0:
            return None, None, None, None, dpsw
        one_each = torch.arange(N + 1, dtype=indices.dtype, device=dev)
        skey, spos = ops.bwd_sort(ts, indices, one_each, N)
        uniq, run_off, nu = ops.bwd_segments(skey)
        ends = torch.cat([offsets[1:].to(torch.int64), torch.tensor([N], dtype=torch.int64, device=dev)])
        pos2bag = torch.repeat_interleave(torch.arange(B, dtype=torch.int32, device=dev), ends - offsets.to(torch.int64),
                                          output_size=N)
        gsum = ops.bwd_coalesce_weighted(grad_out, spos, pos2bag, psw.to(torch.float32), run_off, nu)
        U = int(nu.item())
        ops.check_index_errors(dev)         # the host is synchronised here anyway: a bad index raises like nn.EmbeddingBag
        rows = (uniq[:U].to(torch.int64) & 0xFFFFFFFF)
        if m._fused is not None:
            kind, lr, eps = m._fused.hyper(m.weight)
            ops.rows_update(kind, m.weight.data, m._fused.momentum(m.weight), rows.contiguous(), gsum[:U], lr, eps)
            return None, None, None, None, dpsw
        g = torch.sparse_coo_tensor(rows.view(1, U), gsum[:U], size=tuple(m.weight.shape), is_coalesced=True)
        return g, None, None, None, dpsw

    @staticmethod
    def backward(ctx, grad_out):
        if ctx.has_psw:
            return _BagLookup._backward_weighted(ctx, grad_out)
        indices, offsets = ctx.saved_tensors
        m = ctx.module
        B = offsets.numel()
        grad_out = grad_out.contiguous()
        ts = m.table_set()
        D = ts.D
        skey, sbag = ops.bwd_sort(ts, indices, offsets, B)
        if m._fused is not None:
            kind, lr, eps = m._fused.hyper(m.weight)
            ops.bwd_update(ts, kind, grad_out, D, D, B, skey, sbag, lr, eps)   # [B,1,D]: stride_b = D
            return None, None, None, None, None
        # parity seam: coalesced sparse COO gradient (what grad.coalesce() would return, rwsadagrad.py:98)
        uniq, run_off, nu = ops.bwd_segments(skey)
        gsum = ops.bwd_coalesce(grad_out, D, D, B, D, sbag, run_off, nu)
        U = int(nu.item())
        ops.check_index_errors(indices.device)   # the host is synchronised here anyway: a bad index raises like nn.EmbeddingBag
        rows = (uniq[:U].to(torch.int64) & 0xFFFFFFFF).view(1, U)
        g = torch.sparse_coo_tensor(rows, gsum[:U], size=tuple(m.weight.shape), is_coalesced=True)
        return g, None, None, None, None


class EmbeddingBagSum(nn.Module):
    """`nn.EmbeddingBag(num_embeddings, embedding_dim, mode="sum", sparse=True)` on libdlrmx.

    forward(indices int64|int32 [N], offsets [B], per_sample_weights=None) -> [B, embedding_dim]; the last
    bag runs to N and empty bags give zeros.  `.weight.grad` after backward is a coalesced sparse COO tensor
    unless a fused optimizer is attached (`RWSAdagrad.attach` / `SparseSGD.attach`), in which case the
    rows are updated inside backward and `.grad` stays None.
    """

    def __init__(self, num_embeddings, embedding_dim, mode="sum", sparse=True, _weight=None, device=None):
        super().__init__()
        if mode != "sum":
            raise ValueError("only mode='sum' exists on this path (the reference uses no other, dlrm_s_pytorch.py:301)")
        self.num_embeddings, self.embedding_dim = int(num_embeddings), int(embedding_dim)
        self.mode, self.sparse = mode, sparse
        if _weight is None:
            _weight = torch.empty((self.num_embeddings, self.embedding_dim), dtype=torch.float32, device=device)
            _weight.normal_(0, 1)     # nn.EmbeddingBag's default; create_emb overwrites .data (308)
        self.weight = nn.Parameter(_weight)
        self._fused = None
        self._ts = None

    def table_set(self):
        mom = None if self._fused is None else self._fused.momentum(self.weight)
        key = (self.weight.data_ptr(), None if mom is None else mom.data_ptr())
        if self._ts is None or self._ts_key != key:
            self._ts = ops.TableSet([self.weight.data], None if mom is None else [mom])
            self._ts_key = key
        return self._ts

    def forward(self, input, offsets, per_sample_weights=None):
        if input.dim() != 1 or offsets.dim() != 1:
            raise _lib.DlrmxError("EmbeddingBagSum takes 1-D indices and 1-D offsets")
        return _BagLookup.apply(self.weight, self, input, offsets, per_sample_weights)

    def extra_repr(self):
        return "%d, %d, mode=sum, sparse=True" % (self.num_embeddings, self.embedding_dim)
