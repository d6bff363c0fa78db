"""Thin tensor-level wrappers over the C ABI (include/dlrmx.h).  PyTorch is plumbing here: device
memory, the current stream and autograd hooks; all arithmetic is in libdlrmx.so.  Nothing in this
module has a CPU path -- CPU tensors raise.
"""
import ctypes

import torch

from . import _lib


def _stream():
    return torch.cuda.current_stream().cuda_stream


# Optional per-kernel timing (bench.py): when PROFILE is a dict, every wrapper brackets its launch(es) with
# CUDA events on the launching stream and appends (start, end) under the kernel's name.
PROFILE = None


class _timed:
    def __init__(self, name):
        self.name = name

    def __enter__(self):
        if PROFILE is not None:
            # under CUDA-graph capture: external events become event-record NODES of the graph, so every replay
            # re-records them and elapsed_time() gives the kernel's duration inside the replayed step
            ext = torch.cuda.is_current_stream_capturing()
            self.s = torch.cuda.Event(enable_timing=True, external=True) if ext else torch.cuda.Event(enable_timing=True)
            self.e = torch.cuda.Event(enable_timing=True, external=True) if ext else torch.cuda.Event(enable_timing=True)
            self.s.record()

    def __exit__(self, *a):
        if PROFILE is not None:
            self.e.record()
            PROFILE.setdefault(self.name, []).append((self.s, self.e))
        return False


def profile_summary():
    """{name: (calls, mean ms)} of the events collected so far (synchronises)."""
    torch.cuda.synchronize()
    return {k: (len(v), sum(s.elapsed_time(e) for s, e in v) / len(v)) for k, v in (PROFILE or {}).items()}


def _ptr(t):
    return 0 if t is None else t.data_ptr()


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise _lib.DlrmxError("dlrm_yx_b200 kernels need CUDA tensors (no CPU fallback); got a %s tensor"
                                  % t.device)


class TableSet:
    """A group of embedding tables that are looked up / updated together by one kernel launch.

    Holds the device-side pointer arrays the C ABI takes: table base pointers, row counts, the exclusive
    row prefix (global row ids for the sorted backward) and, for RWSAdagrad, accumulator pointers.
    Tables stay ordinary [rows_t, D] fp32 tensors (state_dict / checkpoint compatible:
    SURVEY.md section 5) -- the arrays are rebuilt whenever a tensor's storage moves.
    """

    def __init__(self, weights, moms=None, acc_in_row=False):
        self.weights = list(weights)
        self.moms = None if moms is None else list(moms)
        # padded rows (row pitch ld > D, e.g. views [:, :D] of a [rows, D+4] allocation): the row-wise Adagrad
        # accumulator of a row may live in the float right after it (acc_in_row) instead of in a separate array
        self.acc_in_row = bool(acc_in_row)
        self._key = None
        self.refresh()

    def refresh(self):
        ws = self.weights
        # only addresses, shapes and pitches are cached (never contents), so in-place updates need no refresh
        key = tuple((w.data_ptr(), tuple(w.shape), w.stride(0)) for w in ws) + \
            tuple(() if self.moms is None else (m.data_ptr() for m in self.moms))
        if key == self._key:
            return
        _need_cuda(*ws)
        D = ws[0].shape[1]
        ld = ws[0].stride(0) if ws[0].shape[0] > 1 else D
        self.bf16 = ws[0].dtype == torch.bfloat16      # bf16 row storage: widened on load, fp32 sums / updates, rounded on store
        for w in ws:
            if w.dtype != ws[0].dtype or w.dtype not in (torch.float32, torch.bfloat16) or w.dim() != 2 or w.shape[1] != D or \
                    w.stride(1) != 1 or (w.shape[0] > 1 and w.stride(0) != ld):
                raise _lib.DlrmxError("tables must be row-major fp32 (or all bf16) [rows, D] with a common D and a common row pitch")
        if self.bf16 and self.acc_in_row:
            raise _lib.DlrmxError("bf16 tables keep their row-wise accumulators in a separate fp32 array (no acc_in_row)")
        if ld < D or (self.acc_in_row and ld <= D):
            raise _lib.DlrmxError("bad table row pitch %d for D=%d (acc_in_row needs padded rows)" % (ld, D))
        self.ld = int(ld)
        dev = ws[0].device
        self.device, self.D, self.T = dev, D, len(ws)
        rows = [int(w.shape[0]) for w in ws]
        base = [0]
        for r in rows:
            base.append(base[-1] + r)
        self.total_rows = base[-1]
        self.rows = rows
        self.table_ptrs = torch.tensor([w.data_ptr() for w in ws], dtype=torch.int64, device=dev)
        self.table_rows = torch.tensor(rows, dtype=torch.int64, device=dev)
        self.row_base = torch.tensor(base, dtype=torch.int64, device=dev)
        self.mom_ptrs = None
        if self.moms is not None:
            _need_cuda(*self.moms)
            self.mom_ptrs = torch.tensor([m.data_ptr() for m in self.moms], dtype=torch.int64, device=dev)
        self._key = key


# ---- index validation ------------------------------------------------------------------------------------------
# One device-side error word per device: the lookup sets bit 0, the backward's key builder bit 1, when an index lies
# outside [0, rows_t) (the row is then skipped / the update dropped -- nothing is corrupted).  nn.EmbeddingBag raises
# on such input; reading the word costs a device synchronisation, so it is read where the host synchronises anyway
# (`check_index_errors()`: end of a step that reads its loss, the parity seam, tests, bench.py, smoke()).
_ERR_FLAGS = {}


def err_flag(device):
    f = _ERR_FLAGS.get(device)
    if f is None:
        f = _ERR_FLAGS[device] = torch.zeros(1, dtype=torch.int32, device=device)
    return f


def check_index_errors(device=None):
    """Raises DlrmxError if any lookup since the last check carried an index outside its table (synchronises)."""
    devs = list(_ERR_FLAGS) if device is None else [torch.device(device)]
    for d in devs:
        f = _ERR_FLAGS.get(d)
        if f is None:
            continue
        v = int(f.item())
        if v:
            f.zero_()
            raise _lib.DlrmxError("embedding index out of range (%s): an index was negative or >= the row count of its "
                                  "table; the offending lookups were skipped" %
                                  " and ".join(n for b, n in ((1, "forward lookup"), (2, "backward sort")) if v & b))


def _idx_suffix(indices, offsets):
    if indices.dtype != offsets.dtype or indices.dtype not in (torch.int32, torch.int64):
        raise _lib.DlrmxError("indices and offsets must both be int32 or both int64")
    return "i32" if indices.dtype == torch.int32 else "i64"


def embbag_forward(tables, indices, offsets, B, out=None, stride_b=None, stride_t=None,
                   per_sample_weights=None, err_flag=None, validate=True):
    """Pooled rows of every bag (t, b).  Default output layout [B, T, D] (the batched module's,
    dlrm_s_pytorch.py:589-591); pass out/stride_b/stride_t to write elsewhere (e.g. features 1.. of the
    [B, 1+T, D] interaction input)."""
    tables.refresh()
    _need_cuda(indices, offsets, per_sample_weights, out)
    indices = indices.contiguous()
    offsets = offsets.contiguous()
    T, D = tables.T, tables.D
    if offsets.numel() < T * B:
        raise _lib.DlrmxError("offsets has %d entries, need at least T*B = %d" % (offsets.numel(), T * B))
    if out is None:
        out = torch.empty((B, T, D), dtype=torch.float32, device=tables.device)
        stride_b, stride_t = T * D, D
    fn = getattr(_lib.load(), "dlrmx_embbag_fwd_" + ("bf16_" if tables.bf16 else "") + _idx_suffix(indices, offsets))
    if err_flag is None and validate:
        err_flag = globals()["err_flag"](tables.device)
    with _timed("embbag_fwd"):
        rc = fn(tables.table_ptrs.data_ptr(), tables.table_rows.data_ptr() if validate else 0,
                indices.data_ptr(), offsets.data_ptr(), indices.numel(), T, B, D, tables.ld, _ptr(per_sample_weights),
                out.data_ptr(), stride_b, stride_t, _ptr(err_flag), _stream())
        _lib.check(rc, "dlrmx_embbag_fwd")
    return out


def embbag_forward_scatter(tables, indices, offsets, B, out_peers, lb, stride_b, stride_t, validate=False):
    """embbag_forward with the exchange fused in: bag (t, b) of the global batch is stored at
    out_peers[b // lb] + (b % lb)*stride_b + t*stride_t (out_peers: device int64 tensor of peer-mapped pointers)."""
    tables.refresh()
    _need_cuda(indices, offsets, out_peers)
    indices = indices.contiguous()
    offsets = offsets.contiguous()
    T, D = tables.T, tables.D
    if offsets.numel() < T * B:
        raise _lib.DlrmxError("offsets has %d entries, need at least T*B = %d" % (offsets.numel(), T * B))
    if out_peers.dtype != torch.int64 or out_peers.numel() * lb != B:
        raise _lib.DlrmxError("out_peers must hold B / lb int64 pointers")
    if tables.bf16:
        raise _lib.DlrmxError("the NVLink scatter lookup takes fp32 tables (bf16 row storage: single-device / NCCL paths)")
    fn = getattr(_lib.load(), "dlrmx_embbag_fwd_scatter_" + _idx_suffix(indices, offsets))
    with _timed("embbag_fwd"):
        rc = fn(tables.table_ptrs.data_ptr(), tables.table_rows.data_ptr() if validate else 0, indices.data_ptr(),
                offsets.data_ptr(), indices.numel(), T, B, D, tables.ld, 0, out_peers.data_ptr(), lb,
                stride_b, stride_t, 0, _stream())
        _lib.check(rc, "dlrmx_embbag_fwd_scatter")


def bwd_sort(tables, indices, offsets, B):
    """(sorted_key uint32-as-int32 tensor [N], sorted_bag int32 [N]): global row ids of all looked-up
    indices, stably sorted, with the bag each one came from."""
    tables.refresh()
    _need_cuda(indices, offsets)
    indices = indices.contiguous()
    offsets = offsets.contiguous()
    N = indices.numel()
    lib = _lib.load()
    dev = tables.device
    skey = torch.empty(N, dtype=torch.int32, device=dev)   # bit pattern is uint32
    sbag = torch.empty(N, dtype=torch.int32, device=dev)
    wsb = lib.dlrmx_bwd_sort_workspace_bytes(N, tables.T * B)
    ws = torch.empty(max(wsb, 1), dtype=torch.uint8, device=dev)
    fn = getattr(lib, "dlrmx_bwd_sort_" + _idx_suffix(indices, offsets))
    with _timed("bwd_sort"):
        rc = fn(indices.data_ptr(), offsets.data_ptr(), N, tables.T, B, tables.row_base.data_ptr(),
                tables.total_rows, skey.data_ptr(), sbag.data_ptr(), err_flag(dev).data_ptr(), ws.data_ptr(), wsb,
                _stream())
        _lib.check(rc, "dlrmx_bwd_sort")
    return skey, sbag


def bwd_segments(skey):
    """(unique_key [N], run_off [N+1], num_unique [1]) on device; only the first num_unique(+1) entries
    are meaningful."""
    _need_cuda(skey)
    N = skey.numel()
    lib = _lib.load()
    dev = skey.device
    uniq = torch.empty(max(N, 1), dtype=torch.int32, device=dev)
    run_off = torch.empty(N + 1, dtype=torch.int32, device=dev)
    nuniq = torch.empty(1, dtype=torch.int32, device=dev)
    wsb = lib.dlrmx_bwd_segments_workspace_bytes(N)
    ws = torch.empty(max(wsb, 1), dtype=torch.uint8, device=dev)
    rc = lib.dlrmx_bwd_segments(skey.data_ptr(), N, uniq.data_ptr(), run_off.data_ptr(), nuniq.data_ptr(),
                                ws.data_ptr(), wsb, _stream())
    _lib.check(rc, "dlrmx_bwd_segments")
    return uniq, run_off, nuniq


def bwd_update(tables, optimizer, grad, stride_b, stride_t, B, skey, sbag, lr, eps=0.0):
    """Fused segmented reduce + sparse optimizer update of the touched table rows (in place)."""
    tables.refresh()
    _need_cuda(grad, skey, sbag)
    N = skey.numel()
    lib = _lib.load()
    if optimizer == _lib.OPT_RWSADAGRAD and tables.mom_ptrs is None and not tables.acc_in_row:
        raise _lib.DlrmxError("RWSAdagrad update needs row-wise accumulators (TableSet(moms=...) or acc_in_row)")
    wsb = lib.dlrmx_bwd_update_workspace_bytes(N, tables.D)
    ws = torch.empty(max(wsb, 1), dtype=torch.uint8, device=tables.device)
    if grad.dtype not in (torch.float32, torch.bfloat16):
        raise _lib.DlrmxError("bwd_update: pooled-row gradients must be fp32 or bf16, got %s" % grad.dtype)
    with _timed("bwd_update"):
        if grad.dtype == torch.bfloat16:        # what the bf16 gradient push / all-to-all delivers (strides in elements)
            rc = lib.dlrmx_bwd_update_ex(optimizer, 1 if tables.bf16 else 0, 1, tables.table_ptrs.data_ptr(),
                                         0 if tables.acc_in_row else _ptr(tables.mom_ptrs), tables.row_base.data_ptr(),
                                         tables.T, B, tables.D, tables.ld, grad.data_ptr(), stride_b, stride_t,
                                         skey.data_ptr(), sbag.data_ptr(), N, tables.total_rows, float(lr), float(eps),
                                         ws.data_ptr(), wsb, _stream())
        else:
            fn = lib.dlrmx_bwd_update_bf16 if tables.bf16 else lib.dlrmx_bwd_update
            rc = fn(optimizer, tables.table_ptrs.data_ptr(),
                    0 if tables.acc_in_row else _ptr(tables.mom_ptrs),
                    tables.row_base.data_ptr(), tables.T, B, tables.D, tables.ld, grad.data_ptr(), stride_b,
                    stride_t, skey.data_ptr(), sbag.data_ptr(), N, tables.total_rows, float(lr), float(eps),
                    ws.data_ptr(), wsb, _stream())
        _lib.check(rc, "dlrmx_bwd_update")


def bwd_coalesce(grad, stride_b, stride_t, B, D, sbag, run_off, nuniq, max_unique=None):
    """Coalesced gradient rows [N, D] (first num_unique valid) for the non-fused parity seam."""
    _need_cuda(grad, sbag, run_off, nuniq)
    N = sbag.numel()
    gsum = torch.empty((max(N, 1), D), dtype=torch.float32, device=grad.device)
    rc = _lib.load().dlrmx_bwd_coalesce(B, D, grad.data_ptr(), stride_b, stride_t, sbag.data_ptr(),
                                        run_off.data_ptr(), nuniq.data_ptr(), N, gsum.data_ptr(), _stream())
    _lib.check(rc, "dlrmx_bwd_coalesce")
    return gsum


def tri_count(F, itself=False):
    return F * (F + 1) // 2 if itself else F * (F - 1) // 2


def interact_dot_fwd(feat, itself=False, out=None):
    """feat [B, F, D] (sample stride may exceed F*D) -> [B, D + F(F-1)/2]."""
    _need_cuda(feat, out)
    B, F, D = feat.shape
    if feat.stride(2) != 1 or feat.stride(1) != D:
        raise _lib.DlrmxError("feat must be [B, F, D] with contiguous F x D blocks")
    if out is None:
        out = torch.empty((B, D + tri_count(F, itself)), dtype=torch.float32, device=feat.device)
    with _timed("interact_fwd"):
        rc = _lib.load().dlrmx_interact_dot_fwd(feat.data_ptr(), feat.stride(0), B, F, D, int(itself),
                                                out.data_ptr(), out.stride(0), _stream())
        _lib.check(rc, "dlrmx_interact_dot_fwd")
    return out


def interact_dot_bwd(feat, dout, itself=False, dfeat=None):
    _need_cuda(feat, dout, dfeat)
    B, F, D = feat.shape
    if dout.stride(1) != 1:
        dout = dout.contiguous()
    if dfeat is None:
        dfeat = torch.empty((B, F, D), dtype=torch.float32, device=feat.device)
    with _timed("interact_bwd"):
        rc = _lib.load().dlrmx_interact_dot_bwd(feat.data_ptr(), feat.stride(0), B, F, D, int(itself),
                                                dout.data_ptr(), dout.stride(0), dfeat.data_ptr(),
                                                dfeat.stride(0), _stream())
        _lib.check(rc, "dlrmx_interact_dot_bwd")
    return dfeat


def interact_mixed_supported(F, D):
    return 2 <= F <= 32 and D % 8 == 0 and D <= 512


def interact_feature_pitch(F, D):
    """Row pitch (floats) the bf16-facing interaction kernels like best for their [B, F, D] input: D + 8 (the pitch of
    their shared-memory tile, so that ONE TMA bulk copy stages a sample) when D % 32 == 0, else D."""
    return D + 8 if (interact_mixed_supported(F, D) and D % 32 == 0 and _lib.get_tunable_py(10) != 1) else D


def alloc_features(B, F, D, device, pitch=None):
    """[B, F, D] fp32 view with row pitch `pitch` (default interact_feature_pitch) of a fresh allocation."""
    pitch = interact_feature_pitch(F, D) if pitch is None else pitch
    return torch.empty((B, F, pitch), dtype=torch.float32, device=device)[:, :, :D]


def interact_dot_fwd_mixed(feat, x_bf16, out_cols, itself=False):
    """feat [B, F, D] fp32 (features 1.. valid), x_bf16 [B, D] (feature 0) -> bf16 [B, out_cols] = [x | tril | 0...]."""
    _need_cuda(feat, x_bf16)
    B, F, D = feat.shape
    if feat.stride(2) != 1 or feat.stride(1) not in (D, D + 8) or not x_bf16.is_contiguous() or \
            x_bf16.dtype != torch.bfloat16:
        raise _lib.DlrmxError("interact_dot_fwd_mixed: feat [B, F, D] fp32 with feature pitch D or D + 8, x contiguous bf16")
    out = torch.empty((B, out_cols), dtype=torch.bfloat16, device=feat.device)
    with _timed("interact_fwd"):
        rc = _lib.load().dlrmx_interact_dot_fwd_mixed(feat.data_ptr(), feat.stride(0), feat.stride(1), x_bf16.data_ptr(), B, F, D,
                                                      int(itself), out.data_ptr(), out.stride(0), _stream())
        _lib.check(rc, "dlrmx_interact_dot_fwd_mixed")
    return out


def interact_dot_bwd_mixed(feat, x_bf16, dout, itself=False, relu_x=False):
    """-> (dfeat fp32 [B, F, D] with features 1.. written, dx bf16 [B, D]); relu_x: dx is gated by (x > 0)."""
    _need_cuda(feat, x_bf16, dout)
    B, F, D = feat.shape
    if dout.dtype != torch.bfloat16 or dout.stride(1) != 1:
        raise _lib.DlrmxError("interact_dot_bwd_mixed: dout must be bf16 with unit column stride")
    dfeat = torch.empty((B, F, D), dtype=torch.float32, device=feat.device)
    dx = torch.empty((B, D), dtype=torch.bfloat16, device=feat.device)
    with _timed("interact_bwd"):
        rc = _lib.load().dlrmx_interact_dot_bwd_mixed(feat.data_ptr(), feat.stride(0), feat.stride(1), x_bf16.data_ptr(), B, F, D,
                                                      int(itself), dout.data_ptr(), dout.stride(0), dfeat.data_ptr(),
                                                      dfeat.stride(0), dx.data_ptr(), int(relu_x), _stream())
        _lib.check(rc, "dlrmx_interact_dot_bwd_mixed")
    return dfeat, dx


def _int_array(values):
    return (ctypes.c_int * len(values))(*[int(v) for v in values])


def a2a_push_grad(dfeat, feat_off, tables_per_rank, lb, D, peer_chunk_bases, bf16=False):
    """dfeat[:, feat_off + t, :] -> rank p's chunk [lb, T_p, D] at peer_chunk_bases[p] (peer-mapped addresses); bf16: the
    owners' buffers are bf16 and the rows are rounded to nearest even on the way."""
    _need_cuda(dfeat)
    bases = (ctypes.c_void_p * len(peer_chunk_bases))(*[int(b) for b in peer_chunk_bases])
    with _timed("a2a_push_grad"):
        fn = _lib.load().dlrmx_a2a_push_grad_bf16 if bf16 else _lib.load().dlrmx_a2a_push_grad
        rc = fn(dfeat.data_ptr(), dfeat.stride(0), dfeat.stride(1), feat_off, len(tables_per_rank),
                _int_array(tables_per_rank), lb, D, bases, _stream())
        _lib.check(rc, "dlrmx_a2a_push_grad")


def peer_allreduce(peer_ptrs, rank, numel, is_bf16):
    """In-place sum over the ranks' peer-mapped buffers (peer_ptrs[p] = rank p's buffer); the caller brackets the call with
    device barriers (extend_distributed.PeerAllReduce does)."""
    if not torch.cuda.is_available():
        raise _lib.DlrmxError("peer_allreduce needs CUDA (no CPU fallback)")
    ptrs = (ctypes.c_void_p * len(peer_ptrs))(*[int(b) for b in peer_ptrs])
    with _timed("peer_allreduce"):
        rc = _lib.load().dlrmx_peer_allreduce(ptrs, len(peer_ptrs), int(rank), int(numel), 1 if is_bf16 else 0, _stream())
        _lib.check(rc, "dlrmx_peer_allreduce")


def a2a_unpack(recv, tables_per_rank, lb, D, dst, feat_off):
    """recv = concat_p [lb, T_p, D]  ->  dst[:, feat_off + t, :] (rank-major table order)."""
    _need_cuda(recv, dst)
    with _timed("a2a_unpack"):
        rc = _lib.load().dlrmx_a2a_unpack(recv.data_ptr(), len(tables_per_rank), _int_array(tables_per_rank), lb, D,
                                          dst.data_ptr(), dst.stride(0), dst.stride(1), feat_off, _stream())
        _lib.check(rc, "dlrmx_a2a_unpack")
    return dst


def a2a_pack_grad(dfeat, feat_off, tables_per_rank, lb, D, send=None):
    _need_cuda(dfeat, send)
    T = sum(int(t) for t in tables_per_rank)
    if send is None:
        send = torch.empty(lb * T * D, dtype=torch.float32, device=dfeat.device)
    with _timed("a2a_pack_grad"):
        rc = _lib.load().dlrmx_a2a_pack_grad(dfeat.data_ptr(), dfeat.stride(0), dfeat.stride(1), feat_off, len(tables_per_rank),
                                             _int_array(tables_per_rank), lb, D, send.data_ptr(), _stream())
        _lib.check(rc, "dlrmx_a2a_pack_grad")
    return send


def rows_update(optimizer, weight, mom, rows, gsum, lr, eps=0.0):
    """Non-fused sparse optimizer step of one table: rows int64 [U] unique, gsum [U, D] coalesced."""
    _need_cuda(weight, mom, rows, gsum)
    if rows.dtype != torch.int64 or not rows.is_contiguous() or not gsum.is_contiguous():
        raise _lib.DlrmxError("rows_update: rows must be contiguous int64, gsum contiguous fp32")
    U, D = gsum.shape
    rc = _lib.load().dlrmx_rows_update(optimizer, weight.data_ptr(), _ptr(mom), rows.data_ptr(), gsum.data_ptr(),
                                       U, D, float(lr), float(eps), _stream())
    _lib.check(rc, "dlrmx_rows_update")


def dense_rows_update(optimizer, weight, mom, gsum, lr, eps=0.0):
    """Replicated tables: optimizer step of every row of weight [rows, D] whose all-reduced gradient row gsum [rows, D]
    is non-zero; clears those rows of gsum."""
    _need_cuda(weight, mom, gsum)
    rows, D = weight.shape
    if gsum.shape != weight.shape or not gsum.is_contiguous() or weight.stride(1) != 1:
        raise _lib.DlrmxError("dense_rows_update: gsum must be contiguous [rows, D] like the table")
    with _timed("dense_rows_update"):
        rc = _lib.load().dlrmx_dense_rows_update(optimizer, weight.data_ptr(), _ptr(mom), gsum.data_ptr(), rows, D,
                                                 weight.stride(0) if rows > 1 else D, float(lr), float(eps), _stream())
        _lib.check(rc, "dlrmx_dense_rows_update")


_RB_COUNTERS = {}


def relu_bwd_bias_grad(dy, y, relu=True, chunks=None, db=None):
    """(dz bf16 [B,N], db_partial fp32 [chunks,N]): dz = dy * (y > 0) and per-chunk column sums of dz, one pass.
    With `db` (fp32 [N]) the kernel also finishes the column sums into it (deterministic last-CTA reduce).
    chunks=None sizes the grid for ~2 CTAs of 256 threads per SM (148 SMs; more chunks only lengthen the last CTA's reduction) with at least 32 rows per chunk."""
    _need_cuda(dy, y, db)
    B, N = dy.shape
    col_blocks = (N + 255) // 256
    if chunks is None:
        chunks = max(1, min((B + 31) // 32, (148 * 2 + col_blocks - 1) // col_blocks))
    dz = torch.empty_like(dy) if relu else dy
    part = torch.empty((chunks, N), dtype=torch.float32, device=dy.device)
    counters = None
    if db is not None:
        key = (dy.device, torch.cuda.current_stream().cuda_stream)
        counters = _RB_COUNTERS.get(key)
        if counters is None or counters.numel() < col_blocks:
            counters = _RB_COUNTERS[key] = torch.zeros(max(col_blocks, 64), dtype=torch.int32, device=dy.device)
    with _timed("relu_bwd_bias_grad"):
        rc = _lib.load().dlrmx_relu_bwd_bias_grad(dy.data_ptr(), y.data_ptr(), dz.data_ptr(), part.data_ptr(), _ptr(db),
                                                  _ptr(counters), B, N, chunks, int(relu), _stream())
        _lib.check(rc, "dlrmx_relu_bwd_bias_grad")
    return dz, part


EPI_NONE, EPI_BIAS_RELU, EPI_BIAS, EPI_RELU_BITS = 0, 1, 3, 4


def relu_bits_empty(M, N, device):
    """Gate-bit matrix of an [M, N] activation: int32 [M, ceil(N / 32)], bit (n & 31) of word [m, n >> 5] = (y[m, n] > 0)."""
    return torch.empty((M, (N + 31) // 32), dtype=torch.int32, device=device)


def gemm_bf16(a, b, b_is_kn=False, epilogue=EPI_NONE, bias=None, relu_mask=None, out=None, relu_bits_out=None):
    """Tensor-core GEMM (tcgen05 / TMEM / TMA): out [M, N] bf16 = epilogue(a [M, K] . op(b)); b is [N, K] (y = x W^t) or,
    with b_is_kn, [K, N] (dx = dz W).  Row-major bf16 operands with unit column stride.  relu_mask: the int32 gate bits of
    relu_bits_empty (EPI_RELU_BITS); relu_bits_out: such a bit matrix, written by the EPI_BIAS_RELU epilogue."""
    _need_cuda(a, b, bias, relu_mask, out, relu_bits_out)
    if a.dtype != torch.bfloat16 or b.dtype != torch.bfloat16 or a.stride(1) != 1 or b.stride(1) != 1:
        raise _lib.DlrmxError("gemm_bf16: bf16 row-major operands with unit column stride")
    M, K = a.shape
    N = b.shape[1] if b_is_kn else b.shape[0]
    if (b.shape[0] if b_is_kn else b.shape[1]) != K:
        raise _lib.DlrmxError("gemm_bf16: contraction sizes differ (%s x %s)" % (tuple(a.shape), tuple(b.shape)))
    if out is None:
        out = torch.empty((M, N), dtype=torch.bfloat16, device=a.device)
    bits = relu_mask if epilogue == EPI_RELU_BITS else relu_bits_out
    if bits is not None and (bits.dtype != torch.int32 or bits.stride(1) != 1 or bits.shape[0] != M):
        raise _lib.DlrmxError("gemm_bf16: gate bits are int32 [M, ceil(N / 32)] (ops.relu_bits_empty)")
    ld_mask = bits.stride(0) if bits is not None else 0
    with _timed("gemm_bf16"):
        rc = _lib.load().dlrmx_gemm_bf16(a.data_ptr(), a.stride(0), b.data_ptr(), b.stride(0), int(b_is_kn), out.data_ptr(),
                                         out.stride(0), M, N, K, int(epilogue), _ptr(bias), _ptr(relu_mask), ld_mask,
                                         _ptr(relu_bits_out), _stream())
        _lib.check(rc, "dlrmx_gemm_bf16")
    return out


def gemm_wgrad_workspace(Ns, Ks, M, device):
    """Workspace for gemm_wgrad_bf16 on layers of these shapes (split-K partial tiles)."""
    n = len(Ns)
    nbytes = _lib.load().dlrmx_gemm_wgrad_workspace_bytes(n, _int_array(Ns), _int_array(Ks), int(M))
    return torch.empty(nbytes, dtype=torch.uint8, device=device)


def gemm_wgrad_bf16(dzs, xs, dws, dbs, workspace):
    """For every layer l: dws[l] [N_l, K_l] fp32 = dzs[l]^t xs[l], dbs[l] [N_l] = column sums of dzs[l] (None: skip) -- all
    layers in ONE split-K tensor-core launch; dzs[l] [M, N_l], xs[l] [M, K_l] bf16 row-major."""
    n = len(dzs)
    _need_cuda(*dzs, *xs, *dws, workspace)
    M = dzs[0].shape[0]
    vp = lambda ts: (ctypes.c_void_p * n)(*[0 if t is None else t.data_ptr() for t in ts])      # noqa: E731
    i64 = lambda vs: (ctypes.c_int64 * n)(*[int(v) for v in vs])                               # noqa: E731
    for dz, x, dw in zip(dzs, xs, dws):
        if dz.dtype != torch.bfloat16 or x.dtype != torch.bfloat16 or dw.dtype != torch.float32 or dz.shape[0] != M or \
                x.shape[0] != M or dz.stride(1) != 1 or x.stride(1) != 1 or dw.stride(1) != 1 or \
                tuple(dw.shape) != (dz.shape[1], x.shape[1]):
            raise _lib.DlrmxError("gemm_wgrad_bf16: dz [M, N] / x [M, K] bf16 row-major, dw [N, K] fp32")
    with _timed("gemm_wgrad"):
        rc = _lib.load().dlrmx_gemm_wgrad_bf16(
            n, vp(dzs), i64([t.stride(0) for t in dzs]), vp(xs), i64([t.stride(0) for t in xs]), vp(dws),
            i64([t.stride(0) for t in dws]), vp(dbs) if dbs is not None else None, _int_array([t.shape[1] for t in dzs]),
            _int_array([t.shape[1] for t in xs]), M, workspace.data_ptr(), workspace.numel(), _stream())
        _lib.check(rc, "dlrmx_gemm_wgrad_bf16")


def head_fwd(h, w, bias, sigmoid=True):
    """p [B, 1] fp32 = sigmoid(h @ w + bias) for a bf16 activation h [B, K] and fp32 w [K], bias [1]."""
    _need_cuda(h, w, bias)
    B, K = h.shape
    p = torch.empty((B, 1), dtype=torch.float32, device=h.device)
    with _timed("head_fwd"):
        rc = _lib.load().dlrmx_head_fwd(h.data_ptr(), w.data_ptr(), bias.data_ptr(), p.data_ptr(), B, K, int(sigmoid),
                                        _stream())
        _lib.check(rc, "dlrmx_head_fwd")
    return p


def head_bwd(dp, p, h, w, gw, gb, sigmoid=True, gb_prev_relu=None):
    """dh bf16 [B, K]; writes gw [K] and gb [1] (fp32) -- backward of head_fwd.  gb_prev_relu [K] fp32: h is the output of
    a Linear + ReLU layer; its ReLU backward and bias gradient are fused in (the result is that layer's dz)."""
    _need_cuda(dp, p, h, w, gw, gb)
    B, K = h.shape
    lib = _lib.load()
    col_blocks = (K + 255) // 256
    # 2 CTAs per SM: the main pass is a few microseconds either way, the last CTA's reduction over the chunks is not
    chunks = max(1, min((B + 31) // 32, (148 * 2 + col_blocks - 1) // col_blocks))
    ws = torch.empty(lib.dlrmx_head_bwd_workspace_bytes(K, chunks) // 4, dtype=torch.float32, device=h.device)
    key = (h.device, torch.cuda.current_stream().cuda_stream, "head")
    counters = _RB_COUNTERS.get(key)
    if counters is None or counters.numel() < col_blocks:
        counters = _RB_COUNTERS[key] = torch.zeros(max(col_blocks, 64), dtype=torch.int32, device=h.device)
    dh = torch.empty_like(h)
    dp = dp.contiguous()
    with _timed("head_bwd"):
        rc = lib.dlrmx_head_bwd(dp.data_ptr(), p.data_ptr(), h.data_ptr(), w.data_ptr(), dh.data_ptr(), gw.data_ptr(),
                                gb.data_ptr(), ws.data_ptr(), counters.data_ptr(), B, K, chunks, int(sigmoid),
                                _ptr(gb_prev_relu), _stream())
        _lib.check(rc, "dlrmx_head_bwd")
    return dh


def dense_opt(mode, params, acc, grads, shadow, lr, eps=0.0, grad_scale=1.0):
    """One launch over a flat fp32 parameter buffer: SGD (mode 0) or Adagrad (mode 1); refreshes the bf16 shadow and
    clears the gradients."""
    _need_cuda(params, acc, grads, shadow)
    with _timed("dense_opt"):
        rc = _lib.load().dlrmx_dense_opt(mode, params.data_ptr(), _ptr(acc), grads.data_ptr(), _ptr(shadow),
                                         params.numel(), float(lr), float(eps), float(grad_scale), _stream())
        _lib.check(rc, "dlrmx_dense_opt")


def criteo_bin_batch(records, max_ind_range=-1, batched=True, index_dtype=None):
    """records: int32 [B, 40] CUDA tensor of raw Criteo-Terabyte binary records -> (X [B,13] fp32, offsets, indices,
    target [B,1] fp32) exactly as CriteoBinDataset.__getitem__ / _transform_features return them
    (data_loader_terabyte.py:68-95, 236-249): batched -> indices int32 [26*B], offsets int32 [26*B+1];
    otherwise lS_o / lS_i int64 [26, B]."""
    _need_cuda(records)
    if records.dtype != torch.int32 or records.dim() != 2 or records.shape[1] != 40 or not records.is_contiguous():
        raise _lib.DlrmxError("criteo_bin_batch: records must be a contiguous int32 [B, 40] tensor")
    B = int(records.shape[0])
    dt = index_dtype or (torch.int32 if batched else torch.int64)
    dev = records.device
    X = torch.empty((B, 13), dtype=torch.float32, device=dev)
    tgt = torch.empty((B, 1), dtype=torch.float32, device=dev)
    indices = torch.empty(26 * B, dtype=dt, device=dev)
    offsets = torch.empty(26 * B + (1 if batched else 0), dtype=dt, device=dev)
    fn = _lib.load().dlrmx_criteo_bin_batch_i32 if dt == torch.int32 else _lib.load().dlrmx_criteo_bin_batch_i64
    with _timed("criteo_bin_batch"):
        rc = fn(records.data_ptr(), B, int(max_ind_range), 1 if batched else 0, X.data_ptr(), indices.data_ptr(),
                offsets.data_ptr(), tgt.data_ptr(), _stream())
    _lib.check(rc, "dlrmx_criteo_bin_batch")
    if batched:
        return X, offsets, indices, tgt
    return X, offsets.view(26, B), indices.view(26, B), tgt


def embbag_psw_grad(weight, indices, offsets, grad_out):
    """dpsw [N] fp32: gradient of EmbeddingBag(mode="sum")'s per_sample_weights for ONE table --
    dpsw[l] = <grad_out[bag(l)], weight[indices[l]]>."""
    _need_cuda(weight, indices, offsets, grad_out)
    indices, offsets, grad_out = indices.contiguous(), offsets.contiguous(), grad_out.contiguous()
    N, B, D = indices.numel(), offsets.numel(), weight.shape[1]
    dpsw = torch.empty(N, dtype=torch.float32, device=weight.device)
    fn = getattr(_lib.load(), "dlrmx_embbag_psw_grad_" + _idx_suffix(indices, offsets))
    with _timed("embbag_psw_grad"):
        rc = fn(weight.data_ptr(), weight.stride(0), weight.shape[0], indices.data_ptr(), offsets.data_ptr(), N, B, D,
                grad_out.data_ptr(), grad_out.stride(0), dpsw.data_ptr(), _stream())
        _lib.check(rc, "dlrmx_embbag_psw_grad")
    return dpsw


def bwd_coalesce_weighted(grad_out, sorted_pos, pos2bag, psw, run_off, nuniq):
    """Weighted coalesced gradient rows [N, D] (first num_unique valid) of ONE table: sorted_pos from bwd_sort run with
    every lookup as its own bag, pos2bag int32 [N], psw fp32 [N] (original order)."""
    _need_cuda(grad_out, sorted_pos, pos2bag, psw, run_off, nuniq)
    grad_out = grad_out.contiguous()
    N, D = sorted_pos.numel(), grad_out.shape[1]
    gsum = torch.empty((N, D), dtype=torch.float32, device=grad_out.device)
    with _timed("bwd_coalesce_weighted"):
        rc = _lib.load().dlrmx_bwd_coalesce_weighted(D, grad_out.data_ptr(), grad_out.stride(0), sorted_pos.data_ptr(),
                                                     pos2bag.contiguous().data_ptr(), psw.contiguous().data_ptr(),
                                                     run_off.data_ptr(), nuniq.data_ptr(), N, gsum.data_ptr(), _stream())
        _lib.check(rc, "dlrmx_bwd_coalesce_weighted")
    return gsum


class _BCEMean(torch.autograd.Function):
    """torch.nn.BCELoss(reduction="mean") as one kernel each way (csrc/loss.cu)."""

    @staticmethod
    def forward(ctx, p, target):
        p, target = p.contiguous(), target.contiguous()
        n = p.numel()
        lib = _lib.load()
        ws = torch.empty(max(lib.dlrmx_bce_mean_workspace_bytes(n) // 4, 1), dtype=torch.float32, device=p.device)
        key = (p.device, torch.cuda.current_stream().cuda_stream, "bce")
        counter = _RB_COUNTERS.get(key)
        if counter is None:
            counter = _RB_COUNTERS[key] = torch.zeros(1, dtype=torch.int32, device=p.device)
        loss = torch.empty((), dtype=torch.float32, device=p.device)
        with _timed("bce_mean"):
            rc = lib.dlrmx_bce_mean_fwd(p.data_ptr(), target.data_ptr(), n, ws.data_ptr(), counter.data_ptr(),
                                        loss.data_ptr(), _stream())
            _lib.check(rc, "dlrmx_bce_mean_fwd")
        ctx.save_for_backward(p, target)
        return loss

    @staticmethod
    def backward(ctx, dloss):
        p, target = ctx.saved_tensors
        dp = torch.empty_like(p)
        dloss = dloss.contiguous().to(torch.float32)
        with _timed("bce_mean"):
            rc = _lib.load().dlrmx_bce_mean_bwd(p.data_ptr(), target.data_ptr(), dloss.data_ptr(), p.numel(), dp.data_ptr(),
                                                _stream())
            _lib.check(rc, "dlrmx_bce_mean_bwd")
        return dp, None


def bce_mean(p, target):
    """Mean binary cross-entropy of predictions p against targets (both fp32 CUDA, same shape)."""
    _need_cuda(p, target)
    if p.dtype != torch.float32 or target.dtype != torch.float32 or p.shape != target.shape:
        raise _lib.DlrmxError("bce_mean: p and target must be fp32 tensors of the same shape")
    return _BCEMean.apply(p, target)
