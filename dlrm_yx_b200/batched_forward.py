"""`torch.ops.batched_forward.forward` -- the torch op of the reference's native extension, re-created on
libdlrmx (boundary B1).

Reference: yx_modfs/batched_forward.cpp:4-13 registers
    forward(Tensor[] weights, Tensor[] table_offsets, Tensor[] offsets, Tensor[] indices,
            int L_max, int BT_block_size, bool shmem) -> Tensor
implemented by batched_embedding_forward_cuda (yx_modfs/table_batched_embeddings_cuda_yx.cu:317-389) and
called as yx_modfs/dlrm.py:108-117 (note: offsets BEFORE indices).  Per device d: weights[d] is the
row-concatenation [sum_t E_t, D] of that device's tables, table_offsets[d] their first rows (int32),
offsets[d] / indices[d] the packed int32 lookups; the result is [B, n_devices*T, D] on weights[0]'s device,
device d's tables occupying feature slots [d*T, (d+1)*T).

Differences from the reference kernel, on purpose (SURVEY.md 2.2 defects): the `t*D` term missing from its
store address is present (so more than one table per device gives correct results), addressing is 64-bit,
no streams are created or leaked and the caller's current stream on each device is used with proper event
ordering, `L_max / BT_block_size / shmem` are accepted and ignored (the reference ignores two of them too).
Importing this module is what registers the op, like `import table_batched_embeddings_yx.batched_forward`.
"""
import torch

from . import _lib

_lib_def = torch.library.Library("batched_forward", "DEF")
_lib_def.define("forward(Tensor[] weights, Tensor[] table_offsets, Tensor[] offsets, Tensor[] indices, "
                "int L_max, int BT_block_size, bool shmem) -> Tensor")

_ptr_cache = {}


def _table_ptrs(w, toff):
    """Device array of per-table base pointers for a concatenated weight matrix (cached per storage)."""
    # _version: table_offsets rewritten in place under the same pointer must not hit a stale entry
    key = (w.data_ptr(), toff.data_ptr(), toff.numel(), w.device.index, toff._version)
    hit = _ptr_cache.get(key)
    if hit is None:
        rows = toff.to(torch.int64)
        hit = (w.data_ptr() + rows * (w.shape[1] * 4)).contiguous()
        if len(_ptr_cache) > 64:
            _ptr_cache.clear()
        _ptr_cache[key] = hit
    return hit


def _forward_cuda(weights, table_offsets, offsets, indices, L_max, BT_block_size, shmem):
    n = len(weights)
    if not (n == len(table_offsets) == len(offsets) == len(indices)) or n == 0:
        raise RuntimeError("batched_forward: need one weights/table_offsets/offsets/indices tensor per device")
    T = table_offsets[0].numel()
    D = weights[0].shape[1]
    B = (offsets[0].numel() - 1) // T
    lib = _lib.load()
    dev0 = weights[0].device
    out = torch.empty((B, n * T, D), dtype=weights[0].dtype, device=dev0)
    if weights[0].dtype != torch.float32:
        raise RuntimeError("batched_forward: fp32 tables only")
    done = []
    ready = torch.cuda.Event()
    ready.record(torch.cuda.current_stream(dev0))          # `out` allocation is stream-ordered on dev0
    for d in range(n):
        w, toff, off, idx = weights[d], table_offsets[d], offsets[d], indices[d]
        if toff.numel() != T or (off.numel() - 1) != T * B or w.shape[1] != D:
            raise RuntimeError("batched_forward: every device must hold the same number of tables and D")
        if idx.dtype != torch.int32 or off.dtype != torch.int32 or toff.dtype != torch.int32:
            raise RuntimeError("batched_forward: table_offsets, offsets and indices are int32")
        with torch.cuda.device(w.device):
            st = torch.cuda.current_stream(w.device)
            st.wait_event(ready)
            if w.device != dev0 and not torch.cuda.can_device_access_peer(w.device.index, dev0.index):
                raise RuntimeError("batched_forward: device %s cannot store to %s over NVLink/P2P" % (w.device, dev0))
            if w.device != dev0:
                # torch enables peer access lazily and per direction: a copy whose SOURCE lives on w.device lets w.device
                # reach dev0's memory, which is what this device's lookup kernel needs to store into `out`
                torch.empty(1, device=dev0).copy_(torch.zeros(1, device=w.device))
            ptrs = _table_ptrs(w, toff)
            slot = out[:, d * T:(d + 1) * T, :]                        # peer stores when d is not dev0
            rc = lib.dlrmx_embbag_fwd_i32(ptrs.data_ptr(), 0, idx.contiguous().data_ptr(),
                                          off.contiguous().data_ptr(), idx.numel(), T, B, D, 0, 0,
                                          slot.data_ptr(), n * T * D, D, 0, st.cuda_stream)
            _lib.check(rc, "dlrmx_embbag_fwd_i32")
            ev = torch.cuda.Event()
            ev.record(st)
            done.append(ev)
    for ev in done:                                                    # no host sync: stream-ordered join
        torch.cuda.current_stream(dev0).wait_event(ev)
    return out


_lib_def.impl("forward", _forward_cuda, "CUDA")
