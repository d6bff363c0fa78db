"""Model-parallel exchange of pooled embeddings with the reference's surface (extend_distributed.py):
module globals `my_size / my_rank / my_local_rank / my_local_size`, `init_distributed`, `get_my_slice`,
`get_split_lengths`, `alltoall(inputs, per_rank_table_splits, batched_emb) -> Request`, `Request.wait()`,
`all_gather`, `barrier`, `DDP`.

One process per GPU; transport is NCCL all-to-all over NVLink (torch.distributed plumbing).  What changes
against the reference (extend_distributed.py:405-508):
  * the batched lookup already writes [B, T_r, D], which IS the send buffer (concatenation over destination
    ranks of [B/W, T_r, D]) -- no torch.cat / flatten copy before the send;
  * `complete_into(recv, feat, feat_off)` (used by dlrm_net.s fast path) drops every source rank's chunk
    straight into the [B/W, 1+T, D] interaction input with ONE kernel (dlrmx_a2a_unpack) instead of W views
    + a cat, and the backward packs the feature-block gradient with dlrmx_a2a_pack_grad;
  * only the NCCL/Gloo all_to_all_single path exists (the scatter / scatter_list fallbacks for backends
    without all-to-all are out of scope).
Autograd behaviour is the reference's: forward batch-scatter / table-gather, backward the exact transpose,
asynchronous in both directions, one request outstanding (the global `myreq`).
"""
import os

import torch
import torch.distributed as dist
from torch.autograd import Function
from torch.autograd.profiler import record_function
from torch.nn.parallel import DistributedDataParallel as DDP  # noqa: F401  (re-export, as the reference does)

from . import ops

my_size = -1
my_rank = -1
my_local_rank = -1
my_local_size = -1
myreq = None


def _env_int(names, default=-1):
    for n in names:
        v = os.environ.get(n)
        if v is not None and int(v) >= 0:
            return int(v)
    return default


def get_my_slice(n):
    """Contiguous slice of n items owned by this rank; the first n % W ranks own one more
    (extend_distributed.py:47-55)."""
    base, extra = divmod(n, my_size)
    lo = my_rank * base + min(my_rank, extra)
    return slice(lo, lo + base + (1 if my_rank < extra else 0), 1)


def get_split_lengths(n):
    """(my length, per-rank lengths or None when n divides evenly) -- extend_distributed.py:58-66."""
    base, extra = divmod(n, my_size)
    if extra == 0:
        return base, None
    splits = [base + (1 if r < extra else 0) for r in range(my_size)]
    return splits[my_rank], splits


def init_distributed(rank=-1, local_rank=-1, size=-1, use_gpu=False, backend=""):
    """Rank discovery from the launcher environment (torchrun / MPI variables) and process-group
    creation: NCCL on GPU, Gloo otherwise (extend_distributed.py:81-207)."""
    global my_size, my_rank, my_local_rank, my_local_size, myreq
    if rank == -1:
        rank = _env_int(["PMI_RANK", "OMPI_COMM_WORLD_RANK", "MV2_COMM_WORLD_RANK", "RANK"], -1)
    if size == -1:
        size = _env_int(["PMI_SIZE", "OMPI_COMM_WORLD_SIZE", "MV2_COMM_WORLD_SIZE", "WORLD_SIZE"], -1)
    if local_rank == -1:
        local_rank = _env_int(["MPI_LOCALRANKID", "OMPI_COMM_WORLD_LOCAL_RANK", "MV2_COMM_WORLD_LOCAL_RANK",
                               "LOCAL_RANK"], 0)
    local_size = _env_int(["MPI_LOCALNRANKS", "OMPI_COMM_WORLD_LOCAL_SIZE", "MV2_COMM_WORLD_LOCAL_SIZE",
                           "LOCAL_WORLD_SIZE"], 1)
    if not backend:
        backend = "nccl" if use_gpu else "gloo"
    if size > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        os.environ["RANK"], os.environ["WORLD_SIZE"] = str(rank), str(size)
        if use_gpu:
            if local_size != torch.cuda.device_count() and local_size > torch.cuda.device_count():
                raise RuntimeError("%d local ranks but %d GPUs" % (local_size, torch.cuda.device_count()))
            torch.cuda.set_device(local_rank)
        if not dist.is_initialized():
            dist.init_process_group(backend, rank=rank, world_size=size)
        my_rank, my_size = dist.get_rank(), dist.get_world_size()
        my_local_rank, my_local_size = local_rank, local_size
    else:
        my_rank, my_size, my_local_rank, my_local_size = 0, 1, 0, 1
    myreq = Request()


class All2AllInfo(object):
    pass


class Request(object):
    """The one in-flight exchange (extend_distributed.py:210-220)."""

    def __init__(self):
        self.req = None
        self.tensor = None
        self.a2a_info = None

    def wait(self):
        """Tuple of W tensors: [B/W, T_p*D] (or [B/W, T_p, D] when batched)."""
        ret = _A2AWait.apply(*self.tensor)
        self.req = None
        self.tensor = None
        return ret


def _split_sizes(info):
    """Element counts per peer: forward receives lb*T_p*D from rank p and sends B_p*T_r*D to it."""
    recv = [info.local_batch_num * t * info.emb_dim for t in info.tables_per_rank]
    if info.global_batch_partition_slices:
        send = [b * info.local_table_num * info.emb_dim for b in info.global_batch_partition_slices]
    else:
        send = None
    return send, recv


class _A2AReq(Function):
    @staticmethod
    def forward(ctx, info, *inputs):
        global myreq
        send_split, recv_split = _split_sizes(info)
        with record_function("DLRM alltoall_req_fwd_single"):
            if info.batched_emb:
                send = inputs[0].contiguous().view(-1)              # [B, T_r, D] already is the send layout
            else:
                send = torch.cat(inputs, dim=1).view(-1)            # T_r x [B, D] -> [B, T_r*D]
            recv = send.new_empty([info.global_table_num * info.local_batch_num * info.emb_dim])
            myreq.req = dist.all_to_all_single(recv, send, recv_split, send_split, async_op=True)
            myreq.tensor = (recv,)
        info.send_split, info.recv_split = send_split, recv_split
        myreq.a2a_info = info
        ctx.info = info
        return myreq.tensor

    @staticmethod
    def backward(ctx, *unused):
        global myreq
        info = ctx.info
        with record_function("DLRM alltoall_req_bwd_single"):
            myreq.req.wait()
            myreq.req = None
            g = myreq.tensor
            myreq.tensor = None
        if g.dtype != torch.float32:           # bf16 on the wire (start_backward_from): exact widening
            g = g.float()
        if info.batched_emb:
            grads = [g.view(info.batch_size, -1, info.emb_dim)]
        else:
            grads = [c.contiguous() for c in g.view(info.batch_size, -1).split(info.emb_dim, dim=1)]
        return (None, *grads)


def _issue_backward(info, send):
    global myreq
    with record_function("DLRM alltoall_wait_bwd_single"):
        recv = send.new_empty([info.batch_size * info.local_table_num * info.emb_dim])
        myreq.req = dist.all_to_all_single(recv, send, info.send_split, info.recv_split, async_op=True)
        myreq.tensor = recv


class _A2AWait(Function):
    @staticmethod
    def forward(ctx, recv):
        global myreq
        info = myreq.a2a_info
        ctx.info = info
        with record_function("DLRM alltoall_wait_fwd_single"):
            myreq.req.wait()
            myreq.req = None
            myreq.tensor = None
        chunks = recv.split(info.recv_split)
        if info.batched_emb:
            return tuple(c.view(info.local_batch_num, -1, info.emb_dim) for c in chunks)
        return tuple(c.view(info.local_batch_num, -1) for c in chunks)

    @staticmethod
    def backward(ctx, *grads):
        send = torch.cat([g.contiguous().view(-1) for g in grads])
        _issue_backward(ctx.info, send)
        return (send,)


def complete_into(recv, feat, feat_off=1):
    """Low-level half of the fast path (no autograd): waits for the forward exchange and writes every
    source rank's [lb, T_p, D] chunk into feat[:, feat_off + t, :] (rank-major table order) with one
    kernel.  Returns the exchange descriptor needed by `start_backward_from`."""
    global myreq
    info = myreq.a2a_info
    with record_function("DLRM alltoall_wait_fwd_single"):
        myreq.req.wait()
        myreq.req = None
        myreq.tensor = None
    ops.a2a_unpack(recv, info.tables_per_rank, info.local_batch_num, info.emb_dim, feat, feat_off)
    return info


def start_backward_from(info, dfeat, feat_off=1, wire_bf16=False):
    """Packs dfeat[:, feat_off:, :] into per-destination chunks (one kernel) and starts the transposed
    exchange; `_A2AReq.backward` completes it.  Returns the send buffer (kept alive by the caller).
    wire_bf16: the chunks cross as bf16 (rounded to nearest even) and are widened on arrival -- the same values the
    peer-memory path's dlrmx_a2a_push_grad_bf16 delivers."""
    send = ops.a2a_pack_grad(dfeat, feat_off, info.tables_per_rank, info.local_batch_num, info.emb_dim)
    if wire_bf16:
        send = send.to(torch.bfloat16)
    _issue_backward(info, send)
    return send


def alltoall(inputs, per_rank_table_splits, batched_emb=False):
    """inputs: this rank's pooled embeddings for the FULL global batch -- a list of T_r tensors [B, D], or
    one [B, T_r, D] tensor when `batched_emb`; per_rank_table_splits[p] = tables held by rank p.
    Starts the exchange and returns the global Request (extend_distributed.py:601-639)."""
    global myreq
    info = All2AllInfo()
    if batched_emb:
        info.batch_size, info.local_table_num, info.emb_dim = inputs[0].size()
    else:
        info.batch_size, info.emb_dim = inputs[0].size()
        info.local_table_num = len(inputs)
    info.local_batch_num, info.global_batch_partition_slices = get_split_lengths(info.batch_size)
    info.tables_per_rank = (list(per_rank_table_splits) if per_rank_table_splits
                            else [info.local_table_num] * my_size)
    info.global_table_wise_partition_slices = per_rank_table_splits
    info.global_table_num = sum(info.tables_per_rank)
    info.batched_emb = batched_emb
    _A2AReq.apply(info, *inputs)
    return myreq


class PeerExchange(object):
    """The alltoall of pooled embeddings / gradients (extend_distributed.py:601-639) WITHOUT a collective: every
    rank's interaction input [lb, 1+T, D] and pooled-row gradient buffer [W*lb, T_r, D] live in peer-mapped
    (symmetric) memory, the lookup kernel stores each pooled row straight into the owner rank's interaction input
    (dlrmx_embbag_fwd_scatter) and the gradient slices are stored straight into the table owners' buffers
    (dlrmx_a2a_push_grad), all over NVLink.  Ranks meet at a device barrier before the first and after the last remote
    store of each direction.  One instance per model; buffers are allocated once.

    tables_per_rank[p] = SHARDED tables held by rank p (rank-major order, as the reference's cat produces);
    extra_features = feature slots behind them that every rank fills locally (replicated tables)."""

    def __init__(self, tables_per_rank, local_batch, emb_dim, device, group=None, feature_pitch=None, extra_features=0,
                 grad_bf16=False):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = dist.group.WORLD if group is None else group
        self.W, self.rank = my_size, my_rank
        self.tpr = [int(t) for t in tables_per_rank]
        self.T = sum(self.tpr)
        self.lb, self.D = int(local_batch), int(emb_dim)
        self.first = [sum(self.tpr[:p]) for p in range(self.W)]
        T_r = self.tpr[self.rank]
        F = self.F = 1 + self.T + int(extra_features)
        # row pitch of the features inside a sample (floats): D, or the interaction kernels' bulk-copy pitch
        # (ops.interact_feature_pitch) -- the remote row stores are addressed with strides, so any pitch works
        P = self.pitch = int(feature_pitch or self.D)
        n_feat = self.lb * F * P
        n_grad = self.W * self.lb * max(max(self.tpr), 1) * self.D     # symmetric: same size on every rank
        self._feat_flat = symm_mem.empty(n_feat, dtype=torch.float32, device=device)
        # grad_bf16 (the towers' mixed-precision mode): the pooled-row gradients cross NVLink and are stored as bf16 --
        # half the bytes of the transfer the sparse update waits for; the update kernel widens them exactly
        self.grad_bf16 = bool(grad_bf16)
        gsize = 2 if self.grad_bf16 else 4
        self._grad_flat = symm_mem.empty(n_grad, dtype=torch.bfloat16 if self.grad_bf16 else torch.float32, device=device)
        self.h_feat = symm_mem.rendezvous(self._feat_flat, self.group)
        self.h_grad = symm_mem.rendezvous(self._grad_flat, self.group)
        self.feat = self._feat_flat.view(self.lb, F, P)[:, :, :self.D]               # my interaction input
        self.grad = self._grad_flat[:self.W * self.lb * T_r * self.D].view(self.W * self.lb, T_r, self.D)
        # forward: where MY tables' rows go inside peer p's interaction input (feature slot 1 + first[rank])
        fptr = [int(q) + (1 + self.first[self.rank]) * P * 4 for q in self.h_feat.buffer_ptrs]
        self.out_peers = torch.tensor(fptr, dtype=torch.int64, device=device)
        # backward: where MY local batch's chunk starts inside owner p's gradient buffer [W*lb, T_p, D]
        self.push_bases = [int(q) + self.rank * self.lb * self.tpr[p] * self.D * gsize
                           for p, q in enumerate(self.h_grad.buffer_ptrs)]
        self.stride_b, self.stride_t = F * P, P

    def lookup_and_scatter(self, table_set, indices, offsets):
        """All local sharded tables x the global batch; every pooled row lands in its sample owner's `feat`.  On return
        (stream-ordered) every peer's rows have landed in MY `feat` as well."""
        with record_function("DLRM alltoall_req_fwd_single"):
            self.h_feat.barrier()                  # peers are done reading the previous contents of their feat
            if self.tpr[self.rank] > 0:
                ops.embbag_forward_scatter(table_set, indices, offsets, self.W * self.lb, self.out_peers, self.lb,
                                           self.stride_b, self.stride_t)

    def wait_rows(self):
        with record_function("DLRM alltoall_wait_fwd_single"):
            self.h_feat.barrier()                  # every peer's rows have landed in my feat

    def push_grad(self, dfeat):
        """dfeat [lb, F, D]: the sharded features 1 .. T go to the table owners' `grad` buffers; returns my buffer
        [W*lb, T_r, D] once every peer's slices have landed (stream-ordered)."""
        with record_function("DLRM alltoall_wait_bwd_single"):       # the reference's names: issue = wait_bwd ...
            self.h_grad.barrier()                  # owners are done with the previous step's gradients
            ops.a2a_push_grad(dfeat, 1, self.tpr, dfeat.shape[0], self.D, self.push_bases, bf16=self.grad_bf16)
        with record_function("DLRM alltoall_req_bwd_single"):        # ... completion = req_bwd (extend_distributed.py:449, 492)
            self.h_grad.barrier()
        return self.grad


class PeerAllReduce(object):
    """Sum of one flat buffer over the data-parallel replicas WITHOUT a collective library: the buffer lives in
    peer-mapped (symmetric) memory, rank r adds slice r of every rank's buffer in rank order and stores the sum into
    every rank's buffer (csrc/peer_allreduce.cu), between two device barriers.  Stands where DistributedDataParallel's
    all-reduce stands for the reference's towers (dlrm_s_pytorch.py:1606-1613).  In place on `self.buf` ([numel], fp32
    or bf16): producers write their operand there, `allreduce()` is stream-ordered, every rank ends with the same bits.
    One instance per buffer; allocate once (symmetric allocations are collective)."""

    def __init__(self, numel, dtype, device, group=None):
        if dtype not in (torch.float32, torch.bfloat16):
            raise ValueError("PeerAllReduce: fp32 or bf16 buffers")
        self.W, self.rank = max(my_size, 1), max(my_rank, 0)
        per16 = 8 if dtype == torch.bfloat16 else 4
        self.numel = int(numel)
        n = (self.numel + per16 - 1) // per16 * per16
        if self.W > 1:
            import torch.distributed._symmetric_memory as symm_mem
            self.group = dist.group.WORLD if group is None else group
            self._flat = symm_mem.empty(n, dtype=dtype, device=device)
            self._flat.zero_()
            self.h = symm_mem.rendezvous(self._flat, self.group)
            self.ptrs = [int(q) for q in self.h.buffer_ptrs]
        else:
            self._flat, self.h = torch.zeros(n, dtype=dtype, device=device), None
            self.ptrs = [self._flat.data_ptr()]
        self.buf = self._flat[:self.numel]

    def allreduce(self):
        with record_function("DLRM peer_allreduce"):
            if self.h is not None:
                self.h.barrier()               # every rank's operand is complete; nobody still reads the previous sum
            ops.peer_allreduce(self.ptrs, self.rank, self._flat.numel(), self._flat.dtype == torch.bfloat16)
            if self.h is not None:
                self.h.barrier()               # every remote store has landed
        return self.buf


def all_gather(input, lengths=None, dim=0):
    """Concatenation of every rank's tensor along `dim` (extend_distributed.py:642-645); `lengths[p]` gives
    rank p's extent when they differ."""
    if my_size <= 1:
        return input
    if dim != 0:
        return all_gather(input.transpose(0, dim).contiguous(), lengths, 0).transpose(0, dim)
    if lengths is None:
        parts = [torch.empty_like(input) for _ in range(my_size)]
    else:
        parts = [input.new_empty((int(n),) + tuple(input.shape[1:])) for n in lengths]
    dist.all_gather(parts, input.contiguous())
    return torch.cat(parts, dim=0)


def barrier():
    if my_size > 1:
        dist.barrier()
