"""MLP towers on the tensor cores, with the reference's construction (create_mlp, dlrm_s_pytorch.py:227-265) and
call shape (apply_mlp: `layers(x)`, 518-524).

`FusedMLP(sequential, ...)` takes over the Linear layers that create_mlp built (same numpy-initialised weights):
  * all weights and biases of the tower live in ONE flat fp32 master buffer (+ one flat gradient buffer, one flat
    Adagrad accumulator, one flat bf16 shadow that the GEMMs read);
  * forward per layer = one tcgen05 GEMM (csrc/gemm_tc.cu: TMA-staged bf16 operands, fp32 accumulators in TMEM),
    fp32 bias + ReLU in the epilogue;
  * backward of a run of layers (`_Tower`) = per layer one tcgen05 dgrad GEMM dz_prev = (dz W) * (y_prev > 0) whose
    epilogue applies the PREVIOUS layer's ReLU mask (no separate ReLU-backward pass), then ONE split-K tcgen05 launch
    for the weight AND bias gradients of every layer of the run (fp32, written straight into the flat gradient buffer);
  * `gemm="cublas"` keeps the round-1 path (cuBLASLt GEMMs through torch + dlrmx_relu_bwd_bias_grad) as the measured
    comparator of tools/gemm_check.py and bench.py --mlp fused-bf16-cublas; it is never a fallback;
  * `step()` = (all-reduce of the flat gradient buffer when data-parallel, the reference wraps the towers in DDP,
    dlrm_s_pytorch.py:1626-1633) + ONE kernel for the whole tower (dlrmx_dense_opt): SGD, or the dense Adagrad branch
    of RWSAdagrad (optim/rwsadagrad.py:115-118), which also refreshes the bf16 shadow and clears the gradients.
Layers whose output width is not a multiple of 8 (the final 256 -> 1 sigmoid layer) run as ordinary fp32 torch ops on
views of the same flat buffers.  Input widths are zero-padded to a multiple of 8 (13 -> 16, 479 -> 480).

Precision: bf16 operands / fp32 accumulation -- looser than the reference's fp32 cuBLAS and stated as such by
bench.py; parity tests use the unfused fp32 towers.
"""
import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch import nn

from . import _lib, ops


def _ceil(x, m):
    return (x + m - 1) // m * m


class _Layer:
    pass


class _LinearAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, h, anchor, mlp, li):
        L = mlp.layers[li]
        hb = h if h.dtype == torch.bfloat16 else h.to(torch.bfloat16)
        if hb.shape[1] != L.K_pad:
            padded = hb.new_zeros((hb.shape[0], L.K_pad))
            padded[:, :L.K].copy_(hb)
            hb = padded
        hb = hb.contiguous()
        if L.act == "relu":
            y = torch._addmm_activation(L.b_shadow, hb, L.w_shadow.t())
        else:
            y = torch.addmm(L.b_shadow, hb, L.w_shadow.t())
        ctx.mlp, ctx.li, ctx.in_cols = mlp, li, h.shape[1]
        # a consumer that fuses this layer's ReLU backward + bias gradient into its own backward (the head) sets the flag
        ctx.flag = {"pre_masked": False}
        mlp._last_relu = (li, ctx.flag) if L.act == "relu" else None
        ctx.save_for_backward(hb, y)
        return y

    @staticmethod
    def backward(ctx, dy):
        hb, y = ctx.saved_tensors
        L = ctx.mlp.layers[ctx.li]
        dyb = dy if dy.dtype == torch.bfloat16 else dy.to(torch.bfloat16)
        dyb = dyb.contiguous()
        acc = ctx.mlp.accumulate       # several forward/backward passes per step (micro-batches): add, don't overwrite
        if ctx.flag["pre_masked"]:
            dz = dyb                   # the head's backward applied this layer's ReLU mask and wrote its bias gradient
        else:
            gb = torch.empty_like(L.gb) if acc else L.gb
            dz, _ = ops.relu_bwd_bias_grad(dyb, y, relu=(L.act == "relu"), db=gb)
            if acc:
                L.gb.add_(gb)
        gW = torch.mm(dz.t(), hb, out_dtype=torch.float32)
        if acc:
            L.gW.add_(gW)
        else:
            L.gW.copy_(gW)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.mm(dz, L.w_shadow)
            if ctx.in_cols != L.K_pad:       # the caller's input was unpadded
                dx = dx[:, :L.K]
        return dx, None, None, None


# ---- deferred weight-gradient launches ------------------------------------------------------------------------------
# The top tower's backward runs first; launched in place, its wgrad kernel (tensor-bound, ~0.15 ms at the Terabyte shape)
# sits between the tower's dgrads and the interaction backward with nothing to overlap.  Deferred, it is launched right
# after the interaction backward, on the same stream, while the sparse update forks onto the side stream from the point
# BEFORE it: one persistent wgrad CTA per SM (200 KB of shared memory, 15 K registers) next to two or three CTAs of the
# update (HBM-bound, 16 bytes of shared memory each).  Under CUDA-graph capture the main chain has the higher stream
# priority (graph.py), so the GEMMs that follow get SM slots as the update's short-lived CTAs retire instead of waiting
# for its whole grid to be dispatched.  (Measured alternatives, Terabyte shape: wgrad on a third, high-priority stream
# with equal-priority main / side streams 1.196 ms -- the bottom tower's GEMMs starve behind the update's grid; the same
# with a high-priority main stream 1.240 ms -- the update starves while two high-priority grids share the SMs.)
_DEFERRED = []
_HOLD = [False]      # set by a consumer that will flush itself, at a point of its choosing (dlrm_net's sparse arch)


def flush_deferred(force=False):
    """Launches the deferred weight-gradient jobs, one split-K launch per tower.  (Both towers merged into ONE launch was
    measured at the Terabyte shape: 39 tiles split 7 ways instead of 34 x 2 + 5 x 14 write 70 MB of partials beside the
    HBM-bound sparse update -- step 1.189 -> 1.240 ms.)"""
    if _HOLD[0] and not force:
        return
    jobs = list(_DEFERRED)
    _DEFERRED.clear()
    for mlp, i0, dzs, xs, dws, dbs in jobs:
        ops.gemm_wgrad_bf16(dzs, xs, dws, dbs, mlp._wgrad_workspace(i0, i0 + len(dzs), dzs[0].shape[0]))
        if i0 == 0:                    # the tower's first layers come last in the backward: its gradients are complete
            mlp._start_allreduce()


class _Tower(torch.autograd.Function):
    """Layers [i0, i1) of a FusedMLP (all Linear + ReLU, widths multiples of 8) as ONE autograd node on the tcgen05 GEMMs."""

    @staticmethod
    def forward(ctx, h, anchor, mlp, i0, i1):
        first = mlp.layers[i0]
        if h.shape[1] != first.K_pad:           # cast and pad in one copy (the copy kernel rounds to nearest even, as .to() does)
            hb = torch.zeros((h.shape[0], first.K_pad), dtype=torch.bfloat16, device=h.device)
            hb[:, :first.K].copy_(h)
        else:
            hb = h if h.dtype == torch.bfloat16 else h.to(torch.bfloat16)
        acts, gates = [hb.contiguous()], []
        for li in range(i0, i1):
            L = mlp.layers[li]
            # every layer but the run's last also leaves its ReLU gates as one bit per element for the next layer's dgrad
            bits = ops.relu_bits_empty(hb.shape[0], L.N, hb.device) if li + 1 < i1 else None
            acts.append(ops.gemm_bf16(acts[-1], L.w_shadow, epilogue=ops.EPI_BIAS_RELU, bias=L.b_master, relu_bits_out=bits))
            gates.append(bits)
        ctx.mlp, ctx.i0, ctx.i1, ctx.in_cols = mlp, i0, i1, h.shape[1]
        # a consumer that fuses the last layer's ReLU backward + bias gradient into its own backward (the head) sets the flag
        ctx.flag = {"pre_masked": False}
        mlp._last_relu = (i1 - 1, ctx.flag)
        ctx.save_for_backward(*acts, *gates[:-1])
        return acts[-1]

    @staticmethod
    def backward(ctx, dy):
        mlp, i0, i1 = ctx.mlp, ctx.i0, ctx.i1
        acts, gates = ctx.saved_tensors[:i1 - i0 + 1], ctx.saved_tensors[i1 - i0 + 1:]
        layers = mlp.layers[i0:i1]
        n = len(layers)
        dyb = (dy if dy.dtype == torch.bfloat16 else dy.to(torch.bfloat16)).contiguous()
        dzs, dbs = [None] * n, [L.gb for L in layers]
        if ctx.flag["pre_masked"] == "nobias":
            dzs[-1] = dyb              # the consumer (interaction backward) applied the last ReLU mask; db comes from the wgrad
        elif ctx.flag["pre_masked"]:
            dzs[-1] = dyb              # the head's backward applied the last ReLU mask and wrote that bias gradient
            dbs[-1] = None
        else:
            dzs[-1], _ = ops.relu_bwd_bias_grad(dyb, acts[-1], relu=True, db=layers[-1].gb)
            dbs[-1] = None
        for j in range(n - 1, 0, -1):  # dz of layer j-1 = (dz_j W_j) masked by layer j-1's output
            dzs[j - 1] = ops.gemm_bf16(dzs[j], layers[j].w_shadow, b_is_kn=True, epilogue=ops.EPI_RELU_BITS, relu_mask=gates[j - 1])
        dx = None
        if ctx.needs_input_grad[0]:
            dx = ops.gemm_bf16(dzs[0], layers[0].w_shadow, b_is_kn=True)
            if ctx.in_cols != layers[0].K_pad:       # the caller's input was unpadded
                dx = dx[:, :layers[0].K]
        xs, dws, ws = list(acts[:-1]), [L.gW for L in layers], mlp._wgrad_workspace(i0, i1, dyb.shape[0])
        if mlp.defer_wgrad:
            # the weight gradients are needed only by step(): their launch waits for flush_deferred() -- called by the sparse
            # arch right after the interaction backward, or at the end of the pass (see there)
            _DEFERRED.append((mlp, i0, dzs, xs, dws, dbs))
            torch.autograd.Variable._execution_engine.queue_callback(flush_deferred)
        else:
            ops.gemm_wgrad_bf16(dzs, xs, dws, dbs, ws)
            if i0 == 0:                # the tower's first layers come last in the backward: its gradients are complete
                mlp._start_allreduce()
        return dx, None, None, None, None


class _Head(torch.autograd.Function):
    """Linear(K -> 1) + Sigmoid on a bf16 activation: one row-wise dot-product kernel each way (fp32 master weights)."""

    @staticmethod
    def forward(ctx, h, anchor, mlp, li):
        L = mlp.layers[li]
        h = h.contiguous()
        p = ops.head_fwd(h, L.w_master.view(-1), L.b_master, sigmoid=(L.act == "sigmoid"))
        ctx.mlp, ctx.li = mlp, li
        # the producer of h is the previous layer's Linear + ReLU: take over its ReLU backward and bias gradient
        last = getattr(mlp, "_last_relu", None)
        ctx.fuse_prev = bool(last is not None and last[0] == li - 1 and mlp.layers[li - 1].fused)
        if ctx.fuse_prev:
            last[1]["pre_masked"] = True
        ctx.save_for_backward(h, p)
        return p

    @staticmethod
    def backward(ctx, dp):
        h, p = ctx.saved_tensors
        L = ctx.mlp.layers[ctx.li]
        P = ctx.mlp.layers[ctx.li - 1] if ctx.fuse_prev else None
        if ctx.mlp.accumulate:
            gW, gb = torch.empty_like(L.gW), torch.empty_like(L.gb)
            gbp = torch.empty_like(P.gb) if P is not None else None
            dh = ops.head_bwd(dp, p, h, L.w_master.view(-1), gW.view(-1), gb, sigmoid=(L.act == "sigmoid"), gb_prev_relu=gbp)
            L.gW.add_(gW)
            L.gb.add_(gb)
            if P is not None:
                P.gb.add_(gbp)
        else:
            dh = ops.head_bwd(dp, p, h, L.w_master.view(-1), L.gW.view(-1), L.gb, sigmoid=(L.act == "sigmoid"),
                              gb_prev_relu=(P.gb if P is not None else None))
        return dh, None, None, None


class FusedMLP(nn.Module):
    def __init__(self, sequential, lr, optimizer="adagrad", eps=1e-10, initial_accumulator_value=0.0, gemm="tc"):
        super().__init__()
        if gemm not in ("tc", "cublas"):
            raise ValueError("gemm must be 'tc' (tcgen05 kernels) or 'cublas' (the library comparator)")
        self.gemm = gemm
        self.defer_wgrad = False    # True: the wgrad launch waits for flush_deferred() (see there)
        self.dp_world = 1           # > 1: data-parallel replicas; the backward pass starts the gradient all-reduce itself
        self._ar_stream = None
        self.peer_ar = None         # ext_dist.PeerAllReduce: the gradient buffer lives in peer-mapped memory (use_peer_allreduce)
        self._ws = {}
        mods = list(sequential)
        if len(mods) % 2 or not all(isinstance(m, nn.Linear) for m in mods[0::2]):
            raise ValueError("FusedMLP expects the Linear/activation pairs that create_mlp builds")
        self.lr, self.eps = float(lr), float(eps)
        self.mode = {"sgd": 0, "adagrad": 1, "rwsadagrad": 1}[optimizer]
        dev = mods[0].weight.device
        if dev.type != "cuda":
            raise _lib.DlrmxError("FusedMLP needs CUDA parameters (no CPU fallback)")
        self.layers, off = [], 0
        for lin, act in zip(mods[0::2], mods[1::2]):
            L = _Layer()
            L.N, L.K = lin.weight.shape
            L.act = "relu" if isinstance(act, nn.ReLU) else ("sigmoid" if isinstance(act, nn.Sigmoid) else None)
            if L.act is None:
                raise ValueError("unsupported activation %r" % act)
            L.fused = (L.N % 8 == 0) and L.act == "relu"
            # single-output head (the tower's last layer): row-wise dot-product kernels instead of a 1-column GEMM
            L.head = (not L.fused) and L.N == 1 and L.K % 8 == 0 and L.act == "sigmoid"
            L.K_pad = _ceil(L.K, 8) if L.fused else L.K
            L.w_off, off = off, _ceil(off + L.N * L.K_pad, 8)
            L.b_off, off = off, _ceil(off + L.N, 8)
            self.layers.append(L)
        self.numel = off
        master = torch.zeros(off, dtype=torch.float32, device=dev)
        for L, lin in zip(self.layers, mods[0::2]):
            master[L.w_off:L.w_off + L.N * L.K_pad].view(L.N, L.K_pad)[:, :L.K].copy_(lin.weight.data)
            master[L.b_off:L.b_off + L.N].copy_(lin.bias.data)
        self.master = nn.Parameter(master, requires_grad=False)       # state_dict: one flat fp32 tensor per tower
        self.register_buffer("acc", torch.full_like(master, float(initial_accumulator_value)))
        self.register_buffer("grad", torch.zeros_like(master), persistent=False)
        self.register_buffer("shadow", master.to(torch.bfloat16), persistent=False)
        self._anchor = torch.zeros(1, device=dev, requires_grad=True)  # makes every layer's backward run
        self.out_bf16 = False       # True: hand the last fused layer's bf16 output on as it is (no fp32 copy)
        self.accumulate = False     # True: the tower runs several times per step; gradients add up until step()
        self._bind()

    def _bind(self):
        for L in self.layers:
            nW = L.N * L.K_pad
            L.w_shadow = self.shadow[L.w_off:L.w_off + nW].view(L.N, L.K_pad)
            L.b_shadow = self.shadow[L.b_off:L.b_off + L.N]
            L.gW = self.grad[L.w_off:L.w_off + nW].view(L.N, L.K_pad)
            L.gb = self.grad[L.b_off:L.b_off + L.N]
            L.w_master = self.master.data[L.w_off:L.w_off + nW].view(L.N, L.K_pad)
            L.b_master = self.master.data[L.b_off:L.b_off + L.N]
            if not L.fused:     # plain fp32 torch layer on views of the flat buffers
                L.w_leaf = self.master.data[L.w_off:L.w_off + nW].view(L.N, L.K_pad).requires_grad_()
                L.b_leaf = self.master.data[L.b_off:L.b_off + L.N].requires_grad_()
                L.w_leaf.grad, L.b_leaf.grad = L.gW, L.gb

    def _apply(self, fn, *args, **kwargs):       # .to()/.cuda() re-allocate the flat buffers: re-create the views
        out = super()._apply(fn, *args, **kwargs)
        self._anchor = torch.zeros(1, device=self.master.device, requires_grad=True)
        self._ws = {}
        self._bind()
        return out

    def weight(self, i):
        """fp32 [N, K] view of layer i's weight (checkpoint / inspection)."""
        L = self.layers[i]
        return self.master.data[L.w_off:L.w_off + L.N * L.K_pad].view(L.N, L.K_pad)[:, :L.K]

    def bias(self, i):
        L = self.layers[i]
        return self.master.data[L.b_off:L.b_off + L.N]

    def _wgrad_workspace(self, i0, i1, M):
        key = (i0, i1, M)
        ws = self._ws.get(key)
        if ws is None:
            ls = self.layers[i0:i1]
            ws = self._ws[key] = ops.gemm_wgrad_workspace([L.N for L in ls], [L.K_pad for L in ls], M, self.master.device)
        return ws

    def forward(self, x):
        h, i, n = x, 0, len(self.layers)
        while i < n:
            L = self.layers[i]
            if L.fused and self.gemm == "tc":
                j = i
                while j < n and self.layers[j].fused:
                    j += 1
                h = _Tower.apply(h, self._anchor, self, i, j)
                i = j
                continue
            if L.fused:
                h = _LinearAct.apply(h, self._anchor, self, i)
            elif L.head and h.dtype == torch.bfloat16:
                h = _Head.apply(h, self._anchor, self, i)
            else:
                z = F.linear(h.float(), L.w_leaf, L.b_leaf)
                h = torch.sigmoid(z) if L.act == "sigmoid" else torch.relu(z)
            i += 1
        return h if (h.dtype == torch.float32 or self.out_bf16) else h.float()

    def use_peer_allreduce(self, par):
        """The flat gradient buffer moves into `par` (ext_dist.PeerAllReduce, fp32, numel = self.numel): the weight-gradient
        kernels write there and the data-parallel sum is the repo's own kernel over NVLink peer memory instead of NCCL.
        (.to() / .cuda() afterwards would move the gradients back into ordinary memory: call this last.)"""
        if par.buf.dtype != torch.float32 or par.buf.numel() != self.numel:
            raise _lib.DlrmxError("use_peer_allreduce: needs an fp32 buffer of %d elements" % self.numel)
        par.buf.copy_(self.grad)
        self.grad = par.buf
        self.peer_ar = par
        self._ws = {}
        self._bind()
        return self

    def _allreduce_grad(self):
        if self.peer_ar is not None:
            self.peer_ar.allreduce()
        else:
            dist.all_reduce(self.grad)

    def _start_allreduce(self):
        """Data-parallel replicas (dp_world > 1): the all-reduce of the flat gradient buffer starts as soon as the tower's
        backward has written it, on a communication stream beside the rest of the backward pass; step() joins it."""
        from . import dlrm_net
        if self.dp_world <= 1 or not dlrm_net.OVERLAP_STREAMS:       # single-stream mode: step() issues it, as before
            return
        cur = torch.cuda.current_stream()
        comm = dlrm_net._comm_stream(self.grad.device)
        if comm != cur:
            comm.wait_stream(cur)
        with torch.cuda.stream(comm):
            self._allreduce_grad()
        self._ar_stream = comm

    @torch.no_grad()
    def step(self, world_size=1):
        scale = 1.0
        if world_size > 1:
            if self._ar_stream is not None:   # started by the backward pass
                if self.dp_world != world_size:
                    raise _lib.DlrmxError("FusedMLP.step: dp_world (%d) != world_size (%d)" % (self.dp_world, world_size))
                torch.cuda.current_stream().wait_stream(self._ar_stream)
                self._ar_stream = None
            else:
                self._allreduce_grad()
            scale = 1.0 / world_size          # DDP averages the replicas' gradients
        ops.dense_opt(self.mode, self.master.data, self.acc if self.mode == 1 else None, self.grad, self.shadow,
                      self.lr, self.eps, scale)
