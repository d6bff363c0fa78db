"""Whole-step CUDA graph: forward + backward + sparse update + tower optimizer captured once, replayed per batch.

The training step of this path is ~80 short kernels (lookup, interaction, a dozen GEMMs, fused element-wise passes,
sort, update) of 10-200 us each; launched one by one from Python the GPU idles between them.  Capturing the step
removes the launch gaps and fixes the cross-stream overlap (sort / sparse update beside the towers) once and for all.

Requirements (checked or documented): static shapes -- the packed index count N must not change between batches
(fixed bag lengths, e.g. the one-index-per-lookup Criteo layout); constant learning rates (they are baked into the
kernel arguments at capture); no host synchronisation inside the step.  The warm-up runs before capture ARE real
training steps on the example batch.
"""
import torch


_CAPTURE = {}


def _capture_stream():
    dev = torch.cuda.current_device()
    s = _CAPTURE.get(dev)
    if s is None:
        s = _CAPTURE[dev] = torch.cuda.Stream(priority=-1)
    return s


class GraphedTrainStep(object):
    def __init__(self, step_fn, example_inputs, warmup=3, pool=None):
        """step_fn(*tensors) -> loss tensor, a full training step using only the tensors it is given (plus module
        state).  example_inputs: device tensors of the shapes / dtypes every later batch will have."""
        self.static = [t.clone() for t in example_inputs]
        # parameters created on the default stream accumulate their gradients on the capture stream here: intended
        quiet = getattr(torch.autograd.graph, "set_warn_on_accumulate_grad_stream_mismatch", None)
        if quiet is not None:
            quiet(False)
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(max(int(warmup), 1)):
                step_fn(*self.static)
        cur.wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        # The step's main chain is captured on a HIGH-priority stream (kernel nodes inherit it); the sparse update forks onto a
        # default-priority side stream.  A CTA of a running grid is never preempted, but whenever a slot frees up the
        # scheduler serves the higher priority first: the tensor-core GEMMs of the main chain (one 226 KB CTA per SM) get SM
        # slots as the update's short-lived CTAs retire instead of waiting for its whole grid to be dispatched.  One stream
        # per device for every capture of the process (as torch's own default capture stream): autograd remembers the stream
        # a leaf first accumulated its gradient on, and a later capture on another stream would depend on that one.
        with torch.cuda.graph(self.graph, pool=pool, stream=_capture_stream()):
            self.loss = step_fn(*self.static)

    def prefetch(self, *host_inputs):
        """Starts the host-to-device copies of the NEXT batch on a copy stream (pinned host tensors), into a staging
        set, so they run beside the step that is executing; `run_prefetched()` then moves them into the graph inputs
        with device-to-device copies (a few microseconds) and replays."""
        if not hasattr(self, "_copy_stream"):
            self._copy_stream = torch.cuda.Stream()
            self._staging = [torch.empty_like(t) for t in self.static]
            self._staged = torch.cuda.Event()
            self._consumed = torch.cuda.Event()
            self._consumed.record()
        self._copy_stream.wait_event(self._consumed)          # the previous staging contents have been consumed
        with torch.cuda.stream(self._copy_stream):
            for dst, src in zip(self._staging, host_inputs):
                dst.copy_(src, non_blocking=True)
            self._staged.record()

    def run_prefetched(self):
        cur = torch.cuda.current_stream()
        cur.wait_event(self._staged)
        for dst, src in zip(self.static, self._staging):
            dst.copy_(src, non_blocking=True)
        self._consumed.record()
        self.graph.replay()
        return self.loss

    def load(self, *inputs):
        """Stream-ordered copies of a batch (device tensors, or pinned host tensors) into the graph's inputs."""
        for dst, src in zip(self.static, inputs):
            if dst.shape != src.shape or dst.dtype != src.dtype:
                raise ValueError("graphed step needs static shapes: got %s %s, captured %s %s"
                                 % (tuple(src.shape), src.dtype, tuple(dst.shape), dst.dtype))
            dst.copy_(src, non_blocking=True)

    def __call__(self, *inputs):
        self.load(*inputs)
        self.graph.replay()
        return self.loss
