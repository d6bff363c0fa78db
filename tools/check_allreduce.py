"""Multi-GPU check of ext_dist.PeerAllReduce (csrc/peer_allreduce.cu) -- launch with torch.distributed.run, one rank per
GPU.  For fp32 and bf16 buffers of several sizes: every rank fills its symmetric buffer with seeded values, all-reduces
in place three times in a row (barrier reuse; each time from fresh operands), eagerly and as a replayed CUDA graph, and
compares bit for bit with the rank-order fp32 sum computed from an NCCL all-gather of the operands; NCCL's own all-reduce
of the same operands must agree within W roundings of the operands' magnitude sum (2^-23 fp32, 2^-8 bf16: a ring rounds
after every hop).  Prints "allreduce parity OK" per rank."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from dlrm_yx_b200 import extend_distributed as ext
    ext.init_distributed(use_gpu=True, backend="nccl")
    rank, W = ext.my_rank, ext.my_size
    dev = torch.device("cuda", torch.cuda.current_device())
    ok = True
    for td, numel in ((torch.float32, 2_400_004), (torch.float32, 12), (torch.bfloat16, 15_500_000), (torch.bfloat16, 8)):
        par = ext.PeerAllReduce(numel, td, dev)
        view = torch.int32 if td == torch.float32 else torch.int16
        operand = torch.empty(numel, dtype=td, device=dev)

        def fill(step):
            g = torch.Generator(device=dev).manual_seed(1000 * step + rank)
            operand.copy_((torch.randn(numel, device=dev, generator=g) * 2).to(td))

        def expected():
            parts = [torch.empty_like(operand) for _ in range(W)]
            dist.all_gather(parts, operand)
            acc, mag = parts[0].float(), parts[0].float().abs()
            for p in parts[1:]:
                acc = acc + p.float()
                mag = mag + p.float().abs()
            return acc.to(td), mag

        graph, static_in = None, torch.empty_like(operand)
        for step in range(6):
            fill(step)
            ref, mag = expected()
            if step < 3:
                par.buf.copy_(operand)
                par.allreduce()
            else:
                static_in.copy_(operand)
                if graph is None:
                    torch.cuda.synchronize()
                    dist.barrier()
                    graph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(graph):
                        par.buf.copy_(static_in)
                        par.allreduce()
                graph.replay()
            torch.cuda.synchronize()
            if not torch.equal(par.buf.view(view), ref.view(view)):
                bad = int((par.buf.view(view) != ref.view(view)).sum())
                print("allreduce FAILED rank %d: %s numel %d step %d: %d elements differ from the rank-order sum" % (rank, td, numel, step, bad),
                      flush=True)
                ok = False
            nccl = operand.clone()
            dist.all_reduce(nccl)
            # NCCL adds in another order and, for bf16, rounds after every hop: W roundings of partial sums bounded by mag
            tol = W * (2.0 ** -23 if td == torch.float32 else 2.0 ** -8)
            err = float(((nccl.float() - ref.float()).abs() / (mag + 1e-30)).max())
            if err > tol:
                print("allreduce FAILED rank %d: %s numel %d step %d: NCCL differs by %g of the magnitude sum (bound %g)"
                      % (rank, td, numel, step, err, tol), flush=True)
                ok = False
        # timing: 20 graph replays
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dist.barrier()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(20):
            graph.replay()
        e1.record()
        torch.cuda.synchronize()
        own = e0.elapsed_time(e1) / 20
        nccl = operand.clone()
        dist.barrier()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(20):
            dist.all_reduce(nccl)
        e1.record()
        torch.cuda.synchronize()
        if rank == 0:
            print("allreduce %s numel %d W %d: own %.1f us (copy-in + 2 barriers + kernel), NCCL %.1f us"
                  % (str(td).split(".")[-1], numel, W, own * 1e3, e0.elapsed_time(e1) / 20 * 1e3), flush=True)
    dist.barrier()
    if ok:
        print("allreduce parity OK (rank %d)" % rank, flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
