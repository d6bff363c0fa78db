"""Probe: copy-engine (cudaMemcpy2DAsync) throughput into a peer-mapped buffer for the exchange's pitched blocks."""
import ctypes, os, time
import torch
import torch.distributed as dist

rank = int(os.environ["RANK"]); W = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
import torch.distributed._symmetric_memory as symm_mem
rt = ctypes.CDLL("libcudart.so.12")
rt.cudaMemcpy2DAsync.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_size_t,
                                 ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
lb, F, D = 32768, 27, 128
feat = symm_mem.empty(lb * F * D, dtype=torch.float32, device="cuda")
hdl = symm_mem.rendezvous(feat, dist.group.WORLD)
for Tr in (3, 13):
    src = torch.randn(W * lb, Tr, D, device="cuda")
    peer = (rank + 1) % W
    dst = hdl.buffer_ptrs[peer] + 1 * D * 4
    s = torch.cuda.Stream()
    for it in range(4):
        hdl.barrier()
        torch.cuda.synchronize()
        t0 = time.time()
        with torch.cuda.stream(s):
            rc = rt.cudaMemcpy2DAsync(dst, F * D * 4, src[peer * lb:].data_ptr(), Tr * D * 4, Tr * D * 4, lb, 4,
                                      s.cuda_stream)   # cudaMemcpyDefault
        s.synchronize()
        dt = time.time() - t0
        hdl.barrier()
        torch.cuda.synchronize()
        if it:
            print(rank, "Tr", Tr, "rc", rc, "2D peer copy %.1f MB in %.3f ms = %.1f GB/s" % (lb * Tr * D * 4 / 1e6, dt * 1e3, lb * Tr * D * 4 / dt / 1e9), flush=True)
    # check it landed
    hdl.barrier()
    mine = feat.view(lb, F, D)
    torch.cuda.synchronize()
dist.destroy_process_group()
