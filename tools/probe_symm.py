"""Probe: does torch symmetric memory (peer-mapped buffers over NVLink) work on this box?  torchrun, 2+ GPUs."""
import os, time
import torch
import torch.distributed as dist

rank = int(os.environ["RANK"]); W = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
import torch.distributed._symmetric_memory as symm_mem
print(rank, "symm_mem imported; backends:", [n for n in dir(symm_mem) if "backend" in n.lower()], flush=True)
n = 64 << 20
t = symm_mem.empty(n, dtype=torch.float32, device="cuda")
hdl = symm_mem.rendezvous(t, dist.group.WORLD)
print(rank, "rendezvous ok", type(hdl).__name__, "ptrs", [hex(p) for p in hdl.buffer_ptrs][:4], "signal", hex(hdl.signal_pad_ptrs[0]), flush=True)
t.fill_(float(rank))
hdl.barrier()
peer = hdl.get_buffer((rank + 1) % W, (n,), torch.float32)
src = torch.full((n,), 100.0 + rank, device="cuda")
torch.cuda.synchronize()
for it in range(3):
    hdl.barrier()
    t0 = time.time()
    peer.copy_(src)            # P2P store into the neighbour's buffer
    torch.cuda.synchronize()
    dt = time.time() - t0
    hdl.barrier()
    torch.cuda.synchronize()
    print(rank, "iter", it, "push %.1f GB/s" % (n * 4 / dt / 1e9), "mine now", t[:2].tolist(), flush=True)
dist.destroy_process_group()
