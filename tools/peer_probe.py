"""Probe: does torch symmetric memory (peer pointers over NVLink) work on this box?
torchrun --nproc-per-node 2 tools/peer_probe.py"""
import os
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
dev = torch.device("cuda", torch.cuda.current_device())
t = symm.empty(1 << 16, dtype=torch.float64, device=dev)
t.fill_(rank + 1.0)
hdl = symm.rendezvous(t, dist.group.WORLD.group_name)
print(rank, "rendezvous ok", "world", hdl.world_size, "ptrs", [hex(p) for p in hdl.buffer_ptrs],
      "signal pad bytes", hdl.signal_pad_size, "multicast", hex(hdl.multicast_ptr) if hdl.multicast_ptr else None, flush=True)
hdl.barrier(channel=0)
peer = hdl.get_buffer((rank + 1) % world, (1 << 16,), torch.float64)
torch.cuda.synchronize()
print(rank, "peer value", float(peer[0].item()), flush=True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
x = torch.ones(16448, dtype=torch.float64, device=dev)
for _ in range(5):
    dist.all_reduce(x)
torch.cuda.synchronize()
e0.record()
for _ in range(50):
    dist.all_reduce(x)
e1.record()
torch.cuda.synchronize()
print(rank, "nccl all_reduce of 131 KB: %.1f us" % (e0.elapsed_time(e1) * 1e3 / 50), flush=True)
dist.destroy_process_group()
