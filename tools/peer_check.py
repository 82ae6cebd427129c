"""NVLink peer reduction (csrc/peer.cu) against NCCL on N GPUs of one box:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/peer_check.py
Checks: equal to the rank-ordered float64 sum bit for bit, identical on all ranks, repeated epochs
(double-buffered staging), and times both routes with CUDA events."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pyfft_b200 import engine  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
dev = torch.device("cuda", torch.cuda.current_device())
n = 2 * 4 * 2049 + 2 * 4 * 256 + 8                      # acc | lin | stats of config 2 (and then some)
pr = engine.PeerReducer.create(n, dev, None)
assert pr, "peer reduction unavailable"
ok = True
for epoch in range(6):
    g = torch.Generator(device=dev).manual_seed(100 * epoch + rank)
    mine = torch.randn(n, dtype=torch.float64, device=dev, generator=g) * (10.0 ** (rank - 1))
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine)
    want = torch.zeros_like(mine)
    for r in range(world):
        want = want + gathered[r]                        # rank order, like the kernel
    state = mine.clone()
    pr.all_reduce(state)
    torch.cuda.synchronize()
    pr.check_status()
    same = bool(torch.equal(state, want))
    ref = mine.clone()
    dist.all_reduce(ref)
    rel = float((state - ref).abs().max() / ref.abs().max())
    ok = ok and same and rel < 1e-14
    if rank == 0:
        print("epoch %d: bit-identical to the rank-ordered sum: %s; vs NCCL rel %.1e" % (epoch, same, rel), flush=True)


def timed(fn, reps=200):
    for _ in range(10):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / reps], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


state = torch.ones(n, dtype=torch.float64, device=dev)
t_peer = timed(lambda: pr.all_reduce(state))
pr.check_status()
t_nccl = timed(lambda: dist.all_reduce(state))
if rank == 0:
    print("%d GPUs, %d doubles (%.0f KB): peer kernel %.1f us, NCCL all_reduce %.1f us per call (max over ranks)%s"
          % (world, n, n * 8 / 1024, t_peer, t_nccl, "" if ok else "  ** MISMATCH **"), flush=True)
dist.destroy_process_group()
sys.exit(0 if ok else 1)
