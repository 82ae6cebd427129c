"""Config 5 on N GPUs of one box: 32 x 32 cross-spectral matrix, nfft 8192, 50 % overlap,
segments sharded in contiguous ranges, every rank holds only its slice (+ read-only halo), ONE
ncclAllReduce of the partial matrices (34 MB of complex128).
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/csm_shard_check.py
First a small record: the all-reduced matrix equals the single-GPU one; then the full size
(2^28 samples per channel in total), time = max over ranks with CUDA events."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pyfft_b200 import engine  # noqa: E402
from pyfft_b200.dist import local_plan  # noqa: E402
from pyfft_b200.windows import windows  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
dev = torch.device("cuda", torch.cuda.current_device())
C, L, hop = 32, 8192, 4096
win = windows("Hanning", nwins=L, verbose=False)


def sharded(x_local, nana, lp, G=None):
    g_lo, g_hi = lp["src"]
    rmap = engine.record_map(nana, 0, 0, g_lo, g_hi - g_lo)
    G = engine.csm_accumulate(x_local, rmap, L, hop, lp["k0"], lp["k1"] - lp["k0"], win, "fp32", G=G)
    dist.all_reduce(torch.view_as_real(G))
    return G


# ---- parity on a small record (same seed on every rank) ---------------------------------------- #
n = 1 << 19
g = torch.Generator(device=dev).manual_seed(55)
full = torch.randn((C, n), generator=g, device=dev)
K = (n - L) // hop + 1
ref = engine.csm_accumulate(full, engine.record_map(n, 0, 0, 0, n), L, hop, 0, K, win, "fp32")
lp = local_plan(K, hop, L, n, 0, 0, rank, world)
a_lo, a_hi = lp["seg_span"]
lp["src"] = (a_lo, max(a_hi, a_lo + 1))                     # no detrend here: hold the segment span only
Gs = sharded(full[:, lp["src"][0]:lp["src"][1]].contiguous(), n, lp)
err = float((Gs - ref).abs().max() / ref.abs().max())
ok = err <= 2e-6
if rank == 0:
    print("sharded + all-reduce vs single GPU on %d ranks: rel. difference %.1e" % (world, err), flush=True)
del full, ref, Gs

# ---- full size: 2^28 samples per channel in total ----------------------------------------------- #
n = 1 << 28
K = (n - L) // hop + 1
lp = local_plan(K, hop, L, n, 0, 0, rank, world)
a_lo, a_hi = lp["seg_span"]
lp["src"] = (a_lo, a_hi)
g = torch.Generator(device=dev).manual_seed(5 + rank)
x = torch.randn((C, a_hi - a_lo), generator=g, device=dev)
G = torch.zeros((L // 2 + 1, C, C), dtype=torch.complex128, device=dev)


def step():
    G.zero_()
    sharded(x, n, lp, G=G)


for _ in range(2):
    step()
ts = []
for _ in range(3):
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    step()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ts.append(float(t.item()))
if rank == 0:
    ms = float(np.median(ts))
    print(json.dumps({"case": "c5 cross-spectral matrix 32 x 2^28, segments sharded over %d GPUs + all-reduce" % world,
                      "n_gpus": world, "ms": round(ms, 3), "msamples_per_s": round(C * n / ms / 1e3, 1),
                      "segments_per_gpu": lp["k1"] - lp["k0"], "collective": "ncclAllReduce of 34 MB (complex128 matrix)"}),
          flush=True)
dist.destroy_process_group()
sys.exit(0 if ok else 1)
