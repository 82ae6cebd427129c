"""Channel-sharded configurations 3 and 4 on N GPUs of one box (no data-path collective):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/channel_shard_check.py
Checks every rank's block against the unsharded result on a small record (and the gathered array),
then times the full-size configurations: every rank transforms its block of the 64 (or 16) channels,
time = max over ranks (CUDA events), one JSON line per case on rank 0."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pyfft_b200 import ccf, engine, hilbert  # noqa: E402
from pyfft_b200.dist import channel_range  # noqa: E402
from pyfft_b200.windows import windows  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
dev = torch.device("cuda", torch.cuda.current_device())
G = dist.group.WORLD

# ---- parity of the sharded calls on a small record ------------------------------------------- #
rng = np.random.default_rng(11)
C, n = 6, 1 << 14
u = rng.standard_normal((C, n)).astype(np.float32)
c0, c1 = channel_range(C, rank, world)
z_all = hilbert(u, return_device=True)
z_loc = hilbert(u, return_device=True, group=G)
z_gat = hilbert(u, return_device=True, group=G, gather=True)
# (which records share a transform depends on the block, so blocks agree to rounding, not bit for bit)
ok = float((z_loc.reshape(-1, n) - z_all[c0:c1]).abs().max()) <= 1e-5 if c1 > c0 else z_loc.numel() == 0
ok = ok and float((z_gat - z_all).abs().max()) <= 1e-5        # pairing of records differs per block: rounding only
_, co_all = ccf(u[0], u[1:], 1.0e6, return_device=True)
_, co_gat = ccf(u[0], u[1:], 1.0e6, return_device=True, group=G, gather=True)
ok = ok and float((co_gat - co_all).abs().max()) <= 1e-5
L, hop = 256, 64
K = (n - L) // hop + 1
win = windows("Hanning", nwins=L, verbose=False)
ud = torch.from_numpy(u).to(dev)
rmap = engine.record_map(n, 0, 0, 0, n)
s_all, _ = engine.stft(ud, rmap, L, hop, 0, K, win, "fp32", nbins=L // 2)
s_gat, _ = engine.stft_sharded(ud, G, rmap, L, hop, 0, K, win, "fp32", gather=True, nbins=L // 2)
ok = ok and torch.equal(s_gat, s_all)
flag = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("sharded == unsharded on %d ranks: %s" % (world, bool(flag.item())), flush=True)


def timed(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dist.barrier()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ts.append(float(t.item()))
    return float(np.median(ts))


# ---- config 3: 64 channels x 2^24, nfft 2048, 75 % overlap, sharded by channel ---------------- #
C, n, L, hop = 64, 1 << 24, 2048, 512
c0, c1 = channel_range(C, rank, world)
g = torch.Generator(device=dev).manual_seed(3 + rank)
x = torch.randn((c1 - c0, n), generator=g, device=dev)          # this rank's channels only
K = (n - L) // hop + 1
win = windows("Hanning", nwins=L, verbose=False)
rmap = engine.record_map(n, 0, 0, 0, n)
hold = {}
ms = timed(lambda: hold.__setitem__("o", engine.stft(x, rmap, L, hop, 0, K, win, "fp32", nbins=L // 2)))
if rank == 0:
    nbytes = C * n * 4 + C * K * (L // 2) * 8
    print(json.dumps({"case": "c3 stft complex64, 64 channels sharded over %d GPUs" % world, "n_gpus": world,
                      "ms": round(ms, 4), "msamples_per_s": round(C * n / ms / 1e3, 1),
                      "algorithmic_gbs_all_gpus": round(nbytes / ms / 1e6, 1), "collective": "none"}), flush=True)
hold.clear(); del x
torch.cuda.empty_cache()

# ---- config 4: Hilbert envelope of 16 channels x 2^25, sharded by channel --------------------- #
C, n = 16, 1 << 25
c0, c1 = channel_range(C, rank, world)
uu = torch.randn((c1 - c0, n), generator=g, device=dev)
ms = timed(lambda: hold.__setitem__("o", engine.hilbert(uu, "fp32", want_analytic=False, want_envelope=True)))
if rank == 0:
    print(json.dumps({"case": "c4 hilbert envelope, 16 channels sharded over %d GPUs" % world, "n_gpus": world,
                      "ms": round(ms, 4), "msamples_per_s": round(C * n / ms / 1e3, 1), "collective": "none"}), flush=True)
dist.destroy_process_group()
sys.exit(0 if bool(flag.item()) else 1)
