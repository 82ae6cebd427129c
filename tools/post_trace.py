"""Where the sharded step's exchange spends its time: globaltimer stamps of CTA 0 of tc_post_kernel (debug hook
pyfft_debug_set_post_trace) on every rank, back-to-back steps as in bench.py.

    python -m torch.distributed.run --nproc-per-node N tools/post_trace.py [--sync]
"""
import ctypes as C
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from pyfft_b200 import _lib, engine  # noqa: E402
from pyfft_b200 import dist as pdist  # noqa: E402
from pyfft_b200.fft_analysis import fftanal  # noqa: E402
from pyfft_b200.windows import windows  # noqa: E402

sync = "--sync" in sys.argv
world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", 1), ("RANK", 0), ("LOCAL_RANK", 0)))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
group = None
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
    group = dist.group.WORLD
NFFT, HOP, FS = bench.NFFT, bench.HOP, bench.FS
n, ntot = 2 ** 26, world * 2 ** 26
K, nana, refl = ntot // HOP - 1, ntot, NFFT - 1
lp = pdist.local_plan(K, HOP, NFFT, nana, 0, refl, rank, world)
g_lo, g_hi = lp["src"]
x = bench.make_slice(torch, dev, g_lo, g_hi, 11, 7.0, 1.0, 0.0)
y = bench.make_slice(torch, dev, g_lo, g_hi, 12, 2.5, 0.5, -np.pi / 4)[None, :].contiguous()
rmap = engine.record_map(nana, 0, refl, g_lo, g_hi - g_lo)
win = windows("Hanning", nwins=NFFT, verbose=False)
S1, S2, _, ENBW = fftanal._getNorms(win, NFFT // 2, FS)
wp = engine.WelchPlan(NFFT, HOP, 1, "fp32", win, dev)
k0, k1 = lp["k0"], lp["k1"]
s_lo, s_hi = lp["stats_span"]
tails = ((s_lo, k0 * HOP), (k1 * HOP, s_hi))


def step():
    return wp.step_tc(x, y, rmap, k0, k1 - k0, tails, K, nana, 1.0 / S1 ** 2 / ENBW, True, group,
                      deferred=(world > 1 and not sync))


trace = torch.zeros(8, dtype=torch.int64, device=dev)
for _ in range(5):
    step()
torch.cuda.synchronize()
run = engine.GraphedStep(step)
_lib.lib.pyfft_debug_set_post_trace.argtypes = [C.c_void_p]
_lib.lib.pyfft_debug_set_post_trace(C.c_void_p(trace.data_ptr()))
runt = engine.GraphedStep(step)                     # same step, its post kernel stamps
_lib.lib.pyfft_debug_set_post_trace(C.c_void_p(0))
acc = np.zeros(8)
reps = 20
for _ in range(reps):
    if group is not None:
        dist.barrier(group=group)
    for _ in range(10):
        run()
    runt()
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    t = trace.cpu().numpy().astype(np.float64)
    acc += t - t[1]
acc /= reps * 1e3
names = ["start", "K1 done", "old collected", "(phase 1)", "reduced", "pushed", "new collected", "end"]
rows = torch.tensor(acc, device=dev)
allr = [torch.zeros_like(rows) for _ in range(world)]
if group is not None:
    dist.all_gather(allr, rows, group=group)
else:
    allr = [rows]
if rank == 0:
    print("mode:", "sync" if sync or world == 1 else "deferred", "world", world, "(us relative to 'K1 done')")
    for r, a in enumerate(allr):
        print("rank %d: " % r + "  ".join("%s %.2f" % (nm, v) for nm, v in zip(names, a.cpu().numpy())))
# step time of the same replay, 200 back to back
for _ in range(3):
    if group is not None:
        dist.barrier(group=group)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(200):
        run()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / 200], device=dev, dtype=torch.float64)
    each = [torch.zeros_like(ms) for _ in range(world)]
    if group is not None:
        dist.all_gather(each, ms, group=group)
    else:
        each = [ms]
    if rank == 0:
        print("ms per step, per rank:", " ".join("%.4f" % float(v) for v in each))
if group is not None:
    wp.flush_tc(force=True)
    torch.cuda.synchronize()
    dist.destroy_process_group()
