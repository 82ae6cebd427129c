"""Where the end-to-end time of fft_pwelch goes (config 2, pinned host inputs): device timeline of the
call's phases reproduced with the engine API (CUDA events), next to the wall time of the real call."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pyfft_b200 import engine, fft_pwelch
from pyfft_b200.windows import windows

dev = torch.device("cuda", 0)
n, L, hop = 2 ** 26, 4096, 2048
xh = torch.randn(n).pin_memory(); yh = torch.randn(n).pin_memory()
tv = np.arange(n) / 1.0e6
kw = dict(Navr=32767, windowoverlap=0.5, windowfunction="Hanning", plotit=False, detrend_style=1, precision="fp32")
for _ in range(3):
    fft_pwelch(tv, xh, yh, **kw)
torch.cuda.synchronize()
ws = []
for _ in range(5):
    t0 = time.perf_counter(); fft_pwelch(tv, xh, yh, **kw); torch.cuda.synchronize(); ws.append((time.perf_counter() - t0) * 1e3)
print("fft_pwelch wall ms (5 calls):", " ".join("%.2f" % w for w in ws))

win = windows("Hanning", nwins=L, verbose=False)
K = 32767
refl = L - 1
rmap = engine.record_map(n, 0, refl, 0, n)
def once():
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
    ev[0].record()
    x = xh.to(dev, non_blocking=True); y = yh.to(dev, non_blocking=True).view(1, n)
    ev[1].record()
    wp = engine.WelchPlan(L, hop, 1, "fp32", win, dev)
    ev[2].record()
    wp.prepare_tc(x, y, rmap, 0, K * hop)
    wp.accumulate_tc(x, y, rmap, 0, K)
    wp.detrend_stats(x, y, rmap, K * hop, n)
    res = wp.finalize(1.0, onesided=True, K=K, correct_nana=n)
    ev[3].record()
    out = [res[k].cpu() for k in ("Pxx", "Pyy", "Pxy", "Cxy", "Cxy2", "phi_xy")]
    ev[4].record()
    torch.cuda.synchronize()
    return [ev[i].elapsed_time(ev[i + 1]) for i in range(4)]
for _ in range(2):
    once()
t0 = time.perf_counter(); r = once(); w = (time.perf_counter() - t0) * 1e3
print("phases ms: copies %.2f  plan (window upload etc.) %.2f  kernels %.2f  device->host %.2f   wall %.2f" % (*r, w))
