"""Times pyfft_welch_accumulate_tc on the bench workload (2 x 2^26, nfft 4096, 50 %): median of 40 back-to-back
calls, CUDA events.  PYFFT_B200_LIB=<path> selects an A/B build of the library."""
import os
import sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pyfft_b200 import engine
from pyfft_b200.windows import windows

dev = torch.device("cuda", 0)
L, n = 4096, 2 ** 26
H = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
win = windows("Hanning", nwins=L, verbose=False)
g = torch.Generator(device=dev).manual_seed(1)
xd = torch.randn(n, device=dev, generator=g) + 7.0
yd = torch.randn(1, n, device=dev, generator=g) + 2.5
K = (n - L) // H + 1
rmap = engine.record_map(n, 0, 0, 0, n)
wp = engine.WelchPlan(L, H, 1, "fp32", win, dev)
wp.prepare_tc(xd, yd, rmap, 0, K * H)
for _ in range(5):
    wp.accumulate_tc(xd, yd, rmap, 0, K)
e = [torch.cuda.Event(enable_timing=True) for _ in range(41)]
e[0].record()
for i in range(40):
    wp.accumulate_tc(xd, yd, rmap, 0, K)
    e[i + 1].record()
torch.cuda.synchronize()
if not os.environ.get("PYFFT_B200_LIB"): assert not wp.tc_failed()
ms = np.array([e[i].elapsed_time(e[i + 1]) for i in range(40)])
print("%s hop %d: accumulate_tc median %.4f ms (min %.4f max %.4f) = %.1f%% of 6542 GB/s" % (
    os.environ.get("PYFFT_B200_LIB", "in-tree"), H, np.median(ms), ms.min(), ms.max(),
    100 * 2 * n * 4 / (np.median(ms) * 1e-3) / 6542.4e9))
