"""Sensitivity of welch_tc_kernel to each role's work (debug switch PYFFT_TC_ABLATE, results invalid):
bit 0 loaders idle, bit 1 twiddle idle, bit 2 epilogue reads D2 without arithmetic, bit 3 epilogue frees
D2 at once.  python tools/tc_ablate.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pyfft_b200 import engine  # noqa: E402
from pyfft_b200.windows import windows  # noqa: E402

dev = torch.device("cuda", 0)
L, hop, n = 4096, 2048, 2 ** 26
K = n // hop - 1
win = windows("Hanning", nwins=L, verbose=False)
g = torch.Generator(device=dev).manual_seed(1)
x = torch.randn(n, generator=g, device=dev) + 7.0
y = (torch.randn(n, generator=g, device=dev) + 2.5).view(1, n)
rmap = engine.record_map(n, 0, L - 1, 0, n)
wp = engine.WelchPlan(L, hop, 1, "fp32", win, dev)
wp.prepare_tc(x, y, rmap, 0, K * hop)
for abl in (0, 1, 2, 4, 8, 3, 5, 6, 9, 10, 7, 11):
    os.environ["PYFFT_TC_ABLATE"] = str(abl)
    ts = []
    for it in range(13):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        wp.accumulate_tc(x, y, rmap, 0, K)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    what = [nm for b, nm in ((1, "loaders idle"), (2, "twiddle idle"), (4, "epilogue loads only"), (8, "epilogue frees at once")) if abl & b]
    print("ablate %2d  %.4f ms  %s%s" % (abl, float(np.median(ts[3:])), ", ".join(what) or "full kernel",
                                       "  [timed out]" if wp.tc_failed() else ""), flush=True)
