"""Times pyfft_welch_accumulate_tc on the bench workload for every misalignment (in floats, against
16 bytes) of the first analysed sample.  python tools/tc_align_bench.py"""
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
xb = torch.randn(n + 8, generator=g, device=dev) + 7.0
yb = torch.randn(n + 8, generator=g, device=dev) + 2.5
for refl in (L - 1, 0):
    for shift in range(4):
        x = xb[shift:shift + n]
        y = yb[shift:shift + n].view(1, n)
        rmap = engine.record_map(n, 0, refl, 0, n)
        wp = engine.WelchPlan(L, hop, 1, "fp32", win, dev)
        wp.prepare_tc(x, y, rmap, 0, K * hop)
        ts = []
        for it in range(13):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            wp.accumulate_tc(x, y, rmap, 0, K)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        assert not wp.tc_failed()
        m = ((x.data_ptr() // 4) - refl) % 4
        print("reflect %4d shift %d -> misalignment %d floats: %.4f ms (median of last 10)" % (refl, shift, m, float(np.median(ts[3:]))))
