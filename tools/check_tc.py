"""GPU check of the tensor-core Welch kernel against the CUDA-core kernel (same acc layout)
and the oracle."""
import sys
import numpy as np
import torch
sys.path.insert(0, '/root/repo')
from pyfft_b200 import _lib, engine
from pyfft_b200.windows import windows
from oracle import signals

dev = torch.device("cuda", 0)
L, H = 4096, 2048
win = windows("Hanning", nwins=L, verbose=False)
for n, dc in ((2 ** 16, (0.0, 0.0)), (2 ** 17, (7.0, 2.5)), (2 ** 20, (7.0, -2.5))):
    t, x, y = signals.c2_cross(n=n, dc=dc, dtype=np.float32)
    xd = torch.from_numpy(x).to(dev)
    yd = torch.from_numpy(y).to(dev)[None, :].contiguous()
    K = (n - L) // H + 1
    rmap = engine.record_map(n, 0, 0, 0, n)
    res = {}
    for mode in ("cuda", "tc"):
        wp = engine.WelchPlan(L, H, 1, "fp32", win, dev)
        if mode == "cuda":
            wp.estimate_mean(xd, yd, rmap, 0, K * H)
            wp.accumulate_mean(xd, yd, rmap, 0, K)
        else:
            assert wp.can_use_tc
            wp.prepare_tc(xd, yd, rmap, 0, K * H)
            wp.accumulate_tc(xd, yd, rmap, 0, K)
            print("  prescale", wp.prescale.cpu().numpy(), "est", wp.est.cpu().numpy()[::2], "tc_failed", wp.tc_failed())
        if K * H < n:
            wp.detrend_stats(xd, yd, rmap, K * H, n)
        wp.mean_correct(n)
        res[mode] = wp.acc.cpu().numpy()[0]
    a, b = res["cuda"], res["tc"]
    for q in range(4):
        den = np.abs(a[q]).max() if q < 2 else np.abs(a[2] + 1j * a[3]).max()
        err = np.abs(a[q] - b[q]).max() / den
        print("n=2^%d K=%d q=%d max rel err vs cuda-core kernel %.3e (argmax bin %d)" % (
            int(np.log2(n)), K, q, err, int(np.argmax(np.abs(a[q] - b[q])))))
