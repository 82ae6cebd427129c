"""Smallest program that launches welch_tc_kernel on the bench workload (2 x 2^26, nfft 4096, 50 %): the
command profiled with ncu (python tools/tc_prof_driver.py [launches])."""
import os
import sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pyfft_b200 import engine
from pyfft_b200.windows import windows

dev = torch.device("cuda", 0)
L, H, n = 4096, 2048, 2 ** 26
win = windows("Hanning", nwins=L, verbose=False)
g = torch.Generator(device=dev).manual_seed(1)
xd = torch.randn(n, device=dev, generator=g) + 7.0
yd = torch.randn(1, n, device=dev, generator=g) + 2.5
K = (n - L) // H + 1
rmap = engine.record_map(n, 0, 0, 0, n)
wp = engine.WelchPlan(L, H, 1, "fp32", win, dev)
wp.prepare_tc(xd, yd, rmap, 0, K * H)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 4):
    wp.accumulate_tc(xd, yd, rmap, 0, K)
torch.cuda.synchronize()
assert not wp.tc_failed()
print("ok")
