import os, sys, time
import numpy as np, torch
sys.path.insert(0, "/root/repo")
import pyfft_b200.fft_analysis as fa
from pyfft_b200 import fft_pwelch
n = 2 ** 26
xh = (torch.randn(n) + 7).pin_memory(); yh = (torch.randn(n) + 2.5).pin_memory()
tv = np.arange(n) / 1.0e6
kw = dict(Navr=2 * (n // 4096) - 1, windowoverlap=0.5, windowfunction="Hanning", plotit=False, detrend_style=1, precision="fp32")
marks = {}
orig = fa._host_many
def hm(t):
    marks["pre"] = time.perf_counter()
    r = orig(t)
    marks["post"] = time.perf_counter()
    return r
fa._host_many = hm
for _ in range(3):
    fft_pwelch(tv, xh, yh, **kw)
rows = []
for _ in range(8):
    torch.cuda.synchronize()
    t0 = time.perf_counter(); out = fft_pwelch(tv, xh, yh, **kw); t1 = time.perf_counter()
    rows.append(((marks["pre"] - t0) * 1e3, (marks["post"] - marks["pre"]) * 1e3, (t1 - marks["post"]) * 1e3, (t1 - t0) * 1e3))
for r in rows:
    print("host until the packed read is issued %.3f ms | waiting for the GPU + D2H %.3f | host bookkeeping after %.3f | total %.3f" % r)
print("tensor_cores", out[6].tensor_cores)
import cProfile, pstats
pr = cProfile.Profile(); pr.enable()
for _ in range(5): fft_pwelch(tv, xh, yh, **kw)
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(18)
