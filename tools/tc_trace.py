"""Pipeline timeline of CTA (0, 0) of welch_tc_kernel (debug hook pyfft_debug_set_tc_trace):
clock64 stamps of every role's hand-shakes, printed per batch relative to the first stamp."""
import ctypes
import sys
import numpy as np
import torch
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
from pyfft_b200 import _lib, engine
from pyfft_b200.windows import windows

dev = torch.device("cuda", 0)
L, H = 4096, 2048
n = 2 ** 26
win = windows("Hanning", nwins=L, verbose=False)
g = torch.Generator(device=dev).manual_seed(1)
xd = torch.randn(n, device=dev, generator=g)
yd = torch.randn(1, n, device=dev, generator=g)
K = (n - L) // H + 1
rmap = engine.record_map(n, 0, 0, 0, n)
wp = engine.WelchPlan(L, H, 1, "fp32", win, dev)
nb = 128
trace = torch.zeros(nb * 16 + 4 * 160, dtype=torch.int64, device=dev)
fn = _lib.lib.pyfft_debug_set_tc_trace
fn.argtypes = [ctypes.c_void_p]
fn.restype = ctypes.c_int
wp.prepare_tc(xd, yd, rmap, 0, K * H)
for rep in range(2):
    wp.accumulate_tc(xd, yd, rmap, 0, K)          # warm
torch.cuda.synchronize()
fn(trace.data_ptr())
wp.accumulate_tc(xd, yd, rmap, 0, K)
torch.cuda.synchronize()
first = trace.cpu().numpy().copy()
e = [torch.cuda.Event(enable_timing=True) for _ in range(41)]
e[0].record()
for i in range(40):
    wp.accumulate_tc(xd, yd, rmap, 0, K)
    e[i + 1].record()
torch.cuda.synchronize()
fn(None)
print("back-to-back accumulate_tc calls, ms each:", " ".join("%.3f" % e[i].elapsed_time(e[i + 1]) for i in range(40)))
print("(first block of numbers below: the 40th back-to-back launch; isolated launch: see the end)")
full = trace.cpu().numpy()
t = full[:nb * 16].reshape(nb, 16)
cta = full[nb * 16:].reshape(160, 4)[:148]
g0 = cta[:, 0].min()
print('CTA start (ns after first): min %d max %d; end: min %d max %d; duration ns: min %d mean %.0f max %d' % (
    (cta[:, 0] - g0).min(), (cta[:, 0] - g0).max(), (cta[:, 1] - g0).min(), (cta[:, 1] - g0).max(),
    (cta[:, 1] - cta[:, 0]).min(), (cta[:, 1] - cta[:, 0]).mean(), (cta[:, 1] - cta[:, 0]).max()))
clk = (cta[:, 3] - cta[:, 2]) / np.maximum(cta[:, 1] - cta[:, 0], 1)
print('SM clock during the kernel (clk/ns): min %.3f mean %.3f max %.3f' % (clk.min(), clk.mean(), clk.max()))
print('CTA 0: first event at clock %d after CTA start; CTA clocks %d' % (t[t > 0].min() - cta[0, 2], cta[0, 3] - cta[0, 2]))
t0 = t[t > 0].min()
names = ["C start", "C raw ok", "C A1 full", "S1 go", "S1 commit", "T got D1", "T B2 full", "S2 go", "S2 commit",
         "E got D2", "E D2 free", "P loaded"]
print("clock cycles relative to the first event; one row per segment")
print("  seg " + " ".join("%10s" % s for s in names))
for b in list(range(0, 8)) + list(range(90, 98)):
    print("%5d " % b + " ".join("%10d" % (t[b, e] - t0 if t[b, e] else -1) for e in range(12)))
d = np.diff(t[30:120, 8])
print("steady state: S2 commit to S2 commit %.0f clk per segment (min %d max %d)" % (d.mean(), d.min(), d.max()))
for e in range(12):
    print("  %-10s minus S2 commit of the same segment: %8.0f   (own period %.0f)" % (
        names[e], (t[30:120, e] - t[30:120, 8]).mean(), np.diff(t[30:120, e]).mean()))

cta1 = first[nb * 16:].reshape(160, 4)[:148]
print('isolated launch: CTA duration ns mean %.0f max %d, SM clock %.3f clk/ns' % (
    (cta1[:, 1] - cta1[:, 0]).mean(), (cta1[:, 1] - cta1[:, 0]).max(),
    ((cta1[:, 3] - cta1[:, 2]) / np.maximum(cta1[:, 1] - cta1[:, 0], 1)).mean()))
