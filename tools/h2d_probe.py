"""Probe: host->device bandwidth from pinned memory, one stream against two concurrent streams."""
import torch
dev = torch.device("cuda", 0)
n = 2 ** 26
hx = torch.randn(n).pin_memory(); hy = torch.randn(n).pin_memory()
dx = torch.empty(n, device=dev); dy = torch.empty(n, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(two, chunks=1):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if two:
        s1.wait_stream(torch.cuda.current_stream()); s2.wait_stream(torch.cuda.current_stream())
        c = n // chunks
        for i in range(chunks):
            with torch.cuda.stream(s1): dx[i*c:(i+1)*c].copy_(hx[i*c:(i+1)*c], non_blocking=True)
            with torch.cuda.stream(s2): dy[i*c:(i+1)*c].copy_(hy[i*c:(i+1)*c], non_blocking=True)
        torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
    else:
        dx.copy_(hx, non_blocking=True); dy.copy_(hy, non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)
for _ in range(3): run(False); run(True)
for two, ch in ((False, 1), (True, 1), (True, 4), (False, 1), (True, 1)):
    ms = min(run(two, ch) for _ in range(5))
    print("two streams" if two else "one stream ", "chunks", ch, "%.2f ms  %.1f GB/s" % (ms, 2 * n * 4 / ms / 1e6))
