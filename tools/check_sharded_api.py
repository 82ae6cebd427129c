"""Parity of the SHARDED public call on N GPUs of one box (the fused step with the NVLink exchange inside its last
kernel) against the same call on ONE GPU and against the float64 oracle:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/check_sharded_api.py

Every rank holds only its slice of the record (dist.local_plan(...)["src"]) and calls
fft_pwelch(group=, shard_offset=); rank 0 also analyses the whole record alone.  Checked: sharded == one GPU within the
fp32 contract (1e-5 of the spectral peak, 1e-4 coherence), identical on all ranks bit for bit, tensor-core route
(K >= 16384) and CUDA-core route (small K), two y channels."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pyfft_b200 import fft_pwelch  # noqa: E402
from pyfft_b200 import dist as pdist  # noqa: E402
from pyfft_b200.fft_analysis import _Plan  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = torch.device("cuda", torch.cuda.current_device())
dist.init_process_group("nccl", device_id=dev)
group = dist.group.WORLD
FS = 1.0e6
bad = 0


class Uniform(object):
    def __init__(self, m):
        self.m = m

    def __len__(self):
        return self.m

    def __getitem__(self, i):
        return (i if i >= 0 else self.m + i) / FS


for tag, n, navr, nch in (("tensor cores, K = 32767", 2 ** 26, 32767, 2), ("CUDA cores, K = 2047", 2 ** 22, 2047, 1)):
    rng = np.random.default_rng(99)
    t = np.arange(n) / FS
    x = (np.sin(2 * np.pi * 9.7e4 * t) + 0.5 * rng.standard_normal(n) + 3.0).astype(np.float32)
    ys = [(0.6 * np.sin(2 * np.pi * 9.7e4 * t + 0.4 * (c + 1)) + 0.4 * rng.standard_normal(n) - 1.0 * c).astype(np.float32)
          for c in range(nch)]
    y = ys[0] if nch == 1 else np.stack(ys, axis=0)                      # channels first (nch, n)
    kw = dict(Navr=navr, windowoverlap=0.5, windowfunction="Hanning", plotit=False, detrend_style=1, precision="fp32")
    plan = _Plan(Uniform(n), n, n, None, navr, 0.5, "Hanning", {})
    hop = plan.nwins - plan.noverlap
    refl = plan.nwins - 1 if plan.reflect else 0
    lp = pdist.local_plan(plan.Navr, hop, plan.nwins, plan.nana, plan.i0, refl, rank, world)
    g_lo, g_hi = lp["src"]
    xs = torch.from_numpy(x[g_lo:g_hi].copy()).to(dev)
    ysl = torch.from_numpy(np.ascontiguousarray(y[..., g_lo:g_hi])).to(dev)
    out = fft_pwelch(Uniform(n), xs, ysl, group=group, shard_offset=g_lo, **kw)
    info = out[6]
    mine = torch.from_numpy(np.concatenate([np.asarray(out[i]).astype(np.complex128).ravel() for i in (1, 2, 3, 4)])).to(dev)
    both = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(both, mine)
    same = all(torch.equal(both[0], b) for b in both)
    if rank == 0:
        one = fft_pwelch(t, torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev), **kw)
        pk = float(np.max(np.abs(one[2])))
        e = [float(np.max(np.abs(np.asarray(out[i]) - np.asarray(one[i])))) for i in (1, 2, 3)]
        ec = float(np.max(np.abs(np.asarray(out[4]) - np.asarray(one[4]))))
        ok = same and max(e) <= 1e-5 * pk and ec <= 1e-4
        bad += 0 if ok else 1
        print("%-26s N=%d nwins %d: sharded vs one GPU  PSD/CSD %.2e of peak, coherence %.2e, identical on all "
              "ranks: %s, tensor cores %s/%s -> %s" % (tag, world, plan.nwins, max(e) / pk, ec, same,
                                                       info.tensor_cores, one[6].tensor_cores, "OK" if ok else "MISMATCH"),
              flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(1 if bad else 0)
