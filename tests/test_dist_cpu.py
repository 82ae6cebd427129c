"""CPU, world_size 2, gloo: the multi-GPU decomposition of the Welch path (segment ranges,
halo spans, disjoint detrend-statistics spans, ONE all-reduce of the partial sums).  The
per-rank arithmetic is done by the numpy oracle here; on the GPU box the same index plan
(pyfft_b200/dist.py) drives the CUDA kernels."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def _worker(rank, world, port, n, L, hop, i0, reflect, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import restatement as R
    from pyfft_b200.dist import local_plan
    rng = np.random.default_rng(9)
    src = rng.standard_normal((n, 2)) + np.array([3.0, -1.0])      # full record, same on each rank
    nana = n - i0 if not reflect else n
    K = (nana - (L - hop)) // hop
    lp = local_plan(K, hop, L, nana, i0, reflect, rank, world)
    g_lo, g_hi = lp["src"]
    local = src[g_lo:g_hi].copy()                                   # the only samples this rank holds

    def ana(j):                                                     # analysed index -> local value
        j = np.asarray(j)
        g = np.where(j < reflect, reflect - j, j - reflect + i0)
        assert g.min() >= g_lo and g.max() < g_hi
        return local[g - g_lo]

    s_lo, s_hi = lp["stats_span"]
    stats = torch.zeros(2, dtype=torch.float64)
    if s_hi > s_lo:
        stats += torch.from_numpy(ana(np.arange(s_lo, s_hi)).sum(0))
    cnt = torch.tensor([s_hi - s_lo], dtype=torch.int64)
    dist.all_reduce(stats)
    dist.all_reduce(cnt)
    assert int(cnt) == nana                                         # disjoint cover of the record
    mean = stats.numpy() / nana
    win = R.window_vector("Hanning", L)
    acc = torch.zeros((3, L), dtype=torch.float64)
    k0, k1 = lp["k0"], lp["k1"]
    if k1 > k0:
        a_lo, a_hi = lp["seg_span"]
        seg = ana(np.arange(a_lo, a_hi)) - mean
        sxx, syy, sxy = R.welch_sums(seg[:, 0], seg[:, 1:2], win, k1 - k0, hop)
        acc[0] += torch.from_numpy(sxx)
        acc[1] += torch.from_numpy(syy[0])
        acc[2] += torch.from_numpy(sxy[0].real)
    dist.all_reduce(acc)                                            # the single collective
    if rank == 0:
        full = src if not reflect else R.reflect_pad(src, L)[:n]
        full = full[i0:] if not reflect else full
        full = full - full.mean(0)
        sxx, syy, sxy = R.welch_sums(full[:, 0], full[:, 1:2], win, K, hop)
        ref = np.stack([sxx, syy[0], sxy[0].real])
        q.put(float(np.max(np.abs(acc.numpy() - ref)) / np.max(np.abs(ref))))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n,L,hop,i0,reflect", [(20000, 256, 128, 0, 255), (9000, 128, 32, 17, 0),
                                                (700, 256, 128, 0, 255)])
def test_two_rank_sharding_equals_single_pass(n, L, hop, i0, reflect):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() + n) % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, L, hop, i0, reflect, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) < 1e-12


def test_local_plan_covers_everything():
    from pyfft_b200.dist import local_plan
    for world in (1, 2, 3, 8):
        for (K, hop, L, tail, i0, refl) in [(32767, 2048, 4096, 0, 0, 4095), (5, 8, 16, 3, 11, 0),
                                            (3, 512, 2048, 100, 0, 2047), (1, 4, 16, 0, 0, 0)]:
            nana = (K - 1) * hop + L + tail
            segs, cover = 0, []
            for r in range(world):
                lp = local_plan(K, hop, L, nana, i0, refl, r, world)
                segs += lp["k1"] - lp["k0"]
                cover.append(lp["stats_span"])
                if lp["k1"] > lp["k0"]:
                    assert lp["seg_span"][1] <= nana
            assert segs == K
            assert cover[0][0] == 0 and cover[-1][1] == nana
            assert all(cover[i][1] == cover[i + 1][0] for i in range(world - 1))


def _channel_worker(rank, world, port, C, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from pyfft_b200.dist import channel_range, gather_channels, shard_channels
    full = torch.arange(C * 6, dtype=torch.float64).reshape(C, 2, 3)       # same on each rank
    mine = shard_channels(full, dist.group.WORLD)
    c0, c1 = channel_range(C, rank, world)
    ok = mine.shape[0] == c1 - c0 and torch.equal(mine, full[c0:c1])
    back = gather_channels(mine * 1.0, C, dist.group.WORLD)                # the one optional collective
    ok = ok and torch.equal(back, full)
    q.put((rank, bool(ok), (c0, c1)))
    dist.destroy_process_group()


@pytest.mark.parametrize("C", [64, 5, 1])
def test_channel_sharding_two_ranks(C):
    """STFT / Hilbert / ccf shard by channel: contiguous blocks that cover every
    channel once, unequal blocks and empty ranks included; gather restores the array."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + C
    procs = [ctx.Process(target=_channel_worker, args=(r, 2, port, C, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert all(r[1] for r in res), res
    assert res[0][2][0] == 0 and res[0][2][1] == res[1][2][0] and res[1][2][1] == C


def test_channel_range_partitions():
    from pyfft_b200.dist import channel_range
    for C in (1, 7, 16, 64):
        for world in (1, 2, 4, 8):
            spans = [channel_range(C, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == C
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            assert max(b - a for a, b in spans) - min(b - a for a, b in spans) <= 1
