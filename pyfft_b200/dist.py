"""Multi-GPU decomposition of the Welch path: contiguous segment ranges per rank, a
read-only halo of ``nfft - hop`` samples at each boundary, no halo exchange, one
all-reduce of the packed partial sums.

Pure index arithmetic lives here so that it can be tested on CPU with gloo
(tests/test_dist_cpu.py); the device work goes through engine.WelchPlan.
"""


def segment_range(K, rank, world):
    """Segments [k0, k1) analysed by `rank` of `world` (contiguous, balanced to +-1)."""
    return K * rank // world, K * (rank + 1) // world


def analysed_span(k0, k1, hop, nfft):
    """Analysed-record indices [j_lo, j_hi) touched by segments [k0, k1)."""
    if k1 <= k0:
        return k0 * hop, k0 * hop
    return k0 * hop, (k1 - 1) * hop + nfft


def stats_span(k0, k1, K, hop, nana, rank, world):
    """Disjoint cover of the whole analysed record [0, nana) used for the global detrend
    sums (fft_analysis.py:356-357): rank r sums j in [k0*hop, k1*hop), the last rank runs
    to the end of the record (tail samples that no segment covers still enter the mean)."""
    j_lo = k0 * hop
    j_hi = nana if rank == world - 1 else k1 * hop
    if rank == 0:
        j_lo = 0
    return j_lo, max(j_lo, min(j_hi, nana))


def source_span(j_lo, j_hi, i0, reflect):
    """Global SOURCE sample indices [g_lo, g_hi) needed for analysed indices [j_lo, j_hi)
    under the record map g(j) = reflect - j (j < reflect) else j - reflect + i0."""
    if j_hi <= j_lo:
        return 0, 0
    lo, hi = [], []
    if j_lo < reflect:                       # mirrored part: g = reflect - j, decreasing in j
        jm = min(j_hi, reflect) - 1
        lo.append(reflect - jm)
        hi.append(reflect - j_lo)
    if j_hi > reflect:
        js = max(j_lo, reflect)
        lo.append(js - reflect + i0)
        hi.append(j_hi - 1 - reflect + i0)
    return min(lo), max(hi) + 1


def local_plan(K, hop, nfft, nana, i0, reflect, rank, world):
    """Everything one rank needs to know: its segments, the analysed span used for the
    detrend sums and the source slice [g_lo, g_hi) it must hold."""
    k0, k1 = segment_range(K, rank, world)
    a_lo, a_hi = analysed_span(k0, k1, hop, nfft)
    s_lo, s_hi = stats_span(k0, k1, K, hop, nana, rank, world)
    lo = min(a_lo, s_lo) if a_hi > a_lo else s_lo
    hi = max(a_hi, s_hi) if a_hi > a_lo else s_hi
    g_lo, g_hi = source_span(lo, hi, i0, reflect)
    return dict(k0=k0, k1=k1, seg_span=(a_lo, a_hi), stats_span=(s_lo, s_hi), src=(g_lo, g_hi))


# --------------------------------------------------------------------------- #
# channel sharding: spectrogram / STFT, Hilbert and ccf: independent
# channels (or pairs), contiguous channel blocks per rank, no data-path collective; an
# all-gather only when the caller asks for the assembled array
# --------------------------------------------------------------------------- #

def channel_range(C, rank, world):
    """Channels [c0, c1) handled by `rank` of `world` (contiguous, balanced to +-1)."""
    return C * rank // world, C * (rank + 1) // world


def group_rank_world(group):
    """(rank, world) of a torch.distributed group, (0, 1) without one."""
    if group is None:
        return 0, 1
    import torch.distributed as dist
    return dist.get_rank(group), dist.get_world_size(group)


def shard_channels(x, group, axis=0):
    """This rank's contiguous channel block of x along `axis` (a view)."""
    rank, world = group_rank_world(group)
    c0, c1 = channel_range(x.shape[axis], rank, world)
    return x.narrow(axis, c0, c1 - c0) if hasattr(x, "narrow") else x.take(range(c0, c1), axis=axis)


def gather_channels(local, C, group):
    """Assembles the (C, ...) array from the per-rank blocks [c0, c1) along axis 0: one
    all_gather of blocks padded to the largest one (the only collective of the channel-sharded
    paths, and only on request)."""
    import torch
    import torch.distributed as dist
    rank, world = group_rank_world(group)
    if world == 1:
        return local
    width = max(channel_range(C, r, world)[1] - channel_range(C, r, world)[0] for r in range(world))
    pad = torch.zeros((width,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[:local.shape[0]] = local
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    out = []
    for r in range(world):
        c0, c1 = channel_range(C, r, world)
        out.append(parts[r][:c1 - c0])
    return torch.cat(out, dim=0)
