"""GPU parity of the tensor-core Welch kernel (pyfft_welch_accumulate_tc, tcgen05) against the fp32
register-FFT kernel (pyfft_welch_accumulate_mean) on the same device buffers, at the engine level:
ragged shards, odd and tiny segment counts, several hops, several y channels, the reflect-padded
zone, and the oracle at a size it finishes in seconds.

The tensor-core path rounds every windowed sample to fp16 once, independently from segment to
segment, so its deviation from the fp32 kernel falls like 1/sqrt(K): measured 2.3e-5 of the
spectral peak at K = 31 and 9e-7 at K = 511.  The bound used here is 6e-4 / sqrt(K) (a factor ~4.5
above the measurement); fft_pwelch only selects the path for K >= 16384, where that is 4.7e-6 and
the 1e-5 contract of BASELINE.json holds with margin (tests/test_gpu_fullsize.py checks it there).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

L = 4096


def _signals(n, nch, seed, dc=True):
    rng = np.random.default_rng(seed)
    t = np.arange(n) / 1.0e5
    x = np.sin(2 * np.pi * 2441.40625 * t) + 0.3 * rng.standard_normal(n)
    ys = []
    for c in range(nch):
        ys.append(0.5 * np.roll(x, 3 + c) + 0.2 * rng.standard_normal(n) + 0.1 * np.sin(2 * np.pi * 7000.0 * t + c))
    y = np.stack(ys)
    if dc:
        x = x + 5.0
        y = y - 2.5 * (1 + np.arange(nch))[:, None]
    return x.astype(np.float32), y.astype(np.float32)


def _both_routes(dev, x, y, hop, rmap_args, seg_first, seg_count, span, shift=(0, 0)):
    """acc (after the mean correction) of the CUDA-core and of the tensor-core route.
    shift: the device records start shift[0] (x) / shift[1] (y) floats past a 16-byte boundary."""
    import torch
    from pyfft_b200 import engine
    from pyfft_b200.windows import windows
    win = windows("Hanning", nwins=L, verbose=False)
    xbuf = torch.zeros(x.size + 8, dtype=torch.float32, device=dev)
    xd = xbuf[shift[0]:shift[0] + x.size]
    xd.copy_(torch.from_numpy(x))
    ybuf = torch.zeros(y.size + 8, dtype=torch.float32, device=dev)
    yd = ybuf[shift[1]:shift[1] + y.size].view(y.shape)
    yd.copy_(torch.from_numpy(y))
    assert xd.data_ptr() % 16 == 4 * shift[0] and yd.data_ptr() % 16 == 4 * shift[1] and yd.is_contiguous()
    rmap = engine.record_map(*rmap_args)
    nana = rmap_args[0]
    out = {}
    for mode in ("cuda", "tc"):
        wp = engine.WelchPlan(L, hop, y.shape[0], "fp32", win, dev)
        assert wp.can_use_tc
        lo, hi = span
        if mode == "cuda":
            wp.estimate_mean(xd, yd, rmap, lo, hi)
            wp.accumulate_mean(xd, yd, rmap, seg_first, seg_count)
        else:
            wp.prepare_tc(xd, yd, rmap, lo, hi)
            wp.accumulate_tc(xd, yd, rmap, seg_first, seg_count)
            assert not wp.tc_failed()
        # samples of the record that the segment range does not cover
        c_lo, c_hi = seg_first * hop, (seg_first + seg_count) * hop
        if c_lo > 0:
            wp.detrend_stats(xd, yd, rmap, 0, c_lo)
        if c_hi < nana:
            wp.detrend_stats(xd, yd, rmap, c_hi, nana)
        wp.mean_correct(nana)
        out[mode] = wp.acc.cpu().numpy()
    return out["cuda"], out["tc"]


def _assert_close(a, b, K):
    tol = 6e-4 / np.sqrt(K)
    for p in range(a.shape[0]):
        den_auto = [np.abs(a[p, 0]).max(), np.abs(a[p, 1]).max()]
        den_x = np.abs(a[p, 2] + 1j * a[p, 3]).max()
        for q, den in ((0, den_auto[0]), (1, den_auto[1]), (2, den_x), (3, den_x)):
            err = np.abs(a[p, q] - b[p, q]).max() / den
            assert err <= tol, (p, q, err, tol)


@pytest.mark.parametrize("hop", [2048, 1024, 512, 4096])
def test_hops(hop, cuda_device):
    n = 2 ** 18
    x, y = _signals(n, 1, 11)
    K = (n - L) // hop + 1
    a, b = _both_routes(cuda_device, x, y, hop, (n, 0, 0, 0, n), 0, K, (0, min(n, 65536)))
    _assert_close(a, b, K)


@pytest.mark.parametrize("seg_first,seg_count", [(0, 1), (0, 2), (0, 3), (5, 7), (3, 148), (1, 149), (0, 297), (17, 1000)])
def test_ragged_shards(seg_first, seg_count, cuda_device):
    """Odd counts leave a half-empty last batch; counts below the SM count leave idle CTAs."""
    hop = 2048
    n = 2 ** 22
    x, y = _signals(n, 1, 12)
    a, b = _both_routes(cuda_device, x, y, hop, (n, 0, 0, 0, n), seg_first, seg_count, (0, 65536))
    _assert_close(a, b, seg_count)


def test_three_y_channels_with_offsets(cuda_device):
    hop = 2048
    n = 2 ** 20
    x, y = _signals(n, 3, 13)
    K = (n - L) // hop + 1
    a, b = _both_routes(cuda_device, x, y, hop, (n, 0, 0, 0, n), 0, K, (0, 65536))
    assert a.shape[0] == 3
    _assert_close(a, b, K)


@pytest.mark.parametrize("seg_first,seg_count", [(0, 600), (0, 1), (1, 2), (2, 301)])
def test_reflect_padded_zone(seg_first, seg_count, cuda_device):
    """fft_pwelch analyses the reflect-padded record (fft_analysis.py:196-205): the first
    nfft - 1 analysed samples are a mirror image; the kernel reads them from a materialised
    copy.  Shards that start inside, at the edge of and behind the zone."""
    hop = 2048
    n = 2 ** 21
    refl = L - 1
    x, y = _signals(n, 2, 14)
    nana = n                                   # analysed length (<= n + refl)
    a, b = _both_routes(cuda_device, x, y, hop, (nana, 0, refl, 0, n), seg_first, seg_count, (0, 65536))
    _assert_close(a, b, seg_count)


def test_sharded_sum_equals_one_call(cuda_device):
    """Two shards of the segment range add up to the single-call result (multi-GPU partition,
    SURVEY section 8(e)): sums of both routes are linear in the segment set."""
    import torch
    from pyfft_b200 import engine
    from pyfft_b200.windows import windows
    dev = cuda_device
    hop, n = 2048, 2 ** 20
    x, y = _signals(n, 1, 15)
    K = (n - L) // hop + 1
    win = windows("Hanning", nwins=L, verbose=False)
    xd, yd = torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev).contiguous()
    rmap = engine.record_map(n, 0, 0, 0, n)

    def run(parts):
        wp = engine.WelchPlan(L, hop, 1, "fp32", win, dev)
        wp.prepare_tc(xd, yd, rmap, 0, 65536)
        for k0, k in parts:
            wp.accumulate_tc(xd, yd, rmap, k0, k)
        assert not wp.tc_failed()
        if K * hop < n:
            wp.detrend_stats(xd, yd, rmap, K * hop, n)
        wp.mean_correct(n)
        return wp.acc.cpu().numpy()

    one = run([(0, K)])
    two = run([(0, 201), (201, K - 201)])
    den = np.abs(one).max()
    # same per-segment arithmetic, different partial-sum order only
    assert np.abs(one - two).max() / den <= 2e-6


def test_against_the_oracle(cuda_device):
    """Tensor-core route through fft_pwelch vs the float64 oracle on a record the oracle finishes in
    seconds (K = 2047 segments: bound 6e-4 / sqrt(K) = 1.3e-5 of the peak; the coherence bound of
    1e-4 at the K >= 16384 where the path is selected scales to 1e-4 * sqrt(16384 / 2047) = 2.8e-4)."""
    from oracle import restatement as R
    from pyfft_b200 import fft_pwelch
    n = 2 ** 22
    x, y = _signals(n, 1, 16)
    t = np.arange(n) / 1.0e5
    kw = dict(Navr=2047, windowoverlap=0.5, windowfunction="Hanning", detrend_style=1)
    ref = R.fft_pwelch(t, x.astype(np.float64), y[0].astype(np.float64), **kw)
    out = fft_pwelch(t, x, y[0], plotit=False, tensor_cores=True, **kw)
    assert out[6].tensor_cores and out[6].nwins == L
    tol = 6e-4 / np.sqrt(2047)
    for i, nm in ((1, "Pxy"), (2, "Pxx"), (3, "Pyy")):
        err = np.max(np.abs(out[i] - ref[nm])) / np.max(np.abs(ref[nm]))
        assert err <= tol, (nm, err)
    assert np.max(np.abs(out[4] - ref["Cxy"])) <= 1e-4 * np.sqrt(16384 / 2047)


@pytest.mark.parametrize("route", ["tc", "cuda"])
def test_rereferenced_shards_one_collective(route, cuda_device):
    """Sharded fused-mean route with ONE collective (pyfft_welch_rereference): two shards analysed
    about different provisional means are moved to the common reference 0, their packed states
    (acc | lin | stats) added -- what the all-reduce does -- and corrected inside finalize; equals the
    single-shard result (mean_correct + finalize) to double-precision rounding of the correction."""
    import torch
    from pyfft_b200 import engine
    from pyfft_b200.windows import windows
    dev = cuda_device
    hop, n = 2048, 2 ** 20
    x, y = _signals(n, 2, 17)
    x = (x + np.linspace(0.0, 3.0, n)).astype(np.float32)       # drift: the shards' provisional means differ
    K = (n - L) // hop + 1
    win = windows("Hanning", nwins=L, verbose=False)
    xd, yd = torch.from_numpy(x).to(dev), torch.from_numpy(y).to(dev).contiguous()
    rmap = engine.record_map(n, 0, 0, 0, n)

    def shard(k0, k1, tail):
        wp = engine.WelchPlan(L, hop, 2, "fp32", win, dev)
        if route == "tc":
            wp.prepare_tc(xd, yd, rmap, k0 * hop, k1 * hop)
            wp.accumulate_tc(xd, yd, rmap, k0, k1 - k0)
            assert not wp.tc_failed()
        else:
            wp.estimate_mean(xd, yd, rmap, k0 * hop, k1 * hop)
            wp.accumulate_mean(xd, yd, rmap, k0, k1 - k0)
        if tail and k1 * hop < n:
            wp.detrend_stats(xd, yd, rmap, k1 * hop, n)
        return wp

    one = shard(0, K, True)
    one.mean_correct(n)
    ref = one.finalize(1.0, onesided=True, K=K)

    a, b = shard(0, 201, False), shard(201, K, True)
    assert abs(float(a.est[0]) - float(b.est[0])) > 0.5
    a.rereference()
    b.rereference()
    a.state += b.state                                          # the all-reduce
    got = a.finalize(1.0, onesided=True, K=K, correct_nana=n)
    for nm in ("Pxx", "Pyy", "Pxy"):
        den = float(ref[nm].abs().max())
        err = float((got[nm] - ref[nm]).abs().max()) / den
        # the re-referencing itself is exact in double.  CUDA-core route: only the order of the fp32
        # partial sums differs; tensor-core route: another provisional mean means another fp16
        # rounding of every sample, i.e. the route's own 1/sqrt(K) noise
        assert err <= (2e-6 if route == "cuda" else 6e-4 / np.sqrt(K)), (nm, err)
    # tensor-core route: both sides carry the route's rounding noise (and the drift is part of the
    # fp16 operand here), hence twice the one-sided coherence bound
    # (bins 0 and 1 hold what the mean removal leaves of the DC line: a ratio of two cancellation
    # residues, compared through the spectra above only)
    assert float((got["Cxy"][2:] - ref["Cxy"][2:]).abs().max()) <= (1e-5 if route == "cuda" else 2e-4 * np.sqrt(16384 / K))


@pytest.mark.parametrize("shift", [(1, 1), (2, 2), (3, 3), (0, 3), (2, 1)])
@pytest.mark.parametrize("refl", [0, L - 1])
def test_records_not_aligned_to_16_bytes(shift, refl, cuda_device):
    """The loaders read 16-byte vectors: every misalignment of the first analysed sample (from the
    buffer address and from the odd reflect padding of nfft - 1 samples), equal or different for
    x and y, including the materialised reflect zone."""
    hop = 2048
    n = 2 ** 20
    x, y = _signals(n, 2, 18)
    K = (n - L) // hop + 1
    a, b = _both_routes(cuda_device, x, y, hop, (n, 0, refl, 0, n), 0, K, (0, 65536), shift=shift)
    _assert_close(a, b, K)


def test_record_that_ends_with_its_allocation(cuda_device):
    """the TMA copies must not read past the last sample: a 256 MiB record is a whole number of allocator
    pages, what follows it may be unmapped (this faulted once)."""
    import torch
    from pyfft_b200 import engine
    from pyfft_b200.windows import windows
    dev = cuda_device
    n, hop = 2 ** 26, 2048
    win = windows("Hanning", nwins=L, verbose=False)
    torch.cuda.empty_cache()
    g = torch.Generator(device=dev).manual_seed(1)
    xd = torch.randn(n, device=dev, generator=g) + 7.0
    yd = torch.randn(1, n, device=dev, generator=g) + 2.5
    K = (n - L) // hop + 1
    rmap = engine.record_map(n, 0, 0, 0, n)
    wp = engine.WelchPlan(L, hop, 1, "fp32", win, dev)
    wp.prepare_tc(xd, yd, rmap, 0, K * hop)
    for _ in range(3):
        wp.accumulate_tc(xd, yd, rmap, 0, K)
    torch.cuda.synchronize()
    assert not wp.tc_failed() and bool(torch.isfinite(wp.acc).all())
