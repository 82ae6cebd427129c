"""GPU parity of K2 (STFT with materialised spectra): spectrogram.stft / fftanal class path /
specgram / lazy per-segment fields of fft_pwelch, against the reference's golden outputs and
the oracle."""
import contextlib
import io

import numpy as np
import pytest

from conftest import load_golden, relerr

pytestmark = pytest.mark.gpu

TOL = {"fp64": 1e-12, "fp32": 2e-6}       # relative to the largest |Xseg| (amplitude spectra)


@pytest.mark.parametrize("precision", ["fp64", "fp32"])
def test_spectrogram_stft_matches_reference_golden(precision, cuda_device):
    from pyfft_b200 import stft
    g = load_golden("stft_c3")
    for c in range(2):
        y = g["in_y"][c] if precision == "fp64" else g["in_y"][c].astype(np.float32)
        o = stft(g["in_t"], y, tper=float(g["tper"]), windowoverlap=0.75, windowfunction="Hanning",
                 verbose=False, precision=precision)
        assert (o.nwins, o.noverlap, o.Navr) == (int(g["nwins"]), int(g["noverlap"]), int(g["Navr"]))
        assert o.Xseg.shape == g["Xseg_%d" % c].shape and o.Xseg.dtype == np.complex128
        tol = TOL[precision]
        assert relerr(o.Xseg, g["Xseg_%d" % c]) <= tol
        assert relerr(o.Xfft, g["Xfft_%d" % c]) <= tol
        assert relerr(o.Xpow, g["Xpow_%d" % c]) <= max(tol, 1e-11)
        assert relerr(o.Pxx, g["Pxx_%d" % c]) <= 10 * tol
        assert relerr(o.tseg, g["tseg"]) < 1e-13 and relerr(o.freq, g["freq"]) < 1e-14
        assert relerr(o.Pxx_seg, (g["Xseg_%d" % c] * np.conj(g["Xseg_%d" % c]))) <= 10 * tol


def test_class_path_two_signals_up_to_pstft(cuda_device):
    from pyfft_b200 import fftanal
    g = load_golden("stft_xy")
    o = fftanal(g["in_t"], g["in_x"], g["in_y"], tper=float(g["tper"]), windowoverlap=0.5, verbose=False)
    o.Xstft(); o.Ystft(); o.Pstft()
    assert o.Navr == int(g["Navr"]) and o.nwins == int(g["nwins"])
    for nm in ("Xseg", "Yseg", "Pxy_seg", "Pxx_seg", "Pyy_seg"):
        assert relerr(getattr(o, nm), g[nm]) <= 1e-12, nm
    w = np.abs(g["Pxy_seg"])
    d = np.abs(np.exp(1j * o.phixy_seg) - np.exp(1j * g["phixy_seg"]))
    assert np.max(w * d) / np.max(w) < 1e-10
    # X conj(Y): the opposite sign convention to fft_pwelch (fft_analysis.py:1960 vs :393)
    assert relerr(o.Pxy_seg, o.Xseg * np.conj(o.Yseg)) < 1e-12
    with pytest.raises(IndexError):                     # the reference's averagewins dies here too
        o.averagewins()


@pytest.mark.parametrize("precision", ["fp64", "fp32"])
def test_specgram_matches_reference_golden(precision, cuda_device):
    from pyfft_b200 import specgram
    g = load_golden("specgram")
    s = g["in_s"] if precision == "fp64" else g["in_s"].astype(np.float32)
    with contextlib.redirect_stdout(io.StringIO()):
        tm, f, S = specgram(g["in_t"], s, wl=512, precision=precision)
        tm2, f2, S2 = specgram(g["in_t"][:10000], g["in_s2"], wl=256, hanning=False, precision="fp64")
    assert S.shape == g["S"].shape and S.dtype == np.float64
    assert relerr(S, g["S"]) <= (1e-12 if precision == "fp64" else 1e-5)
    assert relerr(tm, g["time"]) < 1e-14 and relerr(f, g["freq"]) < 1e-14
    assert relerr(S2, g["S2"]) <= 1e-12 and relerr(tm2, g["time2"]) < 1e-14
    with pytest.raises(TypeError):
        specgram(g["in_t"], s, wl=512, windowAverage=4)


@pytest.mark.parametrize("log2l", [4, 6, 9, 11, 13])
def test_fft_win_options_vs_oracle(log2l, cuda_device):
    """odd segment counts (unpaired last segment), two-sided output, per-segment detrend."""
    from oracle import restatement as R
    from pyfft_b200 import fftanal
    L = 1 << log2l
    rng = np.random.default_rng(log2l)
    K = 7
    n = L // 4 * (K + 3) + 3
    t = np.arange(n) / 2e5
    x = rng.standard_normal(n) + 1.5 + 0.3 * np.sin(0.05 * np.arange(n))
    for onesided, detrendwin in ((True, False), (False, False), (True, True)):
        o = fftanal(t, x, None, tper=L / 2e5 * (1 + 1e-9), windowoverlap=0.75, windowfunction="Hamming",
                    verbose=False, onesided=onesided)
        assert o.nwins == L and o.Navr == K
        tt, fr, X, ps = o.fft_win(x, t, detrendwin=detrendwin)
        rt, rf, RX, rp = R.fft_win(x, t, o.win, o.Navr, o.noverlap, onesided=onesided,
                                   detrend_style=1, detrendwin=detrendwin)
        assert relerr(X, RX) <= 1e-12, (onesided, detrendwin)
        assert relerr(ps, rp) <= 1e-11 and relerr(tt, rt) < 1e-13 and relerr(fr, rf) < 1e-14


def test_lazy_segment_fields_of_fft_pwelch(cuda_device):
    from oracle import restatement as R
    from pyfft_b200 import fft_pwelch
    g = load_golden("welch_c2_cross")
    f, Pxy, Pxx, Pyy, Cxy, phi, info = fft_pwelch(g["in_t"], g["in_x"], g["in_y"], Navr=15,
                                                  windowoverlap=0.5, windowfunction="Hanning",
                                                  plotit=False, detrend_style=1)
    K, L = 15, 4096
    assert info.Xfft_seg.shape == (K, L) and info.Yfft_seg.shape == (1, K, L)
    assert info.Pxx_seg.shape == (K, L // 2) and info.Pxy_seg.shape == (1, K, L // 2)
    assert relerr(info.Pxx_seg.mean(0), g["Pxx"]) < 1e-12
    assert relerr(info.Pxy_seg.mean(1)[0], g["Pxy"]) < 1e-12
    assert relerr(info.Pyy_seg.mean(1)[0], g["Pyy"]) < 1e-12
    # raw transforms against numpy on the reference's own segmentation of the same record
    xp = R.reflect_pad(g["in_x"], L)[:len(g["in_x"])]
    xp = xp - xp.mean()
    ref = np.fft.fft(xp[3 * 2048:3 * 2048 + L] * info.win)
    assert relerr(info.Xfft_seg[3], ref) < 1e-12
    assert info.phixy_seg.shape == info.Pxy_seg.shape and not info.varphi_seg.any()


def test_config3_shape_channels_batch(cuda_device):
    """config 3 cut: many channels in one launch, 75 % overlap, fp32; channel c of the batched
    call equals the single-channel call, and Parseval holds per segment."""
    import torch
    from pyfft_b200 import engine
    from pyfft_b200.windows import windows
    dev = cuda_device
    C, n, L, H = 8, 2 ** 18, 2048, 512
    g = torch.Generator(device=dev).manual_seed(3)
    x = torch.randn((C, n), generator=g, device=dev, dtype=torch.float32)
    win = windows("Hanning", nwins=L, verbose=False)
    K = (n - (L - H)) // H
    rmap = engine.record_map(n, 0, 0, 0, n)
    out, pseg = engine.stft(x, rmap, L, H, 0, K, win, "fp32", bin_first=0, nbins=L // 2, want_pseg=True)
    assert out.shape == (C, K, L // 2) and out.dtype == torch.complex64
    one, _ = engine.stft(x[5:6].contiguous(), rmap, L, H, 0, K, win, "fp32", bin_first=0, nbins=L // 2)
    assert torch.equal(one[0], out[5])
    full, _ = engine.stft(x[:1].contiguous(), rmap, L, H, 0, K, win, "fp32")
    seg = x[0].unfold(0, L, H)[:K].double() * torch.from_numpy(win).to(dev)
    lhs = (full[0].to(torch.complex128).abs() ** 2).sum(1)
    rhs = L * (seg * seg).sum(1)
    assert float(((lhs - rhs).abs() / rhs).max()) < 1e-5
    ref = torch.fft.fft(seg[100])                       # independent transform of one segment
    assert float((full[0, 100].to(torch.complex128) - ref).abs().max() / ref.abs().max()) < 2e-6
