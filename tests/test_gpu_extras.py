"""GPU parity of the SURVEY 8(f) rows and of the API that round 1 shipped without a test:
getNpeaks, fftanal.fftpwelch + crosscorr, mean_angle, fftanal.intspectra, laplace_1d against outputs
of the UNMODIFIED reference (tests/golden/extras.npz, oracle/gen_golden.py extras); integratespectra,
crosscorr_stft, convert2amplitudes, pwelch with two signals and ccf_sh against numpy restatements of
the reference lines (the reference itself raises there, or needs the un-vendored pybaseutils: parity
unpinned); and the ctypes stub of INTEGRATION.md executed verbatim.
"""
import os
import re

import numpy as np
import pytest

from conftest import load_golden, relerr

pytestmark = pytest.mark.gpu

KW = dict(Navr=63, windowoverlap=0.5, windowfunction="Hanning", detrend_style=1)


def _inputs():
    g = load_golden("extras")
    n, fs = int(g["t_n"]), float(g["t_fs"])
    return g, np.arange(n) / fs, g["x"], g["y"]


def test_getNpeaks_matches_reference(cuda_device):
    from pyfft_b200 import getNpeaks
    g, t, x, y = _inputs()
    for npk, key, extra in ((1, "peaks_1", {}), (4, "peaks_4", {}), (3, "peaks_minsep", dict(minsep=2000.0))):
        out = getNpeaks(npk, t, x, y, plotit=False, **KW, **extra)
        got = np.asarray([[float(v) for v in row] for row in out])
        assert got.shape == g[key].shape
        assert np.array_equal(got[:, 1], g[key][:, 1]), key                   # the same bins
        assert np.max(np.abs(got - g[key])) <= 1e-10, key
    # batched over the channels of a 2-D sigy: column c equals the single-channel call
    y2 = np.stack([y, 0.5 * np.roll(y, 3) + 0.1 * x], 1)
    both = getNpeaks(3, t, x, y2, plotit=False, **KW)
    for c in range(2):
        one = getNpeaks(3, t, x, y2[:, c].copy(), plotit=False, **KW)
        assert np.allclose(np.asarray(both[c], dtype=float), np.asarray(one, dtype=float), rtol=0, atol=1e-12)
    # an exclusion window wider than the spectrum on both sides removes NOTHING in the reference
    # (neither branch of :1010-1017 fires): the same peak comes back -- mirrored literally
    rep = np.asarray(getNpeaks(5, t, x, y, plotit=False, minsep=5.0e4, **KW), dtype=float)
    assert np.array_equal(rep, np.repeat(rep[:1], 5, axis=0)) and rep[0, 1] == g["peaks_1"][0, 1]


def test_fftanal_fftpwelch_and_crosscorr_match_reference(cuda_device):
    from pyfft_b200 import fftanal
    g, t, x, y = _inputs()
    o = fftanal(t, x, y, Navr=63, windowoverlap=0.5, verbose=False)
    o.fftpwelch()
    for nm in ("Pxx", "Pyy", "Pxy", "freq"):
        assert relerr(getattr(o, nm), g["cls_" + nm]) <= 1e-12, nm
    o.crosscorr()
    for nm in ("Rxx", "Ryy", "Rxy", "corrcoef", "Ex", "Ey", "lags"):
        a, b = np.asarray(getattr(o, nm)), g["cls_" + nm]
        assert a.shape == b.shape, (nm, a.shape, b.shape)
        assert relerr(a, b) <= 1e-11, nm


def test_mean_angle_matches_reference(cuda_device):
    import torch
    from pyfft_b200 import mean_angle
    g = load_golden("extras")
    ph, vph = g["ma_phi"], g["ma_vphi"]
    for tag, args in (("a", dict()), ("b", dict(vphi=vph)), ("c", dict(vphi=vph, dim=1, angle_range=np.pi))):
        m, v = mean_angle(ph, **args)
        assert m.shape == g["ma_mean_" + tag].shape
        assert np.max(np.abs(m - g["ma_mean_" + tag])) <= 1e-13 and relerr(v, g["ma_var_" + tag]) <= 1e-12, tag
    # device tensors in, NaN segments skipped like nanmean / nanvar
    p2 = ph.copy()
    p2[7, 1] = np.nan
    m, v = mean_angle(torch.from_numpy(p2).to(cuda_device))
    ca, sa = np.cos(p2), np.sin(p2)
    assert np.max(np.abs(m - np.arctan(np.nanmean(sa, 0) / np.nanmean(ca, 0)))) <= 1e-13


def test_intspectra_matches_reference():
    from pyfft_b200 import fftanal
    g = load_golden("extras")
    fr, sp = g["is_freq"], g["is_spec"]
    assert fftanal.intspectra(fr, sp)[0] == pytest.approx(float(g["is_default"]), rel=1e-14)
    assert fftanal.intspectra(fr, sp, ENBW=600.0)[0] == pytest.approx(float(g["is_enbw"]), rel=1e-14)
    assert fftanal.intspectra(fr, sp, ifreq=40)[0] == pytest.approx(float(g["is_whole"]), rel=1e-14)
    with pytest.raises(UnboundLocalError):
        fftanal.intspectra(fr, sp, ifreq=40, ispan=10)


@pytest.mark.parametrize("n", [300, 512])
def test_laplace_1d_matches_reference(n, cuda_device):
    from pyfft_b200 import laplace_1d
    g = load_golden("extras")
    u, sig, ref = g["lp_u_%d" % n], g["lp_sigma_%d" % n], g["lp_out_%d" % n]
    out = laplace_1d(u, sig)
    assert out.shape == ref.shape and out.dtype == np.complex128
    for r in range(len(sig)):                                                 # rows differ by orders of magnitude
        assert relerr(out[r], ref[r]) <= 1e-11, (r, relerr(out[r], ref[r]))
    out32 = laplace_1d(u.astype(np.float32), sig)
    ref32 = np.array([np.fft.rfft((np.exp(s * np.arange(n)) / np.exp(s * np.arange(n)).sum() * u.astype(np.float32)[::-1])[::-1]) for s in sig])
    assert relerr(out32, ref32) <= 2e-6


def test_integratespectra_vs_numpy_restatement(cuda_device):
    """fft_analysis.py:890-903, :931-935 restated with numpy.trapz (pybaseutils.trapz_var is not vendored)."""
    from pyfft_b200 import fft_pwelch, integratespectra
    g, t, x, y = _inputs()
    y2 = np.stack([y, 0.5 * np.roll(y, 3) + 0.1 * x], 1)
    f, Pxy, Pxx, Pyy, Cxy, phi, info = fft_pwelch(t, x, y2, plotit=False, **KW)
    bands = np.array([[2.5e3, 3.5e3], [10e3, 12e3], [0.0, 5e4], [20e3, 20.01e3]])
    out = integratespectra(f, Pxy, Pxx, Pyy, bands, varPxy=info.varPxy, varPxx=info.varPxx, varPyy=info.varPyy)
    trapz = np.trapezoid
    for b, (lo, hi) in enumerate(bands):
        inds = np.where((f >= lo) * (f <= hi))[0]
        assert np.array_equal(out[5].ifrange[b], inds)
        assert relerr(out[0][b], trapz(Pxy[inds], f[inds], axis=0)) <= 1e-13
        assert relerr(out[1][b], np.ones(2) * trapz(Pxx[inds], f[inds])) <= 1e-13
        assert relerr(out[2][b], trapz(Pyy[inds], f[inds], axis=0)) <= 1e-13
        if len(inds) > 1:
            fw = trapz(f[inds][:, None] * np.abs(Pxy[inds]), axis=0) / trapz(np.abs(Pxy[inds]), axis=0)
            assert relerr(out[5].fweighted[b], fw) <= 1e-13
            assert relerr(out[4][b], np.arctan2(out[0][b].imag, out[0][b].real)) <= 1e-14
    one = integratespectra(f, Pxy[:, 0], Pxx, Pyy[:, 0], bands[0])            # the reference's call shape
    assert relerr(one[0], out[0][0][:1]) <= 1e-15 and one[5].Cxy_i.shape == (1,)


def test_class_path_with_two_signals_and_the_methods_the_reference_cannot_reach(cuda_device):
    """fftanal.pwelch() with sigy raises in the reference (Ystft for 2-D sigy, averagewins for 1-D,
    SURVEY 3.2) -- and so do crosscorr_stft and convert2amplitudes after fftpwelch.  Here the 1-D class
    path runs up to Pstft + averagewins' means like the reference's lines say; checked against numpy on the
    per-segment spectra."""
    from pyfft_b200 import fftanal
    g, t, x, y = _inputs()
    o = fftanal(t, x, y, tper=1024 / 1.0e5 * (1 + 1e-9), windowoverlap=0.5, verbose=False)
    o.Xstft()
    o.Ystft()
    o.Pstft()
    assert relerr(o.Pxy_seg, o.Xseg * np.conj(o.Yseg)) <= 1e-13               # X conj(Y) (:1960)
    with pytest.raises(IndexError):                                            # Cxy_Cxy2 on 1-D spectra (:1669)
        o.averagewins()
    assert relerr(o.Pxx, np.mean(o.Pxx_seg, axis=0)) <= 1e-13 and relerr(o.Pxy, np.mean(o.Pxy_seg, axis=0)) <= 1e-13
    # crosscorr_stft on (Navr, F) one-sided segment spectra (:1880-1920)
    o.nch = 1
    with pytest.raises(ValueError):                # the reference's corrcoef_seg line cannot broadcast (:1916)
        o.crosscorr_stft()
    q = o.Pxy_seg.copy()
    q[..., 1:-1] *= 0.5
    r = np.fft.fftshift(np.sqrt(o.nwins) * np.fft.irfft(q, n=o.nwins, axis=-1), axes=-1)
    assert o.Rxy_seg.shape == r.shape and relerr(o.Rxy_seg, r) <= 1e-11
    assert relerr(o.Ex_seg, r[..., 0] * 0 + np.fft.irfft(o.Pxx_seg * np.r_[1, 0.5 * np.ones(o.Pxx_seg.shape[-1] - 2), 1],
                                                         n=o.nwins, axis=-1)[..., 0] * np.sqrt(o.nwins)) <= 1e-11
    # convert2amplitudes (:2022-2048)
    o.varPxx, o.varPyy, o.varPxy = ((p / np.sqrt(o.Navr)) ** 2 for p in (o.Pxx, o.Pyy, o.Pxy))
    o.convert2amplitudes()
    amp = np.sqrt(np.abs(o.ENBW * o.Pxy))
    amp[1:-1] *= np.sqrt(2)
    assert relerr(o.Lxy, amp) <= 1e-14


def test_ccf_sh_vs_window_average_of_ccf(cuda_device):
    from pyfft_b200 import ccf, ccf_sh
    rng = np.random.default_rng(3)
    n, nav = 4096, 512
    a = rng.standard_normal(n)
    b = np.roll(a, 9) + 0.3 * rng.standard_normal(n)
    tau, co = ccf_sh(a, b, 1.0e4, nav)
    ref = np.mean([ccf(a[i:i + nav], b[i:i + nav], 1.0e4)[1] for i in range(0, n - nav + 1, nav)], axis=0)
    assert co.shape == (2 * nav - 1,) and relerr(co, ref) <= 1e-12
    assert tau[np.argmax(co)] == pytest.approx(ccf(a[:nav], b[:nav], 1.0e4)[0][np.argmax(ref)])


def test_integration_md_stub_runs_verbatim(cuda_device, tmp_path):
    """the ctypes stub printed in INTEGRATION.md, executed as written (library path made absolute)."""
    from oracle import restatement as R
    from pyfft_b200 import _lib
    from pyfft_b200.windows import windows
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    md = open(os.path.join(root, "INTEGRATION.md")).read()
    blocks = re.findall(r"```python\n(.*?)```", md, flags=re.S)
    stub = [b for b in blocks if "def welch_sums" in b][0]
    assert 'C.CDLL("libpyfft_b200.so")' in stub
    ns = {}
    exec(compile(stub.replace('C.CDLL("libpyfft_b200.so")', "C.CDLL(%r)" % _lib.LIB_PATH), "INTEGRATION.md", "exec"), ns)
    g, t, x, y = _inputs()
    n, L, K = len(x), 1024, 63
    ref = R.fft_pwelch(t, x, y, **KW)
    win = windows("Hanning", nwins=L, verbose=False)
    S1, S2 = win.sum(), (win ** 2).sum()
    fs = (n - 1) / (t[-1] - t[0])
    out = ns["welch_sums"](x, y[:, None], 0, n, L, L // 2, K, win, 1, True, S1, fs * S2 / S1 ** 2, True)
    assert relerr(out[0], ref["Pxx"]) <= 1e-12 and relerr(out[2][:, 0], ref["Pxy"]) <= 1e-12
    assert np.max(np.abs(out[3][:, 0] - ref["Cxy"])) <= 1e-10
