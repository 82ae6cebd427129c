"""fused per-segment statistics (pyfft_segment_moments / welch_segment_statistics) against
statistics of the per-segment spectra of the UNMODIFIED reference (tests/golden/segstats.npz, written by
`python -m oracle.gen_golden segstats`) and against this package's own materialised *_seg fields at a larger size."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden", "segstats.npz")


def _run(g, tag, precision):
    from pyfft_b200 import welch_segment_statistics
    n, navr, nch, onesided, detrend = (int(v) for v in g[tag + "_cfg"])
    t = np.arange(n) / float(g[tag + "_fs"])
    return welch_segment_statistics(t, g[tag + "_x"], g[tag + "_y"], None, Navr=navr, windowoverlap=0.5,
                                    windowfunction="Hanning", detrend_style=detrend, onesided=bool(onesided),
                                    precision=precision)


@pytest.mark.parametrize("tag", ["a", "b", "c", "d"])
def test_segment_statistics_fp64_vs_reference(tag):
    g = np.load(GOLD)
    out = _run(g, tag, "fp64")
    assert out.Navr == int(g[tag + "_navr"])
    for nm in ("Pxx", "Pyy", "Pxy", "varPxx", "varPyy", "varPxy"):
        ref = g[tag + "_" + nm]
        got = np.asarray(getattr(out, nm)).reshape(ref.shape)
        assert np.max(np.abs(got - ref)) <= 1e-11 * np.max(np.abs(ref)), nm
    # circular mean of the segment phases and its propagated variance (mean_angle)
    ref, got = g[tag + "_phi_mean"], np.asarray(out.phi_mean).reshape(g[tag + "_phi_mean"].shape)
    assert np.max(np.abs(got - ref)) < 1e-9
    ref, got = g[tag + "_varphi_mean"], np.asarray(out.varphi_mean).reshape(g[tag + "_varphi_mean"].shape)
    ok = np.isfinite(ref)
    assert np.array_equal(ok, np.isfinite(got))
    assert np.max(np.abs(got[ok] - ref[ok]) / np.maximum(np.abs(ref[ok]), 1e-12)) < 1e-7


@pytest.mark.parametrize("tag", ["a", "b", "d"])
def test_segment_statistics_fp32_vs_reference(tag):
    g = np.load(GOLD)
    out = _run(g, tag, "fp32")
    for nm, tol in (("Pxx", 1e-5), ("Pyy", 1e-5), ("Pxy", 1e-5), ("varPxx", 1e-4), ("varPyy", 1e-4), ("varPxy", 1e-4)):
        ref = g[tag + "_" + nm]
        got = np.asarray(getattr(out, nm)).reshape(ref.shape)
        assert np.max(np.abs(got - ref)) <= tol * np.max(np.abs(ref)), nm
    # the phase of a segment is well defined where the cross power is: strong bins only in single precision
    ref = g[tag + "_phi_mean"]
    got = np.asarray(out.phi_mean).reshape(ref.shape)
    pxy = np.abs(g[tag + "_Pxy"]).reshape(ref.shape)
    strong = pxy > 1e-2 * pxy.max()
    assert np.max(np.abs(got - ref)[strong]) < 1e-4


def test_segment_statistics_match_materialised_segments():
    """nfft 4096, K = 1023, two y channels: the fused sums against the lazy per-segment fields of fft_pwelch."""
    from pyfft_b200 import fft_pwelch, mean_angle, welch_segment_statistics
    rng = np.random.default_rng(12)
    n, navr = 2 ** 21, 1023
    t = np.arange(n) / 1.0e6
    x = np.sin(2 * np.pi * 5.0e4 * t) + 0.4 * rng.standard_normal(n) + 0.3
    y = np.stack([0.5 * np.sin(2 * np.pi * 5.0e4 * t + 0.7 * (c + 1)) + 0.3 * rng.standard_normal(n) for c in range(2)], 1)
    kw = dict(Navr=navr, windowoverlap=0.5, windowfunction="Hanning", detrend_style=1, precision="fp64")
    out = welch_segment_statistics(t, x, y, None, **kw)
    info = fft_pwelch(t, x, y, None, plotit=False, **kw)[6]
    assert out.nwins == 4096 and info.Pxx_seg.shape == (navr, 2048)
    pk = np.max(np.abs(info.Pxx_seg))
    assert np.max(np.abs(out.varPxx - np.var(info.Pxx_seg, axis=0))) <= 1e-10 * pk ** 2
    assert np.max(np.abs(out.varPyy - np.var(info.Pyy_seg, axis=1).T)) <= 1e-10 * pk ** 2
    v = (np.var(info.Pxy_seg.real, axis=1) + 1j * np.var(info.Pxy_seg.imag, axis=1)).T
    assert np.max(np.abs(out.varPxy - v)) <= 1e-10 * pk ** 2
    mp, vp = mean_angle(info.phixy_seg, dim=1)
    assert np.max(np.abs(out.phi_mean - np.asarray(mp).T)) < 1e-8


def test_segment_statistics_rejects_other_lengths():
    from pyfft_b200 import welch_segment_statistics
    t = np.arange(30000) / 1e5
    x = np.random.default_rng(1).standard_normal(30000)
    with pytest.raises(NotImplementedError):
        welch_segment_statistics(t, x, x.copy(), None, Navr=8)
