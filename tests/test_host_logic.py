"""CPU: host-side mirror of the reference's bookkeeping (segmentation integers, window
factory, result containers, error behaviour) against the golden fixtures and the oracle."""
import json

import numpy as np
import pytest

from conftest import WELCH_CASES, load_golden
from oracle import restatement as R
from pyfft_b200 import fft_analysis as FA
from pyfft_b200.windows import windows


def test_windows_bit_exact_vs_reference_golden():
    g = load_golden("windows")
    for nm in json.loads(str(g["names_json"])):
        for n in (64, 129):
            assert np.array_equal(windows(nm, nwins=n, verbose=False), g["%s_%d" % (nm, n)]), nm
            assert np.array_equal(windows(nm, nwins=n, periodic=False, verbose=False),
                                  g["%s_%d_aper" % (nm, n)]), nm
        assert windows(nm, verbose=False) == float(g["%s_rov" % nm])
    assert np.array_equal(windows("kaiser", nwins=64, beta=8.0, verbose=False), g["kaiser8_64"])


def test_windows_return_conventions(capsys):
    w, (title, fom) = windows("Hanning", nwins=16, msgout=True, verbose=False)
    assert title == "Hanning" and fom.startswith("(SLDR") and w.shape == (16,)
    rov, (title, _) = windows("SFT3F", msgout=True, verbose=False)
    assert rov == 0.667 and title == "Fast-decaying Flattop"
    windows("hamming", nwins=8)
    assert "Using a periodic Hamming window function" in capsys.readouterr().out
    assert windows("no such window", verbose=False) == 0.0           # unmatched -> rectangular
    with pytest.raises(NameError):
        windows("nuttall", nwins=8, verbose=False)                  # family without suffix
    with pytest.raises(KeyError):
        windows("kaiser", nwins=8, verbose=False)


@pytest.mark.parametrize("case", WELCH_CASES)
def test_plan_matches_reference_integers(case):
    g = load_golden(case)
    kw = json.loads(str(g["kwargs_json"]))
    n = len(g["in_t"])
    extra = {k: kw[k] for k in ("tper", "minFreq") if k in kw}
    p = FA._Plan(g["in_t"], g["in_x"].shape[0], n, kw.get("tbounds"), kw.get("Navr"),
                 kw.get("windowoverlap"), kw.get("windowfunction"), extra)
    assert (p.nwins, p.noverlap, p.Navr) == (int(g["nwins"]), int(g["noverlap"]), int(g["Navr"]))
    assert [p.i0, p.i1] == [int(v) for v in g["ibnds"]]
    assert p.Fs == float(g["Fs"])
    S1, S2, NENBW, ENBW = FA.fftanal._getNorms(windows(p.windowfunction, nwins=p.nwins, verbose=False),
                                               p.Nnyquist, p.Fs)
    assert (S1, S2, NENBW, ENBW) == (float(g["S1"]), float(g["S2"]), float(g["NENBW"]), float(g["ENBW"]))


def test_plan_random_sweep_against_oracle():
    rng = np.random.default_rng(0)
    for _ in range(300):
        n = int(rng.integers(200, 50000))
        t = np.arange(n) / 1e6 + float(rng.uniform(0, 1e-3))
        navr = int(rng.integers(1, 40))
        ov = float(rng.choice([0.0, 0.25, 0.5, 0.661, 0.75]))
        part = rng.random() < 0.5
        tb = [t[int(0.1 * n)], t[int(0.8 * n)]] if part else None
        use_tper = rng.random() < 0.3
        kw = {"tper": float(rng.integers(16, n // 4)) / 1e6 * (1 + 1e-9)} if use_tper else {}
        o = R.welch_plan(t, n, tb, None if use_tper else navr, ov, "Hanning", kw.get("tper"))
        p = FA._Plan(t, n, n, tb, None if use_tper else navr, ov, "Hanning", dict(kw))
        assert (p.nwins, p.noverlap, p.Navr, p.i0, p.i1, p.reflect) == \
               (o["nwins"], o["noverlap"], o["Navr"], o["i0"], o["i1"], o["reflect"])


def test_segmentation_statics_and_cola():
    A = FA.fftanal
    assert A._getNwins(2 ** 26, 32767, 0.5) == 4096 and A._getNoverlap(4096, 0.5) == 2048
    assert A._getNavr(2 ** 26, 4096, 2048) == 32767
    assert A._getNwins(100, 1, 0.5) == 100 and A._getNavr(100, 100, 50) == 1
    assert A._getNnyquist(4096) == 2048 and A._getNnyquist(4097) == 2049
    assert A._checkCOLA(2 ** 20, 1024, 512) and not A._checkCOLA(2 ** 20 + 1, 1024, 512)
    assert A._getMINoverlap(4096 * 3, 4096, 2) == 2048
    assert A._getMAXoverlap(4096 * 3, 4096, 2) == 4095
    assert A._detrend_func(1) is FA.detrend_mean and A._detrend_func(-1) is FA.detrend_linear
    assert A._detrend_func(0) is FA.detrend_none and A._detrend_func(None) is FA.detrend_none


def test_detrend_helpers_match_oracle():
    rng = np.random.default_rng(5)
    x = rng.standard_normal((500, 3)) + np.arange(500)[:, None] * 0.01 + 3
    assert np.allclose(FA.detrend_mean(x, axis=0), R.detrend(x, 1), atol=1e-12)
    assert np.allclose(FA.detrend_linear(x, axis=0), R.detrend(x, -1), atol=1e-12)
    assert FA.detrend_none(x) is x


def test_cxy_cxy2_helper_matches_oracle():
    rng = np.random.default_rng(6)
    Pxx = rng.random(32) + 0j
    Pyy = rng.random((32, 2)) + 0j
    Pxy = rng.standard_normal((32, 2)) + 1j * rng.standard_normal((32, 2))
    a = FA.Cxy_Cxy2(Pxx, Pyy, Pxy)
    b = R.cxy_cxy2(Pxx, Pyy, Pxy)
    assert np.allclose(a[0], b[0]) and np.allclose(a[1], b[1])


def test_fftanal_init_mirrors_reference_fields():
    g = load_golden("stft_c3")
    o = FA.fftanal()                                  # prints the reference's hint and returns
    o.init(g["in_t"], g["in_y"][0], tper=float(g["tper"]), windowoverlap=0.75,
           windowfunction="Hanning", verbose=False)
    assert (o.nwins, o.noverlap, o.Navr) == (int(g["nwins"]), int(g["noverlap"]), int(g["Navr"]))
    assert (o.S1, o.S2, o.ENBW) == (float(g["S1"]), float(g["S2"]), float(g["ENBW"]))
    assert o.ibounds == [int(v) for v in g["ibounds"]] and o.onesided and o.nosigy


def test_reference_error_behaviour_is_mirrored():
    t = np.arange(4096) / 1e6
    x = np.zeros(4096)
    with pytest.raises(NotImplementedError):
        FA.fft_pwelch(t, x, x, useMLAB=True, plotit=False)
    with pytest.raises(NotImplementedError):
        FA.fft_pwelch(t, x, x, plotit=True)
    with pytest.raises(NameError):                    # Navr together with tper (fft_analysis.py:102, 208)
        FA._Plan(t, 4096, 4096, None, 4, 0.5, "Hanning", {"tper": 1e-4})
    with pytest.raises(NotImplementedError):          # nT-model branch
        FA._Plan(t, 100, 4096, None, 4, 0.5, "Hanning", {})


def test_fftinfosc_container():
    info = FA.fftinfosc()
    assert info.Pxx.dtype == np.complex128 and info.freq.size == 0
    with pytest.raises(AttributeError):
        info.Pxx_seg                                   # nothing to load lazily yet
    info.Pxx = np.ones(3)
    d = info.dict_from_class()
    assert "Pxx" in d and not any(k.startswith("_") for k in d)
    s = FA.Struct(info)
    assert np.array_equal(s.Pxx, np.ones(3))


def test_peer_reducer_needs_an_nccl_group():
    """The NVLink peer reduction is only taken inside an initialised NCCL group of >= 2 ranks;
    everywhere else create() answers False and the caller keeps its library collective."""
    import torch
    from pyfft_b200 import engine
    assert engine.PeerReducer.create(1024, torch.device("cpu"), None) is False


def test_bind_host_to_gpu_is_harmless_without_a_gpu():
    """engine.bind_host_to_gpu: no NVML / no device -> None and the affinity of the process is untouched."""
    import os
    from pyfft_b200 import engine
    before = os.sched_getaffinity(0)
    got = engine.bind_host_to_gpu("cuda:0")
    assert got is None or set(got) <= before
    if got is None:
        assert os.sched_getaffinity(0) == before


def test_fftinfo_pickles_without_its_loaders():
    """The lazy loaders are closures: pickling computes the derived statistics and drops them."""
    import pickle
    import numpy as np
    from pyfft_b200.fft_analysis import fftinfosc
    f = fftinfosc()
    f._defer_derived(lambda: {"varCxy": np.ones(3), "Lxx": np.zeros(2)})
    f._segment_loader = lambda: {"Pxx_seg": np.zeros(1)}
    g = pickle.loads(pickle.dumps(f))
    assert np.array_equal(g.varCxy, np.ones(3)) and g._segment_loader is None
    try:
        g.Pxx_seg
        raise AssertionError("expected AttributeError")
    except AttributeError:
        pass
    assert np.array_equal(f.Pxx_seg, np.zeros(1))          # the original keeps its loader


def test_cog_weights_reproduce_the_reference_band_selection():
    """doppler._weights (host side of pyfft_spectral_centroid) against the oracle's literal cog, including the
    reference's band selection that pairs the first m shifted bins with the m selected frequencies."""
    import numpy as np
    from oracle.restatement import cog
    from pyfft_b200.doppler import _weights
    rng = np.random.default_rng(3)
    for n in (64, 333, 1000):
        x = rng.standard_normal(n) + 1j * rng.standard_normal(n) + np.exp(2j * np.pi * 0.11 * np.arange(n))
        P = np.abs(np.fft.fft(x)) ** 2
        for fmin, fmax in ((None, None), (5.0e4, 4.0e5), (0.0, 1.0e5), (3.0e6, None)):
            wn, wd = _weights(n, 2.0e6, fmin, fmax)
            got = (P @ wn) / (P @ wd) if wd.sum() > 0 else 0.0
            ref = cog(x, 2.0e6, fmin=fmin, fmax=fmax)
            assert abs(got - ref) <= 1e-9 * 2.0e6, (n, fmin, fmax)


def test_segment_statistics_bookkeeping_against_numpy():
    """segstats.stats_from_moments (host side of pyfft_segment_moments): moments built with numpy from explicit
    per-segment spectra -> np.var over the segments and the reference-style mean_angle, one- and two-sided."""
    import numpy as np
    from pyfft_b200.segstats import stats_from_moments
    from pyfft_b200.fft_analysis import varphi
    rng = np.random.default_rng(4)
    nfft, hop, K, nch, Fs = 64, 32, 37, 2, 1.0e3
    n = (K - 1) * hop + nfft
    x = rng.standard_normal(n) + np.sin(0.7 * np.arange(n))
    y = rng.standard_normal((nch, n)) + 0.5 * x
    w = np.hanning(nfft + 1)[:-1]
    X = np.stack([np.fft.fft(w * x[k * hop:k * hop + nfft]) for k in range(K)])
    Y = np.stack([[np.fft.fft(w * y[c, k * hop:k * hop + nfft]) for k in range(K)] for c in range(nch)])
    nb = nfft // 2 + 1
    m = np.zeros((nch, 12, nb))
    for c in range(nch):
        pxx, pyy, pxy = np.abs(X[:, :nb]) ** 2, np.abs(Y[c, :, :nb]) ** 2, Y[c, :, :nb] * np.conj(X[:, :nb])
        ph = np.angle(pxy)
        for q, v in enumerate((pxx, pyy, pxy.real, pxy.imag, pxx ** 2, pyy ** 2, pxy.real ** 2, pxy.imag ** 2,
                               np.cos(ph), np.sin(ph), np.cos(ph) ** 2, np.sin(ph) ** 2)):
            m[c, q] = v.sum(axis=0)
    scale = 0.37
    for onesided in (True, False):
        out = stats_from_moments(m, K, scale, onesided, nfft, Fs, nfft // 2)
        if onesided:
            wgt = np.ones(nfft // 2)
            wgt[1:-1] = 2.0
            Pxx_seg = (np.abs(X[:, :nfft // 2]) ** 2) * wgt * scale
            Pxy_seg = np.stack([(Y[c] * np.conj(X))[:, :nfft // 2] * wgt * scale for c in range(nch)])
        else:
            Pxx_seg = np.fft.fftshift(np.abs(X) ** 2, axes=-1) * scale
            Pxy_seg = np.stack([np.fft.fftshift(Y[c] * np.conj(X), axes=-1) * scale for c in range(nch)])
        assert np.allclose(out.varPxx, np.var(Pxx_seg, axis=0), rtol=1e-10, atol=1e-12 * np.max(Pxx_seg) ** 2)
        ref = (np.var(Pxy_seg.real, axis=1) + 1j * np.var(Pxy_seg.imag, axis=1)).T
        assert np.allclose(out.varPxy, ref, rtol=1e-9, atol=1e-12 * np.max(np.abs(Pxy_seg)) ** 2)
        assert np.allclose(out.Pxy, Pxy_seg.mean(axis=1).T, rtol=1e-10, atol=1e-12 * np.max(np.abs(Pxy_seg)))
        ph = np.angle(Pxy_seg)
        ca, sa = np.cos(ph), np.sin(ph)
        mp, vp = varphi(Pxy_real=ca.mean(axis=1), Pxy_imag=sa.mean(axis=1), varPxy_real=ca.var(axis=1),
                        varPxy_imag=sa.var(axis=1), angle_range=0.5 * np.pi)
        assert np.allclose(out.phi_mean, mp.T, rtol=0, atol=1e-9)
