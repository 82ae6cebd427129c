"""CPU: pins oracle/restatement.py against OUTPUTS OF THE REFERENCE ITSELF
(tests/golden/*.npz, written by oracle/gen_golden.py from the unmodified
reference in the build container)."""
import json
import os

import numpy as np
import pytest

from conftest import UNIT_SCALE, WELCH_CASES, load_golden, phase_err, relerr
from oracle import restatement as R

FIELDS = ["freq", "Pxy", "Pxx", "Pyy", "Cxy", "phi_xy", "Cxy2", "Lxx", "Lyy", "Lxy", "Rxx", "Ryy",
          "Rxy", "iCxy", "corrcoef", "lags", "Ex", "Ey", "varCxy", "varCxy2", "varPxx", "varPyy",
          "varPxy", "varLxx", "varLyy", "varLxy", "S1", "S2", "ENBW", "NENBW", "win", "minFreq"]


@pytest.mark.parametrize("case", WELCH_CASES)
def test_fft_pwelch_restatement_matches_reference(case):
    g = load_golden(case)
    kw = json.loads(str(g["kwargs_json"]))
    tb = kw.pop("tbounds")
    with np.errstate(all="ignore"):
        o = R.fft_pwelch(g["in_t"], g["in_x"], g["in_y"], tb, **kw)
    for nm in ("nwins", "noverlap", "Navr", "nch"):
        assert int(g[nm]) == int(o[nm]), nm
    assert [int(v) for v in g["ibnds"]] == [o["i0"], o["i1"]]
    for nm in FIELDS:
        if nm == "phi_xy":       # the phase of a (numerically) zero cross-power is noise: weight by |Pxy|
            assert phase_err(o[nm], g[nm], g["Pxy"]) < 1e-11, (case, nm)
        else:
            assert relerr(o[nm], g[nm], UNIT_SCALE.get(nm, 0.0)) < 1e-11, (case, nm)
    ok = np.isfinite(g["varPhxy"])
    # sqrt(1 - |Cxy2|) is ill-conditioned at coherence 1 (auto-spectra): O(1) scale, loose bound
    assert relerr(np.asarray(o["varPhxy"])[ok], g["varPhxy"][ok], 1.0) < 1e-6


def test_port_matches_reference():
    g = load_golden("welch_c2_cross")
    f, Pxy, Pxx, Pyy, Cxy = R.fft_pwelch_port(g["in_t"], g["in_x"], g["in_y"], 15)
    assert relerr(Pxy, g["Pxy"].reshape(Pxy.shape)) < 1e-12
    assert relerr(Pxx, g["Pxx"]) < 1e-12
    assert relerr(Pyy, g["Pyy"].reshape(Pyy.shape)) < 1e-12
    assert relerr(Cxy, g["Cxy"].reshape(Cxy.shape)) < 1e-12
    assert relerr(f, g["freq"]) < 1e-15


def test_windows_restatement_bit_exact():
    g = load_golden("windows")
    for nm in json.loads(str(g["names_json"])):
        for n in (64, 129):
            assert np.array_equal(R.window_vector(nm, n), g["%s_%d" % (nm, n)]), nm
            assert np.array_equal(R.window_vector(nm, n, periodic=False), g["%s_%d_aper" % (nm, n)]), nm
        assert R.window_rov(nm) == float(g["%s_rov" % nm])
    assert np.array_equal(R.window_vector("kaiser", 64, beta=8.0), g["kaiser8_64"])


def test_stft_restatement():
    g = load_golden("stft_c3")
    win = R.window_vector("Hanning", int(g["nwins"]))
    i0, i1 = [int(v) for v in g["ibounds"]]
    for c in range(2):
        tt, fr, X, ps = R.fft_win(g["in_y"][c][i0:i1], g["in_t"][i0:i1], win, int(g["Navr"]),
                                  int(g["noverlap"]))
        assert relerr(X, g["Xseg_%d" % c]) < 1e-12
        assert relerr(ps, g["Xpow_%d" % c]) < 1e-12
        assert relerr(X.mean(0), g["Xfft_%d" % c]) < 1e-12
        assert relerr((X * np.conj(X)).mean(0), g["Pxx_%d" % c]) < 1e-12
    assert relerr(tt, g["tseg"]) < 1e-14 and relerr(fr, g["freq"]) < 1e-14


def test_specgram_restatement():
    g = load_golden("specgram")
    tm, f, S = R.specgram(g["in_t"], g["in_s"], wl=512)
    assert relerr(S, g["S"]) < 1e-13 and relerr(tm, g["time"]) < 1e-14 and relerr(f, g["freq"]) < 1e-14
    tm, f, S = R.specgram(g["in_t"][:10000], g["in_s2"], wl=256, hanning=False)
    assert relerr(S, g["S2"]) < 1e-13 and relerr(tm, g["time2"]) < 1e-14


def test_hilbert_restatement_and_known_answers():
    g = load_golden("hilbert")
    for n in (32, 33, 1024, 1000, 4096):
        z = R.hilbert(g["u_%d" % n])
        assert relerr(z, g["z_%d" % n]) < 1e-13
        assert relerr(z, g["z1d_%d" % n]) < 1e-13
    assert relerr(R.hilbert(g["u_batch"]), g["z_batch"]) < 1e-13
    # analytic known answer: H[cos] = sin  (hilbert.py:117-128 recipe)
    n = 64
    u = np.cos(2 * np.pi * 4 * np.arange(n) / n)
    assert np.allclose(R.hilbert(u).imag, np.sin(2 * np.pi * 4 * np.arange(n) / n), atol=1e-12)
    # even lengths agree with scipy.signal.hilbert
    from scipy.signal import hilbert as sh
    assert relerr(R.hilbert(g["u_1024"]), sh(g["u_1024"])) < 1e-13


def test_ccf_fft_identity_equals_direct_definition():
    # ccf.py cannot be imported on py3; the direct definition (ccf.py:66-77) is the pin
    rng = np.random.default_rng(7)
    fs, n = 1e5, 2048                                   # ccf_test recipe, ccf.py:139-148
    t = np.arange(n) / fs
    x1 = np.sin(2 * np.pi * 1e3 * t) + rng.standard_normal(n)
    x2 = np.sin(2 * np.pi * 1e3 * t + 50 * np.pi / 180) + rng.standard_normal(n)
    ta, ca = R.ccf_direct(x1, x2, fs)
    tb, cb = R.ccf_fft(x1, x2, fs)
    assert np.array_equal(ta, tb)
    assert np.max(np.abs(ca - cb)) < 1e-14
    assert abs(R.ccf_direct(x1, x1, fs)[1][n - 1] - 1.0) < 1e-12     # zero lag of an auto-ccf


def test_analytic_known_answers():
    """identities derivable from the reference code."""
    fs, L = 1e6, 1024
    w = R.window_vector("Hanning", L)
    assert abs(w.sum() - L / 2) < 1e-9 and abs((w ** 2).sum() - 3 * L / 8) < 1e-9
    assert abs(fs * (w ** 2).sum() / w.sum() ** 2 - 1464.84375) < 1e-6
    n = 2 ** 15
    t = np.arange(n) / fs
    amp, kbin = 0.37, 100
    x = amp * np.sin(2 * np.pi * kbin * fs / L * t)
    # tbounds off the record ends -> no reflect padding (which would break the tone's phase)
    o = R.fft_pwelch(t, x, x, [t[1], t[-1]], tper=L / fs * (1 + 1e-9), windowoverlap=0.5,
                     windowfunction="Hanning")
    assert o["nwins"] == L and not o["reflect"]
    assert abs(o["Lxx"][kbin] - amp) < 1e-9                       # bin-centred sine -> Lxx = A
    assert np.allclose(np.abs(o["Cxy"][1:-1]), 1.0) and np.allclose(o["phi_xy"][kbin], 0.0)
    rng = np.random.default_rng(3)
    sig = 1.7
    x = sig * rng.standard_normal(2 ** 18)
    t = np.arange(len(x)) / fs
    o = R.fft_pwelch(t, x, x, Navr=511, windowoverlap=0.5, windowfunction="Hanning")
    assert abs(np.mean(o["Pxx"][5:-5].real) / (2 * sig ** 2 / fs) - 1) < 0.02   # white noise: 2 s^2/Fs


def test_cross_spectral_matrix_rows_are_fft_pwelch_calls():
    rng = np.random.default_rng(11)
    n, C = 2 ** 13, 3
    t = np.arange(n) / 1e6
    s = rng.standard_normal((n, C))
    fr, G = R.cross_spectral_matrix(t, s, 15)
    for i in range(C):
        o = R.fft_pwelch(t, s[:, i], s, Navr=15, windowoverlap=0.5, windowfunction="Hanning")
        assert relerr(G[:, i, :], o["Pxy"]) < 1e-12
    assert np.allclose(G, np.conj(np.transpose(G, (0, 2, 1))))     # Hermitian in (i, j)


def test_fft_deriv_restatement_matches_reference():
    """the spectral derivative (fft_analysis.py:1453-1588) against the
    reference's own outputs (oracle/gen_golden.py fft_deriv)."""
    g = load_golden("fft_deriv")
    for n in (256, 4096):
        x, s = g["x_%d" % n], g["s_%d" % n]
        for tag, kw in (("def", {}), ("naive", dict(modified=False)), ("ham", dict(window=np.hamming)),
                        ("mean", dict(detrend=lambda v: v - v.mean()))):
            d, xo = R.fft_deriv(s, x, **kw)
            assert relerr(d, g["d_%s_%d" % (tag, n)]) <= 1e-12, (tag, n)
            assert relerr(xo, g["xo_%s_%d" % (tag, n)]) <= 1e-14
    d, xo = R.fft_deriv(g["s_256"])
    assert relerr(d, g["d_noxx_256"]) <= 1e-12 and relerr(xo, g["xo_noxx_256"]) <= 1e-14


def test_oracle_cog_matches_reference_fixtures():
    """Doppler.cog restatement against the unmodified reference (tests/golden/cog.npz)."""
    from oracle.restatement import cog
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "cog.npz"))
    fs = 2.0e6
    for tag in "abcd":
        x = g[tag + "_x"]
        assert cog(x, fs) == float(g[tag + "_cog"])
        assert cog(x, fs, fmin=5.0e4, fmax=4.0e5) == float(g[tag + "_cog_band"])
        assert cog(x, fs, fmin=3.0e6) == 0.0
    x = g["s_x"]
    ref = g["s_cog"]
    got = np.asarray([cog(x[i:i + 512], fs) for i in range(0, len(x) - 511, 256)])
    assert np.array_equal(got, ref)
