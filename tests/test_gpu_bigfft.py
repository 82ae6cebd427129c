"""GPU parity of K3b (long transforms): hilbert / hilbert_1d against the reference's golden
outputs, ccf against the direct O(N^2) definition (ccf.py:66-77) and the oracle's FFT identity,
pyfft_fft_c2c against numpy.fft for every pass decomposition (1, 2 and 3 passes)."""
import numpy as np
import pytest

from conftest import load_golden, relerr

pytestmark = pytest.mark.gpu

TOLZ = {"fp64": 1e-12, "fp32": 2e-6}


@pytest.mark.parametrize("precision", ["fp64", "fp32"])
def test_hilbert_matches_reference_golden(precision, cuda_device):
    from pyfft_b200 import hilbert, hilbert_1d
    g = load_golden("hilbert")
    for n in (32, 1024, 4096):
        u = g["u_%d" % n] if precision == "fp64" else g["u_%d" % n].astype(np.float32)
        z = hilbert(u, precision=precision)
        assert z.shape == (n,) and z.dtype == np.complex128
        assert relerr(z, g["z_%d" % n]) <= TOLZ[precision]
        assert relerr(hilbert_1d(u, precision=precision), g["z1d_%d" % n]) <= TOLZ[precision]
    ub = g["u_batch"] if precision == "fp64" else g["u_batch"].astype(np.float32)
    assert relerr(hilbert(ub, precision=precision), g["z_batch"]) <= TOLZ[precision]
    # axis 0 of a time-major array
    assert relerr(hilbert(ub.T.copy(), axes=0, precision=precision), g["z_batch"].T) <= TOLZ[precision]


@pytest.mark.parametrize("precision", ["fp64", "fp32"])
def test_hilbert_any_length_matches_reference(precision, cuda_device):
    """n = 33 (odd: the reference's own mask rule, hilbert.py:47-51, 63-64) and n = 1000 through
    Bluestein on the long-transform kernels, against the outputs of the unmodified reference."""
    from pyfft_b200 import hilbert, hilbert_1d
    g = load_golden("hilbert")
    for n in (33, 1000):
        u = g["u_%d" % n] if precision == "fp64" else g["u_%d" % n].astype(np.float32)
        z = hilbert(u, precision=precision)
        assert z.shape == (n,) and z.dtype == np.complex128
        assert relerr(z, g["z_%d" % n]) <= TOLZ[precision], (n, relerr(z, g["z_%d" % n]))
        assert relerr(hilbert_1d(u, precision=precision), g["z_%d" % n]) <= TOLZ[precision]
        env = hilbert(u, precision=precision, envelope=True)
        assert relerr(env, np.abs(g["z_%d" % n])) <= TOLZ[precision]


@pytest.mark.parametrize("n", [2, 3, 5, 17, 100, 1023, 1025, 4097, 30011, 65537])
def test_fft_any_matches_numpy(n, cuda_device):
    """pyfft_fft_any against numpy.fft for awkward lengths (primes, 2^k +- 1), real and complex rows,
    forward and inverse."""
    import torch
    from pyfft_b200 import engine
    rng = np.random.default_rng(n)
    a = rng.standard_normal((3, n)) + 1j * rng.standard_normal((3, n))
    z = torch.from_numpy(a).to(cuda_device)
    assert relerr(engine.fft_any(z, "fp64").cpu().numpy(), np.fft.fft(a, axis=-1)) <= 1e-12
    assert relerr(engine.fft_any(z, "fp64", inverse=True).cpu().numpy(), np.fft.ifft(a, axis=-1)) <= 1e-12
    r = torch.from_numpy(a.real.copy()).to(cuda_device)
    assert relerr(engine.fft_any(r, "fp64").cpu().numpy(), np.fft.fft(a.real, axis=-1)) <= 1e-12
    z32 = z.to(torch.complex64)
    assert relerr(engine.fft_any(z32, "fp32").cpu().numpy(), np.fft.fft(a.astype(np.complex64).astype(complex), axis=-1)) <= 2e-6


@pytest.mark.parametrize("log2n", [14, 19, 22])
def test_hilbert_long_records_and_envelope(log2n, cuda_device):
    """2- and 3-pass decompositions against the oracle; the fused envelope equals |analytic|."""
    from oracle import restatement as R
    from pyfft_b200 import hilbert
    n = 1 << log2n
    rng = np.random.default_rng(40 + log2n)
    t = np.arange(n) / 1.0e6
    u = (1 + 0.5 * np.sin(2 * np.pi * 1e3 * t)) * np.sin(2 * np.pi * 1e5 * t) + 0.05 * rng.standard_normal(n)
    ref = R.hilbert(u)
    assert relerr(hilbert(u), ref) <= 1e-12
    u32 = u.astype(np.float32)
    ref32 = R.hilbert(u32.astype(np.float64))
    z32 = hilbert(u32)
    assert relerr(z32, ref32) <= 1e-5
    env = hilbert(u32, envelope=True)
    assert env.dtype == np.float64 and relerr(env, np.abs(ref32)) <= 1e-5


@pytest.mark.parametrize("precision", ["fp64", "fp32"])
def test_ccf_matches_direct_definition(precision, cuda_device):
    from oracle import restatement as R
    from pyfft_b200 import ccf
    rng = np.random.default_rng(7)
    fs, n = 1e5, 2048                                   # ccf_test recipe (ccf.py:139-148)
    t = np.arange(n) / fs
    x1 = np.sin(2 * np.pi * 1e3 * t) + rng.standard_normal(n) + 3.0
    x2 = np.sin(2 * np.pi * 1e3 * t - 50 * np.pi / 180) + rng.standard_normal(n) - 1.0
    if precision == "fp32":
        x1, x2 = x1.astype(np.float32), x2.astype(np.float32)
    tau_ref, co_ref = R.ccf_direct(x1.astype(np.float64), x2.astype(np.float64), fs)
    tau, co = ccf(x1, x2, fs, precision=precision)
    assert co.shape == (2 * n - 1,) and np.array_equal(tau, tau_ref)
    tol = 1e-12 if precision == "fp64" else 2e-6
    assert np.max(np.abs(co - co_ref)) <= tol
    # odd length, and several channels against one reference in a single call
    x3 = rng.standard_normal((3, 1001)) + np.array([[0.0], [5.0], [-2.0]])
    x0 = rng.standard_normal(1001)
    _, cm = ccf(x0, x3, fs)
    for c in range(3):
        assert np.max(np.abs(cm[c] - R.ccf_direct(x0, x3[c], fs)[1])) <= 1e-12
    # zero lag of an auto-ccf is 1
    assert abs(ccf(x0, x0, fs)[1][1000] - 1.0) < 1e-12


def test_ccf_long_record_peak_at_known_delay(cuda_device):
    """2^21 samples (padded transform 2^22, three passes): FFT identity oracle + known delay."""
    from oracle import restatement as R
    from pyfft_b200 import ccf
    rng = np.random.default_rng(8)
    n, d = 1 << 21, 37 * 5
    s = rng.standard_normal(n + d)
    x1 = s[d:] + 0.1 * rng.standard_normal(n)
    x2 = s[:n] + 0.1 * rng.standard_normal(n)          # x2 lags x1 by d samples
    tau_ref, co_ref = R.ccf_fft(x1, x2, 1.0e6)
    tau, co = ccf(x1, x2, 1.0e6)
    assert np.max(np.abs(co - co_ref)) <= 1e-12
    assert np.argmax(co) == np.argmax(co_ref)
    _, co32 = ccf(x1.astype(np.float32), x2.astype(np.float32), 1.0e6)
    assert np.max(np.abs(co32 - co_ref)) <= 1e-5


def test_fftcorr_and_cross_correlation_fft(cuda_device):
    from pyfft_b200 import cross_correlation_fft, fftcorr
    rng = np.random.default_rng(9)
    x, y = rng.standard_normal(256), rng.standard_normal(256)
    ref = np.fft.ifft(np.conj(np.fft.fft(x)) * np.fft.fft(y))          # ccf.py:239-248
    assert relerr(fftcorr(x, y), ref) <= 1e-12
    for mode in ("full", "same", "valid"):
        assert relerr(cross_correlation_fft(x, y, mode=mode), np.correlate(x, y, mode)) <= 1e-12
    from pyfft_b200 import convolve_fft, fftconv
    assert relerr(fftconv(x, y), np.fft.ifft(np.fft.fft(x) * np.fft.fft(y))) <= 1e-12      # ccf.py:227-235
    for mode in ("full", "same", "valid"):
        assert relerr(convolve_fft(x, y, mode=mode), np.convolve(x, y, mode)) <= 1e-12


@pytest.mark.parametrize("precision", ["fp64", "fp32"])
@pytest.mark.parametrize("log2n", [4, 7, 9, 10, 13, 18, 19, 21])
def test_fft_c2c_matches_numpy(precision, log2n, cuda_device):
    import torch
    from pyfft_b200 import engine
    n = 1 << log2n
    batch = 3 if log2n <= 13 else 1
    rng = np.random.default_rng(log2n)
    z = rng.standard_normal((batch, n)) + 1j * rng.standard_normal((batch, n))
    if precision == "fp32":
        z = z.astype(np.complex64)
    zd = torch.from_numpy(z).to(cuda_device)
    tol = 1e-12 if precision == "fp64" else 3e-6
    ref = np.fft.fft(z.astype(np.complex128), axis=-1)
    out = engine.fft_c2c(zd, precision).cpu().numpy()
    assert relerr(out, ref) <= tol
    back = engine.fft_c2c(torch.from_numpy(ref.astype(z.dtype)).to(cuda_device), precision, inverse=True)
    assert relerr(back.cpu().numpy(), z) <= tol


@pytest.mark.parametrize("nrec,log2n", [(2, 10), (5, 14), (3, 19)])
def test_hilbert_two_records_per_transform(nrec, log2n, cuda_device):
    """Batches ride one complex transform per two records (odd tail included): every record must
    equal its own single-record result and the oracle; envelope = |analytic|."""
    from oracle import restatement as R
    from pyfft_b200 import hilbert
    n = 1 << log2n
    rng = np.random.default_rng(60 + log2n)
    t = np.arange(n) / 1.0e6
    u = np.stack([(1 + 0.3 * c) * np.sin(2 * np.pi * (5e3 + 1e3 * c) * t) + 0.1 * rng.standard_normal(n) + c
                  for c in range(nrec)])
    z = hilbert(u)
    assert z.shape == (nrec, n)
    for c in range(nrec):
        assert relerr(z[c], R.hilbert(u[c])) <= 1e-12
        assert relerr(z[c], hilbert(u[c])) <= 1e-12
    u32 = u.astype(np.float32)
    z32 = hilbert(u32)
    env = hilbert(u32, envelope=True)
    for c in range(nrec):
        ref = R.hilbert(u32[c].astype(np.float64))
        assert relerr(z32[c], ref) <= 1e-5
        assert relerr(env[c], np.abs(ref)) <= 1e-5


def test_ccf_two_records_per_transform(cuda_device):
    """One reference record against 4 and 5 records (pairs share a transform): each row equals the
    single-pair call."""
    from pyfft_b200 import ccf
    rng = np.random.default_rng(61)
    n = 3000
    x0 = rng.standard_normal(n) + 2.0
    for nrec in (4, 5):
        y = rng.standard_normal((nrec, n)) + np.arange(nrec)[:, None]
        y[1, 17:] += 0.8 * x0[:-17]
        for prec, tol in (("fp64", 1e-12), ("fp32", 3e-6)):
            xs = x0 if prec == "fp64" else x0.astype(np.float32)
            ys = y if prec == "fp64" else y.astype(np.float32)
            _, cm = ccf(xs, ys, 1.0e5, precision=prec)
            assert cm.shape == (nrec, 2 * n - 1)
            for c in range(nrec):
                _, c1 = ccf(xs, ys[c], 1.0e5, precision=prec)
                assert np.max(np.abs(cm[c] - c1)) <= tol


def test_fft_deriv_matches_reference_golden(cuda_device):
    """fft_deriv (fft_analysis.py:1453-1588) against the reference's own outputs
    -- default (modified wavenumber), naive wavenumber, Hamming window, mean detrend, index grid."""
    from pyfft_b200 import detrend_mean, fft_deriv
    g = load_golden("fft_deriv")
    for n in (256, 4096):
        x, s = g["x_%d" % n], g["s_%d" % n]
        for tag, kw in (("def", {}), ("naive", dict(modified=False)), ("ham", dict(window=np.hamming)),
                        ("mean", dict(detrend=detrend_mean))):
            d, xo = fft_deriv(s, x, **kw)
            assert d.dtype == np.float64 and d.shape == (n,)
            assert relerr(d, g["d_%s_%d" % (tag, n)]) <= 1e-11, (tag, n)
            assert relerr(xo, g["xo_%s_%d" % (tag, n)]) <= 1e-14
        d32, _ = fft_deriv(s.astype(np.float32), x)
        assert relerr(d32, g["d_def_%d" % n]) <= 5e-3, relerr(d32, g["d_def_%d" % n])
    d, xo = fft_deriv(g["s_256"])
    assert relerr(d, g["d_noxx_256"]) <= 1e-11 and relerr(xo, g["xo_noxx_256"]) <= 1e-14
    with pytest.raises(NotImplementedError):
        fft_deriv(g["s_256"][:200], g["x_256"][:200])              # not a power of two: no CPU fallback
    with pytest.raises(NotImplementedError):
        fft_deriv(g["s_256"], g["x_256"], lowpass=0.01 / (g["x_256"][1] - g["x_256"][0]))   # would down-sample


@pytest.mark.parametrize("log2n,batch", [(6, 1), (12, 3), (19, 2), (21, 1)])
def test_spectral_derivative_kernel_against_numpy(log2n, batch, cuda_device):
    """pyfft_spectral_derivative (1-, 2- and 3-pass transforms, odd batches: two records per
    transform) against the numpy line it replaces, fft_analysis.py:1528-1553."""
    import torch
    from pyfft_b200 import engine
    n = 1 << log2n
    rng = np.random.default_rng(70 + log2n)
    u = rng.standard_normal((batch, n)).cumsum(axis=1) / np.sqrt(n)
    dx = 1.0 / n * 0.9
    L = n * dx
    k = 2.0 * np.pi * np.fft.fftfreq(n, d=dx / L)
    for modified in (True, False):
        wn = (1j * np.sin(k * dx) / dx if modified else 1j * k) / L
        ref = np.real(np.fft.ifft(wn * np.fft.fft(u, axis=1), axis=1))
        got = engine.spectral_derivative(torch.from_numpy(u).to(cuda_device), dx, "fp64", modified=modified)
        assert relerr(got.cpu().numpy(), ref) <= 1e-11, (modified, relerr(got.cpu().numpy(), ref))


def test_paired_records_of_very_different_amplitude_keep_their_digits(cuda_device):
    """two real records share one complex transform (Hilbert, ccf with a shared first record); each is
    scaled by its own power of two first, so a 10^6 amplitude mismatch between neighbours costs no digits
    (it used to cost log10 of the ratio in fp32)."""
    from oracle import restatement as R
    from pyfft_b200 import ccf, hilbert
    n = 1 << 16
    rng = np.random.default_rng(77)
    t = np.arange(n) / 1.0e5
    base = [(1 + 0.5 * np.sin(2 * np.pi * 40.0 * t + c)) * np.sin(2 * np.pi * (3e3 + 500 * c) * t) + 0.05 * rng.standard_normal(n)
            for c in range(4)]
    amp = np.array([1.0, 1.0e-6, 1.0e5, 3.0e-3])
    u = (np.stack(base) * amp[:, None]).astype(np.float32)
    z = hilbert(u, precision="fp32")
    env = hilbert(u, precision="fp32", envelope=True)
    for c in range(4):
        ref = R.hilbert(u[c].astype(np.float64))
        assert relerr(z[c], ref) <= 1e-5, (c, relerr(z[c], ref))            # the fp32 contract, per record
        assert relerr(env[c], np.abs(ref)) <= 1e-5, c
    x0 = base[0].astype(np.float32)
    _, cm = ccf(x0, u, 1.0e5, precision="fp32")
    for c in range(4):
        _, c1 = ccf(x0.astype(np.float64), u[c].astype(np.float64), 1.0e5, precision="fp64")
        assert np.max(np.abs(cm[c] - c1)) <= 1e-5, (c, np.max(np.abs(cm[c] - c1)))
