"""GPU parity of K4 (multi-channel cross-spectral matrix, BASELINE config 5 shape): against the
oracle's float64 restatement and against row-by-row fft_pwelch calls (how the reference builds
the matrix, HeatPulse_Funcs.py:576-583)."""
import numpy as np
import pytest

from conftest import relerr

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("nch,n,navr,precision", [(4, 2 ** 15, 31, "fp64"), (3, 2 ** 15, 31, "fp64"),
                                                  (8, 2 ** 16, 63, "fp32"), (32, 2 ** 17, 31, "fp32"),
                                                  (16, 2 ** 16, 31, "fp32"), (64, 2 ** 15, 15, "fp32"),
                                                  (32, 2 ** 16, 15, "fp64")])
def test_csm_matches_oracle(nch, n, navr, precision, cuda_device):
    from oracle import restatement as R
    from oracle import signals
    from pyfft_b200 import cross_spectral_matrix
    t, sig = signals.c5_mixture(n=n, nch=nch)
    sig = sig + np.arange(nch)[:, None] * 0.5                    # DC offsets: detrend is exercised
    if precision == "fp32":
        sig = sig.astype(np.float32)
    fr, Gr = R.cross_spectral_matrix(t, sig.astype(np.float64).T, navr, 0.5, "Hanning", 1)
    f, G = cross_spectral_matrix(t, sig, Navr=navr, windowoverlap=0.5, windowfunction="Hanning",
                                 detrend_style=1, precision=precision)
    assert G.shape == Gr.shape and G.dtype == np.complex128
    assert relerr(f, fr) < 1e-14
    tol = 1e-12 if precision == "fp64" else 1e-5
    assert relerr(G, Gr) <= tol
    # Hermitian: G[f, i, j] = conj(G[f, j, i])
    assert relerr(G, np.conj(np.transpose(G, (0, 2, 1)))) <= tol


def test_csm_rows_equal_fft_pwelch_calls(cuda_device):
    from oracle import signals
    from pyfft_b200 import cross_spectral_matrix, fft_pwelch
    t, sig = signals.c5_mixture(n=2 ** 16, nch=8)
    f, G = cross_spectral_matrix(t, sig, Navr=15, windowoverlap=0.5, detrend_style=-1)
    for i in (0, 5):
        fi, Pxy, Pxx, Pyy, Cxy, phi, info = fft_pwelch(t, sig[i], sig, Navr=15, windowoverlap=0.5,
                                                       windowfunction="Hanning", detrend_style=-1, plotit=False)
        assert relerr(G[:, i, :], Pxy) <= 1e-12
        assert relerr(G[:, i, i], Pxx) <= 1e-12


def test_csm_config5_cut(cuda_device):
    """32 channels, nfft 8192, 50 % overlap (config 5 at 2^21 samples per channel)."""
    import torch
    from oracle import restatement as R
    from oracle import signals
    from pyfft_b200 import cross_spectral_matrix
    n = 2 ** 21
    t, sig = signals.c5_mixture(n=n, nch=32)
    sig = sig.astype(np.float32)
    navr = n // 4096 - 1
    fr, Gr = R.cross_spectral_matrix(t, sig.astype(np.float64).T, navr, 0.5, "Hanning", 1)
    f, G = cross_spectral_matrix(t, torch.from_numpy(sig), Navr=navr, windowoverlap=0.5, precision="fp32")
    assert G.shape == (4096, 32, 32)
    assert relerr(G, Gr) <= 1e-5


def test_csm_several_chunks(cuda_device, monkeypatch):
    """The spectra workspace is processed in chunks of segments; a small cap forces several chunks
    (odd tail included) and the sums must not change."""
    from oracle import signals
    from pyfft_b200 import cross_spectral_matrix
    t, sig = signals.c5_mixture(n=2 ** 17, nch=32)
    sig = sig.astype(np.float32)
    f, G1 = cross_spectral_matrix(t, sig, Navr=63, windowoverlap=0.5, precision="fp32")
    monkeypatch.setenv("PYFFT_CSM_CHUNK_BYTES", str(10 * 32 * 2049 * 8))       # 10 segments per chunk
    f, G2 = cross_spectral_matrix(t, sig, Navr=63, windowoverlap=0.5, precision="fp32")
    assert relerr(G2, G1) <= 2e-6
