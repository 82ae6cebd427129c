"""pyfft_b200 -- B200-native (sm_100a) accelerator for PYFFT's segmented spectral
estimation hot path.

Public names follow the reference package (FFT/__init__.py:20-31): ``fft_pwelch``,
``fftanal``, ``detrend_none/mean/linear``, ``stft``, ``specgram``, plus ``windows``,
``hilbert`` and ``ccf``.  Importing this package loads libpyfft_b200.so; it raises if
the CUDA extension has not been built (there is no CPU fallback).
"""
from . import _lib                      # noqa: F401  (fails loudly when the .so is missing)
from .windows import windows            # noqa: F401
from .fft_analysis import (fft_pwelch, fftanal, cross_spectral_matrix, fft_deriv, fftinfosc, Cxy_Cxy2,  # noqa: F401
                           detrend_none, detrend_mean, detrend_linear, getNpeaks, integratespectra,
                           mean_angle, varcoh, varphi)
from .laplace import laplace_1d         # noqa: F401,E402
from .segstats import welch_segment_statistics   # noqa: F401,E402
from .filters import upsample            # noqa: F401,E402
from .doppler import cog, cog_sliding    # noqa: F401,E402

from .spectrogram import stft, specgram  # noqa: F401,E402
from .hilbert import hilbert, hilbert_1d  # noqa: F401,E402
from .ccf import ccf, ccf_sh, fftcorr, fftconv, convolve_fft, cross_correlation_fft  # noqa: F401,E402

__version__ = "0.1.0"
