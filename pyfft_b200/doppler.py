"""Spectral centre of gravity of a (Doppler) signal (Doppler.py:43-81): the
power-weighted mean frequency of one record (``cog``) and along a running window (``cog_sliding``, the first loop
of the reference's ``cogspec``), as per-window FFT powers (K2, pyfft_stft with power output; any window length,
real or complex records) followed by one weighted reduction per window (pyfft_spectral_centroid).

The second half of the reference's ``cogspec`` (Doppler.py:83-96) calls a function ``psd`` that is defined nowhere
in the reference and prints with Python-2 syntax: it cannot run and is not reproduced.  ``cogspec``'s windowing
helper (pybaseutils.utils.sliding_window_1d) is not in the image: windows start every ``floor((1-ov)*win)`` samples
and must fit in the record, as its call site implies (parity of ``cog_sliding`` unpinned; ``cog`` is pinned).
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from . import engine as _eng


def _weights(n, fs, fmin, fmax):
    """wnum, wden (natural FFT bin order) reproducing Doppler.py:51-58 LITERALLY, including its band selection:
    the frequency vector is filtered by fmin <= |f| <= fmax, and the spectrum is then indexed with a mask computed
    from the ALREADY FILTERED frequencies -- all true, of the filtered length -- so the first m entries of the
    fftshift-ed spectrum are paired with the m selected frequencies."""
    if fmax is None:
        fmax = fs
    freq = np.fft.fftshift(np.fft.fftfreq(n, 1.0 / fs))
    nat = (np.arange(n) + n - n // 2) % n                  # natural bin behind entry i of an fftshift-ed array
    wnum, wden = np.zeros(n), np.zeros(n)
    if fmin is not None:
        freq = freq[(np.abs(freq) >= fmin) * (np.abs(freq) <= fmax)]
    m = len(freq)
    wnum[nat[:m]] = freq
    wden[nat[:m]] = 1.0
    return wnum, wden


def _centroids(x, fs, win, hop, nseg, fmin, fmax, precision, device):
    dev = _eng.require_cuda(device)
    prec = _eng.resolve_precision(precision, x)
    with torch.cuda.device(dev):
        xd = _eng.to_device(x, prec, dev, complex_ok=True).reshape(1, -1).contiguous()
        n = xd.shape[1]
        rmap = _eng.record_map(n, 0, 0, 0, n)
        P, _ = _eng.stft(xd, rmap, win, hop, 0, nseg, np.ones(win), prec, power_out=True)
        wnum, wden = _weights(win, fs, fmin, fmax)
        wn = torch.from_numpy(wnum).to(dev)
        wd = torch.from_numpy(wden).to(dev)
        out = torch.empty(nseg, dtype=torch.float64, device=dev)
        P = P.reshape(nseg, win)
        _lib.check(_lib.lib.pyfft_spectral_centroid(_eng._CODE[prec], C.c_void_p(P.data_ptr()), nseg, win,
                                                    C.c_void_p(wn.data_ptr()), C.c_void_p(wd.data_ptr()),
                                                    C.c_void_p(out.data_ptr()),
                                                    C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        return out.cpu().numpy()


def cog(x, fs, fmin=None, fmax=None, precision=None, device=None):
    """Centre of gravity of the power spectrum of x (Doppler.py:43-61)."""
    n = len(x)
    if n == 0:
        return 0.0
    return float(_centroids(x, fs, n, n, 1, fmin, fmax, precision, device)[0])


def cog_sliding(t, x, fs, win=512, ov=0.5, fmin=None, fmax=None, precision=None, device=None):
    """(tcog, coge): mean time and spectral centre of gravity of every window of `win` samples advancing by
    floor((1 - ov) * win) samples (the first loop of cogspec, Doppler.py:70-81), all windows in one batched
    transform."""
    step = int(np.floor((1.0 - ov) * win))
    if step <= 0:
        raise ValueError("ov leaves no step between windows")
    n = len(x)
    nseg = (n - win) // step + 1 if n >= win else 0
    if nseg <= 0:
        return np.zeros(0), np.zeros(0)
    coge = _centroids(x, fs, win, step, nseg, fmin, fmax, precision, device)
    t = np.asarray(t, dtype=np.float64)
    cs = np.concatenate(([0.0], np.cumsum(t)))
    starts = np.arange(nseg) * step
    tcog = (cs[starts + win] - cs[starts]) / win
    return tcog, coge
