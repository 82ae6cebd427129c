"""Drop-in ``stft`` / ``specgram`` (spectrogram.py:49-168) on top of K2."""
import numpy as np
import torch

from . import engine as _eng
from .fft_analysis import fftanal

__all__ = ["specgram", "stft"]


def specgram(t, s, wl=512, hanning=True, overlap=True, windowAverage=None, **kwargs):
    """spectrogram.py:49-134: sqrt(8/3) |FFT(hanning(wl) * block)|^2 / wl for 50 %-overlapped
    blocks (symmetric numpy Hanning, no detrend, two-sided, natural bin order).
    Returns (time, frequency, spectrogram[wl, nWindows]) as float64."""
    precision = kwargs.pop("precision", None)
    device = kwargs.pop("device", None)
    if windowAverage is not None:
        # the reference sizes its output with a float here and fails on Python 3 (:113-116)
        raise TypeError("'float' object cannot be interpreted as an integer")
    is_t = isinstance(s, torch.Tensor)
    s = s.reshape(-1) if is_t else np.asarray(s).flatten()
    n = s.shape[0]
    dt = np.abs(t[1] - t[0])
    nwin = 2 * (n - (n % wl)) // wl - 1 if overlap else (n - (n % wl)) // wl - 1      # :74-78
    print("nWindows = ", nwin)
    if wl & (wl - 1) or wl < 16 or wl > 8192:
        raise NotImplementedError("wl = %d: the GPU path needs a power-of-two window of 16..8192" % wl)
    hop = wl // 2 if overlap else wl
    win = np.hanning(wl) if hanning else np.ones(wl)
    scale = np.full(wl, (np.sqrt(8.0 / 3.0) if hanning else 1.0) / wl)
    precision = _eng.resolve_precision(precision, s)
    dev = _eng.require_cuda(device)
    with torch.cuda.device(dev):
        x = _eng.to_device(s, precision, dev).reshape(1, -1).contiguous()
        rmap = _eng.record_map(n, 0, 0, 0, n)
        out, _ = _eng.stft(x, rmap, wl, hop, 0, max(nwin, 0), win, precision, bin_first=0, nbins=wl,
                           bin_scale=scale, power_out=True)
        spec = out[0].t().to(torch.float64).cpu().numpy()                # (wl, nWindows)
    f = np.fft.fftfreq(wl, dt)
    if not overlap:
        time = np.linspace(t[0] + wl * dt / 2, t[0] + wl * dt * ((nwin - 1) + 1 / 2), num=nwin)
    else:
        time = np.linspace(t[0] + wl * dt / 2, t[0] + wl * dt * ((nwin / 2 - 1) + 1 / 2), num=nwin)
    return time, f, np.ascontiguousarray(spec)


def stft(tt, y_in, tper=None, returnclass=True, **kwargs):
    """spectrogram.py:140-168: fftanal.init(tper=...) + .stft()."""
    if tper is None:
        tper = (tt[-1] - tt[0]) / 20
        if tper < tt[2] - tt[1]:
            print("check your stft window size")
    Ystft = fftanal()
    Ystft.init(tt, y_in, tper=tper, **kwargs)
    Ystft.stft()
    if returnclass:
        return Ystft
    twin = np.linspace(tt[0], tt[-1], num=Ystft.Navr, endpoint=True)
    return twin, Ystft.freq, Ystft.Xseg
