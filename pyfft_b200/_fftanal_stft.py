"""STFT half of the ``fftanal`` object API (fft_analysis.py:1805-2203), on top of K2.

Attached to ``fftanal`` by fft_analysis.py.  Same attribute names and shapes as the
reference: ``Xseg`` (Navr, nf) complex128 in [V/sqrt(Hz)], ``tseg``, ``freq``, ``Xpow``,
``Xfft``; ``Pstft`` uses X conj(Y) (fft_analysis.py:1960 -- the opposite sign convention to
fft_pwelch, mirrored on purpose).
"""
import numpy as np
import torch

from . import _lib
from . import engine as _eng


def _mode(style):
    if style is None or style == 0:
        return _lib.DETREND_NONE
    return _lib.DETREND_MEAN if style > 0 else _lib.DETREND_LINEAR


def _segment_times(tvec, nwins, hop, navr):
    """tt[g] = mean(tvec[g*hop : g*hop + nwins]) (fft_analysis.py:2163) via one cumulative sum."""
    cs = np.concatenate(([0.0], np.cumsum(np.asarray(tvec, dtype=np.float64))))
    a = np.arange(navr) * hop
    return (cs[a + nwins] - cs[a]) / nwins


def fft_win(self, sig, tvec=None, detrendwin=False):
    """fft_analysis.py:2126-2203: windowed, scaled per-segment spectra of one signal.
    Returns (tseg, freq, Xseg (Navr, nf) complex128, pseg (Navr,))."""
    n = len(sig)
    if tvec is None:
        tvec = np.linspace(0.0, 1.0, n)
    nwins, navr, noverlap = self.nwins, self.Navr, self.noverlap
    Fs = self.__Fs__(tvec)
    nfft = nwins
    hop = nwins - noverlap
    if (navr - 1) * hop + nwins > n:
        raise ValueError("operands could not be broadcast together: segment %d runs past the record"
                         % (navr - 1))
    cplx = sig.is_complex() if isinstance(sig, torch.Tensor) else np.iscomplexobj(sig)
    precision = _eng.resolve_precision(getattr(self, "precision", None), sig)
    dev = _eng.require_cuda(getattr(self, "device", None))
    with torch.cuda.device(dev):
        x = _eng.to_device(sig, precision, dev, complex_ok=True).reshape(1, -1).contiguous()
        rmap = _eng.record_map(n, 0, 0, 0, n)
        mode = _mode(self.detrendstyle)
        coef = None
        if mode != _lib.DETREND_NONE and not detrendwin:
            if cplx:
                if mode == _lib.DETREND_LINEAR:
                    raise NotImplementedError("linear detrend of complex records")
                parts = torch.stack((x[0].real, x[0].imag)).contiguous()
                coef = _eng.detrend_coefficients(parts, rmap, mode, precision, n)
            else:
                coef = _eng.detrend_coefficients(x, rmap, mode, precision, n)
        if detrendwin and mode == _lib.DETREND_LINEAR:
            raise NotImplementedError("per-segment linear detrend")
        nf = self.Nnyquist if self.onesided else nfft
        scale = np.full(nf, 1.0 / self.S1 / np.sqrt(self.ENBW))          # :2197, :2202
        if self.onesided:
            scale[1:-1] *= np.sqrt(2)                                     # :2186
            if nfft % 2:
                scale[-1] *= np.sqrt(2)
            first = 0
        else:
            first = nfft - nfft // 2                                      # fftshift order (:2191-2192)
        out, pseg = _eng.stft(x, rmap, nfft, hop, 0, navr, self.win, precision, coef=coef,
                              per_segment=bool(detrendwin and mode != _lib.DETREND_NONE),
                              bin_first=first, nbins=nf, bin_scale=scale, want_pseg=True)
        self._last_seg_dev = out[0]
        Xseg = out[0].to(torch.complex128).cpu().numpy()
        pw = pseg[0].cpu().numpy()
    try:
        self.tper = tvec[nwins] - tvec[0]                                 # :2160-2161
    except IndexError:
        pass
    tt = _segment_times(tvec, nwins, hop, navr)
    freq = np.fft.fftfreq(nfft, 1.0 / Fs)
    freq = freq[:self.Nnyquist] if self.onesided else np.fft.fftshift(freq)
    pw = pw * (1.0 / Fs) / self.S2                                        # trapz step, then /S2 (:2174, :2198)
    return tt, freq, Xseg, pw


def Xstft(self):
    sig = self.__trimsig__(self.sigx, self.ibounds)
    tvec = self.__trimsig__(self.tvec, self.ibounds)
    self.tseg, self.freq, self.Xseg, self.Xpow = self.fft_win(sig, tvec)
    self._Xseg_dev = self._last_seg_dev
    self.Xfft = np.mean(self.Xseg, axis=0)
    return self.freq, self.Xseg


def Ystft(self):
    sig = self.__trimsig__(self.sigy, self.ibounds)
    tvec = self.__trimsig__(self.tvec, self.ibounds)
    self.tseg, self.freq, self.Yseg, self.Ypow = self.fft_win(sig, tvec)
    self._Yseg_dev = self._last_seg_dev
    self.Yfft = np.mean(self.Yseg, axis=0)
    return self.freq, self.Yseg


def _np128(t):
    return t.to(torch.complex128).cpu().numpy()


def Pstft(self):
    """fft_analysis.py:1944-1971.  The per-segment products are elementwise on the
    materialised (Navr, nf) spectra; they are formed on the device tensors K2 left behind."""
    from .fft_analysis import Cxy_Cxy2
    amp = np.sqrt(2) if self.onesided else 1.0
    if hasattr(self, "Xseg"):
        X = self._Xseg_dev
        self.Pxx_seg = _np128(X * X.conj())
        self.Lxx_seg = amp * np.sqrt(np.abs(self.ENBW * self.Pxx_seg))
    if hasattr(self, "Yseg"):
        Y = self._Yseg_dev
        self.Pyy_seg = _np128(Y * Y.conj())
        self.Lyy_seg = amp * np.sqrt(np.abs(self.ENBW * self.Pyy_seg))
    if hasattr(self, "Xseg") and hasattr(self, "Yseg"):
        self.Pxy_seg = _np128(self._Xseg_dev * self._Yseg_dev.conj())    # X conj(Y) (:1960)
        self.Lxy_seg = amp * np.sqrt(np.abs(self.ENBW * self.Pxy_seg))
        self.phixy_seg = np.angle(self.Pxy_seg)
        self.Cxy_seg, self.Cxy2_seg = Cxy_Cxy2(self.Pxx_seg, self.Pyy_seg, self.Pxy_seg)


def averagewins(self):
    """fft_analysis.py:1975-2018: plain means over the segments (double sums on the device)."""
    from .fft_analysis import Cxy_Cxy2
    K = self.Navr
    precision = "fp32" if self._Xseg_dev.dtype == torch.complex64 else "fp64"
    Y = self._Yseg_dev[None] if hasattr(self, "Pyy_seg") else None
    sxx, syy, sxy = _eng.stft_products(self._Xseg_dev, Y, False, precision)
    if hasattr(self, "Pxx_seg"):
        self.Pxx = (sxx.cpu().numpy() / K).astype(np.complex128)
        self.varPxx = (self.Pxx / np.sqrt(K)) ** 2.0
    if Y is not None:
        self.Pyy = (syy[0].cpu().numpy() / K).astype(np.complex128)
        self.varPyy = (self.Pyy / np.sqrt(K)) ** 2.0
        self.Pxy = sxy[0].cpu().numpy() / K
        self.varPxy = (self.Pxy / np.sqrt(K)) ** 2.0
    if hasattr(self, "Pxy"):
        self.phi_xy = np.angle(self.Pxy)
        # the reference's Cxy_Cxy2 needs 2-D Pyy and raises IndexError for this 1-D class path
        # (fft_analysis.py:1669); mirrored by calling the same helper
        self.Cxy, self.Cxy2 = Cxy_Cxy2(self.Pxx, self.Pyy, self.Pxy)
        self.varPhxy = (np.sqrt(1.0 - self.Cxy2) / np.sqrt(2.0 * K * self.Cxy)) ** 2.0
        self.varCxy = ((1 - self.Cxy2) / np.sqrt(2 * K)) ** 2.0
        self.varCxy2 = 4.0 * self.Cxy2 * self.varCxy


def pwelch(self):
    self.Xstft()
    if not self.nosigy:
        self.Ystft()
    self.Pstft()
    self.averagewins()


def stft(self):
    if self.useMLAB:
        raise NotImplementedError("useMLAB=True (scipy.signal.stft backend) is out of scope")
    self.pwelch()


def _inverse(self, spec, halve):
    """sqrt(nfft) * irfft / ifft(ifftshift) along the LAST axis, fftshift-ed (:1845-1870)."""
    dev = _eng.require_cuda(getattr(self, "device", None))
    a = np.asarray(spec, dtype=np.complex128)
    flat = a.reshape(-1, a.shape[-1])                                    # (ncol, F)
    with torch.cuda.device(dev):
        t = torch.from_numpy(np.ascontiguousarray(flat.T)).to(dev)       # [F][ncol]
        wp = _eng.WelchPlan(self.nwins, max(1, self.nwins - self.noverlap), 1, "fp64", self.win, dev)
        r = wp.correlation(t, onesided=bool(self.onesided), halve_interior=halve).cpu().numpy()
    return np.ascontiguousarray(r.T).reshape(a.shape[:-1] + (self.nwins,))


def _crosscorr_generic(self, suffix):
    nfft = self.nwins
    if nfft & (nfft - 1) or nfft < 16 or nfft > 8192:
        raise NotImplementedError("nwins = %d: power-of-two windows of 16..8192 only" % nfft)
    for param in ("Pxx", "Pyy", "Pxy"):
        if hasattr(self, param + suffix):
            tmp = _inverse(self, getattr(self, param + suffix), True)
            zero = tmp[..., nfft // 2].copy()                            # zero lag after the fftshift
            if param == "Pxx":
                setattr(self, "Ex" + suffix, zero)
            if param == "Pyy":
                setattr(self, "Ey" + suffix, zero)
            setattr(self, "R" + param[1:] + suffix, tmp)
    if hasattr(self, "Rxy" + suffix):
        ex, ey = getattr(self, "Ex" + suffix), getattr(self, "Ey" + suffix)
        setattr(self, "corrcoef" + suffix, getattr(self, "Rxy" + suffix).copy()
                / (np.ones((self.nch, 1), dtype=ey.dtype) * np.sqrt(ex * ey)))
    self.lags = (np.asarray(range(1, nfft + 1), dtype=int) - self.Nnyquist) / self.Fs


def crosscorr(self):
    """fft_analysis.py:1840-1878."""
    _crosscorr_generic(self, "")


def crosscorr_stft(self):
    """fft_analysis.py:1880-1920."""
    _crosscorr_generic(self, "_seg")


def convert2amplitudes(self):
    """fft_analysis.py:2022-2048 (O(nfreq) host bookkeeping on averaged spectra)."""
    for param in ("Pxx", "Pyy", "Pxy"):
        if hasattr(self, param):
            tmp = np.sqrt(np.abs(self.ENBW * getattr(self, param)))
            if self.onesided:
                tmp[1:-1] = np.sqrt(2) * tmp[1:-1]
                if self.nwins % 2:
                    tmp[-1] = np.sqrt(2) * tmp[-1]
            setattr(self, "L" + param[1:], tmp)
            setattr(self, "varL" + param[1:],
                    (tmp ** 2) * (getattr(self, "var" + param) / np.abs(getattr(self, param)) ** 2))


def detrend(self, sig):
    from .fft_analysis import fftanal
    return fftanal._detrend_func(detrend_style=self.detrendstyle)(sig)


METHODS = dict(fft_win=fft_win, Xstft=Xstft, Ystft=Ystft, Pstft=Pstft, averagewins=averagewins,
               pwelch=pwelch, stft=stft, crosscorr=crosscorr, crosscorr_stft=crosscorr_stft,
               convert2amplitudes=convert2amplitudes, detrend=detrend)
