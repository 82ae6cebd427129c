"""Drop-in for the reference's ccf.py (which cannot be imported on Python 3: dft.py:200 is a
py2 print statement): normalised cross-covariance and FFT correlation through pyfft_xcorr
(K3b) -- two long forward FFTs, X conj(Y), one long inverse FFT, lag reorder and
normalisation fused into the last pass.
"""
import numpy as np
import torch

from . import dist as _dist
from . import engine


def _dev1(x, prec, dev):
    t = engine.to_device(x, prec, dev)
    if t.dim() != 1:
        raise ValueError("expected a 1-D signal")
    return t


def ccf(x1, x2, fs, precision=None, device=None, return_device=False, group=None, gather=False):
    """(tau, co): normalised cross-covariance at all 2N-1 lags (ccf.py:66-77).
    x2 may be (pairs, N): every row is correlated against x1 (one shared forward FFT).
    ``group``: the rows of x2 are sharded over the ranks of a torch.distributed group in
    contiguous blocks (no collective); each rank gets its rows, or all of them with
    ``gather=True`` (one all_gather)."""
    dev = engine.require_cuda(device)
    if not isinstance(x1, torch.Tensor):
        x1 = np.asarray(x1)
    if not isinstance(x2, torch.Tensor):
        x2 = np.asarray(x2)
    prec = engine.resolve_precision(precision, x1, x2)
    a = _dev1(x1, prec, dev)
    b = engine.to_device(x2, prec, dev)
    single = b.dim() == 1
    if single:
        b = b[None, :]
    npts = a.shape[0]
    if b.shape[1] != npts:
        raise ValueError("operands could not be broadcast together: %d vs %d samples" % (npts, b.shape[1]))
    npairs = b.shape[0]
    if group is not None and not single:
        b = _dist.shard_channels(b, group)
    if b.shape[0] == 0:
        co = torch.empty((0, 2 * npts - 1), dtype=engine._TORCH_REAL[prec], device=b.device)
    else:
        co = engine.xcorr(a, b, prec, normalise=1)
    if group is not None and not single and gather:
        co = _dist.gather_channels(co, npairs, group)
    lags = np.arange(-npts + 1, npts)
    tau = -lags / float(fs)                                      # ccf.py:70
    if single:
        co = co[0]
    if return_device:
        return tau, co
    return tau, co.cpu().numpy().astype(np.float64)


def ccf_sh(x1, x2, fs, nav, ss=None, precision=None, device=None):
    """Average of ccf over sliding windows of `nav` samples (ccf.py:80-104).  The reference
    takes its windows from pybaseutils.sliding_window_1d (not vendored, parity unpinned); they
    are restated as consecutive windows with hop ss (default nav)."""
    dev = engine.require_cuda(device)
    prec = engine.resolve_precision(precision, np.asarray(x1), np.asarray(x2))
    a = _dev1(x1, prec, dev)
    b = _dev1(x2, prec, dev)
    nav = int(nav)
    ss = nav if ss is None else int(ss)
    nwin = (a.shape[0] - nav) // ss + 1
    if nwin < 1:
        raise ValueError("window longer than the record")
    wa = a.unfold(0, nav, ss)[:nwin].contiguous()
    wb = b.unfold(0, nav, ss)[:nwin].contiguous()
    co = engine.xcorr(wa, wb, prec, normalise=1)
    lags = np.arange(-nav + 1, nav)
    return -lags / float(fs), co.mean(dim=0).cpu().numpy().astype(np.float64)


def fftcorr(x, y, precision=None, device=None):
    """Circular FFT correlation 'Rxy <==> X'Y' of two equal power-of-two length sequences
    (ccf.py:239-248, with the mis-unpacked first line read as intended):
    ifft(conj(fft(x)) * fft(y))."""
    dev = engine.require_cuda(device)
    x = np.asarray(x)
    y = np.asarray(y)
    prec = engine.resolve_precision(precision, x, y)
    if x.shape != y.shape or x.ndim != 1:
        raise ValueError("fftcorr expects two 1-D sequences of equal length")
    zx = engine.to_device(x.astype(np.complex128 if prec == "fp64" else np.complex64), prec, dev, True)
    zy = engine.to_device(y.astype(np.complex128 if prec == "fp64" else np.complex64), prec, dev, True)
    X = engine.fft_c2c(zx[None, :], prec)
    Y = engine.fft_c2c(zy[None, :], prec)
    r = engine.fft_c2c(torch.conj(X).resolve_conj() * Y, prec, inverse=True)
    return r[0].cpu().numpy().astype(np.complex128)


def fftconv(x, y, precision=None, device=None):
    """Circular FFT convolution 'x*y <==> XY' of two equal power-of-two length sequences
    (ccf.py:227-235): ifft(fft(x) * fft(y))."""
    dev = engine.require_cuda(device)
    x = np.asarray(x)
    y = np.asarray(y)
    prec = engine.resolve_precision(precision, x, y)
    if x.shape != y.shape or x.ndim != 1:
        raise ValueError("fftconv expects two 1-D sequences of equal length")
    ct = np.complex128 if prec == "fp64" else np.complex64
    zx = engine.to_device(x.astype(ct), prec, dev, True)
    zy = engine.to_device(y.astype(ct), prec, dev, True)
    X = engine.fft_c2c(zx[None, :], prec)
    Y = engine.fft_c2c(zy[None, :], prec)
    r = engine.fft_c2c(X * Y, prec, inverse=True)
    return r[0].cpu().numpy().astype(np.complex128)


def convolve_fft(a, b, mode="valid", precision=None, device=None):
    """numpy.convolve(a, b, mode) through the FFT -- what ccf.py:267-283 intends (its padding
    helper divides by a float, ccf.py:264, and cannot run).  Equal-length real inputs; the
    convolution is the correlation with the second operand reversed (ccf.py:303 states the same
    identity the other way round)."""
    b = np.asarray(b)
    if b.ndim != 1:
        raise ValueError("Can only vectorize vectors")                  # ccf.py:258-259
    return cross_correlation_fft(a, b[::-1].copy(), mode=mode, precision=precision, device=device)


def cross_correlation_fft(a, b, mode="valid", precision=None, device=None):
    """numpy.correlate(a, b, mode) through the FFT -- what ccf.py:285-303 intends ("similar to
    numpy.correlate, but faster"; the reference's own padding helper divides by a float,
    ccf.py:264, and cannot run).  Equal-length real inputs."""
    dev = engine.require_cuda(device)
    a = np.asarray(a)
    b = np.asarray(b)
    if a.ndim != 1 or b.ndim != 1:
        raise ValueError("Can only vectorize vectors")                  # ccf.py:258-259
    if len(a) != len(b):
        raise NotImplementedError("cross_correlation_fft: equal lengths only on the GPU path")
    prec = engine.resolve_precision(precision, a, b)
    full = engine.xcorr(_dev1(a, prec, dev), _dev1(b, prec, dev)[None, :], prec, normalise=0)[0]
    full = full.cpu().numpy().astype(np.float64)
    n = len(a)
    if mode == "full":
        return full
    if mode == "same":
        lo = (n - 1) // 2
        return full[lo:lo + n]
    if mode == "valid":
        return full[n - 1:n]
    raise ValueError("mode must be 'valid', 'same' or 'full'")
