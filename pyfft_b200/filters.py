"""Resampling front-end of the path: ``filters.upsample`` (filters.py:20-34),
the step ``fftanal.resample`` (fft_analysis.py:2363-2373) takes before the spectra when the two signals arrive
on different time bases.

The reference interpolates with ``pybaseutils.utils.interp(xi, yi, ei, xo)``, a package that is not in this image
(and not vendored by the reference): linear interpolation is restated from its call sites and the comment at
filters.py:22 ("cubic" was planned, linear is what runs), evaluated like ``numpy.interp``.  Parity for this entry
is UNPINNED against the reference and pinned against the numpy restatement (oracle/restatement.py:upsample).
``downsample`` / ``downsample_efficient`` (scipy.signal filtfilt pre-filters) are not part of the path's front-end
(fftanal.resample only ever up-samples the shorter record) and are not provided.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from . import engine as _eng


def upsample(u_t, Fs, Fs_new, plotit=False, precision=None, device=None, return_device=False):
    """Up-sample record(s) ``u_t`` (n,) or (n, nch) from Fs to Fs_new by linear interpolation on the device
    (pyfft_resample_linear): tt = arange(n)/Fs, ti = arange(tt[0], tt[-1], 1/Fs_new) (filters.py:24-27)."""
    if plotit:
        raise NotImplementedError("plotit=True: plotting is out of scope")
    nt = int(u_t.shape[0])
    tt_last = (nt - 1) / float(Fs)
    m = int(np.ceil((tt_last - 0.0) / (1.0 / float(Fs_new))))          # len(numpy.arange(0, tt[-1], 1/Fs_new))
    m = max(m, 0)
    dev = _eng.require_cuda(device)
    prec = _eng.resolve_precision(precision, u_t)
    with torch.cuda.device(dev):
        x = _eng.to_device(u_t, prec, dev)
        one_d = x.dim() == 1
        xc = x.reshape(1, nt) if one_d else x.t().contiguous()          # channels first
        nch = xc.shape[0]
        out = torch.empty((nch, m), dtype=xc.dtype, device=dev)
        if m > 0:
            _lib.check(_lib.lib.pyfft_resample_linear(_eng._CODE[prec], C.c_void_p(xc.data_ptr()), xc.stride(0), nch, nt,
                                                      float(Fs), float(Fs_new), m, C.c_void_p(out.data_ptr()),
                                                      out.stride(0),
                                                      C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        res = out.reshape(m) if one_d else out.t()
        if return_device:
            return res
        return res.cpu().numpy()
