"""Per-segment statistics of the Welch path as fused device reductions.

The reference materialises every segment's spectra (``Pxx_seg``, ``Pyy_seg``, ``Pxy_seg``, ``phixy_seg``:
fft_analysis.py:397-460) and leaves their statistics to the caller: the variance over the segments is commented
out and replaced by the stub ``(P / sqrt(Navr))**2`` (:449-458, :503-505); ``mean_angle`` (:1334-1377) averages
``phixy_seg``.  ``welch_segment_statistics`` returns those statistics from ONE pass over the records
(csrc/segstats.cu: twelve running sums per bin and channel pair) without writing anything per-segment --
for BASELINE config 3 that is 17 GB of spectra that never exist.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from . import dist as _dist
from . import engine as _eng
from .windows import windows

_lib_ = _lib.lib


def welch_segment_statistics(tvec, sigx, sigy, tbounds=None, Navr=None, windowoverlap=None, windowfunction=None,
                             detrend_style=None, onesided=None, angle_range=0.5 * np.pi, **kwargs):
    """Segment statistics of ``fft_pwelch(tvec, sigx, sigy, ...)`` (same arguments, same segmentation, same
    normalisation) -> Struct with, per frequency bin (and y channel),

      freq, Navr
      Pxx, Pyy, Pxy          means over the segments (what fft_pwelch returns)
      varPxx, varPyy         np.var(Pxx_seg, axis=0), np.var(Pyy_seg, axis=1).T   (fft_analysis.py:449-451)
      varPxy                 np.var(Pxy_seg.real) + 1j np.var(Pxy_seg.imag)        (:458)
      phi_mean, varphi_mean  mean_angle(phixy_seg, dim=segments, angle_range=...)  (:1334-1377)
      mca, msa, vca, vsa     the component means / variances behind them

    Real records, power-of-two window lengths 64 .. 8192 (the fused kernels); mean / linear / no detrend."""
    from .fft_analysis import _Plan, fftanal

    precision = kwargs.pop("precision", None)
    device = kwargs.pop("device", None)
    group = kwargs.pop("group", None)
    if detrend_style is None:
        detrend_style = 1
    if onesided is None:
        onesided = True
    ntv = len(tvec)
    nx = sigx.shape[0]
    if isinstance(sigy, torch.Tensor):
        ny = sigy.shape[1] if (sigy.dim() == 2 and sigy.shape[1] == ntv) else sigy.shape[0]
    else:
        sy = np.shape(sigy)
        ny = sy[1] if (len(sy) == 2 and sy[1] == ntv) else sy[0]
    plan = _Plan(tvec, nx, ny, tbounds, Navr, windowoverlap, windowfunction, kwargs)
    nwins, nfft, K = plan.nwins, plan.nfft, plan.Navr
    if not _eng.pow2_window(nfft) or nfft < 64 or nfft > 8192:
        raise NotImplementedError("welch_segment_statistics: window length %d; the fused reductions take powers of "
                                  "two in 64 .. 8192 (use the lazy *_seg fields of fft_pwelch otherwise)" % nfft)
    win = windows(plan.windowfunction, nwins=nwins, verbose=False,
                  **({"beta": kwargs["beta"]} if "beta" in kwargs else {}))
    S1, S2, NENBW, ENBW = fftanal._getNorms(win, plan.Nnyquist, plan.Fs)
    precision = _eng.resolve_precision(precision, sigx, sigy)
    dev = _eng.require_cuda(device)
    hop = nwins - plan.noverlap
    if hop <= 0:
        raise ValueError("windowoverlap leaves no hop between segments")
    if (K - 1) * hop + nwins > plan.nana:
        raise ValueError("operands could not be broadcast together: segment %d runs past the analysed record" % (K - 1))
    refl = nwins - 1 if plan.reflect else 0
    rank, world = 0, 1
    if group is not None:
        rank, world = torch.distributed.get_rank(group), torch.distributed.get_world_size(group)
    lp = _dist.local_plan(K, hop, nwins, plan.nana, plan.i0, refl, rank, world)
    mode = _lib.DETREND_MEAN if detrend_style > 0 else (_lib.DETREND_LINEAR if detrend_style < 0 else _lib.DETREND_NONE)
    nbin = nfft // 2 + 1
    with torch.cuda.device(dev):
        x = _eng.to_device(sigx, precision, dev).reshape(-1).contiguous()
        y = _eng.channels_first(_eng.to_device(sigy, precision, dev), ntv)
        nch = y.shape[0]
        rmap = _eng.record_map(plan.nana, plan.i0, refl, 0, x.shape[0])
        wp = _eng.WelchPlan(nfft, hop, nch, precision, win, dev)
        wp.reset()
        if mode != _lib.DETREND_NONE:
            wp.detrend_stats(x, y, rmap, lp["stats_span"][0], lp["stats_span"][1])
            wp.solve_detrend(mode, plan.nana, group)
        seg = _lib.Segmentation(nfft, hop, lp["k0"], lp["k1"] - lp["k0"])
        code = _eng._CODE[precision]
        need = _lib.check(_lib_.pyfft_segment_moments_workspace_bytes(code, C.byref(seg), nch))
        ws = torch.empty(max(int(need), 16), dtype=torch.uint8, device=dev)
        mom = torch.zeros((nch, 12, nbin), dtype=torch.float64, device=dev)
        cx = C.c_void_p(wp.coef.data_ptr()) if wp.coef is not None else C.c_void_p(0)
        cy = C.c_void_p(wp.coef.data_ptr() + 16) if wp.coef is not None else C.c_void_p(0)
        _lib.check(_lib_.pyfft_segment_moments(code, C.c_void_p(x.data_ptr()), C.c_void_p(y.data_ptr()), y.stride(0),
                                               nch, C.byref(rmap), C.byref(seg), C.c_void_p(wp.window.data_ptr()),
                                               cx, cy, C.c_void_p(mom.data_ptr()), C.c_void_p(ws.data_ptr()),
                                               ws.numel(), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        if group is not None and world > 1:
            torch.distributed.all_reduce(mom, group=group)
        m = mom.cpu().numpy()
    out = stats_from_moments(m, K, 1.0 / S1 ** 2 / ENBW, bool(onesided), nfft, plan.Fs, plan.Nnyquist, angle_range)
    out.nwins, out.ENBW = nwins, ENBW
    return out


def stats_from_moments(m, K, scale, onesided, nfft, Fs, nnyq, angle_range=0.5 * np.pi):
    """Host bookkeeping behind welch_segment_statistics: m = (nch, 12, nfft/2 + 1) sums over the K segments as
    written by pyfft_segment_moments -> Struct of means, variances and the circular mean of the segment phase
    (side selection, doubling and scaling as fft_analysis.py:402-446)."""
    from .fft_analysis import Struct, varphi
    nch = m.shape[0]
    if onesided:
        F = nnyq
        sel = np.arange(F)
        wgt = np.ones(F)
        wgt[1:-1] = 2.0                                            # :414-416 (nfft even: the last bin stays x1)
        sign = np.ones(F)
        freq = Fs * np.arange(F) / nfft
    else:
        k = (np.arange(nfft) + nfft - nfft // 2) % nfft            # fftshift order (:423-427)
        sel = np.where(k <= nfft // 2, k, nfft - k)
        sign = np.where(k <= nfft // 2, 1.0, -1.0)                 # Hermitian mirror: Pxy conjugated
        wgt = np.ones(nfft)
        freq = np.fft.fftshift(np.fft.fftfreq(nfft, 1.0 / Fs))
    s = (scale * wgt)[None, :]
    mm = m[:, :, sel] / float(K)                                   # (nch, 12, F) means over the segments
    out = Struct()
    out.Navr, out.freq = K, freq
    out.Pxx = (mm[0, 0] * s[0]).astype(np.complex128)
    out.Pyy = (mm[:, 1] * s).T.astype(np.complex128)
    out.Pxy = ((mm[:, 2] + 1j * sign[None, :] * mm[:, 3]) * s).T
    out.varPxx = (mm[0, 4] - mm[0, 0] ** 2) * s[0] ** 2
    out.varPyy = ((mm[:, 5] - mm[:, 1] ** 2) * s ** 2).T
    out.varPxy = (((mm[:, 6] - mm[:, 2] ** 2) + 1j * (mm[:, 7] - mm[:, 3] ** 2)) * s ** 2).T
    out.mca, out.msa = mm[:, 8].T, (sign[None, :] * mm[:, 9]).T
    out.vca, out.vsa = (mm[:, 10] - mm[:, 8] ** 2).T, (mm[:, 11] - mm[:, 9] ** 2).T
    with np.errstate(all="ignore"):
        out.phi_mean, out.varphi_mean = varphi(Pxy_real=out.mca, Pxy_imag=out.msa, varPxy_real=out.vca,
                                               varPxy_imag=out.vsa, angle_range=angle_range)
    if nch == 1:
        for nm in ("Pyy", "Pxy", "varPyy", "varPxy", "mca", "msa", "vca", "vsa", "phi_mean", "varphi_mean"):
            setattr(out, nm, getattr(out, nm).reshape(-1))
    return out
