"""Drop-in ``fft_pwelch`` / ``fftanal`` on top of the sm_100a kernels.

Same names, positional order, keyword arguments, return tuple and ``fftinfo``
attributes as the reference's fft_analysis.py (``fft_pwelch`` :36-648,
``class fftanal`` :1695-2203, ``Cxy_Cxy2`` :1662-1688, ``fftinfosc`` :794-826).
The host does the O(1)/O(nfft) bookkeeping in float64 exactly like the reference
(segmentation integers, window, S1/S2/ENBW, variances, amplitude spectra); every
O(N) step -- detrend statistics, segment gather, window, FFT, products, segment
means, coherence, correlations -- runs on the GPU through the C ABI
(include/pyfft_b200.h).  There is no CPU fallback.

Extra, opt-in keyword arguments (defaults keep the reference's behaviour):
    precision='fp32'|'fp64'   arithmetic on the GPU (default: float32 inputs -> fp32,
                              anything else -> fp64 like the reference)
    device=...                CUDA device
    group=...                 torch.distributed process group: this rank analyses its
                              share of the segments and the partial sums are combined
                              with one all-reduce (inputs are the full record on
                              every rank unless ``shard`` says otherwise)
    return_device=True        keep the tuple outputs as device tensors

Deliberate differences from the reference (documented in DESIGN.md):
  * ``useMLAB=True`` (matplotlib.mlab backend) and ``plotit=True`` raise
    NotImplementedError -- out of scope for the hot path.
  * the per-segment fields of fftinfo (Pxx_seg, Xfft_seg, ...) are computed lazily
    on first access instead of being materialised on every call.
"""
import os
import threading

import numpy as np
import torch

from . import _lib
from . import dist as _dist
from . import engine as _eng
from .windows import windows

__all__ = ["fft_pwelch", "fftanal", "fftinfosc", "Cxy_Cxy2", "detrend_none", "detrend_mean",
           "detrend_linear", "Struct", "getNpeaks", "integratespectra", "mean_angle", "varcoh", "varphi"]


class Struct(object):
    """Attribute bag with the two methods of pybaseutils.Struct the reference uses
    (fft_analysis.py:22, 1785-1790)."""

    def __init__(self, d=None):
        if d is not None:
            if not isinstance(d, dict):
                d = d.dict_from_class()
            self.__dict__.update(d)

    def dict_from_class(self):
        return dict((k, v) for k, v in self.__dict__.items())


class fftinfosc(Struct):
    """Result container of fft_pwelch (fft_analysis.py:794-826): same attribute names."""

    _LAZY = ("Pxx_seg", "Pyy_seg", "Pxy_seg", "Xfft_seg", "Yfft_seg", "phixy_seg", "varphi_seg")
    # O(nfreq) statistics derived from the averaged spectra (fft_analysis.py:496-540, :585-597): computed on the
    # host at first access (or by dict_from_class) instead of on every call -- twenty numpy passes that most
    # callers never read cost 0.5 ms of an 11 ms end-to-end call
    _LAZY_DERIVED = ("varCxy", "varCxy2", "varPxx", "varPyy", "varPxy", "varPhxy", "Lxx", "Lyy", "Lxy", "varLxx",
                     "varLyy", "varLxy", "Ex", "Ey", "corrcoef", "lags")

    def __init__(self):
        f8, c16 = np.float64, np.complex128
        for name, dt in (("S1", f8), ("S2", f8), ("NENBW", f8), ("ENBW", f8), ("freq", f8),
                         ("Pxx", c16), ("Pyy", c16), ("Pxy", c16), ("Cxy", c16),
                         ("varcoh", c16), ("phi_xy", f8), ("varphi", f8), ("Lxx", c16),
                         ("Lyy", c16), ("Lxy", c16), ("varLxx", c16), ("varLyy", c16),
                         ("varLxy", c16)):
            setattr(self, name, np.array([], dtype=dt))
        self._segment_loader = None
        self._derived_loader = None

    def _defer_derived(self, loader):
        """The derived statistics are produced by loader() -> dict at first access."""
        for nm in fftinfosc._LAZY_DERIVED:
            self.__dict__.pop(nm, None)
        self.__dict__["_derived_loader"] = loader

    def __getattr__(self, name):
        # only reached when normal lookup fails: materialise per-segment spectra / derived statistics on demand
        if name in fftinfosc._LAZY and self.__dict__.get("_segment_loader") is not None:
            self.__dict__.update(self.__dict__["_segment_loader"]())
            self.__dict__["_segment_loader"] = None
            return self.__dict__[name]
        if name in fftinfosc._LAZY_DERIVED and self.__dict__.get("_derived_loader") is not None:
            loader = self.__dict__["_derived_loader"]
            self.__dict__["_derived_loader"] = None
            self.__dict__.update(loader())
            return self.__dict__[name]
        raise AttributeError(name)

    def __getstate__(self):
        """Pickling: the derived statistics are computed, the per-segment loader (a closure over the caller's
        records, which it keeps alive and re-reads at first access -- read the `*_seg` fields before mutating or
        dropping those arrays) is left behind: an unpickled result has no lazy per-segment fields."""
        if self.__dict__.get("_derived_loader") is not None:
            getattr(self, "varCxy")
        d = dict(self.__dict__)
        d["_segment_loader"] = None
        d["_derived_loader"] = None
        return d

    def __setstate__(self, d):
        self.__dict__.update(d)

    def dict_from_class(self):
        if self.__dict__.get("_derived_loader") is not None:
            getattr(self, "varCxy")
        if self.__dict__.get("_segment_loader") is not None:
            getattr(self, "Pxx_seg")
        return dict((k, v) for k, v in self.__dict__.items() if not k.startswith("_"))


# detrend helpers exported under the reference's names (FFT/__init__.py:21); they are
# host utilities for callers -- the hot path uses the device statistics kernel.
def detrend_none(x, axis=None):
    return x


def detrend_mean(x, axis=None):
    x = np.asarray(x)
    if axis is None:
        axis = 0
    return x - x.mean(axis, keepdims=True)


def detrend_linear(x, axis=None):
    x = np.asarray(x)
    if axis is None:
        axis = 0
    n = x.shape[axis]
    tt = np.arange(n, dtype=np.float64) - 0.5 * (n - 1)
    shp = [1] * x.ndim
    shp[axis] = n
    tt = tt.reshape(shp)
    xm = x.mean(axis, keepdims=True)
    b = (tt * (x - xm)).sum(axis, keepdims=True) / (tt * tt).sum()
    return x - xm - b * tt


def Cxy_Cxy2(Pxx, Pyy, Pxy, ibg=None):
    """Complex and mean-squared coherence of averaged spectra (fft_analysis.py:1662-1688).
    O(nfreq) host helper kept for callers; fft_pwelch itself gets both from the finalize
    kernel."""
    Pxx = np.atleast_2d(np.array(Pxx))
    Pyy = np.array(Pyy)
    Pxy = np.array(Pxy)
    if np.size(Pxx, axis=1) != np.size(Pyy, axis=1):
        Pxx = Pxx.T * np.ones((1, np.size(Pyy, axis=1)), dtype=Pxx.dtype)
    norm = np.abs(Pxx) * np.abs(Pyy)
    Cxy2 = Pxy * np.conj(Pxy) / norm
    Cxy = Pxy / np.sqrt(norm)
    if ibg is None:
        return Cxy, Cxy2
    iCxy = np.imag(Cxy) / (1.0 - np.real(Cxy))
    shifted = np.real(Cxy - np.mean(Cxy[:, ibg], axis=-1))
    return iCxy, shifted / (1.0 - shifted)


# --------------------------------------------------------------------------- #
# segmentation bookkeeping (host, float64 / Python ints -- verbatim formulas)
# --------------------------------------------------------------------------- #

def _sample_rate(tvec):
    return (len(tvec) - 1) / (tvec[-1] - tvec[0])                  # fft_analysis.py:136


def _index_bounds(tvec, tbounds, Fs):
    i0 = int(np.floor(Fs * (tbounds[0] - tvec[0])))                # :143
    i1 = int(np.floor(1 + Fs * (tbounds[1] - tvec[0])))            # :144
    return i0, i1


class _Plan(object):
    """Everything fft_pwelch decides before touching the data (fft_analysis.py:102-246)."""

    def __init__(self, tvec, nx, ny, tbounds, Navr, windowoverlap, windowfunction, kwargs):
        if windowfunction is None:
            windowfunction = "Hanning"
        if windowoverlap is None:
            windowoverlap = windows(windowfunction, verbose=False)
        if tbounds is None:
            tbounds = [tvec[0], tvec[-1]]
        self.windowfunction, self.windowoverlap, self.tbounds = windowfunction, windowoverlap, tbounds
        ntv = len(tvec)
        self.Fs = _sample_rate(tvec)
        self.i0, self.i1 = _index_bounds(tvec, tbounds, self.Fs)
        nsig = len(range(ntv)[self.i0:self.i1])                    # numpy slice semantics (:149)
        calc_navr = Navr is None
        if nx != ny:
            raise NotImplementedError("different-length sigx/sigy (the reference's nT-model branch, "
                                      "fft_analysis.py:170-176) is not supported")
        if "minFreq" in kwargs:
            kwargs["tper"] = 2.0 / kwargs["minFreq"]               # :180-181
        if "tper" in kwargs:
            nwins = int(self.Fs * kwargs["tper"])                  # :182-183
            if not calc_navr:
                # the reference leaves `calcNavr` unbound here and dies with NameError (:102, :208)
                raise NameError("calcNavr")
        else:
            if Navr is None:
                Navr = 8                                           # :185-186
            calc_navr = False
            nwins = fftanal._getNwins(nsig, Navr, windowoverlap)   # :189
        noverlap = fftanal._getNoverlap(nwins, windowoverlap)      # :194
        self.reflect = (self.i0 == 0 and self.i1 == ntv)           # :197-198
        if self.reflect:
            nsig = nx + 2 * (nwins - 1)                            # :202-204
        if calc_navr:
            Navr = fftanal._getNavr(nsig, nwins, noverlap)         # :208-209
        if nwins >= nsig:
            Navr, nwins = 1, nsig                                  # :214-216
        self.nwins, self.noverlap, self.Navr = int(nwins), int(noverlap), int(Navr)
        self.nfft = self.nwins
        self.Nnyquist = fftanal._getNnyquist(self.nfft)
        # the analysed record is (reflect-padded sigx)[i0:i1] (:353-354)
        nsrc = nx + 2 * (int(nwins) - 1) if self.reflect else nx
        self.nana = len(range(nsrc)[self.i0:self.i1])


def _host(t):
    return t.cpu().numpy()


def _host_many(tensors):
    """Several small device results in ONE device->host transfer (each .cpu() is a synchronising
    round trip of ~25 us): complex tensors travel as (re, im) float64 pairs."""
    flat = [torch.view_as_real(t).reshape(-1) if t.is_complex() else t.reshape(-1).to(torch.float64)
            for t in tensors]
    host = torch.cat(flat).cpu().numpy()
    out, pos = [], 0
    for t, f in zip(tensors, flat):
        a = host[pos:pos + f.numel()]
        pos += f.numel()
        if t.is_complex():
            a = a.view(np.complex128)
        out.append(a.reshape(tuple(t.shape)).copy())
    return out


_plans = threading.local()


# Data-aware selection of the tensor-core route (fft_pwelch(tensor_cores=None)): after the pass, every
# auto-spectrum must keep min over bins >= TC_DYNAMIC_RANGE / Navr * mean over bins, else the call is
# redone on the fp32 CUDA-core kernel.  The fp16 operands leave a white floor of ~1e-7 of the mean bin
# power; a bin with power P sees a coherence error ~ sqrt(floor / P / Navr) (zero-mean cross term) +
# floor / P (bias).  Calibrated in tests/test_gpu_tc_parity.py.
TC_DYNAMIC_RANGE = 400.0
TC_MIN_RATIO = 4e-3          # the bias term alone: floor / P <= 1e-4 with margin, whatever Navr
# ... and no channel may repeat exactly after 1..8 hops (pyfft_welch_tc_prepare's repetition measure: the
# noise fraction of the record at those lags; 1e-6 already dithers the fp16 roundings)
TC_MIN_REPETITION = 1e-4
TC_MIN_STATIONARITY = 0.5    # ... and the power must be spread over the record (256 sampled runs)
# fp32 fused-mean route (either kernel): the kernel sees the record with a DC offset d = mean - provisional
# mean; its leak d^2 |W[f]|^2 is accumulated in fp32 with the signal and subtracted afterwards, which costs
# ~2^-23 x leak in absolute error.  Tolerated: leak <= 500 x the bin's own power (coherence, 1e-4) and
# <= 10 x the channel's spectral peak (PSD, 1e-5 of peak); above, the call is redone with the exact mean
# first (two passes over HBM).
FUSED_MEAN_MAX_LEAK = 500.0
FUSED_MEAN_MAX_LEAK_PEAK = 10.0


def _welch_plan(nfft, hop, nch, precision, win, dev, group=None):
    """engine.WelchPlan of this thread for these parameters, reset (device buffers, the uploaded
    window and its transform are reused from call to call: ~0.2 ms of a 0.4 ms computation)."""
    cache = getattr(_plans, "cache", None)
    if cache is None:
        cache = _plans.cache = {}
    key = (int(nfft), int(hop), int(nch), precision, str(dev), getattr(group, "group_name", None))
    hit = cache.get(key)
    if hit is not None and np.array_equal(hit[0], win):
        wp = hit[1]
        wp.reset()
        wp.coef = None
        return wp
    wp = _eng.WelchPlan(nfft, hop, nch, precision, win, dev)
    if len(cache) >= 8:
        cache.clear()
    cache[key] = (np.array(win, dtype=np.float64, copy=True), wp)
    return wp


def _make_segment_loader(sigx, sigy, ntv, plan, win, mode, precision, dev, onesided, scale):
    """Per-segment fields of fftinfo (fft_analysis.py:461-481) are produced on first access by
    the STFT kernel instead of on every call: Xfft_seg (Navr, nfft), Yfft_seg (nch, Navr, nfft)
    raw transforms; Pxx_seg (Navr, nf), Pyy_seg / Pxy_seg (nch, Navr, nf) scaled like the
    averaged spectra; phixy_seg; varphi_seg."""
    def load():
        nw, K = plan.nwins, plan.Navr
        hop = nw - plan.noverlap
        with torch.cuda.device(dev):
            cplx = any((a.is_complex() if isinstance(a, torch.Tensor) else np.iscomplexobj(a)) for a in (sigx, sigy))
            x = _eng.to_device(sigx, precision, dev, complex_ok=True).reshape(1, -1).contiguous()
            y = _eng.channels_first(_eng.to_device(sigy, precision, dev, complex_ok=True), ntv)
            if cplx:
                cdt = torch.complex128 if precision == "fp64" else torch.complex64
                x, y = x.to(cdt), y.to(cdt)
            rmap = _eng.record_map(plan.nana, plan.i0, nw - 1 if plan.reflect else 0, 0, x.shape[1])
            out = {}
            specs = []
            for sig in (x, y):
                if cplx and mode != _lib.DETREND_NONE:
                    parts = torch.view_as_real(sig).permute(0, 2, 1).reshape(2 * sig.shape[0], -1).contiguous()
                    coef = _eng.detrend_coefficients(parts, rmap, mode, precision, plan.nana)
                else:
                    coef = _eng.detrend_coefficients(sig, rmap, mode, precision, plan.nana)
                S, _ = _eng.stft(sig, rmap, nw, hop, 0, K, win, precision, coef=coef)
                specs.append(S.to(torch.complex128))
            X, Y = specs[0][0], specs[1]                               # (K, nfft), (nch, K, nfft)
            F = plan.Nnyquist if onesided else nw
            wgt = torch.full((F,), float(scale), dtype=torch.float64, device=dev)
            if onesided:
                wgt[1:-1] *= 2.0                                       # :414-416
                pick = lambda a: a[..., :F]
            else:
                pick = lambda a: torch.fft.fftshift(a, dim=-1)         # index shuffle only (:425-427)
            out["Xfft_seg"] = _host(X)
            out["Yfft_seg"] = _host(Y)
            out["Pxx_seg"] = _host(pick(X * X.conj()) * wgt)
            out["Pyy_seg"] = _host(pick(Y * Y.conj()) * wgt)
            out["Pxy_seg"] = _host(pick(Y * X.conj()[None]) * wgt)     # Y conj(X) (:393)
        out["phixy_seg"] = np.angle(out["Pxy_seg"])
        out["varphi_seg"] = np.zeros_like(out["phixy_seg"])
        return out
    return load


def _pwelch_generic(sigx, sigy, ntv, plan, win, mode, precision, dev, onesided, scale, lp, group, is_cplx):
    """fft_pwelch for any window length and for complex-valued records (fft_analysis.py:359-446 with
    :123-127, :422-428): segment spectra by pyfft_stft_any in chunks, products summed on the device
    (pyfft_stft_products, Y conj(X) :393), then the O(nfft) bookkeeping of :402-446, :489-520,
    :544-597 in numpy on the summed spectra.  -> (nch, [Pxx, Pyy, Pxy, Cxy, Cxy2, phi, Rxx, Ryy, Rxy, iCxy])."""
    nw, K = plan.nwins, plan.Navr
    hop = nw - plan.noverlap
    refl = nw - 1 if plan.reflect else 0
    k0, k1 = lp["k0"], lp["k1"]
    with torch.cuda.device(dev):
        def prep(sig, two_d):
            t = _eng.to_device(sig, precision, dev, complex_ok=True)
            t = _eng.channels_first(t, ntv) if two_d else t.reshape(1, -1).contiguous()
            if is_cplx and not t.is_complex():
                t = t.to(torch.complex128 if precision == "fp64" else torch.complex64)
            return t
        x, y = prep(sigx, False), prep(sigy, True)
        nch = y.shape[0]
        rmap = _eng.record_map(plan.nana, plan.i0, refl, 0, x.shape[1])

        def coefs(sig):
            if mode == _lib.DETREND_NONE:
                return None
            if not is_cplx:
                return _eng.detrend_coefficients(sig, rmap, mode, precision, plan.nana, group=group,
                                                 j_span=lp["stats_span"])
            if mode == _lib.DETREND_LINEAR:
                raise NotImplementedError("linear detrend of complex records")
            parts = torch.view_as_real(sig).permute(0, 2, 1).reshape(2 * sig.shape[0], -1).contiguous()
            return _eng.detrend_coefficients(parts, rmap, mode, precision, plan.nana, group=group,
                                             j_span=lp["stats_span"])
        cx, cy = coefs(x), coefs(y)
        sxx = torch.zeros(nw, dtype=torch.float64, device=dev)
        syy = torch.zeros((nch, nw), dtype=torch.float64, device=dev)
        sxy = torch.zeros((nch, nw), dtype=torch.complex128, device=dev)
        # chunk of segments: the spectra of all channels of a chunk live on the device at once
        per_seg = (nch + 1) * nw * (16 if precision == "fp64" else 8)
        step = int(max(1, min(k1 - k0, (1 << 28) // per_seg))) if k1 > k0 else 1
        for a in range(k0, k1, step):
            kc = min(step, k1 - a)
            X, _ = _eng.stft(x, rmap, nw, hop, a, kc, win, precision, coef=cx)
            Y, _ = _eng.stft(y, rmap, nw, hop, a, kc, win, precision, coef=cy)
            pxx, pyy, pxy = _eng.stft_products(X[0], Y, True, precision)
            sxx += pxx
            syy += pyy
            sxy += pxy
        if group is not None:
            packed = torch.cat((sxx, syy.reshape(-1), torch.view_as_real(sxy).reshape(-1)))
            torch.distributed.all_reduce(packed, group=group)
            sxx = packed[:nw]
            syy = packed[nw:nw + nch * nw].view(nch, nw)
            sxy = torch.view_as_complex(packed[nw + nch * nw:].view(nch, nw, 2))
        hxx, hyy, hxy = _host_many([sxx, syy, sxy])
    nnyq = plan.Nnyquist
    if onesided:                                                   # :402-421
        wgt = np.ones(nnyq)
        wgt[1:-1] = 2.0
        if nw % 2:
            wgt[-1] = 2.0
        hxx, hyy, hxy = hxx[:nnyq] * wgt, hyy[:, :nnyq] * wgt, hxy[:, :nnyq] * wgt
    else:                                                          # :422-428
        hxx, hyy, hxy = (np.fft.fftshift(a, axes=-1) for a in (hxx, hyy, hxy))
    sc = scale / K                                                 # :432-446
    Pxx = (hxx * sc).astype(np.complex128)
    Pyy = (hyy * sc).T.astype(np.complex128)
    Pxy = (hxy * sc).T
    with np.errstate(all="ignore"):
        Cxy, Cxy2 = Cxy_Cxy2(Pxx, Pyy, Pxy)                        # :489, :1662-1680
    phi = np.arctan2(Pxy.imag, Pxy.real)                           # :520
    corr = []
    for P, halve in ((Pxx, True), (Pyy, True), (Pxy, True), (Cxy, False)):   # :544-597
        if onesided:
            q = np.array(P, dtype=np.complex128, copy=True)
            if halve:
                q[1:-1, ...] *= 0.5
                if nw % 2:
                    q[-1, ...] *= 0.5
            r = np.fft.irfft(q, n=nw, axis=0)
        else:
            r = np.fft.ifft(np.fft.ifftshift(P, axes=0), n=nw, axis=0)
        corr.append(np.fft.fftshift(r * np.sqrt(nw), axes=0))
    return nch, [Pxx, Pyy, Pxy, Cxy, Cxy2, phi] + corr


def fft_pwelch(tvec, sigx, sigy, tbounds=None, Navr=None, windowoverlap=None,
               windowfunction=None, useMLAB=None, plotit=None, verbose=None,
               detrend_style=None, onesided=None, **kwargs):
    """Welch auto-/cross-power spectra, coherence and cross-phase of sigx vs sigy.

    Returns ``(freq, Pxy, Pxx, Pyy, Cxy, phi_xy, fftinfo)`` exactly like the reference
    (fft_analysis.py:36-648): power spectral densities in [V^2/Hz] as complex128,
    ``Pxy = <Y conj(X)>`` (:393), one-sided spectra drop the Nyquist bin (:407-416).
    """
    precision = kwargs.pop("precision", None)
    device = kwargs.pop("device", None)
    group = kwargs.pop("group", None)
    return_device = kwargs.pop("return_device", False)
    two_pass = kwargs.pop("two_pass_detrend", False)     # force the K0 + K1 route (testing)
    tensor_cores = kwargs.pop("tensor_cores", None)      # None: automatic; True / False: force
    # sharded calls (group=): sigx / sigy may hold only this rank's slice of the record -- the samples
    # [shard_offset, shard_offset + len) of the global record that tvec describes (dist.local_plan(...)["src"])
    shard_offset = kwargs.pop("shard_offset", None)
    if useMLAB:
        raise NotImplementedError("useMLAB=True (matplotlib.mlab backend) is out of scope")
    if plotit:
        raise NotImplementedError("plotit=True: plotting is out of scope for the accelerated path; "
                                  "pass plotit=False")
    if verbose is None:
        verbose = False
    if detrend_style is None:
        detrend_style = 1                                          # :117-118
    is_cplx = (sigx.is_complex() if isinstance(sigx, torch.Tensor) else np.iscomplexobj(sigx)) or \
              (sigy.is_complex() if isinstance(sigy, torch.Tensor) else np.iscomplexobj(sigy))
    if onesided is None:
        onesided = not is_cplx                                     # :123-127
    # long pinned host records start crossing PCIe NOW, on a side stream, while the host plans the call (window,
    # norms, segmentation: ~0.2 ms that would otherwise sit in front of a 9.7 ms copy)
    early = []
    if group is None and not is_cplx and isinstance(sigx, torch.Tensor) and not sigx.is_cuda:
        dev0 = _eng.require_cuda(device)
        with torch.cuda.device(dev0):
            early.append(_eng.staged_upload(sigx, sigy, _eng.resolve_precision(precision, sigx, sigy), dev0))
    nx = sigx.shape[0]
    ntv = len(tvec)
    if isinstance(sigy, torch.Tensor):
        ny = sigy.shape[1] if (sigy.dim() == 2 and sigy.shape[1] == ntv) else sigy.shape[0]
    else:
        sy = np.shape(sigy)
        ny = sy[1] if (len(sy) == 2 and sy[1] == ntv) else sy[0]
    nloc = nx
    if shard_offset is not None:
        if group is None:
            raise ValueError("shard_offset needs group=")
        nx = ny = ntv                                              # the plan is the global record's
    plan = _Plan(tvec, nx, ny, tbounds, Navr, windowoverlap, windowfunction, kwargs)
    nwins, nfft, K = plan.nwins, plan.nfft, plan.Navr
    if verbose:
        print("using the sm_100a kernels for spectra/coherence calculations")
    win, winparams = windows(plan.windowfunction, nwins=nwins, verbose=verbose, msgout=True,
                             **({"beta": kwargs["beta"]} if "beta" in kwargs else {}))

    fftinfo = fftinfosc()
    fftinfo.win = win
    fftinfo.winparams = winparams
    fftinfo.windowoverlap = plan.windowoverlap
    fftinfo.ibnds = [plan.i0, plan.i1]
    fftinfo.S1, fftinfo.S2, fftinfo.NENBW, fftinfo.ENBW = fftanal._getNorms(win, plan.Nnyquist, plan.Fs)

    precision = _eng.resolve_precision(precision, sigx, sigy)
    dev = _eng.require_cuda(device)
    hop = nwins - plan.noverlap
    if hop <= 0:
        raise ValueError("windowoverlap leaves no hop between segments")
    if (K - 1) * hop + nwins > plan.nana:
        # the reference dies here with a broadcasting ValueError (fft_analysis.py:375)
        raise ValueError("operands could not be broadcast together: segment %d runs past the "
                         "analysed record (%d samples, nwins=%d, hop=%d)" % (K - 1, plan.nana, nwins, hop))
    refl = nwins - 1 if plan.reflect else 0
    rank, world = 0, 1
    if group is not None:
        rank, world = torch.distributed.get_rank(group), torch.distributed.get_world_size(group)
    lp = _dist.local_plan(K, hop, nwins, plan.nana, plan.i0, refl, rank, world)
    nseg = lp["k1"] - lp["k0"]
    mode = _lib.DETREND_MEAN if detrend_style > 0 else (_lib.DETREND_LINEAR if detrend_style < 0
                                                        else _lib.DETREND_NONE)
    scale = 1.0 / (fftinfo.S1 ** 2) / fftinfo.ENBW                 # :432-440

    def run(tc_choice):
        """One pass of the device pipeline -> (wp, res, nch, use_tc, host copies of the results)."""
        with torch.cuda.device(dev):
            # host records: sharded calls upload only the source slice this rank analyses; unsharded
            # pinned records cross PCIe in pieces while the segments behind each piece are accumulated
            staged, src_base = None, 0
            host_in = not (isinstance(sigx, torch.Tensor) and sigx.is_cuda)
            if shard_offset is not None:
                g_lo, g_hi = lp["src"]
                if shard_offset > g_lo or shard_offset + nloc < g_hi:
                    raise ValueError("this rank analyses source samples [%d, %d) but holds [%d, %d)"
                                     % (g_lo, g_hi, shard_offset, shard_offset + nloc))
                x = _eng.to_device(sigx, precision, dev)
                y = _eng.channels_first(_eng.to_device(sigy, precision, dev), nloc)
                src_base = int(shard_offset)
            elif group is not None and world > 1 and host_in:
                g_lo, g_hi = lp["src"]
                g_hi = max(g_hi, g_lo + 1)
                x = _eng.to_device(sigx[g_lo:g_hi], precision, dev)
                time_major = len(sigy.shape) == 1 or sigy.shape[-1] != ntv
                ysl = sigy[g_lo:g_hi] if time_major else sigy[..., g_lo:g_hi]
                y = _eng.channels_first(_eng.to_device(ysl, precision, dev), g_hi - g_lo)
                src_base = g_lo
            else:
                if early:
                    staged = early.pop()                        # started at the top of the call
                else:
                    staged = _eng.staged_upload(sigx, sigy, precision, dev) if group is None else None
                if staged is not None:
                    x, y = staged[0], staged[1]
                else:
                    x = _eng.to_device(sigx, precision, dev)
                    y = _eng.to_device(sigy, precision, dev)
                y = _eng.channels_first(y, ntv)
            nch = y.shape[0]
            if x.dim() != 1:
                x = x.reshape(-1)
            x = x.contiguous()
            rmap = _eng.record_map(plan.nana, plan.i0, refl, src_base, x.shape[0])
            wp = _welch_plan(nfft, hop, nch, precision, win, dev, group)
            fix_nana = None
            fused = mode == _lib.DETREND_MEAN and wp.can_fuse_mean and nseg > 0 and hop <= nwins and not two_pass
            if staged is not None and not fused:
                # only the fused single pass consumes the records piece by piece: wait for all of them
                torch.cuda.current_stream(dev).wait_event(staged[2][-1])
                staged = None
            use_tc = False
            offs = None
            if fused:
                # global mean (:356-357) fused into the single pass over the data: provisional mean
                # from a sample of the shard, exact correction by linearity afterwards
                s_lo, s_hi = lp["stats_span"]
                c_lo, c_hi = lp["k0"] * hop, lp["k1"] * hop          # what the fused kernel sums
                # tensor-core DFT (fp16 operands, fp32 accumulation): its per-segment rounding noise
                # (2^-12 relative, zero mean) averages out as 1/sqrt(K) -- used when enough segments
                # are averaged for the 1e-5-of-peak contract to hold with margin, and checked afterwards
                # against the dynamic range of the result (TC_DYNAMIC_RANGE)
                # (K, not this rank's share of it: the roundings average over all segments of the record, and every
                # rank of a sharded call must take the same route -- the routes end in different exchanges)
                use_tc = wp.can_use_tc and (K >= 16384 if tc_choice is None else bool(tc_choice))
                accumulate = wp.accumulate_tc if use_tc else wp.accumulate_mean
                estimate = wp.prepare_tc if use_tc else wp.estimate_mean
                res = None
                if staged is None and use_tc and K >= world and os.environ.get("PYFFT_UNFUSED_STEP") != "1":
                    # resident records: the whole step in three launches (statistics + head + tail sums | main
                    # kernel | partial sums + NVLink exchange + finalize); None = the group has no peer memory
                    res = wp.step_tc(x, y, rmap, lp["k0"], nseg, ((s_lo, c_lo), (c_hi, s_hi)), K, plan.nana, scale,
                                     bool(onesided), group if world > 1 else None)
                if res is not None:
                    pass
                elif staged is None:
                    estimate(x, y, rmap, c_lo, c_hi)
                    accumulate(x, y, rmap, lp["k0"], nseg)
                else:
                    # the records are still arriving in pieces (engine.staged_upload): the segments that are
                    # complete behind each piece are accumulated while the next piece crosses PCIe.  The
                    # provisional statistics read analysed samples [c_lo, e_hi): wait for the piece that
                    # holds the last source sample they map to (reflect-padded head: sources 1..refl).
                    cur = torch.cuda.current_stream(dev)
                    events, bounds = staged[2], staged[3]
                    e_hi = min(c_hi, c_lo + (1 << 20))
                    need_src = max(e_hi + plan.i0 - refl, min(refl, e_hi) + 1)
                    first = next((i for i, b_i in enumerate(bounds) if b_i >= need_src), len(bounds) - 1)
                    cur.wait_event(events[first])
                    estimate(x, y, rmap, c_lo, e_hi)
                    kdone = lp["k0"]
                    for i, (ev, b_i) in enumerate(zip(events, bounds)):
                        if i < first:
                            continue
                        cur.wait_event(ev)
                        kend = lp["k1"] if i == len(events) - 1 else \
                            max(kdone, min(lp["k1"], (b_i - nwins + refl - plan.i0) // hop + 1))
                        if kend > kdone:
                            accumulate(x, y, rmap, kdone, kend - kdone)
                            kdone = kend
                if res is None:
                    if s_lo < c_lo:
                        wp.detrend_stats(x, y, rmap, s_lo, c_lo)
                    if s_hi > c_hi:
                        wp.detrend_stats(x, y, rmap, c_hi, s_hi)
                fix_nana = plan.nana                        # mean correction folded into finalize
                offs = wp.mean_offset(plan.nana) if (group is None and precision == "fp32") else None
                if group is not None and res is None:
                    wp.rereference()                        # common reference, then ONE collective
                    wp.all_reduce_state(group)
                    wp.K = K
            else:
                if mode != _lib.DETREND_NONE:
                    # global detrend over the whole analysed record (:356-357): each rank sums its slice
                    wp.detrend_stats(x, y, rmap, lp["stats_span"][0], lp["stats_span"][1])
                    wp.solve_detrend(mode, plan.nana, group)
                wp.accumulate(x, y, rmap, lp["k0"], nseg)
            if group is not None and fix_nana is None:
                wp.all_reduce(group)
            fused_step = fused and res is not None
            if not fused_step:
                res = wp.finalize(scale, onesided=bool(onesided), K=K, correct_nana=fix_nana)
            corr = {}
            for nm, key, halve in (("Rxx", "Pxx", True), ("Ryy", "Pyy", True), ("Rxy", "Pxy", True),
                                   ("iCxy", "Cxy", False)):
                corr[nm] = wp.correlation(res[key], onesided=bool(onesided), halve_interior=halve)
            extras = []
            if offs is not None:
                extras.append(offs)
            if use_tc:
                extras += [wp._tc_status, wp.dynamic_range(res, onesided=bool(onesided)), wp.tc_repetition()]
            peer = wp.peer_status() if (group is not None and fused and not fused_step) else None
            if peer is not None:
                extras.append(peer)
            hosted = _host_many([res["Pxx"], res["Pyy"], res["Pxy"], res["Cxy"], res["Cxy2"], res["phi_xy"],
                                 corr["Rxx"], corr["Ryy"], corr["Rxy"], corr["iCxy"]] + extras)
            if peer is not None and hosted[-1].reshape(-1)[0] != 0:
                peer.zero_()
                raise RuntimeError("pyfft_b200: a peer GPU did not reach the NVLink reduction in time; "
                                   "the partial sums of this call are invalid (PYFFT_PEER_REDUCE=0 selects NCCL)")
            return wp, res, nch, use_tc, hosted, offs is not None

    def leak_ratio(hosted, wp):
        """fp32 fused-mean route: the kernel saw the record with a DC offset d = mean - provisional mean,
        which leaks into the lowest bins as d^2 |W[f]|^2 and is removed afterwards by linearity; fp32
        accumulation loses the signal under that leak when it is orders of magnitude larger.  -> max over
        channels and low bins of leak power / bin power."""
        d = hosted[10]
        nl = min(wp.nlin, hosted[0].shape[0])
        w2 = (wp.wlow_host[:nl, 0] ** 2 + wp.wlow_host[:nl, 1] ** 2) * scale
        if onesided:
            w2 = w2 * np.where((np.arange(nl) > 0), 2.0, 1.0)
        lo = 0 if onesided else nfft // 2                      # bin 0 after the fftshift of two-sided spectra
        px = np.abs(hosted[0][lo:lo + nl]).reshape(nl, 1)
        py = np.abs(hosted[1][lo:lo + nl]).reshape(nl, -1)
        tiny = 1e-300
        lx, ly = d[0] ** 2 * w2[:, None], (d[1:] ** 2)[None, :] * w2[:, None]          # leak power per low bin
        pkx = max(float(np.max(np.abs(hosted[0]))), tiny)
        pky = np.maximum(np.max(np.abs(hosted[1]).reshape(hosted[0].shape[0], -1), axis=0), tiny)
        # bins that are (legitimately) empty -- DC after the mean removal of a periodic record -- count with a
        # floor of 1e-4 of the channel's peak: their coherence is undefined, their PSD error is relative to the peak
        per_bin = max(np.max(lx / np.maximum(px, 1e-4 * pkx)), np.max(ly / np.maximum(py, 1e-4 * pky[None, :])))
        per_peak = max(np.max(lx) / pkx, np.max(ly / pky[None, :]))
        return float(per_bin), float(per_peak)

    generic = is_cplx or not _eng.pow2_window(nfft)
    if generic:
        # window lengths that are not powers of two (the reference's default Navr = 8 gives them) and
        # complex-valued records: per-segment spectra by Bluestein (pyfft_stft_any), products on the device
        wp, res, use_tc, has_offs = None, None, False, False
        gx, gy = sigx, sigy
        if early and early[0] is not None:                       # already on its way to the device
            st = early.pop()
            torch.cuda.current_stream(dev).wait_event(st[2][-1])
            gx, gy = st[0], st[1]
        nch, hosted = _pwelch_generic(gx, gy, ntv, plan, win, mode, precision, dev, bool(onesided), scale,
                                      lp, group, is_cplx)
    else:
        wp, res, nch, use_tc, hosted, has_offs = run(tensor_cores)
    fftinfo.tensor_cores = bool(use_tc)
    fftinfo.tc_fallback = None
    fftinfo.mean_fallback = None
    base = 11 if has_offs else 10
    if use_tc:
        flag, rng, rep = int(hosted[base].reshape(-1)[0]), hosted[base + 1], hosted[base + 2]
        fftinfo.tc_checks = dict(status=flag, min_over_mean=float(np.min(rng[:, 0] / np.maximum(rng[:, 1], 1e-300))),
                                 repetition=float(np.min(rep[0])), stationarity=float(np.min(rep[1])))
        why = None
        if flag & 4:
            wp._tc_status.zero_()
            raise RuntimeError("pyfft_b200: a peer GPU did not reach the NVLink exchange of the fused step in time; "
                               "the results of this call are invalid (PYFFT_PEER_REDUCE=0 selects NCCL)")
        if flag & 1:
            why = "a tensor-core pipeline barrier timed out"
        elif flag & 2:
            why = "an fp16 operand overflowed (the record is far louder somewhere than its sampled statistics)"
        elif np.min(rep[0]) < TC_MIN_REPETITION:
            why = ("the record repeats after a few hops (repetition measure %.1e): the fp16 roundings of the "
                   "tensor-core path would not average out over the segments" % float(np.min(rep[0])))
        elif np.min(rep[1]) < TC_MIN_STATIONARITY:
            why = ("the power of the record sits in a fraction %.2g of it: too few effective segments for the "
                   "fp16 roundings of the tensor-core path to average out" % float(np.min(rep[1])))
        elif np.any(rng[:, 0] < max(TC_DYNAMIC_RANGE / K, TC_MIN_RATIO) * rng[:, 1]):
            why = ("the spectrum has bins %.1e below its mean bin power: too close to the fp16 rounding floor "
                   "of the tensor-core path for the coherence contract"
                   % float(np.min(rng[:, 0] / np.maximum(rng[:, 1], 1e-300))))
        if flag:
            wp._tc_status.zero_()
        if group is not None and world > 1 and tensor_cores is None:
            # sharded: the measures are per shard, the rerun is a collective decision
            vote = torch.tensor([0 if why is None else 1], dtype=torch.int32, device=dev)
            torch.distributed.all_reduce(vote, op=torch.distributed.ReduceOp.MAX, group=group)
            if int(vote.item()) and why is None:
                why = "another rank's shard failed the checks of the tensor-core path"
        if why is not None:
            if tensor_cores is None:
                # data-aware selection: rerun on the fp32 CUDA-core kernel (no CPU fallback)
                wp, res, nch, use_tc, hosted, has_offs = run(False)
                fftinfo.tensor_cores = False
                fftinfo.tc_fallback = why
            elif flag:
                raise RuntimeError("pyfft_b200: tensor_cores=True was forced and the kernel failed: " + why)
            else:
                fftinfo.tc_fallback = "forced: " + why
    if has_offs and not two_pass:
        r, rpk = leak_ratio(hosted, wp)
        if not (r <= FUSED_MEAN_MAX_LEAK and rpk <= FUSED_MEAN_MAX_LEAK_PEAK):
            # the provisional mean was too far off for fp32 accumulation: exact mean first (K0), then K1
            two_pass = True
            wp, res, nch, use_tc, hosted, has_offs = run(False)
            fftinfo.tensor_cores = False
            fftinfo.mean_fallback = ("provisional mean off by a leak ratio of %.1e: redone with the two-pass "
                                     "detrend" % r)
    (Pxx, Pyy, Pxy, Cxy, Cxy2, phi_xy, h_rxx, h_ryy, h_rxy, h_icxy) = hosted[:10]

    freq = np.fft.fftfreq(nfft, 1.0 / plan.Fs)                     # :396
    freq = freq[:plan.Nnyquist] if onesided else np.fft.fftshift(freq)

    # ---- O(nfreq) bookkeeping on the averaged spectra, as in the reference (:496-540): at first access ---- #
    fftinfo.Rxx, fftinfo.Ryy = h_rxx, h_ryy
    fftinfo.Rxy, fftinfo.iCxy = h_rxy, h_icxy
    enbw, nnyq, fs_ = fftinfo.ENBW, plan.Nnyquist, plan.Fs

    def derived(Pxx=Pxx, Pyy=Pyy, Pxy=Pxy, Cxy=Cxy, Cxy2=Cxy2, h_rxx=h_rxx, h_ryy=h_ryy, h_rxy=h_rxy):
        d = {}
        with np.errstate(all="ignore"):
            d["varCxy"] = ((1.0 - Cxy * np.conjugate(Cxy)) / np.sqrt(2 * K)) ** 2.0
            d["varCxy2"] = 4.0 * Cxy2 * d["varCxy"]
            d["varPxx"] = (Pxx / np.sqrt(K)) ** 2.0
            d["varPyy"] = (Pyy / np.sqrt(K)) ** 2.0
            d["varPxy"] = (Pxy / np.sqrt(K)) ** 2.0
            d["varPhxy"] = (np.sqrt(1.0 - np.abs(Cxy2))) / np.sqrt(2 * K * np.sqrt(np.abs(Cxy2))) ** 2.0
            for nm, P in (("Lxx", Pxx), ("Lyy", Pyy), ("Lxy", Pxy)):
                amp = np.sqrt(np.abs(enbw * P))
                if onesided:
                    amp[1:-1] = np.sqrt(2) * amp[1:-1]
                    if nfft % 2:
                        amp[-1] = np.sqrt(2) * amp[-1]
                d[nm] = amp
            # zero-lag values sit at index nfft//2 after the fftshift (:585-596)
            d["Ex"] = h_rxx[nfft // 2, ...].copy()
            d["Ey"] = h_ryy[nfft // 2, ...].copy()
            d["corrcoef"] = h_rxy / np.sqrt(np.ones((nfft, 1), dtype=h_rxy.dtype) * (d["Ex"] * d["Ey"]))
            d["lags"] = (np.asarray(range(1, nfft + 1), dtype=int) - nnyq) / fs_
            d["varLxx"] = (d["Lxx"] ** 2) * (d["varPxx"] / np.abs(Pxx) ** 2)
            d["varLyy"] = (d["Lyy"] ** 2) * (d["varPyy"] / np.abs(Pyy) ** 2)
            d["varLxy"] = (d["Lxy"] ** 2) * (d["varPxy"] / np.abs(Pxy) ** 2)
        if nch == 1:                                               # :605-631
            for nm in ("lags", "corrcoef", "Lxx", "Lyy", "Lxy", "varLxx", "varLyy", "varLxy", "varCxy", "varCxy2",
                       "varPxx", "varPyy", "varPxy", "varPhxy"):
                d[nm] = d[nm].flatten()
        return d

    fftinfo._defer_derived(derived)

    if nch == 1:                                                   # :605-631
        Pyy, Pxy, Cxy, Cxy2, phi_xy = (a.flatten() for a in (Pyy, Pxy, Cxy, Cxy2, phi_xy))
        for nm in ("Rxx", "Ryy", "Rxy", "iCxy"):
            setattr(fftinfo, nm, getattr(fftinfo, nm).flatten())

    fftinfo.nch = nch
    fftinfo.Fs = plan.Fs
    fftinfo.Navr = K
    fftinfo.nwins = nwins
    fftinfo.noverlap = plan.noverlap
    fftinfo.overlap = plan.windowoverlap
    fftinfo.window = plan.windowfunction
    fftinfo.minFreq = 2.0 * plan.Fs / nwins
    fftinfo.freq = freq.copy()
    fftinfo.Pxx, fftinfo.Pyy, fftinfo.Pxy = Pxx.copy(), Pyy.copy(), Pxy.copy()
    fftinfo.Cxy, fftinfo.Cxy2, fftinfo.phi_xy = Cxy.copy(), Cxy2.copy(), phi_xy.copy()
    fftinfo.precision = precision
    fftinfo._segment_loader = _make_segment_loader(sigx, sigy, ntv, plan, win, mode, precision, dev,
                                                   bool(onesided), scale)
    if return_device:
        if res is None:
            res = dict((k, torch.from_numpy(np.ascontiguousarray(v)).to(dev)) for k, v in
                       (("Pxy", Pxy), ("Pxx", Pxx), ("Pyy", Pyy), ("Cxy", Cxy), ("phi_xy", phi_xy)))
        return (torch.from_numpy(freq).to(dev), res["Pxy"], res["Pxx"], res["Pyy"], res["Cxy"],
                res["phi_xy"], fftinfo)
    return freq, Pxy, Pxx, Pyy, Cxy, phi_xy, fftinfo


def varcoh(Pxy, varPxy, Pxx, varPxx, Pyy, varPyy, meansquared=True):
    """Coherence with propagated variance (fft_analysis.py:1218-1262), O(nfreq) host arithmetic."""
    ms, mc = np.imag(Pxy), np.real(Pxy)
    vs, vc = np.imag(varPxy), np.real(varPxy)
    with np.errstate(all="ignore"):
        prop = (vc * (2 * mc / (mc ** 2 + ms ** 2)) ** 2 + vs * (2 * ms / (mc ** 2 + ms ** 2)) ** 2
                + varPxx * (1 / Pxx) ** 2 + varPyy * (1 / Pyy) ** 2)
        if meansquared:
            Coh = np.abs(Pxy * np.conj(Pxy)) / (np.abs(Pxx) * np.abs(Pyy))
            return Coh, Coh ** 2 * prop
        Coh = Pxy / np.sqrt(np.abs(Pxx) * np.abs(Pyy))
        varCoh = 0.25 * (Coh ** 2 * prop) / Coh
        return np.sqrt(Coh), varCoh


def varphi(Pxy_real, Pxy_imag, varPxy_real, varPxy_imag, angle_range=np.pi):
    """Phase with propagated variance (fft_analysis.py:1300-1330)."""
    with np.errstate(all="ignore"):
        if angle_range > 0.5 * np.pi:
            ph = np.arctan2(Pxy_imag, Pxy_real)
        else:
            ph = np.arctan(Pxy_imag / Pxy_real)
        tangent = Pxy_imag / Pxy_real
        vartang = (varPxy_imag + varPxy_real * tangent ** 2) / (Pxy_real ** 2)
        return ph, vartang / (1 + tangent ** 2) ** 2


def mean_angle(phi, vphi=None, dim=0, angle_range=0.5 * np.pi, vsyst=None, precision=None, device=None):
    """Circular mean of phase angles along `dim` with propagated variance (fft_analysis.py:1334-1377):
    the reduction over the segments (nanmean / nanvar of cos and sin, the variance sums) runs on the
    device (pyfft_mean_angle); device tensors are accepted and reduced in place."""
    import ctypes as C
    dev = _eng.require_cuda(device)
    prec = _eng.resolve_precision(precision, phi)
    with torch.cuda.device(dev):
        def dev2(a):
            if a is None:
                return None
            t = _eng.to_device(a, prec, dev)
            t = t.movedim(dim, 0) if t.dim() > 1 else t.reshape(-1, 1)
            return t.reshape(t.shape[0], -1).contiguous()
        p = dev2(phi)
        v, sy = dev2(vphi), dev2(vsyst)
        K, ncol = p.shape
        out = torch.empty((ncol, 4), dtype=torch.float64, device=dev)
        _lib.check(_lib.lib.pyfft_mean_angle(_eng._CODE[prec], C.c_void_p(p.data_ptr()),
                                             C.c_void_p(v.data_ptr()) if v is not None else None,
                                             C.c_void_p(sy.data_ptr()) if sy is not None else None, K, ncol,
                                             C.c_void_p(out.data_ptr()),
                                             C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        o = out.cpu().numpy()
    shp = list(phi.shape)
    if len(shp) > 0:
        del shp[dim % len(shp)]
    shp = tuple(shp)
    mca, msa, vca, vsa = (o[:, i].reshape(shp) for i in range(4))
    return varphi(Pxy_real=mca, Pxy_imag=msa, varPxy_real=vca, varPxy_imag=vsa, angle_range=angle_range)


def _cols(a, nfreq):
    """(nfreq,) or (nfreq, nch) -> (nfreq, nch) complex128 (pybaseutils.reshapech: frequency along axis 0)."""
    a = np.asarray(a)
    if a.ndim == 1:
        a = a.reshape(-1, 1)
    elif a.shape[0] != nfreq and a.shape[1] == nfreq:
        a = a.T
    return np.ascontiguousarray(a.astype(np.complex128))


def _trapz_weights(f):
    w = np.zeros(len(f))
    if len(f) > 1:
        d = np.diff(f)
        w[:-1] += 0.5 * d
        w[1:] += 0.5 * d
    return w


def integratespectra(freq, Pxy, Pxx, Pyy, frange, varPxy=None, varPxx=None, varPyy=None, device=None):
    """Band integrals of averaged spectra with propagated variances (fft_analysis.py:835-938): same
    arguments and the same tuple ``(Pxy_i, Pxx_i, Pyy_i, Cxy_i, ph_i, info)``.  ``frange`` may also be an
    (nbands, 2) array: all bands and channels are integrated by ONE kernel (pyfft_band_integrate) and the
    outputs gain a leading band axis.  Variances: the reference calls pybaseutils.trapz_var (un-vendored:
    PARITY UNPINNED) -- restated as sum (w_i^2 var_i) with the trapezoid weights; missing variances count
    as zero (the reference dies on ``np.size_like``)."""
    import ctypes as C
    dev = _eng.require_cuda(device)
    freq = np.ascontiguousarray(np.asarray(freq, dtype=np.float64))
    F = len(freq)
    pxy, pxx, pyy = _cols(Pxy, F), _cols(Pxx, F), _cols(Pyy, F)
    nch = pxy.shape[1]
    fr = np.atleast_2d(np.asarray(frange, dtype=np.float64))
    single = np.ndim(frange) == 1
    with torch.cuda.device(dev):
        td = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        d_f, d_xy, d_xx, d_yy, d_b = td(freq), td(pxy), td(pxx), td(pyy), td(fr)
        out = torch.empty((fr.shape[0], nch, 8), dtype=torch.float64, device=dev)
        p = lambda t: C.c_void_p(t.data_ptr())
        _lib.check(_lib.lib.pyfft_band_integrate(p(d_f), F, p(d_xy), p(d_xx), pxx.shape[1], p(d_yy), nch, p(d_b),
                                                 fr.shape[0], p(out), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        o = out.cpu().numpy()
    Pxy_i = o[..., 0] + 1j * o[..., 1]
    Pxx_i = o[..., 2] + 1j * o[..., 3]
    Pyy_i = o[..., 4] + 1j * o[..., 5]
    var = {}
    for nm, v, ncols in (("xy", varPxy, nch), ("xx", varPxx, pxx.shape[1]), ("yy", varPyy, nch)):
        vv = np.zeros((F, ncols), dtype=np.complex128) if v is None else _cols(v, F)
        acc = np.zeros((fr.shape[0], ncols), dtype=np.complex128)
        for b in range(fr.shape[0]):
            inds = np.where((freq >= fr[b, 0]) * (freq <= fr[b, 1]))[0]
            w2 = _trapz_weights(freq[inds]) ** 2
            acc[b] = (w2[:, None] * vv[inds].real).sum(0) + 1j * (w2[:, None] * vv[inds].imag).sum(0)
        var[nm] = acc
    if pxx.shape[1] == 1 and nch > 1:
        Pxx_i = Pxx_i[:, :1] * np.ones((1, nch))
        var["xx"] = var["xx"][:, :1] * np.ones((1, nch))
    Cxy_i, varCxy_i = varcoh(Pxy_i, var["xy"], Pxx_i, var["xx"], Pyy_i, var["yy"], 0)
    ph_i, varph_i = varphi(Pxy_i.real, Pxy_i.imag, var["xy"].real, var["xy"].imag, np.pi)
    info = Struct()
    pick = (lambda a: a[0]) if single else (lambda a: a)
    info.frange = np.asarray(frange, dtype=np.float64)
    info.ifrange = pick([np.where((freq >= fr[b, 0]) * (freq <= fr[b, 1]))[0] for b in range(fr.shape[0])])
    for nm, a in (("Pxy_i", Pxy_i), ("varPxy_i", var["xy"]), ("Pxx_i", Pxx_i), ("varPxx_i", var["xx"]),
                  ("Pyy_i", Pyy_i), ("varPyy_i", var["yy"]), ("ph_i", ph_i), ("varph_i", varph_i),
                  ("Cxy_i", Cxy_i), ("varCxy_i", varCxy_i), ("fweighted", o[..., 6])):
        setattr(info, nm, pick(a))
    info.angle_range, info.meansquared = np.pi, 0
    return pick(Pxy_i), pick(Pxx_i), pick(Pyy_i), pick(Cxy_i), pick(ph_i), info


def getNpeaks(Npeaks, tvec, sigx, sigy, **kwargs):
    """The Npeaks largest peaks of the cross amplitude spectrum, each as [amplitude of sigy, frequency,
    cross phase] (fft_analysis.py:940-1027): fft_pwelch on the GPU, then the reference's selection loop --
    argmax of Lxy, removal of an equivalent noise bandwidth (or ``minsep``) around it in the compressed
    arrays, with its literal edge rules -- as one kernel (pyfft_find_peaks), batched over the channels of
    sigy (2-D sigy: a tuple per channel is returned; the reference only handles one)."""
    import ctypes as C
    for k in ("tbounds", "Navr", "windowoverlap", "windowfunction", "useMLAB", "plotit", "verbose", "detrend_style"):
        kwargs.setdefault(k, None)
    kwargs.setdefault("onesided", True)
    fmin = kwargs.pop("fmin", None)
    fmax = kwargs.pop("fmax", None)
    minsep = kwargs.pop("minsep", 6)
    freq, Pxy, Pxx, Pyy, Cxy, phi_xy, fftinfo = fft_pwelch(tvec, sigx, sigy, **kwargs)
    nfreq = len(freq)
    ENBW = max((fftinfo.ENBW, minsep))                              # :986-987
    irem = int(2 * nfreq * ENBW / (freq[-1] - freq[0]))             # :989
    fmin = 0.0 if fmin is None else fmin                            # :992-994: the mask below never fires
    fmax = freq[-1] if fmax is None else fmax                       # for fmin < fmax (reference as written)
    keep = np.ones(nfreq, dtype=bool)
    keep[(freq <= fmin) * (freq >= fmax)] = False
    Lxy = np.asarray(fftinfo.Lxy).reshape(nfreq, -1)[keep]
    Lyy = np.asarray(fftinfo.Lyy).reshape(nfreq, -1)[keep]
    ph = np.asarray(phi_xy).reshape(nfreq, -1)[keep]
    fr = freq[keep]
    dev = _eng.require_cuda(kwargs.get("device"))
    with torch.cuda.device(dev):
        L = torch.from_numpy(np.ascontiguousarray(Lxy, dtype=np.float64)).to(dev)
        idx = torch.empty((L.shape[1], int(Npeaks)), dtype=torch.int32, device=dev)
        _lib.check(_lib.lib.pyfft_find_peaks(C.c_void_p(L.data_ptr()), L.shape[0], L.shape[1], int(Npeaks), irem,
                                             C.c_void_p(idx.data_ptr()),
                                             C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        idx = idx.cpu().numpy()
    if (idx < 0).any():
        raise ValueError("attempt to get argmax of an empty sequence")   # numpy in the reference's loop
    outs = []
    for c in range(idx.shape[0]):
        outs.append(tuple([np.copy(Lyy[i, c]), np.copy(fr[i]), np.copy(ph[i, c])] for i in idx[c]))
    return outs[0] if len(outs) == 1 else tuple(outs)


def fft_deriv(sig, xx=None, lowpass=True, Fs_new=None, modified=True, detrend=detrend_none, window=None,
              precision=None, device=None):
    """Spectral derivative of a profile on a uniform grid (fft_analysis.py:1453-1588): same
    arguments and return values ``(dsdx, xx)``.  The grid rescaling, the ``detrend`` and ``window``
    callables and the end-point fix-up run on the host exactly as in the reference
    (:1513-1519, :1543-1563); its numpy.fft line ``real(ifft(wavenumber * fft(sig)))`` (:1553) is
    one call into the long-transform kernels (pyfft_spectral_derivative).  The reference's
    pre-filter (:1493-1511, filters.downsample_efficient) only acts when ``Fs_new < Fs`` -- never
    with the defaults -- and is refused (NotImplementedError); the length must be a power of
    two >= 16 (no CPU fallback)."""
    dev = _eng.require_cuda(device)
    sig = np.asarray(sig)
    precision = _eng.resolve_precision(precision, sig)
    sig = sig.astype(np.float64)
    if xx is None:
        xx = 1.0 * np.asarray(range(0, len(sig)))                             # :1484-1487
    xx = np.asarray(xx, dtype=np.float64)
    if lowpass:                                                               # :1493-1511
        dxo = xx[1] - xx[0]
        lp = 0.5 * 1.0 / dxo if lowpass is True else lowpass
        Fs = 1.0 / dxo
        fsn = min(5.0 * lp, Fs) if Fs_new is None else Fs_new
        if fsn < Fs:
            raise NotImplementedError("fft_deriv: down-sampling before the derivative (Fs_new < Fs) is "
                                      "filters.py territory, outside the GPU path")
    # rescale (:1419-1437)
    slope = np.nanmax(sig) - np.nanmin(sig)
    offset = np.nanmin(sig)
    if slope == 0:
        slope = 1.0
    s = (sig.copy() - offset) / slope
    xslope = np.nanmax(xx) - np.nanmin(xx)
    xoffset = -1e-4
    if xslope == 0:
        xslope = 1.0
    x = (xx.copy() - xoffset) / xslope
    s = detrend(s)                                                            # :1519
    N = len(x)
    if N < 16 or N & (N - 1):
        raise NotImplementedError("fft_deriv: %d points -- the GPU path needs a power of two >= 16" % N)
    dx = x[1] - x[0]
    win = np.ones_like(s) if window is None else window(N)                    # :1543-1547
    s = win * s
    ds0 = (s[1] - s[0]) / (x[1] - x[0])                                       # :1551-1552
    ds1 = (s[-1] - s[-2]) / (x[-1] - x[-2])
    with torch.cuda.device(dev):
        u = _eng.to_device(s[None, :], precision, dev)
        d = _eng.spectral_derivative(u, dx, precision, modified=bool(modified))[0]
        d = d.cpu().numpy().astype(np.float64)
    d /= win                                                                  # :1556
    d[0] = ds0                                                                # :1559-1560
    d[-1] = ds1
    return d * slope / xslope, x * xslope + xoffset                           # unscale, :1439-1451


def cross_spectral_matrix(tvec, sig, tbounds=None, Navr=None, windowoverlap=None, windowfunction=None,
                          detrend_style=None, **kwargs):
    """All-pairs cross-power spectral density of a multi-channel record in ONE pass.

    The reference builds this matrix by calling fft_pwelch once per reference channel with
    every channel as sigy (HeatPulse_Funcs.py:576-583); ``G[f, i, j]`` equals the ``Pxy[f, j]``
    of the call with ``sigx = sig[:, i]`` -- one-sided, Nyquist bin dropped, x2 on interior
    bins, scaled by 1/(S1^2 ENBW), mean over segments (fft_analysis.py:393-446).  Segmentation,
    window, reflect padding and global detrend follow fft_pwelch exactly.

    Returns ``(freq, G)`` with G complex128 of shape (Nnyquist, nch, nch).  ``group``: a
    torch.distributed process group -- segments are sharded across its ranks and the partial
    matrices summed by one all-reduce.
    """
    precision = kwargs.pop("precision", None)
    device = kwargs.pop("device", None)
    group = kwargs.pop("group", None)
    return_device = kwargs.pop("return_device", False)
    if detrend_style is None:
        detrend_style = 1
    ntv = len(tvec)
    precision = _eng.resolve_precision(precision, sig)
    dev = _eng.require_cuda(device)
    with torch.cuda.device(dev):
        x = _eng.channels_first(_eng.to_device(sig, precision, dev), ntv)
        nch, nx = x.shape
        plan = _Plan(tvec, nx, nx, tbounds, Navr, windowoverlap, windowfunction, kwargs)
        nwins, K = plan.nwins, plan.Navr
        if nwins & (nwins - 1) or nwins < 16 or nwins > 8192:
            raise NotImplementedError("nwins = %d: the GPU path needs a power-of-two window of 16..8192 "
                                      "samples" % nwins)
        win = windows(plan.windowfunction, nwins=nwins, verbose=False)
        S1, S2, _, ENBW = fftanal._getNorms(win, plan.Nnyquist, plan.Fs)
        refl = nwins - 1 if plan.reflect else 0
        rmap = _eng.record_map(plan.nana, plan.i0, refl, 0, nx)
        hop = nwins - plan.noverlap
        if (K - 1) * hop + nwins > plan.nana:
            raise ValueError("operands could not be broadcast together: segment %d runs past the "
                             "analysed record" % (K - 1))
        mode = _lib.DETREND_MEAN if detrend_style > 0 else (_lib.DETREND_LINEAR if detrend_style < 0
                                                            else _lib.DETREND_NONE)
        rank, world = 0, 1
        if group is not None:
            rank, world = torch.distributed.get_rank(group), torch.distributed.get_world_size(group)
        lp = _dist.local_plan(K, hop, nwins, plan.nana, plan.i0, refl, rank, world)
        coef = _eng.detrend_coefficients(x, rmap, mode, precision, plan.nana, group=group,
                                         j_span=lp["stats_span"])
        G = _eng.csm_accumulate(x, rmap, nwins, hop, lp["k0"], lp["k1"] - lp["k0"], win, precision, coef=coef)
        if group is not None:
            torch.distributed.all_reduce(torch.view_as_real(G), group=group)
        F = plan.Nnyquist
        wgt = torch.full((F,), 1.0 / (S1 ** 2) / ENBW / K, dtype=torch.float64, device=dev)
        wgt[1:-1] *= 2.0                                           # :414-416
        G = G[:F] * wgt[:, None, None]
    freq = np.fft.fftfreq(nwins, 1.0 / plan.Fs)[:F]
    if return_device:
        return torch.from_numpy(freq).to(dev), G
    return freq, _host(G)


# --------------------------------------------------------------------------- #
# class fftanal -- segmentation statics now, the object API is completed by
# pyfft_b200/stft.py (attached below) once K2 is available.
# --------------------------------------------------------------------------- #

class fftanal(Struct):
    """Object API of the reference (fft_analysis.py:1695-2640)."""

    def __init__(self, tvec=None, sigx=None, sigy=None, **kwargs):
        self.verbose = kwargs.get("verbose", True)
        if tvec is None or sigx is None:
            if self.verbose:
                print("Please give at least a time-vector [s] and a signal vector [a.u.]")
            return
        self.init(tvec, sigx, sigy, **kwargs)

    def init(self, tvec=None, sigx=None, sigy=None, **kwargs):
        """fft_analysis.py:1713-1783."""
        self.nosigy = sigy is None or sigx is sigy
        self.tvec, self.sigx, self.sigy = tvec, sigx, sigy
        self.tbounds = kwargs.get("tbounds", [tvec.min(), tvec.max()])
        self.useMLAB = kwargs.get("useMLAB", False)
        self.plotit = kwargs.get("plotit", False)
        self.verbose = kwargs.get("verbose", True)
        self.Navr = kwargs.get("Navr", None)
        self.window = kwargs.get("windowfunction", "Hanning")
        if self.window is None:
            self.window = "Hanning"
        self.overlap = kwargs.get("windowoverlap", windows(self.window, verbose=False))
        self.tvecy = kwargs.get("tvecy", None)
        self.onesided = kwargs.get("onesided", None)
        self.detrendstyle = kwargs.get("detrend", 1)
        self.frange = kwargs.get("frange", None)
        self.axes = kwargs.get("axes", -1)
        self.precision = kwargs.get("precision", None)
        self.device = kwargs.get("device", None)
        if self.onesided is None:
            self.onesided = not (np.iscomplexobj(sigx) or np.iscomplexobj(sigy))
        if self.tvecy is not None:
            # put the signals on the same time base (fft_analysis.py:1750-1752)
            self.tvec, self.sigx, self.sigy = self.resample(tvec, sigx, self.tvecy, sigy)
            tvec, sigx, sigy = self.tvec, self.sigx, self.sigy
        self.Fs = self.__Fs__(self.tvec)
        self.ibounds = self.__ibounds__(self.tvec, self.tbounds)
        self.nsig = np.size(self.__trimsig__(tvec, self.ibounds))
        calc_navr = False
        if self.Navr is None:
            calc_navr = True
            self.Navr = 8
        if "minFreq" in kwargs:
            kwargs["tper"] = 2.0 / kwargs["minFreq"]
        if "tper" in kwargs:
            self.tper = kwargs["tper"]
            self.nwins = int(self.Fs * self.tper)
        else:
            calc_navr = False
            self.nwins = self.getNwins()
        self.noverlap = self.getNoverlap()
        if calc_navr:
            self.Navr = self.getNavr()
        self.win, self.winparams = self.makewindowfn(self.window, self.nwins, self.verbose)
        self.getNnyquist()
        self.getNorms()

    def update(self, d=None):
        if d is not None:
            if type(d) != dict:
                d = d.dict_from_class()
            super(fftanal, self).__init__(d)

    def fftpwelch(self):
        """fft_analysis.py:1794-1803."""
        self.freq, self.Pxy, self.Pxx, self.Pyy, self.Cxy, self.phi_xy, self.fftinfo = \
            fft_pwelch(self.tvec, self.sigx, self.sigy, self.tbounds, Navr=self.Navr,
                       windowoverlap=self.overlap, windowfunction=self.window,
                       useMLAB=self.useMLAB, plotit=self.plotit, verbose=self.verbose,
                       detrend_style=self.detrendstyle, onesided=self.onesided,
                       precision=self.precision, device=self.device)
        self.update(self.fftinfo)

    # ---- segmentation / normalisation helpers (verbatim formulas) ---------- #
    def getNavr(self):
        self.Navr = fftanal._getNavr(self.nsig, self.nwins, self.noverlap)
        return self.Navr

    def getNwins(self):
        self.nwins = fftanal._getNwins(self.nsig, self.Navr, self.overlap)
        return self.nwins

    def getNoverlap(self):
        self.noverlap = fftanal._getNoverlap(self.nwins, self.overlap)
        return self.noverlap

    def getNnyquist(self):
        self.Nnyquist = self._getNnyquist(self.nwins)
        return self.Nnyquist

    def getNorms(self):
        self.S1, self.S2, self.NENBW, self.ENBW = fftanal._getNorms(self.win, self.Nnyquist, self.Fs)

    @staticmethod
    def resample(tvx, sigx, tvy, sigy):
        """fft_analysis.py:2363-2373: the shorter record is up-sampled (filters.upsample, on the device) to the
        rate of the longer one.  Like the reference, tvec is the longer record's time vector while the up-sampled
        record has numpy.arange(t[0], t[-1], 1/Fs) points."""
        from .filters import upsample
        Fsx = fftanal.__Fs__(tvx)
        Fsy = fftanal.__Fs__(tvy)
        tvec = tvx
        if len(sigx) > len(sigy):
            sigy = upsample(sigy, Fsy, Fsx)
            tvec = tvx
        elif len(sigy) > len(sigx):
            sigx = upsample(sigx, Fsx, Fsy)
            tvec = tvy
        return tvec, sigx, sigy

    @staticmethod
    def __Fs__(tvec):
        return (len(tvec) - 1) / (tvec[-1] - tvec[0])              # :2375-2377

    @staticmethod
    def __ibounds__(tvec, tbounds):
        fs = fftanal.__Fs__(tvec)
        return [int(np.floor((tbounds[0] - tvec[0]) * fs)),
                int(np.floor(1 + (tbounds[1] - tvec[0]) * fs))]    # :2379-2383

    @staticmethod
    def __trimsig__(sigt, ibounds):
        return sigt[ibounds[0]:ibounds[1]]

    @staticmethod
    def makewindowfn(windowfunction, nwins, verbose=True):
        return windows(windowfunction, nwins=nwins, verbose=verbose, msgout=True)

    @staticmethod
    def _getNwins(nsig, Navr, windowoverlap):
        nwins = int(np.floor(nsig * 1.0 / (Navr - Navr * windowoverlap + windowoverlap)))  # :2411-2418
        return nsig if nwins >= nsig else nwins

    @staticmethod
    def _getNoverlap(nwins, windowoverlap):
        return int(np.ceil(windowoverlap * nwins))                 # :2420-2422

    @staticmethod
    def _getNavr(nsig, nwins, noverlap):
        if nwins >= nsig:
            return int(1)
        return (nsig - noverlap) // (nwins - noverlap)            # :2424-2430

    @staticmethod
    def _getMINoverlap(nsig, nwins, Navr):
        noverlap = 1
        while fftanal._checkCOLA(nsig, nwins, noverlap) is False and noverlap < 1e4:
            noverlap += 1
        return noverlap

    @staticmethod
    def _getMAXoverlap(nsig, nwins, Navr):
        noverlap = int(nwins) - 1
        while fftanal._checkCOLA(nsig, nwins, noverlap) is False and noverlap > 0:
            noverlap -= 1
        return noverlap

    @staticmethod
    def _checkCOLA(nsig, nwins, noverlap):
        """Constant-overlap-add test (fft_analysis.py:2466-2468)."""
        return (nsig - nwins) % (nwins - noverlap) == 0

    @staticmethod
    def _getNnyquist(nfft):
        return (nfft + 1) // 2 if nfft % 2 else nfft // 2         # :2470-2484

    @staticmethod
    def _getS1(win):
        return np.sum(win)

    @staticmethod
    def _getS2(win):
        return np.sum(win ** 2.0)

    @staticmethod
    def _getNENBW(Nnyquist, S1, S2):
        return Nnyquist * 1.0 * S2 / (S1 ** 2)

    @staticmethod
    def _getENBW(Fs, S1, S2):
        return Fs * S2 / (S1 ** 2)

    @staticmethod
    def _getNorms(win, Nnyquist, Fs):
        S1, S2 = fftanal._getS1(win), fftanal._getS2(win)
        return S1, S2, fftanal._getNENBW(Nnyquist, S1, S2), fftanal._getENBW(Fs, S1, S2)

    @staticmethod
    def intspectra(freq, sigft, ifreq=None, ispan=None, ENBW=None):
        """Trapezoid integral of a spectrum around its peak (fft_analysis.py:2514-2536); O(nfreq) host
        bookkeeping.  ``ifreq`` together with ``ispan`` leaves the bounds unset in the reference
        (UnboundLocalError) and is refused here the same way."""
        if ifreq is None:
            ifreq = np.argmax(np.abs(sigft), axis=0)
            if ENBW is not None:
                ispan = 2 * np.where(freq >= ENBW)[0][0]
            elif ispan is None:
                ispan = 6
            ilow, ihigh = ifreq - ispan // 2, ifreq + ispan // 2
        elif ispan is None:
            ilow, ihigh = 0, len(sigft)
        else:
            raise UnboundLocalError("cannot access local variable 'ilow' where it is not associated with a value")
        Isig = np.trapezoid(sigft[ilow:ihigh], freq[ilow:ihigh], axis=0)
        return Isig, np.zeros_like(Isig)

    @staticmethod
    def _detrend_func(detrend_style=None):
        if detrend_style is None:
            detrend_style = 0
        if detrend_style > 0:
            return detrend_mean
        if detrend_style < 0:
            return detrend_linear
        return detrend_none


# the STFT half of the object API lives in _fftanal_stft.py
from . import _fftanal_stft as _stft_api  # noqa: E402

for _name, _fn in _stft_api.METHODS.items():
    setattr(fftanal, _name, _fn)


fftanal.fftshift = lambda self, sig, axes=None: np.fft.fftshift(sig, axes=self.axes if axes is None else axes)
fftanal.ifftshift = lambda self, sig, axes=None: np.fft.ifftshift(sig, axes=self.axes if axes is None else axes)
