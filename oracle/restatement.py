"""TEST INFRASTRUCTURE ONLY -- numpy float64 restatement of the reference hot path.

Every function states the reference lines it follows (paths are relative to
/root/reference).  The restatement is pinned by tests/test_oracle_golden.py
against outputs of the reference itself (tests/golden/*.npz, written by
oracle/gen_golden.py in the build container).

Third-party arithmetic that is NOT in /root/reference:
  * numpy.fft (pocketfft)            -- textbook DFT, forward unnormalised,
                                        inverse 1/N; used through numpy here too.
  * pybaseutils.utils.detrend_*      -- un-vendored, unpinned; restated as mean /
                                        least-squares line removal (PARITY UNPINNED).
  * pybaseutils.sliding_window_1d    -- un-vendored, unpinned (ccf_sh only).
"""
import math

import numpy as np

# --------------------------------------------------------------------------- #
# windows.py:57-297
# --------------------------------------------------------------------------- #

_COSINE_SUMS = [
    # (substring, coefficients, ROV)  -- order matters (windows.py:101-217)
    ("3f", (0.26526, -0.5, 0.23474), 0.667),
    ("4f", (0.21706, -0.42103, 0.28294, -0.07897), 0.75),
    ("5f", (0.1881, -0.36923, 0.28702, -0.13077, 0.02488), 0.785),
    ("3m", (0.28235, -0.52105, 0.19659), 0.655),
    ("4m", (0.241906, -0.460841, 0.255381, -0.041872), 0.721),
    ("5m", (0.209671, -0.407331, 0.281225, -0.092669, 0.0091036), 0.760),
    ("3a", (0.40897, -0.5, 0.09103), 0.612),
    ("3b", (0.4243801, -0.4973406, 0.0782793), 0.598),
    ("3", (0.375, -0.5, 0.125), 0.647),
    ("4a", (0.338946, -0.481973, 0.161054, -0.018027), 0.68),
    ("4b", (0.355768, -0.487396, 0.144232, -0.012604), 0.663),
    ("4c", (0.3635819, -0.4891775, 0.1365995, -0.0106411), 0.656),
    ("4", (0.3125, -0.46875, 0.1875, -0.03125), 0.705),
]


def _window_family(name, beta=None):
    """(generator(n) -> float64[n], ROV) following the substring dispatch of
    windows.py:63-272 (first match wins; unmatched -> rectangular)."""
    low = name.lower()

    def cosine_sum(cc):
        def gen(n):
            z = 2.0 * np.pi * np.arange(n) / n        # windows.py:93, 223
            w = np.zeros(n)
            for ii, c in enumerate(cc):
                w += c if ii == 0 else c * np.cos(ii * z)
            return w
        return gen

    if "hann" in low:
        return np.hanning, 0.50                       # windows.py:63-70
    if "hamm" in low:
        return np.hamming, 0.50                       # windows.py:72-81
    if "black" in low:                                # windows.py:83-96
        return cosine_sum((0.35875, -0.48829, 0.14128, -0.01168)), 0.661
    if "nut" in low or "flat" in low or "sft" in low:  # windows.py:98-233
        for key, cc, rov in _COSINE_SUMS:
            if key in low:
                return cosine_sum(cc), rov
        raise NameError("cc")                          # reference leaves cc unbound
    if "kaiser" in low:                               # windows.py:235-240
        return (lambda n: np.kaiser(n, beta)), 2.0 / 3.0
    if "welch" in low:                                # windows.py:242-253
        def gen(n):
            z = 2.0 * np.arange(n) / n
            return 1.0 - (z - 1.0) * (z - 1.0)
        return gen, 0.293
    if "bart" in low:                                 # windows.py:255-262
        return np.bartlett, 0.50
    return (lambda n: np.ones(n)), 0.0                # windows.py:264-271


def window_vector(name, nwins, periodic=True, beta=None):
    """windows.py:274-283: periodic window = func(nwins+1)[:-1]."""
    gen, _ = _window_family(name, beta)
    if periodic:
        return np.asarray(gen(nwins + 1)[:-1], dtype=np.float64)
    return np.asarray(gen(nwins), dtype=np.float64)


def window_rov(name, beta=None):
    """windows.py:286-289: recommended overlap when nwins is absent."""
    return _window_family(name, beta)[1]


# --------------------------------------------------------------------------- #
# segmentation integers: fft_analysis.py:136-221, 2375-2383, 2411-2430, 2470-2484
# --------------------------------------------------------------------------- #

def sample_rate(tvec):
    return (len(tvec) - 1) / (tvec[-1] - tvec[0])            # :136, :2375-2377


def index_bounds(tvec, tbounds):
    fs = sample_rate(tvec)
    i0 = int(np.floor(fs * (tbounds[0] - tvec[0])))           # :143
    i1 = int(np.floor(1 + fs * (tbounds[1] - tvec[0])))       # :144
    return i0, i1


def get_nwins(nsig, navr, ov):
    n = int(np.floor(nsig * 1.0 / (navr - navr * ov + ov)))   # :2411-2418
    return min(n, nsig)


def get_noverlap(nwins, ov):
    return int(np.ceil(ov * nwins))                           # :2420-2422


def get_navr(nsig, nwins, noverlap):
    if nwins >= nsig:
        return 1
    return (nsig - noverlap) // (nwins - noverlap)           # :2424-2430


def get_nnyquist(nfft):
    return (nfft + 1) // 2 if nfft % 2 else nfft // 2        # :2470-2484


def detrend(x, style, axis=0):
    """fft_analysis.py:2538-2549 -> pybaseutils (PARITY UNPINNED)."""
    x = np.asarray(x, dtype=np.result_type(x, np.float64))
    if style is None or style == 0:
        return x
    if style > 0:
        return x - x.mean(axis, keepdims=True)
    n = x.shape[axis]
    t = np.arange(n, dtype=float)
    t -= t.mean()
    sh = [1] * x.ndim
    sh[axis] = n
    t = t.reshape(sh)
    xm = x.mean(axis, keepdims=True)
    slope = (t * (x - xm)).sum(axis, keepdims=True) / (t * t).sum()
    return x - xm - slope * t


def welch_plan(tvec, nx, tbounds=None, Navr=None, windowoverlap=None,
               windowfunction=None, tper=None, minFreq=None):
    """All the scalar bookkeeping of fft_pwelch (fft_analysis.py:102-221)."""
    if windowfunction is None:
        windowfunction = "Hanning"                            # :104-107
    if windowoverlap is None:
        windowoverlap = window_rov(windowfunction)            # :108-110
    if tbounds is None:
        tbounds = [tvec[0], tvec[-1]]                         # :119-120
    fs = sample_rate(tvec)
    i0, i1 = index_bounds(tvec, tbounds)
    nsig = len(tvec[i0:i1])                                   # :149
    calc_navr = Navr is None                                  # :102-103
    if minFreq is not None:
        tper = 2.0 / minFreq                                  # :180-181
    if tper is not None:
        nwins = int(fs * tper)                                # :182-183
    else:
        if Navr is None:
            Navr = 8                                          # :185-186
        calc_navr = False                                     # :188
        nwins = get_nwins(nsig, Navr, windowoverlap)          # :189
    noverlap = get_noverlap(nwins, windowoverlap)             # :194
    reflect = (i0 == 0 and i1 == len(tvec))                   # :197-198
    if reflect:
        nsig = nx + 2 * (nwins - 1)                           # :202-204
    if calc_navr:
        Navr = get_navr(nsig, nwins, noverlap)                # :208-209
    if nwins >= nsig:
        Navr, nwins = 1, nsig                                 # :214-216
    return dict(Fs=fs, i0=i0, i1=i1, nsig=nsig, nwins=nwins, noverlap=noverlap,
                Navr=int(Navr), reflect=reflect, windowoverlap=windowoverlap,
                windowfunction=windowfunction, Nnyquist=get_nnyquist(nwins))


def reflect_pad(x, nwins):
    """fft_analysis.py:202-203 (axis 0)."""
    return np.concatenate((x[nwins - 1:0:-1, ...], x, x[-1:-nwins:-1, ...]), axis=0)


def cxy_cxy2(Pxx, Pyy, Pxy):
    """fft_analysis.py:1662-1680."""
    Pxx = np.atleast_2d(Pxx)
    if Pxx.shape[1] != Pyy.shape[1]:
        Pxx = Pxx.T * np.ones((1, Pyy.shape[1]), dtype=Pxx.dtype)
    den = np.abs(Pxx) * np.abs(Pyy)
    return Pxy / np.sqrt(den), Pxy * np.conj(Pxy) / den


# --------------------------------------------------------------------------- #
# fft_pwelch home-brew path: fft_analysis.py:36-648
# --------------------------------------------------------------------------- #

def _one_sided_weights(nfft, nnyq):
    """x2 on bins 1..F-2 (and the last when nfft is odd): :414-421."""
    wgt = np.ones(nnyq)
    wgt[1:-1] = 2.0
    if nfft % 2:
        wgt[-1] = 2.0
    return wgt


def welch_sums(x_in, y_in, win, navr, hop, chunk=4096):
    """sum over segments of XX*, YY*, YX* (full two-sided, unscaled).

    fft_analysis.py:359-393 with the segment loop batched `chunk` at a time so
    that 2^26-sample records fit in memory.  x_in (n,), y_in (n, nch) already
    detrended.  Note Pxy = Y conj(X) (:393).
    """
    nw = len(win)
    nch = y_in.shape[1]
    cplx = np.iscomplexobj(x_in) or np.iscomplexobj(y_in)
    sxx = np.zeros(nw)
    syy = np.zeros((nch, nw))
    sxy = np.zeros((nch, nw), dtype=complex)
    from numpy.lib.stride_tricks import sliding_window_view
    xv = sliding_window_view(x_in, nw)[::hop][:navr]
    yv = sliding_window_view(y_in, nw, axis=0)[::hop][:navr]      # (K, nch, nw)
    assert xv.shape[0] == navr, (xv.shape, navr)
    for s in range(0, navr, chunk):
        X = np.fft.fft(xv[s:s + chunk] * win, axis=-1)             # :375, :386
        Y = np.fft.fft(yv[s:s + chunk] * win, axis=-1)             # :376, :387
        sxx += (X * np.conj(X)).real.sum(0)                        # :391
        syy += (Y * np.conj(Y)).real.sum(0)                        # :392
        sxy += (Y * np.conj(X)[:, None, :]).sum(0)                 # :393
    del cplx
    return sxx, syy, sxy


def fft_pwelch(tvec, sigx, sigy, tbounds=None, Navr=None, windowoverlap=None,
               windowfunction=None, detrend_style=None, onesided=None,
               tper=None, minFreq=None, chunk=4096, beta=None):
    """Averaged outputs of fft_pwelch(useMLAB=False) -- fft_analysis.py:36-648.

    Returns a dict with the tuple members (freq, Pxy, Pxx, Pyy, Cxy, phi_xy) and
    the fftinfo fields that do not need per-segment storage.
    """
    sigx = np.asarray(sigx)
    sigy = np.atleast_2d(np.asarray(sigy))                        # :163
    if sigy.shape[1] == len(tvec):
        sigy = sigy.T                                             # :164-165
    nch = sigy.shape[1]
    if detrend_style is None:
        detrend_style = 1                                         # :117-118
    if onesided is None:
        onesided = not (np.iscomplexobj(sigx) or np.iscomplexobj(sigy))  # :123-127
    p = welch_plan(tvec, sigx.shape[0], tbounds, Navr, windowoverlap,
                   windowfunction, tper, minFreq)
    nw, nov, K, fs = p["nwins"], p["noverlap"], p["Navr"], p["Fs"]
    i0, i1 = p["i0"], p["i1"]
    if p["reflect"]:
        sigx = reflect_pad(sigx, nw)
        sigy = reflect_pad(sigy, nw)
    nfft = nw                                                     # :219
    nnyq = p["Nnyquist"]
    win = window_vector(p["windowfunction"], nw, beta=beta)      # :231
    S1, S2 = win.sum(), (win ** 2.0).sum()                        # :241-242
    NENBW = nnyq * 1.0 * S2 / S1 ** 2                             # :245
    ENBW = fs * S2 / S1 ** 2                                      # :246
    x_in = detrend(sigx[i0:i1], detrend_style, axis=0)           # :353-357
    y_in = detrend(sigy[i0:i1, :], detrend_style, axis=0)
    sxx, syy, sxy = welch_sums(x_in, y_in, win, K, nw - nov, chunk)
    freq = np.fft.fftfreq(nfft, 1.0 / fs)                         # :396
    if onesided:
        wgt = _one_sided_weights(nfft, nnyq)
        freq = freq[:nnyq]                                        # :407
        sxx, syy, sxy = sxx[:nnyq] * wgt, syy[:, :nnyq] * wgt, sxy[:, :nnyq] * wgt
    else:
        freq = np.fft.fftshift(freq)                              # :423
        sxx = np.fft.fftshift(sxx, axes=-1)
        syy = np.fft.fftshift(syy, axes=-1)
        sxy = np.fft.fftshift(sxy, axes=-1)
    scale = 1.0 / (S1 ** 2) / ENBW / K                            # :432-446
    Pxx = (sxx * scale).astype(complex)
    Pyy = (syy * scale).T.astype(complex)                        # (nf, nch)
    Pxy = (sxy * scale).T
    Cxy, Cxy2 = cxy_cxy2(Pxx, Pyy, Pxy)                           # :489
    out = dict(p)
    out.update(freq=freq, Pxx=Pxx, Pyy=Pyy, Pxy=Pxy, Cxy=Cxy, Cxy2=Cxy2,
               phi_xy=np.arctan2(Pxy.imag, Pxy.real),             # :520
               win=win, S1=S1, S2=S2, NENBW=NENBW, ENBW=ENBW, nch=nch,
               onesided=onesided)
    out["varCxy"] = ((1.0 - Cxy * np.conjugate(Cxy)) / np.sqrt(2 * K)) ** 2.0   # :496
    out["varCxy2"] = 4.0 * Cxy2 * out["varCxy"]                                  # :498
    out["varPxx"] = (Pxx / np.sqrt(K)) ** 2.0                                    # :503-505
    out["varPyy"] = (Pyy / np.sqrt(K)) ** 2.0
    out["varPxy"] = (Pxy / np.sqrt(K)) ** 2.0
    out["varPhxy"] = (np.sqrt(1.0 - np.abs(Cxy2))) / np.sqrt(
        2 * K * np.sqrt(np.abs(Cxy2))) ** 2.0                                    # :514
    L = {}
    for nm, P in (("Lxx", Pxx), ("Lyy", Pyy), ("Lxy", Pxy)):
        a = np.sqrt(np.abs(ENBW * P))                             # :526-528
        if onesided:
            a[1:-1] = np.sqrt(2) * a[1:-1]                        # :533-535
            if nfft % 2:
                a[-1] = np.sqrt(2) * a[-1]
        L[nm] = a
    out.update(L)
    R = {}
    for nm, P in (("Rxx", Pxx), ("Ryy", Pyy), ("Rxy", Pxy), ("iCxy", Cxy)):
        if onesided:
            q = P.copy()
            if nm != "iCxy":                                      # :544-563
                q[1:-1, ...] *= 0.5
                if nfft % 2:
                    q[-1, ...] *= 0.5
            r = np.fft.irfft(q, n=nfft, axis=0)
        else:
            r = np.fft.ifft(np.fft.ifftshift(P, axes=0), n=nfft, axis=0)  # :571-574
        R[nm] = r * np.sqrt(nfft)                                 # :579-582
    Ex, Ey = R["Rxx"][0, ...].copy(), R["Ryy"][0, ...].copy()     # :585-586
    cc = R["Rxy"] / np.sqrt(np.ones((nfft, 1), dtype=R["Rxy"].dtype) * (Ex * Ey))  # :590
    for nm in R:
        R[nm] = np.fft.fftshift(R[nm], axes=0)                    # :592-595
    out.update(R)
    out.update(Ex=Ex, Ey=Ey, corrcoef=np.fft.fftshift(cc, axes=0),
               lags=(np.arange(1, nfft + 1) - nnyq) / fs)          # :596-597
    out["varLxx"] = (out["Lxx"] ** 2) * (out["varPxx"] / np.abs(Pxx) ** 2)      # :601-603
    out["varLyy"] = (out["Lyy"] ** 2) * (out["varPyy"] / np.abs(Pyy) ** 2)
    out["varLxy"] = (out["Lxy"] ** 2) * (out["varPxy"] / np.abs(Pxy) ** 2)
    if nch == 1:                                                  # :605-631
        for nm in ("Pyy", "Pxy", "Cxy", "Cxy2", "phi_xy", "Rxx", "Ryy", "Rxy",
                   "corrcoef", "iCxy", "Lxx", "Lyy", "Lxy", "varLxx", "varLyy",
                   "varLxy", "varCxy", "varCxy2", "varPxx", "varPyy", "varPxy",
                   "varPhxy"):
            out[nm] = out[nm].flatten()
    out["minFreq"] = 2.0 * fs / nw                                # :641
    return out


def fft_pwelch_port(tvec, sigx, sigy, Navr, windowoverlap=0.5,
                    windowfunction="Hanning", detrend_style=1):
    """Literal port of the home-brew hot loop (fft_analysis.py:339-446) with the
    same per-segment Python loop, the same five materialised (nch, K, nfft)
    complex128 temporaries and the same elementwise passes.  This is what
    bench.py times as the CPU baseline ("port"): it has the reference's cost
    structure, not an optimised one.  One-sided, whole-record analysis.
    Returns (freq, Pxy, Pxx, Pyy, Cxy)."""
    sigx = np.asarray(sigx)
    sigy = np.atleast_2d(np.asarray(sigy))
    if sigy.shape[1] == len(tvec):
        sigy = sigy.T
    nch = sigy.shape[1]
    p = welch_plan(tvec, sigx.shape[0], None, Navr, windowoverlap, windowfunction)
    nw, nov, K, fs, nnyq = p["nwins"], p["noverlap"], p["Navr"], p["Fs"], p["Nnyquist"]
    if p["reflect"]:
        sigx = reflect_pad(sigx, nw)
        sigy = reflect_pad(sigy, nw)
    win = window_vector(windowfunction, nw)
    S1, S2 = win.sum(), (win ** 2.0).sum()
    ENBW = fs * S2 / S1 ** 2
    x_in = detrend(sigx[p["i0"]:p["i1"]], detrend_style, axis=0)
    y_in = detrend(sigy[p["i0"]:p["i1"], :], detrend_style, axis=0)
    Pxx_seg = np.zeros((K, nw), dtype=np.complex128)
    Pyy_seg = np.zeros((nch, K, nw), dtype=np.complex128)
    Pxy_seg = np.zeros((nch, K, nw), dtype=np.complex128)
    Xfft = np.zeros((K, nw), dtype=np.complex128)
    Yfft = np.zeros((nch, K, nw), dtype=np.complex128)
    hop = nw - nov
    wcol = np.atleast_2d(win).T * np.ones((1, nch))
    for g in range(K):
        a = g * hop
        Xfft[g, :] = np.fft.fft(win * x_in[a:a + nw], n=nw, axis=0)
        Yfft[:, g, :] = np.fft.fft(wcol * y_in[a:a + nw, :], n=nw, axis=0).T
    Pxx_seg[:] = Xfft * np.conj(Xfft)
    Pyy_seg[:] = Yfft * np.conj(Yfft)
    Pxy_seg[:] = Yfft * (np.ones((nch, 1, 1), dtype=Xfft.dtype) * np.conj(Xfft))
    freq = np.fft.fftfreq(nw, 1.0 / fs)[:nnyq]
    Pxx_seg, Pyy_seg, Pxy_seg = Pxx_seg[:, :nnyq], Pyy_seg[:, :, :nnyq], Pxy_seg[:, :, :nnyq]
    Pxx_seg[:, 1:-1] = 2 * Pxx_seg[:, 1:-1]
    Pyy_seg[:, :, 1:-1] = 2 * Pyy_seg[:, :, 1:-1]
    Pxy_seg[:, :, 1:-1] = 2 * Pxy_seg[:, :, 1:-1]
    if nw % 2:
        Pxx_seg[:, -1] *= 2
        Pyy_seg[:, :, -1] *= 2
        Pxy_seg[:, :, -1] *= 2
    Pxx_seg = (1.0 / S1 ** 2) * Pxx_seg / ENBW
    Pyy_seg = (1.0 / S1 ** 2) * Pyy_seg / ENBW
    Pxy_seg = (1.0 / S1 ** 2) * Pxy_seg / ENBW
    Pxx = np.mean(Pxx_seg, axis=0)
    Pyy = np.mean(Pyy_seg, axis=1).T
    Pxy = np.mean(Pxy_seg, axis=1).T
    Cxy, _ = cxy_cxy2(Pxx, Pyy, Pxy)
    return freq, Pxy, Pxx, Pyy, Cxy


# --------------------------------------------------------------------------- #
# fftanal.fft_win (class STFT path): fft_analysis.py:2126-2203
# --------------------------------------------------------------------------- #

def fft_win(sig, tvec, win, navr, noverlap, onesided=True, detrend_style=1,
            detrendwin=False):
    """Returns (tseg, freq, Xseg (K, nf) complex128, pseg (K,))."""
    sig = np.asarray(sig)
    nw = len(win)
    nnyq = get_nnyquist(nw)
    fs = sample_rate(tvec)
    S1, S2 = win.sum(), (win ** 2.0).sum()
    ENBW = fs * S2 / S1 ** 2
    x_in = sig.copy() if detrendwin else detrend(sig, detrend_style, axis=0)  # :2147-2149
    hop = nw - noverlap
    X = np.zeros((navr, nw), dtype=complex)
    tt = np.zeros(navr)
    pseg = np.zeros(navr)
    for g in range(navr):
        a = g * hop
        seg = x_in[a:a + nw]
        if detrendwin:
            seg = detrend(seg, detrend_style, axis=0)             # :2169-2172
        seg = win * seg
        tt[g] = np.mean(tvec[a:a + nw])                           # :2163
        pseg[g] = np.trapezoid(seg * np.conj(seg), x=tvec[a:a + nw]).real   # :2174
        X[g] = np.fft.fft(seg, n=nw)
    freq = np.fft.fftfreq(nw, 1.0 / fs)
    if onesided:
        freq = freq[:nnyq]
        X = X[:, :nnyq]
        X[:, 1:-1] = np.sqrt(2) * X[:, 1:-1]                      # :2186
        if nw % 2:
            X[:, -1] = np.sqrt(2) * X[:, -1]
    else:
        freq = np.fft.fftshift(freq)
        X = np.fft.fftshift(X, axes=-1)
    X = X / S1                                                    # :2197
    pseg = pseg / S2                                              # :2198
    X = X / np.sqrt(ENBW)                                         # :2202
    return tt, freq, X, pseg


# --------------------------------------------------------------------------- #
# spectrogram.specgram: spectrogram.py:49-134 (hanning=True, overlap=True path)
# --------------------------------------------------------------------------- #

def specgram(t, s, wl=512, hanning=True, overlap=True):
    s = np.asarray(s).flatten()
    n = len(s)
    dt = np.abs(t[1] - t[0])
    nwin = 2 * (n - (n % wl)) // wl - 1 if overlap else (n - (n % wl)) // wl - 1   # :74-78
    out = np.zeros((wl, nwin))
    w = np.hanning(wl)                                            # symmetric (:109)
    for i in range(nwin):
        a = i * wl // 2 if overlap else i * wl
        seg = s[a:a + wl]
        if hanning:
            out[:, i] = np.sqrt(8.0 / 3.0) * np.abs(np.fft.fft(w * seg)) ** 2 / wl
        else:
            out[:, i] = np.abs(np.fft.fft(seg)) ** 2 / wl
    f = np.fft.fftfreq(wl, dt)
    if overlap:
        time = np.linspace(t[0] + wl * dt / 2, t[0] + wl * dt * ((nwin / 2 - 1) + 1 / 2), num=nwin)
    else:
        time = np.linspace(t[0] + wl * dt / 2, t[0] + wl * dt * ((nwin - 1) + 1 / 2), num=nwin)
    return time, f, out


# --------------------------------------------------------------------------- #
# hilbert.py:22-112
# --------------------------------------------------------------------------- #

def hilbert_mask(nfft):
    """The bin weights hilbert.py:47-64 applies (same for hilbert_1d :96-109)."""
    nyq = (nfft + 1) // 2 if nfft % 2 else nfft // 2
    h = np.zeros(nfft)
    h[0] = 1.0
    h[1:nyq] = 2.0
    if nyq < nfft:
        h[nyq] = 1.0
    return h


def hilbert(u, nfft=None, axis=-1):
    u = np.atleast_1d(np.asarray(u))
    if nfft is None:
        nfft = u.shape[axis]
    U = np.fft.fft(u, n=nfft, axis=axis)                          # :54
    sh = [1] * U.ndim
    sh[axis] = nfft
    U = U * hilbert_mask(nfft).reshape(sh)                        # :63-64
    return np.fft.ifft(U, n=nfft, axis=axis).squeeze()           # :67


# --------------------------------------------------------------------------- #
# ccf.py:66-77 (ccf), :239-248 (fftcorr intent)
# --------------------------------------------------------------------------- #

def ccf_direct(x1, x2, fs):
    """Literal ccf.py:66-77 (O(N^2) numpy.correlate)."""
    x1 = np.asarray(x1, float)
    x2 = np.asarray(x2, float)
    npts = len(x1)
    lags = np.arange(-npts + 1, npts)
    tau = -lags / float(fs)
    ccov = np.correlate(x1 - x1.mean(), x2 - x2.mean(), mode="full")
    return tau, ccov / (npts * x1.std() * x2.std())


def ccf_fft(x1, x2, fs):
    """Same numbers as ccf_direct through the FFT identity (ccf.py:239-248:
    'Rxy <==> X'Y' with zero padding to >= 2N-1)."""
    x1 = np.asarray(x1, float)
    x2 = np.asarray(x2, float)
    n = len(x1)
    L = 1 << int(math.ceil(math.log2(2 * n - 1)))
    a = x1 - x1.mean()
    b = x2 - x2.mean()
    r = np.fft.irfft(np.fft.rfft(a, L) * np.conj(np.fft.rfft(b, L)), L)
    full = np.concatenate((r[L - (n - 1):], r[:n]))
    lags = np.arange(-n + 1, n)
    return -lags / float(fs), full / (n * x1.std() * x2.std())


# --------------------------------------------------------------------------- #
# multi-channel cross-spectral matrix (config 5): the reference builds it by
# calling fft_pwelch once per reference channel with all channels as sigy
# (HeatPulse_Funcs.py:576-583) -> G[f, i, j] = Pxy of (sigx = ch i, sigy = ch j)
#                                            = mean_k Y_j conj(X_i) * scale
# --------------------------------------------------------------------------- #

def cross_spectral_matrix(tvec, sig, Navr, windowoverlap=0.5,
                          windowfunction="Hanning", detrend_style=1, chunk=512):
    sig = np.asarray(sig)                                         # (n, C)
    p = welch_plan(tvec, sig.shape[0], None, Navr, windowoverlap, windowfunction)
    nw, nov, K, fs, nnyq = p["nwins"], p["noverlap"], p["Navr"], p["Fs"], p["Nnyquist"]
    if p["reflect"]:
        sig = reflect_pad(sig, nw)
    win = window_vector(windowfunction, nw)
    S1, S2 = win.sum(), (win ** 2.0).sum()
    ENBW = fs * S2 / S1 ** 2
    x = detrend(sig[p["i0"]:p["i1"], :], detrend_style, axis=0)
    from numpy.lib.stride_tricks import sliding_window_view
    v = sliding_window_view(x, nw, axis=0)[::nw - nov][:K]        # (K, C, nw)
    C = sig.shape[1]
    G = np.zeros((nnyq, C, C), dtype=complex)
    for s in range(0, K, chunk):
        X = np.fft.fft(v[s:s + chunk] * win, axis=-1)[:, :, :nnyq]   # (k, C, F)
        G += np.einsum("kjf,kif->fij", X, np.conj(X))
    wgt = _one_sided_weights(nw, nnyq)
    G *= (wgt / (S1 ** 2) / ENBW / K)[:, None, None]
    return np.fft.fftfreq(nw, 1.0 / fs)[:nnyq], G


# --------------------------------------------------------------------------- #
# fft_analysis.py:1419-1588 (rescale / unscale / fft_deriv), SURVEY 8(f) rank 3
# --------------------------------------------------------------------------- #

def fft_deriv(sig, xx=None, lowpass=True, Fs_new=None, modified=True, detrend=None, window=None):
    """Spectral derivative dsig/dxx on a uniform grid: ifft(wavenumber * fft(sig)) with the
    modified wavenumber i sin(k dx)/dx (fft_analysis.py:1453-1588).  The pre-filter branch
    (:1493-1511) only acts when Fs_new < Fs, i.e. never with the defaults (lowpass = Nyquist,
    Fs_new = min(5 lowpass, Fs) = Fs); it belongs to filters.py (out of scope) and is refused."""
    sig = np.asarray(sig, float)
    if xx is None:
        xx = 1.0 * np.arange(len(sig))                                    # :1484-1487
    xx = np.asarray(xx, float)
    if lowpass:                                                           # :1493-1511
        dxo = xx[1] - xx[0]
        lp = 0.5 / dxo if lowpass is True else float(lowpass)
        Fs = 1.0 / dxo
        fsn = min(5.0 * lp, Fs) if Fs_new is None else Fs_new
        if fsn < Fs:
            raise NotImplementedError("fft_deriv: the down-sampling pre-filter is filters.py territory")
    # rescale (:1419-1437): y to [0, 1], x by its span with the fixed offset -1e-4
    slope = np.nanmax(sig) - np.nanmin(sig)
    offset = np.nanmin(sig)
    if slope == 0:
        slope = 1.0
    s = (sig - offset) / slope
    xslope = np.nanmax(xx) - np.nanmin(xx)
    xoffset = -1e-4
    if xslope == 0:
        xslope = 1.0
    x = (xx - xoffset) / xslope
    if detrend is not None:
        s = detrend(s)                                                    # :1519
    N = len(x)
    dx = x[1] - x[0]
    L = N * dx
    k = 2.0 * np.pi * np.fft.fftfreq(N, d=dx / L)                         # :1529-1530
    wavenumber = (1.0j * np.sin(k * dx) / dx if modified else 1.0j * k) / L   # :1531-1540
    win = np.ones_like(s) if window is None else window(N)                # :1543-1547
    s = win * s
    ds0 = (s[1] - s[0]) / (x[1] - x[0])                                   # :1551-1552
    ds1 = (s[-1] - s[-2]) / (x[-1] - x[-2])
    d = np.real(np.fft.ifft(wavenumber * np.fft.fft(s, n=N), n=N))        # :1553
    d = d / win                                                           # :1556
    d[0] = ds0                                                            # :1559-1560
    d[-1] = ds1
    # unscale (:1439-1451)
    return d * slope / xslope, x * xslope + xoffset


# --------------------------------------------------------------------------- #
# SURVEY 8(f) rank 4
# --------------------------------------------------------------------------- #
def cog(x, fs, fmin=None, fmax=None):
    """Doppler.cog (Doppler.py:43-61), line by line -- including the band selection, which indexes the spectrum
    with a mask computed from the already filtered frequency vector (so the first len(freq) entries are kept)."""
    if fmax is None:
        fmax = fs
    n = len(x)
    freq = np.fft.fftshift(np.fft.fftfreq(n, 1 / fs))
    spec = np.fft.fftshift(np.fft.fft(x)) / np.sqrt(n / 2)
    if fmin is not None:
        freq = freq[np.where((np.abs(freq) >= fmin) * (np.abs(freq) <= fmax))]
        spec = spec[np.where((np.abs(freq) >= fmin) * (np.abs(freq) <= fmax))]
    if len(freq) > 0:
        return np.sum(np.abs(np.square(spec)) * freq) / np.sum(np.abs(np.square(spec)))
    return 0.0


def upsample(u_t, Fs, Fs_new):
    """filters.upsample (filters.py:20-34) with numpy.interp standing in for pybaseutils.utils.interp (absent from
    the image and from the reference tree: PARITY UNPINNED for this function)."""
    u_t = np.asarray(u_t)
    nt = len(u_t)
    tt = np.arange(0, nt, 1) / Fs
    ti = np.arange(tt[0], tt[-1], 1 / Fs_new)
    if u_t.ndim == 1:
        return np.interp(ti, tt, u_t)
    return np.stack([np.interp(ti, tt, u_t[:, c]) for c in range(u_t.shape[1])], axis=1)
