"""TEST INFRASTRUCTURE ONLY -- writes tests/golden/*.npz from the UNMODIFIED reference.

Run in the build container (needs /root/reference):

    python -m oracle.gen_golden

Each fixture stores the inputs (so nothing depends on RNG stream stability), the
call's keyword arguments and every averaged output of the reference call.  The
fixtures pin (a) oracle/restatement.py (tests/test_oracle_golden.py, CPU) and
(b) the CUDA path (tests/test_gpu_golden.py, GPU, through the C ABI).
"""
import json
import os
import sys

import numpy as np

from . import ref_shim, signals

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                   "tests", "golden")

WELCH_FIELDS = ["Cxy2", "Lxx", "Lyy", "Lxy", "Rxx", "Ryy", "Rxy", "iCxy", "corrcoef",
                "lags", "Ex", "Ey", "varCxy", "varCxy2", "varPxx", "varPyy", "varPxy",
                "varPhxy", "varLxx", "varLyy", "varLxy", "S1", "S2", "ENBW", "NENBW",
                "win", "minFreq", "nwins", "noverlap", "Navr", "nch", "Fs"]


def _pack_t(t):
    """tvec is stored as (n, fs) when it is exactly arange(n)/fs, else verbatim."""
    n = len(t)
    if np.array_equal(t, np.arange(n) / signals.FS):
        return dict(in_t_n=np.asarray(n), in_t_fs=np.asarray(signals.FS))
    return dict(in_t=t)


def _welch(fa, name, t, x, y, tbounds=None, **kw):
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        r = fa.fft_pwelch(t, x, y, tbounds, useMLAB=False, plotit=False, verbose=False, **kw)
    info = r[6]
    d = dict(in_x=x, in_y=y, freq=r[0], Pxy=r[1], Pxx=r[2], Pyy=r[3], Cxy=r[4],
             phi_xy=r[5])
    for f in WELCH_FIELDS:
        d[f] = np.asarray(getattr(info, f))
    d["ibnds"] = np.asarray(info.ibnds)
    d.update(_pack_t(t))
    meta = dict(kw)
    meta["tbounds"] = None if tbounds is None else [float(v) for v in tbounds]
    d["kwargs_json"] = np.asarray(json.dumps(meta))
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **d)
    print("wrote", name, "K=%d L=%d" % (info.Navr, info.nwins))


def main():
    os.makedirs(OUT, exist_ok=True)
    m = ref_shim.load_reference()
    fa, W, SP, H = m["fft_analysis"], m["windows"], m["spectrogram"], m["hilbert"]
    rng = np.random.default_rng(2026)

    # ---- windows(): every family at two lengths + ROV ---------------------- #
    names = ["Hanning", "Hamming", "Blackman", "Nuttall3", "nut3a", "nut3b", "nut4",
             "nut4a", "nut4b", "nut4c", "SFT3F", "SFT4F", "SFT5F", "SFT3M", "SFT4M",
             "SFT5M", "welch", "bartlett", "box"]
    d = {}
    for nm in names:
        for n in (64, 129):
            d["%s_%d" % (nm, n)] = W.windows(nm, nwins=n, verbose=False)
            d["%s_%d_aper" % (nm, n)] = W.windows(nm, nwins=n, periodic=False, verbose=False)
        d["%s_rov" % nm] = np.asarray(W.windows(nm, verbose=False))
    d["kaiser8_64"] = W.windows("kaiser", nwins=64, beta=8.0, verbose=False)
    d["names_json"] = np.asarray(json.dumps(names))
    np.savez_compressed(os.path.join(OUT, "windows.npz"), **d)

    # ---- fft_pwelch -------------------------------------------------------- #
    t, x = signals.c1_auto(n=2 ** 15)
    _welch(fa, "welch_c1_auto", t, x, x, Navr=63, windowoverlap=0.5,
           windowfunction="Hanning", detrend_style=1)
    t, x = signals.c1_auto(n=2 ** 14, bin_centred=True)
    _welch(fa, "welch_c1_bincentred", t, x, x, Navr=31, windowoverlap=0.5,
           windowfunction="Hanning", detrend_style=1)
    t, x, y = signals.c2_cross(n=2 ** 15)
    _welch(fa, "welch_c2_cross", t, x, y, Navr=15, windowoverlap=0.5,
           windowfunction="Hanning", detrend_style=1)
    t, x, y = signals.c2_cross(n=2 ** 15, dc=(7.0, 2.5))
    _welch(fa, "welch_c2_dc", t, x, y, Navr=15, windowoverlap=0.5,
           windowfunction="Hanning", detrend_style=1)
    # multi-channel sigy, flat-top window at its ROV, linear detrend
    t = signals.tvec(20000)
    x = np.sin(2 * np.pi * 31e3 * t) + 0.2 * rng.standard_normal(len(t)) + 1e-4 * np.arange(len(t))
    y = np.stack([np.sin(2 * np.pi * 31e3 * t - 0.4 * c) + 0.2 * rng.standard_normal(len(t)) + c
                  for c in range(3)], 1)
    _welch(fa, "welch_multi_sft3f_linear", t, x, y, Navr=12, windowfunction="SFT3F",
           detrend_style=-1)
    # defaults everywhere (Navr=8, Hanning ROV, mean detrend)
    _welch(fa, "welch_defaults", t[:4096 * 4], x[:4096 * 4], y[:4096 * 4, 0])
    # two-sided, complex x
    xc = x[:8192] + 1j * rng.standard_normal(8192)
    _welch(fa, "welch_twosided_complex", t[:8192], xc, y[:8192, :2], Navr=15,
           windowoverlap=0.5, windowfunction="Hamming", detrend_style=1)
    # partial record (no reflect padding), explicit tper, no detrend
    _welch(fa, "welch_tbounds_tper", t, x, y[:, :2], [0.002, 0.0121], windowoverlap=0.5,
           windowfunction="Hanning", detrend_style=0, tper=1024 / signals.FS * (1 + 1e-9))
    # odd window length (non power of two)
    _welch(fa, "welch_odd_nfft", t[:12345], x[:12345], y[:12345, 0], Navr=7,
           windowoverlap=0.3, windowfunction="bartlett", detrend_style=1)
    # the reference's own self-test recipe (fft_analysis.py:2899-2944)
    t, x, y = signals.ref_test_fftpwelch()
    _welch(fa, "welch_ref_selftest", t, x, y, [t[1], t[-1]], Navr=8, windowoverlap=0.5,
           windowfunction="hamming", detrend_style=1)

    # ---- fftanal class STFT path (what spectrogram.stft returns) ----------- #
    t, ch = signals.c3_chirps(n=2 ** 15, nch=2)
    tper = 2048 / signals.FS * (1 + 1e-9)
    d = dict(in_y=ch, tper=tper)
    d.update(_pack_t(t))
    for c in range(2):
        o = SP.stft(t, ch[c], tper=tper, windowoverlap=0.75, windowfunction="Hanning",
                    verbose=False)
        d.update({"Xseg_%d" % c: o.Xseg, "Xpow_%d" % c: o.Xpow, "Xfft_%d" % c: o.Xfft,
                  "Pxx_%d" % c: o.Pxx})
    d.update(tseg=o.tseg, freq=o.freq, nwins=o.nwins, noverlap=o.noverlap, Navr=o.Navr,
             S1=o.S1, S2=o.S2, ENBW=o.ENBW, ibounds=np.asarray(o.ibounds))
    np.savez_compressed(os.path.join(OUT, "stft_c3.npz"), **d)
    # two-signal class path up to Pstft (averagewins breaks for 1-D sigy)
    t, ch = t[:2 ** 14], ch[:, :2 ** 14]
    o = fa.fftanal(t, ch[0], ch[1], tper=tper, windowoverlap=0.5, verbose=False)
    o.Xstft(); o.Ystft(); o.Pstft()
    np.savez_compressed(os.path.join(OUT, "stft_xy.npz"), **_pack_t(t), in_x=ch[0], in_y=ch[1],
                        tper=tper, Xseg=o.Xseg, Yseg=o.Yseg, Pxy_seg=o.Pxy_seg,
                        Pxx_seg=o.Pxx_seg, Pyy_seg=o.Pyy_seg, Cxy_seg=o.Cxy_seg,
                        phixy_seg=o.phixy_seg, tseg=o.tseg, freq=o.freq, Navr=o.Navr,
                        nwins=o.nwins, noverlap=o.noverlap)

    # ---- specgram ---------------------------------------------------------- #
    stdout, sys.stdout = sys.stdout, open(os.devnull, "w")
    try:
        tm, f, S = SP.specgram(t, ch[0], wl=512)
        tm2, f2, S2 = SP.specgram(t[:10000], ch[1][:10000], wl=256, hanning=False)
    finally:
        sys.stdout = stdout
    np.savez_compressed(os.path.join(OUT, "specgram.npz"), **_pack_t(t), in_s=ch[0], time=tm,
                        freq=f, S=S, in_s2=ch[1][:10000], time2=tm2, freq2=f2, S2=S2)

    # ---- hilbert ----------------------------------------------------------- #
    d = {}
    for n in (32, 33, 1024, 1000, 4096):
        u = np.sin(2 * np.pi * 3 * np.arange(n) / n) + 0.3 * rng.standard_normal(n)
        d["u_%d" % n] = u
        d["z_%d" % n] = H.hilbert(u)
        d["z1d_%d" % n] = H.hilbert_1d(u)
    u2 = rng.standard_normal((4, 512))
    d["u_batch"] = u2
    d["z_batch"] = H.hilbert(u2)
    np.savez_compressed(os.path.join(OUT, "hilbert.npz"), **d)
    print("golden fixtures in", OUT)


def gen_fft_deriv():
    """fft_analysis.fft_deriv (fft_analysis.py:1453-1588) on the recipes of its own self-test
    (test_fft_deriv, :1591-1650: smooth profiles on a uniform grid), default arguments and the
    unmodified-wavenumber / windowed variants.  Own RNG: does not disturb the other fixtures."""
    m = ref_shim.load_reference()
    fa = m["fft_analysis"]
    rng = np.random.default_rng(1453)
    d = {}
    for n in (256, 4096):
        xx = np.linspace(-1.0, 1.3, n)
        sig = np.exp(-0.5 * (xx / 0.3) ** 2) + 0.2 * np.sin(2 * np.pi * 3 * xx) + 0.01 * rng.standard_normal(n)
        d["x_%d" % n] = xx
        d["s_%d" % n] = sig
        for tag, kw in (("def", {}), ("naive", dict(modified=False)), ("ham", dict(window=np.hamming)),
                        ("mean", dict(detrend=fa.detrend_mean))):
            ds, xo = fa.fft_deriv(sig.copy(), xx.copy(), **kw)
            d["d_%s_%d" % (tag, n)] = ds
            d["xo_%s_%d" % (tag, n)] = xo
    ds, xo = fa.fft_deriv(d["s_256"].copy())                       # xx = None: index grid
    d["d_noxx_256"], d["xo_noxx_256"] = ds, xo
    np.savez_compressed(os.path.join(OUT, "fft_deriv.npz"), **d)
    print("wrote fft_deriv")


def gen_extras():
    """Round-2 fixtures from the unmodified reference (own RNG): getNpeaks (fft_analysis.py:940-1027),
    fftanal.fftpwelch + crosscorr (:1794-1803, :1840-1878; 1-D sigy -- with 2-D sigy, and for crosscorr_stft /
    convert2amplitudes in every case, the reference itself raises), mean_angle (:1334-1377),
    fftanal.intspectra (:2514-2536) and laplace.laplace_1d (laplace.py:85-113)."""
    import importlib.util
    import warnings
    warnings.simplefilter("ignore")
    m = ref_shim.load_reference()
    fa = m["fft_analysis"]
    rng = np.random.default_rng(835)
    n = 2 ** 15
    t = np.arange(n) / 1.0e5
    x = np.sin(2 * np.pi * 3e3 * t) + 0.6 * np.sin(2 * np.pi * 11e3 * t + 0.3) + 0.3 * rng.standard_normal(n)
    y = 0.5 * np.sin(2 * np.pi * 3e3 * t - 0.4) + 0.25 * np.sin(2 * np.pi * 11e3 * t + 1.1) \
        + 0.1 * np.sin(2 * np.pi * 23e3 * t) * 0 + 0.2 * rng.standard_normal(n) + 0.05 * np.roll(x, 7)
    d = dict(t_n=np.asarray(n), t_fs=np.asarray(1.0e5), x=x, y=y)
    kw = dict(Navr=63, windowoverlap=0.5, windowfunction="Hanning", detrend_style=1)
    for npk, tag in ((1, "1"), (4, "4")):
        out = fa.getNpeaks(npk, t, x, y, plotit=False, useMLAB=False, verbose=False, **kw)
        d["peaks_" + tag] = np.asarray([[float(v) for v in row] for row in out])
    out = fa.getNpeaks(3, t, x, y, plotit=False, useMLAB=False, verbose=False, minsep=2000.0, **kw)
    d["peaks_minsep"] = np.asarray([[float(v) for v in row] for row in out])
    o = fa.fftanal(t, x, y, Navr=63, windowoverlap=0.5, verbose=False)
    o.fftpwelch()
    o.crosscorr()
    for nm in ("Pxx", "Pyy", "Pxy", "Rxx", "Ryy", "Rxy", "corrcoef", "Ex", "Ey", "lags", "freq"):
        d["cls_" + nm] = np.asarray(getattr(o, nm))
    ph = rng.uniform(-1.2, 1.2, (50, 3))
    vph = rng.uniform(0.0, 0.05, (50, 3))
    d["ma_phi"], d["ma_vphi"] = ph, vph
    for tag, args in (("a", dict()), ("b", dict(vphi=vph)), ("c", dict(vphi=vph, dim=1, angle_range=np.pi))):
        mp, vp = fa.mean_angle(ph, **args)
        d["ma_mean_" + tag], d["ma_var_" + tag] = np.asarray(mp), np.asarray(vp)
    fr = np.arange(257) * 195.3125
    sp = np.abs(rng.standard_normal(257)) + 20.0 * np.exp(-0.5 * ((np.arange(257) - 90) / 2.0) ** 2)
    d["is_freq"], d["is_spec"] = fr, sp
    d["is_default"] = np.asarray(fa.fftanal.intspectra(fr, sp)[0])
    d["is_enbw"] = np.asarray(fa.fftanal.intspectra(fr, sp, ENBW=600.0)[0])
    d["is_whole"] = np.asarray(fa.fftanal.intspectra(fr, sp, ifreq=40)[0])      # ifreq given, ispan None: whole range
    # (ifreq AND ispan given leaves ilow unbound in the reference: UnboundLocalError, :2528-2534)
    spec = importlib.util.spec_from_file_location("FFT.laplace", os.path.join(ref_shim.REFERENCE_DIR, "laplace.py"))
    lp = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(lp)
    for nl in (300, 512):
        u = np.cos(2 * np.pi * 0.02 * np.arange(nl) - 0.2) * np.exp(-np.arange(nl) / 150.0) + 0.05 * rng.standard_normal(nl)
        sig = np.arange(-0.04, 0.04 + 1e-9, 0.01)
        d["lp_u_%d" % nl], d["lp_sigma_%d" % nl] = u, sig
        d["lp_out_%d" % nl] = lp.laplace_1d(u, sig)
    np.savez_compressed(os.path.join(OUT, "extras.npz"), **d)
    print("wrote extras", d["peaks_4"])


def gen_segstats():
    """Statistics over the segments of the reference's own per-segment spectra (fft_analysis.py:397-460:
    fftinfo.Pxx_seg / Pyy_seg / Pxy_seg / phixy_seg of the unmodified fft_pwelch): np.var (the estimates the
    reference comments out at :449-458) and the reference's mean_angle (:1334-1377) of phixy_seg."""
    import warnings
    warnings.simplefilter("ignore")
    m = ref_shim.load_reference()
    fa = m["fft_analysis"]
    rng = np.random.default_rng(4417)
    d = {}
    cases = (("a", 2 ** 14, 31, 1, True, 1), ("b", 2 ** 14, 63, 2, False, 1), ("c", 3 * 2 ** 12, 23, 1, True, -1),
             ("d", 2 ** 13, 127, 3, True, 0))
    for tag, n, navr, nch, onesided, detrend in cases:
        t = np.arange(n) / 2.0e5
        x = np.sin(2 * np.pi * 9.0e3 * t) + 0.5 * rng.standard_normal(n) + 0.8 + 3.0 * t
        ys = [0.6 * np.sin(2 * np.pi * 9.0e3 * t + 0.5 * (c + 1)) + 0.3 * rng.standard_normal(n) + 0.1 * x - 0.4 * c
              for c in range(nch)]
        y = ys[0] if nch == 1 else np.stack(ys, axis=1)
        out = fa.fft_pwelch(t, x, y, None, Navr=navr, windowoverlap=0.5, windowfunction="Hanning", useMLAB=False,
                            plotit=False, verbose=False, detrend_style=detrend, onesided=onesided)
        info = out[6]
        pxx, pyy, pxy, ph = (np.asarray(getattr(info, k)) for k in ("Pxx_seg", "Pyy_seg", "Pxy_seg", "phixy_seg"))
        d[tag + "_x"], d[tag + "_y"], d[tag + "_fs"] = x, y, np.asarray(2.0e5)
        d[tag + "_cfg"] = np.asarray([n, navr, nch, int(onesided), detrend])
        d[tag + "_navr"] = np.asarray(info.Navr)
        d[tag + "_varPxx"] = np.var(pxx, axis=0)
        d[tag + "_varPyy"] = np.var(pyy, axis=1).T
        d[tag + "_varPxy"] = (np.var(pxy.real, axis=1) + 1j * np.var(pxy.imag, axis=1)).T
        d[tag + "_Pxx"], d[tag + "_Pyy"], d[tag + "_Pxy"] = (np.asarray(out[i]) for i in (2, 3, 1))
        mp, vp = fa.mean_angle(ph, dim=1)
        d[tag + "_phi_mean"], d[tag + "_varphi_mean"] = np.asarray(mp).T, np.asarray(vp).T
    np.savez_compressed(os.path.join(OUT, "segstats.npz"), **d)
    print("wrote segstats", {k: v.shape for k, v in d.items() if k.startswith("b_")})


def gen_cog():
    """Doppler.cog of the unmodified reference (Doppler.py:43-61).  Doppler.py itself cannot be imported as it is
    -- it imports FFT.stft, a module the reference does not have -- so that name is bound to the reference's own
    spectrogram.stft for the import; cog is called untouched."""
    import importlib.util
    import types
    import warnings
    warnings.simplefilter("ignore")
    m = ref_shim.load_reference()
    st = types.ModuleType("FFT.stft")
    st.stft = m["spectrogram"].stft
    sys.modules["FFT.stft"] = st
    mods = {}
    for name in ("filters", "Doppler"):
        spec = importlib.util.spec_from_file_location("FFT." + name, os.path.join(ref_shim.REFERENCE_DIR, name + ".py"))
        mods[name] = importlib.util.module_from_spec(spec)
        sys.modules["FFT." + name] = mods[name]
        spec.loader.exec_module(mods[name])
    cog = mods["Doppler"].cog
    rng = np.random.default_rng(77)
    d = {}
    fs = 2.0e6
    for tag, n, cplx in (("a", 512, True), ("b", 1000, True), ("c", 4096, False), ("d", 333, True)):
        t = np.arange(n) / fs
        f0 = 1.3e5 + 4.0e4 * np.sin(2 * np.pi * 3.0e3 * t)
        ph = 2 * np.pi * np.cumsum(f0) / fs
        x = np.exp(1j * ph) + 0.2 * (rng.standard_normal(n) + 1j * rng.standard_normal(n))
        if not cplx:
            x = x.real
        d[tag + "_x"] = x
        d[tag + "_cog"] = np.asarray(cog(x, fs))
        d[tag + "_cog_band"] = np.asarray(cog(x, fs, fmin=5.0e4, fmax=4.0e5))
        d[tag + "_cog_none"] = np.asarray(cog(x, fs, fmin=3.0e6))           # empty selection -> 0.0
    # sliding windows of one long record, each through the reference's cog
    n, win, step = 2 ** 14, 512, 256
    t = np.arange(n) / fs
    ph = 2 * np.pi * np.cumsum(2.0e5 + 1.0e5 * np.sin(2 * np.pi * 2.0e3 * t)) / fs
    x = np.exp(1j * ph) + 0.3 * (rng.standard_normal(n) + 1j * rng.standard_normal(n))
    d["s_x"], d["s_fs"] = x, np.asarray(fs)
    d["s_cog"] = np.asarray([cog(x[i:i + win], fs) for i in range(0, n - win + 1, step)])
    np.savez_compressed(os.path.join(OUT, "cog.npz"), **d)
    print("wrote cog", d["a_cog"], d["a_cog_band"], d["a_cog_none"], d["s_cog"][:3])


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "cog":
        gen_cog()
    elif len(sys.argv) > 1 and sys.argv[1] == "segstats":
        gen_segstats()
    elif len(sys.argv) > 1 and sys.argv[1] == "fft_deriv":
        gen_fft_deriv()
    elif len(sys.argv) > 1 and sys.argv[1] == "extras":
        gen_extras()
    else:
        main()
