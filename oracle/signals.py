"""TEST INFRASTRUCTURE ONLY -- seeded synthetic inputs for the BASELINE.json configs.

Recipes follow SURVEY.md section 8(d) (which in turn reuses the reference's own
test-signal recipes: fft_analysis.py:2899-2993, spectrogram.py:192-201,
hilbert.py:117-128, ccf.py:139-148).  All arrays are float64; fp32 tests round
them to float32 first and feed the SAME rounded values to the oracle.
"""
import numpy as np

FS = 1.0e6


def tvec(n, fs=FS):
    return np.arange(n) / fs


def c1_auto(n=2 ** 20, nfft=1024, seed=1, bin_centred=False, fs=FS):
    """config 1: sine + 0.1 noise, auto-PSD (sigy = sigx)."""
    rng = np.random.default_rng(seed)
    t = tvec(n, fs)
    f0 = 100 * fs / nfft if bin_centred else 50e3 * (1 + 1.0 / nfft)
    x = np.sin(2 * np.pi * f0 * t) + 0.1 * rng.standard_normal(n)
    return t, x


def c2_cross(n=2 ** 26, seed=2, dc=(0.0, 0.0), fs=FS, f0=97.65625e3, dtype=np.float64):
    """config 2: two channels sharing a tone (y lags by pi/4) + independent unit noise.
    dc=(7.0, 2.5) is the large-DC detrend stress variant (fft_analysis.py:2962-2974)."""
    rng = np.random.default_rng(seed)
    t = tvec(n, fs)
    ph = 2 * np.pi * f0 * t
    x = np.sin(ph)
    x += rng.standard_normal(n)
    x += dc[0]
    y = 0.5 * np.sin(ph - np.pi / 4)
    y += rng.standard_normal(n)
    y += dc[1]
    return t, x.astype(dtype), y.astype(dtype)


def c3_chirps(n=2 ** 24, nch=64, seed=3, fs=FS):
    """config 3: staggered linear chirps 10 -> 200 kHz + 0.1 noise (spectrogram.py:192-201)."""
    rng = np.random.default_rng(seed)
    t = tvec(n, fs)
    out = np.empty((nch, n))
    for c in range(nch):
        f = np.linspace(10e3, 200e3, n) * (1.0 + 0.01 * c)
        out[c] = np.sin(2 * np.pi * np.cumsum(f) / fs) + 0.1 * rng.standard_normal(n)
    return t, out


def c4_am(n=2 ** 25, nch=16, seed=4, fs=FS):
    """config 4: AM carrier with per-channel integer delays 37*c samples + 0.05 noise."""
    rng = np.random.default_rng(seed)
    out = np.empty((nch, n))
    k = np.arange(n + 37 * nch)
    base = (1 + 0.5 * np.sin(2 * np.pi * 1e3 * k / fs)) * np.sin(2 * np.pi * 1e5 * k / fs)
    base += 0.3 * rng.standard_normal(len(k))          # shared broadband part -> sharp ccf peak
    for c in range(nch):
        d = 37 * c
        out[c] = base[nch * 37 - d: nch * 37 - d + n] + 0.05 * rng.standard_normal(n)
    return tvec(n, fs), out


def c5_mixture(n=2 ** 28, nch=32, nsrc=4, seed=5, fs=FS):
    """config 5: nch channels = fixed mixing matrix x nsrc narrow-band sources + unit noise."""
    rng = np.random.default_rng(seed)
    t = tvec(n, fs)
    mix = np.random.default_rng(55).standard_normal((nch, nsrc))
    src = np.stack([np.sin(2 * np.pi * (20e3 + 35e3 * s) * t + 0.3 * s) for s in range(nsrc)])
    out = mix @ src
    out += rng.standard_normal((nch, n))
    return t, out


def ref_test_fftpwelch(n=2 ** 14):
    """The reference's own self-test recipe, fft_analysis.py:2899-2944."""
    from scipy.signal import square
    df = 5.0
    t = (1.0 / df) * np.arange(0.0, 1.0, 1.0 / n)
    x = 0.1 * square(2 * np.pi * 150.0 * t) + 7.0
    y = 0.007 * np.sin(2 * np.pi * 150.0 * t - np.pi / 4) + 2.5
    return t, x, y
