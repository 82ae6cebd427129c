"""Window factory for the spectral-estimation path (host side, O(nfft)).

Drop-in for the reference's ``windows.windows`` (windows.py:57-297): same name
matching (case-insensitive substring tests in the same order, first match wins,
unmatched names give the rectangular window), same recommended overlaps (ROV),
same "periodic = func(nwins+1)[:-1]" construction (windows.py:274-279) and the
same return conventions (``nwins`` given -> window vector, else ROV; ``msgout``
adds the description tuple).  The GPU takes the window as a coefficient vector,
so this stays numpy float64 exactly like the reference.

Window figures of merit in the description strings are from G. Heinzel et al.,
"Spectrum and spectral density estimation by the DFT" (2002), as in the reference.
"""
import numpy as np

__all__ = ["windows", "window_table"]


def _cosine_sum(coeffs):
    """w[n] = sum_m c_m cos(2 pi m n / N), n = 0..N-1 (windows.py:92-95, 222-233)."""
    def gen(n):
        z = 2.0 * np.pi * np.arange(n) / n
        w = np.full(n, coeffs[0], dtype=np.float64)
        for m in range(1, len(coeffs)):
            w += coeffs[m] * np.cos(m * z)
        return w
    return gen


def _welch(n):
    z = 2.0 * np.arange(n) / n
    return 1.0 - (z - 1.0) * (z - 1.0)


def _rect(n):
    return np.ones((n,), dtype=np.float64)


_FLAT = "Fast-decaying Flattop"
_MINSL = "Minimum sidelobe Flattop"
_BH3 = "3-term Blackman-Harris type"
_BH4 = "4-term Blackman-Harris type"

# (suffix tested inside the nuttall/flattop/sft family, title, figures of merit, coefficients, ROV)
# order = order of the tests at windows.py:101-217
_NUTTALL_FAMILY = (
    ("3f", _FLAT, "(SLDR~f**-3, PSLL=-31.7dB, ROV=66.7%, AF=0.998, PF=0.558, OC=-0.029)",
     (0.26526, -0.5, 0.23474), 0.667),
    ("4f", _FLAT, "(SLDR~f**-5, PSLL=-44.7dB, ROV=75.0%, AF=1.000, PF=0.647, OC=0.039)",
     (0.21706, -0.42103, 0.28294, -0.07897), 0.75),
    ("5f", _FLAT, "(SLDR~f**-7, PSLL=-57.3dB, ROV=78.5%, AF=0.969, PF=0.648, OC=0.052)",
     (0.1881, -0.36923, 0.28702, -0.13077, 0.02488), 0.785),
    ("3m", _MINSL, "(SLDR~f**-1, PSLL=-44.2dB, ROV=65.5%, AF=0.949, PF=0.584, OC=-0.005)",
     (0.28235, -0.52105, 0.19659), 0.655),
    ("4m", _MINSL, "(SLDR~f**-1, PSLL=-66.5dB, ROV=72.1%, AF=0.964, PF=0.641, OC=0.044)",
     (0.241906, -0.460841, 0.255381, -0.041872), 0.721),
    ("5m", _MINSL, "(SLDR~f**-1, PSLL=-89.9dB, ROV=76.0%, AF=0.953, PF=0.645, OC=0.053)",
     (0.209671, -0.407331, 0.281225, -0.092669, 0.0091036), 0.760),
    ("3a", _BH3, "(SLDR~f**-3, PSLL=-64.2dB, ROV=61.2%, AF=0.943, PF=0.723, OC=0.227)",
     (0.40897, -0.5, 0.09103), 0.612),
    ("3b", _BH3, "(SLDR~f**-1, PSLL=-71.5dB, ROV=59.8%, AF=0.939, PF=0.721, OC=0.229)",
     (0.4243801, -0.4973406, 0.0782793), 0.598),
    ("3", _BH3, "(SLDR~f**-5, PSLL=-46.7dB, ROV=64.7%, AF=0.969, PF=0.738, OC=0.228)",
     (0.375, -0.5, 0.125), 0.647),
    ("4a", _BH4, "(SLDR~f**-5, PSLL=-82.6dB, ROV=68.0%, AF=0.931, PF=0.721, OC=0.234)",
     (0.338946, -0.481973, 0.161054, -0.018027), 0.68),
    ("4b", _BH4, "(SLDR~f**-3, PSLL=-93.3dB, ROV=66.3%, AF=0.924, PF=0.715, OC=0.233)",
     (0.355768, -0.487396, 0.144232, -0.012604), 0.663),
    ("4c", _BH4, "(SLDR~f**-1, PSLL=-98.1dB, ROV=65.6%, AF=0.923, PF=0.716, OC=0.235)",
     (0.3635819, -0.4891775, 0.1365995, -0.0106411), 0.656),
    ("4", _BH4, "(SLDR~f**-7, PSLL=-60.9dB, ROV=70.5%, AF=0.937, PF=0.723, OC=0.233)",
     (0.3125, -0.46875, 0.1875, -0.03125), 0.705),
)


def window_table(windowfunction, beta=None):
    """Resolve a window name -> (title, figures-of-merit string, generator, ROV)."""
    key = windowfunction.lower()
    if "hann" in key:
        return ("Hanning", "(SLDR~f**-3, PSLL=-31.5dB, ROV=50.0%, AF=1.000, PF=0.707, OC=0.167)",
                np.hanning, 0.50)
    if "hamm" in key:
        return ("Hamming", "(SLDR~f**-1, PSLL=-42.7dB, ROV=50.0%, AF=1.000, PF=0.761, OC=0.234)",
                np.hamming, 0.50)
    if "black" in key:
        return ("Blackman-Harris type",
                "(SLDR~f**-1, PSLL=-92dB, ROV=66.1%, AF=0.926, PF=0.718, OC=0.235)",
                _cosine_sum((0.35875, -0.48829, 0.14128, -0.01168)), 0.661)
    if "nut" in key or "flat" in key or "sft" in key:
        for suffix, title, fom, coeffs, rov in _NUTTALL_FAMILY:
            if suffix in key:
                return title, fom, _cosine_sum(coeffs), rov
        # the reference falls through with its coefficient tuple unbound (windows.py:98-233)
        raise NameError("window family %r needs a 3/4/5 term suffix (e.g. 'SFT3F', 'nuttall4a')"
                        % (windowfunction,))
    if "kaiser" in key:
        if beta is None:
            raise KeyError("beta")                       # windows.py:236
        return ("Kaiser type",
                "(parameters dependent on input shaping parameter %4.3f)" % (beta,),
                (lambda n: np.kaiser(n, beta)), 2.0 / 3.0)
    if "welch" in key:
        return ("Welch", "(SLDR~f**-2, PSLL=-21.3dB, ROV=29.3%, AF=0.828, PF=0.707, OC=0.091)",
                _welch, 0.293)
    if "bart" in key:
        return ("Bartlett", "(SLDR~f**-2, PSLL=-26.5dB, ROV=50.0%, AF=1.000, PF=0.707, OC=0.250)",
                np.bartlett, 0.50)
    return ("Rectangular", "(SLDR~f**-1, PSLL=-13.3dB, ROV=0.0%, AF=0, PF=1, OC=0)", _rect, 0.0)


def windows(windowfunction, **kwargs):
    """windows(name, nwins=None, periodic=True, verbose=True, msgout=False, beta=None)

    With ``nwins``: the float64 window vector (periodic by default, i.e. the
    (nwins+1)-point window with its last point dropped).  Without: the window's
    recommended overlap fraction.  ``msgout=True`` also returns (title, figures).
    """
    verbose = kwargs.setdefault("verbose", True)
    periodic = kwargs.setdefault("periodic", True)
    msgout = kwargs.setdefault("msgout", False)
    title, fom, gen, rov = window_table(windowfunction, kwargs.get("beta"))
    if "nwins" in kwargs:
        nwins = kwargs["nwins"]
        if periodic:
            val = np.asarray(gen(nwins + 1)[:-1], dtype=np.float64)
            kind = "periodic"
        else:
            val = np.asarray(gen(nwins), dtype=np.float64)
            kind = "aperiodic"
        msg = "Using a %s %s window function\n%s" % (kind, title, fom)
    else:
        val = rov
        msg = "Getting recommended overlap for a %s window function" % (title,)
    if verbose:
        print(msg)
    if msgout:
        return val, (title, fom)
    return val
