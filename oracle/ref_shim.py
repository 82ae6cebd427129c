"""TEST INFRASTRUCTURE ONLY -- import shim for the unmodified reference.

Loads ``/root/reference/{windows,fft_analysis,spectrogram,hilbert}.py`` as the
package ``FFT`` without touching them.  The reference hard-imports three things
that are absent here:

* ``matplotlib`` (fft_analysis.py:19-20)             -> empty stub modules
* ``pybaseutils`` (fft_analysis.py:22-24), un-vendored -> ``Struct`` attribute
  bag and ``detrend_none/mean/linear`` restated (PARITY UNPINNED: matplotlib.mlab
  semantics, the ones the reference's own mlab branch passes the same callables
  to, fft_analysis.py:288-290)
* ``numpy.deprecate`` (windows.py:1052), removed in NumPy 2 -> no-op decorator

``ccf.py`` / ``dft.py`` cannot be loaded on Python 3 (dft.py:200 is a py2 print
statement), so they are not offered here.

``/root/reference`` exists in the build container only.  ``__graft_entry__.build()`` places
unmodified copies of the four modules under ``baseline/_ref/`` (git-ignored, it travels to the
GPU box like the built ``.so``); ``bench.py --impl reference`` and the ``cpu_baseline`` leg
load them from there (``load_reference(ref_dir)``) to time the stock CPU path.  Nothing in
``-m gpu`` tests or ``smoke()`` reads either directory.
"""
import importlib.util
import os
import sys
import types

import numpy as np

REFERENCE_DIR = os.environ.get("PYFFT_REFERENCE_DIR", "/root/reference")


MODULES = ("windows", "fft_analysis", "spectrogram", "hilbert")


def available(ref_dir=None):
    d = REFERENCE_DIR if ref_dir is None else ref_dir
    return all(os.path.isfile(os.path.join(d, m + ".py")) for m in MODULES)


class _Struct(object):
    """Stand-in for pybaseutils.Struct.Struct (fft_analysis.py:22, 796, 1785-1790)."""

    def __init__(self, d=None):
        if d:
            self.__dict__.update(d)

    def dict_from_class(self):
        return dict(self.__dict__)


def _detrend_none(x, axis=None):
    return x


def _detrend_mean(x, axis=None):
    x = np.asarray(x)
    if axis is None:
        axis = 0
    return x - x.mean(axis, keepdims=True)


def _detrend_linear(y, axis=None):
    y = np.asarray(y)
    if axis is None:
        axis = 0
    n = y.shape[axis]
    t = np.arange(n, dtype=float)
    t -= t.mean()
    sh = [1] * y.ndim
    sh[axis] = n
    t = t.reshape(sh)
    ym = y.mean(axis, keepdims=True)
    slope = (t * (y - ym)).sum(axis, keepdims=True) / (t * t).sum()
    return y - ym - slope * t


_loaded = {}


def load_reference(ref_dir=None):
    """Return a dict of the reference modules {name: module} (first call decides the directory)."""
    if _loaded:
        return _loaded
    ref_dir = REFERENCE_DIR if ref_dir is None else ref_dir
    if not available(ref_dir):
        raise RuntimeError("reference tree not present at %s" % ref_dir)
    if not hasattr(np, "deprecate"):
        np.deprecate = lambda *a, **k: (lambda f: f)
    saved = {}
    for name in ("matplotlib", "matplotlib.mlab", "matplotlib.pyplot"):
        saved[name] = sys.modules.get(name)
        sys.modules[name] = types.ModuleType(name)
    sys.modules["matplotlib"].mlab = sys.modules["matplotlib.mlab"]
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    pb = types.ModuleType("pybaseutils")
    st = types.ModuleType("pybaseutils.Struct")
    ut = types.ModuleType("pybaseutils.utils")
    st.Struct = _Struct
    ut.detrend_none = _detrend_none
    ut.detrend_mean = _detrend_mean
    ut.detrend_linear = _detrend_linear
    pb.Struct, pb.utils = st, ut
    sys.modules.update({"pybaseutils": pb, "pybaseutils.Struct": st,
                        "pybaseutils.utils": ut})
    pkg = types.ModuleType("FFT")
    pkg.__path__ = [ref_dir]
    sys.modules["FFT"] = pkg
    for name in MODULES:
        spec = importlib.util.spec_from_file_location(
            "FFT." + name, os.path.join(ref_dir, name + ".py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules["FFT." + name] = mod
        spec.loader.exec_module(mod)
        _loaded[name] = mod
    return _loaded
