"""CPU: the C-ABI library loads and exports every symbol include/*.h declares
(no compute calls -- there is no GPU here)."""
import ctypes
import glob
import os
import re

from conftest import ROOT


def declared_symbols():
    names = set()
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        src = open(h).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names.update(re.findall(r"\b(pyfft_[a-z0-9_]+)\s*\(", src))
    return sorted(names)


def test_header_declares_the_expected_entry_points():
    syms = declared_symbols()
    for must in ("pyfft_welch_accumulate", "pyfft_welch_finalize", "pyfft_detrend_stats",
                 "pyfft_stft", "pyfft_hilbert", "pyfft_xcorr", "pyfft_csm_accumulate",
                 "pyfft_welch_correlation", "pyfft_last_error"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(os.path.join(ROOT, "pyfft_b200", "libpyfft_b200.so"))
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, missing


def test_python_binding_covers_the_header():
    from pyfft_b200 import _lib
    assert sorted(_lib.EXPORTS) == declared_symbols()
    assert "sm_100a" in _lib.version()


def test_status_codes_map_to_exceptions():
    """Argument validation happens before any CUDA call, so it can be exercised without a GPU."""
    import pytest
    from pyfft_b200 import _lib
    seg = _lib.Segmentation(1000, 500, 0, 4)          # not a power of two
    with pytest.raises(NotImplementedError):
        _lib.check(_lib.lib.pyfft_welch_workspace_bytes(_lib.F32, ctypes.byref(seg), 1))
    seg = _lib.Segmentation(1024, 0, 0, 4)            # hop 0
    with pytest.raises(ValueError):
        _lib.check(_lib.lib.pyfft_welch_workspace_bytes(_lib.F32, ctypes.byref(seg), 1))
    assert "hop" in _lib.last_error()


def test_library_is_built_for_sm_100a_only():
    import subprocess
    so = os.path.join(ROOT, "pyfft_b200", "libpyfft_b200.so")
    out = subprocess.run(["cuobjdump", "-lelf", so], stdout=subprocess.PIPE, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs
