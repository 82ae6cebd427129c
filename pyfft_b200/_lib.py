"""ctypes binding of libpyfft_b200.so (the C ABI in include/pyfft_b200.h).

There is no fallback: if the shared library is missing or fails to load, importing
this module raises.  Build it with ``python pyfft_b200/build.py`` (or
``__graft_entry__.build()``).
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PYFFT_B200_LIB") or os.path.join(HERE, "libpyfft_b200.so")   # override: A/B builds (tools/)

PYFFT_OK, PYFFT_EINVAL, PYFFT_EUNSUPPORTED, PYFFT_ECUDA = 0, -1, -2, -3
F32, F64 = 0, 1
DETREND_NONE, DETREND_MEAN, DETREND_LINEAR = 0, 1, 2


class RecordMap(C.Structure):
    """pyfft_record_map"""
    _fields_ = [("nsig", C.c_int64), ("i0", C.c_int64), ("reflect", C.c_int64),
                ("src_base", C.c_int64), ("src_len", C.c_int64)]


class Segmentation(C.Structure):
    """pyfft_segmentation"""
    _fields_ = [("nfft", C.c_int32), ("hop", C.c_int32), ("seg_first", C.c_int64),
                ("seg_count", C.c_int64)]


if not os.path.isfile(LIB_PATH):
    raise ImportError(
        "pyfft_b200: %s is missing -- the CUDA extension has not been built. Run "
        "`python pyfft_b200/build.py` (needs nvcc). There is no CPU fallback." % LIB_PATH)

lib = C.CDLL(LIB_PATH)

_vp, _i32, _i64, _dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_double
_pmap, _pseg = C.POINTER(RecordMap), C.POINTER(Segmentation)

_PROTOS = {
    "pyfft_version": (C.c_char_p, []),
    "pyfft_last_error": (C.c_char_p, []),
    "pyfft_shutdown": (C.c_int, []),
    "pyfft_sm_count": (C.c_int, []),
    "pyfft_detrend_stats": (C.c_int, [C.c_int, _vp, _i64, _i32, _pmap, _i64, _i64, _vp, _vp]),
    "pyfft_detrend_stats_xy": (C.c_int, [C.c_int, _vp, _vp, _i64, _i32, _pmap, _i64, _i64, _vp, _vp]),
    "pyfft_detrend_coef": (C.c_int, [C.c_int, _vp, _i32, _i64, _vp, _vp]),
    "pyfft_welch_workspace_bytes": (_i64, [C.c_int, _pseg, _i32]),
    "pyfft_welch_accumulate": (C.c_int, [C.c_int, _vp, _vp, _i64, _i32, _pmap, _pseg, _vp, _vp, _vp,
                                         _vp, _vp, _i64, _vp]),
    "pyfft_welch_static_scratch_bytes": (_i64, []),
    "pyfft_welch_accumulate_static": (C.c_int, [C.c_int, _vp, _vp, _i64, _i32, _pmap, _pseg, _vp, _vp, _vp, _vp, _vp]),
    "pyfft_welch_accumulate_mean": (C.c_int, [C.c_int, _vp, _vp, _i64, _i32, _pmap, _pseg, _vp, _vp, _vp,
                                              _vp, _vp, _vp, _i64, _vp]),
    "pyfft_welch_mean_correct": (C.c_int, [_vp, _vp, _i32, _i32, _i64, _vp, _vp, _i64, _vp, _vp]),
    "pyfft_welch_rereference": (C.c_int, [_vp, _vp, _i32, _i32, _i64, _vp, _vp, _vp]),
    "pyfft_welch_tc_workspace_bytes": (_i64, [_pseg, _i32]),
    "pyfft_welch_tc_prepare": (C.c_int, [_vp, _vp, _i64, _i32, _pmap, _i64, _i64, _i64, _vp, _vp, _vp]),
    "pyfft_welch_accumulate_tc": (C.c_int, [_vp, _vp, _i64, _i32, _pmap, _pseg, _vp, _vp, _vp, _vp, _vp, _vp,
                                            _vp, _i64, _vp, _vp]),
    "pyfft_welch_tc_step_peer_bytes": (_i64, [_i32, _i32]),
    "pyfft_welch_tc_step": (C.c_int, [_vp, _vp, _i64, _i32, _pmap, _pseg, _vp, _vp, _i64, _i64, _i64, _i64, _i64, _i64,
                                      _dbl, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _vp, _i32,
                                      _i32, _i32, _vp]),
    "pyfft_welch_tc_step_flush": (C.c_int, [_i32, _vp, _i64, _i64, _dbl, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp,
                                            _vp, _i32, _i32, _vp]),
    "pyfft_welch_finalize": (C.c_int, [_vp, _i32, _i32, _i64, _dbl, _i32, _vp, _vp, _vp, _vp, _vp,
                                       _vp, _vp]),
    "pyfft_welch_finalize_mean": (C.c_int, [_vp, _vp, _i32, _i32, _i64, _vp, _vp, _i64, _vp, _dbl, _i32,
                                            _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "pyfft_segment_moments_workspace_bytes": (_i64, [C.c_int, _pseg, _i32]),
    "pyfft_segment_moments": (C.c_int, [C.c_int, _vp, _vp, _i64, _i32, _pmap, _pseg, _vp, _vp, _vp, _vp, _vp, _i64,
                                        _vp]),
    "pyfft_welch_dynamic_range": (C.c_int, [_vp, _vp, _i32, _i32, _i32, _vp, _vp]),
    "pyfft_welch_correlation": (C.c_int, [_vp, _i32, _i32, _i32, _i32, _vp, _vp]),
    "pyfft_stft": (C.c_int, [C.c_int, _i32, _vp, _i64, _i32, _pmap, _pseg, _vp, _vp, _i32, _i32, _i32,
                             _vp, _i32, _vp, _vp, _vp]),
    "pyfft_find_peaks": (C.c_int, [_vp, _i32, _i32, _i32, _i32, _vp, _vp]),
    "pyfft_band_integrate": (C.c_int, [_vp, _i32, _vp, _vp, _i32, _vp, _i32, _vp, _i32, _vp, _vp]),
    "pyfft_spectral_centroid": (C.c_int, [C.c_int, _vp, _i64, _i32, _vp, _vp, _vp, _vp]),
    "pyfft_resample_linear": (C.c_int, [C.c_int, _vp, _i64, _i32, _i64, _dbl, _dbl, _i64, _vp, _i64, _vp]),
    "pyfft_mean_angle": (C.c_int, [C.c_int, _vp, _vp, _vp, _i64, _i64, _vp, _vp]),
    "pyfft_anyfft_workspace_bytes": (_i64, [_i64, _i64]),
    "pyfft_laplace_1d": (C.c_int, [C.c_int, _vp, _i64, _vp, _i32, _vp, _vp, _i64, _vp]),
    "pyfft_fft_any": (C.c_int, [C.c_int, _i32, _vp, _i64, _i64, _i32, _i32, _vp, _vp, _i64, _vp]),
    "pyfft_stft_any": (C.c_int, [C.c_int, _i32, _vp, _i64, _i32, _pmap, _pseg, _vp, _vp, _i32, _i32, _i32,
                                 _vp, _i32, _vp, _vp, _vp, _i64, _vp]),
    "pyfft_stft_products": (C.c_int, [C.c_int, _vp, _vp, _i32, _i64, _i32, _i32, _vp, _vp, _vp, _vp]),
    "pyfft_bigfft_workspace_bytes": (_i64, [C.c_int, _i64, _i32]),
    "pyfft_hilbert": (C.c_int, [C.c_int, _vp, _i64, _i32, _vp, _vp, _vp, _i64, _vp]),
    "pyfft_xcorr": (C.c_int, [C.c_int, _vp, _i64, _vp, _i64, _i64, _i32, _i32, _vp, _vp, _i64, _vp]),
    "pyfft_fft_c2c": (C.c_int, [C.c_int, _vp, _vp, _i64, _i32, _i32, _vp, _i64, _vp]),
    "pyfft_spectral_derivative": (C.c_int, [C.c_int, _vp, _i64, _i32, _i32, _dbl, _vp, _vp, _i64, _vp]),
    "pyfft_csm_workspace_bytes": (_i64, [C.c_int, _pseg, _i32]),
    "pyfft_csm_accumulate": (C.c_int, [C.c_int, _vp, _i64, _i32, _pmap, _pseg, _vp, _vp, _vp, _vp,
                                       _i64, _vp]),
    "pyfft_peer_buffer_bytes": (_i64, [_i64, _i32]),
    "pyfft_peer_allreduce": (C.c_int, [_vp, _i64, _vp, _i32, _i32, C.c_uint32, _vp, _vp, _vp]),
}

EXPORTS = tuple(_PROTOS)

for _name, (_res, _args) in _PROTOS.items():
    _fn = getattr(lib, _name)          # AttributeError here = header/library mismatch
    _fn.restype = _res
    _fn.argtypes = _args


def last_error():
    return lib.pyfft_last_error().decode("utf-8", "replace")


def check(rc):
    """Map a status code to the exception the reference's numpy code would raise."""
    if rc >= 0:
        return rc
    msg = last_error()
    if rc == PYFFT_EINVAL:
        raise ValueError(msg)
    if rc == PYFFT_EUNSUPPORTED:
        raise NotImplementedError(msg)
    raise RuntimeError(msg)


def version():
    return lib.pyfft_version().decode()
