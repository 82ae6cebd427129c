"""GPU parity: fft_pwelch on the sm_100a kernels (through the C ABI) vs the golden
fixtures written by the unmodified reference and vs the oracle on seeded inputs.

Tolerances (BASELINE.json north_star):
  fp32: max |P - P_ref| <= 1e-5 * max |P_ref| for PSD/CSD, <= 1e-4 absolute for coherence
  fp64: <= 1e-12 relative to the spectral peak
"""
import json

import numpy as np
import pytest

from conftest import UNIT_SCALE, load_golden, phase_err, relerr

pytestmark = pytest.mark.gpu

POW2_CASES = ["welch_c1_auto", "welch_c1_bincentred", "welch_c2_cross", "welch_c2_dc",
              "welch_tbounds_tper"]
TOL = {"fp64": dict(P=1e-12, C=1e-10, R=1e-11), "fp32": dict(P=1e-5, C=1e-4, R=2e-5)}


def _run(g, precision):
    from pyfft_b200 import fft_pwelch
    kw = json.loads(str(g["kwargs_json"]))
    tb = kw.pop("tbounds")
    x, y = g["in_x"], g["in_y"]
    if precision == "fp32":
        x, y = x.astype(np.float32), y.astype(np.float32)
    return fft_pwelch(g["in_t"], x, y, tb, plotit=False, precision=precision, **kw)


def _check(out, ref, precision, reference_is_fp32_rounded=False):
    tol = TOL[precision]
    freq, Pxy, Pxx, Pyy, Cxy, phi, info = out
    assert relerr(freq, ref["freq"]) < 1e-14
    for nm, a in (("Pxy", Pxy), ("Pxx", Pxx), ("Pyy", Pyy)):
        assert a.dtype == np.complex128
        assert relerr(a, ref[nm]) <= tol["P"], nm
    assert relerr(Cxy, ref["Cxy"], 1.0) <= tol["C"]
    assert relerr(info.Cxy2, ref["Cxy2"], 1.0) <= tol["C"]
    assert phase_err(phi, ref["phi_xy"], ref["Pxy"]) <= 10 * tol["P"]
    for nm in ("Rxx", "Ryy", "Rxy", "Lxx", "Lyy", "Lxy", "Ex", "Ey"):
        assert relerr(getattr(info, nm), ref[nm]) <= tol["R"], nm
    for nm in ("iCxy", "corrcoef"):
        assert relerr(getattr(info, nm), ref[nm], 1.0) <= 10 * tol["C"], nm
    assert relerr(info.lags, ref["lags"]) < 1e-14
    for nm in ("nwins", "noverlap", "Navr", "nch"):
        assert int(getattr(info, nm)) == int(ref[nm])
    for nm in ("S1", "S2", "ENBW", "NENBW", "minFreq"):
        assert float(getattr(info, nm)) == float(ref[nm])


@pytest.mark.parametrize("case", POW2_CASES)
def test_fp64_matches_reference_golden(case, cuda_device):
    g = load_golden(case)
    _check(_run(g, "fp64"), g, "fp64")


@pytest.mark.parametrize("case", POW2_CASES)
def test_fp32_matches_oracle_on_rounded_inputs(case, cuda_device):
    from oracle import restatement as R
    g = load_golden(case)
    kw = json.loads(str(g["kwargs_json"]))
    tb = kw.pop("tbounds")
    x32, y32 = g["in_x"].astype(np.float32), g["in_y"].astype(np.float32)
    with np.errstate(all="ignore"):
        ref = R.fft_pwelch(g["in_t"], x32.astype(np.float64), y32.astype(np.float64), tb, **kw)
    _check(_run(g, "fp32"), ref, "fp32")


ANY_CASES = ["welch_multi_sft3f_linear", "welch_defaults", "welch_odd_nfft", "welch_ref_selftest",
             "welch_twosided_complex"]


@pytest.mark.parametrize("case", ANY_CASES)
def test_arbitrary_window_lengths_and_complex_input_match_reference_golden(case, cuda_device):
    """what the reference does when nobody hand-picks Navr: nwins = 3640, 4289 (three y channels, linear
    detrend), 2374 with a 30 % overlap, its own self-test recipe, and a complex record with two-sided
    spectra -- float64 against the outputs of the unmodified reference."""
    g = load_golden(case)
    _check(_run(g, "fp64"), g, "fp64")


@pytest.mark.parametrize("case", [c for c in ANY_CASES if c != "welch_ref_selftest"])
def test_arbitrary_window_lengths_fp32_inputs_vs_oracle(case, cuda_device):
    from oracle import restatement as R
    g = load_golden(case)
    kw = json.loads(str(g["kwargs_json"]))
    tb = kw.pop("tbounds")
    cplx = np.iscomplexobj(g["in_x"])
    x32 = g["in_x"].astype(np.complex64 if cplx else np.float32)
    y32 = g["in_y"].astype(np.float32)
    with np.errstate(all="ignore"):
        ref = R.fft_pwelch(g["in_t"], x32.astype(np.complex128 if cplx else np.float64), y32.astype(np.float64), tb, **kw)
    from pyfft_b200 import fft_pwelch
    out = fft_pwelch(g["in_t"], x32, y32, tb, plotit=False, precision="fp32", **kw)
    _check(out, ref, "fp32")


@pytest.mark.parametrize("log2l", [4, 5, 6, 7, 8, 9, 10, 11, 12, 13])
@pytest.mark.parametrize("precision", ["fp32", "fp64"])
def test_every_window_length_vs_oracle(log2l, precision, cuda_device):
    from oracle import restatement as R
    from pyfft_b200 import fft_pwelch
    L = 1 << log2l
    rng = np.random.default_rng(100 + log2l)
    K, nch = 9, 2 if log2l % 2 else 3
    n = L // 2 * (K + 1) + 5                           # a few unused tail samples
    t = np.arange(n) / 1e6
    x = np.sin(2 * np.pi * 0.11 * np.arange(n)) + rng.standard_normal(n) + 2.0
    y = np.stack([np.cos(2 * np.pi * 0.11 * np.arange(n) + c) + rng.standard_normal(n) - c
                  for c in range(nch)], 1)
    if precision == "fp32":
        x, y = x.astype(np.float32), y.astype(np.float32)
    det = (1, -1, 0)[log2l % 3]
    tb = [t[3], t[-1]]                                 # no reflect padding here
    kw = dict(windowoverlap=0.5, windowfunction="Hanning", detrend_style=det,
              tper=L / 1e6 * (1 + 1e-9))
    with np.errstate(all="ignore"):
        ref = R.fft_pwelch(t, x.astype(np.float64), y.astype(np.float64), tb, **kw)
    assert ref["nwins"] == L
    out = fft_pwelch(t, x, y, tb, plotit=False, precision=precision, **kw)
    _check(out, ref, precision)


@pytest.mark.parametrize("precision", ["fp32", "fp64"])
def test_two_sided_real_input_and_reflect(precision, cuda_device):
    from oracle import restatement as R
    from pyfft_b200 import fft_pwelch
    rng = np.random.default_rng(42)
    n = 2 ** 14
    t = np.arange(n) / 1e6
    x = rng.standard_normal(n) + 0.5
    y = np.roll(x, 3) + 0.3 * rng.standard_normal(n)
    if precision == "fp32":
        x, y = x.astype(np.float32), y.astype(np.float32)
    kw = dict(Navr=31, windowoverlap=0.5, windowfunction="Hamming", detrend_style=1, onesided=False)
    with np.errstate(all="ignore"):
        ref = R.fft_pwelch(t, x.astype(np.float64), y.astype(np.float64), **kw)
    assert ref["reflect"] and ref["Pxx"].shape == (1024,)
    _check(fft_pwelch(t, x, y, plotit=False, precision=precision, **kw), ref, precision)


def test_device_tensor_inputs_and_identical_signals(cuda_device):
    import torch
    from pyfft_b200 import fft_pwelch
    n = 2 ** 16
    t = np.arange(n) / 1e6
    g = torch.Generator(device=cuda_device).manual_seed(5)
    x = torch.randn(n, generator=g, device=cuda_device, dtype=torch.float32)
    f, Pxy, Pxx, Pyy, Cxy, phi, info = fft_pwelch(t, x, x, Navr=63, windowoverlap=0.5, plotit=False)
    assert info.precision == "fp32" and info.nwins == 2048
    assert np.allclose(np.abs(Cxy), 1.0, atol=1e-5) and np.allclose(phi, 0.0, atol=1e-5)
    assert relerr(Pxx, Pyy) < 1e-6 and relerr(Pxy.real, Pxx.real) < 1e-6


@pytest.mark.parametrize("precision", ["fp32", "fp64"])
@pytest.mark.parametrize("window", ["Hanning", "Hamming", "SFT3F", "Bartlett"])
def test_fused_mean_route_equals_two_pass_and_oracle(precision, window, cuda_device):
    """Mean detrend fused into K1 (provisional mean + exact linear correction): a drifting
    record makes the provisional mean wrong by ~sigma; results must still match the oracle and
    the two-pass (K0 + K1) route.  Windows whose transform leaks beyond nfft/16 bins fall back
    to two passes by themselves."""
    from oracle import restatement as R
    from oracle import signals
    from pyfft_b200 import fft_pwelch
    n = 2 ** 17
    t, x, y = signals.c2_cross(n=n, dc=(50.0, -20.0))
    x = x + np.linspace(0.0, 2.0, n)
    y = y - np.linspace(0.0, 1.0, n) ** 2
    if precision == "fp32":
        x, y = x.astype(np.float32), y.astype(np.float32)
    kw = dict(Navr=63, windowoverlap=0.5, windowfunction=window, detrend_style=1)
    with np.errstate(all="ignore"):
        ref = R.fft_pwelch(t, x.astype(np.float64), y.astype(np.float64), **kw)
    a = fft_pwelch(t, x, y, plotit=False, precision=precision, **kw)
    b = fft_pwelch(t, x, y, plotit=False, precision=precision, two_pass_detrend=True, **kw)
    _check(a, ref, precision)
    _check(b, ref, precision)


@pytest.mark.parametrize("tensor_cores,navr", [(True, 2047), (False, 2047), (None, 1023)])
def test_pinned_host_records_are_uploaded_in_pieces(tensor_cores, navr, cuda_device):
    """Long pinned host records take the staged upload (copy in pieces on a side stream, the complete
    segments of each piece accumulated behind it): same spectra as with device-resident inputs, for
    the tensor-core kernel, the CUDA-core kernel and a two-pass case that has to wait for all pieces."""
    import torch
    from oracle import signals
    from pyfft_b200 import fft_pwelch
    n = 2 ** 22
    t, x, y = signals.c2_cross(n=n, dc=(7.0, 2.5))
    xs, ys = torch.from_numpy(x.astype(np.float32)).pin_memory(), torch.from_numpy(y.astype(np.float32)).pin_memory()
    kw = dict(Navr=navr, windowoverlap=0.5, windowfunction="Hanning", detrend_style=1, plotit=False,
              precision="fp32")
    if tensor_cores is None:
        kw["two_pass_detrend"] = True
    else:
        kw["tensor_cores"] = tensor_cores
    a = fft_pwelch(t, xs, ys, **kw)
    b = fft_pwelch(t, xs.to(cuda_device), ys.to(cuda_device), **kw)
    # the staged route takes its provisional mean / fp16 prescale from the first 2^20 samples (what has
    # arrived), the resident route from the whole record: on the fp32 kernels that cancels exactly; on the
    # forced tensor-core kernel (K = 2047, below the automatic threshold) the two calls round to fp16
    # differently and differ by that kernel's noise at this K (6e-4 / sqrt(K) of peak, coherence 10 x)
    tol = 6e-4 / np.sqrt(navr) * 1.5 if tensor_cores else 2e-6
    for i in (1, 2, 3):
        assert relerr(a[i], b[i]) <= tol, i
    assert np.max(np.abs(a[4] - b[4])) <= (10 * tol if tensor_cores else 2e-6)
    assert relerr(a[6].Rxy, b[6].Rxy) <= tol
    # a (1, n) second record goes the same way
    c = fft_pwelch(t, xs, ys.view(1, n), **kw)
    assert relerr(c[1], a[1]) <= 1e-12


@pytest.mark.parametrize("tensor_cores", [True, False])
def test_pinned_host_records_with_tbounds_in_the_second_half(tensor_cores, cuda_device):
    """staged upload + an analysis window that starts far behind the first piece: the provisional statistics
    must wait for the piece that holds the samples they read (they used to read memory not yet uploaded)."""
    import torch
    from oracle import restatement as R
    from oracle import signals
    from pyfft_b200 import fft_pwelch
    n = 2 ** 23
    t, x, y = signals.c2_cross(n=n, dc=(7.0, 2.5))
    x32, y32 = x.astype(np.float32), y.astype(np.float32)
    tb = [t[n // 2 + 12345], t[n // 2 + 12345 + 2 ** 21 - 1]]      # 2^21 samples from the second half on
    kw = dict(Navr=1023, windowoverlap=0.5, windowfunction="Hanning", detrend_style=1)
    with np.errstate(all="ignore"):
        ref = R.fft_pwelch(t, x32.astype(np.float64), y32.astype(np.float64), tb, **kw)
    for rep in range(3):                                      # the race was nondeterministic
        xs, ys = torch.from_numpy(x32).pin_memory(), torch.from_numpy(y32).pin_memory()
        out = fft_pwelch(t, xs, ys, tb, plotit=False, precision="fp32", tensor_cores=tensor_cores, **kw)
        assert out[6].nwins == int(ref["nwins"]) == 4096 and out[6].Navr == 1023
        tol = 1.5 * 6e-4 / np.sqrt(1023) if tensor_cores else 1e-5
        for i, nm in ((1, "Pxy"), (2, "Pxx"), (3, "Pyy")):
            assert np.isfinite(out[i]).all() and relerr(out[i], ref[nm]) <= tol, (rep, nm, relerr(out[i], ref[nm]))
