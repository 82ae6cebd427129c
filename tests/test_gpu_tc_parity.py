"""GPU parity of the path the headline number is produced on: fft_pwelch with the AUTOMATIC kernel
selection (tensor_cores=None) at Navr >= 16384, against the float64 oracle on the same fp32-rounded
inputs, at the north_star bounds -- 1e-5 of the spectral peak on Pxx / Pyy / Pxy, 1e-4 absolute on
Cxy and Cxy2 -- nothing loosened.

Also the records the fp16 operands of the tensor-core kernel cannot serve, where the selection has
to be data-aware (fftinfo.tc_fallback says why the call was redone on the fp32 CUDA-core kernel):
  * noiseless records that repeat with the hop (the reference's self-test recipe, a bin-centred tone,
    hop-periodic pseudo-noise): identical roundings in every segment, nothing averages out;
  * spectra with bins far below the mean (the same, and any clean narrow-band record);
  * records whose level changes by 10^6 along the way, and a 10^6 burst between the sampled runs
    (fp16 overflow -> status bit -> rerun), never Inf / NaN.
Config 1 at its full size (2^20 samples, nfft 1024, Navr 2047, float64) is here too.
"""
import numpy as np
import pytest

from conftest import relerr

pytestmark = pytest.mark.gpu

L, H, FS = 4096, 2048, 1.0e6
N_SHORT = 2 ** 25 + 4096                     # Navr = 16385: just above the automatic threshold


def _navr(n):
    return n // H - 1


def _oracle(t, x, y, navr):
    from oracle import restatement as R
    with np.errstate(all="ignore"):
        return R.fft_pwelch(t, x.astype(np.float64), y.astype(np.float64), Navr=navr, windowoverlap=0.5,
                            windowfunction="Hanning", detrend_style=1)


def _ours(t, x, y, navr, **kw):
    from pyfft_b200 import fft_pwelch
    return fft_pwelch(t, x, y, Navr=navr, windowoverlap=0.5, windowfunction="Hanning", detrend_style=1,
                      plotit=False, precision="fp32", **kw)


def _assert_contract(out, ref, coh_mask=None):
    """north_star bounds.  coh_mask: bins on which the coherence is defined (both auto-spectra above the
    arithmetic's own noise floor) -- None = every bin."""
    f, Pxy, Pxx, Pyy, Cxy, phi, info = out
    assert info.nwins == L
    for a, nm in ((Pxy, "Pxy"), (Pxx, "Pxx"), (Pyy, "Pyy")):
        assert np.isfinite(a).all(), nm
        assert relerr(a, ref[nm]) <= 1e-5, (nm, relerr(a, ref[nm]))
    m = slice(None) if coh_mask is None else coh_mask
    assert np.isfinite(Cxy[m]).all()
    assert np.max(np.abs(Cxy[m] - ref["Cxy"][m])) <= 1e-4, np.max(np.abs(Cxy[m] - ref["Cxy"][m]))
    assert np.max(np.abs(info.Cxy2[m] - ref["Cxy2"][m])) <= 1e-4


def _sine_noise(n, seed, dc=(7.0, 2.5)):
    rng = np.random.default_rng(seed)
    ph = 2 * np.pi * (400.0 / L) * np.arange(n)
    x = (np.sin(ph) + rng.standard_normal(n) + dc[0]).astype(np.float32)
    y = (0.5 * np.sin(ph - np.pi / 4) + rng.standard_normal(n) + dc[1]).astype(np.float32)
    return np.arange(n) / FS, x, y


@pytest.mark.parametrize("n", [N_SHORT, 2 ** 26])
def test_auto_selected_tensor_core_path_vs_oracle(n, cuda_device):
    """config 2's record (sine + unit white noise + DC) -- at 2^26 this IS the bench workload."""
    t, x, y = _sine_noise(n, 2)
    out = _ours(t, x, y, _navr(n))
    info = out[6]
    assert info.Navr == _navr(n) >= 16384
    assert info.tensor_cores and info.tc_fallback is None, info.tc_fallback
    assert info.tc_checks["repetition"] > 0.3 and info.tc_checks["min_over_mean"] > 0.3
    _assert_contract(out, _oracle(t, x, y, _navr(n)))


def test_coloured_noise_with_30_dB_of_dynamic_range_stays_on_tensor_cores(cuda_device):
    """a spectrum that is not flat (first-order low-pass noise, ~30 dB from DC to Nyquist, a weak coherent
    part): the tensor-core path must hold the contract in the quiet bins too."""
    n = N_SHORT
    rng = np.random.default_rng(5)
    from scipy.signal import lfilter
    a = 0.94
    s = lfilter([1.0 - a], [1.0, -a], rng.standard_normal(n))
    x = (s + 3.0).astype(np.float32)
    y = (0.3 * np.roll(s, 5) + 0.05 * rng.standard_normal(n) - 1.0).astype(np.float32)
    t = np.arange(n) / FS
    out = _ours(t, x, y, _navr(n))
    assert out[6].tensor_cores, out[6].tc_fallback
    _assert_contract(out, _oracle(t, x, y, _navr(n)))


def _selftest_recipe(n):
    """fft_analysis.py:2899-2944 (square + DC vs phase-shifted sine, NO noise), resampled so that the
    period (64 samples) divides the hop."""
    from scipy.signal import square
    ph = 2 * np.pi * np.arange(n) / 64.0
    x = (0.1 * square(ph) + 7.0).astype(np.float32)
    y = (0.007 * np.sin(ph - np.pi / 4) + 2.5).astype(np.float32)
    return np.arange(n) / FS, x, y


def _bin_centred_tone(n):
    ph = 2 * np.pi * (400.0 / L) * np.arange(n)
    return np.arange(n) / FS, np.sin(ph).astype(np.float32), (0.5 * np.sin(ph - 0.7)).astype(np.float32)


def _hop_periodic_noise(n):
    """pseudo-noise that repeats every hop: a full, flat spectrum and identical segments."""
    rng = np.random.default_rng(9)
    bx, by = rng.standard_normal(H), rng.standard_normal(H)
    x = np.tile(bx, n // H + 1)[:n].astype(np.float32)
    y = (0.5 * np.tile(bx, n // H + 1)[:n] + np.tile(by, n // H + 1)[:n]).astype(np.float32)
    return np.arange(n) / FS, x, y


@pytest.mark.parametrize("make", [_selftest_recipe, _bin_centred_tone, _hop_periodic_noise])
def test_noiseless_periodic_records_are_detected_and_meet_the_contract(make, cuda_device):
    n = N_SHORT
    t, x, y = make(n)
    out = _ours(t, x, y, _navr(n))
    info = out[6]
    assert not info.tensor_cores and info.tc_fallback, "selection must be data-aware"
    assert info.tc_checks["repetition"] < 1e-4
    ref = _oracle(t, x, y, _navr(n))
    # coherence is 0/0 where both spectra are rounding noise: compare it where the record has power
    pk = max(np.max(np.abs(ref["Pxx"])), np.max(np.abs(ref["Pyy"])))
    mask = (np.abs(ref["Pxx"]) > 1e-6 * np.max(np.abs(ref["Pxx"]))) & (np.abs(ref["Pyy"]) > 1e-6 * np.max(np.abs(ref["Pyy"])))
    assert mask.any() and pk > 0
    _assert_contract(out, ref, coh_mask=mask)


def test_forced_tensor_cores_on_a_repeating_record_shows_why_the_check_exists(cuda_device):
    """documentation of the failure mode, not a contract: with the selection overridden the hop-periodic
    record comes out an order of magnitude outside 1e-5 (identical roundings in every segment)."""
    n = N_SHORT
    t, x, y = _hop_periodic_noise(n)
    out = _ours(t, x, y, _navr(n), tensor_cores=True)
    assert out[6].tensor_cores and "forced" in (out[6].tc_fallback or "")
    ref = _oracle(t, x, y, _navr(n))
    err = relerr(out[2], ref["Pxx"])
    assert np.isfinite(err) and 1e-5 < err < 1e-2, err


@pytest.mark.parametrize("factor", [1.0e6, 1.0e-6])
def test_level_step_of_1e6_after_the_head(factor, cuda_device):
    """x 10^6 (and / 10^6) from sample 65 536 on: the prescale comes from runs spread over the whole
    record, so the loud part is in range; the quiet part underflows to fp16 subnormals but carries
    10^-12 of the power.  Either route is fine -- the contract is not."""
    n = N_SHORT
    t, x, y = _sine_noise(n, 11, dc=(0.5, -0.25))
    g = np.ones(n, dtype=np.float32)
    g[65536:] = np.float32(factor)
    x, y = x * g, y * g
    out = _ours(t, x, y, _navr(n))
    _assert_contract(out, _oracle(t, x, y, _navr(n)))


def test_burst_between_the_sampled_runs_overflows_fp16_and_is_rerun(cuda_device):
    n = N_SHORT
    t, x, y = _sine_noise(n, 12)
    # a glitch common to both channels, not in any of the 256 sampled runs of 256 samples (stride ~ 131.6 k samples)
    x[100000:100512] *= np.float32(3.0e6)
    y[100000:100512] *= np.float32(3.0e6)
    out = _ours(t, x, y, _navr(n))
    info = out[6]
    assert not info.tensor_cores and "overflow" in info.tc_fallback
    _assert_contract(out, _oracle(t, x, y, _navr(n)))


def test_config1_at_full_size_fp64(cuda_device):
    """BASELINE.json configs[0]: 2^20 samples, nfft 1024, Hanning, 50 %, Navr 2047, float64, <= 1e-12."""
    from oracle import restatement as R
    from oracle import signals
    from pyfft_b200 import fft_pwelch
    for centred in (False, True):
        t, x = signals.c1_auto(n=2 ** 20, bin_centred=centred)
        kw = dict(Navr=2047, windowoverlap=0.5, windowfunction="Hanning", detrend_style=1)
        ref = R.fft_pwelch(t, x, x, **kw)
        f, Pxy, Pxx, Pyy, Cxy, phi, info = fft_pwelch(t, x, x, plotit=False, **kw)
        assert info.nwins == 1024 and info.Navr == 2047 and info.precision == "fp64"
        for a, nm in ((Pxy, "Pxy"), (Pxx, "Pxx"), (Pyy, "Pyy")):
            assert relerr(a, ref[nm]) <= 1e-12, nm
        assert np.max(np.abs(Cxy - ref["Cxy"])) <= 1e-10
