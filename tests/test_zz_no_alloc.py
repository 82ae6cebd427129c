"""pyfft_welch_accumulate_static -- the allocation-free variant of pyfft_welch_accumulate (scratch and
twiddles in a file-scope `static __device__` array of the library) -- against the existing entry
point, through the C ABI, on two calls in a row that reuse the shared array with different shapes.

Collected last on purpose: it must pass alone in a fresh process and
after every other test.  Tolerances are the ones the Welch tests use (tests/test_gpu_welch.py): fp64
1e-12, fp32 1e-5 of the spectral peak.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _both(dev, precision, nfft, hop, n, nch, seed):
    """acc of pyfft_welch_accumulate (caller workspace) and of pyfft_welch_accumulate_static."""
    import torch
    from pyfft_b200 import _lib, engine
    from pyfft_b200._lib import Segmentation, check, lib
    from pyfft_b200.windows import windows
    rng = np.random.default_rng(seed)
    dt = torch.float64 if precision == "fp64" else torch.float32
    t = np.arange(n)
    x = np.sin(2 * np.pi * 0.0371 * t) + 0.3 * rng.standard_normal(n) + 3.0
    y = np.stack([0.5 * np.roll(x, 2 + c) + 0.2 * rng.standard_normal(n) - 1.5 for c in range(nch)])
    xd = torch.from_numpy(x).to(dev, dt)
    yd = torch.from_numpy(y).to(dev, dt).contiguous()
    win = torch.from_numpy(windows("Hanning", nwins=nfft, verbose=False)).to(dev, dt)
    K = (n - nfft) // hop + 1
    rmap = engine.record_map(n, 0, 0, 0, n)
    seg = Segmentation(nfft, hop, 0, K)
    code = _lib.F64 if precision == "fp64" else _lib.F32
    # mean-removal coefficients (a, b) per channel, x first: exercised like the two-pass route does
    coef = torch.tensor([[3.0, 0.0]] + [[-1.5, 0.0]] * nch, dtype=torch.float64, device=dev).reshape(-1)
    nbin = nfft // 2 + 1
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda a: C.c_void_p(a.data_ptr())
    acc_a = torch.zeros((nch, 4, nbin), dtype=torch.float64, device=dev)
    need = check(lib.pyfft_welch_workspace_bytes(code, C.byref(seg), nch))
    ws = torch.empty(need, dtype=torch.uint8, device=dev)
    check(lib.pyfft_welch_accumulate(code, p(xd), p(yd), yd.stride(0), nch, C.byref(rmap), C.byref(seg), p(win),
                                     p(coef), C.c_void_p(coef.data_ptr() + 16), p(acc_a), p(ws), need, st))
    acc_b = torch.zeros_like(acc_a)
    check(lib.pyfft_welch_accumulate_static(code, p(xd), p(yd), yd.stride(0), nch, C.byref(rmap), C.byref(seg),
                                            p(win), p(coef), C.c_void_p(coef.data_ptr() + 16), p(acc_b), st))
    torch.cuda.synchronize()
    return acc_a.cpu().numpy(), acc_b.cpu().numpy(), need


def _assert_close(a, b, tol):
    assert np.isfinite(b).all()
    for q in range(4):
        scale = np.max(np.abs(a[:, :3])) if q == 3 else np.max(np.abs(a[:, q]))
        assert np.max(np.abs(a[:, q] - b[:, q])) <= tol * scale, (q, np.max(np.abs(a[:, q] - b[:, q])) / scale)


def test_static_scratch_entry_matches_the_workspace_entry_twice_in_a_row(cuda_device):
    from pyfft_b200._lib import lib
    cap = lib.pyfft_welch_static_scratch_bytes()
    assert cap == 16 << 20
    # call 1: fp64, nfft 1024, two y channels; call 2 (same array, other shape and type): fp32, nfft 4096
    a, b, need = _both(cuda_device, "fp64", 1024, 512, 2 ** 17, 2, 1)
    assert need <= cap
    _assert_close(a, b, 1e-12)
    a, b, need = _both(cuda_device, "fp32", 4096, 2048, 2 ** 20, 1, 2)
    assert need <= cap
    _assert_close(a, b, 1e-5)
    # and once more with the first shape: stale contents of the larger call must not matter
    a, b, _ = _both(cuda_device, "fp64", 1024, 512, 2 ** 17, 2, 3)
    _assert_close(a, b, 1e-12)


def test_static_scratch_entry_rejects_what_does_not_fit(cuda_device):
    import torch
    from pyfft_b200 import _lib, engine
    from pyfft_b200._lib import Segmentation, lib
    nfft, hop, n, nch = 8192, 4096, 2 ** 23, 2
    seg = Segmentation(nfft, hop, 0, (n - nfft) // hop + 1)
    assert lib.pyfft_welch_workspace_bytes(_lib.F64, C.byref(seg), nch) > lib.pyfft_welch_static_scratch_bytes()
    x = torch.zeros(n, dtype=torch.float64, device=cuda_device)
    y = torch.zeros((nch, n), dtype=torch.float64, device=cuda_device)
    win = torch.ones(nfft, dtype=torch.float64, device=cuda_device)
    acc = torch.zeros((nch, 4, nfft // 2 + 1), dtype=torch.float64, device=cuda_device)
    rmap = engine.record_map(n, 0, 0, 0, n)
    p = lambda a: C.c_void_p(a.data_ptr())
    rc = lib.pyfft_welch_accumulate_static(_lib.F64, p(x), p(y), y.stride(0), nch, C.byref(rmap), C.byref(seg), p(win),
                                           None, None, p(acc), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == _lib.PYFFT_EUNSUPPORTED
    assert "static array holds" in _lib.last_error()
    torch.cuda.synchronize()
    assert float(acc.abs().max()) == 0.0
