"""GPU: BASELINE.json config 2 at FULL size (2 channels x 2^26 samples, nfft 4096, 50 %
overlap, fp32) checked through size-independent properties, because the float64 oracle
needs far more than seconds at this size:

  * Parseval: sum_f |X_k[f]|^2 = L * sum_n |w x_k[n]|^2, summed over all segments, with the
    right-hand side evaluated by plain torch elementwise ops on an unfolded view
    (independent of every kernel under test)
  * associativity: accumulating the segments in two shards equals one pass
  * linearity: P(a x) = a^2 P(x)
  * a 2^22-sample prefix of the same record against the oracle
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

N, L, H = 2 ** 26, 4096, 2048


def _signals(dev):
    import torch
    g = torch.Generator(device=dev).manual_seed(2)
    ph = torch.arange(N, device=dev, dtype=torch.float64) * (2 * np.pi * 400.0 / L)
    x = (torch.sin(ph) + torch.randn(N, generator=g, device=dev, dtype=torch.float64) + 7.0).float()
    y = (0.5 * torch.sin(ph - np.pi / 4) + torch.randn(N, generator=g, device=dev, dtype=torch.float64)
         + 2.5).float()
    del ph
    return x, y


def test_config2_full_size_properties(cuda_device):
    import torch
    from pyfft_b200 import _lib, engine
    from pyfft_b200.windows import windows
    dev = cuda_device
    x, y = _signals(dev)
    yc = y[None, :].contiguous()
    win = windows("Hanning", nwins=L, verbose=False)
    K = (N - H) // H                                   # whole record, no reflect (plain map)
    rmap = engine.record_map(N, 0, 0, 0, N)

    def run(xx, yy, shards):
        wp = engine.WelchPlan(L, H, 1, "fp32", win, dev)
        wp.detrend_stats(xx, yy, rmap, 0, N)
        wp.solve_detrend(_lib.DETREND_MEAN, N)
        for k0, k1 in shards:
            wp.accumulate(xx, yy, rmap, k0, k1 - k0)
        return wp

    wp = run(x, yc, [(0, K)])
    acc = wp.acc.cpu().numpy()[0]                       # (4, L/2+1) unscaled sums
    coef = wp.coef.cpu().numpy()
    assert abs(coef[0] - x.double().mean().item()) < 1e-9 and coef[1] == 0.0
    assert abs(coef[2] - y.double().mean().item()) < 1e-9

    # Parseval with an independent right-hand side
    w64 = torch.from_numpy(win).to(dev)
    for q, (sig, m) in enumerate(((x, coef[0]), (y, coef[2]))):
        rhs = 0.0
        view = sig.unfold(0, L, H)[:K]
        for s in range(0, K, 2048):
            blk = (view[s:s + 2048].double() - m) * w64
            rhs += float((blk * blk).sum())
        full = 2.0 * acc[q].sum() - acc[q][0] - acc[q][L // 2]    # Hermitian: bins 1..L/2-1 twice
        assert abs(full / (L * rhs) - 1.0) < 5e-6, (q, full, L * rhs)

    # associativity over segment shards (what the multi-GPU path relies on)
    wp2 = run(x, yc, [(0, K // 3), (K // 3, K)])
    a2 = wp2.acc.cpu().numpy()[0]
    assert np.max(np.abs(a2 - acc)) <= 2e-6 * np.max(np.abs(acc))
    assert wp2.K == K

    # linearity
    wp3 = run(x * 3.0, yc, [(0, K)])
    a3 = wp3.acc.cpu().numpy()[0]
    assert np.max(np.abs(a3[0] - 9.0 * acc[0])) <= 1e-5 * np.max(acc[0])
    assert np.max(np.abs(a3[2] - 3.0 * acc[2])) <= 1e-5 * np.max(np.abs(acc[2]))

    # coherence / phase of the shared tone
    res = wp.finalize(1.0, onesided=True)
    cxy = res["Cxy"].cpu().numpy()[:, 0]
    phi = res["phi_xy"].cpu().numpy()[:, 0]
    assert abs(abs(cxy[400]) - 1.0) < 0.02 and abs(phi[400] + np.pi / 4) < 0.01   # Y conj(X): -pi/4
    assert np.median(np.abs(cxy[1000:2000])) < 0.02


def test_config2_prefix_against_oracle(cuda_device):
    from oracle import restatement as R
    from pyfft_b200 import fft_pwelch
    from conftest import relerr
    x, y = _signals(cuda_device)
    n = 2 ** 22
    xs, ys = x[:n].cpu().numpy(), y[:n].cpu().numpy()
    t = np.arange(n) / 1e6
    kw = dict(Navr=2047, windowoverlap=0.5, windowfunction="Hanning", detrend_style=1)
    ref = R.fft_pwelch(t, xs.astype(np.float64), ys.astype(np.float64), **kw)
    f, Pxy, Pxx, Pyy, Cxy, phi, info = fft_pwelch(t, xs, ys, plotit=False, **kw)
    assert info.nwins == L and info.Navr == 2047 and info.precision == "fp32"
    for a, nm in ((Pxy, "Pxy"), (Pxx, "Pxx"), (Pyy, "Pyy")):
        assert relerr(a, ref[nm]) <= 1e-5, nm
    assert relerr(Cxy, ref["Cxy"], 1.0) <= 1e-4


def test_config2_tensor_core_path_vs_cuda_core_path(cuda_device):
    """The tcgen05 DFT-as-GEMM kernel against the fp32 register-FFT kernel at full size, on the
    hardest input for it (white noise only: every bin is as high as the peak), plus the
    sine + noise record of config 2.  Contract: 1e-5 of the spectral peak, 1e-4 on coherence."""
    import torch
    from pyfft_b200 import fft_pwelch
    dev = cuda_device
    t = np.arange(N) / 1.0e6
    g = torch.Generator(device=dev).manual_seed(77)
    xs, ys = _signals(dev)
    cases = {"sine+noise": (xs, ys)}
    xn = torch.randn(N, generator=g, device=dev, dtype=torch.float32) * 3.0 + 11.0
    yn = 0.6 * xn + torch.randn(N, generator=g, device=dev, dtype=torch.float32) * 0.1 - 4.0
    cases["noise only, coherent pair"] = (xn, yn)
    kw = dict(Navr=N // H - 1, windowoverlap=0.5, windowfunction="Hanning", detrend_style=1, plotit=False)
    for name, (x, y) in cases.items():
        a = fft_pwelch(t, x, y, tensor_cores=True, **kw)
        b = fft_pwelch(t, x, y, tensor_cores=False, **kw)
        assert a[6].tensor_cores and not b[6].tensor_cores and a[6].nwins == L
        for i, nm in ((1, "Pxy"), (2, "Pxx"), (3, "Pyy")):
            err = np.max(np.abs(a[i] - b[i])) / np.max(np.abs(b[i]))
            assert err <= 1e-5, (name, nm, err)
        assert np.max(np.abs(a[4] - b[4])) <= 1e-4, name
