"""BASELINE.json configurations 3, 4 and 5 at their FULL sizes on one GPU, through properties that
do not need a full-size oracle run: sampled units against the float64 numpy definition
(fft_analysis.py:2152-2176, hilbert.py:47-64, ccf.py:66-77), exactness of what must be exact,
known delays / envelopes of the synthetic recipes, and additivity over segment
ranges (what the multi-GPU all-reduce relies on)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_config3_full_size_sampled_segments(cuda_device):
    """64 channels x 2^24 samples, nfft 2048, 75 % overlap: 32 765 segments per channel; 96 sampled
    (channel, segment) spectra against numpy's FFT of the same windowed samples, power output
    equal to |X|^2 of the complex output, and linearity in the record."""
    import torch
    from pyfft_b200 import engine
    from pyfft_b200.windows import windows
    dev = cuda_device
    C, n, L, hop = 64, 2 ** 24, 2048, 512
    K = (n - L) // hop + 1
    g = torch.Generator(device=dev).manual_seed(3)
    t = torch.arange(n, device=dev, dtype=torch.float32) / 1.0e6
    x = torch.empty((C, n), dtype=torch.float32, device=dev)
    for c in range(C):
        f0 = 10e3 + 190e3 * c / C
        x[c] = torch.sin(2 * np.pi * (f0 * t + 0.5e6 * t * t)) + 0.1 * torch.randn(n, generator=g, device=dev)
    del t
    win = windows("Hanning", nwins=L, verbose=False)
    rmap = engine.record_map(n, 0, 0, 0, n)
    out, _ = engine.stft(x, rmap, L, hop, 0, K, win, "fp32", nbins=L // 2)
    assert out.shape == (C, K, L // 2) and out.dtype == torch.complex64
    rng = np.random.default_rng(33)
    picks = [(int(rng.integers(C)), int(k)) for k in rng.integers(0, K, 90)] + [(0, 0), (C - 1, K - 1), (5, K - 2),
                                                                                 (7, 1), (63, 0), (31, K // 2)]
    worst = 0.0
    for c, k in picks:
        seg = x[c, k * hop:k * hop + L].cpu().numpy().astype(np.float64)
        ref = np.fft.fft(win * seg)[:L // 2]
        got = out[c, k].cpu().numpy().astype(np.complex128)
        worst = max(worst, float(np.max(np.abs(got - ref)) / np.max(np.abs(ref))))
    assert worst <= 2e-6, worst
    # power output = |complex output|^2 (same kernel, other epilogue), on a slice of channels
    pw, _ = engine.stft(x[:4], rmap, L, hop, 0, K, win, "fp32", nbins=L // 2, power_out=True)
    ref_pw = out[:4].real.double() ** 2 + out[:4].imag.double() ** 2
    assert float((pw.double() - ref_pw).abs().max() / ref_pw.max()) <= 1e-6
    del pw, ref_pw
    # linearity: STFT(2.5 x) = 2.5 STFT(x) (power-of-two-free factor, fp32 rounding only)
    o2, _ = engine.stft(2.5 * x[60:], rmap, L, hop, 0, K, win, "fp32", nbins=L // 2)
    d = (o2 - 2.5 * out[60:]).abs().max() / out[60:].abs().max()
    assert float(d) <= 2e-6


def test_config4_full_size_hilbert_and_ccf(cuda_device):
    """16 channels x 2^25 samples: the real part of the analytic signal IS the record (exact), the
    envelope of the amplitude-modulated carrier follows 1 + 0.5 sin(2 pi 1e3 t), the analytic
    spectrum has no negative frequencies (checked with an independent FFT on one channel), and
    the ccf of channel 0 against the delayed channels peaks at the known delays."""
    import torch
    from pyfft_b200 import engine
    dev = cuda_device
    C, n, fs = 16, 2 ** 25, 1.0e6
    g = torch.Generator(device=dev).manual_seed(4)
    t = torch.arange(n, device=dev, dtype=torch.float64) / fs
    base = ((1 + 0.5 * torch.sin(2 * np.pi * 1e3 * t)) * torch.sin(2 * np.pi * 1e5 * t)).float()
    amp = (1 + 0.5 * torch.sin(2 * np.pi * 1e3 * t)).float()
    del t
    u = torch.empty((C, n), dtype=torch.float32, device=dev)
    for c in range(C):
        u[c] = torch.roll(base, 37 * c) + 0.05 * torch.randn(n, generator=g, device=dev)
    z, _ = engine.hilbert(u, "fp32", want_analytic=True, want_envelope=False)
    assert z.shape == (C, n) and z.dtype == torch.complex64
    assert torch.equal(z.real, u)                               # hilbert.py:66: real part = input
    _, env = engine.hilbert(u, "fp32", want_analytic=False, want_envelope=True)
    assert float((env - z.abs()).abs().max()) <= 1e-5           # fused envelope = |analytic|
    # envelope of the noiseless carrier, away from the ends (circular transform) and for channel 0
    zb, _ = engine.hilbert(torch.stack([base, base]), "fp32", want_analytic=True, want_envelope=False)
    inner = slice(n // 64, n - n // 64)
    assert float((zb[0].abs()[inner] - amp[inner]).abs().max()) <= 2e-3
    assert float((zb[0] - zb[1]).abs().max()) <= 2e-5           # both halves of a pair: H(a) = Im A - b, H(b) = a - Re A
    del zb, amp, base
    # no power at negative frequencies: independent FFT of one analytic record
    Z = torch.fft.fft(z[3].to(torch.complex128))
    neg = Z[n // 2 + 1:].abs().max()
    assert float(neg / Z.abs().max()) <= 1e-6
    del Z, z, env
    torch.cuda.empty_cache()
    # ccf of channel 0 against channels 1..15 (padded transforms of 2^26 points, two records each)
    co = engine.xcorr(u[0], u[1:], "fp32", normalise=1)
    assert co.shape == (C - 1, 2 * n - 1)
    peak = torch.argmax(co, dim=1).cpu().numpy()
    # numpy.correlate(x0, xc, 'full') with xc = x0 delayed by d samples peaks at index n - 1 - d
    want = np.array([n - 1 - 37 * c for c in range(1, C)])
    assert np.array_equal(peak, want), (peak - (n - 1))
    pk = co[torch.arange(C - 1, device=dev), torch.from_numpy(want).to(dev)]
    assert float(pk.min()) > 0.95 and float(co.abs().max()) <= 1.0 + 1e-5      # normalised: |ccf| <= 1
    a0 = engine.xcorr(u[0], u[0:1], "fp32", normalise=1)
    assert abs(float(a0[0, n - 1]) - 1.0) <= 1e-5               # zero lag of an auto-ccf


def test_config5_full_size_matrix_properties(cuda_device):
    """32 channels x 2^28 samples, nfft 8192, 50 % overlap (65 535 segments, spectra processed in
    16 GB chunks): Hermitian matrix with real non-negative diagonal, Cauchy-Schwarz, Parseval
    against the sample variance, and additivity over segment ranges (the property the all-reduce
    of the sharded path relies on)."""
    import torch
    from pyfft_b200 import engine
    from pyfft_b200.windows import windows
    dev = cuda_device
    C, n, L, hop = 32, 2 ** 28, 8192, 4096
    K = (n - L) // hop + 1
    g = torch.Generator(device=dev).manual_seed(5)
    x = torch.empty((C, n), dtype=torch.float32, device=dev)
    src = torch.randn((4, n), generator=g, device=dev)
    mix = torch.randn((C, 4), generator=g, device=dev)
    for c in range(C):                                           # row by row: no 34 GB temporary
        x[c] = mix[c] @ src
        x[c] += torch.randn(n, generator=g, device=dev)
    del src
    win = windows("Hanning", nwins=L, verbose=False)
    rmap = engine.record_map(n, 0, 0, 0, n)
    G = engine.csm_accumulate(x, rmap, L, hop, 0, K, win, "fp32")
    assert G.shape == (L // 2 + 1, C, C) and G.dtype == torch.complex128
    scale = float(G.abs().max())
    assert float((G - G.conj().transpose(1, 2)).abs().max()) <= 1e-12 * scale        # mirrored at the flush
    dg = torch.diagonal(G, dim1=1, dim2=2)
    assert float(dg.imag.abs().max()) == 0.0 and float(dg.real.min()) >= 0.0
    cs = G.abs() ** 2 - (dg.real[:, :, None] * dg.real[:, None, :])
    assert float(cs.max()) <= 1e-9 * scale * scale                                    # |G_ij|^2 <= G_ii G_jj
    # Parseval: sum over ALL bins of |X|^2 = L * sum (w x)^2; with white-ish data the segment average
    # of sum (w x)^2 is S2 * var -> one-sided sum (interior bins doubled) / (K L S2) ~ variance
    S2 = float(np.sum(win ** 2))
    one = dg.real.clone()
    one[1:-1] *= 2.0
    var_est = one.sum(dim=0) / (K * L * S2)
    var = x.double().var(dim=1) if False else torch.stack([x[c].double().var() for c in range(C)])
    assert float(((var_est - var) / var).abs().max()) <= 5e-3
    # additivity over segment ranges (two "ranks"): G = G[0:K1) + G[K1:K)
    K1 = K // 2 + 1
    Ga = engine.csm_accumulate(x, rmap, L, hop, 0, K1, win, "fp32")
    engine.csm_accumulate(x, rmap, L, hop, K1, K - K1, win, "fp32", G=Ga)
    assert float((Ga - G).abs().max()) <= 2e-6 * scale
