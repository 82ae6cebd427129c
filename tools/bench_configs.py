"""Device-resident timings of the other BASELINE.json configurations:
c3 STFT (K2), c4 Hilbert / ccf (K3b), c5 cross-spectral matrix (K4), each against its
compulsory-bytes HBM roofline.  One JSON line per case; run on the GPU box:

    python tools/bench_configs.py > gpurun_out/configs.jsonl

These are not bench.py lines (bench.py measures the headline metric on config 2 only); they
feed DESIGN.md section 6.
"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def peak_gbs():
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            d = json.load(f)
        for k in ("hbm_gbs", "hbm_gbs_burst", "hbm_copy_gbs"):
            if k in d:
                return float(d[k]), "MEASURED_PEAKS.json:" + k
    except Exception:
        pass
    return 6459.3, "fallback (B200_PROFILING.md)"


def main():
    import torch
    from pyfft_b200 import engine
    from pyfft_b200.windows import windows

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    peak, peak_src = peak_gbs()
    only = set(sys.argv[1:])

    def timed(fn, reps=5, warm=3):
        if os.environ.get("PYFFT_BENCH_ONCE") == "1":          # under ncu: one launch of everything is enough
            reps, warm = 1, 1
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        return float(np.median(ts))

    def emit(case, ms, nbytes, samples, **extra):
        gbs = nbytes / ms / 1e6
        line = dict(case=case, ms=round(ms, 4), algorithmic_bytes=int(nbytes), achieved_gbs=round(gbs, 1),
                    peak_gbs=peak, peak_source=peak_src, frac=round(gbs / peak, 4),
                    msamples_per_s=round(samples / ms / 1e3, 1))
        line.update(extra)
        print(json.dumps(line), flush=True)

    def chirps(C, n, seed):
        g = torch.Generator(device=dev).manual_seed(seed)
        t = torch.arange(n, device=dev, dtype=torch.float32) / 1.0e6
        out = torch.empty((C, n), dtype=torch.float32, device=dev)
        for c in range(C):
            f0 = 10e3 + 190e3 * c / max(C, 1)
            out[c] = torch.sin(2 * np.pi * (f0 * t + 0.5 * 1e6 * t * t)) + 0.1 * torch.randn(n, generator=g, device=dev)
        return out

    # ---- c3: 64 channels x 2^24, nfft 2048, 75 % overlap, complex64 spectrogram output --------- #
    if not only or "c3" in only:
        C, n, L, hop = int(os.environ.get("PYFFT_C3_CH", "64")), 2 ** 24, 2048, 512
        x = chirps(C, n, 3)
        K = (n - L) // hop + 1
        win = windows("Hanning", nwins=L, verbose=False)
        rmap = engine.record_map(n, 0, 0, 0, n)
        F = L // 2
        hold = {}

        def run_c():
            hold["o"] = engine.stft(x, rmap, L, hop, 0, K, win, "fp32", nbins=F)

        ms = timed(run_c)
        emit("c3 stft complex64 (K2)", ms, C * n * 4 + C * K * F * 8, C * n, C=C, N=n, nfft=L, hop=hop, K=K)
        hold.clear()

        def run_p():
            hold["o"] = engine.stft(x, rmap, L, hop, 0, K, win, "fp32", nbins=F, power_out=True)

        ms = timed(run_p)
        emit("c3 stft power fp32 (K2)", ms, C * n * 4 + C * K * F * 4, C * n, C=C, N=n, nfft=L, hop=hop, K=K)
        hold.clear()
        del x
        torch.cuda.empty_cache()

    # ---- c4: 16 channels x 2^25, Hilbert envelope / analytic signal; ccf 0 vs 1..15 ------------ #
    if not only or "c4" in only:
        C, n = int(os.environ.get("PYFFT_C4_CH", "16")), 2 ** 25
        g = torch.Generator(device=dev).manual_seed(4)
        t = torch.arange(n, device=dev, dtype=torch.float32) / 1.0e6
        u = torch.empty((C, n), dtype=torch.float32, device=dev)
        base = (1 + 0.5 * torch.sin(2 * np.pi * 1e3 * t)) * torch.sin(2 * np.pi * 1e5 * t)
        for c in range(C):
            u[c] = torch.roll(base, 37 * c) + 0.05 * torch.randn(n, generator=g, device=dev)
        del base, t
        hold = {}

        def run_env():
            hold["o"] = engine.hilbert(u, "fp32", want_analytic=False, want_envelope=True)

        ms = timed(run_env)
        emit("c4 hilbert envelope (K3b)", ms, C * n * 4 * 2, C * n, C=C, N=n)
        hold.clear()

        def run_an():
            hold["o"] = engine.hilbert(u, "fp32", want_analytic=True, want_envelope=False)

        ms = timed(run_an)
        emit("c4 hilbert analytic complex64 (K3b)", ms, C * n * 4 * 3, C * n, C=C, N=n)
        hold.clear()
        torch.cuda.empty_cache()

        def run_ccf():
            hold["o"] = engine.xcorr(u[0], u[1:], "fp32", normalise=1)

        ms = timed(run_ccf, reps=3, warm=2)
        emit("c4 ccf 0 vs 1..%d (K3b)" % (C - 1), ms, C * n * 4 + (C - 1) * (2 * n - 1) * 4, C * n, C=C, N=n, fft_len=2 ** 26)
        hold.clear()
        del u
        torch.cuda.empty_cache()

    # ---- c5: 32 channels, nfft 8192, 50 % overlap, 32 x 32 cross-spectral matrix -------------- #
    if not only or "c5" in only:
        C, L, hop = 32, 8192, 4096
        n = 2 ** int(os.environ.get("PYFFT_C5_LOG2N", "25"))        # per-GPU share of 2^28 at 8 GPUs
        g = torch.Generator(device=dev).manual_seed(5)
        src = torch.randn((4, n), generator=g, device=dev)
        mix = torch.randn((C, 4), generator=g, device=dev)
        x = (mix @ src + torch.randn((C, n), generator=g, device=dev)).contiguous()
        del src
        K = (n - L) // hop + 1
        win = windows("Hanning", nwins=L, verbose=False)
        rmap = engine.record_map(n, 0, 0, 0, n)
        G = torch.zeros((L // 2 + 1, C, C), dtype=torch.complex128, device=dev)

        def run_csm():
            engine.csm_accumulate(x, rmap, L, hop, 0, K, win, "fp32", G=G)

        ms = timed(run_csm, reps=3, warm=2)
        flops = K * (L // 2 + 1) * C * C * 8.0
        emit("c5 cross-spectral matrix (K2 + K4)", ms, C * n * 4, C * n, C=C, N=n, nfft=L, hop=hop, K=K,
             product_tflops=round(flops / ms / 1e9, 2))


if __name__ == "__main__":
    main()
