#!/usr/bin/env python
"""Benchmark of the north-star path: two-channel Welch cross-PSD + coherence, nfft 4096,
50 % overlap, Hanning, global mean detrend, fp32 (BASELINE.json configs[1]).

    python bench.py --gpus N --steps K --warmup W            (torchrun for N > 1)
    python bench.py --impl reference ...                     the unmodified reference on the host cores (baseline/_ref)

One "step" = one full pass of the hot path over this rank's resident record slice:
detrend statistics -> (all-reduce) -> fused gather/window/FFT/accumulate -> (all-reduce of the
partial sums) -> finalize (scale, mean, coherence, phase).  Weak scaling: every rank holds
2 x 2^26 fresh samples of one long record of N * 2^26 samples per channel, sharded by
contiguous segment ranges with a read-only halo.  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NFFT, HOP, NCH = 4096, 2048, 2
N_PER_GPU = 2 ** 26
FS = 1.0e6
METRIC = "Welch cross-PSD Msamples/s (nfft 4096)"
UNIT = "Msamples/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--samples", type=int, default=N_PER_GPU, help="samples per channel per GPU")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: --samples per channel PER GPU; strong: --samples per channel in TOTAL")
    ap.add_argument("--no-graph", action="store_true", help="launch the step's kernels one by one (no CUDA graph)")
    ap.add_argument("--no-affinity", action="store_true",
                    help="do not bind the process to the CPU cores next to its GPU (engine.bind_host_to_gpu)")
    ap.add_argument("--deferred-exchange", action="store_true",
                    help="N > 1: a step's exchange is collected by the next step's last kernel (a stream of records: no "
                         "rank waits for a slower one inside a step; the timed region ends with the flush that completes "
                         "its last step) instead of being waited for inside the step.  Within +-1.5 %% of the default on "
                         "NVSwitch (profiles/r02f_multi_gpu.txt); the default line reports it beside its own number")
    ap.add_argument("--unfused", action="store_true",
                    help="the step as separate calls (prepare, accumulate, tail sums, [rereference, reduce,] finalize) "
                         "instead of the three launches of pyfft_welch_tc_step")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    return ap.parse_args()


# --------------------------------------------------------------------------- #
# CPU arm: literal port of the reference hot loop on a bounded sample
# --------------------------------------------------------------------------- #

def cpu_port_msamples(n=2 ** 22, repeats=1):
    """Times oracle.restatement.fft_pwelch_port (same per-segment loop, same materialised
    complex128 temporaries as fft_analysis.py:339-446) on 2 channels x n samples."""
    from oracle import restatement as R
    from oracle import signals
    t, x, y = signals.c2_cross(n=n, dtype=np.float32)
    x = x.astype(np.float64)
    y = y.astype(np.float64)
    navr = n // HOP - 1
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        R.fft_pwelch_port(t, x, y, navr, 0.5, "Hanning", 1)
        best = min(best, time.perf_counter() - t0)
    return NCH * n / best / 1e6, best


REF_DIR = os.path.join(ROOT, "baseline", "_ref")


def cpu_reference_msamples(n=2 ** 23, repeats=1):
    """Times the UNMODIFIED reference (baseline/_ref/fft_analysis.py, placed there by
    __graft_entry__.build(), loaded behind oracle/ref_shim.py's import shim) through its own public
    call fft_pwelch(..., useMLAB=False, plotit=False) on 2 channels x n float64 samples.
    -> (Msamples/s, seconds) or None when the reference modules are not on this machine."""
    from oracle import ref_shim, signals
    if not ref_shim.available(REF_DIR):
        return None
    ref = ref_shim.load_reference(REF_DIR)["fft_analysis"]
    t, x, y = signals.c2_cross(n=n, dtype=np.float32)
    x = x.astype(np.float64)
    y = y.astype(np.float64)
    navr = n // HOP - 1
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        out = ref.fft_pwelch(t, x, y, None, Navr=navr, windowoverlap=0.5, windowfunction="Hanning",
                             useMLAB=False, plotit=False, verbose=False, detrend_style=1)
        best = min(best, time.perf_counter() - t0)
    assert out[6].nwins == NFFT and out[2].shape == (NFFT // 2,)
    return NCH * n / best / 1e6, best


def run_reference(args):
    """CPU arm: the reference's own fft_pwelch on the box's host cores, a bounded cut of the workload
    per step (numpy's pocketfft and the Python segment loop are single-threaded: 1 core)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    lg = 23 if args.steps + args.warmup <= 12 else 22      # keeps the whole run within a few minutes
    n = 2 ** lg
    stock = cpu_reference_msamples(n) is not None
    fn = cpu_reference_msamples if stock else cpu_port_msamples
    vals = []
    for i in range(args.warmup + args.steps):
        v, sec = fn(n)
        if i >= args.warmup:
            vals.append((v, sec))
    ms = 1e3 * sum(s for _, s in vals) / len(vals)
    value = NCH * n / (ms / 1e3) / 1e6
    port_v, _ = cpu_port_msamples(n)
    sample = "2 ch x 2^%d samples per step (1/%d of the 2^26 workload), float64 like the reference" % (lg, 2 ** (26 - lg))
    cfg = workload_config(args.gpus, N_PER_GPU)
    cfg["workload"] += " -- CPU arm times a 2 x 2^%d-sample cut of it per step" % lg
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": cfg,
           "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": "reference" if stock else "port",
                            "sample": sample, "port_value": port_v},
           "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0,
           "note": ("unmodified reference fft_analysis.fft_pwelch(useMLAB=False, plotit=False) from baseline/_ref "
                    "behind oracle/ref_shim.py" if stock else
                    "numpy port of fft_pwelch's home-brew loop (baseline/_ref absent on this machine)") +
                   "; numpy's pocketfft and the Python loop are single-threaded, so 1 of %d host cores is used; "
                   "port_value = oracle.restatement.fft_pwelch_port on the same cut" % (os.cpu_count() or 1)}
    print(json.dumps(out), flush=True)


def workload_config(world, n_per_gpu):
    return {"workload": "two-channel cross-PSD + coherence, %d x 2^%d samples per channel, nfft=4096, "
                        "50%% overlap, Hanning, mean detrend, fp32" % (world, int(np.log2(n_per_gpu))),
            "channels": NCH, "samples_per_channel": world * n_per_gpu, "nfft": NFFT, "hop": HOP,
            "parallelism": "segment ranges x%d + 1 all-reduce" % world if world > 1 else "single GPU",
            "cache": "inputs (%.0f MB per GPU) exceed the 126 MB L2; no flush needed"
                     % (NCH * n_per_gpu * 4 / 1e6)}


# --------------------------------------------------------------------------- #
# clocks sampler
# --------------------------------------------------------------------------- #

class Clocks(object):
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
        rows = self.rows[-8:]                    # the samples taken under load at the end
        sm = sorted(int(r[0]) for r in rows if r and r[0].isdigit())
        mx = max([int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()] or [0])
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [nm for i, nm in enumerate(names)
                   if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None,
                "reasons": reasons, "samples": len(sm)}


# --------------------------------------------------------------------------- #
# GPU arm
# --------------------------------------------------------------------------- #

def make_slice(torch, dev, g_lo, g_hi, seed, dc, amp, phase):
    """Samples g_lo..g_hi-1 of one long synthetic record: tone + unit white noise + DC.
    Noise comes from per-block generators keyed by the GLOBAL block index, so the halo
    samples two ranks share are identical."""
    blk = 1 << 20
    out = torch.empty(g_hi - g_lo, dtype=torch.float32, device=dev)
    f0 = 400.0 / NFFT                                          # bin-centred tone, cycles/sample
    for b in range(g_lo // blk, (g_hi - 1) // blk + 1):
        lo, hi = max(g_lo, b * blk), min(g_hi, (b + 1) * blk)
        gen = torch.Generator(device=dev).manual_seed(seed * 100003 + b)
        noise = torch.randn(blk, generator=gen, device=dev, dtype=torch.float32)
        g = torch.arange(lo, hi, device=dev, dtype=torch.float64)
        tone = (amp * torch.sin(2 * np.pi * f0 * g + phase)).float()
        out[lo - g_lo:hi - g_lo] = tone + noise[lo - b * blk:hi - b * blk] + dc
    return out


def traffic_of(kernel, n):
    """DRAM bytes (read + write) per launch of the dominant kernel from the committed ncu --set full
    capture (profiles/traffic.json), valid for the single-GPU workload size it was taken on; else None."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[kernel]
    except (OSError, KeyError, ValueError):
        return None
    return t["dram_bytes_read"] + t["dram_bytes_write"] if n == 2 ** 26 else None


def run_ours(args):
    import torch
    import torch.distributed as dist
    from pyfft_b200 import _lib, engine, fft_pwelch
    from pyfft_b200 import dist as pdist
    from pyfft_b200.fft_analysis import fftanal
    from pyfft_b200.windows import windows

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        group = dist.group.WORLD

    # host side of e2e: this rank's pinned buffers live on the NUMA node of its GPU
    cores = None if args.no_affinity else engine.bind_host_to_gpu(dev)
    clocks = Clocks(local)                                     # sampled over the whole run (100 ms period)
    if rank == 0:
        clocks.start()
    if args.scaling == "strong":
        n, ntot = args.samples // world, args.samples          # per-rank share (segments are split, + halo)
    else:
        n, ntot = args.samples, world * args.samples
    navr = ntot // HOP - 1                                     # -> nwins = 4096 (fft_analysis.py:2411-2418)
    assert fftanal._getNwins(ntot, navr, 0.5) == NFFT
    K, nana, refl = navr, ntot, NFFT - 1                       # whole record => reflect map (:196-205)
    lp = pdist.local_plan(K, HOP, NFFT, nana, 0, refl, rank, world)
    g_lo, g_hi = lp["src"]
    x = make_slice(torch, dev, g_lo, g_hi, 11, 7.0, 1.0, 0.0)
    y = make_slice(torch, dev, g_lo, g_hi, 12, 2.5, 0.5, -np.pi / 4)[None, :].contiguous()
    rmap = engine.record_map(nana, 0, refl, g_lo, g_hi - g_lo)
    win = windows("Hanning", nwins=NFFT, verbose=False)
    S1, S2, _, ENBW = fftanal._getNorms(win, NFFT // 2, FS)
    scale = 1.0 / S1 ** 2 / ENBW
    wp = engine.WelchPlan(NFFT, HOP, 1, "fp32", win, dev)
    k0, k1 = lp["k0"], lp["k1"]
    s_lo, s_hi = lp["stats_span"]

    c_lo, c_hi = k0 * HOP, k1 * HOP                           # span the fused kernel sums itself
    assert wp.can_fuse_mean

    use_tc = wp.can_use_tc and os.environ.get("PYFFT_NO_TC") != "1"

    tails = ((s_lo, c_lo), (c_hi, s_hi))
    fused_step = use_tc and not args.unfused
    defer = world > 1 and args.deferred_exchange

    def step():
        # ONE pass over HBM: provisional mean (+ scale) from 64 Ki samples, fused Welch
        # accumulation (tensor-core DFT when available), exact mean correction by linearity
        # after the (all-reduced) raw sums are known
        if fused_step:
            # three launches: statistics + head + tail sums | main kernel | partial sums + NVLink exchange + finalize
            # sharded: the exchange of a step is collected by the next one (a stream of records; no rank waits for
            # a slower one inside a step); the timed region ends with the flush that completes its last step
            out = wp.step_tc(x, y, rmap, k0, k1 - k0, tails, K, nana, scale, True, group, deferred=defer)
            if out is not None or wp._deferred_args is not None:
                return out
        wp.reset()
        if use_tc:
            wp.prepare_tc(x, y, rmap, c_lo, c_hi)
            wp.accumulate_tc(x, y, rmap, k0, k1 - k0)
        else:
            wp.estimate_mean(x, y, rmap, c_lo, c_hi)
            wp.accumulate_mean(x, y, rmap, k0, k1 - k0)
        if s_lo < c_lo:
            wp.detrend_stats(x, y, rmap, s_lo, c_lo)
        if s_hi > c_hi:
            wp.detrend_stats(x, y, rmap, c_hi, s_hi)
        if group is not None:
            wp.rereference()                       # shards to a common reference, then ONE collective
            wp.all_reduce_state(group)
        return wp.finalize(scale, onesided=True, K=K, correct_nana=nana)         # correction folded in

    def step_two_pass():
        wp.reset()
        wp.detrend_stats(x, y, rmap, s_lo, s_hi)
        wp.solve_detrend(_lib.DETREND_MEAN, nana, group)
        wp.accumulate(x, y, rmap, k0, k1 - k0)
        if group is not None:
            dist.all_reduce(wp.acc, group=group)
        return wp.finalize(scale, onesided=True, K=K)

    def barrier():
        if group is not None:
            dist.barrier(group=group)
        torch.cuda.synchronize()

    gate = {"red": None, "buf": None}

    def device_gate():
        """N > 1: after the host barrier the ranks' CPUs reach the timed loop tens of microseconds apart, and coupled
        steps (one exchange per step) charge that start skew to the timed region of every rank.  A 2 ms device-side
        sleep lets every rank enqueue its whole timed region, then a rendezvous kernel over NVLink peer memory (else
        an NCCL all-reduce) releases all GPUs within a few microseconds of each other; the start event follows it."""
        if group is None:
            return
        if gate["buf"] is None:
            gate["buf"] = torch.zeros(8, dtype=torch.float64, device=dev)
            gate["red"] = engine.PeerReducer.create(8, dev, group)
        torch.cuda._sleep(4000000)
        if gate["red"]:
            gate["red"].all_reduce(gate["buf"])
        else:
            dist.all_reduce(gate["buf"], group=group)

    def timed(fn, steps, warmup, last=None):
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        device_gate()
        e0.record()
        for _ in range(steps):
            fn()
        if last is not None:
            last()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if group is not None:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX, group=group)
        return float(ms.item()) / steps

    # the step's launches are always the same: captured once into a CUDA graph and replayed (the five small
    # kernels around K1 then cost their device time, not a Python + ctypes round trip each)
    step()
    torch.cuda.synchronize()
    step_is_fused = fused_step and wp._aux is not None
    step_run = step if args.no_graph else engine.GraphedStep(step)
    deferred_on = step_is_fused and defer and wp._deferred_args is not None
    flush = (lambda: wp.flush_tc(force=True)) if deferred_on else None
    ms_step = timed(step_run, args.steps, max(args.warmup, 3), flush)

    def run_complete():
        """One more step and, when the exchange is deferred, the flush that completes it -> its spectra."""
        r = step_run()
        return wp.flush_tc(force=True) if deferred_on else r

    ms_sync = None
    if deferred_on:
        # the same step with its own exchange waited for inside the step (what one isolated call costs)
        def step_sync():
            return wp.step_tc(x, y, rmap, k0, k1 - k0, tails, K, nana, scale, True, group, deferred=False)
        ms_sync = timed(step_sync if args.no_graph else engine.GraphedStep(step_sync), args.steps, 3)
    ms_deferred = None
    if step_is_fused and world > 1 and not deferred_on and wp._peer is None:
        # for the record: the exchange of a step collected by the next one (no rank waits inside a step)
        def step_def():
            return wp.step_tc(x, y, rmap, k0, k1 - k0, tails, K, nana, scale, True, group, deferred=True)
        step_def()
        if wp._deferred_args is not None:
            run_def = step_def if args.no_graph else engine.GraphedStep(step_def)
            ms_deferred = timed(run_def, args.steps, 3, lambda: wp.flush_tc(force=True))
            wp.flush_tc(force=True)
    # dominant kernel alone (welch_pair_kernel + its partial-sum reductions, ~3 us)
    wp.prepare_tc(x, y, rmap, c_lo, c_hi) if use_tc else wp.estimate_mean(x, y, rmap, c_lo, c_hi)
    ms_k1 = timed(lambda: (wp.accumulate_tc if use_tc else wp.accumulate_mean)(x, y, rmap, k0, k1 - k0),
                  args.steps, 3)
    wp.estimate_mean(x, y, rmap, c_lo, c_hi)
    ms_k1_cuda = timed(lambda: wp.accumulate_mean(x, y, rmap, k0, k1 - k0), args.steps, 3)
    ms_two = timed(step_two_pass, args.steps, 3)
    r1, r2 = run_complete(), step_two_pass()
    dev_fused = float((r1["Pxy"] - r2["Pxy"]).abs().max() / r2["Pxy"].abs().max())
    assert dev_fused < 1e-5 and not wp.tc_failed(), dev_fused
    dev_coh = float((r1["Cxy"] - r2["Cxy"]).abs().max())
    # keep the same step running for ~0.8 s so that the 100 ms clock sampler sees the load.  The number of steps is
    # derived from the (max-over-ranks) step time, NOT from each rank's own clock: the steps of a sharded run are
    # coupled by their exchange, and a rank that ran one batch more than its peers would wait for them until the
    # bounded wait of the exchange expires
    for _ in range(max(1, int(np.ceil(800.0 / (50.0 * ms_step))))):
        for _ in range(50):
            step_run()
        torch.cuda.synchronize()
    clk = clocks.stop() if rank == 0 else None
    res = run_complete()
    pk = int(torch.argmax(res["Pxx"].real))
    assert pk == 400 and bool(torch.isfinite(res["Cxy"].real).all()), "bench sanity check failed"

    value = NCH * ntot / (ms_step / 1e3) / 1e6
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    alg_bytes = NCH * n * 4                                    # every input sample read once
    achieved = alg_bytes / (ms_k1 / 1e3) / 1e9

    # ---- end to end: pinned host buffers -> public API -> host result ------------ #
    e2e = None
    if not args.no_e2e:
        xh = x.cpu().pin_memory()
        yh = y[0].cpu().pin_memory()
        if world == 1:
            def full(h):                                       # plain n-sample host record
                o = torch.empty(n, dtype=torch.float32).pin_memory()
                o[:h.numel()] = h
                o[h.numel():] = h[:n - h.numel()]
                return o
            xh, yh = full(xh), full(yh)
            tv = np.arange(n) / FS

            def e2e_step():
                out = fft_pwelch(tv, xh, yh, Navr=navr, windowoverlap=0.5, windowfunction="Hanning",
                                 plotit=False, detrend_style=1, precision="fp32")
                return out[1]
            h2d = 2 * n * 4
        else:
            class Uniform(object):                             # tvec of the global record without 8 bytes per sample
                def __init__(self, m):
                    self.m = m

                def __len__(self):
                    return self.m

                def __getitem__(self, i):
                    return (i if i >= 0 else self.m + i) / FS
            tv = Uniform(ntot)

            def e2e_step():                                    # public API, this rank's slice of the record
                out = fft_pwelch(tv, xh, yh, Navr=navr, windowoverlap=0.5, windowfunction="Hanning",
                                 plotit=False, detrend_style=1, precision="fp32", group=group,
                                 shard_offset=g_lo)
                return out[1]
            h2d = 2 * (g_hi - g_lo) * 4
        es = max(3, min(args.steps, 10))
        ms_e2e = timed(e2e_step, es, 2)

        # the ceiling of e2e on this box: the same host -> device copies alone (all ranks at once)
        xd_, yd_ = torch.empty_like(xh, device=dev), torch.empty_like(yh, device=dev)

        def copies_only():
            xd_.copy_(xh, non_blocking=True)
            yd_.copy_(yh, non_blocking=True)
        ms_h2d = timed(copies_only, es, 2)
        del xd_, yd_
        e2e = {"value": NCH * ntot / (ms_e2e / 1e3) / 1e6, "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": 6 * (NFFT // 2) * 16 + 4 * NFFT * 8, "ms_per_step": ms_e2e,
               "h2d_only_ms_per_step": ms_h2d, "h2d_only_gbs_per_gpu": h2d / (ms_h2d / 1e3) / 1e9,
               "ceiling_from_h2d": NCH * ntot / (ms_h2d / 1e3) / 1e6,
               "steps": es, "api": "pyfft_b200.fft_pwelch" if world == 1 else "pyfft_b200.fft_pwelch(group=, shard_offset=)"}

    if rank != 0:
        if group is not None:
            dist.destroy_process_group()
        return
    cpu = None
    if world == 1 and not args.no_cpu:
        r = cpu_reference_msamples(2 ** 23, repeats=2)
        pv, psec = cpu_port_msamples(2 ** 23, repeats=1)
        v, sec = r if r is not None else (pv, psec)
        cpu = {"value": v, "unit": UNIT, "cores": 1, "kind": "reference" if r is not None else "port",
               "sample": "2 ch x 2^23 samples (1/8 of the workload), float64, best of 2 (%.2f s)" % sec,
               "port_value": pv}
    elif world > 1:
        cpu = {"value": None, "unit": UNIT, "cores": 1, "kind": "reference",
               "sample": "not timed on N > 1 lines (rank 0 of the N = 1 line carries it)"}
    kname = "welch_tc_kernel (tcgen05)" if use_tc else "welch_pair_kernel<float,12,fused mean>"
    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
           "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
           "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": workload_config(world, n),
           "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                        "frac": achieved / peak, "traffic": traffic_of(kname, n),
                        "kernel": kname,
                        "algorithmic_bytes": alg_bytes, "kernel_ms": ms_k1,
                        "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s",
                        "whole_step_frac": alg_bytes / (ms_step / 1e3) / 1e9 / peak,
                        "two_pass_step_ms": ms_two, "cuda_core_kernel_ms": ms_k1_cuda,
                        "vs_two_pass_cuda_core_max_rel": dev_fused, "vs_two_pass_cuda_core_coherence_abs": dev_coh},
           "cpu_baseline": cpu, "e2e": e2e, "clocks": clk,
           # kernels of this library per step; sharded: + rereference + the NVLink peer reduction (csrc/peer.cu)
           "gpu_launches": (3 if step_is_fused else (6 if world == 1 else (8 if wp._peer else 7))) * args.steps,
           "step_kernels": ("tc_pre_kernel, welch_tc_kernel, tc_post_kernel" if step_is_fused else "unfused sequence"),
           "exchange": (None if world == 1 else
                        ("deferred: collected by the next step's tc_post_kernel, the timed region ends with the flush "
                         "of its last step" if deferred_on else "waited for inside every step")),
           "sync_exchange_ms_per_step": ms_sync, "deferred_exchange_ms_per_step": ms_deferred,
           "launch_mode": "per-kernel launches" if args.no_graph else "one CUDA graph replay per step",
           "host_affinity": (None if cores is None else "%d cores next to the GPU (NVML)" % len(cores))}
    if world > 1:
        out["config"]["collective"] = (
            "inside tc_post_kernel: per-item pushes over NVLink peer memory, 131 KB per rank" if step_is_fused else
            ("one kernel over NVLink peer memory (pyfft_peer_allreduce), 131 KB" if wp._peer
             else "one NCCL all-reduce, 131 KB"))
    print(json.dumps(out), flush=True)
    if group is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
