"""The fused tensor-core step (pyfft_welch_tc_step: tc_pre_kernel -> welch_tc_kernel -> tc_post_kernel) against
the unfused sequence of calls it replaces, one-sided and two-sided, with and without a reflect-padded head, and
the sharded exchange of tc_post_kernel with two ranks driven from one process (two streams, plain device
buffers as peer memory)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

NFFT, HOP = 4096, 2048


def _setup(n, nch, seed=3):
    from pyfft_b200 import engine
    from pyfft_b200.windows import windows
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(seed)
    t = torch.arange(n, dtype=torch.float64) / 1e6
    x = (torch.sin(2 * np.pi * 97656.25 * t) + 0.5 * torch.randn(n, generator=g, dtype=torch.float64) + 3.0).float()
    y = torch.stack([(0.7 * torch.sin(2 * np.pi * 97656.25 * t + 0.3 * (c + 1)) +
                      0.4 * torch.randn(n, generator=g, dtype=torch.float64) - 1.5 * c).float() for c in range(nch)])
    win = windows("Hanning", nwins=NFFT, verbose=False)
    return engine, dev, x.to(dev), y.to(dev).contiguous(), win


def _unfused(engine, wp, x, y, rmap, k0, k1, tails, K, nana, scale, onesided):
    wp.reset()
    wp.prepare_tc(x, y, rmap, k0 * HOP, k1 * HOP)
    wp.accumulate_tc(x, y, rmap, k0, k1 - k0)
    for a, b in tails:
        if b > a:
            wp.detrend_stats(x, y, rmap, a, b)
    return wp.finalize(scale, onesided=onesided, K=K, correct_nana=nana)


@pytest.mark.parametrize("onesided", [True, False])
@pytest.mark.parametrize("reflect", [0, NFFT - 1])
@pytest.mark.parametrize("nch", [1, 3])
def test_fused_step_matches_unfused(onesided, reflect, nch):
    n = 1 << 21
    engine, dev, x, y, win = _setup(n, nch)
    nana = n - 100                                    # a tail behind the last segment
    K = (nana + reflect - NFFT) // HOP + 1
    rmap = engine.record_map(nana + reflect, 0 if reflect == 0 else 0, reflect, 0, n)
    nana_t = nana + reflect
    tails = ((0, 0), (K * HOP, nana_t))
    scale = 1.7e-3
    wp = engine.WelchPlan(NFFT, HOP, nch, "fp32", win, dev)
    assert wp.can_use_tc
    ref = _unfused(engine, wp, x, y, rmap, 0, K, tails, K, nana_t, scale, onesided)
    ref = {k: v.clone() for k, v in ref.items()}
    acc_ref, stats_ref = wp.acc.clone(), wp.stats.clone()
    wp2 = engine.WelchPlan(NFFT, HOP, nch, "fp32", win, dev)
    wp2.state.fill_(7.0)                              # the step writes the state, it does not add to it
    out = wp2.step_tc(x, y, rmap, 0, K, tails, K, nana_t, scale, onesided)
    torch.cuda.synchronize()
    assert not wp2.tc_failed()
    pk = float(ref["Pxx"].abs().max())
    for k in ("Pxx", "Pyy", "Pxy"):
        assert float((out[k] - ref[k]).abs().max()) <= 1e-9 * max(pk, float(ref[k].abs().max())), k
    assert float((out["Cxy"] - ref["Cxy"]).abs().max()) < 1e-8
    assert float((out["Cxy2"] - ref["Cxy2"]).abs().max()) < 1e-8
    strong = ref["Pxy"].abs() > 1e-3 * ref["Pxy"].abs().max()
    assert float((out["phi_xy"] - ref["phi_xy"])[strong].abs().max()) < 1e-7
    assert float((wp2.acc - acc_ref).abs().max()) <= 1e-9 * float(acc_ref.abs().max())
    assert torch.allclose(wp2.stats[0::2], stats_ref[0::2], rtol=1e-12, atol=1e-6)
    assert float(wp2.lin[:, :, 8:].abs().max()) == 0.0
    # the data-aware measures of the prepare step are where the host looks for them
    rep = wp2.tc_repetition()
    assert rep.shape == (2, nch + 1) and bool((rep[0] > 1e-2).all())


@pytest.mark.parametrize("world", [2, 3])
def test_fused_step_ranks_in_one_process(world, monkeypatch):
    """world = 2, 3 on one GPU: each 'rank' analyses its share of the segments on its own stream, the peer buffers
    are plain device tensors; all must end with the whole-record result, bit-identical."""
    from pyfft_b200 import _lib
    # every rank's exchange kernel spins until all ranks have pushed: on ONE GPU they must leave SMs for the other
    # ranks' main kernels (a real rank has a GPU to itself and runs one CTA per SM)
    monkeypatch.setenv("PYFFT_POST_BLOCKS", "32")
    nch, n = 2, 1 << 21
    engine, dev, x, y, win = _setup(n, nch, seed=5)
    reflect = NFFT - 1
    nana = n + reflect
    K = (nana - NFFT) // HOP + 1
    rmap = engine.record_map(nana, 0, reflect, 0, n)
    scale = 2.5e-3
    cuts = [K * r // world for r in range(world + 1)]
    spans = [(cuts[r], cuts[r + 1], ((0, 0), (K * HOP, nana) if r == world - 1 else (0, 0))) for r in range(world)]
    # the same decomposition through the unfused calls: every shard about its own provisional statistics, moved
    # to the common reference, the states added, one finalize with est = 0
    total = None
    for k0, k1, tails in spans:
        wp = engine.WelchPlan(NFFT, HOP, nch, "fp32", win, dev)
        wp.reset()
        wp.prepare_tc(x, y, rmap, k0 * HOP, k1 * HOP)
        wp.accumulate_tc(x, y, rmap, k0, k1 - k0)
        for a, b in tails:
            if b > a:
                wp.detrend_stats(x, y, rmap, a, b)
        wp.rereference()
        total = wp.state.clone() if total is None else total + wp.state
    wp.state.copy_(total)
    ref = wp.finalize(scale, onesided=True, K=K, correct_nana=nana)
    ref = {k: v.clone() for k, v in ref.items()}
    nbytes = int(_lib.lib.pyfft_welch_tc_step_peer_bytes(nch, world))
    bufs = [torch.zeros((nbytes + 7) // 8, dtype=torch.float64, device=dev) for _ in range(world)]
    base = torch.tensor([b.data_ptr() for b in bufs], dtype=torch.int64, device=dev)
    plans = [engine.WelchPlan(NFFT, HOP, nch, "fp32", win, dev) for _ in range(world)]
    streams = [torch.cuda.Stream(device=dev) for _ in range(world)]
    outs = [None] * world
    torch.cuda.synchronize()
    for rep in range(3):                              # epochs advance on the device
        for r in range(world):
            with torch.cuda.stream(streams[r]):
                k0, k1, tails = spans[r]
                outs[r] = plans[r].step_tc(x, y, rmap, k0, k1 - k0, tails, K, nana, scale, True,
                                           peer_override=(base, r, world))
        torch.cuda.synchronize()
        for r in range(world):
            assert int(plans[r]._tc_status.item()) == 0
            pk = float(ref["Pxx"].abs().max())
            for k in ("Pxx", "Pyy", "Pxy"):
                assert float((outs[r][k] - ref[k]).abs().max()) <= 1e-9 * max(pk, float(ref[k].abs().max())), (rep, r, k)
            assert float((outs[r]["Cxy"] - ref["Cxy"]).abs().max()) < 1e-8
        for k in ("Pxx", "Pyy", "Pxy", "Cxy", "phi_xy"):
            for r in range(1, world):
                assert torch.equal(outs[0][k], outs[r][k]), k   # rank-ordered sums: bit-identical on every rank


def test_fused_step_deferred_exchange_two_ranks(monkeypatch):
    """Deferred exchange: step i's sums are collected by step i + 1 (or the flush); records that change from step
    to step must come out in order, one step late, and bit-identical on both ranks."""
    from pyfft_b200 import _lib
    monkeypatch.setenv("PYFFT_POST_BLOCKS", "48")
    nch, n = 1, 1 << 20
    reflect = NFFT - 1
    nana = n + reflect
    K = (nana - NFFT) // HOP + 1
    kh = K // 2
    spans = [(0, kh, ((0, 0), (0, 0))), (kh, K, ((0, 0), (K * HOP, nana)))]
    scale = 1.0e-3
    recs, refs = [], []
    for seed in (21, 22, 23):
        engine, dev, x, y, win = _setup(n, nch, seed=seed)
        rmap = engine.record_map(nana, 0, reflect, 0, n)
        total = None
        for k0, k1, tails in spans:
            wp = engine.WelchPlan(NFFT, HOP, nch, "fp32", win, dev)
            wp.reset()
            wp.prepare_tc(x, y, rmap, k0 * HOP, k1 * HOP)
            wp.accumulate_tc(x, y, rmap, k0, k1 - k0)
            for a, b in tails:
                if b > a:
                    wp.detrend_stats(x, y, rmap, a, b)
            wp.rereference()
            total = wp.state.clone() if total is None else total + wp.state
        wp.state.copy_(total)
        refs.append({k: v.clone() for k, v in wp.finalize(scale, onesided=True, K=K, correct_nana=nana).items()})
        recs.append((x, y))
    nbytes = int(_lib.lib.pyfft_welch_tc_step_peer_bytes(nch, 2))
    bufs = [torch.zeros((nbytes + 7) // 8, dtype=torch.float64, device=dev) for _ in range(2)]
    base = torch.tensor([b.data_ptr() for b in bufs], dtype=torch.int64, device=dev)
    plans = [engine.WelchPlan(NFFT, HOP, nch, "fp32", win, dev) for _ in range(2)]
    streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
    torch.cuda.synchronize()

    def check(outs, ref, tag):
        pk = float(ref["Pxx"].abs().max())
        for r in (0, 1):
            assert int(plans[r]._tc_status.item()) == 0
            for k in ("Pxx", "Pyy", "Pxy"):
                assert float((outs[r][k] - ref[k]).abs().max()) <= 1e-9 * max(pk, float(ref[k].abs().max())), (tag, r, k)
        for k in ("Pxx", "Pyy", "Pxy", "Cxy", "phi_xy"):
            assert torch.equal(outs[0][k], outs[1][k]), (tag, k)

    for i, (x, y) in enumerate(recs):
        outs = [None, None]
        for r in (0, 1):
            with torch.cuda.stream(streams[r]):
                k0, k1, tails = spans[r]
                outs[r] = plans[r].step_tc(x, y, rmap, k0, k1 - k0, tails, K, nana, scale, True,
                                           peer_override=(base, r, 2), deferred=True)
        torch.cuda.synchronize()
        if i == 0:
            assert outs[0] is None and outs[1] is None
        else:
            check(outs, refs[i - 1], i)
    outs = [None, None]
    for r in (0, 1):
        with torch.cuda.stream(streams[r]):
            outs[r] = plans[r].flush_tc()
    torch.cuda.synchronize()
    check(outs, refs[-1], "flush")
    assert plans[0].flush_tc() is None
