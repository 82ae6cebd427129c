"""the steps either side of the path -- Doppler.cog / the sliding centre of gravity (fixtures
from the UNMODIFIED reference's cog, tests/golden/cog.npz, `python -m oracle.gen_golden cog`) and the up-sampling
front-end of fftanal (against the numpy restatement: pybaseutils.utils.interp is not in the image, parity unpinned)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden", "cog.npz")
FS = 2.0e6


@pytest.mark.parametrize("tag", ["a", "b", "c", "d"])
def test_cog_vs_reference(tag):
    from pyfft_b200 import cog
    g = np.load(GOLD)
    x = g[tag + "_x"]
    assert abs(cog(x, FS, precision="fp64") - float(g[tag + "_cog"])) <= 1e-9 * FS
    # the reference's band selection pairs the first m bins of the shifted spectrum with the m selected frequencies
    assert abs(cog(x, FS, fmin=5.0e4, fmax=4.0e5, precision="fp64") - float(g[tag + "_cog_band"])) <= 1e-9 * FS
    assert cog(x, FS, fmin=3.0e6, precision="fp64") == 0.0 == float(g[tag + "_cog_none"])
    assert abs(cog(x, FS, precision="fp32") - float(g[tag + "_cog"])) <= 1e-5 * FS


def test_cog_sliding_vs_reference_windows():
    from pyfft_b200 import cog_sliding
    g = np.load(GOLD)
    x = g["s_x"]
    t = np.arange(len(x)) / FS
    tc, c = cog_sliding(t, x, FS, win=512, ov=0.5, precision="fp64")
    assert c.shape == g["s_cog"].shape
    assert np.max(np.abs(c - g["s_cog"])) <= 1e-9 * FS
    assert np.allclose(tc, [t[i:i + 512].mean() for i in range(0, len(x) - 511, 256)], rtol=0, atol=1e-15)
    # a window length that is not a power of two (Bluestein route), checked against the oracle's cog
    from oracle.restatement import cog as cog_port
    tc, c = cog_sliding(t, x, FS, win=300, ov=0.25, precision="fp64")
    ref = np.asarray([cog_port(x[i:i + 300], FS) for i in range(0, len(x) - 299, 225)])
    assert c.shape == ref.shape and np.max(np.abs(c - ref)) <= 1e-8 * FS


@pytest.mark.parametrize("shape", [(1000,), (4097, 3)])
def test_upsample_vs_numpy_restatement(shape):
    from pyfft_b200 import upsample
    from oracle.restatement import upsample as port
    rng = np.random.default_rng(5)
    u = rng.standard_normal(shape)
    for fs, fs_new in ((1.0e5, 2.5e5), (1.0e3, 1.0e3 * np.pi), (48000.0, 96000.0)):
        ref = port(u, fs, fs_new)
        got = upsample(u, fs, fs_new, precision="fp64")
        assert got.shape == ref.shape
        assert np.max(np.abs(got - ref)) <= 1e-12
    got32 = upsample(u.astype(np.float32), 1.0e5, 2.5e5)
    assert got32.dtype == np.float32 and np.max(np.abs(got32 - port(u, 1.0e5, 2.5e5))) <= 1e-6


def test_fftanal_resamples_the_shorter_record():
    """tvecy given: sigy on a slower time base is up-sampled to sigx's (fft_analysis.py:1750-1752, :2363-2373)."""
    from pyfft_b200 import fftanal
    from oracle.restatement import upsample as port
    rng = np.random.default_rng(8)
    n, fsx, fsy = 2 ** 15, 1.0e5, 2.5e4
    tx = np.arange(n) / fsx
    ty = np.arange(n // 4) / fsy
    x = np.sin(2 * np.pi * 3.0e3 * tx) + 0.2 * rng.standard_normal(n)
    y = np.sin(2 * np.pi * 3.0e3 * ty + 0.4) + 0.2 * rng.standard_normal(n // 4)
    o = fftanal(verbose=False)
    tvec, sx, sy = o.resample(tx, x, ty, y)
    assert tvec is tx and sx is x
    assert np.max(np.abs(sy - port(y, fsy, fsx))) <= 1e-12
