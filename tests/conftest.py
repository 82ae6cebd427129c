import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    """npz fixture -> dict; rebuilds tvec when it was stored as (n, fs)."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    d = {k: z[k] for k in z.files}
    if "in_t" not in d and "in_t_n" in d:
        d["in_t"] = np.arange(int(d["in_t_n"])) / float(d["in_t_fs"])
    return d


WELCH_CASES = ["welch_c1_auto", "welch_c1_bincentred", "welch_c2_cross", "welch_c2_dc",
               "welch_multi_sft3f_linear", "welch_defaults", "welch_twosided_complex",
               "welch_tbounds_tper", "welch_odd_nfft", "welch_ref_selftest"]


def relerr(a, b, floor=0.0):
    """max |a - b| / max(max |b|, floor).  `floor` gives the natural scale of quantities that
    can be numerically zero (e.g. 1 - |Cxy|^2 of an auto-spectrum)."""
    a = np.asarray(a)
    b = np.asarray(b)
    assert a.shape == b.shape, (a.shape, b.shape)
    if not b.size:
        return 0.0
    den = max(float(np.max(np.abs(b))), floor)
    return float(np.max(np.abs(a - b)) / (den if den > 0 else 1.0))


# fields whose natural scale is O(1) even when their value is ~0
UNIT_SCALE = {"varCxy": 1.0, "varCxy2": 1.0, "Cxy": 1.0, "Cxy2": 1.0, "iCxy": 1.0, "corrcoef": 1.0}


def phase_err(phi_a, phi_b, pxy):
    """max |Pxy| |e^{i phi_a} - e^{i phi_b}| / max |Pxy|: phase error where there is power."""
    a = np.abs(np.asarray(pxy)).reshape(np.shape(phi_b))
    d = np.abs(np.exp(1j * np.asarray(phi_a)) - np.exp(1j * np.asarray(phi_b)))
    return float(np.max(a * d) / np.max(a))


@pytest.fixture(scope="session")
def cuda_device():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("test marked gpu but no CUDA device is visible")
    return torch.device("cuda", 0)
