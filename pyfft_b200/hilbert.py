"""Drop-in for the reference's hilbert.py: the analytic signal through one long forward FFT,
the bin mask and one long inverse FFT, all inside pyfft_hilbert (K3b).

Same names, arguments and return values as hilbert.py:22-112; the extra keywords
(``precision``, ``device``, ``return_device``, ``envelope``) default to the reference's
behaviour.  Lengths that are not powers of two (or shorter than 16) take Bluestein's identity on the
same long-transform kernels (pyfft_fft_any: forward, the bin mask of hilbert.py:47-64 including its
odd-n form, inverse); there is no CPU fallback.
"""
import numpy as np
import torch

from . import dist as _dist
from . import engine


def _prepare(uin, nfft, axes, precision, device):
    dev = engine.require_cuda(device)
    if not isinstance(uin, torch.Tensor):
        uin = np.atleast_1d(np.asarray(uin))
    prec = engine.resolve_precision(precision, uin)
    u = engine.to_device(uin, prec, dev)
    if u.dim() == 0:
        u = u[None]
    if u.dim() > 2:
        raise NotImplementedError("hilbert: at most 2-D input")
    ax = axes % u.dim()
    if u.dim() == 2 and ax == 0:
        u = u.t()
    u2 = u.reshape(-1, u.shape[-1])
    n = u2.shape[1]
    if nfft is None:
        nfft = n
    nfft = int(nfft)
    if nfft < n:                                    # numpy.fft.fft(u, n=nfft) truncates ...
        u2 = u2[:, :nfft]
    elif nfft > n:                                  # ... or zero-pads (hilbert.py:54)
        u2 = torch.nn.functional.pad(u2, (0, nfft - n))
    return u2, prec, (u.dim() == 2 and ax == 0), u.dim()


def _analytic(u2, prec, want_analytic=True, want_envelope=False):
    """(analytic | None, envelope | None) of the rows of u2: the fused K3b kernel for powers of two,
    else forward transform of any length, mask, inverse transform."""
    n = u2.shape[1]
    if engine._is_pow2(n) and n >= 16:
        return engine.hilbert(u2, prec, want_analytic=want_analytic, want_envelope=want_envelope)
    U = engine.fft_any(u2, prec)                                   # hilbert.py:54
    nyq = (n + 1) // 2 if n % 2 else n // 2                        # :47-51
    mask = torch.ones(n, dtype=U.real.dtype, device=U.device)
    mask[nyq + 1:] = 0.0                                           # :63
    mask[1:nyq] = 2.0                                              # :64
    z = engine.fft_any(U * mask, prec, inverse=True)               # :67
    return (z if want_analytic else None), (z.abs() if want_envelope else None)


def hilbert(uin, nfft=None, axes=-1, precision=None, device=None, return_device=False, envelope=False,
            group=None, gather=False):
    """Analytic signal u + j H[u] (hilbert.py:22-67).  ``envelope=True`` returns |z| instead
    (fused into the last pass; the complex signal is never written).  ``group``: a
    torch.distributed process group -- the records of a 2-D input are sharded over its ranks in
    contiguous blocks (no collective); each rank gets its block back, or the assembled array
    with ``gather=True`` (one all_gather)."""
    u2, prec, transposed, ndim = _prepare(uin, nfft, axes, precision, device)
    nrec = u2.shape[0]
    if group is not None and ndim == 2:
        u2 = _dist.shard_channels(u2, group)
    if u2.shape[0] == 0:
        z = torch.empty((0, u2.shape[1]), dtype=engine._TORCH_CPLX[prec], device=u2.device)
        env = torch.empty((0, u2.shape[1]), dtype=engine._TORCH_REAL[prec], device=u2.device)
    else:
        z, env = _analytic(u2, prec, want_analytic=not envelope, want_envelope=envelope)
    out = env if envelope else z
    if group is not None and ndim == 2 and gather:
        out = _dist.gather_channels(out, nrec, group)
    if transposed:
        out = out.t()
    if ndim == 1:
        out = out[0]
    if return_device:
        return out.squeeze()
    res = out.cpu().numpy()
    res = res.astype(np.float64 if envelope else np.complex128)
    return res.squeeze()                             # hilbert.py:67


def hilbert_1d(uin, nfft=None, precision=None, device=None, return_device=False):
    """hilbert.py:70-112 -- same mask written as an array; 1-D input."""
    if not isinstance(uin, torch.Tensor):
        uin = np.atleast_1d(np.asarray(uin))
    if uin.ndim != 1:
        raise ValueError("hilbert_1d expects a 1-D signal")
    u2, prec, _, _ = _prepare(uin, nfft, -1, precision, device)
    z, _ = _analytic(u2, prec)
    if return_device:
        return z[0]
    return z[0].cpu().numpy().astype(np.complex128)
