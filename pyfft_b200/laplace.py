"""Drop-in for the reference's laplace.py: ``laplace_1d`` -- the "brute-force"
Laplace transform as a batch of real FFTs of the exponentially weighted signal, one row per value of
the real part sigma (laplace.py:85-113).  The weights, the transform of ANY length (Bluestein on the
long-transform kernels) and the rfft bin selection are one call, pyfft_laplace_1d; no CPU fallback.
"""
import ctypes as C

import numpy as np
import torch

from . import engine
from ._lib import check, lib


def laplace():
    pass


def laplace_1d(uin, real_sigma_interval=np.arange(-1, 1 + 0.001, 0.001), nfft=None, precision=None, device=None,
               return_device=False):
    """Rows: the real range (sigma), columns: the imaginary range (rfft bins 0..n/2); complex128.
    ``nfft`` is accepted and ignored exactly like in the reference (it is only assigned there)."""
    dev = engine.require_cuda(device)
    u = np.atleast_1d(np.asarray(uin)) if not isinstance(uin, torch.Tensor) else uin
    prec = engine.resolve_precision(precision, u)
    sig = np.ascontiguousarray(np.asarray(real_sigma_interval, dtype=np.float64).reshape(-1))
    with torch.cuda.device(dev):
        ud = engine.to_device(u, prec, dev).reshape(-1).contiguous()
        n = ud.shape[0]
        sd = torch.from_numpy(sig).to(dev)
        out = torch.empty((len(sig), n // 2 + 1), dtype=engine._TORCH_CPLX[prec], device=dev)
        rows = engine._any_rows(n, len(sig))
        wsb = check(lib.pyfft_anyfft_workspace_bytes(n, rows))
        ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        for r0 in range(0, len(sig), rows):
            rc = min(rows, len(sig) - r0)
            check(lib.pyfft_laplace_1d(engine._CODE[prec], C.c_void_p(ud.data_ptr()), n,
                                       C.c_void_p(sd.data_ptr() + 8 * r0), rc, C.c_void_p(out[r0].data_ptr()),
                                       C.c_void_p(ws.data_ptr()), wsb, st))
        if return_device:
            return out
        return out.cpu().numpy().astype(np.complex128)
