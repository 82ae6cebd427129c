"""Device-side plumbing between the drop-in Python API and the C ABI.

torch is used for device memory, streams and torch.distributed only; every O(N)
computation goes through libpyfft_b200.so.  All methods are asynchronous on the
current torch CUDA stream; nothing here synchronises except explicit ``.cpu()``
conversions done by the callers.
"""
import ctypes as C
import os

import numpy as np
import torch

from . import _lib
from ._lib import RecordMap, Segmentation, check, lib

_TORCH_REAL = {"fp32": torch.float32, "fp64": torch.float64}
_TORCH_CPLX = {"fp32": torch.complex64, "fp64": torch.complex128}
_CODE = {"fp32": _lib.F32, "fp64": _lib.F64}


def bind_host_to_gpu(device=None):
    """Restricts this process to the CPU cores next to `device` (NVML's CPU affinity of the GPU), so that the pinned
    staging buffers it allocates afterwards are first-touched on the GPU's own NUMA node: with one process per GPU,
    eight host->device streams otherwise cross the socket interconnect.  Call it before allocating pinned memory.
    -> sorted list of the cores, or None when NVML / the affinity call is not available (nothing changed)."""
    try:
        import pynvml
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(dev).uuid)
        handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        words = (max(os.cpu_count() or 64, 64) + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(handle, words)
        cores = set(64 * i + b for i, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1)
        cores &= set(os.sched_getaffinity(0))
        if not cores:
            return None
        os.sched_setaffinity(0, cores)
        return sorted(cores)
    except Exception:
        return None


def require_cuda(device=None):
    if not torch.cuda.is_available():
        raise RuntimeError("pyfft_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    if device is None:
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device(device)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def resolve_precision(precision, *arrays):
    """'fp32' | 'fp64' | None.  None: float32 inputs -> fp32, anything else -> fp64 (the
    reference computes in float64 throughout)."""
    if precision is not None:
        p = str(precision).lower()
        p = {"float32": "fp32", "single": "fp32", "f32": "fp32", "float64": "fp64",
             "double": "fp64", "f64": "fp64"}.get(p, p)
        if p not in ("fp32", "fp64"):
            raise ValueError("precision must be 'fp32' or 'fp64', got %r" % (precision,))
        return p
    for a in arrays:
        dt = a.dtype
        if dt in (np.float32, torch.float32, np.complex64, torch.complex64):
            continue
        return "fp64"
    return "fp32"


def to_device(a, precision, device, complex_ok=False):
    """numpy array / torch tensor -> contiguous-on-demand device tensor of the compute dtype."""
    if isinstance(a, torch.Tensor):
        t = a
    else:
        a = np.asarray(a)
        t = torch.from_numpy(a if a.flags.writeable else a.copy())
    if t.is_complex():
        if not complex_ok:
            raise ValueError("complex input not supported on this path")
        want = _TORCH_CPLX[precision]
    else:
        want = _TORCH_REAL[precision]
    return t.to(device=device, dtype=want, non_blocking=True)


_upload_streams = {}


def staged_upload(sigx, sigy, precision, device, nchunks=4, min_samples=1 << 22):
    """Host -> device copy of two long pinned records in `nchunks` pieces on a side stream, with an
    event after each piece, so that the caller can start on the segments of a piece while the next
    one is still crossing PCIe.  -> (x, y, events, bounds) with bounds[i] = samples per channel that
    are on the device once events[i] has fired; None when the inputs do not qualify (not pinned host
    tensors of the compute dtype, short, or time-major 2-D) -- the caller then uses to_device."""
    if not (isinstance(sigx, torch.Tensor) and isinstance(sigy, torch.Tensor)):
        return None
    want = _TORCH_REAL[precision]
    if sigx.is_cuda or sigy.is_cuda or sigx.dtype != want or sigy.dtype != want:
        return None
    if not (sigx.is_pinned() and sigy.is_pinned()) or sigx.dim() != 1:
        return None
    n = sigx.shape[0]
    if n < min_samples or sigy.dim() not in (1, 2) or sigy.shape[-1] != n or not sigy.is_contiguous() \
            or not sigx.is_contiguous():
        return None
    cur = torch.cuda.current_stream(device)
    key = str(device)
    side = _upload_streams.get(key)
    if side is None:
        side = _upload_streams[key] = torch.cuda.Stream(device=device)
    x = torch.empty(n, dtype=want, device=device)
    y = torch.empty(tuple(sigy.shape), dtype=want, device=device)
    side.wait_stream(cur)
    events, bounds = [], []
    step = (n + nchunks - 1) // nchunks
    step = (step + 1023) // 1024 * 1024
    with torch.cuda.stream(side):
        a = 0
        while a < n:
            b = min(n, a + step)
            x[a:b].copy_(sigx[a:b], non_blocking=True)
            y[..., a:b].copy_(sigy[..., a:b], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(side)
            events.append(ev)
            bounds.append(b)
            a = b
    x.record_stream(side)
    y.record_stream(side)
    return x, y, events, bounds


def channels_first(y, n):
    """(n,), (n, C) or (C, n) device tensor -> contiguous (C, n).  Mirrors the reference's
    orientation rule: a 2-D sigy whose second axis has the record length is transposed
    (fft_analysis.py:163-167)."""
    if y.dim() == 1:
        y = y[None, :]
    elif y.shape[1] != n:
        y = y.t()            # (n, C) time-major, the reference's native layout
    return y.contiguous()


class PeerReducer:
    """Sum over the ranks of a process group of one packed double buffer, by ONE kernel over
    NVLink peer memory (csrc/peer.cu) instead of a library collective.  The symmetric buffer
    (two staging copies + flags) comes from torch.distributed._symmetric_memory, which peer-maps
    it into every rank of the group; PYFFT_PEER_REDUCE=0 disables it."""

    _cache = {}

    def __init__(self, n, device, group):
        import torch.distributed._symmetric_memory as symm
        dist = torch.distributed
        group = dist.group.WORLD if group is None else group
        self.n = int(n)
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        nbytes = check(lib.pyfft_peer_buffer_bytes(self.n, self.world))
        self.buf = symm.empty((nbytes + 7) // 8, dtype=torch.float64, device=device)
        self.buf.zero_()
        self.hdl = symm.rendezvous(self.buf, group.group_name)
        self.peer_base = torch.tensor([int(p) for p in self.hdl.buffer_ptrs], dtype=torch.int64, device=device)
        self.ticket = torch.zeros(2, dtype=torch.int32, device=device)     # [1]: device-side epoch counter
        self.status = torch.zeros(1, dtype=torch.int32, device=device)
        self.epoch = 0
        torch.cuda.synchronize(device)
        dist.barrier(group=group)                      # every rank's buffer is zeroed before the first epoch

    @classmethod
    def create(cls, n, device, group):
        """-> PeerReducer, or False when peer memory is not available (single process, CPU group,
        no peer access): the caller keeps using NCCL."""
        if os.environ.get("PYFFT_PEER_REDUCE", "1") == "0":
            return False
        dist = torch.distributed
        if not (dist.is_available() and dist.is_initialized()):
            return False
        g = dist.group.WORLD if group is None else group
        if dist.get_world_size(g) < 2 or dist.get_backend(g) != "nccl":
            return False
        key = (int(n), str(device), g.group_name)
        if key not in cls._cache:
            try:
                ok = cls(n, device, g)
            except Exception as exc:                   # no peer mapping on this machine
                import warnings
                warnings.warn("pyfft_b200: NVLink peer reduction unavailable (%s); using NCCL" % (exc,))
                ok = False
            # all ranks must take the same route
            flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=g)
            cls._cache[key] = ok if int(flag.item()) == 1 else False
        return cls._cache[key]

    def all_reduce(self, state):
        # the epoch is kept on the device (ticket[1]) and advanced by the kernel: the launch is the same
        # every step and can be replayed from a CUDA graph
        self.epoch += 1
        check(lib.pyfft_peer_allreduce(_ptr(state), self.n, _ptr(self.peer_base), self.rank, self.world,
                                       0, _ptr(self.ticket), _ptr(self.status), _stream()))

    def check_status(self):
        """Synchronising check of the bounded waits (tests / debugging)."""
        if int(self.status.item()) != 0:
            raise RuntimeError("pyfft_b200: a peer did not reach the NVLink reduction in time")


class PeerStepBuffer:
    """Symmetric staging buffer of the fused sharded step (pyfft_welch_tc_step, world > 1): two staging copies per
    rank and per 32-bin work item, per-item flags and epochs; peer-mapped like PeerReducer's."""

    _cache = {}

    def __init__(self, nch, device, group):
        import torch.distributed._symmetric_memory as symm
        dist = torch.distributed
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        nbytes = check(lib.pyfft_welch_tc_step_peer_bytes(int(nch), self.world))
        self.buf = symm.empty((nbytes + 7) // 8, dtype=torch.float64, device=device)
        self.buf.zero_()
        self.hdl = symm.rendezvous(self.buf, group.group_name)
        self.peer_base = torch.tensor([int(p) for p in self.hdl.buffer_ptrs], dtype=torch.int64, device=device)
        torch.cuda.synchronize(device)
        dist.barrier(group=group)                      # every rank's flags are zero before the first epoch

    @classmethod
    def create(cls, nch, device, group):
        """-> PeerStepBuffer, or False when peer memory is not available (the caller keeps the unfused sequence)."""
        if os.environ.get("PYFFT_PEER_REDUCE", "1") == "0":
            return False
        dist = torch.distributed
        if not (dist.is_available() and dist.is_initialized()):
            return False
        g = dist.group.WORLD if group is None else group
        if dist.get_world_size(g) < 2 or dist.get_backend(g) != "nccl":
            return False
        key = (int(nch), str(device), g.group_name)
        if key not in cls._cache:
            try:
                ok = cls(nch, device, g)
            except Exception as exc:                   # no peer mapping on this machine
                import warnings
                warnings.warn("pyfft_b200: NVLink peer exchange unavailable (%s); using the unfused step" % (exc,))
                ok = False
            flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=g)          # all ranks must take the same route
            cls._cache[key] = ok if int(flag.item()) == 1 else False
        return cls._cache[key]


class WelchPlan(object):
    """Streaming Welch accumulator for one x channel against nch y channels.

    Mirrors the structure SURVEY.md section 5 asks for: partial sums plus a segment
    count, so long records can be fed shard by shard (or rank by rank) and
    finalised once.
    """

    def __init__(self, nfft, hop, nch, precision, window, device=None):
        self.device = require_cuda(device)
        self.nfft, self.hop, self.nch = int(nfft), int(hop), int(nch)
        self.precision = precision
        self.code = _CODE[precision]
        self.nbin = self.nfft // 2 + 1
        win = np.ascontiguousarray(np.asarray(window, dtype=np.float64))
        if win.shape != (self.nfft,):
            raise ValueError("window must have nfft = %d points" % self.nfft)
        self.window = torch.from_numpy(win).to(self.device, _TORCH_REAL[precision])
        # acc | lin | stats share one buffer: one fill resets them, one all-reduce adds shards
        self.nlin = self.nfft // 16
        n_acc, n_lin, n_st = self.nch * 4 * self.nbin, self.nch * 4 * self.nlin, 2 * (self.nch + 1)
        self._peer = None                              # PeerReducer, False (unavailable) or None (not tried)
        self._peer_group = None
        self.state = torch.zeros(n_acc + n_lin + n_st, dtype=torch.float64, device=self.device)
        self.acc = self.state[:n_acc].view(self.nch, 4, self.nbin)
        self.lin = self.state[n_acc:n_acc + n_lin].view(self.nch, 4, self.nlin)
        self.stats = self.state[n_acc + n_lin:]
        self.coef = None
        self.K = 0
        self._ws = None
        # mean-fused route (pyfft_welch_accumulate_mean): usable when the window's transform
        # is negligible beyond the nfft/16 bins the kernel keeps linear sums for
        W = np.fft.fft(win)
        p = np.abs(W) ** 2
        leak = float(p[self.nlin:self.nfft // 2 + 1].max() / (win ** 2).sum()) if self.nfft >= 1024 else 1.0
        self.can_fuse_mean = self.nfft >= 1024 and leak < (1e-9 if precision == "fp32" else 1e-18)
        self.wlow_host = np.ascontiguousarray(np.stack([W.real[:self.nlin], W.imag[:self.nlin]], 1))
        self.wlow = torch.from_numpy(self.wlow_host).to(self.device)
        self.est = None
        self.K_local = 0
        # tensor-core route (pyfft_welch_accumulate_tc): nfft 4096, fp32, hop 512 / 1024 / 2048 / 4096, window
        # transform negligible beyond the 8 bins it keeps linear sums for
        leak8 = float(p[8:self.nfft // 2 + 1].max() / (win ** 2).sum()) if self.nfft == 4096 else 1.0
        self.can_use_tc = (self.nfft == 4096 and precision == "fp32" and self.hop in (512, 1024, 2048, 4096)
                           and leak8 < 1e-9)
        self.prescale = None
        self._tc_status = None
        self._zero_est = None
        self._tc_est = None
        self._aux = None
        self._deferred = None
        self._deferred_args = None

    def reset(self):
        """Zeroes acc, lin, stats and the segment counts (one fill)."""
        self.state.zero_()
        self.K = self.K_local = 0

    # -- detrend ------------------------------------------------------------ #
    def detrend_stats(self, x, y, rmap, j_begin, j_end):
        """Adds sum x(j), sum j x(j) for j in [j_begin, j_end) (x first, then the y channels)."""
        check(lib.pyfft_detrend_stats_xy(self.code, _ptr(x), _ptr(y), y.stride(0), self.nch, C.byref(rmap),
                                         j_begin, j_end, _ptr(self.stats), _stream()))

    def solve_detrend(self, mode, nsig, group=None):
        if mode == _lib.DETREND_NONE:
            self.coef = None
            return
        if group is not None:
            torch.distributed.all_reduce(self.stats, group=group)
        self.coef = torch.empty_like(self.stats)
        check(lib.pyfft_detrend_coef(mode, _ptr(self.stats), self.nch + 1, nsig, _ptr(self.coef),
                                     _stream()))

    # -- mean detrend fused into the accumulation ---------------------------- #
    def estimate_mean(self, x, y, rmap, j_begin, j_end, nsample=65536):
        """Provisional means from the first `nsample` analysed samples of this shard (any
        estimate works: the exact global mean is restored by mean_correct)."""
        j1 = min(j_end, j_begin + nsample)
        tmp = torch.zeros_like(self.stats)
        check(lib.pyfft_detrend_stats(self.code, _ptr(x), 0, 1, C.byref(rmap), j_begin, j1, _ptr(tmp),
                                      _stream()))
        check(lib.pyfft_detrend_stats(self.code, _ptr(y), y.stride(0), self.nch, C.byref(rmap), j_begin, j1,
                                      C.c_void_p(tmp.data_ptr() + 16), _stream()))
        self.est = torch.empty_like(tmp)
        check(lib.pyfft_detrend_coef(_lib.DETREND_MEAN, _ptr(tmp), self.nch + 1, j1 - j_begin,
                                     _ptr(self.est), _stream()))

    def accumulate_mean(self, x, y, rmap, seg_first, seg_count):
        """K1 with the mean removal fused: adds to acc, lin and the raw sums in stats."""
        seg = Segmentation(self.nfft, self.hop, seg_first, seg_count)
        need = check(lib.pyfft_welch_workspace_bytes(self.code, C.byref(seg), self.nch))
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        check(lib.pyfft_welch_accumulate_mean(self.code, _ptr(x), _ptr(y), y.stride(0), self.nch,
                                              C.byref(rmap), C.byref(seg), _ptr(self.window), _ptr(self.est),
                                              _ptr(self.acc), _ptr(self.lin), _ptr(self.stats),
                                              _ptr(self._ws), self._ws.numel(), _stream()))
        self.K += int(seg_count)
        self.K_local += int(seg_count)

    # -- tensor-core route -------------------------------------------------- #
    def prepare_tc(self, x, y, rmap, j_begin, j_end):
        """Provisional means and power-of-two scales from 256 runs of 256 samples spread evenly over
        the analysed indices [j_begin, j_end) of this shard."""
        j1 = j_end
        if self._tc_est is None:                       # zeroed by the entry point itself
            self._tc_est = torch.empty(2 * (self.nch + 1), dtype=torch.float64, device=self.device)
            self.prescale = torch.empty(16 * (self.nch + 1), dtype=torch.float64, device=self.device)
        self.est = self._tc_est
        check(lib.pyfft_welch_tc_prepare(_ptr(x), _ptr(y), y.stride(0), self.nch, C.byref(rmap), j_begin, j1,
                                         self.hop, _ptr(self.est), _ptr(self.prescale), _stream()))

    def tc_repetition(self):
        """(2, nch+1) device view: repetition measure and stationarity of every channel from the last
        prepare_tc (include/pyfft_b200.h)."""
        return self.prescale[self.nch + 1:3 * (self.nch + 1)].view(2, self.nch + 1)

    def mean_offset(self, nana):
        """(nch+1,) device tensor: exact mean of the analysed record minus the provisional mean the fused
        kernels subtracted (stats must cover the whole record; call before rereference)."""
        return self.stats[0::2] / float(nana) - self.est[0::2]

    def accumulate_tc(self, x, y, rmap, seg_first, seg_count):
        """K1 on the tensor cores (tcgen05): adds to acc, lin and the raw sums in stats."""
        seg = Segmentation(self.nfft, self.hop, seg_first, seg_count)
        need = check(lib.pyfft_welch_tc_workspace_bytes(C.byref(seg), self.nch))
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        if self._tc_status is None:
            self._tc_status = torch.zeros(1, dtype=torch.int32, device=self.device)
        check(lib.pyfft_welch_accumulate_tc(_ptr(x), _ptr(y), y.stride(0), self.nch, C.byref(rmap),
                                            C.byref(seg), _ptr(self.window), _ptr(self.est),
                                            _ptr(self.prescale), _ptr(self.acc), _ptr(self.lin),
                                            _ptr(self.stats), _ptr(self._ws), self._ws.numel(),
                                            _ptr(self._tc_status), _stream()))
        self.K += int(seg_count)
        self.K_local += int(seg_count)

    def step_tc(self, x, y, rmap, seg_first, seg_count, tails, K, nana, scale, onesided=True, group=None,
                peer_override=None, deferred=False):
        """The whole fused tensor-core step over resident records in three launches (pyfft_welch_tc_step):
        provisional statistics + head + tail sums, the main kernel, and ONE kernel that sums the partials,
        exchanges them with the other ranks of `group` over NVLink peer memory, writes `state` and the
        finalised spectra.  tails = ((a0, a1), (b0, b1)): analysed-index spans outside [seg_first*hop,
        (seg_first+seg_count)*hop) whose raw sums the exact mean needs.  -> the dict of finalize(), or None when
        the group has no peer memory (nothing was launched: use the unfused sequence).
        peer_override = (peer_base int64 device tensor, rank, world): tests drive several ranks from one process.
        deferred (sharded streams of steps): the exchange pushed by this call is collected by the NEXT deferred call
        or by flush_tc(), a whole main kernel later -- no rank waits for a slower one inside a step.  The dict
        returned then holds the PREVIOUS deferred step's spectra once this call's kernels have run (None for the
        first call of a stream); flush_tc() returns the last step's."""
        world, rank, peer = 1, 0, None
        if peer_override is not None:
            class _P(object):
                pass
            peer = _P()
            peer.peer_base, rank, world = peer_override
        elif group is not None and torch.distributed.get_world_size(group) > 1:
            peer = PeerStepBuffer.create(self.nch, self.device, group)
            if not peer:
                return None
            world, rank = peer.world, peer.rank
        n1 = self.nch + 1
        if self._aux is None:
            self._aux = torch.empty(21 * n1, dtype=torch.float64, device=self.device)
        self.est, self.prescale = self._aux[:2 * n1], self._aux[2 * n1:18 * n1]
        seg = Segmentation(self.nfft, self.hop, seg_first, seg_count)
        need = check(lib.pyfft_welch_tc_workspace_bytes(C.byref(seg), self.nch))
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        if self._tc_status is None:
            self._tc_status = torch.zeros(1, dtype=torch.int32, device=self.device)
        F = self.nfft // 2 if onesided else self.nfft
        dev = self.device
        out = dict(Pxx=torch.empty(F, dtype=torch.complex128, device=dev),
                   Pyy=torch.empty((F, self.nch), dtype=torch.complex128, device=dev),
                   Pxy=torch.empty((F, self.nch), dtype=torch.complex128, device=dev),
                   Cxy=torch.empty((F, self.nch), dtype=torch.complex128, device=dev),
                   Cxy2=torch.empty((F, self.nch), dtype=torch.complex128, device=dev),
                   phi_xy=torch.empty((F, self.nch), dtype=torch.float64, device=dev))
        (a0, a1), (b0, b1) = tails
        check(lib.pyfft_welch_tc_step(_ptr(x), _ptr(y), y.stride(0), self.nch, C.byref(rmap), C.byref(seg),
                                      _ptr(self.window), _ptr(self.wlow), int(a0), int(a1), int(b0), int(b1),
                                      int(K), int(nana), float(scale), 1 if onesided else 0, _ptr(self._aux),
                                      _ptr(self.state), _ptr(out["Pxx"]), _ptr(out["Pyy"]), _ptr(out["Pxy"]),
                                      _ptr(out["Cxy"]), _ptr(out["Cxy2"]), _ptr(out["phi_xy"]), _ptr(self._ws),
                                      self._ws.numel(), _ptr(self._tc_status),
                                      _ptr(peer.peer_base) if peer else C.c_void_p(0), rank, world,
                                      1 if (deferred and world > 1) else 0, _stream()))
        self.K = int(K)
        self.K_local = int(seg_count)
        self.coef = None
        if deferred and world > 1:
            first = self._deferred is None
            self._deferred = self._deferred_args = (peer, rank, world, int(K), int(nana), float(scale), bool(onesided))
            return None if first else out
        if self._deferred is not None:
            raise RuntimeError("step_tc: a deferred step is pending; call flush_tc() before a plain step")
        return out

    def flush_tc(self, force=False):
        """Collects the exchange of the last deferred step_tc -> its dict of spectra (None if nothing is pending).
        force: the deferred steps were replayed from a CUDA graph, which this object does not see -- launch the
        flush with the arguments of the last deferred call (the device keeps the pending flags itself)."""
        if self._deferred is None and not (force and self._deferred_args is not None):
            return None
        peer, rank, world, K, nana, scale, onesided = self._deferred or self._deferred_args
        self._deferred = None
        F = self.nfft // 2 if onesided else self.nfft
        dev = self.device
        out = dict(Pxx=torch.empty(F, dtype=torch.complex128, device=dev),
                   Pyy=torch.empty((F, self.nch), dtype=torch.complex128, device=dev),
                   Pxy=torch.empty((F, self.nch), dtype=torch.complex128, device=dev),
                   Cxy=torch.empty((F, self.nch), dtype=torch.complex128, device=dev),
                   Cxy2=torch.empty((F, self.nch), dtype=torch.complex128, device=dev),
                   phi_xy=torch.empty((F, self.nch), dtype=torch.float64, device=dev))
        check(lib.pyfft_welch_tc_step_flush(self.nch, _ptr(self.wlow), K, nana, scale, 1 if onesided else 0,
                                            _ptr(self.state), _ptr(out["Pxx"]), _ptr(out["Pyy"]), _ptr(out["Pxy"]),
                                            _ptr(out["Cxy"]), _ptr(out["Cxy2"]), _ptr(out["phi_xy"]),
                                            _ptr(self._tc_status), _ptr(peer.peer_base), rank, world, _stream()))
        return out

    def tc_failed(self):
        """True when a tensor-core barrier wait timed out (synchronises)."""
        return self._tc_status is not None and int(self._tc_status.item()) != 0

    def mean_correct(self, nana, group=None):
        """stats must cover the whole analysed record (all-reduced here when sharded)."""
        if group is not None:
            torch.distributed.all_reduce(self.stats, group=group)
        check(lib.pyfft_welch_mean_correct(_ptr(self.acc), _ptr(self.lin), self.nfft, self.nch,
                                           self.K_local, _ptr(self.est), _ptr(self.stats), int(nana),
                                           _ptr(self.wlow), _stream()))
        self.coef = None

    def rereference(self):
        """Sharded fused-mean route: moves this shard's sums to the common reference 0, so ONE
        all-reduce of `state` (all_reduce_state) followed by finalize(..., correct_nana=nana)
        replaces the pair (all-reduce stats + mean_correct, all-reduce acc)."""
        check(lib.pyfft_welch_rereference(_ptr(self.acc), _ptr(self.lin), self.nfft, self.nch, self.K_local,
                                          _ptr(self.est), _ptr(self.wlow), _stream()))
        if self._zero_est is None:
            self._zero_est = torch.zeros(2 * (self.nch + 1), dtype=torch.float64, device=self.device)
        self.est = self._zero_est

    def all_reduce_state(self, group=None):
        """The one collective of the sharded fused-mean route: acc | lin | stats in one buffer.
        Over NVLink peer memory (PeerReducer) when the group's GPUs can map each other's memory,
        else NCCL.  peer_status() is the device flag of the bounded waits (0 = fine)."""
        gname = getattr(group, "group_name", None)
        if self._peer is None or self._peer_group != gname:
            self._peer = PeerReducer.create(self.state.numel(), self.device, group)
            self._peer_group = gname
        if self._peer:
            self._peer.all_reduce(self.state)
        else:
            torch.distributed.all_reduce(self.state, group=group)

    def peer_status(self):
        """Device int32 tensor raised by a timed-out NVLink reduction, or None (NCCL route / unsharded)."""
        return self._peer.status if self._peer else None

    def dynamic_range(self, res, onesided=True):
        """-> (nch+1, 2) double device tensor: min and mean bin power of Pxx and of every Pyy column
        (pyfft_welch_dynamic_range) -- the data-aware check of the tensor-core route."""
        out = torch.empty((self.nch + 1, 2), dtype=torch.float64, device=self.device)
        check(lib.pyfft_welch_dynamic_range(_ptr(res["Pxx"]), _ptr(res["Pyy"]), res["Pxx"].shape[0], self.nch,
                                            1 if onesided else 0, _ptr(out), _stream()))
        return out

    # -- accumulate --------------------------------------------------------- #
    def accumulate(self, x, y, rmap, seg_first, seg_count):
        seg = Segmentation(self.nfft, self.hop, seg_first, seg_count)
        need = check(lib.pyfft_welch_workspace_bytes(self.code, C.byref(seg), self.nch))
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        cx = C.c_void_p(self.coef.data_ptr()) if self.coef is not None else C.c_void_p(0)
        cy = C.c_void_p(self.coef.data_ptr() + 16) if self.coef is not None else C.c_void_p(0)
        check(lib.pyfft_welch_accumulate(self.code, _ptr(x), _ptr(y), y.stride(0), self.nch,
                                         C.byref(rmap), C.byref(seg), _ptr(self.window), cx, cy,
                                         _ptr(self.acc), _ptr(self._ws), self._ws.numel(), _stream()))
        self.K += int(seg_count)

    def all_reduce(self, group=None):
        """One collective over the packed partial sums."""
        torch.distributed.all_reduce(self.acc, group=group)
        k = torch.tensor([self.K], dtype=torch.int64, device=self.device)
        torch.distributed.all_reduce(k, group=group)
        self.K = int(k.item())

    # -- finalize ----------------------------------------------------------- #
    def finalize(self, scale, onesided=True, K=None, correct_nana=None):
        """-> dict of device tensors Pxx (F,), Pyy/Pxy/Cxy/Cxy2 (F, nch) complex128, phi (F, nch).
        correct_nana: unsharded fused-mean route -- apply the mean correction (instead of a separate
        mean_correct call) while acc is read; stats must cover the whole analysed record."""
        K = self.K if K is None else int(K)
        F = self.nfft // 2 if onesided else self.nfft
        dev = self.device
        out = dict(Pxx=torch.empty(F, dtype=torch.complex128, device=dev),
                   Pyy=torch.empty((F, self.nch), dtype=torch.complex128, device=dev),
                   Pxy=torch.empty((F, self.nch), dtype=torch.complex128, device=dev),
                   Cxy=torch.empty((F, self.nch), dtype=torch.complex128, device=dev),
                   Cxy2=torch.empty((F, self.nch), dtype=torch.complex128, device=dev),
                   phi_xy=torch.empty((F, self.nch), dtype=torch.float64, device=dev))
        if correct_nana is not None:
            check(lib.pyfft_welch_finalize_mean(_ptr(self.acc), _ptr(self.lin), self.nfft, self.nch, K,
                                                _ptr(self.est), _ptr(self.stats), int(correct_nana),
                                                _ptr(self.wlow), float(scale), 1 if onesided else 0,
                                                _ptr(out["Pxx"]), _ptr(out["Pyy"]), _ptr(out["Pxy"]),
                                                _ptr(out["Cxy"]), _ptr(out["Cxy2"]), _ptr(out["phi_xy"]),
                                                _stream()))
            self.coef = None
            return out
        check(lib.pyfft_welch_finalize(_ptr(self.acc), self.nfft, self.nch, K, float(scale),
                                       1 if onesided else 0, _ptr(out["Pxx"]), _ptr(out["Pyy"]),
                                       _ptr(out["Pxy"]), _ptr(out["Cxy"]), _ptr(out["Cxy2"]),
                                       _ptr(out["phi_xy"]), _stream()))
        return out

    def correlation(self, spec, onesided=True, halve_interior=True):
        """sqrt(nfft) * irfft / ifft(ifftshift) of a (F,) or (F, ncol) complex128 device tensor,
        fftshift-ed along axis 0 (fft_analysis.py:544-595)."""
        ncol = 1 if spec.dim() == 1 else spec.shape[1]
        shape = (self.nfft,) if spec.dim() == 1 else (self.nfft, ncol)
        out = torch.empty(shape, dtype=torch.float64 if onesided else torch.complex128,
                          device=self.device)
        spec = spec.contiguous()
        check(lib.pyfft_welch_correlation(_ptr(spec), self.nfft, ncol, 1 if onesided else 0,
                                          1 if halve_interior else 0, _ptr(out), _stream()))
        return out


class GraphedStep(object):
    """A fixed sequence of launches (a bench / serving step over resident buffers) captured ONCE into a CUDA
    graph and replayed: the small kernels around K1 then cost their device time only, not a Python + ctypes
    round trip each.  fn must allocate nothing new per call except its outputs (kept alive by the graph's
    pool) and must have run before (twiddle / constant tables are uploaded on first use)."""

    def __init__(self, fn, warmup=3):
        dev = torch.cuda.current_device()
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):
                fn()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.out = fn()

    def __call__(self):
        self.graph.replay()
        return self.out


def record_map(nsig, i0, reflect, src_base, src_len):
    return RecordMap(int(nsig), int(i0), int(reflect), int(src_base), int(src_len))


# --------------------------------------------------------------------------- #
# K2: STFT with materialised per-segment spectra
# --------------------------------------------------------------------------- #

def detrend_coefficients(x, rmap, mode, precision, nsig, group=None, j_span=None):
    """Global trend coefficients (a, b) per channel of a (C, n) device tensor -> double[2C]
    device tensor or None.  x(j) - (a + b j) is what pybaseutils.detrend_mean/linear remove
    (fft_analysis.py:2147-2149)."""
    if mode == _lib.DETREND_NONE:
        return None
    nch = x.shape[0]
    stats = torch.zeros(2 * nch, dtype=torch.float64, device=x.device)
    j0, j1 = (0, nsig) if j_span is None else j_span
    check(lib.pyfft_detrend_stats(_CODE[precision], _ptr(x), x.stride(0), nch, C.byref(rmap), j0, j1,
                                  _ptr(stats), _stream()))
    if group is not None:
        torch.distributed.all_reduce(stats, group=group)
    coef = torch.empty_like(stats)
    check(lib.pyfft_detrend_coef(mode, _ptr(stats), nch, nsig, _ptr(coef), _stream()))
    return coef


def pow2_window(nfft):
    """True when the in-CTA power-of-two transforms (K1 / K2 / K4) take this window length."""
    nfft = int(nfft)
    return 16 <= nfft <= 8192 and (nfft & (nfft - 1)) == 0


_ANY_BUDGET = 1 << 30          # bytes of Bluestein workspace per call of the arbitrary-length route


def _any_rows(n, want):
    """Rows per pyfft_stft_any / pyfft_fft_any call so that its workspace stays near _ANY_BUDGET."""
    per_row = check(lib.pyfft_anyfft_workspace_bytes(int(n), 2)) - check(lib.pyfft_anyfft_workspace_bytes(int(n), 1))
    return int(max(1, min(want, 65535, _ANY_BUDGET // max(per_row, 1))))


def stft(x, rmap, nfft, hop, seg_first, seg_count, window, precision, coef=None,
         per_segment=False, bin_first=0, nbins=None, bin_scale=None, power_out=False,
         want_pseg=False):
    """x: (C, n) real device tensor (or complex, interleaved) -> (out [C][K][nbins], pseg [C][K] | None).

    out is complex64/128 (or float32/64 when power_out) following `precision`.  Window lengths that are
    powers of two in 16..8192 run the fused in-CTA transform (pyfft_stft); every other length goes
    through Bluestein's identity on the long-transform kernels (pyfft_stft_any, float64 arithmetic)."""
    dev = x.device
    cplx = x.is_complex()
    nch = x.shape[0]
    nbins = nfft if nbins is None else int(nbins)
    odt = _TORCH_REAL[precision] if power_out else _TORCH_CPLX[precision]
    out = torch.empty((nch, int(seg_count), nbins), dtype=odt, device=dev)
    pseg = torch.zeros((nch, int(seg_count)), dtype=torch.float64, device=dev) if want_pseg else None
    scale = None
    if bin_scale is not None:
        scale = torch.from_numpy(np.ascontiguousarray(np.asarray(bin_scale, dtype=np.float64))).to(dev)
        if scale.numel() != nbins:
            raise ValueError("bin_scale needs nbins entries")
    xr = torch.view_as_real(x) if cplx else x
    stride = x.stride(0) * (2 if cplx else 1)                # real elements between channels
    if pow2_window(nfft):
        win = torch.from_numpy(np.ascontiguousarray(np.asarray(window, dtype=np.float64))).to(
            dev, _TORCH_REAL[precision])
        seg = Segmentation(int(nfft), int(hop), int(seg_first), int(seg_count))
        check(lib.pyfft_stft(_CODE[precision], 1 if cplx else 0, _ptr(xr), stride, nch,
                             C.byref(rmap), C.byref(seg), _ptr(win), _ptr(coef), 1 if per_segment else 0,
                             int(bin_first), nbins, _ptr(scale), 1 if power_out else 0, _ptr(out), _ptr(pseg),
                             _stream()))
        return out, pseg
    # arbitrary window length: one channel and a chunk of segments per call, written in place
    win = torch.from_numpy(np.ascontiguousarray(np.asarray(window, dtype=np.float64))).to(dev)
    rows = _any_rows(nfft, int(seg_count))
    wsb = check(lib.pyfft_anyfft_workspace_bytes(int(nfft), rows))
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    esz = xr.element_size()
    cstep = 4 if cplx else 2                                  # doubles of coef per channel
    for c in range(nch):
        for k0 in range(0, int(seg_count), rows):
            kc = min(rows, int(seg_count) - k0)
            seg = Segmentation(int(nfft), int(hop), int(seg_first) + k0, kc)
            o = out[c, k0:k0 + kc]
            check(lib.pyfft_stft_any(_CODE[precision], 1 if cplx else 0, C.c_void_p(xr.data_ptr() + c * stride * esz),
                                     stride, 1, C.byref(rmap), C.byref(seg), _ptr(win),
                                     C.c_void_p(coef.data_ptr() + 8 * cstep * c) if coef is not None else C.c_void_p(0),
                                     1 if per_segment else 0, int(bin_first), nbins, _ptr(scale),
                                     1 if power_out else 0, C.c_void_p(o.data_ptr()),
                                     C.c_void_p(pseg.data_ptr() + 8 * (c * int(seg_count) + k0)) if pseg is not None
                                     else C.c_void_p(0), _ptr(ws), wsb, _stream()))
    return out, pseg


def fft_any(z, precision, inverse=False):
    """(batch, n) real or complex device tensor, ANY n -> numpy.fft.fft / ifft along the last axis
    (complex of `precision`); powers of two >= 16 with complex input take pyfft_fft_c2c directly."""
    batch, n = z.shape
    z = z.contiguous()
    if z.is_complex() and _is_pow2(n) and n >= 16:
        return fft_c2c(z, precision, inverse=inverse)
    out = torch.empty((batch, n), dtype=_TORCH_CPLX[precision], device=z.device)
    rows = _any_rows(n, batch)
    wsb = check(lib.pyfft_anyfft_workspace_bytes(int(n), rows))
    ws = torch.empty(wsb, dtype=torch.uint8, device=z.device)
    zr = torch.view_as_real(z) if z.is_complex() else z
    for r0 in range(0, batch, rows):
        rc = min(rows, batch - r0)
        check(lib.pyfft_fft_any(_CODE[precision], 1 if z.is_complex() else 0, _ptr(zr[r0]), n, n, rc,
                                1 if inverse else 0, C.c_void_p(out[r0].data_ptr()), _ptr(ws), wsb, _stream()))
    return out


def stft_sharded(x, group, rmap, nfft, hop, seg_first, seg_count, window, precision, gather=False, **kw):
    """Channel-sharded stft: rank r of `group` transforms its contiguous
    block of the rows of x and keeps the result (no collective); gather=True assembles
    [C][K][nbins] on every rank with one all_gather."""
    from . import dist as _dist
    xl = _dist.shard_channels(x, group)
    if xl.shape[0] == 0:
        nb = int(kw.get("nbins") or nfft)
        odt = _TORCH_REAL[precision] if kw.get("power_out") else _TORCH_CPLX[precision]
        out, pseg = torch.empty((0, int(seg_count), nb), dtype=odt, device=x.device), None
    else:
        out, pseg = stft(xl, rmap, nfft, hop, seg_first, seg_count, window, precision, **kw)
    if gather:
        out = _dist.gather_channels(out, x.shape[0], group)
        if pseg is not None:
            pseg = _dist.gather_channels(pseg, x.shape[0], group)
    return out, pseg


def stft_products(X, Y, conj_x, precision):
    """X [K][nb], Y [nch][K][nb] (or None) complex device tensors -> double sums
    (sxx [nb], syy [nch][nb], sxy [nch][nb] complex128) over the K segments."""
    K, nb = X.shape
    dev = X.device
    sxx = torch.empty(nb, dtype=torch.float64, device=dev)
    if Y is None:
        check(lib.pyfft_stft_products(_CODE[precision], _ptr(X), None, 1, K, nb, 1 if conj_x else 0,
                                      _ptr(sxx), None, None, _stream()))
        return sxx, None, None
    nch = Y.shape[0]
    syy = torch.empty((nch, nb), dtype=torch.float64, device=dev)
    sxy = torch.empty((nch, nb), dtype=torch.complex128, device=dev)
    check(lib.pyfft_stft_products(_CODE[precision], _ptr(X), _ptr(Y), nch, K, nb, 1 if conj_x else 0,
                                  _ptr(sxx), _ptr(syy), _ptr(sxy), _stream()))
    return sxx, syy, sxy


# --------------------------------------------------------------------------- #
# K3b: long transforms
# --------------------------------------------------------------------------- #

def _is_pow2(n):
    return n >= 1 and (n & (n - 1)) == 0


def hilbert(u, precision, want_analytic=True, want_envelope=False):
    """u: (batch, n) real device tensor, n a power of two -> (analytic complex | None,
    envelope real | None) device tensors (hilbert.py:22-67)."""
    batch, n = u.shape
    dev = u.device
    u = u.contiguous()
    z = torch.empty((batch, n), dtype=_TORCH_CPLX[precision], device=dev) if want_analytic else None
    env = torch.empty((batch, n), dtype=_TORCH_REAL[precision], device=dev) if want_envelope else None
    # two records ride one complex transform (bigfft.cu): the work buffer is a workspace of
    # ceil(batch / 2) transforms; a single record works in place in the analytic output
    ws, wsb = None, 0
    if batch >= 2 or z is None:
        wsb = check(lib.pyfft_bigfft_workspace_bytes(_CODE[precision], n, (batch + 1) // 2))
        ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    check(lib.pyfft_hilbert(_CODE[precision], _ptr(u), n, batch, _ptr(z), _ptr(env), _ptr(ws), wsb,
                            _stream()))
    return z, env


def xcorr(x, y, precision, normalise=1):
    """x: (n,) or (pairs, n); y: (pairs, n) real device tensors -> (pairs, 2n-1) real device
    tensor in numpy.correlate 'full' order (ccf.py:66-77)."""
    pairs, n = y.shape
    dev = y.device
    y = y.contiguous()
    x = x.contiguous()
    shared = x.dim() == 1 or x.shape[0] == 1
    if not shared and x.shape != y.shape:
        raise ValueError("x must be one record or one record per pair")
    L = 16
    while L < 2 * n - 1:
        L *= 2
    wsb = check(lib.pyfft_bigfft_workspace_bytes(_CODE[precision], L, pairs + (1 if shared else pairs)))
    ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
    out = torch.empty((pairs, 2 * n - 1), dtype=_TORCH_REAL[precision], device=dev)
    check(lib.pyfft_xcorr(_CODE[precision], _ptr(x), 0 if shared else n, _ptr(y), n, n, pairs,
                          int(normalise), _ptr(out), _ptr(ws), wsb, _stream()))
    return out


def spectral_derivative(u, dx, precision, modified=True):
    """u: (batch, n) real device tensor on a uniform grid of step dx, n a power of two ->
    real(ifft(wavenumber * fft(u))) per record (fft_analysis.py:1528-1553)."""
    batch, n = u.shape
    u = u.contiguous()
    out = torch.empty_like(u)
    wsb = check(lib.pyfft_bigfft_workspace_bytes(_CODE[precision], n, (batch + 1) // 2))
    ws = torch.empty(wsb, dtype=torch.uint8, device=u.device)
    check(lib.pyfft_spectral_derivative(_CODE[precision], _ptr(u), n, batch, 1 if modified else 0, float(dx),
                                        _ptr(out), _ptr(ws), wsb, _stream()))
    return out


def fft_c2c(z, precision, inverse=False):
    """(batch, n) complex device tensor, n a power of two -> numpy.fft.fft / ifft along the last axis."""
    batch, n = z.shape
    z = z.contiguous()
    out = torch.empty_like(z)
    wsb = check(lib.pyfft_bigfft_workspace_bytes(_CODE[precision], n, batch))
    ws = torch.empty(wsb, dtype=torch.uint8, device=z.device)
    check(lib.pyfft_fft_c2c(_CODE[precision], _ptr(z), _ptr(out), n, batch, 1 if inverse else 0, _ptr(ws),
                            wsb, _stream()))
    return out


# --------------------------------------------------------------------------- #
# K4: multi-channel cross-spectral matrix
# --------------------------------------------------------------------------- #

def csm_accumulate(x, rmap, nfft, hop, seg_first, seg_count, window, precision, coef=None, G=None):
    """x: (C, n) real device tensor -> G (nfft/2+1, C, C) complex128 device tensor with the
    unscaled sums  G[f, i, j] += sum_k X_j conj(X_i)  (fft_analysis.py:393 per channel pair)."""
    dev = x.device
    nch = x.shape[0]
    nb = nfft // 2 + 1
    if G is None:
        G = torch.zeros((nb, nch, nch), dtype=torch.complex128, device=dev)
    win = torch.from_numpy(np.ascontiguousarray(np.asarray(window, dtype=np.float64))).to(
        dev, _TORCH_REAL[precision])
    seg = Segmentation(int(nfft), int(hop), int(seg_first), int(seg_count))
    need = check(lib.pyfft_csm_workspace_bytes(_CODE[precision], C.byref(seg), nch))
    ws = torch.empty(need, dtype=torch.uint8, device=dev)
    check(lib.pyfft_csm_accumulate(_CODE[precision], _ptr(x), x.stride(0), nch, C.byref(rmap), C.byref(seg),
                                   _ptr(win), _ptr(coef), _ptr(torch.view_as_real(G)), _ptr(ws), need,
                                   _stream()))
    return G
