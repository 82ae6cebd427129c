"""Thread/register-level numpy model of the kernel's Stockham pass structure.

Validates the index math used in pyfft_b200/csrc/fft_core.cuh (slot <-> element
mapping, write positions, twiddle indices, partner bins) against numpy.fft, and
counts shared-memory bank conflicts of the padded layout.  Development aid only.
"""
import numpy as np


def radices(log2l):
    r = []
    m = log2l
    while m >= 4:
        r.append(16); m -= 4
    if m:
        r.append(1 << m)
    return r


def phys(a):
    return a + (a >> 4)


def model_fft(x):
    L = len(x); log2l = L.bit_length() - 1
    TPS = max(L // 16, 1)
    PPT = L // TPS
    tw = np.exp(-2j * np.pi * np.arange(L) / L)
    regs = np.zeros((TPS, PPT), complex)
    for t in range(TPS):
        for s in range(PPT):
            regs[t, s] = x[t + TPS * s]
    NS = 1
    rads = radices(log2l)
    conflicts = 0
    for pi, R in enumerate(rads):
        last = pi == len(rads) - 1
        NB = PPT // R
        smem = np.zeros(phys(L) + 1, complex)
        newregs = np.zeros_like(regs)
        wr_addr = {}
        for t in range(TPS):
            for m in range(NB):
                j = t + TPS * m
                u = np.array([regs[t, m + r * NB] for r in range(R)])
                if NS > 1:
                    k = j % NS
                    for r in range(1, R):
                        idx = r * k * (L // (NS * R))
                        assert idx < L
                        u[r] *= tw[idx]
                u = np.fft.fft(u)
                idxD = (j // NS) * NS * R + (j % NS)
                for r in range(R):
                    o = idxD + r * NS
                    if last:
                        assert o == t + TPS * (m + r * NB), (o, t, m, r)
                        newregs[t, m + r * NB] = u[r]
                    else:
                        smem[phys(o)] = u[r]
                        wr_addr[(t, m, r)] = phys(o)
        if not last:
            # bank conflicts for writes: per (m, r), per half-warp of 16 consecutive t (float2 = 2 words)
            for m in range(NB):
                for r in range(R):
                    for h0 in range(0, TPS, 16):
                        banks = [(2 * wr_addr[(t, m, r)]) % 32 for t in range(h0, min(h0 + 16, TPS))]
                        conflicts += len(banks) - len(set(banks))
            for t in range(TPS):
                for s in range(PPT):
                    newregs[t, s] = smem[phys(t + TPS * s)]
            for s in range(PPT):
                for h0 in range(0, TPS, 16):
                    banks = [(2 * phys(t + TPS * s)) % 32 for t in range(h0, min(h0 + 16, TPS))]
                    conflicts += len(banks) - len(set(banks))
        regs = newregs
        NS *= R
    out = np.zeros(L, complex)
    for t in range(TPS):
        for s in range(PPT):
            out[t + TPS * s] = regs[t, s]
    return out, conflicts


if __name__ == "__main__":
    rng = np.random.default_rng(0)
    for lg in range(4, 14):
        L = 1 << lg
        x = rng.standard_normal(L) + 1j * rng.standard_normal(L)
        y, c = model_fft(x)
        err = np.max(np.abs(y - np.fft.fft(x))) / np.max(np.abs(y))
        print(L, radices(lg), "err %.2e" % err, "bank conflicts", c)
