// Register/shared-memory Stockham FFT building blocks for sm_100a.
//
// A transform of length L = 2^m (16 <= L <= 8192) is carried by TPS = L/16
// threads holding 16 complex points each.  Invariant kept across every pass:
//     thread t, register slot s  <->  element index  t + TPS*s
// both for the inputs of a pass and for the natural-order result of the last
// pass.  Radix-16 passes come first, then one pass of radix 2/4/8 if m % 4 != 0.
// Between passes the data is exchanged through shared memory in the padded
// layout phys(a) = a + (a >> 4), which makes every 64-bit access of a half-warp
// hit 16 distinct bank pairs (checked by tools/model_stockham.py).
//
// Numerical contract (numpy.fft, the reference's only FFT: fft_analysis.py:386,
// 2105, 2116): forward e^{-2 pi i nk/L} unnormalised; inverse conjugate kernel,
// normalisation left to the caller.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pyfft {

template <typename T> struct __align__(2 * sizeof(T)) cpx { T x, y; };

template <typename T> __device__ __forceinline__ cpx<T> mk(T a, T b) { cpx<T> r; r.x = a; r.y = b; return r; }
template <typename T> __device__ __forceinline__ cpx<T> operator+(cpx<T> a, cpx<T> b) { return mk<T>(a.x + b.x, a.y + b.y); }
template <typename T> __device__ __forceinline__ cpx<T> operator-(cpx<T> a, cpx<T> b) { return mk<T>(a.x - b.x, a.y - b.y); }
#ifndef PYFFT_NO_F32X2
// fp32: one packed FADD2 per complex add / subtract (sm_100 f32x2 pipe: the same FP32 throughput from half the
// issue slots -- the register FFT is issue-bound and 42 % of its instructions are FADD, profiles/r01a)
__device__ __forceinline__ cpx<float> operator+(cpx<float> a, cpx<float> b) {
    const float2 r = __fadd2_rn(make_float2(a.x, a.y), make_float2(b.x, b.y));
    return mk<float>(r.x, r.y);
}
__device__ __forceinline__ cpx<float> operator-(cpx<float> a, cpx<float> b) {
    const float2 r = __fadd2_rn(make_float2(a.x, a.y), make_float2(-b.x, -b.y));
    return mk<float>(r.x, r.y);
}
#endif
template <typename T> __device__ __forceinline__ cpx<T> cmul(cpx<T> a, cpx<T> b) {
    return mk<T>(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
template <typename T> __device__ __forceinline__ cpx<T> cconj(cpx<T> a) { return mk<T>(a.x, -a.y); }

// multiply by -i (forward) / +i (inverse)
template <typename T, bool INV> __device__ __forceinline__ cpx<T> rot90(cpx<T> a) {
    return INV ? mk<T>(-a.y, a.x) : mk<T>(a.y, -a.x);
}
// multiply by a compile-time twiddle (c, -s) forward, (c, +s) inverse
template <typename T, bool INV> __device__ __forceinline__ cpx<T> twc(cpx<T> a, T c, T s) {
    return INV ? mk<T>(a.x * c - a.y * s, a.y * c + a.x * s)
               : mk<T>(a.x * c + a.y * s, a.y * c - a.x * s);
}

__host__ __device__ constexpr int phys(int a) { return a + (a >> 4); }
__host__ __device__ constexpr int smem_elems(int L) { return L + (L >> 4) + 1; }

template <typename T, bool INV>
__device__ __forceinline__ void fft2(cpx<T>& a, cpx<T>& b) {
    cpx<T> t = a;
    a = t + b;
    b = t - b;
}

template <typename T, bool INV>
__device__ __forceinline__ void fft4(cpx<T>& a0, cpx<T>& a1, cpx<T>& a2, cpx<T>& a3) {
    cpx<T> t0 = a0 + a2, t1 = a0 - a2, t2 = a1 + a3, t3 = rot90<T, INV>(a1 - a3);
    a0 = t0 + t2; a1 = t1 + t3; a2 = t0 - t2; a3 = t1 - t3;
}

// In-register DFT of R points, natural-order result left in u[0..R-1].
template <typename T, int R, bool INV> struct SmallFFT;

template <typename T, bool INV> struct SmallFFT<T, 2, INV> {
    static __device__ __forceinline__ void run(cpx<T>* u) { fft2<T, INV>(u[0], u[1]); }
};
template <typename T, bool INV> struct SmallFFT<T, 4, INV> {
    static __device__ __forceinline__ void run(cpx<T>* u) { fft4<T, INV>(u[0], u[1], u[2], u[3]); }
};
template <typename T, bool INV> struct SmallFFT<T, 8, INV> {
    // n = i + 4m, k = q + 2p:  X[k] = sum_i W8^{iq} W4^{ip} sum_m v[i+4m] W2^{mq}
    static __device__ __forceinline__ void run(cpx<T>* u) {
        const T h = T(0.70710678118654752440);
#pragma unroll
        for (int i = 0; i < 4; ++i) fft2<T, INV>(u[i], u[i + 4]);
        u[5] = twc<T, INV>(u[5], h, h);
        u[6] = rot90<T, INV>(u[6]);
        u[7] = twc<T, INV>(u[7], -h, h);
        fft4<T, INV>(u[0], u[1], u[2], u[3]);
        fft4<T, INV>(u[4], u[5], u[6], u[7]);
        // result X[q + 2p] sits in u[4q + p]
        cpx<T> o[8];
#pragma unroll
        for (int q = 0; q < 2; ++q)
#pragma unroll
            for (int p = 0; p < 4; ++p) o[q + 2 * p] = u[4 * q + p];
#pragma unroll
        for (int i = 0; i < 8; ++i) u[i] = o[i];
    }
};
template <typename T, bool INV> struct SmallFFT<T, 16, INV> {
    // n = i + 4m, k = q + 4p:  X[k] = sum_i W16^{iq} W4^{ip} sum_m v[i+4m] W4^{mq}
    static __device__ __forceinline__ void run(cpx<T>* u) {
        const T h = T(0.70710678118654752440);
        const T c1 = T(0.92387953251128675613);   // cos(pi/8)
        const T s1 = T(0.38268343236508977173);   // sin(pi/8)
#pragma unroll
        for (int i = 0; i < 4; ++i) fft4<T, INV>(u[i], u[i + 4], u[i + 8], u[i + 12]);
        // u[i + 4q] *= W16^{iq}
        u[5] = twc<T, INV>(u[5], c1, s1);       // W16^1
        u[6] = twc<T, INV>(u[6], h, h);         // W16^2
        u[7] = twc<T, INV>(u[7], s1, c1);       // W16^3
        u[9] = twc<T, INV>(u[9], h, h);         // W16^2
        u[10] = rot90<T, INV>(u[10]);           // W16^4
        u[11] = twc<T, INV>(u[11], -h, h);      // W16^6
        u[13] = twc<T, INV>(u[13], s1, c1);     // W16^3
        u[14] = twc<T, INV>(u[14], -h, h);      // W16^6
        u[15] = twc<T, INV>(u[15], -c1, -s1);   // W16^9
#pragma unroll
        for (int q = 0; q < 4; ++q) fft4<T, INV>(u[4 * q], u[4 * q + 1], u[4 * q + 2], u[4 * q + 3]);
        // result X[q + 4p] sits in u[4q + p]
        cpx<T> o[16];
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int p = 0; p < 4; ++p) o[q + 4 * p] = u[4 * q + p];
#pragma unroll
        for (int i = 0; i < 16; ++i) u[i] = o[i];
    }
};

// ---- shared-memory exchange through explicit 32-bit shared-window addresses ---- //
// (keeps every access an LDS/STS; generic pointers selected at run time made the
// compiler fall back to generic LD/ST)
template <typename T> struct SharedIO;
template <> struct SharedIO<float> {
    static __device__ __forceinline__ cpx<float> ld(uint32_t a) {
        cpx<float> r;
        asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(r.x), "=f"(r.y) : "r"(a) : "memory");
        return r;
    }
    static __device__ __forceinline__ void st(uint32_t a, cpx<float> v) {
        asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(a), "f"(v.x), "f"(v.y) : "memory");
    }
};
template <> struct SharedIO<double> {
    static __device__ __forceinline__ cpx<double> ld(uint32_t a) {
        cpx<double> r;
        asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(r.x), "=d"(r.y) : "r"(a) : "memory");
        return r;
    }
    static __device__ __forceinline__ void st(uint32_t a, cpx<double> v) {
        asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(a), "d"(v.x), "d"(v.y) : "memory");
    }
};

// Two exchange buffers used alternately: every exchange is
//     write cur;  __syncthreads();  read cur;  flip()
// A thread can only start writing a buffer again after it passed the barrier of the
// exchange in between, which every thread reaches after finishing its reads of that
// buffer -- so ONE barrier per exchange is enough, across passes, epilogue exchanges and
// consecutive segments alike.
// SINGLE = true (two buffers do not fit in shared memory): cur == other and every write
// is preceded by an extra barrier after the last reads.
template <typename T, bool SINGLE_ = false> struct Exchange {
    static constexpr bool SINGLE = SINGLE_;
    uint32_t cur, other;         // byte addresses in the shared window
    __device__ __forceinline__ void flip() { uint32_t x = cur; cur = other; other = x; }
    __device__ __forceinline__ void before_write() const { if (SINGLE) __syncthreads(); }
    // phys(base + r*STRIDE) == phys(base) + r*(STRIDE + STRIDE/16) when STRIDE % 16 == 0, and
    // == phys(base) + r when STRIDE == 1, base % 16 == 0, r < 16: one address per thread plus
    // compile-time immediates instead of a shift/add per access.
    template <int STRIDE> static __device__ __forceinline__ constexpr uint32_t lin(int r) {
        return (uint32_t)(r * (STRIDE % 16 == 0 ? STRIDE + STRIDE / 16 : STRIDE)) * (uint32_t)sizeof(cpx<T>);
    }
    static __device__ __forceinline__ uint32_t addr(uint32_t buf, int idx) {
        return buf + (uint32_t)phys(idx) * (uint32_t)sizeof(cpx<T>);
    }
    __device__ __forceinline__ cpx<T> ld(uint32_t buf, int idx) const {
        return SharedIO<T>::ld(buf + (uint32_t)phys(idx) * (uint32_t)sizeof(cpx<T>));
    }
    __device__ __forceinline__ void st(uint32_t buf, int idx, cpx<T> v) const {
        SharedIO<T>::st(buf + (uint32_t)phys(idx) * (uint32_t)sizeof(cpx<T>), v);
    }
};

// Per-thread base twiddles, one per twiddled pass p = 1, 2, 3:
//     w[p-1] = exp(-+2 pi i (t mod NS_p) / (NS_p R_p))
// loaded once per kernel from the long-double table; the powers w^r a radix-R butterfly
// needs are formed in registers by a multiplication tree of depth <= 4 (no per-segment
// table traffic: the table lookups were what saturated L1 in the first version).
template <typename T> struct Twiddles { cpx<T> w[3]; };

template <typename T, int LOG2L, bool INV>
__device__ __forceinline__ Twiddles<T> load_twiddles(const cpx<T>* __restrict__ tw, int t) {
    constexpr int L = 1 << LOG2L;
    Twiddles<T> r;
    int ns = 16;
#pragma unroll
    for (int p = 0; p < 3; ++p) {
        r.w[p] = mk<T>(T(1), T(0));
        if (ns < L) {
            const int rem = L / ns;
            const int R = rem >= 16 ? 16 : rem;
            cpx<T> w = tw[(t & (ns - 1)) * (L / (ns * R))];
            if (INV) w.y = -w.y;
            r.w[p] = w;
            ns *= R;
        }
    }
    return r;
}

// exp(-+2 pi i m / 16), m = 0..7, as compile-time constants
template <typename T, bool INV>
__device__ __forceinline__ cpx<T> w16_power(int m) {
    const T h = T(0.70710678118654752440), c1 = T(0.92387953251128675613), s1 = T(0.38268343236508977173);
    T re = T(1), im = T(0);
    switch (m) {
        case 1: re = c1; im = -s1; break;
        case 2: re = h; im = -h; break;
        case 3: re = s1; im = -c1; break;
        case 4: re = T(0); im = T(-1); break;
        case 5: re = -s1; im = -c1; break;
        case 6: re = -h; im = -h; break;
        case 7: re = -c1; im = -s1; break;
        default: break;
    }
    return mk<T>(re, INV ? -im : im);
}

// u[r] *= w^r, r = 1..R-1
template <typename T, int R>
__device__ __forceinline__ void apply_twiddle_powers(cpx<T>* u, cpx<T> w1) {
    u[1] = cmul(u[1], w1);
    if (R > 2) {
        const cpx<T> w2 = cmul(w1, w1);
        const cpx<T> w3 = cmul(w2, w1);
        u[2] = cmul(u[2], w2);
        u[3] = cmul(u[3], w3);
        if (R > 4) {
            const cpx<T> w4 = cmul(w2, w2);
            const cpx<T> w5 = cmul(w4, w1), w6 = cmul(w4, w2), w7 = cmul(w4, w3);
            u[4] = cmul(u[4], w4); u[5] = cmul(u[5], w5); u[6] = cmul(u[6], w6); u[7] = cmul(u[7], w7);
            if (R > 8) {
                const cpx<T> w8 = cmul(w4, w4);
                u[8] = cmul(u[8], w8);
                u[9] = cmul(u[9], cmul(w8, w1));
                u[10] = cmul(u[10], cmul(w8, w2));
                u[11] = cmul(u[11], cmul(w8, w3));
                u[12] = cmul(u[12], cmul(w8, w4));
                u[13] = cmul(u[13], cmul(w8, w5));
                u[14] = cmul(u[14], cmul(w8, w6));
                u[15] = cmul(u[15], cmul(w8, w7));
            }
        }
    }
}

// One Stockham pass of radix R with NS = product of the previous radices.
//  FIRST: inputs are already in v[] (no exchange read, no twiddles)
//  LAST : results stay in v[] in natural order (no exchange write)
//  w1   : this thread's base twiddle for the pass (butterfly m = 0); a remainder pass
//         (R < 16, always the last one) runs 16/R butterflies whose base twiddles differ
//         by the compile-time rotation exp(-+2 pi i m TPS / L)
template <typename T, int L, int NS, int R, bool INV, bool FIRST, bool LAST, bool SINGLE>
__device__ __forceinline__ void stockham_pass(cpx<T> (&v)[16], const Exchange<T, SINGLE>& ex, uint32_t buf_in,
                                              uint32_t buf_out, cpx<T> w1, int t) {
    constexpr int TPS = (L / 16) > 0 ? (L / 16) : 1;
    constexpr int NB = 16 / R;
    if (!FIRST) {
        if (TPS % 16 == 0) {
            const uint32_t a = ex.addr(buf_in, t);
#pragma unroll
            for (int s = 0; s < 16; ++s) v[s] = SharedIO<T>::ld(a + ex.template lin<TPS>(s));
        } else {
#pragma unroll
            for (int s = 0; s < 16; ++s) v[s] = ex.ld(buf_in, t + TPS * s);
        }
        if (SINGLE && !LAST) __syncthreads();
    }
#pragma unroll
    for (int m = 0; m < NB; ++m) {
        const int j = t + TPS * m;
        cpx<T> u[R];
#pragma unroll
        for (int r = 0; r < R; ++r) u[r] = v[m + r * NB];
        if (NS > 1) {
            cpx<T> wm = w1;
            if (m > 0) wm = cmul(w1, w16_power<T, INV>(m));   // m TPS / (NS R) = m / 16 here
            apply_twiddle_powers<T, R>(u, wm);
        }
        SmallFFT<T, R, INV>::run(u);
        if (LAST) {
#pragma unroll
            for (int r = 0; r < R; ++r) v[m + r * NB] = u[r];
        } else {
            const int base = (j / NS) * (NS * R) + (j & (NS - 1));
            if (NS % 16 == 0 || (NS == 1 && R == 16)) {
                const uint32_t a = ex.addr(buf_out, base);
#pragma unroll
                for (int r = 0; r < R; ++r) SharedIO<T>::st(a + ex.template lin<NS>(r), u[r]);
            } else {
#pragma unroll
                for (int r = 0; r < R; ++r) ex.st(buf_out, base + r * NS, u[r]);
            }
        }
    }
}

template <typename T, int L, int NS, int REMAIN, int P, bool INV, bool SINGLE> struct PassChain {
    // REMAIN = L / NS still to be transformed; P = index of this pass (>= 1)
    static constexpr int R = REMAIN >= 16 ? 16 : REMAIN;
    static constexpr bool LAST = (REMAIN == R);
    static __device__ __forceinline__ void run(cpx<T> (&v)[16], Exchange<T, SINGLE>& ex,
                                               const Twiddles<T>& tw, int t) {
        __syncthreads();                 // the previous pass wrote ex.cur
        const uint32_t in = ex.cur;
        ex.flip();
        stockham_pass<T, L, NS, R, INV, false, LAST, SINGLE>(v, ex, in, ex.cur, tw.w[P - 1], t);
        if (!LAST) PassChain<T, L, NS * R, REMAIN / R, P + 1, INV, SINGLE>::run(v, ex, tw, t);
    }
};
template <typename T, int L, int NS, int P, bool INV, bool SINGLE> struct PassChain<T, L, NS, 1, P, INV, SINGLE> {
    static __device__ __forceinline__ void run(cpx<T> (&)[16], Exchange<T, SINGLE>&, const Twiddles<T>&, int) {}
};

// Full transform of one segment slot.  v[] in: element t + TPS*s; out: bin t + TPS*s.
// All threads of the CTA must call this together (it contains barriers).
template <typename T, int LOG2L, bool INV, bool SINGLE>
__device__ __forceinline__ void fft_block(cpx<T> (&v)[16], Exchange<T, SINGLE>& ex,
                                          const Twiddles<T>& tw, int t) {
    constexpr int L = 1 << LOG2L;
    static_assert(LOG2L >= 4 && LOG2L <= 13, "16 <= L <= 8192");
    if (L == 16) {
        stockham_pass<T, L, 1, 16, INV, true, true, SINGLE>(v, ex, 0u, 0u, mk<T>(T(1), T(0)), t);
    } else {
        ex.before_write();
        stockham_pass<T, L, 1, 16, INV, true, false, SINGLE>(v, ex, 0u, ex.cur, mk<T>(T(1), T(0)), t);
        PassChain<T, L, 16, L / 16, 1, INV, SINGLE>::run(v, ex, tw, t);
    }
}

}  // namespace pyfft
