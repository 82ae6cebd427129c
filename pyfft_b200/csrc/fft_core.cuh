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

// One Stockham pass of radix R with NS = product of the previous radices.
//  FIRST: inputs are already in v[] (no exchange read, no twiddles)
//  LAST : results stay in v[] in natural order (no exchange write)
// `tw` = exp(-2 pi i k / L), k = 0..L-1.
//  SINGLE: buf_in == buf_out (one exchange buffer only): barrier between this
//          pass's reads and its writes
template <typename T, int L, int NS, int R, bool INV, bool FIRST, bool LAST, bool SINGLE = false>
__device__ __forceinline__ void stockham_pass(cpx<T> (&v)[16], const cpx<T>* buf_in, cpx<T>* buf_out,
                                              const cpx<T>* __restrict__ tw, int t) {
    constexpr int TPS = (L / 16) > 0 ? (L / 16) : 1;
    constexpr int NB = 16 / R;
    if (!FIRST) {
#pragma unroll
        for (int s = 0; s < 16; ++s) v[s] = buf_in[phys(t + TPS * s)];
        if (SINGLE && !LAST) __syncthreads();
    }
#pragma unroll
    for (int m = 0; m < NB; ++m) {
        const int j = t + TPS * m;
        cpx<T> u[R];
#pragma unroll
        for (int r = 0; r < R; ++r) u[r] = v[m + r * NB];
        if (NS > 1) {
            const int k = j & (NS - 1);
            constexpr int STEP = L / (NS * R);
#pragma unroll
            for (int r = 1; r < R; ++r) {
                cpx<T> w = tw[r * k * STEP];
                if (INV) w.y = -w.y;
                u[r] = cmul(u[r], w);
            }
        }
        SmallFFT<T, R, INV>::run(u);
        if (LAST) {
#pragma unroll
            for (int r = 0; r < R; ++r) v[m + r * NB] = u[r];
        } else {
            const int base = (j / NS) * (NS * R) + (j & (NS - 1));
#pragma unroll
            for (int r = 0; r < R; ++r) buf_out[phys(base + r * NS)] = u[r];
        }
    }
}

// Two exchange buffers used alternately: every exchange is
//     write bufs[cur];  __syncthreads();  read bufs[cur];  cur ^= 1
// A thread can only start writing bufs[c] again after it passed the barrier of
// the exchange in between, which every thread reaches after finishing its reads
// of bufs[c] -- so ONE barrier per exchange is enough, across passes, epilogue
// exchanges and consecutive segments alike.
//
// SINGLE = true (used when two buffers do not fit in shared memory): b[0] == b[1]
// and every write is preceded by an extra barrier after the last reads.
template <typename T, bool SINGLE_ = false> struct Exchange {
    static constexpr bool SINGLE = SINGLE_;
    cpx<T>* b[2];
    int cur;
    __device__ __forceinline__ cpx<T>* buf() const { return b[cur]; }
    __device__ __forceinline__ void flip() { cur ^= 1; }
    // call before writing buf() when the previous reads of it were NOT already
    // separated from this write by a barrier (first pass of a segment, epilogue)
    __device__ __forceinline__ void before_write() const { if (SINGLE) __syncthreads(); }
};

template <typename T, int L, int NS, int REMAIN, bool INV, bool SINGLE> struct PassChain {
    // REMAIN = L / NS still to be transformed; this is never the first pass.
    static constexpr int R = REMAIN >= 16 ? 16 : REMAIN;
    static constexpr bool LAST = (REMAIN == R);
    static __device__ __forceinline__ void run(cpx<T> (&v)[16], Exchange<T, SINGLE>& ex,
                                               const cpx<T>* __restrict__ tw, int t) {
        // the previous pass wrote ex.buf(); barrier, then read it
        __syncthreads();
        const cpx<T>* in = ex.buf();
        ex.flip();
        stockham_pass<T, L, NS, R, INV, false, LAST, SINGLE>(v, in, ex.buf(), tw, t);
        if (!LAST) PassChain<T, L, NS * R, REMAIN / R, INV, SINGLE>::run(v, ex, tw, t);
    }
};
template <typename T, int L, int NS, bool INV, bool SINGLE> struct PassChain<T, L, NS, 1, INV, SINGLE> {
    static __device__ __forceinline__ void run(cpx<T> (&)[16], Exchange<T, SINGLE>&, const cpx<T>*, int) {}
};

// Full transform of one segment slot.  v[] in: element t + TPS*s; out: bin t + TPS*s.
// All threads of the CTA must call this together (it contains barriers).
template <typename T, int LOG2L, bool INV, bool SINGLE>
__device__ __forceinline__ void fft_block(cpx<T> (&v)[16], Exchange<T, SINGLE>& ex,
                                          const cpx<T>* __restrict__ tw, int t) {
    constexpr int L = 1 << LOG2L;
    static_assert(LOG2L >= 4 && LOG2L <= 13, "16 <= L <= 8192");
    if (L == 16) {
        stockham_pass<T, L, 1, 16, INV, true, true>(v, nullptr, nullptr, tw, t);
    } else {
        ex.before_write();
        stockham_pass<T, L, 1, 16, INV, true, false>(v, nullptr, ex.buf(), tw, t);
        PassChain<T, L, 16, L / 16, INV, SINGLE>::run(v, ex, tw, t);
    }
}

}  // namespace pyfft
