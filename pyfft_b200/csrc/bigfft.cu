// K3b: long power-of-two transforms (2^4 .. 2^27 points) as one to three in-place passes of
// register/shared-memory Stockham FFTs, with the work of the callers fused into the first
// load and the last store of the pass chain:
//   pyfft_hilbert  (hilbert.py:22-112)   real load -> forward -> bin mask -> inverse -> analytic / |z|
//   pyfft_xcorr    (ccf.py:66-77, 239-303) mean removal + zero padding -> forward x, forward y,
//                                         X conj(Y) -> inverse -> lag reorder + 1/(n std std)
//   pyfft_fft_c2c  (fft_analysis.py:2096-2116) numpy.fft.fft / ifft contract, natural order
//
// Two real records per complex transform.  Both callers are LINEAR in the record, so records
// a and b ride one transform as z = a + i b and no spectral un-mixing is needed:
//   Hilbert:  A(z) = A(a) + i A(b) with A(u) = u + i H(u), hence H(a) = Im A(z) - b and
//             H(b) = a - Re A(z); the last pass re-reads a and b and writes both analytic signals;
//   ccf with a shared first record x:  ifft(X conj(Z)) = c(x, a) - i c(x, b).
// Half the transforms, half the HBM traffic of one transform per record.
//
// Decomposition N = N1*N2*N3 (decimation in frequency): pass i transforms digit i
// (length N_i, element stride B_i = prod_{j>i} N_j) for every other index and multiplies the
// result by W_{N_i B_i}^{k_i b}.  The spectrum is left in digit-transposed order
//     position p = k1*N2*N3 + k2*N3 + k3   <->   frequency k = k1 + N1*k2 + N1*N2*k3,
// which element-wise work (mask, X conj(Y)) does not care about; the inverse runs the passes
// backwards (decimation in time, twiddle before the butterflies) and lands in natural order,
// so no transposition pass ever touches HBM.  Only pyfft_fft_c2c, whose contract is a
// natural-order spectrum, transposes -- inside its last pass, through shared memory.
#include "common.cuh"

#ifndef PYFFT_BIG_STRIDED_THREADS
#define PYFFT_BIG_STRIDED_THREADS 256
#endif

namespace pyfft {

enum { IN_CPLX = 0, IN_REAL = 1, IN_REAL2 = 2 };
enum { OUT_CPLX = 0, OUT_MASK = 1, OUT_MULCONJ = 2, OUT_NATURAL = 3, OUT_HILBERT = 4, OUT_XCORR = 5,
       OUT_HILBERT2 = 6, OUT_XCORR2 = 7, OUT_DERIV = 8, OUT_REAL2 = 9 };
enum { TW_NONE = 0, TW_POST = 1, TW_PRE = 2 };

template <typename T> struct BigPass {
    const void* in;
    const void* in2;
    void* out;
    void* out2;
    int64_t in_bs, in2_bs, out_bs, out2_bs;   // per-batch strides in elements of the respective arrays
    int64_t A, B, N;                          // this pass: index = a*R*B + r*B + b; N = whole length
    int64_t N1, N2;                           // leading digit sizes (frequency decoding, natural store)
    int tw_mode;
    const cpx<T>* tw_lo;                      // W_M^j = tw_hi[j >> lo_bits] * tw_lo[j & mask], M = R*B
    const cpx<T>* tw_hi;
    int lo_bits;
    const cpx<T>* tw;                         // R-point table for the butterflies
    int in_mode;
    int64_t n_valid;                          // IN_REAL: samples beyond n_valid read as zero
    const double* mean;                       // IN_REAL: per-batch mean to remove (or null)
    int out_mode;
    double scale;
    const double* bscale;                     // OUT_XCORR: per-batch factor (or null)
    int64_t nsig;                             // OUT_XCORR: record length n (output has 2n-1 lags)
    int nreal;                                // IN_REAL2 / OUT_*2: number of real records (two per transform)
    const double* gscale;                     // IN_REAL2 / OUT_*2: power-of-two scale per real record (or null): two
                                              // records of very different amplitude share one transform, and the
                                              // smaller one would lose log10(ratio) digits to the rounding of the larger
    double d_dx;                              // OUT_DERIV: grid step (the record spans N * d_dx)
    int d_modified;                           // OUT_DERIV: modified wavenumber i sin(k dx)/dx instead of i k
};

template <typename T>
__device__ __forceinline__ cpx<T> big_twiddle(const BigPass<T>& p, int64_t j, bool inv) {
    const cpx<T> hi = p.tw_hi[j >> p.lo_bits];
    const cpx<T> lo = p.tw_lo[j & ((int64_t(1) << p.lo_bits) - 1)];
    cpx<T> w = cmul(hi, lo);
    if (inv) w.y = -w.y;
    return w;
}

// strided passes move runs of TB neighbouring columns: 512 threads double the run (256 bytes for
// the 256-point passes of a 2^25 transform), which HBM rewards; contiguous passes keep 256 threads
template <int LOG2R, bool WIDE = false> struct BigGeom {
    static constexpr int R = 1 << LOG2R;
    static constexpr int TPS = R / 16;
    static constexpr int THREADS = WIDE ? PYFFT_BIG_STRIDED_THREADS : 256;
    static constexpr int TB = THREADS / TPS;       // transforms per CTA
    static constexpr int E = smem_elems(R);
};

// wide CTAs for strided fp32 passes only (fp64 runs are 256 bytes already)
template <typename T, bool STRIDED> struct BigWide { static constexpr bool value = STRIDED && sizeof(T) == 4; };

template <typename T, int LOG2R, bool STRIDED>
constexpr size_t big_smem_bytes() {
    using G = BigGeom<LOG2R, BigWide<T, STRIDED>::value>;
    return (size_t)G::TB * G::E * sizeof(cpx<T>);
}

// v[s] *= base * step^s
template <typename T>
__device__ __forceinline__ void twiddle_column(cpx<T> (&v)[16], cpx<T> base, cpx<T> step) {
    apply_twiddle_powers<T, 16>(v, step);
#pragma unroll
    for (int s = 0; s < 16; ++s) v[s] = cmul(v[s], base);
}

template <typename T, int LOG2R, bool INV, bool STRIDED>
__global__ void __launch_bounds__(BigGeom<LOG2R, BigWide<T, STRIDED>::value>::THREADS)
bigfft_pass_kernel(const BigPass<T> p) {
    using G = BigGeom<LOG2R, BigWide<T, STRIDED>::value>;
    constexpr int R = G::R, TPS = G::TPS, TB = G::TB, E = G::E;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cpx<T>* sm = reinterpret_cast<cpx<T>*>(smem_raw);
    const int tid = threadIdx.x;
    // STRIDED: neighbouring threads own neighbouring columns b (contiguous in memory);
    // contiguous rows (B == 1): neighbouring threads own neighbouring elements of one row
    const int slot = STRIDED ? tid % TB : tid / TPS;
    const int t = STRIDED ? tid / TB : tid % TPS;
    const int64_t bat = blockIdx.y;

    int64_t a, b;
    bool active;
    int64_t k1 = 0, k2 = 0;
    if (p.out_mode == OUT_NATURAL) {
        // rows a = k1*N2 + k2 with TB consecutive k1: the transposed store is contiguous in k1
        k2 = blockIdx.x % p.N2;
        k1 = (blockIdx.x / p.N2) * TB + slot;
        active = k1 < p.N1;
        a = k1 * p.N2 + k2;
        b = 0;
    } else {
        const int64_t q = (int64_t)blockIdx.x * TB + slot;
        active = q < p.A * p.B;
        a = q / p.B;
        b = q - a * p.B;
    }
    const int64_t base_addr = a * R * p.B + b;

    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(sm);
    // one exchange buffer per transform (a second one would only save a barrier in the 512-point pass):
    // the passes are bound by the latency of their strided loads, so resident CTAs matter more
    Exchange<T, true> ex;
    ex.cur = sbase + (uint32_t)(slot * E * sizeof(cpx<T>));
    ex.other = ex.cur;
    const Twiddles<T> tws = load_twiddles<T, LOG2R, INV>(p.tw, t);

    cpx<T> v[16];
    if (active) {
        if (p.in_mode == IN_REAL) {
            const T* src = static_cast<const T*>(p.in) + bat * p.in_bs;
            const T m = p.mean ? (T)p.mean[bat] : T(0);
#pragma unroll
            for (int s = 0; s < 16; ++s) {
                const int64_t i = base_addr + (int64_t)(t + TPS * s) * p.B;
                v[s] = mk<T>(i < p.n_valid ? src[i] - m : T(0), T(0));
            }
        } else if (p.in_mode == IN_REAL2) {
            const T* sa = static_cast<const T*>(p.in) + 2 * bat * p.in_bs;
            const bool hb = 2 * bat + 1 < p.nreal;
            const T* sb = hb ? sa + p.in_bs : sa;
            const T ma = p.mean ? (T)p.mean[2 * bat] : T(0);
            const T mb = (p.mean && hb) ? (T)p.mean[2 * bat + 1] : T(0);
            const T ga = p.gscale ? (T)p.gscale[2 * bat] : T(1);
            const T gb = (p.gscale && hb) ? (T)p.gscale[2 * bat + 1] : T(1);
#pragma unroll
            for (int s = 0; s < 16; ++s) {
                const int64_t i = base_addr + (int64_t)(t + TPS * s) * p.B;
                const bool ok = i < p.n_valid;
                v[s] = mk<T>(ok ? (sa[i] - ma) * ga : T(0), (ok && hb) ? (sb[i] - mb) * gb : T(0));
            }
        } else {
            const cpx<T>* src = static_cast<const cpx<T>*>(p.in) + bat * p.in_bs;
#pragma unroll
            for (int s = 0; s < 16; ++s) v[s] = src[base_addr + (int64_t)(t + TPS * s) * p.B];
        }
    } else {
#pragma unroll
        for (int s = 0; s < 16; ++s) v[s] = mk<T>(T(0), T(0));
    }
    cpx<T> wb = mk<T>(T(1), T(0)), ws = wb;
    if (p.tw_mode != TW_NONE && active) {
        wb = big_twiddle<T>(p, (int64_t)t * b, INV);
        ws = big_twiddle<T>(p, (int64_t)TPS * b, INV);
    }
    if (p.tw_mode == TW_PRE) twiddle_column<T>(v, wb, ws);
    fft_block<T, LOG2R, INV, true>(v, ex, tws, t);
    if (p.tw_mode == TW_POST) twiddle_column<T>(v, wb, ws);

    const T sc = (T)p.scale;
    if (p.out_mode == OUT_NATURAL) {
        // transpose through shared memory: [slot][idx] -> thread (slot' fastest) stores k1-contiguous
        const uint32_t e = ex.cur;
        ex.before_write();
#pragma unroll
        for (int s = 0; s < 16; ++s) ex.st(e, t + TPS * s, v[s]);
        __syncthreads();
        const int slot2 = tid % TB, t2 = tid / TB;
        const int64_t k1b = (blockIdx.x / p.N2) * TB + slot2;
        if (k1b < p.N1) {
            cpx<T>* dst = static_cast<cpx<T>*>(p.out) + bat * p.out_bs;
            const uint32_t e2 = (uint32_t)((int)e + (slot2 - slot) * (int)(E * sizeof(cpx<T>)));
#pragma unroll
            for (int s = 0; s < 16; ++s) {
                const int idx = t2 + TPS * s;
                const cpx<T> z = ex.ld(e2, idx);
                dst[k1b + p.N1 * k2 + p.N1 * p.N2 * (int64_t)idx] = mk<T>(z.x * sc, z.y * sc);
            }
        }
        return;
    }
    if (!active) return;
    if (p.out_mode == OUT_CPLX) {
        cpx<T>* dst = static_cast<cpx<T>*>(p.out) + bat * p.out_bs;
#pragma unroll
        for (int s = 0; s < 16; ++s)
            dst[base_addr + (int64_t)(t + TPS * s) * p.B] = mk<T>(v[s].x * sc, v[s].y * sc);
    } else if (p.out_mode == OUT_MASK) {
        // hilbert.py:47-64 for even N: bin 0 x1, 1..N/2-1 x2, N/2 x1, above x0
        cpx<T>* dst = static_cast<cpx<T>*>(p.out) + bat * p.out_bs;
        const int64_t klow = (a / p.N2) + p.N1 * (a % p.N2);
        const int64_t nyq = p.N / 2;
#pragma unroll
        for (int s = 0; s < 16; ++s) {
            const int64_t k = klow + p.N1 * p.N2 * (int64_t)(t + TPS * s);
            const T wgt = (k == 0 || k == nyq) ? T(1) : (k < nyq ? T(2) : T(0));
            dst[base_addr + (int64_t)(t + TPS * s) * p.B] = mk<T>(v[s].x * wgt, v[s].y * wgt);
        }
    } else if (p.out_mode == OUT_DERIV) {
        // fft_analysis.py:1528-1540: k = 2 pi fftfreq(N, d = dx/L) = 2 pi m (m the signed bin index, L = N dx),
        // wavenumber = i sin(k dx)/dx / L (modified) or i k / L; (re + i im) * (i w) = (-im w, re w)
        cpx<T>* dst = static_cast<cpx<T>*>(p.out) + bat * p.out_bs;
        const int64_t klow = (a / p.N2) + p.N1 * (a % p.N2);
        const double len = (double)p.N * p.d_dx;
#pragma unroll
        for (int s = 0; s < 16; ++s) {
            const int64_t k = klow + p.N1 * p.N2 * (int64_t)(t + TPS * s);
            const double m = (double)(k < (p.N + 1) / 2 ? k : k - p.N);
            const double wd = p.d_modified ? sinpi(2.0 * m * p.d_dx) / p.d_dx : 6.283185307179586476925286766559 * m;
            // the Nyquist bin of a real record is real, so i w X[N/2] only feeds the imaginary part that
            // the reference drops (:1553 real()): zero it, which also keeps two packed records apart
            const T w = 2 * k == p.N ? T(0) : (T)(wd / len);
            dst[base_addr + (int64_t)(t + TPS * s) * p.B] = mk<T>(-v[s].y * w, v[s].x * w);
        }
    } else if (p.out_mode == OUT_REAL2) {
        // real results of two records that shared the transform: Re -> record 2 bat, Im -> 2 bat + 1
        const bool hb = 2 * bat + 1 < p.nreal;
        T* da = static_cast<T*>(p.out) + 2 * bat * p.out_bs;
        const T sa_ = p.gscale ? (T)(p.scale / p.gscale[2 * bat]) : sc;
        const T sb_ = (p.gscale && hb) ? (T)(p.scale / p.gscale[2 * bat + 1]) : sc;
#pragma unroll
        for (int s = 0; s < 16; ++s) {
            const int64_t i = base_addr + (int64_t)(t + TPS * s) * p.B;
            da[i] = v[s].x * sa_;
            if (hb) da[p.out_bs + i] = v[s].y * sb_;
        }
    } else if (p.out_mode == OUT_MULCONJ) {
        cpx<T>* dst = static_cast<cpx<T>*>(p.out) + bat * p.out_bs;
        const cpx<T>* oth = static_cast<const cpx<T>*>(p.in2) + bat * p.in2_bs;
#pragma unroll
        for (int s = 0; s < 16; ++s) {
            const int64_t i = base_addr + (int64_t)(t + TPS * s) * p.B;
            dst[i] = cmul(oth[i], cconj(v[s]));
        }
    } else if (p.out_mode == OUT_HILBERT) {
        cpx<T>* za = p.out ? static_cast<cpx<T>*>(p.out) + bat * p.out_bs : nullptr;
        T* env = p.out2 ? static_cast<T*>(p.out2) + bat * p.out2_bs : nullptr;
#pragma unroll
        for (int s = 0; s < 16; ++s) {
            const int64_t i = base_addr + (int64_t)(t + TPS * s) * p.B;
            const T re = v[s].x * sc, im = v[s].y * sc;
            if (za) za[i] = mk<T>(re, im);
            if (env) env[i] = sqrt(re * re + im * im);
        }
    } else if (p.out_mode == OUT_HILBERT2) {
        // v = A(ga a + i gb b): ga H(a) = Im v - gb b, gb H(b) = ga a - Re v; records 2 bat and 2 bat + 1
        const bool hb = 2 * bat + 1 < p.nreal;
        const T* ua = static_cast<const T*>(p.in2) + 2 * bat * p.in2_bs;
        const T* ub = hb ? ua + p.in2_bs : ua;
        cpx<T>* za = p.out ? static_cast<cpx<T>*>(p.out) + 2 * bat * p.out_bs : nullptr;
        T* env = p.out2 ? static_cast<T*>(p.out2) + 2 * bat * p.out2_bs : nullptr;
        const T ga2 = p.gscale ? (T)p.gscale[2 * bat] : T(1), gb2 = (p.gscale && hb) ? (T)p.gscale[2 * bat + 1] : T(1);
        const T iga = T(1) / ga2, igb = T(1) / gb2;               // powers of two: exact
#pragma unroll
        for (int s = 0; s < 16; ++s) {
            const int64_t i = base_addr + (int64_t)(t + TPS * s) * p.B;
            const T a = ua[i], b = hb ? ub[i] : T(0);
            const T ha = (v[s].y * sc - gb2 * b) * iga, hbv = (ga2 * a - v[s].x * sc) * igb;
            if (za) za[i] = mk<T>(a, ha);
            if (env) env[i] = sqrt(a * a + ha * ha);
            if (hb) {
                if (za) za[p.out_bs + i] = mk<T>(b, hbv);
                if (env) env[p.out2_bs + i] = sqrt(b * b + hbv * hbv);
            }
        }
    } else if (p.out_mode == OUT_XCORR2) {
        // v = c(x, a) - i c(x, b): records 2 bat and 2 bat + 1 of the second operand
        const bool hb = 2 * bat + 1 < p.nreal;
        T* da = static_cast<T*>(p.out) + 2 * bat * p.out_bs;
        const T fa = (T)(p.scale * (p.bscale ? p.bscale[2 * bat] : 1.0) / (p.gscale ? p.gscale[2 * bat] : 1.0));
        const T fb = hb ? -(T)(p.scale * (p.bscale ? p.bscale[2 * bat + 1] : 1.0) / (p.gscale ? p.gscale[2 * bat + 1] : 1.0)) : T(0);
        const int64_t n = p.nsig, tail = p.N - (n - 1);
#pragma unroll
        for (int s = 0; s < 16; ++s) {
            const int64_t i = base_addr + (int64_t)(t + TPS * s) * p.B;
            int64_t o = -1;
            if (i < n) o = i + n - 1;
            else if (i >= tail) o = i - tail;
            if (o >= 0) {
                da[o] = v[s].x * fa;
                if (hb) da[p.out_bs + o] = v[s].y * fb;
            }
        }
    } else if (p.out_mode == OUT_XCORR) {
        // numpy.correlate(..., 'full') order: lags -(n-1) .. n-1  (ccf.py:69-74)
        T* dst = static_cast<T*>(p.out) + bat * p.out_bs;
        const T f = (T)(p.scale * (p.bscale ? p.bscale[bat] : 1.0));
        const int64_t n = p.nsig, tail = p.N - (n - 1);
#pragma unroll
        for (int s = 0; s < 16; ++s) {
            const int64_t i = base_addr + (int64_t)(t + TPS * s) * p.B;
            if (i < n) dst[i + n - 1] = v[s].x * f;
            else if (i >= tail) dst[i - tail] = v[s].x * f;
        }
    }
}

// sums[2b] += sum (x - shift_b), sums[2b+1] += sum (x - shift_b)^2   (double accumulation)
template <typename T>
__global__ void __launch_bounds__(256) moments_kernel(const T* __restrict__ x, int64_t stride, int64_t n,
                                                      const double* __restrict__ shift, double* __restrict__ sums) {
    const int bat = blockIdx.y;
    const T* p = x + (int64_t)bat * stride;
    const double sh = shift ? shift[bat] : 0.0;
    double s0 = 0.0, s1 = 0.0;
    const int64_t step = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += step) {
        const double v = (double)p[i] - sh;
        s0 += v;
        s1 += v * v;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s0 += __shfl_xor_sync(0xffffffffu, s0, o);
        s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    }
    __shared__ double sh0[8], sh1[8];
    if ((threadIdx.x & 31) == 0) { sh0[threadIdx.x >> 5] = s0; sh1[threadIdx.x >> 5] = s1; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int i = 0; i < 8; ++i) { a += sh0[i]; b += sh1[i]; }
        atomicAdd(&sums[2 * bat], a);
        atomicAdd(&sums[2 * bat + 1], b);
    }
}

// mean[b] = sums[2b] / n
// power-of-two scale per real record from a sample of <= 65 536 values: rms * g in [1/2, 1)
// (any g is exact mathematically -- it only balances the two records that share a transform).
// 256 runs of 256 consecutive values spread evenly over the record, one CTA of 1024 threads per record, 16
// independent coalesced loads in flight per thread (a 2 KB-strided walk by 256 threads was a serial chain of
// TLB misses: 1.1 ms for 16 records of 2^25 points)
template <typename T>
__global__ void __launch_bounds__(1024) record_scale_kernel(const T* __restrict__ x, int64_t stride, int64_t n,
                                                            const double* __restrict__ mean, double* __restrict__ g) {
    const int rec = blockIdx.x;
    const T* p = x + (int64_t)rec * stride;
    const double m = mean ? mean[rec] : 0.0;
    double s = 0.0, c = 0.0;
    if (n <= 65536) {
        for (int64_t i = threadIdx.x; i < n; i += 1024) {
            const double v = (double)p[i] - m;
            s += v * v;
            c += 1.0;
        }
    } else {
        const int64_t run_stride = (n - 256) / 255;
#pragma unroll 16
        for (int k = 0; k < 64; ++k) {
            const int j = k * 1024 + (int)threadIdx.x;
            const double v = (double)p[(int64_t)(j >> 8) * run_stride + (j & 255)] - m;
            s += v * v;
        }
        c = 64.0;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); c += __shfl_xor_sync(0xffffffffu, c, o); }
    __shared__ double sh[2][32];
    if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = s; sh[1][threadIdx.x >> 5] = c; }
    __syncthreads();
    if (threadIdx.x == 0) {
        s = c = 0.0;
        for (int w = 0; w < 32; ++w) { s += sh[0][w]; c += sh[1][w]; }
        const double rms = c > 0.0 ? sqrt(s / c) : 0.0;
        int e = 0;
        if (rms > 0.0 && isfinite(rms)) frexp(rms, &e);
        g[rec] = ldexp(1.0, -e);
    }
}

__global__ void mean_from_sums_kernel(const double* sums, int nb, int64_t n, double* mean) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nb) mean[i] = sums[2 * i] / (double)n;
}

// norm[p] = 1 / (n * std_x * std_y) from the centred second moments (numpy.std: population, ccf.py:76)
__global__ void xcorr_norm_kernel(const double* sx, int x_shared, const double* sy, int pairs, int64_t n,
                                  double* norm) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= pairs) return;
    const double vx = sx[2 * (x_shared ? 0 : i) + 1] / (double)n;
    const double vy = sy[2 * i + 1] / (double)n;
    norm[i] = 1.0 / ((double)n * sqrt(vx) * sqrt(vy));
}

// ------------------------------------------------------------------------- //
// host side: pass planning and launch
// ------------------------------------------------------------------------- //
struct BigPlan {
    int npass;
    int bits[3];
};

static BigPlan plan_bits(int m) {
    BigPlan pl;
    if (m <= 9) { pl.npass = 1; pl.bits[0] = m; pl.bits[1] = pl.bits[2] = 0; }
    // the longest pass goes LAST (innermost digit, contiguous rows): a strided pass of length R
    // moves runs of 256 / (R / 16) columns, i.e. 128-byte runs for R = 256 but only 64 for R = 512
    else if (m <= 18) { pl.npass = 2; pl.bits[0] = m / 2; pl.bits[1] = m - m / 2; pl.bits[2] = 0; }
    else {
        pl.npass = 3;
        pl.bits[2] = (m + 2) / 3;
        pl.bits[1] = (m - pl.bits[2] + 1) / 2;
        pl.bits[0] = m - pl.bits[2] - pl.bits[1];
    }
    return pl;
}

template <typename T, int LOG2R, bool INV, bool STRIDED>
static int launch_pass_t(const BigPass<T>& p, int64_t tiles, int batch, cudaStream_t st) {
    constexpr size_t smem = big_smem_bytes<T, LOG2R, STRIDED>();
    auto kern = bigfft_pass_kernel<T, LOG2R, INV, STRIDED>;
    PYFFT_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<dim3((unsigned)tiles, (unsigned)batch), BigGeom<LOG2R, BigWide<T, STRIDED>::value>::THREADS, smem, st>>>(p);
    PYFFT_LAUNCH_OK("bigfft_pass_kernel");
    return PYFFT_OK;
}

template <typename T>
static int launch_pass(int log2r, bool inv, BigPass<T> p, int batch, cudaStream_t st) {
    const int R = 1 << log2r;
    int status;
    const int dt = sizeof(T) == 8 ? PYFFT_F64 : PYFFT_F32;
    p.tw = reinterpret_cast<const cpx<T>*>(twiddle_table(dt, R, &status));
    if (!p.tw) return status;
    if (p.tw_mode != TW_NONE) {
        const int64_t M = (int64_t)R * p.B;
        const int mb = ilog2(M);
        p.lo_bits = (mb + 1) / 2;
        p.tw_lo = reinterpret_cast<const cpx<T>*>(twiddle_table_ex(dt, M, 1, int64_t(1) << p.lo_bits, &status));
        if (!p.tw_lo) return status;
        p.tw_hi = reinterpret_cast<const cpx<T>*>(
            twiddle_table_ex(dt, M, int64_t(1) << p.lo_bits, M >> p.lo_bits, &status));
        if (!p.tw_hi) return status;
    }
    const bool strided = p.B > 1;
    const int TB = ((strided && sizeof(T) == 4) ? PYFFT_BIG_STRIDED_THREADS : 256) / (R / 16);
    int64_t tiles;
    if (p.out_mode == OUT_NATURAL) tiles = p.N2 * ((p.N1 + TB - 1) / TB);
    else tiles = (p.A * p.B + TB - 1) / TB;
#define PYFFT_CASE(N)                                                                             \
    case N:                                                                                       \
        if (inv) return strided ? launch_pass_t<T, N, true, true>(p, tiles, batch, st)            \
                                : launch_pass_t<T, N, true, false>(p, tiles, batch, st);          \
        return strided ? launch_pass_t<T, N, false, true>(p, tiles, batch, st)                    \
                       : launch_pass_t<T, N, false, false>(p, tiles, batch, st);
    switch (log2r) {
        PYFFT_CASE(4) PYFFT_CASE(5) PYFFT_CASE(6) PYFFT_CASE(7) PYFFT_CASE(8) PYFFT_CASE(9)
    }
#undef PYFFT_CASE
    return fail(PYFFT_EUNSUPPORTED, "pass length 2^%d not supported", log2r);
}

template <typename T> static BigPass<T> blank_pass(int64_t N) {
    BigPass<T> p;
    p.in = p.in2 = nullptr; p.out = p.out2 = nullptr;
    p.in_bs = p.in2_bs = p.out_bs = p.out2_bs = 0;
    p.A = 1; p.B = 1; p.N = N; p.N1 = 1; p.N2 = 1;
    p.tw_mode = TW_NONE; p.tw_lo = p.tw_hi = nullptr; p.lo_bits = 0; p.tw = nullptr;
    p.in_mode = IN_CPLX; p.n_valid = N; p.mean = nullptr;
    p.out_mode = OUT_CPLX; p.scale = 1.0; p.bscale = nullptr; p.nsig = 0; p.nreal = 0;
    p.d_dx = 1.0; p.d_modified = 0; p.gscale = nullptr;
    return p;
}

// Forward (or, with inv, conjugate-kernel) decimation-in-frequency chain: `first` describes the
// load of pass 0 (in, in_bs, in_mode, n_valid, mean), `last` the store of the final pass
// (out_mode, out, ...).  Intermediate data lives in buf (complex [batch][N], stride N).
template <typename T>
static int run_dif(int m, bool inv, const BigPass<T>& first, const BigPass<T>& last, void* buf, int batch,
                   cudaStream_t st) {
    const BigPlan pl = plan_bits(m);
    const int64_t N = int64_t(1) << m;
    int64_t A = 1;
    for (int i = 0; i < pl.npass; ++i) {
        const int64_t R = int64_t(1) << pl.bits[i];
        BigPass<T> p = blank_pass<T>(N);
        p.A = A; p.B = N / (A * R);
        p.N1 = int64_t(1) << pl.bits[0];
        p.N2 = pl.npass == 3 ? int64_t(1) << pl.bits[1] : 1;
        if (pl.npass == 1) p.N1 = 1;
        p.tw_mode = p.B > 1 ? TW_POST : TW_NONE;
        p.in = buf; p.in_bs = N; p.out = buf; p.out_bs = N;
        if (i == 0) {
            p.in = first.in; p.in_bs = first.in_bs; p.in_mode = first.in_mode;
            p.n_valid = first.n_valid; p.mean = first.mean; p.nreal = first.nreal; p.gscale = first.gscale;
        }
        if (i == pl.npass - 1) {
            p.out_mode = last.out_mode; p.scale = last.scale;
            p.d_dx = last.d_dx; p.d_modified = last.d_modified;
            p.in2 = last.in2; p.in2_bs = last.in2_bs;
            if (last.out) { p.out = last.out; p.out_bs = last.out_bs; }
        }
        int rc = launch_pass<T>(pl.bits[i], inv, p, batch, st);
        if (rc) return rc;
        A *= R;
    }
    return PYFFT_OK;
}

// Inverse decimation-in-time chain over a digit-transposed spectrum held in buf; `last`
// describes the store of the final pass (natural order).
template <typename T>
static int run_dit_inverse(int m, const BigPass<T>& last, void* buf, int batch, cudaStream_t st) {
    const BigPlan pl = plan_bits(m);
    const int64_t N = int64_t(1) << m;
    int64_t A = N;
    for (int i = pl.npass - 1; i >= 0; --i) {
        const int64_t R = int64_t(1) << pl.bits[i];
        // digit i: A = prod_{j<i} N_j, B = prod_{j>i} N_j
        int64_t Ai = 1;
        for (int j = 0; j < i; ++j) Ai <<= pl.bits[j];
        BigPass<T> p = blank_pass<T>(N);
        p.A = Ai; p.B = N / (Ai * R);
        p.tw_mode = p.B > 1 ? TW_PRE : TW_NONE;
        p.in = buf; p.in_bs = N; p.out = buf; p.out_bs = N;
        if (i == 0) {
            p.out_mode = last.out_mode; p.out = last.out; p.out_bs = last.out_bs;
            p.out2 = last.out2; p.out2_bs = last.out2_bs; p.scale = last.scale;
            p.bscale = last.bscale; p.nsig = last.nsig;
            p.in2 = last.in2; p.in2_bs = last.in2_bs; p.nreal = last.nreal; p.gscale = last.gscale;
        }
        int rc = launch_pass<T>(pl.bits[i], true, p, batch, st);
        if (rc) return rc;
        (void)A;
    }
    return PYFFT_OK;
}

static int check_pow2(int64_t n, int lo, int hi, int* m) {
    if (n <= 0) return fail(PYFFT_EINVAL, "transform length must be positive");
    const int l = ilog2(n);
    if ((int64_t(1) << l) != n) return fail(PYFFT_EUNSUPPORTED, "length %lld is not a power of two", (long long)n);
    if (l < lo || l > hi)
        return fail(PYFFT_EUNSUPPORTED, "length 2^%d outside the supported range 2^%d .. 2^%d", l, lo, hi);
    *m = l;
    return PYFFT_OK;
}

// scales of `nrec` real records into the workspace slack at byte offset `used` (null when there is no room)
template <typename T>
static double* pair_scales(const void* x, int64_t stride, int64_t n, int nrec, const double* mean, void* ws, int64_t used,
                           int64_t wsb, cudaStream_t st) {
    const int64_t off = (used + 15) & ~(int64_t)15;
    if (!ws || wsb < off + 8 * (int64_t)nrec) return nullptr;
    double* g = reinterpret_cast<double*>(static_cast<char*>(ws) + off);
    record_scale_kernel<T><<<nrec, 1024, 0, st>>>(static_cast<const T*>(x), stride, n, mean, g);
    return g;
}

template <typename T>
static int hilbert_impl(const void* in, int64_t n, int batch, void* analytic, void* envelope, void* ws,
                        int64_t wsb, cudaStream_t st) {
    int m = 0;
    int rc = check_pow2(n, 4, 27, &m);
    if (rc) return rc;
    const int npair = (batch + 1) / 2;
    if (batch >= 2 && ws && wsb >= (int64_t)npair * n * (int64_t)sizeof(cpx<T>)) {
        // two records per transform (see the top of the file); the work buffer is the workspace
        BigPass<T> first = blank_pass<T>(n), last = blank_pass<T>(n);
        first.in = in; first.in_bs = n; first.in_mode = IN_REAL2; first.n_valid = n; first.nreal = batch;
        // per-record scales live in the slack pyfft_bigfft_workspace_bytes adds behind the work buffer
        double* gs = pair_scales<T>(in, n, n, batch, nullptr, ws, (int64_t)npair * n * (int64_t)sizeof(cpx<T>), wsb, st);
        first.gscale = gs;
        last.out_mode = OUT_MASK;
        if ((rc = run_dif<T>(m, false, first, last, ws, npair, st))) return rc;
        BigPass<T> fin = blank_pass<T>(n);
        fin.out_mode = OUT_HILBERT2; fin.out = analytic; fin.out_bs = n; fin.out2 = envelope; fin.out2_bs = n;
        fin.in2 = in; fin.in2_bs = n; fin.nreal = batch; fin.gscale = gs;
        fin.scale = 1.0 / (double)n;
        return run_dit_inverse<T>(m, fin, ws, npair, st);
    }
    void* buf = analytic;                        // the analytic output doubles as the work buffer
    if (!buf) {
        if (!ws || wsb < (int64_t)batch * n * (int64_t)sizeof(cpx<T>))
            return fail(PYFFT_EINVAL, "hilbert: workspace of %lld bytes needed when analytic is NULL",
                        (long long)((int64_t)batch * n * sizeof(cpx<T>)));
        buf = ws;
    }
    BigPass<T> first = blank_pass<T>(n), last = blank_pass<T>(n);
    first.in = in; first.in_bs = n; first.in_mode = IN_REAL; first.n_valid = n;
    last.out_mode = OUT_MASK;
    if ((rc = run_dif<T>(m, false, first, last, buf, batch, st))) return rc;
    BigPass<T> fin = blank_pass<T>(n);
    fin.out_mode = OUT_HILBERT; fin.out = analytic; fin.out_bs = n; fin.out2 = envelope; fin.out2_bs = n;
    fin.scale = 1.0 / (double)n;
    return run_dit_inverse<T>(m, fin, buf, batch, st);
}

// fft_analysis.py:1553: real(ifft(wavenumber * fft(sig))) for `batch` real records on one grid
template <typename T>
static int deriv_impl(const void* in, int64_t n, int batch, int modified, double dx, void* out, void* ws,
                      int64_t wsb, cudaStream_t st) {
    int m = 0;
    int rc = check_pow2(n, 4, 27, &m);
    if (rc) return rc;
    const int npair = (batch + 1) / 2;
    if (!ws || wsb < (int64_t)npair * n * (int64_t)sizeof(cpx<T>))
        return fail(PYFFT_EINVAL, "spectral_derivative: workspace of %lld bytes needed",
                    (long long)((int64_t)npair * n * sizeof(cpx<T>)));
    BigPass<T> first = blank_pass<T>(n), last = blank_pass<T>(n);
    first.in = in; first.in_bs = n; first.in_mode = IN_REAL2; first.n_valid = n; first.nreal = batch;
    double* gs = pair_scales<T>(in, n, n, batch, nullptr, ws, (int64_t)npair * n * (int64_t)sizeof(cpx<T>), wsb, st);
    first.gscale = gs;
    last.out_mode = OUT_DERIV; last.d_dx = dx; last.d_modified = modified;
    if ((rc = run_dif<T>(m, false, first, last, ws, npair, st))) return rc;
    BigPass<T> fin = blank_pass<T>(n);
    fin.out_mode = OUT_REAL2; fin.out = out; fin.out_bs = n; fin.nreal = batch; fin.gscale = gs;
    fin.scale = 1.0 / (double)n;
    return run_dit_inverse<T>(m, fin, ws, npair, st);
}

template <typename T>
static int xcorr_impl(const void* x, int64_t xs, const void* y, int64_t ys, int64_t n, int pairs, int normalise,
                      void* out, void* ws, int64_t wsb, cudaStream_t st) {
    if (n < 2) return fail(PYFFT_EINVAL, "xcorr needs n >= 2");
    int m = ilog2(2 * n - 1);
    if (m < 4) m = 4;
    if (m > 27) return fail(PYFFT_EUNSUPPORTED, "xcorr: padded length 2^%d exceeds 2^27", m);
    const int64_t N = int64_t(1) << m;
    const int nx = xs == 0 ? 1 : pairs;
    const int64_t cb = (int64_t)sizeof(cpx<T>);
    const int64_t need = ((int64_t)nx + pairs) * N * cb + 4096 + 64 * (int64_t)(nx + pairs);
    if (!ws || wsb < need)
        return fail(PYFFT_EINVAL, "xcorr: workspace too small: %lld < %lld bytes", (long long)wsb, (long long)need);
    char* w = static_cast<char*>(ws);
    void* bufx = w;
    void* bufy = w + (int64_t)nx * N * cb;
    double* sums_x = reinterpret_cast<double*>(w + ((int64_t)nx + pairs) * N * cb);   // [nx][2]
    double* sums_y = sums_x + 2 * nx;                                               // [pairs][2]
    double* mean_x = sums_y + 2 * pairs;
    double* mean_y = mean_x + nx;
    double* norm = mean_y + pairs;
    const size_t small_bytes = (size_t)(3 * nx + 4 * pairs) * sizeof(double);
    PYFFT_CUDA_OK(cudaMemsetAsync(sums_x, 0, small_bytes, st));
    const T* xp = static_cast<const T*>(x);
    const T* yp = static_cast<const T*>(y);
    int gx = (int)((n + 256 * 16 - 1) / (256 * 16));
    if (gx > 1184) gx = 1184;
    // means (pass 1), then centred second moments (pass 2): numpy's mean()/std() (ccf.py:74-76)
    if (normalise) {
        moments_kernel<T><<<dim3(gx, nx), 256, 0, st>>>(xp, xs, n, nullptr, sums_x);
        moments_kernel<T><<<dim3(gx, pairs), 256, 0, st>>>(yp, ys, n, nullptr, sums_y);
        mean_from_sums_kernel<<<(nx + pairs + 127) / 128, 128, 0, st>>>(sums_x, nx + pairs, n, mean_x);
        PYFFT_CUDA_OK(cudaMemsetAsync(sums_x, 0, (size_t)(2 * (nx + pairs)) * sizeof(double), st));
        moments_kernel<T><<<dim3(gx, nx), 256, 0, st>>>(xp, xs, n, mean_x, sums_x);
        moments_kernel<T><<<dim3(gx, pairs), 256, 0, st>>>(yp, ys, n, mean_y, sums_y);
        xcorr_norm_kernel<<<(pairs + 127) / 128, 128, 0, st>>>(sums_x, xs == 0, sums_y, pairs, n, norm);
        PYFFT_LAUNCH_OK("xcorr statistics");
    }

    BigPass<T> first = blank_pass<T>(N), last = blank_pass<T>(N);
    first.in = x; first.in_bs = xs; first.in_mode = IN_REAL; first.n_valid = n; first.mean = normalise ? mean_x : nullptr;
    int rc;
    if ((rc = run_dif<T>(m, false, first, last, bufx, nx, st))) return rc;
    if (xs == 0 && pairs >= 2) {
        // shared first record: two second records per transform (see the top of the file)
        const int npair = (pairs + 1) / 2;
        first.in = y; first.in_bs = ys; first.in_mode = IN_REAL2; first.nreal = pairs;
        first.mean = normalise ? mean_y : nullptr;
        double* gs = norm + pairs;                                           // inside the 64 (nx + pairs) bytes of slack
        record_scale_kernel<T><<<pairs, 1024, 0, st>>>(yp, ys, n, normalise ? mean_y : nullptr, gs);
        first.gscale = gs;
        last.out_mode = OUT_MULCONJ; last.in2 = bufx; last.in2_bs = 0;
        if ((rc = run_dif<T>(m, false, first, last, bufy, npair, st))) return rc;
        BigPass<T> fin2 = blank_pass<T>(N);
        fin2.out_mode = OUT_XCORR2; fin2.out = out; fin2.out_bs = 2 * n - 1; fin2.scale = 1.0 / (double)N;
        fin2.bscale = normalise == 1 ? norm : nullptr; fin2.nsig = n; fin2.nreal = pairs; fin2.gscale = gs;
        return run_dit_inverse<T>(m, fin2, bufy, npair, st);
    }
    first.in = y; first.in_bs = ys; first.mean = normalise ? mean_y : nullptr;
    last.out_mode = OUT_MULCONJ; last.in2 = bufx; last.in2_bs = xs == 0 ? 0 : N;
    if ((rc = run_dif<T>(m, false, first, last, bufy, pairs, st))) return rc;
    BigPass<T> fin = blank_pass<T>(N);
    fin.out_mode = OUT_XCORR; fin.out = out; fin.out_bs = 2 * n - 1; fin.scale = 1.0 / (double)N;
    fin.bscale = normalise == 1 ? norm : nullptr; fin.nsig = n;
    return run_dit_inverse<T>(m, fin, bufy, pairs, st);
}

template <typename T>
static int c2c_impl(const void* in, void* out, int64_t n, int batch, int inverse, void* ws, int64_t wsb,
                    cudaStream_t st) {
    int m = 0;
    int rc = check_pow2(n, 4, 27, &m);
    if (rc) return rc;
    const BigPlan pl = plan_bits(m);
    BigPass<T> first = blank_pass<T>(n), last = blank_pass<T>(n);
    first.in = in; first.in_bs = n;
    last.scale = inverse ? 1.0 / (double)n : 1.0;
    if (pl.npass == 1) {
        // one contiguous pass over batch rows: natural order falls out directly
        BigPass<T> p = blank_pass<T>(n);
        p.in = in; p.out = out; p.A = batch; p.scale = last.scale;
        return launch_pass<T>(m, inverse != 0, p, 1, st);
    }
    if (!ws || wsb < (int64_t)batch * n * (int64_t)sizeof(cpx<T>))
        return fail(PYFFT_EINVAL, "fft_c2c: workspace of %lld bytes needed",
                    (long long)((int64_t)batch * n * sizeof(cpx<T>)));
    last.out_mode = OUT_NATURAL; last.out = out; last.out_bs = n;
    return run_dif<T>(m, inverse != 0, first, last, ws, batch, st);
}

}  // namespace pyfft

using namespace pyfft;

extern "C" {

int64_t pyfft_bigfft_workspace_bytes(int dtype, int64_t nfft, int32_t batch) {
    if (nfft <= 0 || batch <= 0) return fail(PYFFT_EINVAL, "bigfft_workspace_bytes: bad arguments");
    const int64_t cb = dtype == PYFFT_F64 ? 16 : 8;
    return (int64_t)batch * nfft * cb + 4096 + 64 * (int64_t)batch;
}

int pyfft_hilbert(int dtype, const void* in, int64_t n, int32_t batch, void* analytic, void* envelope,
                  void* workspace, int64_t workspace_bytes, void* stream) {
    if (!in || batch <= 0 || (!analytic && !envelope))
        return fail(PYFFT_EINVAL, "hilbert: NULL input, batch <= 0 or no output requested");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == PYFFT_F32) return hilbert_impl<float>(in, n, batch, analytic, envelope, workspace, workspace_bytes, st);
    if (dtype == PYFFT_F64) return hilbert_impl<double>(in, n, batch, analytic, envelope, workspace, workspace_bytes, st);
    return fail(PYFFT_EINVAL, "unknown dtype %d", dtype);
}

int pyfft_xcorr(int dtype, const void* x, int64_t x_stride, const void* y, int64_t y_stride, int64_t n,
                int32_t pairs, int32_t normalise, void* out, void* workspace, int64_t workspace_bytes,
                void* stream) {
    if (!x || !y || !out || pairs <= 0) return fail(PYFFT_EINVAL, "xcorr: NULL pointer or pairs <= 0");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == PYFFT_F32)
        return xcorr_impl<float>(x, x_stride, y, y_stride, n, pairs, normalise, out, workspace, workspace_bytes, st);
    if (dtype == PYFFT_F64)
        return xcorr_impl<double>(x, x_stride, y, y_stride, n, pairs, normalise, out, workspace, workspace_bytes, st);
    return fail(PYFFT_EINVAL, "unknown dtype %d", dtype);
}

int pyfft_spectral_derivative(int dtype, const void* in, int64_t n, int32_t batch, int32_t modified, double dx,
                              void* out, void* workspace, int64_t workspace_bytes, void* stream) {
    if (!in || !out || batch <= 0) return fail(PYFFT_EINVAL, "spectral_derivative: NULL pointer or batch <= 0");
    if (!(dx > 0.0)) return fail(PYFFT_EINVAL, "spectral_derivative: the grid step must be positive");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == PYFFT_F32) return deriv_impl<float>(in, n, batch, modified, dx, out, workspace, workspace_bytes, st);
    if (dtype == PYFFT_F64) return deriv_impl<double>(in, n, batch, modified, dx, out, workspace, workspace_bytes, st);
    return fail(PYFFT_EINVAL, "unknown dtype %d", dtype);
}

int pyfft_fft_c2c(int dtype, const void* in, void* out, int64_t n, int32_t batch, int32_t inverse,
                  void* workspace, int64_t workspace_bytes, void* stream) {
    if (!in || !out || batch <= 0) return fail(PYFFT_EINVAL, "fft_c2c: NULL pointer or batch <= 0");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == PYFFT_F32) return c2c_impl<float>(in, out, n, batch, inverse, workspace, workspace_bytes, st);
    if (dtype == PYFFT_F64) return c2c_impl<double>(in, out, n, batch, inverse, workspace, workspace_bytes, st);
    return fail(PYFFT_EINVAL, "unknown dtype %d", dtype);
}

}  // extern "C"
