// K2: short-time Fourier transform with the per-segment spectra written out, and the
// spectral products of materialised STFTs.
//
// Reference lines replaced: fftanal.fft_win (fft_analysis.py:2126-2203), the loop of
// spectrogram.specgram (spectrogram.py:91-112), fftanal.Pstft + averagewins means
// (fft_analysis.py:1944-1990).
//
// Real input: TWO consecutive segments of one channel ride one complex FFT (segment 2p in
// the real part, 2p+1 in the imaginary part) and are separated by Hermitian symmetry, so a
// transform costs half of what the reference pays per segment.
#include "common.cuh"

namespace pyfft {

template <typename T> struct StftParams {
    const T* x;             // channel c at x + c*ch_stride (complex input: interleaved re,im)
    int64_t ch_stride;
    RecMap map;
    int64_t hop, seg_first, seg_count, chunk;   // chunk = FFTs (segment pairs for real input) per slot
    const T* window;
    const double* coef;     // [nch][2] or null
    int per_segment;        // remove each segment's own mean instead of coef
    int bin_first, nbins;
    const double* bin_scale;
    int power_out;
    void* out;              // [nch][seg_count][nbins] complex<T> or T (power)
    double* pseg;           // [nch][seg_count] or null
    const cpx<T>* tw;
};

template <int LOG2L> struct StftGeom {
    static constexpr int L = 1 << LOG2L;
    static constexpr int TPS = L / 16;
    static constexpr int THREADS = TPS > 256 ? TPS : 256;
    static constexpr int SLOTS = THREADS / TPS;
    static constexpr int E = smem_elems(L);
};

template <typename T, int LOG2L, bool SINGLE>
constexpr size_t stft_smem_bytes() {
    return (size_t)(SINGLE ? 1 : 2) * StftGeom<LOG2L>::SLOTS * StftGeom<LOG2L>::E * sizeof(cpx<T>) +
           (size_t)StftGeom<LOG2L>::L * sizeof(T) + (size_t)StftGeom<LOG2L>::SLOTS * 4 * sizeof(double);
}

// sum over the TPS threads of one slot (TPS <= 32: shuffles; else shared atomics on `red`)
template <int TPS>
__device__ __forceinline__ double slot_sum(double v, double* red, int t) {
    constexpr int W = TPS < 32 ? TPS : 32;
#pragma unroll
    for (int o = W / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o, W);
    if (TPS <= 32) return v;
    if ((t & 31) == 0) atomicAdd(red, v);
    return 0.0;    // caller reads *red after a barrier
}


// Fast path of the real-input STFT (bin_first == 0, nbins <= L/2 + 1: the one-sided spectra every
// caller on the path asks for).  Same arithmetic as stft_kernel; what differs is the integer work
// around it: segments outside the mirrored head are read through one base pointer with
// compile-time offsets, the output rows are written through two row pointers with compile-time
// offsets, the per-bin scale (with the 1/2 of the Hermitian split folded in) sits in registers,
// and only the upper half of the packed spectrum is exchanged for the split.
template <typename T, int LOG2L, bool SINGLE>
__global__ void __launch_bounds__(StftGeom<LOG2L>::THREADS,
                                  (sizeof(T) == 4 && StftGeom<LOG2L>::THREADS == 256) ? 2 : 1)
stft_real_kernel(const StftParams<T> p) {
    using G = StftGeom<LOG2L>;
    constexpr int L = G::L, TPS = G::TPS, SLOTS = G::SLOTS, E = G::E;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cpx<T>* sm = reinterpret_cast<cpx<T>*>(smem_raw);
    T* swin = reinterpret_cast<T*>(sm + (size_t)(SINGLE ? 1 : 2) * SLOTS * E);
    double* sred = reinterpret_cast<double*>(swin + L);     // [SLOTS][4]
    for (int i = threadIdx.x; i < L; i += G::THREADS) swin[i] = p.window[i];
    if (threadIdx.x < SLOTS * 4) sred[threadIdx.x] = 0.0;
    __syncthreads();

    const int slot = threadIdx.x / TPS;
    const int t = threadIdx.x - slot * TPS;
    const int c = blockIdx.y;
    const int part = blockIdx.x * SLOTS + slot;
    const int64_t nfft_total = (p.seg_count + 1) / 2;
    const int64_t q0 = (int64_t)part * p.chunk;
    const int64_t q1 = (q0 + p.chunk < nfft_total) ? q0 + p.chunk : nfft_total;
    const T* __restrict__ src = p.x + (int64_t)c * p.ch_stride;
    const double ca = p.coef ? p.coef[2 * c] : 0.0, cb = p.coef ? p.coef[2 * c + 1] : 0.0;
    const bool linear = cb != 0.0;
    const T cbf = (T)cb;
    double* red = sred + slot * 4;

    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(sm);
    Exchange<T, SINGLE> ex;
    ex.cur = sbase + (uint32_t)(slot * E * sizeof(cpx<T>));
    ex.other = SINGLE ? ex.cur : sbase + (uint32_t)((SLOTS + slot) * E * sizeof(cpx<T>));
    const Twiddles<T> tws = load_twiddles<T, LOG2L, false>(p.tw, t);

    // sc[s]: scale of bin t + TPS*s (s = 8: the Nyquist bin, thread 0 only), 0 for bins not asked for
    T sc[9];
    const T half = p.power_out ? T(0.25) : T(0.5);
#pragma unroll
    for (int s = 0; s < 9; ++s) {
        const int f = t + TPS * s;
        const bool ok = f < p.nbins && (s < 8 || t == 0);
        sc[s] = ok ? half * (p.bin_scale ? (T)p.bin_scale[f] : T(1)) : T(0);
    }
    const bool full = p.nbins >= L / 2;                      // every bin below Nyquist is stored
    const bool want_nyq = t == 0 && p.nbins > L / 2;
    const int64_t plain0 = p.map.i0 - p.map.reflect - p.map.src_base + t;

    for (int64_t it = 0; it < p.chunk; ++it) {               // same trip count for every slot (barriers inside)
        const int64_t q = q0 + it;
        const bool active = q < q1;
        const int64_t ka = 2 * q;
        const bool has_b = active && (ka + 1 < p.seg_count);
        cpx<T> v[16];
        T c0 = T(0), c1 = T(0);
        if (active) {
            const int64_t j0 = (p.seg_first + ka) * p.hop;
            c0 = (T)(ca + cb * (double)j0);
            c1 = (T)(ca + cb * (double)(j0 + p.hop));
            if (j0 >= p.map.reflect) {
                const T* __restrict__ sa = src + (j0 + plain0);
#pragma unroll
                for (int s = 0; s < 16; ++s) v[s].x = sa[TPS * s];
                if (has_b) {
                    const T* __restrict__ sb = sa + p.hop;
#pragma unroll
                    for (int s = 0; s < 16; ++s) v[s].y = sb[TPS * s];
                } else {
#pragma unroll
                    for (int s = 0; s < 16; ++s) v[s].y = c1;
                }
            } else {
                // segments touching the mirrored head of the record (fft_analysis.py:196-205)
#pragma unroll
                for (int s = 0; s < 16; ++s) {
                    const int64_t n = j0 + t + TPS * s;
                    v[s].x = src[src_index(p.map, n)];
                    v[s].y = has_b ? src[src_index(p.map, n + p.hop)] : c1;
                }
            }
            if (linear) {
#pragma unroll
                for (int s = 0; s < 16; ++s) {
                    const T n = (T)(t + TPS * s);
                    v[s].x -= c0 + cbf * n;
                    v[s].y -= c1 + cbf * n;
                }
                c0 = T(0);
                c1 = T(0);
                if (!has_b) {
#pragma unroll
                    for (int s = 0; s < 16; ++s) v[s].y = T(0);
                }
            }
        } else {
#pragma unroll
            for (int s = 0; s < 16; ++s) v[s] = mk<T>(T(0), T(0));
        }
        // the reductions below contain barriers / full-warp shuffles: every thread of the CTA runs
        // them (p.per_segment and p.pseg are uniform), inactive slots on zeros
        if (p.per_segment) {                                 // detrendwin=True (fft_analysis.py:2169-2172)
            double sx = 0.0, sy = 0.0;
#pragma unroll
            for (int s = 0; s < 16; ++s) { sx += (double)(v[s].x - c0); sy += (double)(v[s].y - c1); }
            sx = slot_sum<TPS>(sx, red + 0, t);
            sy = slot_sum<TPS>(sy, red + 1, t);
            if (TPS > 32) {
                __syncthreads();
                sx = red[0]; sy = red[1];
                __syncthreads();
                if (t == 0) { red[0] = 0.0; red[1] = 0.0; }
            }
            c0 += (T)(sx / L);
            c1 += (T)(sy / L);
        }
#pragma unroll
        for (int s = 0; s < 16; ++s) {
            const T wv = swin[t + TPS * s];
            v[s] = mk<T>((v[s].x - c0) * wv, (v[s].y - c1) * wv);
        }
        if (p.pseg) {
            // trapezoid rule core of fft_analysis.py:2174: sum g - (g_0 + g_{L-1})/2
            double pa = 0.0, pb = 0.0;
#pragma unroll
            for (int s = 0; s < 16; ++s) {
                const bool edge = (s == 0 && t == 0) || (s == 15 && t == TPS - 1);
                const double wgt = edge ? 0.5 : 1.0;
                pa += wgt * (double)v[s].x * v[s].x;
                pb += wgt * (double)v[s].y * v[s].y;
            }
            pa = slot_sum<TPS>(pa, red + 2, t);
            pb = slot_sum<TPS>(pb, red + 3, t);
            if (TPS > 32) {
                __syncthreads();
                pa = red[2]; pb = red[3];
                __syncthreads();
                if (t == 0) { red[2] = 0.0; red[3] = 0.0; }
            }
            if (t == 0 && active) {
                p.pseg[(int64_t)c * p.seg_count + ka] = pa;
                if (has_b) p.pseg[(int64_t)c * p.seg_count + ka + 1] = pb;
            }
        }
        fft_block<T, LOG2L, false, SINGLE>(v, ex, tws, t);

        // the split needs Z[L - f] for the bins f = t + TPS*s, s < 8: partner thread TPS - t, slot
        // 15 - s (thread 0 pairs with itself, slot 16 - s) -- only the upper half is exchanged, in
        // the unpadded layout [slot - 8][t] (ascending writes, descending reads: conflict-free)
        ex.before_write();
        const uint32_t e = ex.cur;
        constexpr uint32_t CB = (uint32_t)sizeof(cpx<T>);
        if (TPS % 16 == 0) {
            const uint32_t a = e + (uint32_t)t * CB;
#pragma unroll
            for (int s = 8; s < 16; ++s) SharedIO<T>::st(a + (uint32_t)((s - 8) * TPS) * CB, v[s]);
        } else {
#pragma unroll
            for (int s = 0; s < 16; ++s) ex.st(e, t + TPS * s, v[s]);
        }
        __syncthreads();
        ex.flip();
        if (!active) continue;
        const int64_t row = ((int64_t)c * p.seg_count + ka) * p.nbins + t;
        const uint32_t pa = e + (uint32_t)(TPS - t) * CB;
        // 2U = A + conj B (segment ka), 2V = (A - conj B)/i (segment ka + 1); the 1/2 lives in sc[]
        auto split = [&](int s, cpx<T>& U, cpx<T>& V) {
            const cpx<T> A = v[s];
            cpx<T> B;
            if (TPS % 16 == 0) {
                B = SharedIO<T>::ld(pa + (uint32_t)((7 - s) * TPS) * CB);
                if (s == 0 && t == 0) B = A;                 // DC pairs with itself
            } else {
                B = ex.ld(e, (L - (t + TPS * s)) & (L - 1));
            }
            U = mk<T>(A.x + B.x, A.y - B.y);
            V = mk<T>(A.y + B.y, B.x - A.x);
        };
        if (p.power_out) {
            T* __restrict__ oa = static_cast<T*>(p.out) + row;
            T* __restrict__ ob = oa + p.nbins;
#pragma unroll
            for (int s = 0; s < 8; ++s) {
                cpx<T> U, V;
                split(s, U, V);
                if (full || t + TPS * s < p.nbins) {
                    oa[TPS * s] = (U.x * U.x + U.y * U.y) * sc[s];
                    if (has_b) ob[TPS * s] = (V.x * V.x + V.y * V.y) * sc[s];
                }
            }
            if (want_nyq) {
                oa[L / 2] = T(4) * v[8].x * v[8].x * sc[8];
                if (has_b) ob[L / 2] = T(4) * v[8].y * v[8].y * sc[8];
            }
        } else {
            cpx<T>* __restrict__ oa = static_cast<cpx<T>*>(p.out) + row;
            cpx<T>* __restrict__ ob = oa + p.nbins;
#pragma unroll
            for (int s = 0; s < 8; ++s) {
                cpx<T> U, V;
                split(s, U, V);
                if (full || t + TPS * s < p.nbins) {
                    oa[TPS * s] = mk<T>(U.x * sc[s], U.y * sc[s]);
                    if (has_b) ob[TPS * s] = mk<T>(V.x * sc[s], V.y * sc[s]);
                }
            }
            if (want_nyq) {
                oa[L / 2] = mk<T>(T(2) * v[8].x * sc[8], T(0));
                if (has_b) ob[L / 2] = mk<T>(T(2) * v[8].y * sc[8], T(0));
            }
        }
    }
}

template <typename T, int LOG2L, bool CPLX, bool SINGLE>
__global__ void __launch_bounds__(StftGeom<LOG2L>::THREADS) stft_kernel(const StftParams<T> p) {
    using G = StftGeom<LOG2L>;
    constexpr int L = G::L, TPS = G::TPS, SLOTS = G::SLOTS, E = G::E;
    constexpr int SPF = CPLX ? 1 : 2;                       // segments per FFT
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cpx<T>* sm = reinterpret_cast<cpx<T>*>(smem_raw);
    T* swin = reinterpret_cast<T*>(sm + (size_t)(SINGLE ? 1 : 2) * SLOTS * E);
    double* sred = reinterpret_cast<double*>(swin + L);     // [SLOTS][4]
    for (int i = threadIdx.x; i < L; i += G::THREADS) swin[i] = p.window[i];
    if (threadIdx.x < SLOTS * 4) sred[threadIdx.x] = 0.0;
    __syncthreads();

    const int slot = threadIdx.x / TPS;
    const int t = threadIdx.x - slot * TPS;
    const int c = blockIdx.y;
    const int part = blockIdx.x * SLOTS + slot;
    const int64_t nfft_total = (p.seg_count + SPF - 1) / SPF;
    const int64_t q0 = (int64_t)part * p.chunk;
    const int64_t q1 = (q0 + p.chunk < nfft_total) ? q0 + p.chunk : nfft_total;
    const T* __restrict__ src = p.x + (int64_t)c * p.ch_stride;
    const double ca = p.coef ? p.coef[2 * c] : 0.0, cb = p.coef ? p.coef[2 * c + 1] : 0.0;
    double* red = sred + slot * 4;

    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(sm);
    Exchange<T, SINGLE> ex;
    ex.cur = sbase + (uint32_t)(slot * E * sizeof(cpx<T>));
    ex.other = SINGLE ? ex.cur : sbase + (uint32_t)((SLOTS + slot) * E * sizeof(cpx<T>));
    const Twiddles<T> tws = load_twiddles<T, LOG2L, false>(p.tw, t);

    for (int64_t it = 0; it < p.chunk; ++it) {
        const int64_t q = q0 + it;
        const bool active = q < q1;
        const int64_t ka = q * SPF;                          // first segment of this FFT (local index)
        const bool has_b = !CPLX && active && (ka + 1 < p.seg_count);
        cpx<T> v[16];
#pragma unroll
        for (int s = 0; s < 16; ++s) v[s] = mk<T>(T(0), T(0));
        if (active) {
#pragma unroll
            for (int h = 0; h < SPF; ++h) {
                if (h == 1 && !has_b) break;
                const int64_t j0 = (p.seg_first + ka + h) * p.hop;
                const T tr0 = (T)(ca + cb * (double)j0), trb = (T)cb;
#pragma unroll
                for (int s = 0; s < 16; ++s) {
                    const int n = t + TPS * s;
                    const int64_t e = src_index(p.map, j0 + n);
                    if (CPLX) {
                        v[s] = mk<T>(src[2 * e], src[2 * e + 1]);
                    } else {
                        const T trend = tr0 + trb * (T)n;
                        if (h == 0) v[s].x = src[e] - trend; else v[s].y = src[e] - trend;
                    }
                }
            }
            if (CPLX && p.coef) {
                // complex records are detrended by their (complex) mean: coef holds re/im means
                // as two "channels" 2c, 2c+1 (slope unused)
                const T mr = (T)p.coef[4 * c], mi = (T)p.coef[4 * c + 2];
#pragma unroll
                for (int s = 0; s < 16; ++s) { v[s].x -= mr; v[s].y -= mi; }
            }
        }
        if (p.per_segment) {                                 // detrendwin=True (fft_analysis.py:2169-2172)
            double sx = 0.0, sy = 0.0;
#pragma unroll
            for (int s = 0; s < 16; ++s) { sx += (double)v[s].x; sy += (double)v[s].y; }
            sx = slot_sum<TPS>(sx, red + 0, t);
            sy = slot_sum<TPS>(sy, red + 1, t);
            if (TPS > 32) {
                __syncthreads();
                sx = red[0]; sy = red[1];
                __syncthreads();
                if (t == 0) { red[0] = 0.0; red[1] = 0.0; }
            }
            const T mx = (T)(sx / L), my = (T)(sy / L);
#pragma unroll
            for (int s = 0; s < 16; ++s) { v[s].x -= mx; v[s].y -= my; }
        }
        double pa = 0.0, pb = 0.0;
#pragma unroll
        for (int s = 0; s < 16; ++s) {
            const T wv = swin[t + TPS * s];
            v[s].x *= wv; v[s].y *= wv;
            if (p.pseg) {
                // trapezoid rule core of fft_analysis.py:2174: sum g - (g_0 + g_{L-1})/2
                const bool edge = (s == 0 && t == 0) || (s == 15 && t == TPS - 1);
                const double wgt = edge ? 0.5 : 1.0;
                if (CPLX) pa += wgt * ((double)v[s].x * v[s].x + (double)v[s].y * v[s].y);
                else { pa += wgt * (double)v[s].x * v[s].x; pb += wgt * (double)v[s].y * v[s].y; }
            }
        }
        if (p.pseg) {
            pa = slot_sum<TPS>(pa, red + 2, t);
            pb = slot_sum<TPS>(pb, red + 3, t);
            if (TPS > 32) {
                __syncthreads();
                pa = red[2]; pb = red[3];
                __syncthreads();
                if (t == 0) { red[2] = 0.0; red[3] = 0.0; }
            }
            if (t == 0 && active) {
                p.pseg[(int64_t)c * p.seg_count + ka] = pa;
                if (has_b) p.pseg[(int64_t)c * p.seg_count + ka + 1] = pb;
            }
        }
        fft_block<T, LOG2L, false, SINGLE>(v, ex, tws, t);

        if (CPLX) {
            if (active) {
#pragma unroll
                for (int s = 0; s < 16; ++s) {
                    const int f = t + TPS * s;
                    const int o = (f - p.bin_first) & (L - 1);
                    if (o < p.nbins) {
                        const T sc = p.bin_scale ? (T)p.bin_scale[o] : T(1);
                        const int64_t idx = ((int64_t)c * p.seg_count + ka) * p.nbins + o;
                        if (p.power_out) static_cast<T*>(p.out)[idx] = (v[s].x * v[s].x + v[s].y * v[s].y) * sc;
                        else static_cast<cpx<T>*>(p.out)[idx] = mk<T>(v[s].x * sc, v[s].y * sc);
                    }
                }
            }
            continue;
        }
        // real input: split the two packed segments with the partner bins Z[L - f]
        ex.before_write();
        const uint32_t e = ex.cur;
#pragma unroll
        for (int s = 0; s < 16; ++s) ex.st(e, t + TPS * s, v[s]);
        __syncthreads();
        ex.flip();
        if (active) {
#pragma unroll
            for (int s = 0; s < 9; ++s) {
                if (s == 8 && t != 0) break;                 // Nyquist bin: slot 8 of thread 0 only
                const int f = t + TPS * s;
                const cpx<T> A = v[s];
                const cpx<T> B = ex.ld(e, (L - f) & (L - 1));
                // U = (A + conj B)/2 (segment ka), V = (A - conj B)/(2i) (segment ka + 1)
                const cpx<T> U = mk<T>(T(0.5) * (A.x + B.x), T(0.5) * (A.y - B.y));
                const cpx<T> V = mk<T>(T(0.5) * (A.y + B.y), T(0.5) * (B.x - A.x));
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    if (h == 1 && !has_b) break;
                    const cpx<T> X = h == 0 ? U : V;
                    const int64_t row = ((int64_t)c * p.seg_count + ka + h) * p.nbins;
#pragma unroll
                    for (int m = 0; m < 2; ++m) {            // bin f and its mirror L - f (conjugate)
                        if (m == 1 && (f == 0 || f == L / 2)) break;
                        const int fb = m == 0 ? f : L - f;
                        const int o = (fb - p.bin_first) & (L - 1);
                        if (o < p.nbins) {
                            const T sc = p.bin_scale ? (T)p.bin_scale[o] : T(1);
                            if (p.power_out) static_cast<T*>(p.out)[row + o] = (X.x * X.x + X.y * X.y) * sc;
                            else static_cast<cpx<T>*>(p.out)[row + o] = mk<T>(X.x * sc, (m == 0 ? X.y : -X.y) * sc);
                        }
                    }
                }
            }
        }
    }
}

// sums over segments of X conj(X), Y conj(Y), Y conj(X) (or X conj(Y)) from materialised STFTs
template <typename T>
__global__ void __launch_bounds__(256) stft_products_kernel(const cpx<T>* __restrict__ X, const cpx<T>* __restrict__ Y,
                                                            int64_t K, int nb, int conj_x, double* __restrict__ sxx,
                                                            double* __restrict__ syy, double* __restrict__ sxy) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    const int c = blockIdx.y;
    if (b >= nb) return;
    double axx = 0.0, ayy = 0.0, ar = 0.0, ai = 0.0;
    const cpx<T>* yc = Y ? Y + (int64_t)c * K * nb : nullptr;
    for (int64_t k = 0; k < K; ++k) {
        const cpx<T> x = X[k * nb + b];
        const double xr = x.x, xi = x.y;
        axx += xr * xr + xi * xi;
        if (yc) {
            const cpx<T> y = yc[k * nb + b];
            const double yr = y.x, yi = y.y;
            ayy += yr * yr + yi * yi;
            ar += yr * xr + yi * xi;                        // Re Y conj(X) == Re X conj(Y)
            ai += conj_x ? (yi * xr - yr * xi) : (xi * yr - xr * yi);
        }
    }
    if (c == 0 && sxx) sxx[b] = axx;
    if (yc) {
        syy[(int64_t)c * nb + b] = ayy;
        sxy[2 * ((int64_t)c * nb + b)] = ar;
        sxy[2 * ((int64_t)c * nb + b) + 1] = ai;
    }
}

// grid.x per channel so that the CTAs fill the resident slots of the device a whole number of
// times (a partial last wave costs as much as a full one): best fill among 1..4 waves
static int64_t stft_parts(int resident, int nch, int64_t maxb) {
    int64_t best = 1;
    double best_fill = 0.0;
    for (int w = 1; w <= 4; ++w) {
        const int64_t gx = ((int64_t)resident * w) / nch;
        if (gx < 1) continue;
        const double fill = (double)(gx * nch) / ((double)resident * w);
        if (fill > best_fill + 0.02) { best_fill = fill; best = gx; }
    }
    if (best > maxb) best = maxb;
    return best < 1 ? 1 : best;
}

template <typename T, int LOG2L, bool CPLX>
static int launch_stft(StftParams<T> p, int nch, cudaStream_t st) {
    using G = StftGeom<LOG2L>;
    constexpr bool SINGLE = stft_smem_bytes<T, LOG2L, false>() > 200 * 1024;
    constexpr size_t smem = stft_smem_bytes<T, LOG2L, SINGLE>();
    int status;
    p.tw = reinterpret_cast<const cpx<T>*>(twiddle_table(sizeof(T) == 8 ? PYFFT_F64 : PYFFT_F32, G::L, &status));
    if (!p.tw) return status;
    const int64_t nfft_total = (p.seg_count + (CPLX ? 0 : 1)) / (CPLX ? 1 : 2);
    int sms = device_sm_count();
    if (sms <= 0) sms = 148;
    const int64_t maxb = (nfft_total + G::SLOTS - 1) / G::SLOTS;
    const bool fast = !CPLX && p.bin_first == 0 && p.nbins <= G::L / 2 + 1;
    if (fast) {
        auto kern = stft_real_kernel<T, LOG2L, SINGLE>;
        PYFFT_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int per_sm = 0;
        PYFFT_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, G::THREADS, smem));
        if (per_sm < 1) per_sm = 1;
        const int64_t gx = stft_parts(per_sm * sms, nch, maxb);
        p.chunk = (nfft_total + gx * G::SLOTS - 1) / (gx * G::SLOTS);
        kern<<<dim3((unsigned)gx, (unsigned)nch), G::THREADS, smem, st>>>(p);
        PYFFT_LAUNCH_OK("stft_real_kernel");
        return PYFFT_OK;
    }
    int64_t gx = ((int64_t)4 * sms + nch - 1) / nch;
    if (gx > maxb) gx = maxb;
    if (gx < 1) gx = 1;
    p.chunk = (nfft_total + gx * G::SLOTS - 1) / (gx * G::SLOTS);
    auto kern = stft_kernel<T, LOG2L, CPLX, SINGLE>;
    PYFFT_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<dim3((unsigned)gx, (unsigned)nch), G::THREADS, smem, st>>>(p);
    PYFFT_LAUNCH_OK("stft_kernel");
    return PYFFT_OK;
}

template <typename T>
static int dispatch_stft(int log2l, bool cplx, const StftParams<T>& p, int nch, cudaStream_t st) {
#define PYFFT_CASE(N) \
    case N: return cplx ? launch_stft<T, N, true>(p, nch, st) : launch_stft<T, N, false>(p, nch, st);
    switch (log2l) {
        PYFFT_CASE(4) PYFFT_CASE(5) PYFFT_CASE(6) PYFFT_CASE(7) PYFFT_CASE(8)
        PYFFT_CASE(9) PYFFT_CASE(10) PYFFT_CASE(11) PYFFT_CASE(12) PYFFT_CASE(13)
    }
#undef PYFFT_CASE
    return fail(PYFFT_EUNSUPPORTED, "nfft = 2^%d not supported", log2l);
}

template <typename T>
static int stft_entry(int32_t complex_in, const void* x, int64_t ch_stride, int32_t nch, const RecMap& m,
                      const pyfft_segmentation* seg, int log2l, const void* window, const double* coef,
                      int32_t per_segment, int32_t bin_first, int32_t nbins, const double* bin_scale,
                      int32_t power_out, void* out, double* pseg, cudaStream_t st) {
    StftParams<T> p;
    p.x = static_cast<const T*>(x);
    p.ch_stride = ch_stride;
    p.map = m;
    p.hop = seg->hop;
    p.seg_first = seg->seg_first;
    p.seg_count = seg->seg_count;
    p.chunk = 0;
    p.window = static_cast<const T*>(window);
    p.coef = coef;
    p.per_segment = per_segment;
    p.bin_first = bin_first;
    p.nbins = nbins;
    p.bin_scale = bin_scale;
    p.power_out = power_out;
    p.out = out;
    p.pseg = pseg;
    p.tw = nullptr;
    return dispatch_stft<T>(log2l, complex_in != 0, p, nch, st);
}

}  // namespace pyfft

using namespace pyfft;

extern "C" {

int pyfft_stft(int dtype, int32_t complex_in, const void* x, int64_t ch_stride, int32_t nch,
               const pyfft_record_map* map, const pyfft_segmentation* seg, const void* window,
               const double* coef, int32_t detrend_per_segment, int32_t bin_first, int32_t nbins,
               const double* bin_scale, int32_t power_out, void* out, double* pseg, void* stream) {
    RecMap m;
    int rc = check_map(map, &m);
    if (rc) return rc;
    int log2l;
    if ((rc = check_seg(seg, 4, 13, &log2l))) return rc;
    if (!x || !window || !out || nch <= 0) return fail(PYFFT_EINVAL, "stft: NULL pointer or nch <= 0");
    if (nbins <= 0 || nbins > seg->nfft || bin_first < 0 || bin_first >= seg->nfft)
        return fail(PYFFT_EINVAL, "stft: need 0 < nbins <= nfft and 0 <= bin_first < nfft");
    if (seg->seg_count == 0) return PYFFT_OK;
    const int64_t j_lo = seg->seg_first * (int64_t)seg->hop;
    const int64_t j_hi = (seg->seg_first + seg->seg_count - 1) * (int64_t)seg->hop + seg->nfft;
    if ((rc = check_range(m, j_lo, j_hi))) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == PYFFT_F32)
        return stft_entry<float>(complex_in, x, ch_stride, nch, m, seg, log2l, window, coef, detrend_per_segment,
                                 bin_first, nbins, bin_scale, power_out, out, pseg, st);
    if (dtype == PYFFT_F64)
        return stft_entry<double>(complex_in, x, ch_stride, nch, m, seg, log2l, window, coef, detrend_per_segment,
                                  bin_first, nbins, bin_scale, power_out, out, pseg, st);
    return fail(PYFFT_EINVAL, "unknown dtype %d", dtype);
}

int pyfft_stft_products(int dtype, const void* X, const void* Y, int32_t nch, int64_t K, int32_t nb,
                        int32_t conj_x, double* sxx, double* syy, double* sxy, void* stream) {
    if (!X || K <= 0 || nb <= 0 || nch <= 0) return fail(PYFFT_EINVAL, "stft_products: bad arguments");
    if (Y && (!syy || !sxy)) return fail(PYFFT_EINVAL, "stft_products: syy/sxy required when Y is given");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    dim3 grid((nb + 255) / 256, nch);
    if (dtype == PYFFT_F32)
        stft_products_kernel<float><<<grid, 256, 0, st>>>(static_cast<const cpx<float>*>(X),
            static_cast<const cpx<float>*>(Y), K, nb, conj_x, sxx, syy, sxy);
    else if (dtype == PYFFT_F64)
        stft_products_kernel<double><<<grid, 256, 0, st>>>(static_cast<const cpx<double>*>(X),
            static_cast<const cpx<double>*>(Y), K, nb, conj_x, sxx, syy, sxy);
    else
        return fail(PYFFT_EINVAL, "unknown dtype %d", dtype);
    PYFFT_LAUNCH_OK("stft_products_kernel");
    return PYFFT_OK;
}

}  // extern "C"
