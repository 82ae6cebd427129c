// K0 (detrend statistics), K1 (fused Welch auto/cross accumulation), the
// partial-sum reduction, the finalize kernel and the correlation epilogue.
//
// Reference lines replaced: fft_analysis.py:353-393 (detrend, segment loop,
// window, FFT, products), :402-446 (side selection, scaling, mean),
// :1662-1680 (coherence), :520 (phase), :544-597 (correlations).
#include <map>
#include <mutex>

#include "common.cuh"

namespace pyfft {

// ------------------------------------------------------------------------- //
// K0: per-channel sum x(j) and sum j*x(j) over analysed indices
// ------------------------------------------------------------------------- //
template <typename T>
__global__ void __launch_bounds__(256) detrend_stats_kernel(const T* __restrict__ first, const T* __restrict__ ptr,
                                                            int64_t ch_stride, RecMap map, int64_t j_begin,
                                                            int64_t j_end, double* __restrict__ stats) {
    // channel c of the stats: `first` (when given) is channel 0 and ptr holds the following ones
    const int c = blockIdx.y;
    const T* p = first ? (c == 0 ? first : ptr + (int64_t)(c - 1) * ch_stride) : ptr + (int64_t)c * ch_stride;
    double s0 = 0.0, s1 = 0.0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    // mirrored head of the record (at most nwins-1 samples): element-wise map
    const int64_t j_mid = j_begin > map.reflect ? j_begin : (map.reflect < j_end ? map.reflect : j_end);
    for (int64_t j = j_begin + tid; j < j_mid; j += stride) {
        const double v = (double)p[src_index(map, j)];
        s0 += v;
        s1 += v * (double)j;
    }
    // plain part: contiguous in memory, 8 independent coalesced loads in flight per thread
    const T* __restrict__ q = p + (map.i0 - map.reflect - map.src_base);
    int64_t j = j_mid + tid;
    for (; j + 7 * stride < j_end; j += 8 * stride) {
        T r[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) r[i] = q[j + i * stride];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const double v = (double)r[i];
            s0 += v;
            s1 += v * (double)(j + i * stride);
        }
    }
    for (; j < j_end; j += stride) {
        const double v = (double)q[j];
        s0 += v;
        s1 += v * (double)j;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s0 += __shfl_xor_sync(0xffffffffu, s0, o);
        s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    }
    __shared__ double sh[2][8];
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) { sh[0][w] = s0; sh[1][w] = s1; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int i = 0; i < 8; ++i) { a += sh[0][i]; b += sh[1][i]; }
        atomicAdd(&stats[2 * c + 0], a);
        atomicAdd(&stats[2 * c + 1], b);
    }
}

__global__ void detrend_coef_kernel(int mode, const double* __restrict__ stats, int nch, int64_t nsig,
                                    double* __restrict__ coef) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= nch) return;
    const double n = (double)nsig;
    const double mean = stats[2 * c] / n;
    double a = 0.0, b = 0.0;
    if (mode == PYFFT_DETREND_MEAN) {
        a = mean;
    } else if (mode == PYFFT_DETREND_LINEAR) {
        // t_j = j - (n-1)/2 ; slope = sum t_j (x_j - mean) / sum t_j^2 ; sum t_j = 0
        const double tc = 0.5 * (n - 1.0);
        const double num = stats[2 * c + 1] - tc * stats[2 * c];
        const double den = n * (n * n - 1.0) / 12.0;
        b = den > 0.0 ? num / den : 0.0;
        a = mean - b * tc;
    }
    coef[2 * c] = a;
    coef[2 * c + 1] = b;
}

// ------------------------------------------------------------------------- //
// K1: fused gather -> detrend -> window -> FFT(u + i v) -> split -> accumulate
// ------------------------------------------------------------------------- //
template <typename T> struct WelchParams {
    const T* x;
    const T* y;
    int64_t y_stride;
    RecMap map;
    int64_t hop, seg_first, seg_count, chunk;
    const T* window;
    const double* coef_x;
    const double* coef_y;
    const cpx<T>* tw;
    T* partial;     // [nch][nparts][4][L/2+1]
    int nparts;
    // mean-fused mode (FUSE): coef holds a PROVISIONAL mean; the kernel also returns what the
    // finalize step needs to move the result to the exact global mean by linearity
    T* lin_partial;     // [nch][nparts][4][L/16]: sum_k U_k, V_k (re, im) for bins f < L/16
    double* dev_sums;   // [nch+1]: sum of (sample - provisional mean) over the first `hop`
                        // samples of every segment of this call (x first, then the y channels)
};

template <int LOG2L> struct WelchGeom {
    static constexpr int L = 1 << LOG2L;
    static constexpr int TPS = L / 16;
    static constexpr int THREADS = TPS > 256 ? TPS : 256;
    static constexpr int SLOTS = THREADS / TPS;
    static constexpr int NBIN = L / 2 + 1;
    static constexpr int E = smem_elems(L);
};

template <typename T, int LOG2L, bool SINGLE>
constexpr size_t welch_smem_bytes() {
    return (size_t)(SINGLE ? 1 : 2) * WelchGeom<LOG2L>::SLOTS * WelchGeom<LOG2L>::E * sizeof(cpx<T>) +
           (size_t)WelchGeom<LOG2L>::L * sizeof(T);
}

template <typename T, int LOG2L, bool SINGLE, bool FUSE>
__global__ void __launch_bounds__(WelchGeom<LOG2L>::THREADS,
                                  (sizeof(T) == 4 && WelchGeom<LOG2L>::THREADS == 256) ? 2 : 1)
welch_pair_kernel(const WelchParams<T> p) {
    using G = WelchGeom<LOG2L>;
    constexpr int L = G::L, TPS = G::TPS, SLOTS = G::SLOTS, NBIN = G::NBIN, E = G::E;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cpx<T>* sm = reinterpret_cast<cpx<T>*>(smem_raw);
    T* swin = reinterpret_cast<T*>(sm + (size_t)(SINGLE ? 1 : 2) * SLOTS * E);
    for (int i = threadIdx.x; i < L; i += G::THREADS) swin[i] = p.window[i];
    __syncthreads();

    const int slot = threadIdx.x / TPS;
    const int t = threadIdx.x - slot * TPS;
    const int pair = blockIdx.y;
    const int part = blockIdx.x * SLOTS + slot;
    const int64_t k0 = (int64_t)part * p.chunk;
    const int64_t k1 = (k0 + p.chunk < p.seg_count) ? k0 + p.chunk : p.seg_count;
    const T* __restrict__ u = p.x;
    const T* __restrict__ w = p.y + (int64_t)pair * p.y_stride;
    const double ax = p.coef_x ? p.coef_x[0] : 0.0, bx = p.coef_x ? p.coef_x[1] : 0.0;
    const double ay = p.coef_y ? p.coef_y[2 * pair] : 0.0, by = p.coef_y ? p.coef_y[2 * pair + 1] : 0.0;
    const T bxf = (T)bx, byf = (T)by;
    const bool linear = (bx != 0.0) || (by != 0.0);

    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(sm);
    Exchange<T, SINGLE> ex;
    ex.cur = sbase + (uint32_t)(slot * E * sizeof(cpx<T>));
    ex.other = SINGLE ? ex.cur : sbase + (uint32_t)((SLOTS + slot) * E * sizeof(cpx<T>));
    const Twiddles<T> tws = load_twiddles<T, LOG2L, false>(p.tw, t);

    T acc[8][4];
    T nyq[4] = {T(0), T(0), T(0), T(0)};
    T lin[4] = {T(0), T(0), T(0), T(0)};
    T devx = T(0), devy = T(0);
    const int hs = (int)(p.hop / TPS), hrem = (int)(p.hop - (int64_t)hs * TPS);
#pragma unroll
    for (int s = 0; s < 8; ++s)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[s][q] = T(0);

    for (int64_t it = 0; it < p.chunk; ++it) {       // same trip count for every slot (barriers inside)
        const int64_t k = k0 + it;
        const bool active = k < k1;
        cpx<T> v[16];
        if (active) {
            const int64_t j0 = (p.seg_first + k) * p.hop;
            T cx0 = (T)(ax + bx * (double)j0), cy0 = (T)(ay + by * (double)j0);
            if (j0 >= p.map.reflect) {
                // plain segment: one base offset, 32-bit element offsets, coalesced 4-byte loads
                const int64_t e0 = j0 - p.map.reflect + p.map.i0 - p.map.src_base;
                const T* __restrict__ us = u + e0 + t;
                const T* __restrict__ ws = w + e0 + t;
#pragma unroll
                for (int s = 0; s < 16; ++s) v[s] = mk<T>(us[TPS * s], ws[TPS * s]);
                // pull the NEW samples of the next segment of this slot into L2 (one 128-byte line
                // per thread) so that its loads one iteration from now are L2 hits, not DRAM misses
                if (k + 1 < k1) {
                    constexpr int LINE = 128 / (int)sizeof(T);
                    const int64_t hop_lines = (p.hop + LINE - 1) / LINE;
                    if (t < 2 * hop_lines) {
                        const T* base = (t < hop_lines ? u : w) + e0 + L;
                        const int64_t ln = t < hop_lines ? t : t - hop_lines;
                        asm volatile("prefetch.global.L1 [%0];" ::"l"(base + ln * LINE));
                    }
                }
            } else {
                // segment touching the mirrored head of the record (fft_analysis.py:196-205)
#pragma unroll
                for (int s = 0; s < 16; ++s) {
                    const int64_t e = src_index(p.map, j0 + t + TPS * s);
                    v[s] = mk<T>(u[e], w[e]);
                }
            }
            if (linear) {
#pragma unroll
                for (int s = 0; s < 16; ++s) {
                    const T n = (T)(t + TPS * s);
                    v[s].x -= cx0 + bxf * n;
                    v[s].y -= cy0 + byf * n;
                }
                cx0 = T(0);
                cy0 = T(0);
            }
            if (FUSE) {
                // deviations of the first `hop` samples of the segment: every analysed sample of
                // [seg_first*hop, (seg_first+seg_count)*hop) is counted exactly once
                if (hs == 8 && hrem == 0) {             // 50 % overlap: no predicates
                    const T sx = ((v[0].x + v[1].x) + (v[2].x + v[3].x)) + ((v[4].x + v[5].x) + (v[6].x + v[7].x));
                    const T sy = ((v[0].y + v[1].y) + (v[2].y + v[3].y)) + ((v[4].y + v[5].y) + (v[6].y + v[7].y));
                    devx += sx - T(8) * cx0;
                    devy += sy - T(8) * cy0;
                } else {
#pragma unroll
                    for (int s = 0; s < 16; ++s)
                        if (s < hs || (s == hs && t < hrem)) { devx += v[s].x - cx0; devy += v[s].y - cy0; }
                }
            }
#pragma unroll
            for (int s = 0; s < 16; ++s) {
                const T wv = swin[t + TPS * s];
                v[s] = mk<T>((v[s].x - cx0) * wv, (v[s].y - cy0) * wv);
            }
        } else {
#pragma unroll
            for (int s = 0; s < 16; ++s) v[s] = mk<T>(T(0), T(0));
        }
        fft_block<T, LOG2L, false, SINGLE>(v, ex, tws, t);

        // exchange the upper half of the spectrum so that every thread can fetch Z[L-f] for its
        // bins f = t + TPS*s, s < 8: the partner is thread TPS - t, slot 15 - s (thread 0 pairs with
        // itself: slot 16 - s, taken from its own registers).  Unpadded layout [slot-8][t]: ascending
        // writes and descending reads of 16 consecutive 8-byte words are both conflict-free.
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
        if (active) {
            const uint32_t pa = e + (uint32_t)(TPS - t) * CB;
#pragma unroll
            for (int s = 0; s < 8; ++s) {
                const int f = t + TPS * s;
                const cpx<T> A = v[s];
                cpx<T> B;
                if (TPS % 16 == 0) {
                    B = SharedIO<T>::ld(pa + (uint32_t)((7 - s) * TPS) * CB);
                    // thread 0 pairs with itself (bin TPS*s with TPS*(16-s)): its slot 16-s sits in row
                    // 8-s of column 0, exactly where this address points -- only DC (s = 0) is special
                    if (s == 0 && t == 0) B = A;
                } else {
                    B = ex.ld(e, (L - f) & (L - 1));
                }
                // U = (A + conj B)/2, V = (A - conj B)/(2i); the factors 1/2 are applied at the end
                const T ur = A.x + B.x, ui = A.y - B.y, vr = A.y + B.y, vi = B.x - A.x;
                acc[s][0] += ur * ur + ui * ui;
                acc[s][1] += vr * vr + vi * vi;
                acc[s][2] += vr * ur + vi * ui;      // Re V conj(U)   (fft_analysis.py:393: Y conj(X))
                acc[s][3] += vi * ur - vr * ui;      // Im V conj(U)
                if (FUSE && s == 0) { lin[0] += ur; lin[1] += ui; lin[2] += vr; lin[3] += vi; }
            }
            if (t == 0) {                             // Nyquist bin f = L/2 = slot 8 of thread 0; self-paired
                const T ur = v[8].x + v[8].x, vr = v[8].y + v[8].y;
                nyq[0] += ur * ur;
                nyq[1] += vr * vr;
                nyq[2] += vr * ur;
            }
        }
    }

    T* out = p.partial + ((size_t)pair * p.nparts + part) * 4 * NBIN;
#pragma unroll
    for (int s = 0; s < 8; ++s)
#pragma unroll
        for (int q = 0; q < 4; ++q) out[q * NBIN + t + TPS * s] = T(0.25) * acc[s][q];
    if (t == 0) {
#pragma unroll
        for (int q = 0; q < 4; ++q) out[q * NBIN + L / 2] = T(0.25) * nyq[q];
    }
    if (FUSE) {
        T* lo = p.lin_partial + ((size_t)pair * p.nparts + part) * 4 * TPS;
#pragma unroll
        for (int q = 0; q < 4; ++q) lo[q * TPS + t] = T(0.5) * lin[q];
        double sx = (double)devx, sy = (double)devy;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            sx += __shfl_xor_sync(0xffffffffu, sx, o);
            sy += __shfl_xor_sync(0xffffffffu, sy, o);
        }
        if ((threadIdx.x & 31) == 0) {
            if (pair == 0) atomicAdd(&p.dev_sums[0], sx);
            atomicAdd(&p.dev_sums[1 + pair], sy);
        }
    }
}

// stats[2c] += dev_sums[c] + count * est[2c]   (raw sum of the samples the fused kernel covered)
__global__ void welch_raw_sums_kernel(const double* __restrict__ dev_sums, const double* __restrict__ est,
                                      int n, double count, double* __restrict__ stats) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < n) stats[2 * c] += dev_sums[c] + count * est[2 * c];
}

// Moves sums accumulated about a provisional mean to the exact mean m = est + delta:
// U = U~ - dx W, V = V~ - dy W with W = FFT(window)  (bins f < nlin; negligible elsewhere,
// checked by the caller against the window's leakage)
__global__ void __launch_bounds__(256) welch_mean_correct_kernel(double* __restrict__ acc, const double* __restrict__ lin,
                                                                 int nlin, int nbin, int nch, double K,
                                                                 const double* __restrict__ est,
                                                                 const double* __restrict__ stats, double nana,
                                                                 const double* __restrict__ W) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= nch * nlin) return;
    const int c = idx / nlin, f = idx - c * nlin;
    const double dx = stats[0] / nana - est[0];
    const double dy = stats[2 * (1 + c)] / nana - est[2 * (1 + c)];
    const double wr = W[2 * f], wi = W[2 * f + 1], w2 = wr * wr + wi * wi;
    const double* l = lin + (size_t)c * 4 * nlin;
    const double sur = l[f], sui = l[nlin + f], svr = l[2 * nlin + f], svi = l[3 * nlin + f];
    double* a = acc + (size_t)c * 4 * nbin;
    a[f] += -2.0 * dx * (wr * sur + wi * sui) + K * dx * dx * w2;
    a[nbin + f] += -2.0 * dy * (wr * svr + wi * svi) + K * dy * dy * w2;
    a[2 * nbin + f] += -dx * (wr * svr + wi * svi) - dy * (wr * sur + wi * sui) + K * dx * dy * w2;
    a[3 * nbin + f] += -dx * (wr * svi - wi * svr) - dy * (wi * sur - wr * sui);
}

// Moves the sums of ONE shard from its provisional means est to the common reference 0 (the
// spectra of the raw windowed samples, bins f < nlin), lin included, so that shards analysed
// about different provisional means can be added (one all-reduce of acc | lin | stats) and
// corrected once with est = 0:  U0 = U~ + ex W,  V0 = V~ + ey W.
__global__ void __launch_bounds__(256) welch_rereference_kernel(double* __restrict__ acc, double* __restrict__ lin,
                                                                int nlin, int nbin, int nch, double K,
                                                                const double* __restrict__ est,
                                                                const double* __restrict__ W) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= nch * nlin) return;
    const int c = idx / nlin, f = idx - c * nlin;
    const double dx = -est[0], dy = -est[2 * (1 + c)];
    const double wr = W[2 * f], wi = W[2 * f + 1], w2 = wr * wr + wi * wi;
    double* l = lin + (size_t)c * 4 * nlin;
    const double sur = l[f], sui = l[nlin + f], svr = l[2 * nlin + f], svi = l[3 * nlin + f];
    double* a = acc + (size_t)c * 4 * nbin;
    a[f] += -2.0 * dx * (wr * sur + wi * sui) + K * dx * dx * w2;
    a[nbin + f] += -2.0 * dy * (wr * svr + wi * svi) + K * dy * dy * w2;
    a[2 * nbin + f] += -dx * (wr * svr + wi * svi) - dy * (wr * sur + wi * sui) + K * dx * dy * w2;
    a[3 * nbin + f] += -dx * (wr * svi - wi * svr) - dy * (wi * sur - wr * sui);
    l[f] = sur - K * dx * wr;
    l[nlin + f] = sui - K * dx * wi;
    l[2 * nlin + f] = svr - K * dy * wr;
    l[3 * nlin + f] = svi - K * dy * wi;
}

// acc[pair][q][f] += sum_part partial[pair][part][q][f]
// 32 consecutive outputs per block, the parts split over the 8 warps (8 loads in flight each),
// combined through shared memory: the part loop is short enough not to be latency-bound
template <typename T>
__global__ void __launch_bounds__(256) welch_reduce_kernel(const T* __restrict__ partial, int nparts,
                                                           int per_pair, double* __restrict__ acc) {
    const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
    const int i = blockIdx.x * 32 + lane;
    const int pair = blockIdx.y;
    __shared__ double sh[8][33];
    double s = 0.0;
    if (i < per_pair) {
        const T* p = partial + (size_t)pair * nparts * per_pair + i;
        int k = g;
        for (; k + 56 < nparts; k += 64) {
            T r[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) r[u] = p[(size_t)(k + 8 * u) * per_pair];
#pragma unroll
            for (int u = 0; u < 8; ++u) s += (double)r[u];
        }
        for (; k < nparts; k += 8) s += (double)p[(size_t)k * per_pair];
    }
    sh[g][lane] = s;
    __syncthreads();
    if (g == 0 && i < per_pair) {
        double t = 0.0;
#pragma unroll
        for (int u = 0; u < 8; ++u) t += sh[u][lane];
        acc[(size_t)pair * per_pair + i] += t;
    }
}

// ------------------------------------------------------------------------- //
// finalize
// ------------------------------------------------------------------------- //
// mean correction of welch_mean_correct_kernel applied while reading acc (lin == NULL: none)
struct MeanFix {
    const double* lin;
    const double* est;
    const double* stats;
    const double* W;
    int nlin;
    double K, nana;
};

__global__ void __launch_bounds__(256) welch_finalize_kernel(const double* __restrict__ acc, MeanFix fix, int nfft, int nch,
                                                             double scale_over_k, int onesided, int F,
                                                             double* __restrict__ Pxx, double* __restrict__ Pyy,
                                                             double* __restrict__ Pxy, double* __restrict__ Cxy,
                                                             double* __restrict__ Cxy2, double* __restrict__ phi) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= F * nch) return;
    const int o = idx / nch, c = idx - o * nch;     // output bin, channel
    const int nbin = nfft / 2 + 1;
    int f;                 // source bin in [0, nfft/2]
    bool conj = false;
    double wgt = 1.0;
    if (onesided) {
        f = o;
        // x2 on bins 1..F-2 only (fft_analysis.py:414-416); nfft is even here so the last bin stays x1
        if (o >= 1 && o <= F - 2) wgt = 2.0;
    } else {
        // fftshift order: output o <-> frequency index (o + nfft/2) mod nfft  (fft_analysis.py:423-427)
        const int k = (o + nfft - nfft / 2) % nfft;     // numpy fftshift: shift by nfft//2
        if (k <= nfft / 2) f = k; else { f = nfft - k; conj = true; }
    }
    const double* a = acc + (size_t)c * 4 * nbin;
    const double s = scale_over_k * wgt;
    double a0 = a[0 * nbin + f], a1 = a[1 * nbin + f], a2 = a[2 * nbin + f], a3 = a[3 * nbin + f];
    if (fix.lin && f < fix.nlin) {
        const int nlin = fix.nlin;
        const double dx = fix.stats[0] / fix.nana - fix.est[0];
        const double dy = fix.stats[2 * (1 + c)] / fix.nana - fix.est[2 * (1 + c)];
        const double wr = fix.W[2 * f], wi = fix.W[2 * f + 1], w2 = wr * wr + wi * wi;
        const double* l = fix.lin + (size_t)c * 4 * nlin;
        const double sur = l[f], sui = l[nlin + f], svr = l[2 * nlin + f], svi = l[3 * nlin + f];
        a0 += -2.0 * dx * (wr * sur + wi * sui) + fix.K * dx * dx * w2;
        a1 += -2.0 * dy * (wr * svr + wi * svi) + fix.K * dy * dy * w2;
        a2 += -dx * (wr * svr + wi * svi) - dy * (wr * sur + wi * sui) + fix.K * dx * dy * w2;
        a3 += -dx * (wr * svi - wi * svr) - dy * (wi * sur - wr * sui);
    }
    const double pxx = a0 * s, pyy = a1 * s;
    const double pr = a2 * s, pi = (conj ? -a3 : a3) * s;
    if (Pxx && c == 0) { Pxx[2 * o] = pxx; Pxx[2 * o + 1] = 0.0; }
    if (Pyy) { Pyy[2 * idx] = pyy; Pyy[2 * idx + 1] = 0.0; }
    if (Pxy) { Pxy[2 * idx] = pr; Pxy[2 * idx + 1] = pi; }
    const double den = fabs(pxx) * fabs(pyy);                 // fft_analysis.py:1672, 1677
    if (Cxy) { const double r = 1.0 / sqrt(den); Cxy[2 * idx] = pr * r; Cxy[2 * idx + 1] = pi * r; }
    if (Cxy2) { Cxy2[2 * idx] = (pr * pr + pi * pi) / den; Cxy2[2 * idx + 1] = 0.0 / den; }
    if (phi) phi[idx] = atan2(pi, pr);                        // fft_analysis.py:520
}

// ------------------------------------------------------------------------- //
// K3a: correlation epilogue -- one CTA per column, double precision
// ------------------------------------------------------------------------- //
template <int LOG2L>
__global__ void __launch_bounds__(WelchGeom<LOG2L>::TPS)
welch_correlation_kernel(const double* __restrict__ spec, int ncol, int onesided, int halve,
                         const cpx<double>* __restrict__ tw, double* __restrict__ out) {
    // launched with exactly TPS = L/16 threads: every thread carries 16 points
    constexpr int L = 1 << LOG2L, TPS = L / 16;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cpx<double>* sm = reinterpret_cast<cpx<double>*>(smem_raw);
    const int col = blockIdx.x;
    const int t = threadIdx.x;
    const int F = onesided ? L / 2 : L;
    Exchange<double, true> ex;
    ex.cur = ex.other = (uint32_t)__cvta_generic_to_shared(sm);
    const Twiddles<double> tws = load_twiddles<double, LOG2L, true>(tw, t);
    cpx<double> v[16];
#pragma unroll
    for (int s = 0; s < 16; ++s) {
        const int k = t + TPS * s;                 // frequency index 0..L-1
        cpx<double> z = mk<double>(0.0, 0.0);
        if (onesided) {
            // numpy irfft(q, n=L): q has L/2 bins, the Nyquist bin is absent (= 0) and the
            // imaginary part of q[0] is ignored; negative frequencies by Hermitian symmetry
            const int f = k <= L / 2 ? k : L - k;
            if (f < F) {
                z.x = spec[2 * ((size_t)f * ncol + col)];
                z.y = spec[2 * ((size_t)f * ncol + col) + 1];
                if (k > L / 2) z.y = -z.y;
                if (f == 0) z.y = 0.0;
                if (halve && f >= 1 && f <= F - 2) { z.x *= 0.5; z.y *= 0.5; }
            }
        } else {
            // ifft(ifftshift(P)): frequency index k sits at input position (k + L/2) mod L
            const int o = (k + L / 2) % L;
            z.x = spec[2 * ((size_t)o * ncol + col)];
            z.y = spec[2 * ((size_t)o * ncol + col) + 1];
        }
        v[s] = z;
    }
    fft_block<double, LOG2L, true, true>(v, ex, tws, t);
    const double sc = sqrt((double)L) / (double)L;     // sqrt(nfft) * (1/nfft of the inverse)
#pragma unroll
    for (int s = 0; s < 16; ++s) {
        const int n = t + TPS * s;                     // lag index before fftshift
        const int o = (n + L / 2) % L;                 // fftshift (fft_analysis.py:592-595)
        if (onesided) {
            out[(size_t)o * ncol + col] = v[s].x * sc;
        } else {
            out[2 * ((size_t)o * ncol + col)] = v[s].x * sc;
            out[2 * ((size_t)o * ncol + col) + 1] = v[s].y * sc;
        }
    }
}

// ------------------------------------------------------------------------- //
// host side
// ------------------------------------------------------------------------- //
static int welch_grid_x(int64_t seg_count, int slots, int nch) {
    int sms = device_sm_count();
    if (sms <= 0) sms = 148;
    int64_t want = ((int64_t)2 * sms + nch - 1) / nch;            // ~2 CTAs per SM over all pairs
    int64_t maxb = (seg_count + slots - 1) / slots;
    if (want > maxb) want = maxb;
    if (want < 1) want = 1;
    return (int)want;
}

template <typename T, int LOG2L>
static int launch_welch(const void* x, const void* y, int64_t y_stride, int nch, const RecMap& map,
                        const pyfft_segmentation* seg, const void* window, const double* coef_x,
                        const double* coef_y, double* acc, double* lin, double* stats, void* workspace,
                        int64_t workspace_bytes, cudaStream_t st) {
    using G = WelchGeom<LOG2L>;
    constexpr bool SINGLE = welch_smem_bytes<T, LOG2L, false>() > 200 * 1024;
    constexpr size_t smem = welch_smem_bytes<T, LOG2L, SINGLE>();
    const bool fuse = lin != nullptr;
    int status;
    const cpx<T>* tw = reinterpret_cast<const cpx<T>*>(
        twiddle_table(sizeof(T) == 8 ? PYFFT_F64 : PYFFT_F32, G::L, &status));
    if (!tw) return status;
    const int gx = welch_grid_x(seg->seg_count, G::SLOTS, nch);
    const int nparts = gx * G::SLOTS;
    const int64_t main_bytes = (int64_t)nch * nparts * 4 * G::NBIN * (int64_t)sizeof(T);
    const int64_t lin_bytes = (int64_t)nch * nparts * 4 * G::TPS * (int64_t)sizeof(T);
    const int64_t need = main_bytes + lin_bytes + 8 * (int64_t)(nch + 1);
    if (workspace_bytes < need || !workspace)
        return fail(PYFFT_EINVAL, "workspace too small: %lld < %lld bytes", (long long)workspace_bytes,
                    (long long)need);
    WelchParams<T> p;
    p.x = static_cast<const T*>(x);
    p.y = static_cast<const T*>(y);
    p.y_stride = y_stride;
    p.map = map;
    p.hop = seg->hop;
    p.seg_first = seg->seg_first;
    p.seg_count = seg->seg_count;
    p.chunk = (seg->seg_count + nparts - 1) / nparts;
    p.window = static_cast<const T*>(window);
    p.coef_x = coef_x;
    p.coef_y = coef_y;
    p.tw = tw;
    p.partial = static_cast<T*>(workspace);
    p.nparts = nparts;
    p.lin_partial = reinterpret_cast<T*>(static_cast<char*>(workspace) + main_bytes);
    p.dev_sums = reinterpret_cast<double*>(static_cast<char*>(workspace) + main_bytes + lin_bytes);
    const int per_pair = 4 * G::NBIN;
    if (fuse) {
        PYFFT_CUDA_OK(cudaMemsetAsync(p.dev_sums, 0, sizeof(double) * (size_t)(nch + 1), st));
        auto kern = welch_pair_kernel<T, LOG2L, SINGLE, true>;
        PYFFT_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<dim3(gx, nch), G::THREADS, smem, st>>>(p);
        PYFFT_LAUNCH_OK("welch_pair_kernel");
        welch_reduce_kernel<T><<<dim3((4 * G::TPS + 31) / 32, nch), 256, 0, st>>>(p.lin_partial, nparts, 4 * G::TPS, lin);
        // coef_x and coef_y are one contiguous (a, b) array: x first, then the y channels
        welch_raw_sums_kernel<<<(nch + 1 + 127) / 128, 128, 0, st>>>(p.dev_sums, coef_x, nch + 1,
                                                                     (double)seg->seg_count * (double)seg->hop, stats);
    } else {
        auto kern = welch_pair_kernel<T, LOG2L, SINGLE, false>;
        PYFFT_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<dim3(gx, nch), G::THREADS, smem, st>>>(p);
        PYFFT_LAUNCH_OK("welch_pair_kernel");
    }
    welch_reduce_kernel<T><<<dim3((per_pair + 31) / 32, nch), 256, 0, st>>>(p.partial, nparts, per_pair, acc);
    PYFFT_LAUNCH_OK("welch_reduce_kernel");
    return PYFFT_OK;
}

template <typename T>
static int dispatch_welch(int log2l, const void* x, const void* y, int64_t y_stride, int nch,
                          const RecMap& map, const pyfft_segmentation* seg, const void* window,
                          const double* coef_x, const double* coef_y, double* acc, double* lin, double* stats,
                          void* ws, int64_t wsb, cudaStream_t st) {
#define PYFFT_CASE(N) \
    case N: return launch_welch<T, N>(x, y, y_stride, nch, map, seg, window, coef_x, coef_y, acc, lin, stats, ws, wsb, st);
    switch (log2l) {
        PYFFT_CASE(4) PYFFT_CASE(5) PYFFT_CASE(6) PYFFT_CASE(7) PYFFT_CASE(8)
        PYFFT_CASE(9) PYFFT_CASE(10) PYFFT_CASE(11) PYFFT_CASE(12) PYFFT_CASE(13)
    }
#undef PYFFT_CASE
    return fail(PYFFT_EUNSUPPORTED, "nfft = 2^%d not supported", log2l);
}

// min and mean over the bins of the averaged auto-spectra (one block per channel: 0 = x from Pxx, 1.. = y
// from Pyy[F][nch]); one-sided spectra: the un-doubled end bins are counted twice so that all bins are alike
__global__ void __launch_bounds__(256) welch_dynamic_range_kernel(const double* __restrict__ Pxx, const double* __restrict__ Pyy,
                                                                  int F, int nch, int onesided, double* __restrict__ out) {
    const int c = blockIdx.x;
    double mn = 1e300, sum = 0.0;
    for (int f = threadIdx.x; f < F; f += 256) {
        double v = c == 0 ? Pxx[2 * f] : Pyy[2 * ((size_t)f * nch + (c - 1))];
        if (onesided && (f == 0 || f == F - 1)) v *= 2.0;
        if (!(v >= 0.0)) v = 0.0;                     // NaN counts as an empty bin
        mn = v < mn ? v : mn;
        sum += v;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double a = __shfl_xor_sync(0xffffffffu, mn, o);
        mn = a < mn ? a : mn;
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
    }
    __shared__ double sh[2][8];
    if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = mn; sh[1][threadIdx.x >> 5] = sum; }
    __syncthreads();
    if (threadIdx.x == 0) {
        mn = sh[0][0]; sum = sh[1][0];
        for (int w = 1; w < 8; ++w) { mn = sh[0][w] < mn ? sh[0][w] : mn; sum += sh[1][w]; }
        out[2 * c] = mn;
        out[2 * c + 1] = sum / (double)F;
    }
}

int welch_reduce_launch_f32(const float* partial, int nparts, int per_pair, int nch, double* acc, cudaStream_t st) {
    welch_reduce_kernel<float><<<dim3((per_pair + 31) / 32, nch), 256, 0, st>>>(partial, nparts, per_pair, acc);
    PYFFT_LAUNCH_OK("welch_reduce_kernel");
    return PYFFT_OK;
}

int welch_raw_sums_launch(const double* dev_sums, const double* est, int n, double count, double* stats,
                          cudaStream_t st) {
    welch_raw_sums_kernel<<<(n + 127) / 128, 128, 0, st>>>(dev_sums, est, n, count, stats);
    PYFFT_LAUNCH_OK("welch_raw_sums_kernel");
    return PYFFT_OK;
}

// ------------------------------------------------------------------------- //
// allocation-free variant of pyfft_welch_accumulate (pyfft_welch_accumulate_static): the per-CTA partial
// sums and the twiddle table live in ONE file-scope device array instead of a caller workspace and the
// cudaMalloc'ed twiddle cache.  Every GPU that loads the library holds the array until exit.  All calls
// share it: calls that overlap on the GPU (different streams) race.
// ------------------------------------------------------------------------- //
constexpr int64_t STATIC_TW_BYTES = 8192 * 16;                       // twiddles of one call (nfft <= 8192, complex double)
constexpr int64_t STATIC_WS_BYTES = 16ll << 20;                      // partial sums: 16 MiB
static __device__ __align__(256) unsigned char g_static_scratch[STATIC_TW_BYTES + STATIC_WS_BYTES];
static std::mutex g_static_mutex;
static std::map<int, void*> g_static_addr;                           // device ordinal -> device address of the array

// the array's NAME is a host-side stand-in; its device address comes from cudaGetSymbolAddress, once per
// device, with that device current
static void* static_scratch_address(int* status) {
    *status = PYFFT_OK;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { *status = fail(PYFFT_ECUDA, "cudaGetDevice failed"); return nullptr; }
    std::lock_guard<std::mutex> lock(g_static_mutex);
    auto it = g_static_addr.find(dev);
    if (it != g_static_addr.end()) return it->second;
    void* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_scratch);
    if (e != cudaSuccess || !ptr) {
        *status = fail(PYFFT_ECUDA, "cudaGetSymbolAddress(static scratch) failed: %s", cudaGetErrorString(e));
        return nullptr;
    }
    g_static_addr[dev] = ptr;
    return ptr;
}

// tw[k] = exp(-2 pi i k / L), evaluated on the device in double (sincospi) and rounded to T
template <typename T>
__global__ void static_twiddle_kernel(cpx<T>* __restrict__ tw, int L) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= L) return;
    double s, c;
    sincospi(-2.0 * (double)k / (double)L, &s, &c);
    tw[k].x = (T)c;
    tw[k].y = (T)s;
}

template <typename T, int LOG2L>
static int launch_welch_static(const void* x, const void* y, int64_t y_stride, int nch, const RecMap& map,
                               const pyfft_segmentation* seg, const void* window, const double* coef_x,
                               const double* coef_y, double* acc, unsigned char* scratch, cudaStream_t st) {
    using G = WelchGeom<LOG2L>;
    constexpr bool SINGLE = welch_smem_bytes<T, LOG2L, false>() > 200 * 1024;
    constexpr size_t smem = welch_smem_bytes<T, LOG2L, SINGLE>();
    const int gx = welch_grid_x(seg->seg_count, G::SLOTS, nch);
    const int nparts = gx * G::SLOTS;
    const int64_t main_bytes = (int64_t)nch * nparts * 4 * G::NBIN * (int64_t)sizeof(T);
    const int64_t lin_bytes = (int64_t)nch * nparts * 4 * G::TPS * (int64_t)sizeof(T);
    const int64_t need = main_bytes + lin_bytes + 8 * (int64_t)(nch + 1);
    if (need > STATIC_WS_BYTES)
        return fail(PYFFT_EUNSUPPORTED,
                    "welch_accumulate_static: this call needs %lld bytes of scratch, the static array holds %lld "
                    "(use pyfft_welch_accumulate with a caller workspace)", (long long)need, (long long)STATIC_WS_BYTES);
    cpx<T>* tw = reinterpret_cast<cpx<T>*>(scratch);
    unsigned char* ws = scratch + STATIC_TW_BYTES;
    static_twiddle_kernel<T><<<(G::L + 255) / 256, 256, 0, st>>>(tw, G::L);
    PYFFT_LAUNCH_OK("static_twiddle_kernel");
    WelchParams<T> p;
    p.x = static_cast<const T*>(x);
    p.y = static_cast<const T*>(y);
    p.y_stride = y_stride;
    p.map = map;
    p.hop = seg->hop;
    p.seg_first = seg->seg_first;
    p.seg_count = seg->seg_count;
    p.chunk = (seg->seg_count + nparts - 1) / nparts;
    p.window = static_cast<const T*>(window);
    p.coef_x = coef_x;
    p.coef_y = coef_y;
    p.tw = tw;
    p.partial = reinterpret_cast<T*>(ws);                  // every CTA writes all of its partial sums: nothing to clear
    p.nparts = nparts;
    p.lin_partial = reinterpret_cast<T*>(ws + main_bytes);
    p.dev_sums = reinterpret_cast<double*>(ws + main_bytes + lin_bytes);
    auto kern = welch_pair_kernel<T, LOG2L, SINGLE, false>;
    PYFFT_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<dim3(gx, nch), G::THREADS, smem, st>>>(p);
    PYFFT_LAUNCH_OK("welch_pair_kernel");
    const int per_pair = 4 * G::NBIN;
    welch_reduce_kernel<T><<<dim3((per_pair + 31) / 32, nch), 256, 0, st>>>(p.partial, nparts, per_pair, acc);
    PYFFT_LAUNCH_OK("welch_reduce_kernel");
    return PYFFT_OK;
}

template <typename T>
static int dispatch_welch_static(int log2l, const void* x, const void* y, int64_t y_stride, int nch, const RecMap& map,
                                 const pyfft_segmentation* seg, const void* window, const double* coef_x,
                                 const double* coef_y, double* acc, unsigned char* scratch, cudaStream_t st) {
#define PYFFT_CASE(N) \
    case N: return launch_welch_static<T, N>(x, y, y_stride, nch, map, seg, window, coef_x, coef_y, acc, scratch, st);
    switch (log2l) {
        PYFFT_CASE(4) PYFFT_CASE(5) PYFFT_CASE(6) PYFFT_CASE(7) PYFFT_CASE(8)
        PYFFT_CASE(9) PYFFT_CASE(10) PYFFT_CASE(11) PYFFT_CASE(12) PYFFT_CASE(13)
    }
#undef PYFFT_CASE
    return fail(PYFFT_EUNSUPPORTED, "nfft = 2^%d not supported", log2l);
}

}  // namespace pyfft

using namespace pyfft;

extern "C" {

static int detrend_stats_launch(int dtype, const void* first, const void* ptr, int64_t ch_stride, int32_t nch,
                                const pyfft_record_map* map, int64_t j_begin, int64_t j_end, double* stats,
                                void* stream);

int pyfft_detrend_stats(int dtype, const void* ptr, int64_t ch_stride, int32_t nch,
                        const pyfft_record_map* map, int64_t j_begin, int64_t j_end,
                        double* stats, void* stream) {
    if (!ptr || !stats || nch <= 0) return fail(PYFFT_EINVAL, "detrend_stats: NULL pointer or nch <= 0");
    return detrend_stats_launch(dtype, nullptr, ptr, ch_stride, nch, map, j_begin, j_end, stats, stream);
}

int pyfft_detrend_stats_xy(int dtype, const void* x, const void* y, int64_t y_stride, int32_t nch,
                           const pyfft_record_map* map, int64_t j_begin, int64_t j_end,
                           double* stats, void* stream) {
    if (!x || !y || !stats || nch <= 0) return fail(PYFFT_EINVAL, "detrend_stats_xy: NULL pointer or nch <= 0");
    return detrend_stats_launch(dtype, x, y, y_stride, nch + 1, map, j_begin, j_end, stats, stream);
}

static int detrend_stats_launch(int dtype, const void* first, const void* ptr, int64_t ch_stride, int32_t nch,
                                const pyfft_record_map* map, int64_t j_begin, int64_t j_end, double* stats,
                                void* stream) {
    RecMap m;
    int rc = check_map(map, &m);
    if (rc) return rc;
    if ((rc = check_range(m, j_begin, j_end))) return rc;
    if (j_begin == j_end) return PYFFT_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int sms = device_sm_count();
    if (sms <= 0) sms = 148;
    int64_t blocks = (j_end - j_begin + 256 * 8 - 1) / (256 * 8);
    int64_t cap = ((int64_t)sms * 8 + nch - 1) / nch;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    dim3 grid((unsigned)blocks, (unsigned)nch);
    if (dtype == PYFFT_F32)
        detrend_stats_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(first), static_cast<const float*>(ptr),
                                                          ch_stride, m, j_begin, j_end, stats);
    else if (dtype == PYFFT_F64)
        detrend_stats_kernel<double><<<grid, 256, 0, st>>>(static_cast<const double*>(first), static_cast<const double*>(ptr),
                                                           ch_stride, m, j_begin, j_end, stats);
    else
        return fail(PYFFT_EINVAL, "unknown dtype %d", dtype);
    PYFFT_LAUNCH_OK("detrend_stats_kernel");
    return PYFFT_OK;
}

int pyfft_detrend_coef(int mode, const double* stats, int32_t nch, int64_t nsig, double* coef, void* stream) {
    if (!stats || !coef || nch <= 0 || nsig <= 0) return fail(PYFFT_EINVAL, "detrend_coef: bad arguments");
    if (mode < PYFFT_DETREND_NONE || mode > PYFFT_DETREND_LINEAR)
        return fail(PYFFT_EINVAL, "unknown detrend mode %d", mode);
    detrend_coef_kernel<<<(nch + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(mode, stats, nch, nsig, coef);
    PYFFT_LAUNCH_OK("detrend_coef_kernel");
    return PYFFT_OK;
}

int64_t pyfft_welch_workspace_bytes(int dtype, const pyfft_segmentation* seg, int32_t nch) {
    int log2l;
    int rc = check_seg(seg, 4, 13, &log2l);
    if (rc) return rc;
    if (nch <= 0) return fail(PYFFT_EINVAL, "nch must be positive");
    const int L = 1 << log2l;
    const int tps = L / 16;
    const int threads = tps > 256 ? tps : 256;
    const int slots = threads / tps;
    const int gx = welch_grid_x(seg->seg_count, slots, nch);
    const int64_t esz = dtype == PYFFT_F64 ? 8 : 4;
    int64_t b = (int64_t)nch * gx * slots * 4 * (L / 2 + 1 + tps) * esz + 8 * (int64_t)(nch + 1);
    return b > 0 ? b : 16;
}

static int welch_accumulate_common(int dtype, const void* x, const void* y, int64_t y_stride, int32_t nch,
                                   const pyfft_record_map* map, const pyfft_segmentation* seg,
                                   const void* window, const double* coef_x, const double* coef_y,
                                   double* acc, double* lin, double* stats, void* workspace,
                                   int64_t workspace_bytes, void* stream) {
    RecMap m;
    int rc = check_map(map, &m);
    if (rc) return rc;
    int log2l;
    if ((rc = check_seg(seg, 4, 13, &log2l))) return rc;
    if (!x || !y || !window || !acc || nch <= 0) return fail(PYFFT_EINVAL, "welch_accumulate: NULL pointer or nch <= 0");
    if (seg->seg_count == 0) return PYFFT_OK;
    const int64_t j_lo = seg->seg_first * (int64_t)seg->hop;
    const int64_t j_hi = (seg->seg_first + seg->seg_count - 1) * (int64_t)seg->hop + seg->nfft;
    if ((rc = check_range(m, j_lo, j_hi))) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == PYFFT_F32)
        return dispatch_welch<float>(log2l, x, y, y_stride, nch, m, seg, window, coef_x, coef_y, acc, lin, stats, workspace, workspace_bytes, st);
    if (dtype == PYFFT_F64)
        return dispatch_welch<double>(log2l, x, y, y_stride, nch, m, seg, window, coef_x, coef_y, acc, lin, stats, workspace, workspace_bytes, st);
    return fail(PYFFT_EINVAL, "unknown dtype %d", dtype);
}

int pyfft_welch_accumulate(int dtype, const void* x, const void* y, int64_t y_stride, int32_t nch,
                           const pyfft_record_map* map, const pyfft_segmentation* seg,
                           const void* window, const double* coef_x, const double* coef_y,
                           double* acc, void* workspace, int64_t workspace_bytes, void* stream) {
    return welch_accumulate_common(dtype, x, y, y_stride, nch, map, seg, window, coef_x, coef_y, acc, nullptr,
                                   nullptr, workspace, workspace_bytes, stream);
}

int64_t pyfft_welch_static_scratch_bytes(void) { return STATIC_WS_BYTES; }

int pyfft_welch_accumulate_static(int dtype, const void* x, const void* y, int64_t y_stride, int32_t nch,
                                  const pyfft_record_map* map, const pyfft_segmentation* seg,
                                  const void* window, const double* coef_x, const double* coef_y,
                                  double* acc, void* stream) {
    RecMap m;
    int rc = check_map(map, &m);
    if (rc) return rc;
    int log2l;
    if ((rc = check_seg(seg, 4, 13, &log2l))) return rc;
    if (!x || !y || !window || !acc || nch <= 0)
        return fail(PYFFT_EINVAL, "welch_accumulate_static: NULL pointer or nch <= 0");
    if (seg->seg_count == 0) return PYFFT_OK;
    const int64_t j_lo = seg->seg_first * (int64_t)seg->hop;
    const int64_t j_hi = (seg->seg_first + seg->seg_count - 1) * (int64_t)seg->hop + seg->nfft;
    if ((rc = check_range(m, j_lo, j_hi))) return rc;
    int status;
    unsigned char* scratch = static_cast<unsigned char*>(static_scratch_address(&status));
    if (!scratch) return status;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == PYFFT_F32)
        return dispatch_welch_static<float>(log2l, x, y, y_stride, nch, m, seg, window, coef_x, coef_y, acc, scratch, st);
    if (dtype == PYFFT_F64)
        return dispatch_welch_static<double>(log2l, x, y, y_stride, nch, m, seg, window, coef_x, coef_y, acc, scratch, st);
    return fail(PYFFT_EINVAL, "unknown dtype %d", dtype);
}

int pyfft_welch_accumulate_mean(int dtype, const void* x, const void* y, int64_t y_stride, int32_t nch,
                                const pyfft_record_map* map, const pyfft_segmentation* seg,
                                const void* window, const double* est, double* acc, double* lin,
                                double* stats, void* workspace, int64_t workspace_bytes, void* stream) {
    if (!est || !lin || !stats) return fail(PYFFT_EINVAL, "welch_accumulate_mean: est, lin and stats are required");
    if (seg && seg->hop > seg->nfft) return fail(PYFFT_EINVAL, "welch_accumulate_mean: hop > nfft leaves gaps in the mean");
    return welch_accumulate_common(dtype, x, y, y_stride, nch, map, seg, window, est, est + 2, acc, lin, stats,
                                   workspace, workspace_bytes, stream);
}

int pyfft_welch_mean_correct(double* acc, const double* lin, int32_t nfft, int32_t nch, int64_t K,
                             const double* est, const double* stats, int64_t nana, const double* wlow,
                             void* stream) {
    if (!acc || !lin || !est || !stats || !wlow || nfft < 16 || nch <= 0 || K < 0 || nana <= 0)
        return fail(PYFFT_EINVAL, "welch_mean_correct: bad arguments");
    const int nlin = nfft / 16, nbin = nfft / 2 + 1;
    welch_mean_correct_kernel<<<(nch * nlin + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        acc, lin, nlin, nbin, nch, (double)K, est, stats, (double)nana, wlow);
    PYFFT_LAUNCH_OK("welch_mean_correct_kernel");
    return PYFFT_OK;
}

int pyfft_welch_rereference(double* acc, double* lin, int32_t nfft, int32_t nch, int64_t K,
                            const double* est, const double* wlow, void* stream) {
    if (!acc || !lin || !est || !wlow || nfft < 16 || nch <= 0 || K < 0)
        return fail(PYFFT_EINVAL, "welch_rereference: bad arguments");
    const int nlin = nfft / 16, nbin = nfft / 2 + 1;
    welch_rereference_kernel<<<(nch * nlin + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        acc, lin, nlin, nbin, nch, (double)K, est, wlow);
    PYFFT_LAUNCH_OK("welch_rereference_kernel");
    return PYFFT_OK;
}

int pyfft_welch_finalize(const double* acc, int32_t nfft, int32_t nch, int64_t K, double scale,
                         int32_t onesided, double* Pxx, double* Pyy, double* Pxy, double* Cxy,
                         double* Cxy2, double* phi, void* stream) {
    if (!acc || nfft < 2 || (nfft & 1) || nch <= 0 || K <= 0)
        return fail(PYFFT_EINVAL, "welch_finalize: bad arguments (even nfft, nch > 0, K > 0 required)");
    const int F = onesided ? nfft / 2 : nfft;
    const int total = F * nch;
    MeanFix fix{};
    welch_finalize_kernel<<<(total + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        acc, fix, nfft, nch, scale / (double)K, onesided, F, Pxx, Pyy, Pxy, Cxy, Cxy2, phi);
    PYFFT_LAUNCH_OK("welch_finalize_kernel");
    return PYFFT_OK;
}

int pyfft_welch_finalize_mean(const double* acc, const double* lin, int32_t nfft, int32_t nch, int64_t K,
                              const double* est, const double* stats, int64_t nana, const double* wlow,
                              double scale, int32_t onesided, double* Pxx, double* Pyy, double* Pxy,
                              double* Cxy, double* Cxy2, double* phi, void* stream) {
    if (!acc || !lin || !est || !stats || !wlow || nfft < 16 || (nfft & 1) || nch <= 0 || K <= 0 || nana <= 0)
        return fail(PYFFT_EINVAL, "welch_finalize_mean: bad arguments");
    const int F = onesided ? nfft / 2 : nfft;
    const int total = F * nch;
    MeanFix fix{lin, est, stats, wlow, nfft / 16, (double)K, (double)nana};
    welch_finalize_kernel<<<(total + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        acc, fix, nfft, nch, scale / (double)K, onesided, F, Pxx, Pyy, Pxy, Cxy, Cxy2, phi);
    PYFFT_LAUNCH_OK("welch_finalize_kernel");
    return PYFFT_OK;
}

int pyfft_welch_dynamic_range(const double* Pxx, const double* Pyy, int32_t nfreq, int32_t nch, int32_t onesided,
                              double* out, void* stream) {
    if (!Pxx || !Pyy || !out || nfreq <= 0 || nch <= 0) return fail(PYFFT_EINVAL, "welch_dynamic_range: bad arguments");
    welch_dynamic_range_kernel<<<nch + 1, 256, 0, static_cast<cudaStream_t>(stream)>>>(Pxx, Pyy, nfreq, nch, onesided, out);
    PYFFT_LAUNCH_OK("welch_dynamic_range_kernel");
    return PYFFT_OK;
}

int pyfft_welch_correlation(const double* spec, int32_t nfft, int32_t ncol, int32_t onesided,
                            int32_t halve_interior, double* out, void* stream) {
    if (!spec || !out || ncol <= 0) return fail(PYFFT_EINVAL, "welch_correlation: bad arguments");
    pyfft_segmentation seg = {nfft, 1, 0, 1};
    int log2l;
    int rc = check_seg(&seg, 4, 13, &log2l);
    if (rc) return rc;
    int status;
    const cpx<double>* tw = reinterpret_cast<const cpx<double>*>(twiddle_table(PYFFT_F64, nfft, &status));
    if (!tw) return status;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t smem = (size_t)smem_elems(nfft) * sizeof(cpx<double>);
#define PYFFT_CASE(N)                                                                                   \
    case N: {                                                                                           \
        auto kern = welch_correlation_kernel<N>;                                                        \
        PYFFT_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        kern<<<ncol, (1 << N) / 16, smem, st>>>(spec, ncol, onesided, halve_interior, tw, out);         \
        break;                                                                                          \
    }
    switch (log2l) {
        PYFFT_CASE(4) PYFFT_CASE(5) PYFFT_CASE(6) PYFFT_CASE(7) PYFFT_CASE(8)
        PYFFT_CASE(9) PYFFT_CASE(10) PYFFT_CASE(11) PYFFT_CASE(12) PYFFT_CASE(13)
    }
#undef PYFFT_CASE
    PYFFT_LAUNCH_OK("welch_correlation_kernel");
    return PYFFT_OK;
}

}  // extern "C"
