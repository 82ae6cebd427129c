// per-segment statistics of the Welch path as fused reductions.
//
// The reference keeps every segment's spectra (Pxx_seg, Pyy_seg, Pxy_seg, phixy_seg: fft_analysis.py:397-460,
// :1955-1970) and leaves the consumers to reduce them: the "real" segment variances are commented out and replaced
// by the stub (P / sqrt(Navr))^2 (:449-458, :503-505) because materialising K x nfft spectra per channel is what
// costs; mean_angle (:1334-1377) averages phixy_seg over the segments.  Here the same statistics come out of ONE pass
// over the records, the K1 structure (two real channels per complex FFT) with twelve running sums per bin and
// channel pair instead of four; nothing per-segment is written:
//   m0..m3   sum Pxx, Pyy, Re Pxy, Im Pxy            (Pxy = Y conj(X), :393)
//   m4..m7   sum Pxx^2, Pyy^2, (Re Pxy)^2, (Im Pxy)^2  -> np.var over the segments
//   m8..m11  sum cos phi, sin phi, cos^2 phi, sin^2 phi with phi = angle(Pxy_seg)  -> mean_angle
// of the UNSCALED products of the windowed, detrended segments (the host applies 1/S1^2/ENBW and the one-sided
// doubling, like pyfft_welch_finalize).
#include "common.cuh"

namespace pyfft {

constexpr int SEG_NMOM = 12;

template <int LOG2L> struct SegGeom {
    static constexpr int L = 1 << LOG2L;
    static constexpr int TPS = L / 16;
    static constexpr int THREADS = TPS > 128 ? TPS : 128;
    static constexpr int SLOTS = THREADS / TPS;
    static constexpr int NBIN = L / 2 + 1;
    static constexpr int E = smem_elems(L);
};

template <typename T> struct SegParams {
    const T* x;
    const T* y;
    int64_t y_stride;
    RecMap map;
    int64_t hop, seg_first, seg_count, chunk;
    const T* window;
    const double* coef_x;
    const double* coef_y;
    const cpx<T>* tw;
    T* partial;     // [nch][nparts][NM][NBIN]
    int nparts;
};

template <typename T, int LOG2L, bool SINGLE>
constexpr size_t seg_smem_bytes() {
    return (size_t)(SINGLE ? 1 : 2) * SegGeom<LOG2L>::SLOTS * SegGeom<LOG2L>::E * sizeof(cpx<T>) +
           (size_t)SegGeom<LOG2L>::L * sizeof(T);
}

// the sums of one bin of one segment: u = 2 X, v = 2 Y (the factors 1/2 of the split are applied here)
template <typename T, int M0, int NM>
__device__ __forceinline__ void seg_accumulate(T (&a)[NM], T ur, T ui, T vr, T vi) {
    const T pxx = T(0.25) * (ur * ur + ui * ui), pyy = T(0.25) * (vr * vr + vi * vi);
    const T pr = T(0.25) * (vr * ur + vi * ui), pi = T(0.25) * (vi * ur - vr * ui);
    if (M0 == 0) {
        a[0] += pxx;
        a[1] += pyy;
        a[2] += pr;
        a[3] += pi;
        a[4] += pxx * pxx;
        a[5] += pyy * pyy;
        a[6] += pr * pr;
        a[7] += pi * pi;
    }
    if (M0 + NM == SEG_NMOM) {
        // exp(i angle(p)) = p / |p|; angle(0) = 0 (numpy)
        const T m2 = pr * pr + pi * pi;
        T c = T(1), s = T(0);
        if (m2 > T(0)) {
            const T r = T(1) / sqrt(m2);
            c = pr * r;
            s = pi * r;
        }
        T* b = a + (NM - 4);
        b[0] += c;
        b[1] += s;
        b[2] += c * c;
        b[3] += s * s;
    }
}

// moments [M0, M0 + NM) of SEG_NMOM: float keeps all twelve in registers (8 bins x 12 per thread), double takes
// two launches (0..7, 8..11) to stay inside the register file
template <typename T, int LOG2L, bool SINGLE, int M0, int NM>
__global__ void __launch_bounds__(SegGeom<LOG2L>::THREADS) segment_moments_kernel(const SegParams<T> p) {
    using G = SegGeom<LOG2L>;
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

    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(sm);
    Exchange<T, SINGLE> ex;
    ex.cur = sbase + (uint32_t)(slot * E * sizeof(cpx<T>));
    ex.other = SINGLE ? ex.cur : sbase + (uint32_t)((SLOTS + slot) * E * sizeof(cpx<T>));
    const Twiddles<T> tws = load_twiddles<T, LOG2L, false>(p.tw, t);

    T acc[8][NM];
    T nyq[NM];
#pragma unroll
    for (int q = 0; q < NM; ++q) {
        nyq[q] = T(0);
#pragma unroll
        for (int s = 0; s < 8; ++s) acc[s][q] = T(0);
    }

    for (int64_t it = 0; it < p.chunk; ++it) {       // same trip count for every slot (barriers inside)
        const int64_t k = k0 + it;
        const bool active = k < k1;
        cpx<T> v[16];
        if (active) {
            const int64_t j0 = (p.seg_first + k) * p.hop;
            const T cx0 = (T)(ax + bx * (double)j0), cy0 = (T)(ay + by * (double)j0);
#pragma unroll
            for (int s = 0; s < 16; ++s) {
                const int n = t + TPS * s;
                const int64_t e = src_index(p.map, j0 + n);
                const T wv = swin[n];
                v[s] = mk<T>((u[e] - (cx0 + bxf * (T)n)) * wv, (w[e] - (cy0 + byf * (T)n)) * wv);
            }
        } else {
#pragma unroll
            for (int s = 0; s < 16; ++s) v[s] = mk<T>(T(0), T(0));
        }
        fft_block<T, LOG2L, false, SINGLE>(v, ex, tws, t);
        // every thread fetches Z[L - f] for its bins f = t + TPS s, s < 8 (see welch_pair_kernel)
        ex.before_write();
        const uint32_t e = ex.cur;
#pragma unroll
        for (int s = 0; s < 16; ++s) ex.st(e, t + TPS * s, v[s]);
        __syncthreads();
        ex.flip();
        if (active) {
#pragma unroll
            for (int s = 0; s < 8; ++s) {
                const int f = t + TPS * s;
                const cpx<T> A = v[s];
                const cpx<T> B = ex.ld(e, (L - f) & (L - 1));
                // U = (A + conj B)/2, V = (A - conj B)/(2i)
                seg_accumulate<T, M0, NM>(acc[s], A.x + B.x, A.y - B.y, A.y + B.y, B.x - A.x);
            }
            if (t == 0) seg_accumulate<T, M0, NM>(nyq, v[8].x + v[8].x, T(0), v[8].y + v[8].y, T(0));   // f = L/2, self-paired
        }
    }

    T* out = p.partial + ((size_t)pair * p.nparts + part) * NM * NBIN;
#pragma unroll
    for (int s = 0; s < 8; ++s)
#pragma unroll
        for (int q = 0; q < NM; ++q) out[q * NBIN + t + TPS * s] = acc[s][q];
    if (t == 0) {
#pragma unroll
        for (int q = 0; q < NM; ++q) out[q * NBIN + L / 2] = nyq[q];
    }
}

// out[pair][M0 + q][f] = sum over parts of partial[pair][part][q][f]
template <typename T>
__global__ void __launch_bounds__(256) segment_reduce_kernel(const T* __restrict__ partial, int nparts, int nm, int nbin,
                                                             int m0, double* __restrict__ out) {
    const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
    const int i = blockIdx.x * 32 + lane;           // (q, f) flattened
    const int pair = blockIdx.y;
    const int per = nm * nbin;
    __shared__ double sh[8][33];
    double s = 0.0;
    if (i < per) {
        const T* p = partial + (size_t)pair * nparts * per + i;
        for (int k = g; k < nparts; k += 8) s += (double)p[(size_t)k * per];
    }
    sh[g][lane] = s;
    __syncthreads();
    if (g == 0 && i < per) {
        double t = 0.0;
#pragma unroll
        for (int u = 0; u < 8; ++u) t += sh[u][lane];
        const int q = i / nbin, f = i - q * nbin;
        out[((size_t)pair * SEG_NMOM + m0 + q) * nbin + f] = t;
    }
}

static int seg_grid_x(int64_t seg_count, int slots, int nch) {
    int sms = device_sm_count();
    if (sms <= 0) sms = 148;
    int64_t want = ((int64_t)2 * sms + nch - 1) / nch;
    const int64_t maxb = (seg_count + slots - 1) / slots;
    if (want > maxb) want = maxb;
    if (want < 1) want = 1;
    return (int)want;
}

template <typename T, int LOG2L, int M0, int NM>
static int launch_moments(SegParams<T> p, int nch, double* out, void* workspace, int64_t workspace_bytes,
                          cudaStream_t st) {
    using G = SegGeom<LOG2L>;
    constexpr bool SINGLE = seg_smem_bytes<T, LOG2L, false>() > 200 * 1024;
    constexpr size_t smem = seg_smem_bytes<T, LOG2L, SINGLE>();
    const int gx = seg_grid_x(p.seg_count, G::SLOTS, nch);
    const int nparts = gx * G::SLOTS;
    const int64_t need = (int64_t)nch * nparts * NM * G::NBIN * (int64_t)sizeof(T);
    if (!workspace || workspace_bytes < need)
        return fail(PYFFT_EINVAL, "segment_moments: workspace too small: %lld < %lld bytes", (long long)workspace_bytes,
                    (long long)need);
    p.chunk = (p.seg_count + nparts - 1) / nparts;
    p.partial = static_cast<T*>(workspace);
    p.nparts = nparts;
    auto kern = segment_moments_kernel<T, LOG2L, SINGLE, M0, NM>;
    PYFFT_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<dim3(gx, nch), G::THREADS, smem, st>>>(p);
    PYFFT_LAUNCH_OK("segment_moments_kernel");
    segment_reduce_kernel<T><<<dim3((NM * G::NBIN + 31) / 32, nch), 256, 0, st>>>(p.partial, nparts, NM, G::NBIN, M0, out);
    PYFFT_LAUNCH_OK("segment_reduce_kernel");
    return PYFFT_OK;
}

template <int LOG2L>
static int launch_moments_any(int dtype, const void* x, const void* y, int64_t y_stride, int nch, const RecMap& map,
                              const pyfft_segmentation* seg, const void* window, const double* coef_x,
                              const double* coef_y, double* out, void* ws, int64_t wsb, cudaStream_t st) {
    int status;
    if (dtype == PYFFT_F32) {
        SegParams<float> p;
        p.x = static_cast<const float*>(x);
        p.y = static_cast<const float*>(y);
        p.y_stride = y_stride;
        p.map = map;
        p.hop = seg->hop;
        p.seg_first = seg->seg_first;
        p.seg_count = seg->seg_count;
        p.window = static_cast<const float*>(window);
        p.coef_x = coef_x;
        p.coef_y = coef_y;
        p.tw = reinterpret_cast<const cpx<float>*>(twiddle_table(PYFFT_F32, SegGeom<LOG2L>::L, &status));
        if (!p.tw) return status;
        return launch_moments<float, LOG2L, 0, SEG_NMOM>(p, nch, out, ws, wsb, st);
    }
    SegParams<double> p;
    p.x = static_cast<const double*>(x);
    p.y = static_cast<const double*>(y);
    p.y_stride = y_stride;
    p.map = map;
    p.hop = seg->hop;
    p.seg_first = seg->seg_first;
    p.seg_count = seg->seg_count;
    p.window = static_cast<const double*>(window);
    p.coef_x = coef_x;
    p.coef_y = coef_y;
    p.tw = reinterpret_cast<const cpx<double>*>(twiddle_table(PYFFT_F64, SegGeom<LOG2L>::L, &status));
    if (!p.tw) return status;
    const int rc = launch_moments<double, LOG2L, 0, 8>(p, nch, out, ws, wsb, st);
    if (rc) return rc;
    return launch_moments<double, LOG2L, 8, 4>(p, nch, out, ws, wsb, st);
}

}  // namespace pyfft

using namespace pyfft;

extern "C" {

int64_t pyfft_segment_moments_workspace_bytes(int dtype, const pyfft_segmentation* seg, int32_t nch) {
    if (!seg || nch <= 0 || seg->nfft < 64 || seg->nfft > 8192)
        return fail(PYFFT_EINVAL, "segment_moments_workspace_bytes: bad arguments");
    const int tps = seg->nfft / 16;
    const int threads = tps > 128 ? tps : 128;
    const int slots = threads / tps;
    const int gx = seg_grid_x(seg->seg_count, slots, nch);
    const int64_t nm = dtype == PYFFT_F64 ? 8 : SEG_NMOM;
    return (int64_t)nch * gx * slots * nm * (seg->nfft / 2 + 1) * (dtype == PYFFT_F64 ? 8 : 4);
}

int pyfft_segment_moments(int dtype, const void* x, const void* y, int64_t y_stride, int32_t nch,
                          const pyfft_record_map* map, const pyfft_segmentation* seg, const void* window,
                          const double* coef_x, const double* coef_y, double* out, void* workspace,
                          int64_t workspace_bytes, void* stream) {
    RecMap m;
    int rc = check_map(map, &m);
    if (rc) return rc;
    int log2l;
    if ((rc = check_seg(seg, 6, 13, &log2l))) return rc;
    if (!x || !y || !window || !out || nch <= 0) return fail(PYFFT_EINVAL, "segment_moments: NULL pointer or nch <= 0");
    if (dtype != PYFFT_F32 && dtype != PYFFT_F64) return fail(PYFFT_EINVAL, "unknown dtype %d", dtype);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (seg->seg_count == 0) {
        PYFFT_CUDA_OK(cudaMemsetAsync(out, 0, sizeof(double) * (size_t)nch * SEG_NMOM * (seg->nfft / 2 + 1), st));
        return PYFFT_OK;
    }
    const int64_t j_lo = seg->seg_first * (int64_t)seg->hop;
    const int64_t j_hi = (seg->seg_first + seg->seg_count - 1) * (int64_t)seg->hop + seg->nfft;
    if ((rc = check_range(m, j_lo, j_hi))) return rc;
#define PYFFT_CASE(N) \
    case N: return launch_moments_any<N>(dtype, x, y, y_stride, nch, m, seg, window, coef_x, coef_y, out, workspace, \
                                         workspace_bytes, st);
    switch (log2l) {
        PYFFT_CASE(6) PYFFT_CASE(7) PYFFT_CASE(8) PYFFT_CASE(9) PYFFT_CASE(10) PYFFT_CASE(11) PYFFT_CASE(12) PYFFT_CASE(13)
    }
#undef PYFFT_CASE
    return fail(PYFFT_EUNSUPPORTED, "segment_moments: nfft = 2^%d not supported", log2l);
}

}  // extern "C"
