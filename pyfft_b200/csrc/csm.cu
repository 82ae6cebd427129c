// K4: multi-channel cross-spectral matrix  G[f][i][j] += sum_k X_j[k][f] conj(X_i[k][f])
//
// Reference lines replaced: the per-reference-channel loop over fft_pwelch in
// HeatPulse_Funcs.py:576-583 (each call = fft_analysis.py:359-393 with sigx = channel i and
// sigy = all channels; Pxy[:, j] = mean_k Y_j conj(X_i), :393).
//
// Segments are processed in chunks: K2 (pyfft_stft) writes the chunk's spectra
// X[c][k][f] (complex, compute precision) into the workspace, then csm_product_kernel forms
// the nch x nch rank-1 updates per bin on the CUDA cores and adds them to the double
// accumulator.  A CTA owns FB consecutive bins and all nch rows; thread (i, fb) keeps row i of
// G[f0+fb] (nch complex sums) in registers and streams the segments of the chunk through
// shared memory.
#include "common.cuh"

namespace pyfft {

constexpr int CSM_FB = 8;        // bins per CTA
constexpr int CSM_KB = 4;        // segments staged per barrier

template <typename T, int C>
__global__ void __launch_bounds__(C * CSM_FB)
csm_product_kernel(const cpx<T>* __restrict__ X, int64_t K, int nb, int ksplit, double* __restrict__ G) {
    // X: [C][K][nb]; G: [nb][C][C][2]
    constexpr int FB = CSM_FB, KB = CSM_KB;
    __shared__ __align__(16) cpx<T> sx[2][KB][C][FB];
    const int fb = threadIdx.x % FB;
    const int i = threadIdx.x / FB;
    const int f = blockIdx.x * FB + fb;
    const bool fok = f < nb;
    const int64_t kper = (K + ksplit - 1) / ksplit;
    const int64_t k0 = (int64_t)blockIdx.y * kper;
    const int64_t k1 = k0 + kper < K ? k0 + kper : K;
    T ar[C], ai[C];
#pragma unroll
    for (int j = 0; j < C; ++j) { ar[j] = T(0); ai[j] = T(0); }
    const cpx<T>* src = X + (int64_t)i * K * nb + (fok ? f : 0);
    int buf = 0;
    for (int64_t kb = k0; kb < k1; kb += KB) {
#pragma unroll
        for (int q = 0; q < KB; ++q) {
            const int64_t k = kb + q;
            sx[buf][q][i][fb] = (fok && k < k1) ? src[k * nb] : mk<T>(T(0), T(0));
        }
        __syncthreads();                    // two buffers: one barrier per stage is enough
#pragma unroll
        for (int q = 0; q < KB; ++q) {
            const cpx<T> xi = sx[buf][q][i][fb];
#pragma unroll
            for (int j = 0; j < C; ++j) {
                const cpx<T> xj = sx[buf][q][j][fb];
                // X_j conj(X_i)
                ar[j] += xj.x * xi.x + xj.y * xi.y;
                ai[j] += xj.y * xi.x - xj.x * xi.y;
            }
        }
        buf ^= 1;
    }
    if (!fok) return;
    double* g = G + ((int64_t)f * C + i) * C * 2;
    if (ksplit == 1) {
#pragma unroll
        for (int j = 0; j < C; ++j) { g[2 * j] += (double)ar[j]; g[2 * j + 1] += (double)ai[j]; }
    } else {
#pragma unroll
        for (int j = 0; j < C; ++j) { atomicAdd(&g[2 * j], (double)ar[j]); atomicAdd(&g[2 * j + 1], (double)ai[j]); }
    }
}

// generic channel count (any nch <= 64): one thread per (i, j, fb) would be wasteful; instead
// thread (i, fb) loops j with accumulators in local memory only for small nch -- used for
// channel counts without a specialised instance
template <typename T>
__global__ void __launch_bounds__(256)
csm_product_generic_kernel(const cpx<T>* __restrict__ X, int64_t K, int nb, int nch, double* __restrict__ G) {
    // one thread per (f, i, j)
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t total = (int64_t)nb * nch * nch;
    if (idx >= total) return;
    const int j = (int)(idx % nch);
    const int i = (int)((idx / nch) % nch);
    const int f = (int)(idx / ((int64_t)nch * nch));
    const cpx<T>* xi = X + (int64_t)i * K * nb + f;
    const cpx<T>* xj = X + (int64_t)j * K * nb + f;
    double sr = 0.0, si = 0.0;
    for (int64_t k = 0; k < K; ++k) {
        const cpx<T> a = xi[k * nb], b = xj[k * nb];
        sr += (double)b.x * a.x + (double)b.y * a.y;
        si += (double)b.y * a.x - (double)b.x * a.y;
    }
    G[2 * idx] += sr;
    G[2 * idx + 1] += si;
}

static int64_t csm_chunk(int64_t seg_count, int nch, int nb, int64_t esz) {
    // ~512 MB of spectra per chunk, an even number of segments (K2 packs segment pairs)
    int64_t kc = (int64_t(512) << 20) / ((int64_t)nch * nb * esz);
    kc &= ~int64_t(1);
    if (kc < 2) kc = 2;
    if (kc > seg_count) kc = seg_count;
    return kc;
}

template <typename T>
static int csm_impl(int dtype, const void* x, int64_t ch_stride, int nch, const pyfft_record_map* map,
                    const pyfft_segmentation* seg, const void* window, const double* coef, double* G,
                    void* ws, int64_t wsb, cudaStream_t st) {
    const int nb = seg->nfft / 2 + 1;
    const int64_t esz = 2 * (int64_t)sizeof(T);
    const int64_t kc = csm_chunk(seg->seg_count, nch, nb, esz);
    const int64_t need = (int64_t)nch * kc * nb * esz;
    if (!ws || wsb < need)
        return fail(PYFFT_EINVAL, "csm: workspace too small: %lld < %lld bytes", (long long)wsb, (long long)need);
    int sms = device_sm_count();
    if (sms <= 0) sms = 148;
    for (int64_t s0 = 0; s0 < seg->seg_count; s0 += kc) {
        const int64_t cnt = s0 + kc <= seg->seg_count ? kc : seg->seg_count - s0;
        pyfft_segmentation sub = {seg->nfft, seg->hop, seg->seg_first + s0, cnt};
        int rc = pyfft_stft(dtype, 0, x, ch_stride, nch, map, &sub, window, coef, 0, 0, nb, nullptr, 0, ws,
                            nullptr, st);
        if (rc) return rc;
        const cpx<T>* X = static_cast<const cpx<T>*>(ws);
        const int gx = (nb + CSM_FB - 1) / CSM_FB;
        int ksplit = (4 * sms + gx - 1) / gx;
        if (ksplit < 1) ksplit = 1;
        if ((int64_t)ksplit * CSM_KB * 4 > cnt) ksplit = 1;
        switch (nch) {
#define PYFFT_CASE(C)                                                                              \
            case C:                                                                                \
                csm_product_kernel<T, C><<<dim3(gx, ksplit), C * CSM_FB, 0, st>>>(X, cnt, nb, ksplit, G); \
                break;
            PYFFT_CASE(2) PYFFT_CASE(4) PYFFT_CASE(8) PYFFT_CASE(16) PYFFT_CASE(32)
#undef PYFFT_CASE
            default: {
                const int64_t total = (int64_t)nb * nch * nch;
                csm_product_generic_kernel<T><<<(unsigned)((total + 255) / 256), 256, 0, st>>>(X, cnt, nb, nch, G);
            }
        }
        PYFFT_LAUNCH_OK("csm_product_kernel");
    }
    return PYFFT_OK;
}

}  // namespace pyfft

using namespace pyfft;

extern "C" {

int64_t pyfft_csm_workspace_bytes(int dtype, const pyfft_segmentation* seg, int32_t nch) {
    int log2l;
    int rc = check_seg(seg, 4, 13, &log2l);
    if (rc) return rc;
    if (nch <= 0 || nch > 64) return fail(PYFFT_EINVAL, "csm: 1 <= nch <= 64 required");
    const int nb = seg->nfft / 2 + 1;
    const int64_t esz = dtype == PYFFT_F64 ? 16 : 8;
    const int64_t kc = csm_chunk(seg->seg_count > 0 ? seg->seg_count : 1, nch, nb, esz);
    return (int64_t)nch * kc * nb * esz;
}

int pyfft_csm_accumulate(int dtype, const void* x, int64_t ch_stride, int32_t nch,
                         const pyfft_record_map* map, const pyfft_segmentation* seg, const void* window,
                         const double* coef, double* G, void* workspace, int64_t workspace_bytes, void* stream) {
    int log2l;
    int rc = check_seg(seg, 4, 13, &log2l);
    if (rc) return rc;
    if (!x || !window || !G || nch <= 0 || nch > 64)
        return fail(PYFFT_EINVAL, "csm_accumulate: NULL pointer or nch outside 1..64");
    if (seg->seg_count == 0) return PYFFT_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (dtype == PYFFT_F32)
        return csm_impl<float>(dtype, x, ch_stride, nch, map, seg, window, coef, G, workspace, workspace_bytes, st);
    if (dtype == PYFFT_F64)
        return csm_impl<double>(dtype, x, ch_stride, nch, map, seg, window, coef, G, workspace, workspace_bytes, st);
    return fail(PYFFT_EINVAL, "unknown dtype %d", dtype);
}

}  // extern "C"
