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
#include <stdlib.h>

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

// Register-tiled product for fp32 spectra and nch = 16 / 32 / 64: a thread owns an 8 x 8 block
// (rows 8a.., columns 8b.., b >= a: G is Hermitian, the lower triangle is mirrored when the sums
// are flushed -- also inside the diagonal blocks, so G is Hermitian bit for bit with a real diagonal) of one bin's matrix = 128 fp32 accumulators and 256 FFMA per segment for 8
// 16-byte shared-memory loads.  A CTA owns FB consecutive bins (25 for 32 channels: 8 warps, one
// 200-byte run per channel and segment) and a range of segments; the spectra X[c][k][f] are transposed on the fly into
// [k][bin][c] by 8-byte cp.async copies, in a ring of 3 stages of 4 segments.  Inside a row
// channel 8a + 2h + p sits in 16-byte slot (h*NB + a) ^ csm_swz(bin), half p.
// XOR applied to the 16-byte slot index (h*NB + block) of a row: neighbouring bins land in
// opposite 64-byte halves, bins 2 apart in different slots -- the 16-byte loads of a quarter warp
// (one h, the row blocks of at most two bins) and the 8-byte cp.async writes of consecutive bins
// both spread over the banks
__host__ __device__ constexpr int csm_swz(int fb) { return ((fb & 1) << 2) | ((fb >> 1) & 3); }

template <int C> struct CsmTile {
    static constexpr int NB = C / 8, NT = NB * (NB + 1) / 2, KB = 4;
    static constexpr int FB = 256 / NT > 32 ? 32 : 256 / NT;     // bins per CTA: 32 / 25 / 7
    static constexpr int THREADS = (FB * NT + 31) / 32 * 32;     // 96 / 256 / 256 (the last lanes idle)
    static constexpr int MINB = THREADS >= 256 ? 1 : 2;          // 8 x 8 tiles want ~250 registers
    static constexpr int ROW = C * 8;                    // bytes per (segment, bin) row, 16-byte slots XOR-swizzled by the bin
    static constexpr int STAGE = KB * FB * ROW;
    static constexpr int NSTAGE = 3;
    static constexpr int SMEM = NSTAGE * STAGE;
};

template <int C>
__global__ void __launch_bounds__(CsmTile<C>::THREADS, CsmTile<C>::MINB)
csm_tile_kernel(const cpx<float>* __restrict__ X, int64_t K, int nb, int ksplit, double* __restrict__ G) {
    using Tl = CsmTile<C>;
    constexpr int NB = Tl::NB, NT = Tl::NT, FB = Tl::FB, KB = Tl::KB, THREADS = Tl::THREADS, ROW = Tl::ROW;
    extern __shared__ __align__(16) unsigned char csm_smem[];
    const int tid = threadIdx.x;
    const bool lane_on = tid < FB * NT;
    const int fb = lane_on ? tid / NT : 0;
    int tile = lane_on ? tid - fb * NT : 0, a = 0;
    while (tile >= NB - a) { tile -= NB - a; ++a; }
    const int b = a + tile;
    const int f0 = blockIdx.x * FB;
    const int64_t kper = (K + ksplit - 1) / ksplit;
    const int64_t k0 = (int64_t)blockIdx.y * kper;
    const int64_t k1 = k0 + kper < K ? k0 + kper : K;
    if (k0 >= k1) return;
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(csm_smem);
    const int swz = csm_swz(fb);
    const int shi = swz & ~(NB - 1) & (4 * NB - 1), alo = a ^ (swz & (NB - 1)), blo = b ^ (swz & (NB - 1));

    float ar[8][8], ai[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) { ar[i][j] = 0.f; ai[i][j] = 0.f; }

    // staging: thread-invariant (channel, bin) slots, NSLOT per thread and segment; only the
    // segment index changes from copy to copy
    constexpr int NSLOT = (C * FB + THREADS - 1) / THREADS;
    uint32_t goff[NSLOT], soff[NSLOT];                       // element offset in X (without k), byte offset in a stage
    uint32_t okmask = 0;
#pragma unroll
    for (int i = 0; i < NSLOT; ++i) {
        const int e = tid + THREADS * i;
        const int c = e / FB, fbb = e - c * FB;
        const bool ok = e < C * FB && f0 + fbb < nb;
        goff[i] = ok ? (uint32_t)((int64_t)c * K * nb + f0 + fbb) : 0u;
        soff[i] = (uint32_t)(fbb * ROW + 16 * ((((c >> 1) & 3) * NB + (c >> 3)) ^ csm_swz(fbb)) + (c & 1) * 8);
        if (e < C * FB) okmask |= (ok ? 1u : 0u) << (2 * i) | 2u << (2 * i);   // bit 2i: copy, bit 2i+1: slot exists
    }
    auto issue = [&](int stage, int64_t kb) {
        const uint32_t dst0 = sbase + (uint32_t)(stage * Tl::STAGE);
#pragma unroll
        for (int q = 0; q < KB; ++q) {
            const int64_t k = kb + q;
            const bool kok = k < k1;
            const cpx<float>* srow = X + (kok ? k : k0) * nb;
#pragma unroll
            for (int i = 0; i < NSLOT; ++i) {
                if (okmask >> (2 * i + 1) & 1u) {
                    const int nbytes = (kok && (okmask >> (2 * i) & 1u)) ? 8 : 0;      // 0: zero fill
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(dst0 + q * (FB * ROW) + soff[i]),
                                 "l"(srow + goff[i]), "r"(nbytes) : "memory");
                }
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    // NSTAGE-deep ring: the copies of stage s + NSTAGE - 1 are in flight while stage s is consumed
    constexpr int NSTAGE = Tl::NSTAGE;
    int64_t knext = k0;
#pragma unroll
    for (int sidx = 0; sidx < NSTAGE - 1; ++sidx) {
        if (knext < k1) issue(sidx, knext);
        else asm volatile("cp.async.commit_group;" ::: "memory");
        knext += KB;
    }
    int stage = 0, fill = NSTAGE - 1;
    for (int64_t kb = k0; kb < k1; kb += KB) {
        asm volatile("cp.async.wait_group %0;" ::"n"(NSTAGE - 2) : "memory");
        __syncthreads();                                     // stage `stage` landed; stage `fill` is free again
        if (knext < k1) issue(fill, knext);
        else asm volatile("cp.async.commit_group;" ::: "memory");
        knext += KB;
        fill = fill + 1 == NSTAGE ? 0 : fill + 1;
        const unsigned char* row = csm_smem + stage * Tl::STAGE + fb * ROW;
#pragma unroll
        for (int q = 0; q < KB; ++q) {
            const unsigned char* pq = row + q * (FB * ROW);
            float4 xi[4], xj[4];
#pragma unroll
            for (int h = 0; h < 4; ++h) {
                xi[h] = *reinterpret_cast<const float4*>(pq + (((h * NB) ^ shi) | alo) * 16);
                xj[h] = *reinterpret_cast<const float4*>(pq + (((h * NB) ^ shi) | blo) * 16);
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float ir = (i & 1) ? xi[i >> 1].z : xi[i >> 1].x;
                const float ii = (i & 1) ? xi[i >> 1].w : xi[i >> 1].y;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float jr = (j & 1) ? xj[j >> 1].z : xj[j >> 1].x;
                    const float ji = (j & 1) ? xj[j >> 1].w : xj[j >> 1].y;
                    // X_j conj(X_i)
                    ar[i][j] = fmaf(jr, ir, ar[i][j]);
                    ar[i][j] = fmaf(ji, ii, ar[i][j]);
                    ai[i][j] = fmaf(ji, ir, ai[i][j]);
                    ai[i][j] = fmaf(-jr, ii, ai[i][j]);
                }
            }
        }
        stage = stage + 1 == NSTAGE ? 0 : stage + 1;
    }
    const int f = f0 + fb;
    if (f >= nb || !lane_on) return;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int gi = 8 * a + i, gj = 8 * b + j;
            if (gj < gi) continue;                      // diagonal blocks: upper triangle only, mirrored below
            double* g = G + (((int64_t)f * C + gi) * C + gj) * 2;
            atomicAdd(g, (double)ar[i][j]);
            if (gi == gj) continue;                     // G_ii is real: |X_i|^2
            atomicAdd(g + 1, (double)ai[i][j]);
            {
                double* gt = G + (((int64_t)f * C + gj) * C + gi) * 2;
                atomicAdd(gt, (double)ar[i][j]);
                atomicAdd(gt + 1, -(double)ai[i][j]);
            }
        }
    }
}

template <int C>
static int launch_csm_tile(const cpx<float>* X, int64_t cnt, int nb, double* G, int sms, cudaStream_t st) {
    using Tl = CsmTile<C>;
    auto kern = csm_tile_kernel<C>;
    PYFFT_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Tl::SMEM));
    int per_sm = 0;
    PYFFT_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, Tl::THREADS, Tl::SMEM));
    if (per_sm < 1) per_sm = 1;
    const int gx = (nb + Tl::FB - 1) / Tl::FB;
    // about six waves of resident CTAs, at least 16 segments per CTA
    int64_t ksplit = ((int64_t)6 * per_sm * sms + gx - 1) / gx;
    if (ksplit > cnt / 16) ksplit = cnt / 16;
    if (ksplit < 1) ksplit = 1;
    kern<<<dim3(gx, (unsigned)ksplit), Tl::THREADS, Tl::SMEM, st>>>(X, cnt, nb, (int)ksplit, G);
    PYFFT_LAUNCH_OK("csm_tile_kernel");
    return PYFFT_OK;
}

template <typename T>
static int csm_tile_dispatch(const cpx<T>*, int64_t, int, int, double*, int, cudaStream_t, bool* done) {
    *done = false;
    return PYFFT_OK;
}
template <>
int csm_tile_dispatch<float>(const cpx<float>* X, int64_t cnt, int nb, int nch, double* G, int sms, cudaStream_t st,
                             bool* done) {
    *done = true;
    switch (nch) {
        case 16: return launch_csm_tile<16>(X, cnt, nb, G, sms, st);
        case 32: return launch_csm_tile<32>(X, cnt, nb, G, sms, st);
        case 64: return launch_csm_tile<64>(X, cnt, nb, G, sms, st);
    }
    *done = false;
    return PYFFT_OK;
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
    // up to 16 GB of spectra per chunk (180 GB of HBM: config 5's per-GPU share is one chunk, so
    // the accumulators are flushed once), an even number of segments (K2 packs segment pairs)
    int64_t cap = int64_t(16) << 30;
    if (const char* e = getenv("PYFFT_CSM_CHUNK_BYTES")) {     // tests: force several chunks on small records
        const long long v = atoll(e);
        if (v > 0) cap = v;
    }
    int64_t kc = cap / ((int64_t)nch * nb * esz);
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
        bool done = false;
        rc = csm_tile_dispatch<T>(X, cnt, nb, nch, G, sms, st, &done);
        if (rc) return rc;
        if (done) continue;
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
