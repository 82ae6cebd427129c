// K1-TC: the Welch cross-spectral accumulation for nfft = 4096 on the 5th-generation tensor
// cores (tcgen05.mma, accumulators in TMEM), replacing the register/shared-memory FFT of
// welch_pair_kernel when fp32 data and mean detrend are asked for.
//
// Reference lines replaced: fft_analysis.py:353-393 (global mean removal, segment loop, window,
// np.fft.fft, products) and the segment sum of :444-446 -- same as pyfft_welch_accumulate_mean.
//
// DFT as two real GEMMs (n = 64 n1 + n2, f = k1 + 64 k2, W_N = exp(-2 pi i / N)):
//   stage 1   A1[k1, n2] = sum_n1 x~[64 n1 + n2] W_64^(n1 k1)             k1 = 0..32 only (x real)
//             D1[(ch, n2), j] = sum_n1 A1op[(ch, n2), n1] * G[j, n1]      M = 128, N = 64, K = 64
//             j <= 32: Re A1[k1 = j],  j >= 33: Im A1[k1 = j - 32]
//   twiddle   B[k1, n2] = A1[k1, n2] * W_4096^(n2 k1)                     CUDA cores, out of TMEM
//   stage 2   X[k1 + 64 k2] = sum_n2 B[k1, n2] W_64^(n2 k2)
//             D2[(k2, re|im), (seg, ch, k1)] = F2[(k2, c'), (n2, c)] * B2[(seg, ch, k1), (n2, c)]
//                                                                         M = 128, N = 224, K = 128
// Lane (k2, c) of D2 holds one component of X and Y of 33 bins for every segment of the batch,
// so the cross-spectral sums are thread-local (one shuffle for the mixed Im term) and every
// lane owns distinct bins: 132 accumulator registers, no redundancy.  Rows k1 = 33..63 of
// stage 1 are never computed: bins with f mod 64 > 32 come from lanes k2 >= 32 by Hermitian
// symmetry X[4096 - f] = conj X[f].
// Precision: operands are fp16 with fp32 accumulation.  The DFT matrices are split in two fp16
// terms (hi + lo, 22 significant bits, two MMAs per product); the data is scaled by a power of
// two to unit-ish rms and rounded ONCE to fp16 at each stage -- a zero-mean relative error of
// 2^-12 per element that averages out over the segments (use for Navr >~ 10^4).
#include <cuda_fp16.h>

#include <map>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "tc_core.cuh"

namespace pyfft {

namespace tcw {
constexpr int L = 4096, NBIN = L / 2 + 1;
constexpr int SEGS = 3;                      // segments per batch (per team)
constexpr int COLS_PER_SEG = 72;             // D2 columns per segment: X at +0 (33), Y at +36 (33)
constexpr int YOFF = 36;                     // (multiples of 4 and of 8*k+4: cheap swizzle arithmetic)
constexpr int N2 = 224;                      // stage-2 MMA N (3 * 72 = 216, padded to a multiple of 16)
constexpr int NLIN = 8;                      // bins with linear sums (mean correction)
// constant block (global -> shared), byte offsets
constexpr int OFF_GHI = 0, OFF_GLO = 8192, OFF_F2HI = 16384, OFF_F2LO = 49152, OFF_TW = 81920;
constexpr int CONST_BYTES = 81920 + 33 * 64 * 8;           // 98816
constexpr int OFF_WIN = CONST_BYTES;                         // float[4096]
constexpr int OFF_TEAM = 115712;                             // 1024-aligned
constexpr int TEAM_BYTES = N2 * 128 * 2;                     // B2: 224 rows x 128 halves (2 K atoms) = 57344
constexpr int SMEM_BYTES = OFF_TEAM + 2 * TEAM_BYTES + 1024; // + alignment slack
}  // namespace tcw

struct TcParams {
    const float* x;
    const float* y;
    int64_t y_stride;
    RecMap map;
    int64_t hop, seg_first, seg_count, chunk;
    const float* window;
    const double* est;        // (mean, -) pairs: x first, then the y channels
    const double* prescale;   // power-of-two scale per channel, same order
    const unsigned char* consts;
    float* partial;           // [nch][nparts][4][NBIN]
    int nparts;
    float* lin_partial;       // [nch][nparts][4][L/16]  (zeroed by the host; bins < NLIN written)
    double* dev_sums;         // [nch+1]
    int* status;
};

__device__ __forceinline__ void team_sync(int team) { asm volatile("bar.sync %0, 128;" ::"r"(team + 1) : "memory"); }

__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

__global__ void __launch_bounds__(256, 1) welch_tc_kernel(const TcParams p) {
    using namespace tcw;
    using namespace tc;
    extern __shared__ unsigned char smem_raw[];
    __shared__ __align__(8) unsigned long long mbar_store[2];
    __shared__ uint32_t tmem_slot;

    const uint32_t raw = (uint32_t)__cvta_generic_to_shared(smem_raw);
    const uint32_t pad = (1024u - (raw & 1023u)) & 1023u;
    unsigned char* sm = smem_raw + pad;
    const uint32_t sm_addr = raw + pad;

    const int tid = threadIdx.x;
    const int team = tid >> 7, tt = tid & 127, warp_t = tt >> 5;
    const int pair = blockIdx.y;

    // ---- one-time setup: constants, window, barriers, tensor memory ---------------------- //
    {
        const uint4* src = reinterpret_cast<const uint4*>(p.consts);
        uint4* dst = reinterpret_cast<uint4*>(sm);
        for (int i = tid; i < CONST_BYTES / 16; i += 256) dst[i] = src[i];
        float* sw = reinterpret_cast<float*>(sm + OFF_WIN);
        for (int i = tid; i < L; i += 256) sw[i] = p.window[i];
    }
    const uint32_t mbar = (uint32_t)__cvta_generic_to_shared(&mbar_store[team]);
    if (tt == 0) { mbar_init(mbar, 1); fence_mbar_init(); }
    if (tid < 32) tmem_alloc((uint32_t)__cvta_generic_to_shared(&tmem_slot), 512);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tcol = tmem_slot + (uint32_t)team * 256u;                // this team's column base
    const uint32_t tlane = tcol + ((uint32_t)(warp_t * 32) << 16);          // + this warp's lane quarter

    // ---- per-thread roles ------------------------------------------------------------------ //
    // loader / twiddle stage: row r = tt = (ch, n2) of the stage-1 operand
    const int ch = tt >> 6, n2 = tt & 63;
    const int cidx = ch == 0 ? 0 : 1 + pair;
    const float* __restrict__ src = ch == 0 ? p.x : p.y + (int64_t)pair * p.y_stride;
    const float mean_f = (float)p.est[2 * cidx];
    const float sc = (float)p.prescale[cidx];
    const float nms = -mean_f * sc;                                         // exact: sc is a power of two
    const float* __restrict__ swin = reinterpret_cast<const float*>(sm + OFF_WIN);
    const float2* __restrict__ stw = reinterpret_cast<const float2*>(sm + OFF_TW);
    unsigned char* __restrict__ tbuf = sm + OFF_TEAM + team * TEAM_BYTES;
    const uint32_t tbuf_addr = sm_addr + OFF_TEAM + team * TEAM_BYTES;
    const uint32_t a1_row = (uint32_t)((tt >> 3) * 1024 + (tt & 7) * 128);  // own row of a 128-row K atom
    const int hb = (int)(p.hop >> 9);                                       // 8-sample chunks inside the hop
    // stage-2 operand: 16-byte chunk of this thread's (n2, re|im) pair inside row r, XOR (r & 7);
    // r & 7 = (k1 + 4 ch) & 7 because the row bases are multiples of 8 (x) or 8m + 4 (y)
    uint32_t xorv[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) xorv[i] = (uint32_t)((((n2 & 31) >> 2) ^ ((i + 4 * ch) & 7)) << 4);

    // epilogue: lane (k2, c) of the stage-2 accumulator
    float acc[33][4];
    float linx[NLIN], liny[NLIN];
#pragma unroll
    for (int k = 0; k < 33; ++k)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[k][q] = 0.f;
#pragma unroll
    for (int k = 0; k < NLIN; ++k) { linx[k] = 0.f; liny[k] = 0.f; }
    float devsum = 0.f;

    const int part = blockIdx.x * 2 + team;
    const int64_t c0 = (int64_t)blockIdx.x * p.chunk;
    const int64_t c1 = (c0 + p.chunk < p.seg_count) ? c0 + p.chunk : p.seg_count;
    uint32_t parity = 0;
    bool ok = true;

    for (int64_t kb = c0 + (int64_t)team * SEGS; kb < c1; kb += 2 * SEGS) {
        // ================= P1: gather -> detrend -> window -> fp16 -> stage-1 operand ======== //
#pragma unroll 1
        for (int sg = 0; sg < SEGS; ++sg) {
            const int64_t k = kb + sg;
            unsigned char* rowp = tbuf + sg * 16384 + a1_row;
            if (k >= c1) {
#pragma unroll
                for (int c = 0; c < 8; ++c) *reinterpret_cast<uint4*>(rowp + (c << 4)) = make_uint4(0, 0, 0, 0);
                continue;
            }
            const int64_t j0 = (p.seg_first + k) * p.hop;
            const bool plain = j0 >= p.map.reflect;
            const float* __restrict__ q = src + (j0 - p.map.reflect + p.map.i0 - p.map.src_base) + n2;
#pragma unroll
            for (int hf = 0; hf < 2; ++hf) {                   // 32 loads in flight per thread
                float v[32];
                if (plain) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = q[64 * (32 * hf + i)];
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = src[src_index(p.map, j0 + 64 * (32 * hf + i) + n2)];
                }
#pragma unroll
                for (int cc = 0; cc < 4; ++cc) {
                    const int c = 4 * hf + cc;
                    float* u = v + 8 * cc;
                    float wv[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) wv[i] = swin[64 * (8 * c + i) + n2];
#pragma unroll
                    for (int i = 0; i < 8; ++i) u[i] = fmaf(u[i], sc, nms);               // (x - mean) * scale
                    if (c < hb) devsum += ((u[0] + u[1]) + (u[2] + u[3])) + ((u[4] + u[5]) + (u[6] + u[7]));
#pragma unroll
                    for (int i = 0; i < 8; ++i) u[i] *= wv[i];
                    uint4 w;
                    w.x = pack_h2(u[0], u[1]); w.y = pack_h2(u[2], u[3]);
                    w.z = pack_h2(u[4], u[5]); w.w = pack_h2(u[6], u[7]);
                    *reinterpret_cast<uint4*>(rowp + ((c ^ (tt & 7)) << 4)) = w;
                }
            }
        }
        // pull this team's next batch into L2 while the tensor core works on this one
        {
            const int64_t kn = kb + 2 * SEGS;
            const int64_t jn = (p.seg_first + kn) * p.hop;
            if (kn < c1 && jn >= p.map.reflect) {
                const float* nb = src + (jn - p.map.reflect + p.map.i0 - p.map.src_base);
                const int64_t nsamp = (int64_t)(SEGS - 1) * p.hop + L;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int64_t off = ((int64_t)n2 * 4 + i) * 32;
                    if (off < nsamp) asm volatile("prefetch.global.L2 [%0];" ::"l"(nb + off));
                }
            }
        }
        fence_async_smem();
        team_sync(team);

        // ================= P2: stage-1 MMAs (one thread) ===================================== //
        if (tt == 0) {
            tc_fence_after();
            constexpr uint32_t id1 = idesc_f16(128, 64);
#pragma unroll 1
            for (int sg = 0; sg < SEGS; ++sg) {
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks)
                        mma_f16(tcol + sg * 64, sw128_desc(tbuf_addr + sg * 16384 + ks * 32),
                                sw128_desc(sm_addr + (h ? OFF_GLO : OFF_GHI) + ks * 32), id1, (h | ks) ? 1u : 0u);
            }
            mma_commit(mbar);
        }
        ok = mbar_wait(mbar, parity) && ok;
        parity ^= 1;
        tc_fence_after();

        // ================= P3: twiddle, fp16, stage-2 operand ================================= //
#pragma unroll 1
        for (int sg = 0; sg < SEGS; ++sg) {
            float d0[32], d1[32];
            tmem_ld32(tlane + sg * 64, d0);
            tmem_ld32(tlane + sg * 64 + 32, d1);
            tmem_ld_wait();
            // d0[k1] = Re A1[k1] (k1 < 32), d1[0] = Re A1[32], d1[k1] = Im A1[k1] (k1 = 1..31)
            // row r = sg*72 + ch*36 + k1 of the stage-2 operand; 8-row groups are contiguous, so the
            // row offset is r*128 and only the 16-byte chunk index is XOR-ed with (r & 7)
            unsigned char* bp = tbuf + (n2 >> 5) * (N2 * 128) + (n2 & 3) * 4 + (sg * COLS_PER_SEG + ch * YOFF) * 128;
#pragma unroll
            for (int g = 0; g < 5; ++g) {                        // k1 = 8g .. 8g+7 (g = 4: k1 = 32 only)
                float2 t[8];
#pragma unroll
                for (int i = 0; i < 8; ++i)
                    if (8 * g + i <= 32) t[i] = stw[(8 * g + i) * 64 + n2];      // twiddles of the group first
                uint32_t w[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int k1 = 8 * g + i;
                    if (k1 <= 32) {
                        const float re = k1 < 32 ? d0[k1] : d1[0];
                        const float im = (k1 >= 1 && k1 <= 31) ? d1[k1] : 0.f;
                        w[i] = pack_h2(re * t[i].x - im * t[i].y, re * t[i].y + im * t[i].x);
                    }
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int k1 = 8 * g + i;
                    if (k1 <= 32) *reinterpret_cast<uint32_t*>(bp + k1 * 128 + xorv[k1 & 7]) = w[i];
                }
            }
        }
        tc_fence_before();
        fence_async_smem();
        team_sync(team);

        // ================= P4: stage-2 MMAs (one thread) ===================================== //
        if (tt == 0) {
            tc_fence_after();
            constexpr uint32_t id2 = idesc_f16(128, N2);
#pragma unroll
            for (int h = 0; h < 2; ++h)
#pragma unroll
                for (int ks = 0; ks < 8; ++ks)
                    mma_f16(tcol, sw128_desc(sm_addr + (h ? OFF_F2LO : OFF_F2HI) + (ks >> 2) * (128 * 128) + (ks & 3) * 32),
                            sw128_desc(tbuf_addr + (ks >> 2) * (N2 * 128) + (ks & 3) * 32), id2, (h | ks) ? 1u : 0u);
            mma_commit(mbar);
        }
        ok = mbar_wait(mbar, parity) && ok;
        parity ^= 1;
        tc_fence_after();

        // ================= P5: cross-spectral accumulation out of TMEM ======================== //
#pragma unroll 1
        for (int sg = 0; sg < SEGS; ++sg) {
            float xv[32], yv[32], x32, y32, pad0, pad1;
            tmem_ld32(tlane + sg * COLS_PER_SEG, xv);
            tmem_ld2(tlane + sg * COLS_PER_SEG + 32, x32, pad0);
            tmem_ld32(tlane + sg * COLS_PER_SEG + YOFF, yv);
            tmem_ld2(tlane + sg * COLS_PER_SEG + YOFF + 32, y32, pad1);
            tmem_ld_wait();
#pragma unroll
            for (int k1 = 0; k1 <= 32; ++k1) {
                const float xx = k1 < 32 ? xv[k1] : x32;
                const float yy = k1 < 32 ? yv[k1] : y32;
                const float xo = __shfl_xor_sync(0xffffffffu, xx, 1);   // the other component of X
                acc[k1][0] = fmaf(xx, xx, acc[k1][0]);
                acc[k1][1] = fmaf(yy, yy, acc[k1][1]);
                acc[k1][2] = fmaf(yy, xx, acc[k1][2]);
                acc[k1][3] = fmaf(yy, xo, acc[k1][3]);
                if (k1 < NLIN) { linx[k1] += xx; liny[k1] += yy; }
            }
        }
        tc_fence_before();
    }

    // ---- write the team's partial sums ------------------------------------------------------- //
    {
        const float sx = (float)p.prescale[0], sy = (float)p.prescale[1 + pair];
        const float i00 = 1.f / (sx * sx), i11 = 1.f / (sy * sy), i01 = 1.f / (sx * sy);
        const int k2 = tt >> 1, c = tt & 1;
        float* out = p.partial + ((size_t)pair * p.nparts + part) * 4 * NBIN;
#pragma unroll
        for (int k1 = 0; k1 <= 32; ++k1) {
            const float t0 = acc[k1][0] + __shfl_xor_sync(0xffffffffu, acc[k1][0], 1);
            const float t1 = acc[k1][1] + __shfl_xor_sync(0xffffffffu, acc[k1][1], 1);
            const float t2 = acc[k1][2] + __shfl_xor_sync(0xffffffffu, acc[k1][2], 1);
            const float o3 = __shfl_xor_sync(0xffffffffu, acc[k1][3], 1);
            // even lane: acc3 = sum Yr Xi, partner's = sum Yi Xr  ->  Im(Y conj X) = Yi Xr - Yr Xi
            float t3 = o3 - acc[k1][3];
            if (c == 0) {
                int f = k1 + 64 * k2;
                bool wr = true;
                if (k2 >= 32) {
                    if (f == L / 2) { /* Nyquist */ }
                    else if (k1 >= 1 && k1 <= 31) { f = L - f; t3 = -t3; }
                    else wr = false;
                }
                if (wr) {
                    out[f] = t0 * i00;
                    out[NBIN + f] = t1 * i11;
                    out[2 * NBIN + f] = t2 * i01;
                    out[3 * NBIN + f] = t3 * i01;
                }
            }
        }
        if (k2 == 0) {
            float* lo = p.lin_partial + ((size_t)pair * p.nparts + part) * 4 * (L / 16);
#pragma unroll
            for (int k1 = 0; k1 < NLIN; ++k1) {
                lo[(0 + c) * (L / 16) + k1] = linx[k1] / sx;       // rows 0/1: Re/Im sum U
                lo[(2 + c) * (L / 16) + k1] = liny[k1] / sy;       // rows 2/3: Re/Im sum V
            }
        }
        double ds = (double)devsum / (double)sc;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ds += __shfl_xor_sync(0xffffffffu, ds, o);
        if ((tt & 31) == 0 && (ch == 1 || pair == 0)) atomicAdd(&p.dev_sums[cidx], ds);
    }
    if (!ok && tt == 0) atomicExch(p.status, 1);

    tc_fence_before();
    __syncthreads();
    if (tid < 32) tmem_free(tmem_slot, 512);
}

// ------------------------------------------------------------------------- //
// provisional statistics: mean (rounded to float) and a power-of-two scale that brings the
// rms of (x - mean) to ~1/8, from the first samples of the span
// ------------------------------------------------------------------------- //
// pass 1: est[2c] += sum (x - x0), est[2c+1] += sum (x - x0)^2 with x0 = the first sample of the span
__global__ void __launch_bounds__(256) tc_moments_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                         int64_t y_stride, RecMap map, int64_t j_begin, int64_t j_end,
                                                         double* __restrict__ est) {
    const int c = blockIdx.y;                       // 0 = x, 1.. = y channels
    const float* p = c == 0 ? x : y + (int64_t)(c - 1) * y_stride;
    const double x0 = (double)p[src_index(map, j_begin)];
    double s0 = 0.0, s1 = 0.0;
    const int64_t step = (int64_t)gridDim.x * 256;
    for (int64_t j = j_begin + (int64_t)blockIdx.x * 256 + threadIdx.x; j < j_end; j += step) {
        const double v = (double)p[src_index(map, j)] - x0;
        s0 += v;
        s1 += v * v;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s0 += __shfl_xor_sync(0xffffffffu, s0, o);
        s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&est[2 * c], s0);
        atomicAdd(&est[2 * c + 1], s1);
    }
}

// pass 2: est[2c] = mean (rounded to fp32), est[2c+1] = 0, prescale[c] = 2^k with rms * 2^k in [1/8, 1/4)
__global__ void tc_prepare_finish_kernel(const float* __restrict__ x, const float* __restrict__ y, int64_t y_stride,
                                         RecMap map, int64_t j_begin, int64_t j_end, int nch1,
                                         double* __restrict__ est, double* __restrict__ prescale) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= nch1) return;
    const float* p = c == 0 ? x : y + (int64_t)(c - 1) * y_stride;
    const double x0 = (double)p[src_index(map, j_begin)];
    const double n = (double)(j_end - j_begin);
    const double md = est[2 * c] / n;
    double var = est[2 * c + 1] / n - md * md;
    if (!(var > 0.0)) var = 0.0;
    const double rms = sqrt(var);
    int e = 0;
    if (rms > 0.0 && isfinite(rms)) frexp(rms, &e);   // rms = f * 2^e, f in [0.5, 1)
    int sh2 = -e - 2;
    if (sh2 > 100) sh2 = 100;
    if (sh2 < -100) sh2 = -100;
    est[2 * c] = (double)(float)(x0 + md);            // exactly representable in fp32
    est[2 * c + 1] = 0.0;
    prescale[c] = ldexp(1.0, sh2);
}

// ------------------------------------------------------------------------- //
// host side
// ------------------------------------------------------------------------- //
static std::mutex g_tc_mutex;
static std::map<int, unsigned char*> g_tc_consts;

static void put_half(std::vector<unsigned char>& buf, size_t off, __half h) {
    memcpy(buf.data() + off, &h, 2);
}

static const unsigned char* tc_constants(int* status) {
    using namespace tcw;
    *status = PYFFT_OK;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { *status = fail(PYFFT_ECUDA, "cudaGetDevice failed"); return nullptr; }
    std::lock_guard<std::mutex> lock(g_tc_mutex);
    auto it = g_tc_consts.find(dev);
    if (it != g_tc_consts.end()) return it->second;
    std::vector<unsigned char> buf(CONST_BYTES, 0);
    const long double two_pi = 6.283185307179586476925286766559005768L;
    auto split = [&](long double v, size_t off_hi, size_t off_lo) {
        const __half hi = __float2half_rn((float)v);
        const __half lo = __float2half_rn((float)(v - (long double)__half2float(hi)));
        put_half(buf, off_hi, hi);
        put_half(buf, off_lo, lo);
    };
    // stage 1, B operand: G[j][n1], K-major rows j
    for (int j = 0; j < 64; ++j)
        for (int n1 = 0; n1 < 64; ++n1) {
            const int k1 = j <= 32 ? j : j - 32;
            const long double a = two_pi * (long double)((n1 * k1) & 63) / 64.0L;
            const long double v = j <= 32 ? cosl(a) : -sinl(a);
            const uint32_t o = tc::sw128_offset(64, j, n1);
            split(v, OFF_GHI + o, OFF_GLO + o);
        }
    // stage 2, A operand: F2[(k2, co)][(n2, ci)]
    for (int k2 = 0; k2 < 64; ++k2)
        for (int n2 = 0; n2 < 64; ++n2) {
            const long double a = two_pi * (long double)((n2 * k2) & 63) / 64.0L;
            const long double cs = cosl(a), sn = sinl(a);
            const long double m[2][2] = {{cs, sn}, {-sn, cs}};     // [co][ci]
            for (int co = 0; co < 2; ++co)
                for (int ci = 0; ci < 2; ++ci) {
                    const uint32_t o = tc::sw128_offset(128, 2 * k2 + co, 2 * n2 + ci);
                    split(m[co][ci], OFF_F2HI + o, OFF_F2LO + o);
                }
        }
    // twiddles W_4096^(n2 k1) = (cos, -sin), [k1][n2]
    for (int k1 = 0; k1 <= 32; ++k1)
        for (int n2 = 0; n2 < 64; ++n2) {
            const long double a = two_pi * (long double)(n2 * k1) / 4096.0L;
            const float t[2] = {(float)cosl(a), (float)(-sinl(a))};
            memcpy(buf.data() + OFF_TW + (size_t)(k1 * 64 + n2) * 8, t, 8);
        }
    unsigned char* d = nullptr;
    cudaError_t e = cudaMalloc(&d, buf.size());
    if (e == cudaSuccess) e = cudaMemcpy(d, buf.data(), buf.size(), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        *status = fail(PYFFT_ECUDA, "tensor-core constant upload failed: %s", cudaGetErrorString(e));
        if (d) cudaFree(d);
        return nullptr;
    }
    g_tc_consts[dev] = d;
    return d;
}

static int tc_grid_x(int64_t seg_count, int nch) {
    int sms = device_sm_count();
    if (sms <= 0) sms = 148;
    int64_t want = (sms + nch - 1) / nch;
    const int64_t maxb = (seg_count + 2 * tcw::SEGS - 1) / (2 * tcw::SEGS);
    if (want > maxb) want = maxb;
    if (want < 1) want = 1;
    return (int)want;
}

}  // namespace pyfft

using namespace pyfft;

extern "C" {

int64_t pyfft_welch_tc_workspace_bytes(const pyfft_segmentation* seg, int32_t nch) {
    if (!seg || nch <= 0) return fail(PYFFT_EINVAL, "welch_tc_workspace_bytes: bad arguments");
    const int gx = tc_grid_x(seg->seg_count, nch);
    return (int64_t)nch * gx * 2 * 4 * (tcw::NBIN + tcw::L / 16) * 4 + 8 * (int64_t)(nch + 1) + 16;
}

int pyfft_welch_tc_prepare(const void* x, const void* y, int64_t y_stride, int32_t nch,
                           const pyfft_record_map* map, int64_t j_begin, int64_t j_end, double* est,
                           double* prescale, void* stream) {
    RecMap m;
    int rc = check_map(map, &m);
    if (rc) return rc;
    if (!x || !y || !est || !prescale || nch <= 0) return fail(PYFFT_EINVAL, "welch_tc_prepare: bad arguments");
    if ((rc = check_range(m, j_begin, j_end))) return rc;
    if (j_end <= j_begin) return fail(PYFFT_EINVAL, "welch_tc_prepare: empty sample span");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    PYFFT_CUDA_OK(cudaMemsetAsync(est, 0, sizeof(double) * 2 * (size_t)(nch + 1), st));
    int gx = (int)((j_end - j_begin + 2047) / 2048);
    if (gx > 64) gx = 64;
    tc_moments_kernel<<<dim3(gx, nch + 1), 256, 0, st>>>(static_cast<const float*>(x), static_cast<const float*>(y),
                                                        y_stride, m, j_begin, j_end, est);
    tc_prepare_finish_kernel<<<(nch + 1 + 63) / 64, 64, 0, st>>>(static_cast<const float*>(x),
                                                               static_cast<const float*>(y), y_stride, m, j_begin,
                                                               j_end, nch + 1, est, prescale);
    PYFFT_LAUNCH_OK("tc_prepare_kernel");
    return PYFFT_OK;
}

int pyfft_welch_accumulate_tc(const void* x, const void* y, int64_t y_stride, int32_t nch,
                              const pyfft_record_map* map, const pyfft_segmentation* seg, const void* window,
                              const double* est, const double* prescale, double* acc, double* lin, double* stats,
                              void* workspace, int64_t workspace_bytes, int32_t* status_flag, void* stream) {
    using namespace tcw;
    RecMap m;
    int rc = check_map(map, &m);
    if (rc) return rc;
    if (!seg || seg->nfft != L) return fail(PYFFT_EUNSUPPORTED, "the tensor-core path handles nfft = 4096 only");
    if (seg->hop <= 0 || seg->hop > L || (seg->hop & 511))
        return fail(PYFFT_EUNSUPPORTED, "the tensor-core path needs a hop that is a multiple of 512 (got %d)", seg->hop);
    if (!x || !y || !window || !est || !prescale || !acc || !lin || !stats || !status_flag || nch <= 0)
        return fail(PYFFT_EINVAL, "welch_accumulate_tc: NULL pointer or nch <= 0");
    if (seg->seg_count == 0) return PYFFT_OK;
    const int64_t j_lo = seg->seg_first * (int64_t)seg->hop;
    const int64_t j_hi = (seg->seg_first + seg->seg_count - 1) * (int64_t)seg->hop + seg->nfft;
    if ((rc = check_range(m, j_lo, j_hi))) return rc;
    int status;
    const unsigned char* consts = tc_constants(&status);
    if (!consts) return status;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int gx = tc_grid_x(seg->seg_count, nch);
    const int nparts = gx * 2;
    const int64_t main_bytes = (int64_t)nch * nparts * 4 * NBIN * 4;
    const int64_t lin_bytes = (int64_t)nch * nparts * 4 * (L / 16) * 4;
    if (!workspace || workspace_bytes < main_bytes + lin_bytes + 8 * (int64_t)(nch + 1))
        return fail(PYFFT_EINVAL, "welch_accumulate_tc: workspace too small");
    TcParams p;
    p.x = static_cast<const float*>(x);
    p.y = static_cast<const float*>(y);
    p.y_stride = y_stride;
    p.map = m;
    p.hop = seg->hop;
    p.seg_first = seg->seg_first;
    p.seg_count = seg->seg_count;
    // chunk: a multiple of one round of both teams (6 segments) so that batches stay full
    int64_t chunk = (seg->seg_count + gx - 1) / gx;
    chunk = (chunk + 2 * SEGS - 1) / (2 * SEGS) * (2 * SEGS);
    p.chunk = chunk;
    p.window = static_cast<const float*>(window);
    p.est = est;
    p.prescale = prescale;
    p.consts = consts;
    p.partial = static_cast<float*>(workspace);
    p.nparts = nparts;
    p.lin_partial = reinterpret_cast<float*>(static_cast<char*>(workspace) + main_bytes);
    p.dev_sums = reinterpret_cast<double*>(static_cast<char*>(workspace) + main_bytes + lin_bytes);
    p.status = status_flag;
    // teams that get no segment still write zeros; the linear sums are sparse -> clear them
    PYFFT_CUDA_OK(cudaMemsetAsync(p.partial, 0, (size_t)(main_bytes + lin_bytes) + sizeof(double) * (size_t)(nch + 1), st));
    PYFFT_CUDA_OK(cudaFuncSetAttribute(welch_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    welch_tc_kernel<<<dim3(gx, nch), 256, SMEM_BYTES, st>>>(p);
    PYFFT_LAUNCH_OK("welch_tc_kernel");
    if ((rc = welch_reduce_launch_f32(p.lin_partial, nparts, 4 * (L / 16), nch, lin, st))) return rc;
    if ((rc = welch_raw_sums_launch(p.dev_sums, est, nch + 1, (double)seg->seg_count * (double)seg->hop, stats, st)))
        return rc;
    return welch_reduce_launch_f32(p.partial, nparts, 4 * NBIN, nch, acc, st);
}

}  // extern "C"
