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
//                                                                         M = 128, N = 144, K = 128
// Lane (k2, c) of D2 holds one component of X and Y of 33 bins for both segments of a batch,
// so the cross-spectral sums are thread-local (one shuffle for the mixed Im term) and every
// lane owns distinct bins: 132 accumulator registers, no redundancy.  Rows k1 = 33..63 of
// stage 1 are never computed: bins with f mod 64 > 32 come from lanes k2 >= 32 by Hermitian
// symmetry X[4096 - f] = conj X[f].
// Precision: operands are fp16 with fp32 accumulation.  The DFT matrices are split in two fp16
// terms (hi + lo, 22 significant bits, two MMAs per product); the data is scaled by a power of
// two to unit-ish rms and rounded ONCE to fp16 at each stage -- a zero-mean relative error of
// 2^-12 per element that averages out over the segments (use for Navr >~ 10^4).
//
// Execution: one persistent CTA per SM, warp-specialised, all hand-shakes through mbarriers:
//   warpgroup 0  epilogue   D2 (TMEM) -> register accumulators                 (128 threads)
//   warpgroup 1  twiddle    D1 (TMEM) -> twiddle -> fp16 -> B2 operand (smem)   (128 threads)
//   warpgroup 2-5 loaders   global -> detrend -> window -> fp16 -> A1 operand; set g (8 warps)
//                           takes the segments of parity g and owns A1 slot g   (2 x 256 threads)
//   warps 24, 25 MMA issue  one thread each for stage 1 and stage 2 (independent instruction
//                           streams; the stages meet only through the mbarriers)
// A1: 2 slots (segment parity); D1: 3 slots (segment index % 3); B2, D2: 2 slots (batch parity).  Registers
// are moved from the twiddle and MMA warpgroups to the epilogue group with setmaxnreg.
#include <cuda_fp16.h>

#include <map>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "tc_core.cuh"

namespace pyfft {

namespace tcw {
constexpr int L = 4096, NBIN = L / 2 + 1;
constexpr int SEGS = 2;                      // segments per batch (one B2 slot, one twiddle pass)
constexpr int YOFF = 36;                     // D2 columns of a segment: X at +0 (33 bins), Y at +36 (33 bins); rows of B2 alike
constexpr int COLS_PER_SEG = 72;             // D2 columns / B2 rows per segment
constexpr int N2 = SEGS * COLS_PER_SEG;      // stage-2 MMA N = 144 (a multiple of 16)
constexpr int NLIN = 8;                      // bins with linear sums (mean correction)
constexpr int NTHREADS = 896;
constexpr int HEAD = 3 * L;                  // samples per channel of the materialised reflect zone
// constant block in global memory: [shared part][F2 hi, plain rows][F2 lo, plain rows]
// shared part (copied to shared memory), byte offsets
constexpr int OFF_GHI = 0, OFF_GLO = 8192, OFF_TW = 16384;
constexpr int CONST_SMEM_BYTES = 16384 + 33 * 64 * 8;        // 33280
// stage-2 matrix F2[(k2, c')][(n2, c)] as 128 rows of 64 packed half2, loaded into tensor memory
constexpr int OFF_F2T_HI = CONST_SMEM_BYTES, OFF_F2T_LO = OFF_F2T_HI + 128 * 64 * 4;
constexpr int CONST_BYTES = OFF_F2T_LO + 128 * 64 * 4;       // 98816
constexpr int WIN_LD = 68;                                   // window stored [n2][n1], padded: conflict-free LDS.128
constexpr int OFF_WIN = CONST_SMEM_BYTES;                    // float[64][WIN_LD] = 17408 bytes
constexpr int OFF_A1 = 51200;                                // 1024-aligned; 2 slots of 128 rows x 64 halves
constexpr int A1_BYTES = 16384;
constexpr int OFF_B2 = OFF_A1 + 2 * A1_BYTES;                // 2 slots of N2 rows x 128 halves (2 K atoms)
constexpr int B2_BYTES = N2 * 128 * 2;                       // 36864
constexpr int B2_SEG_BYTES = COLS_PER_SEG * 128;             // row offset of segment 1 inside a K atom
constexpr int SMEM_BYTES = OFF_B2 + 2 * B2_BYTES + 1024;     // + alignment slack = 158720
static_assert(OFF_WIN + 64 * WIN_LD * 4 <= OFF_A1, "window table overlaps the operand slots");
// tensor memory columns (480 of 512)
constexpr int TM_F2HI = 0, TM_F2LO = 64;                     // stage-2 A operand: 128 lanes x 64 columns of half2
constexpr int TM_D1 = 128;                                   // 3 slots x 64 columns (segment index % 3)
constexpr int TM_D2 = 320;                                   // 1 slot of N2 columns
// mbarriers
enum { A1_FULL = 0, S1_DONE = 2, D1_FREE = 5, B2_FULL = 8, B2_FREE = 10, S2_DONE = 12, D2_FREE = 14, NBARS = 16 };
// registers per thread after reallocation (launch: 72; released 512*16 + 128*16 + 128*24 = requested 128*104)
constexpr int REGS_EPI = 176, REGS_TWD = 56, REGS_LDR = 56, REGS_MMA = 48;
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
    const float* head;        // [nch+1][HEAD]: analysed samples j < head_len materialised (reflect zone)
    int64_t head_len;         // segments that start before map.reflect read from head instead of x / y
    float* partial;           // [nch][nparts][4][NBIN]
    int nparts;
    float* lin_partial;       // [nch][nparts][4][L/16]  (zeroed by the host; bins < NLIN written)
    double* dev_sums;         // [nch+1]
    int* status;
    long long* trace;         // debug: [batch][16] clock64 stamps of CTA (0, 0), or NULL
};

__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

#define PYFFT_TC_TRACE(EV, B) \
    if (trace_on) p.trace[(B) * 16 + (EV)] = clock64();

// one cross-spectral update of a lane's accumulators for bin slot k1
#define PYFFT_TC_ACC(K1, XX, YY)                                              \
    {                                                                         \
        const float xo_ = __shfl_xor_sync(0xffffffffu, (XX), 1);              \
        acc[K1][0] = fmaf((XX), (XX), acc[K1][0]);                            \
        acc[K1][1] = fmaf((YY), (YY), acc[K1][1]);                            \
        acc[K1][2] = fmaf((YY), (XX), acc[K1][2]);                            \
        acc[K1][3] = fmaf((YY), xo_, acc[K1][3]);                             \
    }

__global__ void __launch_bounds__(tcw::NTHREADS, 1) welch_tc_kernel(const TcParams p) {
    using namespace tcw;
    using namespace tc;
    extern __shared__ unsigned char smem_raw[];
    __shared__ __align__(8) unsigned long long bars[NBARS];
    __shared__ uint32_t tmem_slot;
    __shared__ int abort_store;

    const uint32_t raw = (uint32_t)__cvta_generic_to_shared(smem_raw);
    const uint32_t pad = (1024u - (raw & 1023u)) & 1023u;
    unsigned char* sm = smem_raw + pad;
    const uint32_t sm_addr = raw + pad;
    const uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(&bars[0]);
    volatile int* abort_flag = &abort_store;

    const int tid = threadIdx.x, warp = tid >> 5, wg = warp >> 2;
    const int pair = blockIdx.y;
    const bool trace_on = p.trace != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && (tid & 31) == 0 &&
                          ((tid & 127) == 0 || tid == 800);
    if (p.trace != nullptr && tid == 0) {
        unsigned long long t_;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
        p.trace[2048 + 4 * (blockIdx.y * gridDim.x + blockIdx.x)] = (long long)t_;
        p.trace[2048 + 4 * (blockIdx.y * gridDim.x + blockIdx.x) + 2] = clock64();
    }

    // ---- one-time setup: constants, window, barriers, tensor memory ---------------------- //
    {
        const uint4* src = reinterpret_cast<const uint4*>(p.consts);
        uint4* dst = reinterpret_cast<uint4*>(sm);
        for (int i = tid; i < CONST_SMEM_BYTES / 16; i += NTHREADS) dst[i] = src[i];
        float* sw = reinterpret_cast<float*>(sm + OFF_WIN);
        for (int n = tid; n < L; n += NTHREADS) sw[(n & 63) * WIN_LD + (n >> 6)] = p.window[n];
    }
    if (tid == 0) {
        abort_store = 0;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar0 + 8 * (A1_FULL + i), 256);
            mbar_init(bar0 + 8 * (B2_FULL + i), 128);
            mbar_init(bar0 + 8 * (B2_FREE + i), 1);
            mbar_init(bar0 + 8 * (S2_DONE + i), 1);
            mbar_init(bar0 + 8 * (D2_FREE + i), 128);
        }
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            mbar_init(bar0 + 8 * (S1_DONE + i), 1);
            mbar_init(bar0 + 8 * (D1_FREE + i), 128);
        }
        fence_mbar_init();
    }
    if (warp == 0) tmem_alloc((uint32_t)__cvta_generic_to_shared(&tmem_slot), 512);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tbase = tmem_slot;
    if (tid < 128) {
        // stage-2 matrix into tensor memory: thread = lane = row (k2, c'), 64 columns of half2 each for hi / lo
        const uint32_t tl = tbase + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
        for (int part = 0; part < 2; ++part) {
            const uint4* row = reinterpret_cast<const uint4*>(p.consts + (part ? OFF_F2T_LO : OFF_F2T_HI) + tid * 256);
#pragma unroll 1
            for (int c = 0; c < 64; c += 32) {            // 8 independent 16-byte loads in flight
                uint4 u[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) u[i] = row[c / 4 + i];
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    tmem_st8(tl + (part ? TM_F2LO : TM_F2HI) + c + 8 * i, u[2 * i].x, u[2 * i].y, u[2 * i].z, u[2 * i].w,
                             u[2 * i + 1].x, u[2 * i + 1].y, u[2 * i + 1].z, u[2 * i + 1].w);
            }
        }
        tmem_st_wait();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();

    const int64_t c0 = (int64_t)blockIdx.x * p.chunk;
    const int64_t c1 = (c0 + p.chunk < p.seg_count) ? c0 + p.chunk : p.seg_count;
    const int nb = c1 > c0 ? (int)((c1 - c0 + 1) >> 1) : 0;                  // batches of two segments

    if (wg == 0) {
        // ================= epilogue: D2 -> cross-spectral sums in registers =================== //
        setmaxnreg_inc<REGS_EPI>();
        const uint32_t tl = tbase + ((uint32_t)(warp * 32) << 16) + TM_D2;   // this warp's lane quarter
        float acc[33][4];
        float linx[NLIN], liny[NLIN];
#pragma unroll
        for (int k = 0; k < 33; ++k)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[k][q] = 0.f;
#pragma unroll
        for (int k = 0; k < NLIN; ++k) { linx[k] = 0.f; liny[k] = 0.f; }

        bool live = true;
#pragma unroll 1
        for (int b = 0; b < nb && live; ++b) {
#pragma unroll 1                                             // keeps the loop body small for the instruction cache
            for (int sg = 0; sg < SEGS; ++sg) {
                if (sg == 0) {
                    if (!mbar_wait_hot(bar0 + 8 * S2_DONE, (uint32_t)b & 1u, abort_flag)) { live = false; break; }
                    tc_fence_after();
                    PYFFT_TC_TRACE(10, b)
                }
                const uint32_t col = tl + sg * COLS_PER_SEG;
                // software pipeline over chunks of 8 bins: the next chunk is in flight during the FMAs
                float xa[8], ya[8], xb[8], yb[8];
                tmem_ld8(col, xa);
                tmem_ld8(col + YOFF, ya);
                tmem_ld_wait();
                tmem_ld8(col + 8, xb);
                tmem_ld8(col + YOFF + 8, yb);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    PYFFT_TC_ACC(i, xa[i], ya[i]);
                    if (i < NLIN) { linx[i] += xa[i]; liny[i] += ya[i]; }
                }
                tmem_ld_wait();
                tmem_ld8(col + 16, xa);
                tmem_ld8(col + YOFF + 16, ya);
#pragma unroll
                for (int i = 0; i < 8; ++i) PYFFT_TC_ACC(8 + i, xb[i], yb[i]);
                tmem_ld_wait();
                tmem_ld8(col + 24, xb);
                tmem_ld8(col + YOFF + 24, yb);
#pragma unroll
                for (int i = 0; i < 8; ++i) PYFFT_TC_ACC(16 + i, xa[i], ya[i]);
                tmem_ld_wait();
                float x32, y32, d0, d1;
                tmem_ld2(col + 32, x32, d0);
                tmem_ld2(col + YOFF + 32, y32, d1);
#pragma unroll
                for (int i = 0; i < 8; ++i) PYFFT_TC_ACC(24 + i, xb[i], yb[i]);
                tmem_ld_wait();
                if (sg == SEGS - 1) {                                        // D2 is in registers now
                    tc_fence_before();
                    mbar_arrive(bar0 + 8 * D2_FREE);
                    PYFFT_TC_TRACE(11, b)
                }
                PYFFT_TC_ACC(32, x32, y32);
            }
        }

        // ---- write the CTA's partial sums ---------------------------------------------------- //
        const float sx = (float)p.prescale[0], sy = (float)p.prescale[1 + pair];
        const float i00 = 1.f / (sx * sx), i11 = 1.f / (sy * sy), i01 = 1.f / (sx * sy);
        const int k2 = tid >> 1, c = tid & 1;
        float* out = p.partial + ((size_t)pair * p.nparts + blockIdx.x) * 4 * NBIN;
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
            float* lo = p.lin_partial + ((size_t)pair * p.nparts + blockIdx.x) * 4 * (L / 16);
#pragma unroll
            for (int k1 = 0; k1 < NLIN; ++k1) {
                lo[(0 + c) * (L / 16) + k1] = linx[k1] / sx;       // rows 0/1: Re/Im sum U
                lo[(2 + c) * (L / 16) + k1] = liny[k1] / sy;       // rows 2/3: Re/Im sum V
            }
        }
    } else if (wg == 1) {
        // ================= twiddle: D1 -> W_4096^(n2 k1) -> fp16 -> stage-2 operand ========== //
        // Both segments of a batch in lock-step (each twiddle factor is read once for the two), in
        // groups of 4 k1 with the next group's TMEM loads in flight during the arithmetic.
        setmaxnreg_dec<REGS_TWD>();
        const int ln = tid - 128;                                            // TMEM lane = row (ch, n2) of D1
        const int ch = ln >> 6, n2 = ln & 63;
        const uint32_t tl = tbase + ((uint32_t)((warp & 3) * 32) << 16) + TM_D1;
        const float2* __restrict__ tw = reinterpret_cast<const float2*>(sm + OFF_TW) + n2;
        // row r = sg*72 + ch*36 + k1 of the operand; 8-row groups are contiguous, so the row offset
        // is r*128 and only the 16-byte chunk index is XOR-ed with r & 7 = (k1 + 4 ch) & 7
        const uint32_t thread_off = (uint32_t)((n2 >> 5) * (N2 * 128) + (n2 & 3) * 4 + ch * YOFF * 128);   // inside a segment block
        const uint32_t xbase = (uint32_t)((((n2 & 31) >> 2) ^ (4 * ch)) << 4);   // chunk ^ (k1 & 7) = xbase ^ (i << 4)

#pragma unroll 1
        for (int b = 0; b < nb; ++b) {
            const int s = b & 1;
            const int s0 = 2 * b, s1 = 2 * b + 1;                            // segment indices; D1 slot = index % 3
            const int d0 = s0 % 3, d1 = s1 % 3;
            // the stage-2 MMAs of batch b-2 have finished reading this B2 slot
            if (b >= 2 && !mbar_wait_hot(bar0 + 8 * (B2_FREE + s), ((uint32_t)(b >> 1) & 1u) ^ 1u, abort_flag)) break;
            // stage 1 of the batch's second segment done implies the first one (same issuing thread,
            // the tensor pipe completes in order)
            PYFFT_TC_TRACE(7, b)
            if (!mbar_wait_hot(bar0 + 8 * (S1_DONE + d1), (uint32_t)(s1 / 3) & 1u, abort_flag)) break;
            tc_fence_after();
            PYFFT_TC_TRACE(8, b)
            const uint32_t ca = tl + d0 * 64, cb = tl + d1 * 64;
            unsigned char* bp = sm + OFF_B2 + s * B2_BYTES + thread_off;     // segment 1: + B2_SEG_BYTES
            // group buffers: [0..3] Re seg 0, [4..7] Im seg 0, [8..11] Re seg 1, [12..15] Im seg 1
            float ga[16], gb[16];
            tmem_ld4(ca, &ga[0]);
            tmem_ld4(ca + 32, &ga[4]);                    // first value = Re A1[32]; Im A1[0] = 0
            tmem_ld4(cb, &ga[8]);
            tmem_ld4(cb + 32, &ga[12]);
            tmem_ld_wait();
            const float re32a = ga[4], re32b = ga[12];
            ga[4] = 0.f;
            ga[12] = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) {                 // k1 = 4j .. 4j+3
                const float* cur = (j & 1) ? gb : ga;
                float* nxt = (j & 1) ? ga : gb;
                if (j < 7) {
                    tmem_ld4(ca + 4 * (j + 1), &nxt[0]);
                    tmem_ld4(ca + 32 + 4 * (j + 1), &nxt[4]);
                    tmem_ld4(cb + 4 * (j + 1), &nxt[8]);
                    tmem_ld4(cb + 32 + 4 * (j + 1), &nxt[12]);
                }
                float2 t[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) t[i] = tw[(4 * j + i) * 64];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t off = (uint32_t)((4 * j + i) * 128) + (xbase ^ (uint32_t)(((4 * j + i) & 7) << 4));
                    *reinterpret_cast<uint32_t*>(bp + off) =
                        pack_h2(cur[i] * t[i].x - cur[4 + i] * t[i].y, cur[i] * t[i].y + cur[4 + i] * t[i].x);
                    *reinterpret_cast<uint32_t*>(bp + B2_SEG_BYTES + off) =
                        pack_h2(cur[8 + i] * t[i].x - cur[12 + i] * t[i].y, cur[8 + i] * t[i].y + cur[12 + i] * t[i].x);
                }
                if (j < 7) tmem_ld_wait();
                if (j == 6) {                             // the last D1 columns are in registers: slots free
                    tc_fence_before();
                    mbar_arrive(bar0 + 8 * (D1_FREE + d0));
                    mbar_arrive(bar0 + 8 * (D1_FREE + d1));
                }
            }
            {
                const float2 t32 = tw[32 * 64];
                *reinterpret_cast<uint32_t*>(bp + 32 * 128 + xbase) = pack_h2(re32a * t32.x, re32a * t32.y);
                *reinterpret_cast<uint32_t*>(bp + B2_SEG_BYTES + 32 * 128 + xbase) = pack_h2(re32b * t32.x, re32b * t32.y);
            }
            fence_async_smem();
            mbar_arrive(bar0 + 8 * (B2_FULL + s));
            PYFFT_TC_TRACE(9, b)
        }
    } else if (wg <= 5) {
        // ================= loaders: gather -> detrend -> window -> fp16 -> stage-1 operand ==== //
        // Two sets of 8 warps; set g takes the segments of parity g and owns A1 slot g.  Thread
        // (quarter, n2) converts the 16 samples n = 64 (16 quarter + i) + n2 of BOTH channels, so each
        // window value is read once for the two.  The loads of the set's next segment are issued
        // before it waits for the slot: their latency overlaps the stage-1 MMAs of the previous batch
        // and the other set's conversion.
        setmaxnreg_dec<REGS_LDR>();
        const int lt = tid - 256;
        const int g = lt >> 8;                                               // set = segment parity = A1 slot
        const int qt = (lt >> 6) & 3, n2 = lt & 63;
        const float* __restrict__ ysrc = p.y + (int64_t)pair * p.y_stride;
        const float scx = (float)p.prescale[0], scy = (float)p.prescale[1 + pair];
        const float nmx = -(float)p.est[0] * scx, nmy = -(float)p.est[2 * (1 + pair)] * scy;   // exact: powers of two
        const float4* __restrict__ wrow =
            reinterpret_cast<const float4*>(sm + OFF_WIN + n2 * (WIN_LD * 4) + qt * 64);
        // rows n2 (x) and 64 + n2 (y) of the slot, 16-byte chunks 2 qt and 2 qt + 1
        unsigned char* rowx = sm + OFF_A1 + g * A1_BYTES + (n2 >> 3) * 1024 + (n2 & 7) * 128;
        const uint32_t ck0 = (uint32_t)(((2 * qt) ^ (n2 & 7)) << 4), ck1 = (uint32_t)(((2 * qt + 1) ^ (n2 & 7)) << 4);
        const int hb = (int)(p.hop >> 9) - 2 * qt;                           // chunks of this quarter inside the hop
        const float in0 = hb > 0 ? 1.f : 0.f, in1 = hb > 1 ? 1.f : 0.f;
        // segments that start in the reflect-padded zone read the materialised head of the record
        const int64_t poff = p.map.i0 - p.map.reflect - p.map.src_base + n2 + 1024 * qt;
        const float* __restrict__ hx = p.head + n2 + 1024 * qt;
        const float* __restrict__ hy = hx + (int64_t)(1 + pair) * HEAD;
        // L2 prefetch two segments of this set ahead: 128 lines of 128 bytes per channel; threads
        // (qt, n2 < 32) take line qt*32 + n2 of x, threads n2 >= 32 the same line of y
        const float* pf = (n2 < 32 ? p.x : ysrc) + (p.map.i0 - p.map.reflect - p.map.src_base) + (qt * 32 + (n2 & 31)) * 32;
        const int64_t first = p.seg_first + c0 + g;
        float sumx = 0.f, sumy = 0.f;

#pragma unroll 1
        for (int b = 0; b < nb; ++b) {
            const bool have = c0 + 2 * b + g < c1;
            float vx[16], vy[16];
            if (have) {
                const int64_t j0 = (first + 2 * b) * p.hop;
                const bool zone = j0 < p.map.reflect;
                const float* __restrict__ qx = zone ? hx + j0 : p.x + (j0 + poff);
                const float* __restrict__ qy = zone ? hy + j0 : ysrc + (j0 + poff);
#pragma unroll
                for (int i = 0; i < 16; ++i) vx[i] = qx[64 * i];
#pragma unroll
                for (int i = 0; i < 16; ++i) vy[i] = qy[64 * i];
                if (c0 + 2 * b + g + 4 < c1) {
                    const int64_t jn = (first + 2 * b + 4) * p.hop;
                    if (jn >= p.map.reflect) asm volatile("prefetch.global.L2 [%0];" ::"l"(pf + jn));
                }
            } else {
#pragma unroll
                for (int i = 0; i < 16; ++i) { vx[i] = 0.f; vy[i] = 0.f; }
            }
            // the stage-1 MMAs of this set's previous segment have finished reading the A1 slot
            if (b >= 1) {
                const int sp = 2 * (b - 1) + g;
                if (!mbar_wait_abortable<200>(bar0 + 8 * (S1_DONE + sp % 3), (uint32_t)(sp / 3) & 1u, abort_flag)) break;
            }
            if (qt == 0) { PYFFT_TC_TRACE(2 * g, b) }
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                float* ux = vx + 8 * c;
                float* uy = vy + 8 * c;
                const float4 w0 = wrow[2 * c], w1 = wrow[2 * c + 1];
#pragma unroll
                for (int i = 0; i < 8; ++i) { ux[i] = fmaf(ux[i], scx, nmx); uy[i] = fmaf(uy[i], scy, nmy); }   // (x - mean) * scale
                if (have) {
                    const float in = c ? in1 : in0;
                    sumx = fmaf(in, ((ux[0] + ux[1]) + (ux[2] + ux[3])) + ((ux[4] + ux[5]) + (ux[6] + ux[7])), sumx);
                    sumy = fmaf(in, ((uy[0] + uy[1]) + (uy[2] + uy[3])) + ((uy[4] + uy[5]) + (uy[6] + uy[7])), sumy);
                }
                uint4 ox, oy;
                ox.x = pack_h2(ux[0] * w0.x, ux[1] * w0.y); ox.y = pack_h2(ux[2] * w0.z, ux[3] * w0.w);
                ox.z = pack_h2(ux[4] * w1.x, ux[5] * w1.y); ox.w = pack_h2(ux[6] * w1.z, ux[7] * w1.w);
                oy.x = pack_h2(uy[0] * w0.x, uy[1] * w0.y); oy.y = pack_h2(uy[2] * w0.z, uy[3] * w0.w);
                oy.z = pack_h2(uy[4] * w1.x, uy[5] * w1.y); oy.w = pack_h2(uy[6] * w1.z, uy[7] * w1.w);
                if (!have) { ox = make_uint4(0, 0, 0, 0); oy = ox; }
                *reinterpret_cast<uint4*>(rowx + (c ? ck1 : ck0)) = ox;
                *reinterpret_cast<uint4*>(rowx + 8192 + (c ? ck1 : ck0)) = oy;
            }
            fence_async_smem();
            mbar_arrive(bar0 + 8 * (A1_FULL + g));
            if (qt == 0) { PYFFT_TC_TRACE(2 * g + 1, b) }
        }
        double dx = (double)sumx / (double)scx, dy = (double)sumy / (double)scy;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            dx += __shfl_xor_sync(0xffffffffu, dx, o);
            dy += __shfl_xor_sync(0xffffffffu, dy, o);
        }
        if ((tid & 31) == 0) {
            if (pair == 0) atomicAdd(&p.dev_sums[0], dx);
            atomicAdd(&p.dev_sums[1 + pair], dy);
        }
    } else {
        // ================= MMA issue: one thread per GEMM stage ================================ //
        // Two warps so that the two serial instruction streams (waits, descriptors, issue) run on
        // different schedulers; the stages only meet through the mbarriers.
        setmaxnreg_dec<REGS_MMA>();
        if (tid == 768) {
            // stage 1: one 128 x 64 x 64 product (hi + lo) per segment
            constexpr uint32_t id1 = idesc_f16(128, 64);
            const uint32_t g_lo = sw128_desc_lo(sm_addr + OFF_GHI);
            const int nseg = 2 * nb;
#pragma unroll 1
            for (int si = 0; si < nseg; ++si) {
                const int g = si & 1, b = si >> 1, d = si % 3;                // A1 slot, batch, D1 slot
                if (g == 0) { PYFFT_TC_TRACE(12, b) }
                bool live = mbar_wait_hot(bar0 + 8 * (A1_FULL + g), (uint32_t)b & 1u, abort_flag);
                if (si >= 3)
                    live = live && mbar_wait_hot(bar0 + 8 * (D1_FREE + d), ((uint32_t)(si / 3) & 1u) ^ 1u, abort_flag);
                if (!live) break;
                tc_fence_after();
                if (g == 0) { PYFFT_TC_TRACE(13, b) }
                const uint32_t a1 = sw128_desc_lo(sm_addr + OFF_A1 + g * A1_BYTES);
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks)
                        mma_f16_lo(tbase + TM_D1 + d * 64, a1 + ks * 2,
                                   g_lo + (h ? (OFF_GLO - OFF_GHI) / 16 : 0) + ks * 2, id1, (h | ks) ? 1u : 0u);
                mma_commit(bar0 + 8 * (S1_DONE + d));
                PYFFT_TC_TRACE(4 + g, b)
            }
        } else if (tid == 800) {
            // stage 2: 128 x 144 x 128 (hi + lo) per batch; the matrix comes from tensor memory, only
            // the data operand is read from shared memory
            constexpr uint32_t id2 = idesc_f16(128, N2);
#pragma unroll 1
            for (int bb = 0; bb < nb; ++bb) {
                const int s = bb & 1;
                PYFFT_TC_TRACE(14, bb)
                const uint32_t b2 = sw128_desc_lo(sm_addr + OFF_B2 + s * B2_BYTES);
                // the operand is ready long before the epilogue has drained the single D2 slot
                bool live = mbar_wait_hot(bar0 + 8 * (B2_FULL + s), (uint32_t)(bb >> 1) & 1u, abort_flag);
                if (bb >= 1) live = live && mbar_wait_hot(bar0 + 8 * D2_FREE, (uint32_t)(bb - 1) & 1u, abort_flag);
                if (!live) break;
                tc_fence_after();
                PYFFT_TC_TRACE(15, bb)
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int ks = 0; ks < 8; ++ks)
                        mma_f16_ts_lo(tbase + TM_D2, tbase + (h ? TM_F2LO : TM_F2HI) + ks * 8,
                                      b2 + ((ks >> 2) * (N2 * 128) + (ks & 3) * 32) / 16, id2, (h | ks) ? 1u : 0u);
                mma_commit(bar0 + 8 * S2_DONE);
                mma_commit(bar0 + 8 * (B2_FREE + s));
                PYFFT_TC_TRACE(6, bb)
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (tid == 0 && *abort_flag) atomicExch(p.status, 1);
    if (p.trace != nullptr && tid == 0) {
        unsigned long long t_;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
        p.trace[2048 + 4 * (blockIdx.y * gridDim.x + blockIdx.x) + 1] = (long long)t_;
        p.trace[2048 + 4 * (blockIdx.y * gridDim.x + blockIdx.x) + 3] = clock64();
    }
    if (warp == 0) tmem_free(tbase, 512);
}
#undef PYFFT_TC_ACC
#undef PYFFT_TC_TRACE

// ------------------------------------------------------------------------- //
// provisional statistics: mean (rounded to float) and a power-of-two scale that brings the
// rms of (x - mean) to ~1/8, from the first samples of the span
// ------------------------------------------------------------------------- //
// est[2c] += sum (x - x0), est[2c+1] += sum (x - x0)^2 with x0 = the first sample of the span; the
// last block of a channel to finish (ticket in prescale[c], zeroed by the host) turns the sums into
// est[2c] = mean (rounded to fp32), est[2c+1] = 0, prescale[c] = 2^k with rms * 2^k in [1/8, 1/4)
__global__ void __launch_bounds__(256) tc_prepare_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                         int64_t y_stride, RecMap map, int64_t j_begin, int64_t j_end,
                                                         double* __restrict__ est, double* __restrict__ prescale) {
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
    __shared__ double sh[2][8];
    if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = s0; sh[1][threadIdx.x >> 5] = s1; }
    __syncthreads();
    if (threadIdx.x != 0) return;
    s0 = s1 = 0.0;
#pragma unroll
    for (int w = 0; w < 8; ++w) { s0 += sh[0][w]; s1 += sh[1][w]; }
    atomicAdd(&est[2 * c], s0);
    atomicAdd(&est[2 * c + 1], s1);
    __threadfence();
    if (atomicAdd(&prescale[c], 1.0) != (double)(gridDim.x - 1)) return;     // not the last block of the channel
    __threadfence();
    const double n = (double)(j_end - j_begin);
    const double md = atomicAdd(&est[2 * c], 0.0) / n;                         // read through L2
    double var = atomicAdd(&est[2 * c + 1], 0.0) / n - md * md;
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

// after the main kernel, in one launch: acc += sum of the per-CTA partial spectra, lin += sum of the
// partial linear sums (bins < NLIN), stats[2c] += raw sum of the covered samples
__global__ void __launch_bounds__(256) tc_finish_kernel(const float* __restrict__ partial,
                                                        const float* __restrict__ lin_partial, int nparts,
                                                        const double* __restrict__ dev_sums,
                                                        const double* __restrict__ est, double count, int nch1,
                                                        double* __restrict__ acc, double* __restrict__ lin,
                                                        double* __restrict__ stats) {
    using namespace tcw;
    constexpr int PER_PAIR = 4 * NBIN, MAIN_BLOCKS = (PER_PAIR + 31) / 32;
    const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
    const int pair = blockIdx.y;
    const bool is_lin = blockIdx.x == MAIN_BLOCKS;
    if (is_lin && pair == 0 && threadIdx.x >= 32)
        for (int c = threadIdx.x - 32; c < nch1; c += 224) stats[2 * c] += dev_sums[c] + count * est[2 * c];
    // 32 consecutive outputs per block, the parts split over the 8 warps, combined through shared memory
    const int per = is_lin ? 4 * (L / 16) : PER_PAIR;
    int i = blockIdx.x * 32 + lane;
    if (is_lin) i = (lane >> 3) * (L / 16) + (lane & 7);         // row (re/im of U, V) and bin < NLIN = 8
    const float* src = is_lin ? lin_partial : partial;
    __shared__ double sh[8][33];
    double s = 0.0;
    if (i < per) {
        const float* p = src + (size_t)pair * nparts * per + i;
        int k = g;
        for (; k + 56 < nparts; k += 64) {
            float r[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) r[u] = p[(size_t)(k + 8 * u) * per];
#pragma unroll
            for (int u = 0; u < 8; ++u) s += (double)r[u];
        }
        for (; k < nparts; k += 8) s += (double)p[(size_t)k * per];
    }
    sh[g][lane] = s;
    __syncthreads();
    if (g == 0 && i < per) {
        double t = 0.0;
#pragma unroll
        for (int u = 0; u < 8; ++u) t += sh[u][lane];
        (is_lin ? lin : acc)[(size_t)pair * per + i] += t;
    }
}
static_assert(tcw::NLIN == 8, "tc_finish_kernel maps 32 lanes onto 4 rows x 8 bins");

// materialises the analysed samples j < n of every channel (the reflect-padded zone and the
// segments that reach into it), so that the loaders never evaluate the record map
__global__ void __launch_bounds__(256) tc_head_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                      int64_t y_stride, int pair0, RecMap map, int64_t n,
                                                      float* __restrict__ head) {
    const int c = blockIdx.y;                       // 0 = x, 1.. = y channels
    const float* p = c == 0 ? x : y + (int64_t)(c - 1) * y_stride;
    const int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (j < n) head[(int64_t)c * tcw::HEAD + j] = p[src_index(map, j)];
    (void)pair0;
}

// ------------------------------------------------------------------------- //
// host side
// ------------------------------------------------------------------------- //
static std::mutex g_tc_mutex;
static long long* g_tc_trace = nullptr;     // debug hook, see pyfft_debug_set_tc_trace
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
                    // plain rows: row (k2, co), halves (n2, ci) -> column n2 of the tensor-memory copy
                    const size_t o = ((size_t)(2 * k2 + co) * 128 + 2 * n2 + ci) * 2;
                    split(m[co][ci], OFF_F2T_HI + o, OFF_F2T_LO + o);
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
    const int64_t maxb = (seg_count + tcw::SEGS - 1) / tcw::SEGS;
    if (want > maxb) want = maxb;
    if (want < 1) want = 1;
    return (int)want;
}

}  // namespace pyfft

using namespace pyfft;

extern "C" {

// debug: device buffer of 16 long long per batch that CTA (0, 0) of the next launches fills with
// clock64() stamps of its pipeline events (NULL switches it off); used by tools/tc_trace.py
int pyfft_debug_set_tc_trace(void* buf) {
    g_tc_trace = static_cast<long long*>(buf);
    return PYFFT_OK;
}

int64_t pyfft_welch_tc_workspace_bytes(const pyfft_segmentation* seg, int32_t nch) {
    if (!seg || nch <= 0) return fail(PYFFT_EINVAL, "welch_tc_workspace_bytes: bad arguments");
    const int gx = tc_grid_x(seg->seg_count, nch);
    return (int64_t)nch * gx * 4 * (tcw::NBIN + tcw::L / 16) * 4 + 8 * (int64_t)(nch + 1) + 16 +
           (int64_t)(nch + 1) * tcw::HEAD * 4;
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
    PYFFT_CUDA_OK(cudaMemsetAsync(prescale, 0, sizeof(double) * (size_t)(nch + 1), st));
    int gx = (int)((j_end - j_begin + 2047) / 2048);
    if (gx > 64) gx = 64;
    tc_prepare_kernel<<<dim3(gx, nch + 1), 256, 0, st>>>(static_cast<const float*>(x), static_cast<const float*>(y),
                                                        y_stride, m, j_begin, j_end, est, prescale);
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
    const int nparts = gx;                        // one partial per CTA
    const int64_t main_bytes = (int64_t)nch * nparts * 4 * NBIN * 4;
    const int64_t lin_bytes = (int64_t)nch * nparts * 4 * (L / 16) * 4;
    const int64_t sums_bytes = (8 * (int64_t)(nch + 1) + 15) & ~(int64_t)15;
    const int64_t head_bytes = (int64_t)(nch + 1) * HEAD * 4;
    if (!workspace || workspace_bytes < main_bytes + lin_bytes + sums_bytes + head_bytes)
        return fail(PYFFT_EINVAL, "welch_accumulate_tc: workspace too small");
    // segments that start before map.reflect read a materialised copy of the head of the record
    int64_t head_len = 0;
    if (j_lo < m.reflect) {
        const int64_t last = (m.reflect - 1) / seg->hop;                        // last segment that starts in the zone
        const int64_t last_here = last < seg->seg_first + seg->seg_count - 1 ? last : seg->seg_first + seg->seg_count - 1;
        head_len = last_here * seg->hop + L;
        if (head_len > HEAD)
            return fail(PYFFT_EUNSUPPORTED, "the tensor-core path supports reflect padding up to %d samples", 2 * L - 1);
    }
    TcParams p;
    p.x = static_cast<const float*>(x);
    p.y = static_cast<const float*>(y);
    p.y_stride = y_stride;
    p.map = m;
    p.hop = seg->hop;
    p.seg_first = seg->seg_first;
    p.seg_count = seg->seg_count;
    // chunk: a multiple of the batch size so that only the very last batch can be ragged
    int64_t chunk = (seg->seg_count + gx - 1) / gx;
    chunk = (chunk + SEGS - 1) / SEGS * SEGS;
    p.chunk = chunk;
    p.window = static_cast<const float*>(window);
    p.est = est;
    p.prescale = prescale;
    p.consts = consts;
    float* head = reinterpret_cast<float*>(static_cast<char*>(workspace) + main_bytes + lin_bytes + sums_bytes);
    p.head = head;
    p.head_len = head_len;
    p.partial = static_cast<float*>(workspace);
    p.nparts = nparts;
    p.lin_partial = reinterpret_cast<float*>(static_cast<char*>(workspace) + main_bytes);
    p.dev_sums = reinterpret_cast<double*>(static_cast<char*>(workspace) + main_bytes + lin_bytes);
    p.status = status_flag;
    p.trace = g_tc_trace;
    // every CTA (also one that gets no segment) writes all of its partial sums and the NLIN bins of
    // the linear sums that tc_finish_kernel reads: only the raw-sum accumulators need clearing
    PYFFT_CUDA_OK(cudaMemsetAsync(p.dev_sums, 0, sizeof(double) * (size_t)(nch + 1), st));
    if (head_len > 0)
        tc_head_kernel<<<dim3((unsigned)((head_len + 255) / 256), nch + 1), 256, 0, st>>>(
            p.x, p.y, y_stride, 0, m, head_len, head);
    PYFFT_CUDA_OK(cudaFuncSetAttribute(welch_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    welch_tc_kernel<<<dim3(gx, nch), NTHREADS, SMEM_BYTES, st>>>(p);
    PYFFT_LAUNCH_OK("welch_tc_kernel");
    tc_finish_kernel<<<dim3((4 * NBIN + 31) / 32 + 1, nch), 256, 0, st>>>(
        p.partial, p.lin_partial, nparts, p.dev_sums, est, (double)seg->seg_count * (double)seg->hop, nch + 1, acc, lin,
        stats);
    PYFFT_LAUNCH_OK("tc_finish_kernel");
    return PYFFT_OK;
}

}  // extern "C"
