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
//             D2[(k2, re|im), (ch, k1)] = F2[(k2, c'), (n2, c)] * B2[(ch, k1), (n2, c)]
//                                                                         M = 128, N = 80, K = 128
// Lane (k2, c) of D2 holds one component of X and Y of 33 bins of a segment, so the cross-spectral
// sums are thread-local (one shuffle for the mixed Im term) and every lane owns distinct bins:
// 132 accumulator registers, no redundancy.  Rows k1 = 33..63 of stage 1 are never computed: bins
// with f mod 64 > 32 come from lanes k2 >= 32 by Hermitian symmetry X[4096 - f] = conj X[f].
// Precision: operands are fp16 with fp32 accumulation.  The DFT matrices are split in two fp16
// terms (hi + lo, 22 significant bits, two MMAs per product); the data is scaled by a power of
// two to unit-ish rms and rounded ONCE to fp16 at each stage -- a zero-mean relative error of
// 2^-12 per element that averages out over the segments when the record carries broadband noise
// (the host checks the dynamic range of the result, see pyfft_welch_dynamic_range).
//
// Data path (round 2): BOTH operands that stream are kept out of the shared-memory read port.
//   * the raw samples arrive by TMA (cp.async.bulk, one copy of hop samples per channel and segment: a
//     bulk copy costs its issuing thread ~170 clocks whatever its size, tools/bulk_rate.cu) in a ring of
//     16 K samples per channel; with 50 % overlap every sample crosses L2 -> SM once and is converted
//     from shared memory by both segments that contain it;
//   * the stage-1 data operand A1op lives in TENSOR memory (tcgen05.st by the converter warps, thread =
//     row (ch, n2), 32 columns of packed half2), so stage 1 reads only the 8 KB matrix tile from shared
//     memory and runs at the tensor pipe's N/2 = 32 clocks per MMA instead of the 48 clocks that the
//     operand fetch of a 128 x 16 shared-memory A tile costs (tools/mma_rate.cu);
//   * stage 2 as before: the matrix F2 (hi, lo) in tensor memory, the twiddled data from shared memory.
// Everything is per segment and double-buffered (A1, D1, B2, D2 x 2): no role waits for a slot that its
// consumer is still draining.  One MMA-issuing thread sustains one tcgen05.mma per ~51 clocks whatever
// its shape (tools/mma_rate.cu), so three threads issue: stage 1, stage 2 of even and of odd segments.
//
// Execution: one persistent CTA of 20 warps per SM, warp-specialised, all hand-shakes through mbarriers:
//   warps 0-3    epilogue    D2 (TMEM) -> register accumulators, packed f32x2 FMAs         (176 regs)
//   warps 4-7    twiddle     D1 (TMEM) -> twiddle (factors in registers) -> fp16 -> B2     (120 regs)
//   warps 8-15   converters  ring (smem) -> (x - mean) * window * scale -> fp16 -> A1 (TMEM); warp
//                            (q, h): TMEM lane quarter q = rows (ch, n2), n1 half h; the 32 window
//                            values a thread needs stay in registers                         (72 regs)
//   warp 16      TMA producer (one thread)      warp 17      stage-1 MMA issue
//   warps 18, 19 stage-2 MMA issue for even / odd segments                                   (40 regs)
#include <cuda_fp16.h>

#include <map>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "tc_core.cuh"

namespace pyfft {

namespace tcw {
constexpr int L = 4096, NBIN = L / 2 + 1;
constexpr int UNIT = 512;                    // samples per unit: a converter thread reads 8 samples (n1) of a unit
constexpr int NU = 32;                       // ring depth in units; the ring is filled in BLOCKS of one hop (m units, m | 8)
constexpr int RING_CH_FLOATS = NU * (UNIT + 4);   // one channel's ring: NU / m blocks of m * 512 + 4 floats (worst case m = 1);
                                             // + 16 bytes per block: the copy starts at the 16-byte boundary below the block
constexpr int RING_BYTES = 2 * RING_CH_FLOATS * 4;   // x then y: 132096
constexpr int YOFF = 36;                     // D2 columns of a segment: X at +0 (33 bins), Y at +36 (33 bins); rows of B2 alike
constexpr int NCOL = 80;                     // stage-2 MMA N (a multiple of 16 >= 69), D2 columns / B2 rows per segment
constexpr int NLIN = 8;                      // bins with linear sums (mean correction)
constexpr int NTHREADS = 640;
constexpr int HEAD = 3 * L;                  // samples per channel of the materialised reflect zone
constexpr int HEADP = HEAD + 4;              // pitch of a channel in that buffer (+ the alignment shift)
// constant block in global memory: [G hi][G lo][twiddles][F2 hi, plain rows][F2 lo, plain rows]
constexpr int OFF_GHI = 0, OFF_GLO = 8192, OFF_TW = 16384;
constexpr int CONST_SMEM_BYTES = 16384 + 33 * 64 * 8;        // 33280 (offset of the F2 rows)
constexpr int OFF_F2T_HI = CONST_SMEM_BYTES, OFF_F2T_LO = OFF_F2T_HI + 128 * 64 * 4;
constexpr int CONST_BYTES = OFF_F2T_LO + 128 * 64 * 4;       // 98816
// shared memory (1024-aligned): G hi | G lo | B2 x 2 | ring
constexpr int OFF_B2 = 16384;
constexpr int B2_ATOM = NCOL * 128;                          // bytes of one K atom (64 halves) of all rows: 10240
constexpr int B2_BYTES = 2 * B2_ATOM;                        // K = 128: 20480
constexpr int OFF_RING = OFF_B2 + 2 * B2_BYTES;              // 57344
constexpr int SMEM_BYTES = OFF_RING + RING_BYTES + 1024;     // + alignment slack = 190464
static_assert(B2_ATOM % 1024 == 0 && OFF_RING % 16 == 0 && (RING_CH_FLOATS * 4) % 16 == 0, "operand alignment");
// tensor memory columns (480 of 512)
constexpr int TM_F2HI = 0, TM_F2LO = 64;                     // stage-2 A operand: 128 lanes x 64 columns of half2
constexpr int TM_A1 = 128;                                   // 2 slots x 32 columns of half2 (stage-1 A operand)
constexpr int TM_D1 = 192;                                   // 2 slots x 64 columns
constexpr int TM_D2 = 320;                                   // 2 slots x NCOL columns
// mbarriers
enum { RAW_FULL = 0, A1_FULL = NU, S1_DONE = NU + 2, D1_FREE = NU + 4, B2_FULL = NU + 6, S2_DONE = NU + 8,
       D2_FREE = NU + 10, NBARS = NU + 12 };
// registers per thread after reallocation; the launch allocation is 96 x 640 (ptxas -v must say 96: a
// smaller launch allocation makes setmaxnreg.inc wait for ever)
constexpr int REGS_LAUNCH = 96, REGS_EPI = 176, REGS_TWD = 120, REGS_CNV = 72, REGS_MISC = 40;
static_assert(128 * REGS_EPI + 128 * REGS_TWD + 256 * REGS_CNV + 128 * REGS_MISC <= NTHREADS * REGS_LAUNCH,
              "register reallocation exceeds the launch allocation");
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
    const float* head;        // [nch+1][HEADP]: analysed samples j < head_len at [shift + j] (reflect zone)
    int64_t head_len;         // blocks that start before head_len are copied from head instead of x / y
    float* partial;           // [nch][nparts][4][NBIN]
    int nparts;
    float* lin_partial;       // [nch][nparts][4][L/16]  (bins < NLIN written)
    double* dev_sums;         // [nch+1]
    int* status;
    long long* trace;         // debug: per-CTA start / end stamps and the time line of CTA (0, 0), or NULL
};

// programmatic dependent launch (pyfft_welch_tc_step): a kernel launched with the programmatic-serialization
// attribute may start while its predecessor in the stream still runs; pdl_wait() returns once the predecessor
// grid has completed and its writes are visible (at once when the kernel was launched the ordinary way),
// pdl_launch_dependents() lets the successor be scheduled as soon as resources free up
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

// float offset (0..3) of the sample at `p` inside its 16-byte line
__host__ __device__ __forceinline__ int align_shift(const float* p) { return (int)((reinterpret_cast<uintptr_t>(p) & 15) >> 2); }

// source sample of the analysed index j (outside the reflect zone) of a channel
__host__ __device__ __forceinline__ const float* direct_ptr(const float* base, const RecMap& m, int64_t j) {
    return base + (j - m.reflect + m.i0 - m.src_base);
}

__device__ __forceinline__ void bulk_load(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t mbar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst_smem), "l"(src), "r"(bytes), "r"(mbar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t mbar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
          "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
        : "memory");
}

// timing experiments only (results invalid): bit 0 converters neither wait for nor read the ring, bit 1 the
// twiddle role does not store, bit 2 the epilogue only loads, bit 3 no TMA copies (with bit 0), bit 4 only the
// hi halves of the DFT matrices (12 of 24 MMAs per segment), bit 5 / bit 6 the epilogue / twiddle role frees its
// D2 / D1 slot before reading it (what more slots could give at best)
#ifndef PYFFT_TC_ABLATE
#define PYFFT_TC_ABLATE 0
#endif

// debug time line (tools/tc_trace.py): event EV of segment S of CTA (0, 0), stamped by one lane of a role
#define PYFFT_TC_TRACE(EV, S) \
    if (trace_on && (S) < 128) p.trace[(S) * 16 + (EV)] = clock64();

__global__ void __launch_bounds__(tcw::NTHREADS, 1) welch_tc_kernel(const TcParams p) {
    using namespace tcw;
    using namespace tc;
    extern __shared__ unsigned char smem_raw[];
    __shared__ __align__(8) unsigned long long bars[NBARS];
    __shared__ uint32_t tmem_slot;
    __shared__ int abort_store;
    __shared__ int conv_prog[8];             // segments each converter warp has finished reading

    const uint32_t raw = (uint32_t)__cvta_generic_to_shared(smem_raw);
    const uint32_t pad = (1024u - (raw & 1023u)) & 1023u;
    unsigned char* sm = smem_raw + pad;
    const uint32_t sm_addr = raw + pad;
    const uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(&bars[0]);
    volatile int* abort_flag = &abort_store;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int pair = blockIdx.y;
    const bool trace_on = p.trace != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && lane == 0 &&
                          (warp == 0 || warp == 4 || warp == 8 || warp >= 16);
    if (p.trace != nullptr && tid == 0) {
        unsigned long long t_;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
        p.trace[2048 + 4 * (blockIdx.y * gridDim.x + blockIdx.x)] = (long long)t_;
        p.trace[2048 + 4 * (blockIdx.y * gridDim.x + blockIdx.x) + 2] = clock64();
    }

    pdl_launch_dependents();
    // ---- one-time setup: stage-1 matrix, barriers, tensor memory --------------------------- //
    {
        const uint4* src = reinterpret_cast<const uint4*>(p.consts);
        uint4* dst = reinterpret_cast<uint4*>(sm);
        for (int i = tid; i < 16384 / 16; i += NTHREADS) dst[i] = src[i];
        uint4* b2 = reinterpret_cast<uint4*>(sm + OFF_B2);                     // rows 33..35, 69..79 stay zero
        for (int i = tid; i < 2 * B2_BYTES / 16; i += NTHREADS) b2[i] = make_uint4(0, 0, 0, 0);
    }
    if (tid == 0) {
        abort_store = 0;
        for (int i = 0; i < NU; ++i) mbar_init(bar0 + 8 * (RAW_FULL + i), 1);
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar0 + 8 * (A1_FULL + i), 256);
            mbar_init(bar0 + 8 * (S1_DONE + i), 1);
            mbar_init(bar0 + 8 * (D1_FREE + i), 128);
            mbar_init(bar0 + 8 * (B2_FULL + i), 128);
            mbar_init(bar0 + 8 * (S2_DONE + i), 1);
            mbar_init(bar0 + 8 * (D2_FREE + i), 128);
        }
        fence_mbar_init();
    }
    if (tid < 8) conv_prog[tid] = 0;
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
    pdl_wait();            // est, prescale, head and the zeroed raw sums come from the kernels before this one

    const int64_t c0 = (int64_t)blockIdx.x * p.chunk;
    const int64_t c1 = (c0 + p.chunk < p.seg_count) ? c0 + p.chunk : p.seg_count;
    const int nseg = c1 > c0 ? (int)(c1 - c0) : 0;
    const int m = (int)(p.hop >> 9);                                           // units per hop: 1, 2, 4 or 8
    const int lm = 31 - __clz(m);

    if (warp < 4) {
        // ================= epilogue: D2 -> cross-spectral sums in registers =================== //
        setmaxnreg_inc<REGS_EPI>();
        const uint32_t tl = tbase + ((uint32_t)(warp * 32) << 16) + TM_D2;     // this warp's lane quarter
        // pairs of neighbouring bins k1 = 2i, 2i+1 in one 64-bit register pair: packed FMAs
        float2 axx[16], ayy[16], ayx[16], ayo[16];
        float a32[4] = {0.f, 0.f, 0.f, 0.f};
        float2 linx[NLIN / 2], liny[NLIN / 2];
#pragma unroll
        for (int i = 0; i < 16; ++i) axx[i] = ayy[i] = ayx[i] = ayo[i] = make_float2(0.f, 0.f);
#pragma unroll
        for (int i = 0; i < NLIN / 2; ++i) linx[i] = liny[i] = make_float2(0.f, 0.f);

#pragma unroll 1
        for (int s = 0; s < nseg; ++s) {
            const int slot = s & 1;
            if (!mbar_wait_hot(bar0 + 8 * (S2_DONE + slot), (uint32_t)(s >> 1) & 1u, abort_flag)) break;
            tc_fence_after();
            PYFFT_TC_TRACE(9, s)
            if (PYFFT_TC_ABLATE & 32) { tc_fence_before(); mbar_arrive(bar0 + 8 * (D2_FREE + slot)); }   // "free D2 at once"
            const uint32_t col = tl + slot * NCOL;
            float xa[8], ya[8], xb[8], yb[8];
            tmem_ld8(col, xa);
            tmem_ld8(col + YOFF, ya);
            tmem_ld_wait();
            tmem_ld8(col + 8, xb);
            tmem_ld8(col + YOFF + 8, yb);
#define PYFFT_TC_ACC4(BASE, XV, YV)                                                                      \
    if (!(PYFFT_TC_ABLATE & 4)) _Pragma("unroll") for (int i = 0; i < 4; ++i) {                                                      \
        const float2 x2 = make_float2(XV[2 * i], XV[2 * i + 1]), y2 = make_float2(YV[2 * i], YV[2 * i + 1]); \
        const float2 o2 = make_float2(__shfl_xor_sync(0xffffffffu, XV[2 * i], 1),                        \
                                      __shfl_xor_sync(0xffffffffu, XV[2 * i + 1], 1));                   \
        axx[BASE + i] = __ffma2_rn(x2, x2, axx[BASE + i]);                                               \
        ayy[BASE + i] = __ffma2_rn(y2, y2, ayy[BASE + i]);                                               \
        ayx[BASE + i] = __ffma2_rn(y2, x2, ayx[BASE + i]);                                               \
        ayo[BASE + i] = __ffma2_rn(y2, o2, ayo[BASE + i]);                                               \
    }
            PYFFT_TC_ACC4(0, xa, ya)
            if (warp == 0) {                                  // bins f < 8 live in lanes k2 = 0 of warp 0
#pragma unroll
                for (int i = 0; i < NLIN / 2; ++i) {
                    linx[i] = __fadd2_rn(linx[i], make_float2(xa[2 * i], xa[2 * i + 1]));
                    liny[i] = __fadd2_rn(liny[i], make_float2(ya[2 * i], ya[2 * i + 1]));
                }
            }
            tmem_ld_wait();
            tmem_ld8(col + 16, xa);
            tmem_ld8(col + YOFF + 16, ya);
            PYFFT_TC_ACC4(4, xb, yb)
            tmem_ld_wait();
            tmem_ld8(col + 24, xb);
            tmem_ld8(col + YOFF + 24, yb);
            PYFFT_TC_ACC4(8, xa, ya)
            tmem_ld_wait();
            float x32, y32, d0, d1;
            tmem_ld2(col + 32, x32, d0);
            tmem_ld2(col + YOFF + 32, y32, d1);
            PYFFT_TC_ACC4(12, xb, yb)
            tmem_ld_wait();
            tc_fence_before();                                // D2 is in registers now
            if (!(PYFFT_TC_ABLATE & 32)) mbar_arrive(bar0 + 8 * (D2_FREE + slot));
            PYFFT_TC_TRACE(10, s)
            {
                const float xo = __shfl_xor_sync(0xffffffffu, x32, 1);
                a32[0] = fmaf(x32, x32, a32[0]);
                a32[1] = fmaf(y32, y32, a32[1]);
                a32[2] = fmaf(y32, x32, a32[2]);
                a32[3] = fmaf(y32, xo, a32[3]);
            }
        }
#undef PYFFT_TC_ACC4

        // ---- write the CTA's partial sums ---------------------------------------------------- //
        const float sx = (float)p.prescale[0], sy = (float)p.prescale[1 + pair];
        const float i00 = 1.f / (sx * sx), i11 = 1.f / (sy * sy), i01 = 1.f / (sx * sy);
        const int k2 = tid >> 1, c = tid & 1;
        float* out = p.partial + ((size_t)pair * p.nparts + blockIdx.x) * 4 * NBIN;
#pragma unroll
        for (int k1 = 0; k1 <= 32; ++k1) {
            const float v0 = k1 == 32 ? a32[0] : ((k1 & 1) ? axx[k1 >> 1 & 15].y : axx[k1 >> 1 & 15].x);
            const float v1 = k1 == 32 ? a32[1] : ((k1 & 1) ? ayy[k1 >> 1 & 15].y : ayy[k1 >> 1 & 15].x);
            const float v2 = k1 == 32 ? a32[2] : ((k1 & 1) ? ayx[k1 >> 1 & 15].y : ayx[k1 >> 1 & 15].x);
            const float v3 = k1 == 32 ? a32[3] : ((k1 & 1) ? ayo[k1 >> 1 & 15].y : ayo[k1 >> 1 & 15].x);
            const float t0 = v0 + __shfl_xor_sync(0xffffffffu, v0, 1);
            const float t1 = v1 + __shfl_xor_sync(0xffffffffu, v1, 1);
            const float t2 = v2 + __shfl_xor_sync(0xffffffffu, v2, 1);
            const float o3 = __shfl_xor_sync(0xffffffffu, v3, 1);
            // even lane: v3 = sum Yr Xi, partner's = sum Yi Xr  ->  Im(Y conj X) = Yi Xr - Yr Xi
            float t3 = o3 - v3;
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
                lo[(0 + c) * (L / 16) + k1] = ((k1 & 1) ? linx[k1 >> 1].y : linx[k1 >> 1].x) / sx;   // rows 0/1: Re/Im sum U
                lo[(2 + c) * (L / 16) + k1] = ((k1 & 1) ? liny[k1 >> 1].y : liny[k1 >> 1].x) / sy;   // rows 2/3: Re/Im sum V
            }
        }
    } else if (warp < 8) {
        // ================= twiddle: D1 -> W_4096^(n2 k1) -> fp16 -> stage-2 operand ========== //
        // Thread = TMEM lane = row (ch, n2) of D1; its 33 twiddle factors stay in registers as pairs of
        // neighbouring k1, so a pair of outputs costs four packed f32x2 operations.
        setmaxnreg_inc<REGS_TWD>();
        const int ln = tid - 128;                                            // TMEM lane = row (ch, n2) of D1
        const int ch = ln >> 6, n2 = ln & 63;
        const uint32_t tl = tbase + ((uint32_t)((warp & 3) * 32) << 16) + TM_D1;
        float2 wc[16], ws[16];                                               // (cos, cos), (-sin, -sin) of k1 = 2i, 2i+1
        float2 w32;
        {
            const float2* __restrict__ tw = reinterpret_cast<const float2*>(p.consts + OFF_TW) + n2;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float2 t0 = tw[(2 * i) * 64], t1 = tw[(2 * i + 1) * 64];
                wc[i] = make_float2(t0.x, t1.x);
                ws[i] = make_float2(t0.y, t1.y);
            }
            w32 = tw[32 * 64];
        }
        // row r = ch*36 + k1 of the operand; 8-row groups are contiguous, so the row offset is r*128 and
        // only the 16-byte chunk index is XOR-ed with r & 7 = (k1 + 4 ch) & 7
        const uint32_t thread_off = (uint32_t)((n2 >> 5) * B2_ATOM + (n2 & 3) * 4 + ch * YOFF * 128);
        const uint32_t xbase = (uint32_t)((((n2 & 31) >> 2) ^ (4 * ch)) << 4);   // chunk ^ (k1 & 7) = xbase ^ ((k1 & 7) << 4)

#pragma unroll 1
        for (int s = 0; s < nseg; ++s) {
            const int slot = s & 1;
            // the stage-2 MMAs of segment s-2 have finished reading this B2 slot
            if (s >= 2 && !mbar_wait_hot(bar0 + 8 * (S2_DONE + slot), ((uint32_t)(s >> 1) & 1u) ^ 1u, abort_flag)) break;
            if (!mbar_wait_hot(bar0 + 8 * (S1_DONE + slot), (uint32_t)(s >> 1) & 1u, abort_flag)) break;
            tc_fence_after();
            PYFFT_TC_TRACE(5, s)
            if (PYFFT_TC_ABLATE & 64) { tc_fence_before(); mbar_arrive(bar0 + 8 * (D1_FREE + slot)); }   // "free D1 at once"
            const uint32_t ca = tl + slot * 64;
            unsigned char* bp = sm + OFF_B2 + slot * B2_BYTES + thread_off;
            float ra[8], ia[8], rb[8], ib[8];
            tmem_ld8(ca, ra);
            tmem_ld8(ca + 32, ia);                        // first value = Re A1[32]; Im A1[0] = 0
            tmem_ld_wait();
            const float re32 = ia[0];
            ia[0] = 0.f;
#define PYFFT_TC_TWD(J, RV, IV)                                                                          \
    _Pragma("unroll") for (int i = 0; i < 4; ++i) {                                                      \
        const float2 a2 = make_float2(RV[2 * i], RV[2 * i + 1]), b2 = make_float2(IV[2 * i], IV[2 * i + 1]); \
        const float2 t2 = __fmul2_rn(b2, ws[4 * (J) + i]);                                               \
        const float2 re = __ffma2_rn(a2, wc[4 * (J) + i], make_float2(-t2.x, -t2.y));                    \
        const float2 im = __ffma2_rn(a2, ws[4 * (J) + i], __fmul2_rn(b2, wc[4 * (J) + i]));              \
        const int k1 = 8 * (J) + 2 * i;                                                                  \
        if (!(PYFFT_TC_ABLATE & 2)) {                                                                    \
        *reinterpret_cast<uint32_t*>(bp + k1 * 128 + (xbase ^ (uint32_t)((k1 & 7) << 4))) = pack_h2(re.x, im.x);           \
        *reinterpret_cast<uint32_t*>(bp + (k1 + 1) * 128 + (xbase ^ (uint32_t)(((k1 + 1) & 7) << 4))) = pack_h2(re.y, im.y); \
        }                                                                                                \
    }
            tmem_ld8(ca + 8, rb);
            tmem_ld8(ca + 40, ib);
            PYFFT_TC_TWD(0, ra, ia)
            tmem_ld_wait();
            tmem_ld8(ca + 16, ra);
            tmem_ld8(ca + 48, ia);
            PYFFT_TC_TWD(1, rb, ib)
            tmem_ld_wait();
            tmem_ld8(ca + 24, rb);
            tmem_ld8(ca + 56, ib);
            PYFFT_TC_TWD(2, ra, ia)
            tmem_ld_wait();
            tc_fence_before();                            // the last D1 columns are in registers: slot free
            if (!(PYFFT_TC_ABLATE & 64)) mbar_arrive(bar0 + 8 * (D1_FREE + slot));
            PYFFT_TC_TWD(3, rb, ib)
            *reinterpret_cast<uint32_t*>(bp + 32 * 128 + xbase) = pack_h2(re32 * w32.x, re32 * w32.y);
            fence_async_smem();
            mbar_arrive(bar0 + 8 * (B2_FULL + slot));
            PYFFT_TC_TRACE(6, s)
        }
#undef PYFFT_TC_TWD
    } else if (warp < 16) {
        // ================= converters: ring -> detrend -> window -> fp16 -> stage-1 operand (TMEM) ==== //
        // Warp (q, h): TMEM lane quarter q (rows (ch, n2): ch = q >> 1, n2 = 32 (q & 1) + lane) and the half
        // n1 = 32 h .. 32 h + 31 of the row = units 4 h .. 4 h + 3 of the segment = TMEM columns 16 h .. 16 h + 15.
        setmaxnreg_dec<REGS_CNV>();
        const int cw = warp - 8, q = cw & 3, h = cw >> 2;
        const int ch = q >> 1, n2 = 32 * (q & 1) + lane;
        const float* __restrict__ xsrc = ch ? p.y + (int64_t)pair * p.y_stride : p.x;
        const int csel = ch ? 1 + pair : 0;
        const float sc = (float)p.prescale[csel];
        const float nm = -(float)p.est[2 * csel];
        const int dsh = align_shift(direct_ptr(xsrc, p.map, 0));             // the same for every block (hops are multiples of 4)
        float2 w2[16];                                                       // window * scale of n1 = 32 h + 2 i, + 1
#pragma unroll
        for (int i = 0; i < 16; ++i)
            w2[i] = make_float2(p.window[64 * (32 * h + 2 * i) + n2] * sc, p.window[64 * (32 * h + 2 * i + 1) + n2] * sc);
        const float2 nm2 = make_float2(nm, nm);
        const float* ring = reinterpret_cast<const float*>(sm + OFF_RING) + ch * RING_CH_FLOATS + dsh + n2;
        const uint32_t ta = tbase + ((uint32_t)(q * 32) << 16) + TM_A1 + 16 * h;
        const int bpitch = m * UNIT + 4;                                     // floats per block of the ring
        const int nblk = NU >> lm;                                           // blocks in the ring
        float2 sum2 = make_float2(0.f, 0.f);

#pragma unroll 1
        for (int s = 0; s < nseg; ++s) {
            const int slot = s & 1;
            PYFFT_TC_TRACE(0, s)
            // the stage-1 MMAs of segment s-2 have finished reading this A1 slot (long ago in steady state)
            if (s >= 2 && !mbar_wait_abortable<32>(bar0 + 8 * (S1_DONE + slot), ((uint32_t)(s >> 1) & 1u) ^ 1u, abort_flag)) break;
            tc_fence_after();
            // software pipeline over the four units: the loads of unit a + 1 are in flight while unit a is
            // converted.  m >= 4 (hop 2048 / 4096): the warp's four units sit in ONE block of the ring -- one
            // barrier wait and one base pointer per segment; otherwise per unit.
            bool live = true;
            int prev_blk = -1;
            float v[2][8];
            uint32_t pk[8];
            const float* ub = ring;
            if (m >= 4) {
                const int U0 = s * m + 4 * h;
                const int blk = U0 >> lm, bs = blk & (nblk - 1);
                if (!(PYFFT_TC_ABLATE & 1))
                    live = mbar_wait_abortable<32>(bar0 + 8 * (RAW_FULL + bs), (uint32_t)(blk >> (5 - lm)) & 1u, abort_flag);
                ub = ring + bs * bpitch + (U0 & (m - 1)) * UNIT;
            }
#define PYFFT_TC_LOAD(A, BUF)                                                                               \
    if (m >= 4) {                                                                                           \
        _Pragma("unroll") for (int b = 0; b < 8; ++b)                                                       \
            v[BUF][b] = (PYFFT_TC_ABLATE & 1) ? (float)(s + b) : ub[(A) * UNIT + 64 * b];                   \
    } else {                                                                                                \
        const int U = s * m + 4 * h + (A);                     /* unit index inside the CTA's span */       \
        const int blk = U >> lm, bs = blk & (nblk - 1);        /* its block and the block's ring slot */    \
        if (blk != prev_blk) {                                                                              \
            if (!mbar_wait_abortable<32>(bar0 + 8 * (RAW_FULL + bs), (uint32_t)(blk >> (5 - lm)) & 1u, abort_flag)) live = false; \
            prev_blk = blk;                                                                                 \
        }                                                                                                   \
        const float* up = ring + bs * bpitch + (U & (m - 1)) * UNIT;                                        \
        _Pragma("unroll") for (int b = 0; b < 8; ++b) v[BUF][b] = up[64 * b];                               \
    }
#define PYFFT_TC_CONV(A, BUF)                                                                               \
    {                                                                                                       \
        float2 t[4];                                                                                        \
        _Pragma("unroll") for (int b = 0; b < 4; ++b)                                                       \
            t[b] = __fadd2_rn(make_float2(v[BUF][2 * b], v[BUF][2 * b + 1]), nm2);                          \
        if (4 * h + (A) < m)                                   /* the part of the segment inside its hop */ \
            sum2 = __fadd2_rn(sum2, __fadd2_rn(__fadd2_rn(t[0], t[1]), __fadd2_rn(t[2], t[3])));            \
        _Pragma("unroll") for (int b = 0; b < 4; ++b) {                                                     \
            const float2 o = __fmul2_rn(t[b], w2[4 * (A) + b]);                                             \
            pk[4 * ((A) & 1) + b] = pack_h2(o.x, o.y);                                                      \
        }                                                                                                   \
    }
            PYFFT_TC_LOAD(0, 0)
            PYFFT_TC_LOAD(1, 1)
            PYFFT_TC_CONV(0, 0)
            PYFFT_TC_LOAD(2, 0)
            PYFFT_TC_CONV(1, 1)
            tmem_st8(ta + 32 * slot, pk[0], pk[1], pk[2], pk[3], pk[4], pk[5], pk[6], pk[7]);
            PYFFT_TC_LOAD(3, 1)
            PYFFT_TC_CONV(2, 0)
            PYFFT_TC_CONV(3, 1)
            tmem_st8(ta + 32 * slot + 8, pk[0], pk[1], pk[2], pk[3], pk[4], pk[5], pk[6], pk[7]);
#undef PYFFT_TC_LOAD
#undef PYFFT_TC_CONV
            if (!live) break;
            PYFFT_TC_TRACE(1, s)
            tmem_st_wait();
            tc_fence_before();
            mbar_arrive(bar0 + 8 * (A1_FULL + slot));
            PYFFT_TC_TRACE(2, s)
            __syncwarp();
            if (lane == 0) *reinterpret_cast<volatile int*>(&conv_prog[cw]) = s + 1;
        }
        double dsum = (double)sum2.x + (double)sum2.y;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) dsum += __shfl_xor_sync(0xffffffffu, dsum, o);
        // sum2 = sum of (x - provisional mean), unscaled: the window carries the power-of-two scale
        if (lane == 0 && (ch == 1 || pair == 0)) atomicAdd(&p.dev_sums[csel], dsum);
    } else {
        setmaxnreg_dec<REGS_MISC>();
        if (tid == 512) {
            // ================= TMA producer: one block (hop samples) of both channels per segment ====== //
            const float* __restrict__ ysrc = p.y + (int64_t)pair * p.y_stride;
            const int64_t jc = (p.seg_first + c0) * p.hop;                     // first analysed sample of the CTA
            const int nblocks = nseg > 0 ? nseg - 1 + (8 >> lm) : 0;           // one block = one hop = m units
            const int nblk = NU >> lm;                                         // blocks in the ring
            const uint32_t bpitchb = (uint32_t)(m * UNIT + 4) * 4;             // bytes between blocks of a channel's ring
            const int shx = align_shift(direct_ptr(p.x, p.map, 0));
            const int shy = align_shift(direct_ptr(ysrc, p.map, 0));
            // a copy starts at the 16-byte boundary at or below the block's first sample and ends at the one at or
            // above its last: it never leaves the 16-byte granules that hold valid samples (no read past the
            // end of the caller's allocation)
            const uint32_t bx = (uint32_t)(m * UNIT) * 4 + (shx ? 16u : 0u), by = (uint32_t)(m * UNIT) * 4 + (shy ? 16u : 0u);
            const float* hx = p.head;
            const float* hy = p.head + (int64_t)(1 + pair) * HEADP;
            int done_seg = 0;                                                  // segments all converter warps have read
            // running pointers: a lone thread runs ~4 clocks per dependent instruction
            const float* dx_ = direct_ptr(p.x, p.map, jc) - shx;
            const float* dy_ = direct_ptr(ysrc, p.map, jc) - shy;
            int64_t j = jc;
            const int tb = (8 >> lm) - 1;                                      // block B completes segment B - tb
#pragma unroll 1
            for (int B = 0; B < nblocks; ++B, j += p.hop, dx_ += p.hop, dy_ += p.hop) {
                if (B >= nblk) {
                    // block B - nblk is last read by segment B - nblk
                    const int need = B - nblk + 1 < nseg ? B - nblk + 1 : nseg;
                    bool live = true;
                    while (done_seg < need) {
                        const volatile int* cp = conv_prog;
                        int mn = cp[0];
#pragma unroll
                        for (int i = 1; i < 8; ++i) { const int v = cp[i]; mn = v < mn ? v : mn; }
                        done_seg = mn;
                        if (done_seg < need) {
                            if (*abort_flag) { live = false; break; }
                            __nanosleep(64);
                        }
                    }
                    if (!live) break;
                }
                const bool in_head = j < p.head_len;
                const float* sx_ = in_head ? hx + j : dx_;
                const float* sy_ = in_head ? hy + j : dy_;
                const int bs = B & (nblk - 1);
                const uint32_t bar = bar0 + 8 * (RAW_FULL + bs);
                const uint32_t dst = sm_addr + OFF_RING + bs * bpitchb;
                if (!(PYFFT_TC_ABLATE & 8)) {
                    mbar_expect_tx(bar, bx + by);
                    bulk_load(dst, sx_, bx, bar);
                    bulk_load(dst + RING_CH_FLOATS * 4, sy_, by, bar);
                }
                if (B >= tb) { PYFFT_TC_TRACE(11, B - tb) }
            }
        } else if (tid == 544) {
            // ================= stage 1: 128 x 64 x 64 (hi + lo) per segment, A from tensor memory == //
            constexpr uint32_t id1 = idesc_f16(128, 64);
            const uint32_t g_lo = sw128_desc_lo(sm_addr + OFF_GHI);
#pragma unroll 1
            for (int s = 0; s < nseg; ++s) {
                const int slot = s & 1;
                bool live = mbar_wait_hot(bar0 + 8 * (A1_FULL + slot), (uint32_t)(s >> 1) & 1u, abort_flag);
                if (s >= 2)
                    live = live && mbar_wait_hot(bar0 + 8 * (D1_FREE + slot), ((uint32_t)(s >> 1) & 1u) ^ 1u, abort_flag);
                if (!live) break;
                tc_fence_after();
                PYFFT_TC_TRACE(3, s)
#pragma unroll
                for (int hh = 0; hh < ((PYFFT_TC_ABLATE & 16) ? 1 : 2); ++hh)
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks)
                        mma_f16_ts_lo(tbase + TM_D1 + slot * 64, tbase + TM_A1 + slot * 32 + ks * 8,
                                      g_lo + (hh ? (OFF_GLO - OFF_GHI) / 16 : 0) + ks * 2, id1, (hh | ks) ? 1u : 0u);
                mma_commit(bar0 + 8 * (S1_DONE + slot));
                PYFFT_TC_TRACE(4, s)
            }
        } else if (tid == 576 || tid == 608) {
            // ================= stage 2: 128 x 80 x 128 (hi + lo) per segment; one thread per slot ===== //
            // Measured alternatives (profiles/r02b_tc_experiments.txt): ONE thread for both slots 0.264 ms (its
            // 16 x ~86 clocks of issue per segment become the period); the hi and the lo products of a segment
            // from two threads 0.254 ms; the issuing warps converged with elect.sync (back-to-back UTCHMMA in
            // SASS) 0.36-0.38 ms; waits parked in hardware instead of hot polls: no change; one polling lane per
            // warp (+ shuffle) instead of every thread polling: 0.299 ms.
            constexpr uint32_t id2 = idesc_f16(128, NCOL);
            const int slot = tid == 608 ? 1 : 0;
            const uint32_t b2 = sw128_desc_lo(sm_addr + OFF_B2 + slot * B2_BYTES);
#pragma unroll 1
            for (int s = slot; s < nseg; s += 2) {
                bool live = mbar_wait_hot(bar0 + 8 * (B2_FULL + slot), (uint32_t)(s >> 1) & 1u, abort_flag);
                if (s >= 2)
                    live = live && mbar_wait_hot(bar0 + 8 * (D2_FREE + slot), ((uint32_t)(s >> 1) & 1u) ^ 1u, abort_flag);
                if (!live) break;
                tc_fence_after();
                PYFFT_TC_TRACE(7, s)
#pragma unroll
                for (int hh = 0; hh < ((PYFFT_TC_ABLATE & 16) ? 1 : 2); ++hh)
#pragma unroll
                    for (int ks = 0; ks < 8; ++ks)
                        mma_f16_ts_lo(tbase + TM_D2 + slot * NCOL, tbase + (hh ? TM_F2LO : TM_F2HI) + ks * 8,
                                      b2 + ((ks >> 2) * B2_ATOM + (ks & 3) * 32) / 16, id2, (hh | ks) ? 1u : 0u);
                mma_commit(bar0 + 8 * (S2_DONE + slot));
                PYFFT_TC_TRACE(8, s)
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (tid == 0 && *abort_flag) atomicOr(p.status, 1);
    if (p.trace != nullptr && tid == 0) {
        unsigned long long t_;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
        p.trace[2048 + 4 * (blockIdx.y * gridDim.x + blockIdx.x) + 1] = (long long)t_;
        p.trace[2048 + 4 * (blockIdx.y * gridDim.x + blockIdx.x) + 3] = clock64();
    }
    if (warp == 0) tmem_free(tbase, 512);
}
#undef PYFFT_TC_TRACE

// ------------------------------------------------------------------------- //
// provisional statistics: mean (rounded to float) and a power-of-two scale that brings the
// rms of (x - mean) to ~1/8, from PREP_BLOCKS runs of PREP_RUN samples spread evenly over the span
// (a record whose level changes along the way is scaled by its overall rms, not by its head)
// ------------------------------------------------------------------------- //
constexpr int PREP_BLOCKS = 256, PREP_RUN = 256;     // one sample (and its 8 lagged partners) per thread: all loads independent
__host__ __device__ __forceinline__ int64_t prep_stride(int64_t span) {
    const int64_t st = span > PREP_RUN ? (span - PREP_RUN) / (PREP_BLOCKS - 1) : 0;
    return st < PREP_RUN ? PREP_RUN : st;
}
// est[2c] += sum (x - x0), est[2c+1] += sum (x - x0)^2 with x0 = the first sample of the span; the
// last block of a channel to finish (ticket in prescale[c], zeroed by the host) turns the sums into
// est[2c] = mean (rounded to fp32), est[2c+1] = 0, prescale[c] = 2^k with rms * 2^k in [1/8, 1/4).
// Segment-to-segment repetition: the fp16 roundings of the tensor-core path average out over the segments
// only if they differ from segment to segment.  A record that repeats EXACTLY after a few hops (a noiseless
// synthetic signal whose period divides c hops) rounds identically every c segments and nothing averages.
// prescale[nch1 + c] = min over lags l = 1..PREP_LAGS hops of  sum (x[j + l hop] - x[j])^2 / (2 sum (x - mean)^2)
// over the sampled runs: ~1 for noise, the noise fraction for tone + noise, 0 for an exact repetition (a noise
// fraction of 1e-6 already dithers the roundings; the host refuses the path below 1e-4).
// Stationarity: the 1/sqrt(K) averaging also needs the power to be spread over the segments.
// prescale[2 nch1 + c] = (sum_b v_b)^2 / (B sum_b v_b^2) with v_b the variance of run b: 1 for a stationary
// record, 1/256 when one run of the 256 carries all the power (the effective number of segments is that
// fraction of K); the host refuses the path below 1/2.
constexpr int PREP_LAGS = 8;
// block bx of nblk (= PREP_BLOCKS) of channel c
__device__ __forceinline__ void tc_prepare_block(const float* __restrict__ x, const float* __restrict__ y,
                                                 int64_t y_stride, const RecMap& map, int64_t j_begin, int64_t j_end,
                                                 int64_t hop, int nch1, double* __restrict__ est,
                                                 double* __restrict__ prescale, int bx, int nblk, int c) {
    const float* p = c == 0 ? x : y + (int64_t)(c - 1) * y_stride;
    const double x0 = (double)p[src_index(map, j_begin)];
    double s[2 + PREP_LAGS];
#pragma unroll
    for (int i = 0; i < 2 + PREP_LAGS; ++i) s[i] = 0.0;
    const int64_t stride = prep_stride(j_end - j_begin);
    const int64_t r0 = j_begin + (int64_t)bx * stride;
    const bool lags_ok = hop > 0 && (j_end - j_begin) >= 2 * PREP_LAGS * hop + PREP_RUN;   // room to look 8 hops forward or back
    for (int64_t j = r0 + threadIdx.x; j < r0 + PREP_RUN && j < j_end; j += 256) {
        const float v0 = p[src_index(map, j)];
        const double v = (double)v0 - x0;
        s[0] += v;
        s[1] += v * v;
        if (lags_ok) {
            // look forward, or backward near the end of the span
            const int64_t dir = j + PREP_LAGS * hop < j_end ? hop : -hop;
#pragma unroll
            for (int l = 1; l <= PREP_LAGS; ++l) {
                const double d = (double)p[src_index(map, j + l * dir)] - (double)v0;
                s[1 + l] += d * d;
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 2 + PREP_LAGS; ++i)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s[i] += __shfl_xor_sync(0xffffffffu, s[i], o);
    __shared__ double sh[2 + PREP_LAGS][8];
    if ((threadIdx.x & 31) == 0)
#pragma unroll
        for (int i = 0; i < 2 + PREP_LAGS; ++i) sh[i][threadIdx.x >> 5] = s[i];
    __syncthreads();
    if (threadIdx.x != 0) return;
    double* lagsum = prescale + 3 * nch1 + PREP_LAGS * c;
    double* runvar = prescale + (3 + PREP_LAGS) * nch1 + 2 * c;
    {
        double t0 = 0.0, t1 = 0.0;
        for (int w = 0; w < 8; ++w) { t0 += sh[0][w]; t1 += sh[1][w]; }
        int64_t nb = r0 + PREP_RUN < j_end ? PREP_RUN : j_end - r0;
        if (nb > 1) {
            const double mb = t0 / (double)nb;
            double vb = t1 / (double)nb - mb * mb;
            if (!(vb > 0.0)) vb = 0.0;
            atomicAdd(&runvar[0], vb);
            atomicAdd(&runvar[1], vb * vb);
        }
    }
#pragma unroll
    for (int i = 0; i < 2 + PREP_LAGS; ++i) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += sh[i][w];
        if (i < 2) atomicAdd(&est[2 * c + i], t);
        else atomicAdd(&lagsum[i - 2], t);
    }
    __threadfence();
    if (atomicAdd(&prescale[c], 1.0) != (double)(nblk - 1)) return;     // not the last block of the channel
    __threadfence();
    // samples the runs covered / runs with at least two samples, in closed form (this thread is the critical
    // path of the step: the main kernel waits for it); block b covers [b stride, b stride + PREP_RUN) of the span
    const int64_t span = j_end - j_begin;
    int64_t nfull = span >= PREP_RUN ? (span - PREP_RUN) / stride + 1 : 0;
    if (nfull > PREP_BLOCKS) nfull = PREP_BLOCKS;
    int64_t cnt = nfull * PREP_RUN;
    if (nfull < PREP_BLOCKS && nfull * stride < span) cnt += span - nfull * stride;
    int64_t runs = span > 1 ? (span - 1 + stride - 1) / stride : 0;
    if (runs > PREP_BLOCKS) runs = PREP_BLOCKS;
    const double n = (double)cnt;
    // the sums, read through L2 with independent loads
    double sums[4 + PREP_LAGS];
    sums[0] = __ldcg(&est[2 * c]);
    sums[1] = __ldcg(&est[2 * c + 1]);
    sums[2] = __ldcg(&runvar[0]);
    sums[3] = __ldcg(&runvar[1]);
#pragma unroll
    for (int l = 0; l < PREP_LAGS; ++l) sums[4 + l] = __ldcg(&lagsum[l]);
    const double md = sums[0] / n;
    double var = sums[1] / n - md * md;
    if (!(var > 0.0)) var = 0.0;
    double rep = 1.0;
    if (lags_ok && var > 0.0)
#pragma unroll
        for (int l = 0; l < PREP_LAGS; ++l) {
            const double r = sums[4 + l] / (2.0 * var * n);
            rep = r < rep ? r : rep;
        }
    prescale[nch1 + c] = rep;
    prescale[2 * nch1 + c] = (sums[3] > 0.0 && runs > 0) ? sums[2] * sums[2] / ((double)runs * sums[3]) : 1.0;
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

__global__ void __launch_bounds__(256) tc_prepare_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                         int64_t y_stride, RecMap map, int64_t j_begin, int64_t j_end,
                                                         int64_t hop, int nch1, double* __restrict__ est,
                                                         double* __restrict__ prescale) {
    tc_prepare_block(x, y, y_stride, map, j_begin, j_end, hop, nch1, est, prescale, (int)blockIdx.x, (int)gridDim.x,
                     (int)blockIdx.y);             // channel 0 = x, 1.. = y channels
}

// after the main kernel, in one launch: acc += sum of the per-CTA partial spectra, lin += sum of the
// partial linear sums (bins < NLIN), stats[2c] += raw sum of the covered samples
__global__ void __launch_bounds__(256) tc_finish_kernel(const float* __restrict__ partial,
                                                        const float* __restrict__ lin_partial, int nparts,
                                                        const double* __restrict__ dev_sums,
                                                        const double* __restrict__ est, double count, int nch1,
                                                        double* __restrict__ acc, double* __restrict__ lin,
                                                        double* __restrict__ stats, int* __restrict__ status) {
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
        // an fp16 operand out of range (a record far louder than the sampled statistics said) shows as Inf / NaN
        if (!isfinite(t)) atomicOr(status, 2);
    }
}
static_assert(tcw::NLIN == 8, "tc_finish_kernel maps 32 lanes onto 4 rows x 8 bins");

// materialises the analysed samples j < n of every channel (the reflect-padded zone, whole blocks), so that
// the TMA producer copies plain 16-byte-aligned runs: sample j of channel c sits at head[c][shift_c + j]
// with shift_c = the offset of the channel's direct samples inside their 16-byte line (the converters use
// one shift per channel for every block)
__global__ void __launch_bounds__(256) tc_head_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                      int64_t y_stride, RecMap map, int64_t n,
                                                      float* __restrict__ head) {
    const int c = blockIdx.y;                       // 0 = x, 1.. = y channels
    const float* p = c == 0 ? x : y + (int64_t)(c - 1) * y_stride;
    const int sh = align_shift(direct_ptr(p, map, 0));
    const int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (j < n) head[(int64_t)c * tcw::HEADP + sh + j] = p[src_index(map, j)];
}

// ------------------------------------------------------------------------- //
// The whole fused step in three launches (pyfft_welch_tc_step): tc_pre_kernel -> welch_tc_kernel ->
// tc_post_kernel.  A step over resident records is launch-bound around the main kernel: every boundary
// between two small kernels costs ~2 us of a 0.26 ms step, and the sharded step had three of them around
// its exchange.
// ------------------------------------------------------------------------- //
constexpr int TAIL_BLOCKS = 4;

// grid.x = PREP_BLOCKS (provisional statistics) + head_blocks (reflect-padded head of the record) + TAIL_BLOCKS
// (raw sums of the analysed samples outside the segments' own span: tail[2c] += sum x, tail[2c+1] += sum j x)
__global__ void __launch_bounds__(256) tc_pre_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                     int64_t y_stride, RecMap map, int64_t j_begin, int64_t j_end,
                                                     int64_t hop, int nch1, double* __restrict__ est,
                                                     double* __restrict__ prescale, int64_t head_len,
                                                     float* __restrict__ head, int head_blocks, int64_t ta0,
                                                     int64_t ta1, int64_t tb0, int64_t tb1, double* __restrict__ tail) {
    pdl_launch_dependents();
    const int c = blockIdx.y;
    int bx = blockIdx.x;
    if (bx < PREP_BLOCKS) {
        tc_prepare_block(x, y, y_stride, map, j_begin, j_end, hop, nch1, est, prescale, bx, PREP_BLOCKS, c);
        return;
    }
    bx -= PREP_BLOCKS;
    const float* p = c == 0 ? x : y + (int64_t)(c - 1) * y_stride;
    if (bx < head_blocks) {
        const int sh = align_shift(direct_ptr(p, map, 0));
        const int64_t j = (int64_t)bx * 256 + threadIdx.x;
        if (j < head_len) head[(int64_t)c * tcw::HEADP + sh + j] = p[src_index(map, j)];
        return;
    }
    bx -= head_blocks;
    if (ta1 <= ta0 && tb1 <= tb0) return;
    double s0 = 0.0, s1 = 0.0;
    for (int64_t j = ta0 + (int64_t)bx * 256 + threadIdx.x; j < ta1; j += TAIL_BLOCKS * 256) {
        const double v = (double)p[src_index(map, j)];
        s0 += v;
        s1 += v * (double)j;
    }
    for (int64_t j = tb0 + (int64_t)bx * 256 + threadIdx.x; j < tb1; j += TAIL_BLOCKS * 256) {
        const double v = (double)p[src_index(map, j)];
        s0 += v;
        s1 += v * (double)j;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s0 += __shfl_xor_sync(0xffffffffu, s0, o);
        s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    }
    __shared__ double sh2[2][8];
    if ((threadIdx.x & 31) == 0) { sh2[0][threadIdx.x >> 5] = s0; sh2[1][threadIdx.x >> 5] = s1; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int i = 0; i < 8; ++i) { a += sh2[0][i]; b += sh2[1][i]; }
        atomicAdd(&tail[2 * c], a);
        atomicAdd(&tail[2 * c + 1], b);
    }
}

// One work item = 16 bins of one channel pair (129 items per pair: every SM takes one):
//   A. the sums over the main kernel's per-CTA partials (4 rows x 16 bins; the item of bins 0..15 also the linear
//      sums of bins < NLIN and the raw sample sums),
//   B. sharded (world > 1): the shard's sums are moved to the common reference 0 (welch_rereference_kernel),
//      WRITTEN into slot [rank] of every peer's staging buffer over NVLink, published with one release store per
//      peer, and -- once the item's flags of all ranks have arrived -- summed from the local staging buffer in rank
//      order (bit-identical on every GPU).  Flags are per item: no CTA waits for another CTA of its own grid.
//   C. state (acc | lin | stats) is written and the 32 bins are finalised (welch_finalize_kernel with the mean
//      correction).
// Peer buffer of one rank: stage[2][world][items][POST_SLOT] values of 16 bytes | uint32 flags[world][items] |
// uint32 epochs[items] | uint32 pending[items]; zero-filled once.  A value travels as two 8-byte words (32 data bits
// | epoch), valid by themselves on arrival (the protocol NCCL calls LL); PYFFT_POST_FENCE=1 selects plain stores +
// system fence + one flag per item and rank instead.  A rank can be at most one epoch ahead of a peer (it needs
// the peer's words of epoch e to finish e, which the peer sends after it has read epoch e-1), hence two copies.
constexpr int POST_BINS = 16;                     // bins per work item
constexpr int POST_SLOT = 4 * POST_BINS + 4 * tcw::NLIN + 8;        // 64 sums + 32 linear sums + 2 raw sums, padded
constexpr int POST_NBB = (tcw::NBIN + POST_BINS - 1) / POST_BINS;   // items per pair: 129
constexpr int POST_THREADS = 512, POST_WARPS = POST_THREADS / 32;
struct TcPostParams {
    const float* partial;
    const float* lin_partial;
    int nparts;
    const double* dev_sums;
    const double* tail;
    const double* est;
    double count, K_local, K_total, nana, scale_over_k;
    int nch, onesided, F;
    const double* W;
    double* acc;
    double* lin;
    double* stats;
    double* Pxx;
    double* Pyy;
    double* Pxy;
    double* Cxy;
    double* Cxy2;
    double* phi;
    int* status;
    const int64_t* peer_base;
    int rank, world;
    int mode;                 // 0 = this step, exchange waited for; 1 = deferred exchange; 2 = flush
    int ll;                   // exchange protocol: 1 = flagged 8-byte words (default), 0 = plain stores + fence + flag
    long long* trace;         // debug: globaltimer stamps of CTA 0 (pyfft_debug_set_post_trace), or NULL
};

// mode 0: reduce -> [exchange: push, wait, sum] -> finalize, all for THIS step (what pyfft_welch_tc_step returns
//         is complete when the kernel ends; every step waits for the slowest rank);
// mode 1: deferred exchange of a stream of steps: an item first collects and finalises the sums that it pushed in
//         the PREVIOUS deferred step (they arrived a whole main kernel ago: no rank waits for another unless it is
//         a full step ahead), then reduces and pushes this step's; outputs and state are the previous step's;
// mode 2: flush -- collect and finalise what is pending, push nothing.
// per item on the device: epochs[item] (last epoch pushed) and pending[item] (pushed, not yet collected).
#define PYFFT_POST_STAMP(I)                                                              \
    if (p.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {                     \
        unsigned long long t_;                                                           \
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                           \
        p.trace[I] = (long long)t_;                                                      \
    }
__global__ void __launch_bounds__(POST_THREADS) tc_post_kernel(const TcPostParams p) {
    using namespace tcw;
    constexpr int PER_PAIR = 4 * NBIN, LROW = L / 16, NB = POST_BINS;
    const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
    const int lb = lane & (NB - 1), half = lane / NB;       // bin of the item and which half of the parts
    const int items = p.nch * POST_NBB;
    __shared__ double sh[POST_WARPS][5][33];
    __shared__ double sa[4][NB], sl[4][NLIN], sraw[2];
    __shared__ uint32_t s_epoch, s_pending;
    __shared__ int s_ok;
    const size_t stage_doubles = (size_t)4 * p.world * items * POST_SLOT;     // 16 bytes per value (two flagged words)
    uint32_t* const my_words = p.world > 1 ? reinterpret_cast<uint32_t*>(reinterpret_cast<double*>(p.peer_base[p.rank]) + stage_doubles)
                                           : nullptr;      // flags[world][items] | epochs[items] | pending[items]
    PYFFT_POST_STAMP(0)
    bool waited = false;                                     // for the main kernel (programmatic dependent launch)
    for (int item = blockIdx.x; item < items; item += gridDim.x) {
        const int pair = item / POST_NBB, bb = item - pair * POST_NBB;
        const int f = bb * NB + lb;
        const int nval = bb == 0 ? 4 * NB + 4 * NLIN + 2 : 4 * NB;     // the item's sums (+ linear sums + 2 raw sums)
        if (threadIdx.x == 0) {
            s_ok = 1;
            s_epoch = p.world > 1 ? my_words[(size_t)p.world * items + item] : 0u;
            s_pending = p.world > 1 ? my_words[(size_t)(p.world + 1) * items + item] : 0u;
        }
        __syncthreads();
        const uint32_t epoch_prev = s_epoch;
        // what this launch does with the item: collect + finalise what the previous deferred step pushed / reduce
        // this step's partials / push them / collect and finalise them
        const bool do_old = p.world > 1 && p.mode != 0 && s_pending != 0u;
        const bool do_reduce = p.mode != 2;
        const bool do_push = p.world > 1 && p.mode != 2;
        const bool do_now = p.mode == 0 || p.world == 1;    // finalise this step's sums inside this launch
        // this step's partials: the two halves of a warp take the even / odd parts of every 16th pair of parts, 5 parts
        // x 4 rows (+ 10 linear sums in the item of bins 0..15) per thread, all independent, issued in one go and left
        // in registers while the previous step is finalised
        float rp[5][4], rl[10];
#define PYFFT_POST_LOADS                                                                                         \
        {                                                                                                            \
            if (!waited) {                                                                                           \
                pdl_wait();                                                                                          \
                waited = true;                                                                                       \
            }                                                                                                        \
            PYFFT_POST_STAMP(1)                                                                                      \
            const float* pp = p.partial + (size_t)pair * p.nparts * PER_PAIR + f;                                    \
            const float* lp = p.lin_partial + (size_t)pair * p.nparts * 4 * LROW + (lane >> 3) * LROW + (lane & 7);  \
            const bool main_ok = do_reduce && f < NBIN, lin_ok = do_reduce && bb == 0;                               \
            _Pragma("unroll") for (int u = 0; u < 5; ++u) {                                                          \
                const int k = 2 * (g + POST_WARPS * u) + half;                                                       \
                _Pragma("unroll") for (int q = 0; q < 4; ++q)                                                        \
                    rp[u][q] = (main_ok && k < p.nparts) ? pp[(size_t)k * PER_PAIR + q * NBIN] : 0.f;                \
            }                                                                                                        \
            _Pragma("unroll") for (int u = 0; u < 10; ++u) {                                                         \
                const int k = g + POST_WARPS * u;                                                                    \
                rl[u] = (lin_ok && k < p.nparts) ? lp[(size_t)k * 4 * LROW] : 0.f;                                   \
            }                                                                                                        \
        }
#pragma unroll 1
        for (int round = do_old ? 0 : 1; round < 2; ++round) {
            // round 0: the previous deferred step (sums about the reference 0); round 1: this step
            double ex = 0.0, ey = 0.0, Kc = p.K_total;
            uint32_t epoch = epoch_prev;
            if (round == 1) {
                if (!do_old) PYFFT_POST_LOADS
                PYFFT_POST_STAMP(3)
                if (!do_reduce) break;
                // ---- A: reduce the partials of the main kernel ---------------------------------- //
                double s[4] = {0.0, 0.0, 0.0, 0.0}, t = 0.0;
#pragma unroll
                for (int u = 0; u < 5; ++u)
#pragma unroll
                    for (int q = 0; q < 4; ++q) s[q] += (double)rp[u][q];
#pragma unroll
                for (int u = 0; u < 10; ++u) t += (double)rl[u];
                for (int k = 10 * POST_WARPS + 2 * g + half; k < p.nparts; k += 2 * POST_WARPS)      // > 160 parts: never on 148 SMs
                    if (f < NBIN)
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            s[q] += (double)p.partial[((size_t)pair * p.nparts + k) * PER_PAIR + q * NBIN + f];
                for (int k = 10 * POST_WARPS + g; k < p.nparts; k += POST_WARPS)
                    if (bb == 0)
                        t += (double)p.lin_partial[((size_t)pair * p.nparts + k) * 4 * LROW + (lane >> 3) * LROW + (lane & 7)];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    s[q] += __shfl_xor_sync(0xffffffffu, s[q], NB);
                    if (half == 0) sh[g][q][lb] = s[q];
                }
                sh[g][4][lane] = t;
                __syncthreads();
                if (g < 5) {
                    double v = 0.0;
#pragma unroll
                    for (int u = 0; u < POST_WARPS; ++u) v += sh[u][g][g < 4 ? lb : lane];
                    if (g < 4) {
                        if (half == 0) sa[g][lb] = v;
                        // an fp16 operand out of range (a record far louder than the sampled statistics said) shows as Inf / NaN
                        if (f < NBIN && !isfinite(v)) atomicOr(p.status, 2);
                    } else if (bb == 0) {
                        sl[lane >> 3][lane & 7] = v;
                    }
                } else if (threadIdx.x == 160 && bb == 0) {
                    sraw[0] = p.tail[0] + p.dev_sums[0] + p.count * p.est[0];
                    sraw[1] = p.tail[2 * (1 + pair)] + p.dev_sums[1 + pair] + p.count * p.est[2 * (1 + pair)];
                }
                __syncthreads();
                PYFFT_POST_STAMP(4)
                ex = p.est[0];
                ey = p.est[2 * (1 + pair)];
                Kc = p.K_local;
                if (do_push) {
                    // ---- B1: to the common reference 0 (welch_rereference_kernel), then into every peer --- //
                    if (bb == 0 && threadIdx.x < NLIN) {
                        const int b = threadIdx.x;
                        const double dx = -ex, dy = -ey;
                        const double wr = p.W[2 * b], wi = p.W[2 * b + 1], w2 = wr * wr + wi * wi;
                        const double sur = sl[0][b], sui = sl[1][b], svr = sl[2][b], svi = sl[3][b];
                        sa[0][b] += -2.0 * dx * (wr * sur + wi * sui) + Kc * dx * dx * w2;
                        sa[1][b] += -2.0 * dy * (wr * svr + wi * svi) + Kc * dy * dy * w2;
                        sa[2][b] += -dx * (wr * svr + wi * svi) - dy * (wr * sur + wi * sui) + Kc * dx * dy * w2;
                        sa[3][b] += -dx * (wr * svi - wi * svr) - dy * (wi * sur - wr * sui);
                        sl[0][b] = sur - Kc * dx * wr;
                        sl[1][b] = sui - Kc * dx * wi;
                        sl[2][b] = svr - Kc * dy * wr;
                        sl[3][b] = svi - Kc * dy * wi;
                    }
                    __syncthreads();
                    epoch = epoch_prev + 1u;
                    const size_t slot0 = ((size_t)(epoch & 1u) * p.world * items + item) * POST_SLOT;   // + r * items * POST_SLOT
                    if ((int)threadIdx.x < nval) {
                        const int i = threadIdx.x;
                        const double v = i < 4 * NB ? sa[i / NB][i % NB]
                                                    : (i < 4 * NB + 4 * NLIN ? sl[(i - 4 * NB) >> 3][(i - 4 * NB) & 7]
                                                                             : sraw[i - 4 * NB - 4 * NLIN]);
                        if (p.ll) {
                            // every double travels as two 8-byte words (32 data bits | epoch): an 8-byte store arrives
                            // whole, so a word that carries the epoch is valid by itself -- no fence, no separate flag,
                            // and the latency of the exchange is one NVLink crossing
                            const unsigned long long bits = (unsigned long long)__double_as_longlong(v);
                            const unsigned long long w0 = (bits & 0xffffffffull) | ((unsigned long long)epoch << 32);
                            const unsigned long long w1 = (bits >> 32) | ((unsigned long long)epoch << 32);
                            for (int r = 0; r < p.world; ++r) {
                                unsigned long long* d = reinterpret_cast<unsigned long long*>(p.peer_base[r]) +
                                                        2 * (slot0 + (size_t)p.rank * items * POST_SLOT + i);
                                asm volatile("st.relaxed.sys.global.v2.u64 [%0], {%1, %2};" ::"l"(d), "l"(w0), "l"(w1) : "memory");
                            }
                        } else {
                            for (int r = 0; r < p.world; ++r)
                                (reinterpret_cast<double*>(p.peer_base[r]) + slot0 + (size_t)p.rank * items * POST_SLOT)[i] = v;
                        }
                    }
                    if (!p.ll) {
                        __threadfence_system();
                        __syncthreads();
                        if ((int)threadIdx.x < p.world) {
                            uint32_t* theirs = reinterpret_cast<uint32_t*>(reinterpret_cast<double*>(p.peer_base[threadIdx.x]) + stage_doubles);
                            st_release_sys(theirs + (size_t)p.rank * items + item, epoch);
                        }
                    }
                    if (threadIdx.x == 0) {
                        my_words[(size_t)p.world * items + item] = epoch;
                        my_words[(size_t)(p.world + 1) * items + item] = do_now ? 0u : 1u;
                    }
                    PYFFT_POST_STAMP(5)
                    if (!do_now) break;                       // collected by the next deferred step or the flush
                    ex = 0.0;
                    ey = 0.0;
                    Kc = p.K_total;
                    __syncthreads();
                }
            }
            if (p.world > 1) {
                // ---- B2: the item's values of all ranks, summed from the local staging copy in rank order ---- //
                const size_t slot0 = ((size_t)(epoch & 1u) * p.world * items + item) * POST_SLOT;
                if (!p.ll) {
                    if ((int)threadIdx.x < p.world) {
                        bool ok = false;
                        for (uint32_t it = 0; it < (1u << 24); ++it) {
                            if ((int32_t)(ld_acquire_sys(my_words + (size_t)threadIdx.x * items + item) - epoch) >= 0) { ok = true; break; }
                            __nanosleep(20);
                        }
                        if (!ok) { s_ok = 0; atomicOr(p.status, 4); }
                    }
                    __syncthreads();
                }
                if ((int)threadIdx.x < nval) {
                    const int i = threadIdx.x;
                    double s = 0.0;
                    if (p.ll) {
                        const unsigned long long* d0 = reinterpret_cast<const unsigned long long*>(p.peer_base[p.rank]) + 2 * (slot0 + i);
                        const size_t rstride = (size_t)2 * items * POST_SLOT;
                        for (int r0 = 0; r0 < p.world && s_ok; r0 += 8) {           // eight ranks' words in flight
                            unsigned long long w0[8], w1[8];
#pragma unroll
                            for (int q = 0; q < 8; ++q)
                                if (r0 + q < p.world)
                                    asm volatile("ld.relaxed.sys.global.v2.u64 {%0, %1}, [%2];"
                                                 : "=l"(w0[q]), "=l"(w1[q]) : "l"(d0 + (size_t)(r0 + q) * rstride) : "memory");
#pragma unroll
                            for (int q = 0; q < 8; ++q) {
                                if (r0 + q >= p.world) break;
                                uint32_t it = 0;
                                while ((uint32_t)(w0[q] >> 32) != epoch || (uint32_t)(w1[q] >> 32) != epoch) {
                                    if (++it > (1u << 24)) { s_ok = 0; atomicOr(p.status, 4); break; }
                                    __nanosleep(20);
                                    asm volatile("ld.relaxed.sys.global.v2.u64 {%0, %1}, [%2];"
                                                 : "=l"(w0[q]), "=l"(w1[q]) : "l"(d0 + (size_t)(r0 + q) * rstride) : "memory");
                                }
                                // rank order: the same sum on every GPU
                                s += __longlong_as_double((long long)((w0[q] & 0xffffffffull) | (w1[q] << 32)));
                            }
                        }
                    } else if (s_ok) {
                        const double* st = reinterpret_cast<const double*>(p.peer_base[p.rank]) + slot0 + i;
                        for (int r = 0; r < p.world; ++r) s += ld_relaxed_sys(st + (size_t)r * items * POST_SLOT);
                    }
                    if (!s_ok) s = __longlong_as_double(0x7ff8000000000000ll);    // a peer never arrived: fail loudly
                    if (i < 4 * NB) sa[i / NB][i % NB] = s;
                    else if (i < 4 * NB + 4 * NLIN) sl[(i - 4 * NB) >> 3][(i - 4 * NB) & 7] = s;
                    else sraw[i - 4 * NB - 4 * NLIN] = s;
                }
                if (round == 0 && threadIdx.x == 0 && !do_push) my_words[(size_t)(p.world + 1) * items + item] = 0u;
                __syncthreads();
            }
            PYFFT_POST_STAMP(round == 0 ? 2 : 6)
            // what the previous step pushed does not depend on this step's main kernel: it was collected before
            // waiting for it, and is finalised while this step's partials are on their way
            if (round == 0) PYFFT_POST_LOADS
            // ---- C: state and the finalised spectra (4 x NB threads: one group of outputs each) ---------- //
            if (threadIdx.x < 4 * NB) {
                const int q = threadIdx.x / NB, b = threadIdx.x % NB, fb = bb * NB + b;
                if (fb < NBIN) p.acc[((size_t)pair * 4 + q) * NBIN + fb] = sa[q][b];
            }
            if (bb == 0 && threadIdx.x >= 128) {
                for (int i = threadIdx.x - 128; i < 4 * LROW; i += POST_THREADS - 128)
                    p.lin[(size_t)pair * 4 * LROW + i] = (i & (LROW - 1)) < NLIN ? sl[i / LROW][i & (LROW - 1)] : 0.0;
                if (threadIdx.x == 128) {
                    if (pair == 0) { p.stats[0] = sraw[0]; p.stats[1] = 0.0; }
                    p.stats[2 * (1 + pair)] = sraw[1];
                    p.stats[2 * (1 + pair) + 1] = 0.0;
                }
            }
            if (threadIdx.x < 4 * NB && bb * NB + (int)(threadIdx.x % NB) < NBIN) {
                const int part = threadIdx.x / NB, b = threadIdx.x % NB;     // part 0: Pxx, Pyy, Pxy   1: Cxy   2: Cxy2   3: phi
                const int fb = bb * NB + b;
                double a0 = sa[0][b], a1 = sa[1][b], a2 = sa[2][b], a3 = sa[3][b];
                if (bb == 0 && b < NLIN) {
                    const double dx = sraw[0] / p.nana - ex, dy = sraw[1] / p.nana - ey;
                    const double wr = p.W[2 * fb], wi = p.W[2 * fb + 1], w2 = wr * wr + wi * wi;
                    const double sur = sl[0][fb], sui = sl[1][fb], svr = sl[2][fb], svi = sl[3][fb];
                    a0 += -2.0 * dx * (wr * sur + wi * sui) + Kc * dx * dx * w2;
                    a1 += -2.0 * dy * (wr * svr + wi * svi) + Kc * dy * dy * w2;
                    a2 += -dx * (wr * svr + wi * svi) - dy * (wr * sur + wi * sui) + Kc * dx * dy * w2;
                    a3 += -dx * (wr * svi - wi * svr) - dy * (wi * sur - wr * sui);
                }
                const int nch = p.nch;
                // outputs of bin fb: one-sided o = fb (< F = L/2), x2 on bins 1..F-2 (fft_analysis.py:414-416);
                // two-sided, fftshift order (:423-427): o = (fb + L/2) mod L and, mirrored, o = L/2 - fb with Pxy conjugated
                for (int side = 0; side < 2; ++side) {
                    int o;
                    double wgt = 1.0;
                    bool conj = false;
                    if (p.onesided) {
                        if (side == 1 || fb >= p.F) break;
                        o = fb;
                        if (o >= 1 && o <= p.F - 2) wgt = 2.0;
                    } else if (side == 0) {
                        o = (fb + L / 2) & (L - 1);
                    } else {
                        if (fb == 0 || fb == L / 2) break;
                        o = L / 2 - fb;
                        conj = true;
                    }
                    const double s = p.scale_over_k * wgt;
                    const double pxx = a0 * s, pyy = a1 * s, pr = a2 * s, pi = (conj ? -a3 : a3) * s;
                    const size_t idx = (size_t)o * nch + pair;
                    const double den = fabs(pxx) * fabs(pyy);                 // fft_analysis.py:1672, 1677
                    if (part == 0) {
                        if (p.Pxx && pair == 0) { p.Pxx[2 * o] = pxx; p.Pxx[2 * o + 1] = 0.0; }
                        if (p.Pyy) { p.Pyy[2 * idx] = pyy; p.Pyy[2 * idx + 1] = 0.0; }
                        if (p.Pxy) { p.Pxy[2 * idx] = pr; p.Pxy[2 * idx + 1] = pi; }
                    } else if (part == 1) {
                        if (p.Cxy) { const double r = 1.0 / sqrt(den); p.Cxy[2 * idx] = pr * r; p.Cxy[2 * idx + 1] = pi * r; }
                    } else if (part == 2) {
                        if (p.Cxy2) { p.Cxy2[2 * idx] = (pr * pr + pi * pi) / den; p.Cxy2[2 * idx + 1] = 0.0 / den; }
                    } else {
                        if (p.phi) p.phi[idx] = atan2(pi, pr);                // fft_analysis.py:520
                    }
                }
            }
            __syncthreads();
        }
#undef PYFFT_POST_LOADS
        __syncthreads();
    }
    PYFFT_POST_STAMP(7)
}
#undef PYFFT_POST_STAMP

// ------------------------------------------------------------------------- //
// host side
// ------------------------------------------------------------------------- //
static std::mutex g_tc_mutex;
static long long* g_tc_trace = nullptr;     // debug hook, see pyfft_debug_set_tc_trace
static long long* g_post_trace = nullptr;   // debug hook, see pyfft_debug_set_post_trace
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
    if (want > seg_count) want = seg_count;
    if (want < 1) want = 1;
    return (int)want;
}

// argument checks and kernel parameters shared by pyfft_welch_accumulate_tc and pyfft_welch_tc_step
// (dev_sums is placed by the caller)
static int tc_setup(const void* x, const void* y, int64_t y_stride, int32_t nch, const pyfft_record_map* map,
                    const pyfft_segmentation* seg, const void* window, const double* est, const double* prescale,
                    void* workspace, int64_t workspace_bytes, int32_t* status_flag, TcParams* out, int* gx_out) {
    using namespace tcw;
    RecMap m;
    int rc = check_map(map, &m);
    if (rc) return rc;
    if (!seg || seg->nfft != L) return fail(PYFFT_EUNSUPPORTED, "the tensor-core path handles nfft = 4096 only");
    if (seg->hop != 512 && seg->hop != 1024 && seg->hop != 2048 && seg->hop != 4096)
        return fail(PYFFT_EUNSUPPORTED, "the tensor-core path needs a hop of 512, 1024, 2048 or 4096 samples (got %d)", seg->hop);
    if (!x || !y || !window || !est || !prescale || !status_flag || nch <= 0)
        return fail(PYFFT_EINVAL, "tensor-core Welch: NULL pointer or nch <= 0");
    if (seg->seg_count <= 0) return fail(PYFFT_EINVAL, "tensor-core Welch: no segments");
    const int64_t j_lo = seg->seg_first * (int64_t)seg->hop;
    const int64_t j_hi = (seg->seg_first + seg->seg_count - 1) * (int64_t)seg->hop + seg->nfft;
    if ((rc = check_range(m, j_lo, j_hi))) return rc;
    int status;
    const unsigned char* consts = tc_constants(&status);
    if (!consts) return status;
    const int gx = tc_grid_x(seg->seg_count, nch);
    const int nparts = gx;                        // one partial per CTA
    const int64_t main_bytes = (int64_t)nch * nparts * 4 * NBIN * 4;
    const int64_t lin_bytes = (int64_t)nch * nparts * 4 * (L / 16) * 4;
    const int64_t sums_bytes = (8 * (int64_t)(nch + 1) + 15) & ~(int64_t)15;
    const int64_t head_bytes = (int64_t)(nch + 1) * HEADP * 4;
    if (!workspace || workspace_bytes < main_bytes + lin_bytes + sums_bytes + head_bytes)
        return fail(PYFFT_EINVAL, "tensor-core Welch: workspace too small");
    // blocks (one hop of analysed samples) that reach into the reflect-padded zone are copied from a
    // materialised head of the record; every other block is a plain run of the source
    int64_t head_len = 0;
    if (j_lo < m.reflect) {
        head_len = (m.reflect + seg->hop - 1) / seg->hop * seg->hop;
        if (head_len > m.nsig) head_len = m.nsig;
        if (head_len > HEAD)
            return fail(PYFFT_EUNSUPPORTED, "the tensor-core path supports reflect padding up to %d samples", HEAD);
    }
    if (((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) | (uintptr_t)(y_stride * 4)) & 3) != 0)
        return fail(PYFFT_EINVAL, "tensor-core Welch: x / y must be 4-byte aligned");
    TcParams p;
    p.x = static_cast<const float*>(x);
    p.y = static_cast<const float*>(y);
    p.y_stride = y_stride;
    p.map = m;
    p.hop = seg->hop;
    p.seg_first = seg->seg_first;
    p.seg_count = seg->seg_count;
    p.chunk = (seg->seg_count + gx - 1) / gx;
    p.window = static_cast<const float*>(window);
    p.est = est;
    p.prescale = prescale;
    p.consts = consts;
    p.head = reinterpret_cast<float*>(static_cast<char*>(workspace) + main_bytes + lin_bytes + sums_bytes);
    p.head_len = head_len;
    p.partial = static_cast<float*>(workspace);
    p.nparts = nparts;
    p.lin_partial = reinterpret_cast<float*>(static_cast<char*>(workspace) + main_bytes);
    p.dev_sums = reinterpret_cast<double*>(static_cast<char*>(workspace) + main_bytes + lin_bytes);
    p.status = status_flag;
    p.trace = g_tc_trace;
    *out = p;
    *gx_out = gx;
    return PYFFT_OK;
}

static int tc_post_launch(const TcPostParams& q_in, bool pdl, cudaStream_t st) {
    int sms = device_sm_count();
    if (sms <= 0) sms = 148;
    TcPostParams q = q_in;
    q.ll = getenv("PYFFT_POST_FENCE") == nullptr ? 1 : 0;
    q.trace = g_post_trace;
    int64_t blocks = (int64_t)q.nch * POST_NBB;
    if (blocks > (int64_t)sms) blocks = (int64_t)sms;             // one CTA per SM, items in the same order on every GPU
    if (const char* cap = getenv("PYFFT_POST_BLOCKS")) {           // tests: several ranks driven on ONE GPU must all fit
        const int64_t c = atoll(cap);
        if (c >= 1 && c < blocks) blocks = c;
    }      // co-resident: a CTA of the sharded step may wait for its peers
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)blocks);
    cfg.blockDim = dim3(POST_THREADS);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    PYFFT_CUDA_OK(cudaLaunchKernelEx(&cfg, tc_post_kernel, q));
    return PYFFT_OK;
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

// debug: device buffer of 8 long long that CTA 0 of the next tc_post_kernel launches fills with globaltimer
// stamps (0 start, 1 predecessor done, 2 deferred sums collected, 3 deferred step finalised, 4 partials reduced,
// 5 pushed, 6 this step's sums collected, 7 end); NULL switches it off
int pyfft_debug_set_post_trace(void* buf) {
    g_post_trace = static_cast<long long*>(buf);
    return PYFFT_OK;
}

int64_t pyfft_welch_tc_workspace_bytes(const pyfft_segmentation* seg, int32_t nch) {
    if (!seg || nch <= 0) return fail(PYFFT_EINVAL, "welch_tc_workspace_bytes: bad arguments");
    const int gx = tc_grid_x(seg->seg_count, nch);
    return (int64_t)nch * gx * 4 * (tcw::NBIN + tcw::L / 16) * 4 + 8 * (int64_t)(nch + 1) + 16 +
           (int64_t)(nch + 1) * tcw::HEADP * 4;
}

int pyfft_welch_tc_prepare(const void* x, const void* y, int64_t y_stride, int32_t nch,
                           const pyfft_record_map* map, int64_t j_begin, int64_t j_end, int64_t hop, double* est,
                           double* prescale, void* stream) {
    RecMap m;
    int rc = check_map(map, &m);
    if (rc) return rc;
    if (!x || !y || !est || !prescale || nch <= 0) return fail(PYFFT_EINVAL, "welch_tc_prepare: bad arguments");
    if ((rc = check_range(m, j_begin, j_end))) return rc;
    if (j_end <= j_begin) return fail(PYFFT_EINVAL, "welch_tc_prepare: empty sample span");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    PYFFT_CUDA_OK(cudaMemsetAsync(est, 0, sizeof(double) * 2 * (size_t)(nch + 1), st));
    PYFFT_CUDA_OK(cudaMemsetAsync(prescale, 0, sizeof(double) * (size_t)(nch + 1) * (5 + PREP_LAGS), st));
    tc_prepare_kernel<<<dim3(PREP_BLOCKS, nch + 1), 256, 0, st>>>(static_cast<const float*>(x), static_cast<const float*>(y),
                                                        y_stride, m, j_begin, j_end, hop, nch + 1, est, prescale);
    PYFFT_LAUNCH_OK("tc_prepare_kernel");
    return PYFFT_OK;
}

int pyfft_welch_accumulate_tc(const void* x, const void* y, int64_t y_stride, int32_t nch,
                              const pyfft_record_map* map, const pyfft_segmentation* seg, const void* window,
                              const double* est, const double* prescale, double* acc, double* lin, double* stats,
                              void* workspace, int64_t workspace_bytes, int32_t* status_flag, void* stream) {
    using namespace tcw;
    if (!acc || !lin || !stats) return fail(PYFFT_EINVAL, "welch_accumulate_tc: NULL pointer");
    if (seg && seg->seg_count == 0 && seg->nfft == L) return PYFFT_OK;
    TcParams p;
    int gx;
    const int rc = tc_setup(x, y, y_stride, nch, map, seg, window, est, prescale, workspace, workspace_bytes, status_flag,
                            &p, &gx);
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int nparts = p.nparts;
    const int64_t head_len = p.head_len;
    float* head = const_cast<float*>(p.head);
    const RecMap m = p.map;
    // every CTA (also one that gets no segment) writes all of its partial sums and the NLIN bins of
    // the linear sums that tc_finish_kernel reads: only the raw-sum accumulators need clearing
    PYFFT_CUDA_OK(cudaMemsetAsync(p.dev_sums, 0, sizeof(double) * (size_t)(nch + 1), st));
    if (head_len > 0)
        tc_head_kernel<<<dim3((unsigned)((head_len + 255) / 256), nch + 1), 256, 0, st>>>(
            p.x, p.y, y_stride, m, head_len, head);
    PYFFT_CUDA_OK(cudaFuncSetAttribute(welch_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    welch_tc_kernel<<<dim3(gx, nch), NTHREADS, SMEM_BYTES, st>>>(p);
    PYFFT_LAUNCH_OK("welch_tc_kernel");
    tc_finish_kernel<<<dim3((4 * NBIN + 31) / 32 + 1, nch), 256, 0, st>>>(
        p.partial, p.lin_partial, nparts, p.dev_sums, est, (double)seg->seg_count * (double)seg->hop, nch + 1, acc, lin,
        stats, status_flag);
    PYFFT_LAUNCH_OK("tc_finish_kernel");
    return PYFFT_OK;
}

int64_t pyfft_welch_tc_step_peer_bytes(int32_t nch, int32_t world) {
    if (nch <= 0 || world <= 0) return fail(PYFFT_EINVAL, "welch_tc_step_peer_bytes: bad arguments");
    const int64_t items = (int64_t)nch * POST_NBB;
    return 4 * (int64_t)world * items * POST_SLOT * 8 + 4 * ((int64_t)world * items + 2 * items) + 64;
}

int pyfft_welch_tc_step(const void* x, const void* y, int64_t y_stride, int32_t nch, const pyfft_record_map* map,
                        const pyfft_segmentation* seg, const void* window, const double* wlow, int64_t tail_a0,
                        int64_t tail_a1, int64_t tail_b0, int64_t tail_b1, int64_t K_total, int64_t nana, double scale,
                        int32_t onesided, double* aux, double* state, double* Pxx, double* Pyy, double* Pxy,
                        double* Cxy, double* Cxy2, double* phi, void* workspace, int64_t workspace_bytes,
                        int32_t* status_flag, const int64_t* peer_base, int32_t rank, int32_t world, int32_t deferred,
                        void* stream) {
    using namespace tcw;
    if (!aux || !state || !wlow) return fail(PYFFT_EINVAL, "welch_tc_step: NULL pointer");
    if (K_total <= 0 || nana <= 0) return fail(PYFFT_EINVAL, "welch_tc_step: K_total and nana must be positive");
    if (world < 1 || world > 256 || rank < 0 || rank >= world || (world > 1 && !peer_base))
        return fail(PYFFT_EINVAL, "welch_tc_step: bad rank / world / peer_base");
    const int nch1 = nch + 1;
    double* est = aux;
    double* prescale = aux + 2 * nch1;
    double* tail = aux + 18 * nch1;
    double* dev_sums = aux + 20 * nch1;
    TcParams p;
    int gx;
    int rc = tc_setup(x, y, y_stride, nch, map, seg, window, est, prescale, workspace, workspace_bytes, status_flag, &p,
                      &gx);
    if (rc) return rc;
    p.dev_sums = dev_sums;
    const int64_t c_lo = seg->seg_first * (int64_t)seg->hop, c_hi = c_lo + seg->seg_count * (int64_t)seg->hop;
    if (tail_a1 > tail_a0 && (rc = check_range(p.map, tail_a0, tail_a1))) return rc;
    if (tail_b1 > tail_b0 && (rc = check_range(p.map, tail_b0, tail_b1))) return rc;
    if ((rc = check_range(p.map, c_lo, c_hi))) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    PYFFT_CUDA_OK(cudaMemsetAsync(aux, 0, sizeof(double) * 21 * (size_t)nch1, st));
    const int head_blocks = (int)((p.head_len + 255) / 256);
    tc_pre_kernel<<<dim3(PREP_BLOCKS + head_blocks + TAIL_BLOCKS, nch1), 256, 0, st>>>(
        p.x, p.y, y_stride, p.map, c_lo, c_hi, seg->hop, nch1, est, prescale, p.head_len, const_cast<float*>(p.head),
        head_blocks, tail_a0, tail_a1, tail_b0, tail_b1, tail);
    PYFFT_LAUNCH_OK("tc_pre_kernel");
    PYFFT_CUDA_OK(cudaFuncSetAttribute(welch_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    // the main kernel and the post kernel are launched programmatically dependent: their CTAs are scheduled as
    // the predecessor's drain away, the main kernel's set-up (constants, barriers, tensor memory) overlaps
    // tc_pre_kernel, and both wait (griddepcontrol.wait) before touching what the predecessor writes
    const bool pdl = getenv("PYFFT_NO_PDL") == nullptr;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(gx, nch);
        cfg.blockDim = dim3(NTHREADS);
        cfg.dynamicSmemBytes = SMEM_BYTES;
        cfg.stream = st;
        cfg.attrs = attr;
        cfg.numAttrs = pdl ? 1 : 0;
        PYFFT_CUDA_OK(cudaLaunchKernelEx(&cfg, welch_tc_kernel, p));
    }
    TcPostParams q;
    q.partial = p.partial;
    q.lin_partial = p.lin_partial;
    q.nparts = p.nparts;
    q.dev_sums = dev_sums;
    q.tail = tail;
    q.est = est;
    q.count = (double)seg->seg_count * (double)seg->hop;
    q.K_local = (double)seg->seg_count;
    q.K_total = (double)K_total;
    q.nana = (double)nana;
    q.scale_over_k = scale / (double)K_total;
    q.nch = nch;
    q.onesided = onesided;
    q.F = onesided ? L / 2 : L;
    q.W = wlow;
    q.acc = state;
    q.lin = state + (size_t)nch * 4 * NBIN;
    q.stats = q.lin + (size_t)nch * 4 * (L / 16);
    q.Pxx = Pxx;
    q.Pyy = Pyy;
    q.Pxy = Pxy;
    q.Cxy = Cxy;
    q.Cxy2 = Cxy2;
    q.phi = phi;
    q.status = status_flag;
    q.peer_base = peer_base;
    q.rank = rank;
    q.world = world;
    q.mode = (deferred && world > 1) ? 1 : 0;
    return tc_post_launch(q, pdl, st);
}

int pyfft_welch_tc_step_flush(int32_t nch, const double* wlow, int64_t K_total, int64_t nana, double scale,
                              int32_t onesided, double* state, double* Pxx, double* Pyy, double* Pxy, double* Cxy,
                              double* Cxy2, double* phi, int32_t* status_flag, const int64_t* peer_base, int32_t rank,
                              int32_t world, void* stream) {
    using namespace tcw;
    if (!state || !wlow || !status_flag || nch <= 0) return fail(PYFFT_EINVAL, "welch_tc_step_flush: NULL pointer or nch <= 0");
    if (K_total <= 0 || nana <= 0) return fail(PYFFT_EINVAL, "welch_tc_step_flush: K_total and nana must be positive");
    if (world < 2 || world > 256 || rank < 0 || rank >= world || !peer_base)
        return fail(PYFFT_EINVAL, "welch_tc_step_flush: bad rank / world / peer_base (only sharded steps are deferred)");
    TcPostParams q = {};
    q.K_total = (double)K_total;
    q.nana = (double)nana;
    q.scale_over_k = scale / (double)K_total;
    q.nch = nch;
    q.onesided = onesided;
    q.F = onesided ? L / 2 : L;
    q.W = wlow;
    q.acc = state;
    q.lin = state + (size_t)nch * 4 * NBIN;
    q.stats = q.lin + (size_t)nch * 4 * (L / 16);
    q.Pxx = Pxx;
    q.Pyy = Pyy;
    q.Pxy = Pxy;
    q.Cxy = Cxy;
    q.Cxy2 = Cxy2;
    q.phi = phi;
    q.status = status_flag;
    q.peer_base = peer_base;
    q.rank = rank;
    q.world = world;
    q.mode = 2;
    return tc_post_launch(q, false, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
