// Back-to-back tcgen05.mma throughput on one SM (B200): cycles per MMA for M = 128, K = 16 as a function of N,
// with the A operand in shared memory (SS) or in tensor memory (TS), issued by ONE thread or by TWO threads of
// different warps (bursts of NMMA instructions, one commit each).  Also the tcgen05.st rate.
// tools/tc_probe.cu commits after every product and therefore measures latency, not throughput.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/mma_rate tools/mma_rate.cu
#include <cstdio>
#include <cstdlib>

#include "../pyfft_b200/csrc/tc_core.cuh"

using namespace pyfft::tc;

constexpr int NMMA = 256;

// NT issuing threads (lane 0 of warps 0..NT-1), each into its own accumulator
template <int N, bool TS, int NT>
__global__ void __launch_bounds__(128) rate_kernel(long long* cycles, int* status) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) unsigned long long mbar[2];
    __shared__ uint32_t tmem_base_slot;
    unsigned char* sa = smem + ((1024u - ((uint32_t)__cvta_generic_to_shared(smem) & 1023u)) & 1023u);   // A: 128 rows x 64 halves
    unsigned char* sb = sa + 128 * 128;                                                                  // B: 256 rows x 64 halves
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (128 * 128 + 256 * 128) / 4; i += 128) reinterpret_cast<uint32_t*>(sa)[i] = 0x3c003c00u;   // 1.0h
    const uint32_t mb = (uint32_t)__cvta_generic_to_shared(&mbar[0]);
    if (tid == 0) { mbar_init(mb, 1); mbar_init(mb + 8, 1); fence_mbar_init(); }
    if (warp == 0) tmem_alloc((uint32_t)__cvta_generic_to_shared(&tmem_base_slot), 512);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tbase = tmem_base_slot;
    for (int c = 0; c < 32; c += 8)                       // A in tensor memory: columns 480..511
        tmem_st8(tbase + ((uint32_t)(warp * 32) << 16) + 480 + c, 0x3c003c00u, 0x3c003c00u, 0x3c003c00u, 0x3c003c00u,
                 0x3c003c00u, 0x3c003c00u, 0x3c003c00u, 0x3c003c00u);
    tmem_st_wait();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t a_lo = sw128_desc_lo((uint32_t)__cvta_generic_to_shared(sa));
    const uint32_t b_lo = sw128_desc_lo((uint32_t)__cvta_generic_to_shared(sb));
    constexpr uint32_t idesc = idesc_f16(128, N);
    long long best = 1ll << 60;
    bool ok = true;
    for (int rep = 0; rep < 5 && ok; ++rep) {
        __syncthreads();
        const long long t0 = clock64();
        if ((tid & 31) == 0 && warp < NT) {
            const uint32_t d = tbase + warp * 160;
#pragma unroll 4
            for (int i = 0; i < NMMA; ++i) {
                const int ks = i & 3;
                if (TS) mma_f16_ts_lo(d, tbase + 480 + ks * 8, b_lo + ks * 2, idesc, (i & 7) ? 1u : 0u);
                else mma_f16_lo(d, a_lo + ks * 2, b_lo + ks * 2, idesc, (i & 7) ? 1u : 0u);
            }
            mma_commit(mb + 8 * warp);
        }
        ok = mbar_wait(mb, (uint32_t)rep & 1u) && (NT < 2 || mbar_wait(mb + 8, (uint32_t)rep & 1u));
        if (tid == 0) {
            const long long dt = clock64() - t0;
            if (dt < best) best = dt;
        }
    }
    if (tid == 0) { cycles[0] = best; if (!ok) status[0] = 1; }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_free(tbase, 512);
}

// The kernel's own mix: thread 0 issues N = 64 bursts of 8 (stage 1), threads 1 and 2 N = 80 bursts of 16
// (stage 2, alternating segments), each burst committed and waited for by its issuer -- NSEG segments.
// Nominal tensor time per segment: 8 x 32 + 16 x 40 = 896 clocks.
__global__ void __launch_bounds__(128) mix_kernel(long long* cycles, int* status, int nseg, int wait_each) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) unsigned long long mbar[3];
    __shared__ uint32_t tmem_base_slot;
    unsigned char* sa = smem + ((1024u - ((uint32_t)__cvta_generic_to_shared(smem) & 1023u)) & 1023u);
    unsigned char* sb = sa + 128 * 128;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (128 * 128 + 256 * 128) / 4; i += 128) reinterpret_cast<uint32_t*>(sa)[i] = 0x3c003c00u;
    const uint32_t mb = (uint32_t)__cvta_generic_to_shared(&mbar[0]);
    if (tid == 0) { for (int i = 0; i < 3; ++i) mbar_init(mb + 8 * i, 1); fence_mbar_init(); }
    if (warp == 0) tmem_alloc((uint32_t)__cvta_generic_to_shared(&tmem_base_slot), 512);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tbase = tmem_base_slot;
    for (int c = 0; c < 32; c += 8)
        tmem_st8(tbase + ((uint32_t)(warp * 32) << 16) + 480 + c, 0x3c003c00u, 0x3c003c00u, 0x3c003c00u, 0x3c003c00u,
                 0x3c003c00u, 0x3c003c00u, 0x3c003c00u, 0x3c003c00u);
    tmem_st_wait();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t b_lo = sw128_desc_lo((uint32_t)__cvta_generic_to_shared(sb));
    __syncthreads();
    const long long t0 = clock64();
    bool ok = true;
    if ((tid & 31) == 0 && warp < 3) {
        const uint32_t d = tbase + warp * 128;
        const int per = warp == 0 ? 8 : 16;
        const uint32_t idesc = warp == 0 ? idesc_f16(128, 64) : idesc_f16(128, 80);
        const int mine = warp == 0 ? nseg : nseg / 2;
        for (int s = 0; s < mine && ok; ++s) {
            for (int i = 0; i < per; ++i)
                mma_f16_ts_lo(d, tbase + 480 + (i & 3) * 8, b_lo + (i & 3) * 2, idesc, i ? 1u : 0u);
            mma_commit(mb + 8 * warp);
            if (wait_each) ok = mbar_wait(mb + 8 * warp, (uint32_t)s & 1u);
        }
        if (!wait_each) for (int s = 0; s < mine && ok; ++s) ok = mbar_wait(mb + 8 * warp, (uint32_t)s & 1u);
    }
    __syncthreads();
    if (tid == 0) cycles[0] = clock64() - t0;
    if (!ok) status[0] = 1;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_free(tbase, 512);
}

static void run_mix(int wait_each) {
    long long* dC; int* dS;
    cudaMalloc(&dC, 8); cudaMalloc(&dS, 4); cudaMemset(dS, 0, 4);
    const int smem = 128 * 128 + 256 * 128 + 1024, nseg = 64;
    cudaFuncSetAttribute(mix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    mix_kernel<<<1, 128, smem>>>(dC, dS, nseg, wait_each);
    mix_kernel<<<1, 128, smem>>>(dC, dS, nseg, wait_each);
    cudaError_t e = cudaDeviceSynchronize();
    long long c = 0; int s = 0;
    cudaMemcpy(&c, dC, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&s, dS, 4, cudaMemcpyDeviceToHost);
    printf("kernel mix (1 x N64 x 8 + 2 x N80 x 16 per segment), %s: %.0f clk per segment (nominal 896; status %d, %s)\n",
           wait_each ? "issuer waits for its own commit before the next burst" : "no waits between bursts", (double)c / nseg, s,
           cudaGetErrorString(e));
    cudaFree(dC); cudaFree(dS);
}

// NW warps store x16 columns each into tensor memory: bytes per clock
template <int NW>
__global__ void __launch_bounds__(NW * 32) st_rate_kernel(long long* cycles) {
    __shared__ uint32_t tmem_base_slot;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (warp == 0) tmem_alloc((uint32_t)__cvta_generic_to_shared(&tmem_base_slot), 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t t = tmem_base_slot + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 32;
    uint32_t v = tid;
    __syncthreads();
    const long long t0 = clock64();
    for (int r = 0; r < 256; ++r) {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1};"
                     ::"r"(t + (r & 1) * 16), "r"(v) : "memory");
        if ((r & 3) == 3) tmem_st_wait();
    }
    tmem_st_wait();
    __syncthreads();
    if (tid == 0) cycles[0] = clock64() - t0;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_free(tmem_base_slot, 512);
}

template <int N, bool TS, int NT>
static void run() {
    long long* dC; int* dS;
    cudaMalloc(&dC, 8); cudaMalloc(&dS, 4); cudaMemset(dS, 0, 4);
    const int smem = 128 * 128 + 256 * 128 + 1024;
    cudaFuncSetAttribute(rate_kernel<N, TS, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    rate_kernel<N, TS, NT><<<1, 128, smem>>>(dC, dS);
    cudaError_t e = cudaDeviceSynchronize();
    long long c = 0; int s = 0;
    cudaMemcpy(&c, dC, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&s, dS, 4, cudaMemcpyDeviceToHost);
    printf("%d issuing thread(s), %s M=128 N=%3d K=16: %7.1f clk per MMA overall (%d x %d MMAs in %lld clk, status %d, %s)\n", NT,
           TS ? "TS" : "SS", N, (double)c / (NT * NMMA), NT, NMMA, c, s, cudaGetErrorString(e));
    cudaFree(dC); cudaFree(dS);
}

template <int NW>
static void run_st() {
    long long* dC; cudaMalloc(&dC, 8);
    st_rate_kernel<NW><<<1, NW * 32>>>(dC);
    cudaError_t e = cudaDeviceSynchronize();
    long long c = 0; cudaMemcpy(&c, dC, 8, cudaMemcpyDeviceToHost);
    printf("tcgen05.st x16, %2d warps: %.1f B/clk (%lld clk for %d x16 stores per warp, %s)\n", NW,
           256.0 * NW * 32 * 64 / (double)c, c, 256, cudaGetErrorString(e));
    cudaFree(dC);
}

int main() {
    run<16, false, 1>(); run<64, false, 1>(); run<80, false, 1>(); run<96, false, 1>(); run<128, false, 1>();
    run<144, false, 1>(); run<160, false, 1>(); run<256, false, 1>();
    run<16, true, 1>(); run<64, true, 1>(); run<80, true, 1>(); run<96, true, 1>(); run<128, true, 1>();
    run<144, true, 1>(); run<160, true, 1>(); run<256, true, 1>();
    run<16, true, 2>(); run<64, true, 2>(); run<80, true, 2>(); run<128, true, 2>(); run<144, true, 2>();
    run<64, false, 2>(); run<128, false, 2>();
    run_mix(1);
    run_st<4>(); run_st<8>(); run_st<16>();
    return 0;
}
