// Probe: tcgen05.mma with an MN-major ("transposed") A operand in shared memory.
// D[128][N] = A[128][K] * B[N][K]^T, B K-major SW128 as everywhere else; A stored with M as the
// contiguous direction in the canonical SWIZZLE_128B MN-major layout:
//   element (m, k) at  (m >> 6) * MBLK + (k >> 3) * KGRP + (k & 7) * 128
//                      + ((((m & 63) >> 3) ^ (k & 7)) << 4) + (m & 7) * 2
// i.e. a 128-byte line holds 64 consecutive m of one k; eight k make a 1024-byte atom.
// Tries both assignments of (MBLK, KGRP) to the descriptor's leading/stride byte offsets and
// reports which one the hardware uses.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/tc_probe_mn tools/tc_probe_mn.cu
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../pyfft_b200/csrc/tc_core.cuh"

using namespace pyfft::tc;

constexpr int MBLK = 8192, KGRP = 1024;       // bytes: K = 64 -> 8 k-groups of 1024 B per 64-row block

__device__ __host__ inline uint32_t mn_offset(int m, int k) {
    return (uint32_t)((m >> 6) * MBLK + (k >> 3) * KGRP + (k & 7) * 128 + ((((m & 63) >> 3) ^ (k & 7)) << 4) + (m & 7) * 2);
}

__device__ __forceinline__ uint64_t mn_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((addr & 0x3FFFFu) >> 4);
    d |= (uint64_t)(lbo_bytes >> 4) << 16;
    d |= (uint64_t)(sbo_bytes >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

template <int N, int K>
__global__ void __launch_bounds__(128) probe_mn_kernel(const __half* A, const __half* B, float* D, int* status, int variant,
                                                      int reps, long long* cycles) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) unsigned long long mbar;
    __shared__ uint32_t tmem_base_slot;
    unsigned char* sa = smem + ((1024u - ((uint32_t)__cvta_generic_to_shared(smem) & 1023u)) & 1023u);
    unsigned char* sb = sa + 128 * K * 2;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 128 * K; i += 128) {
        const int r = i / K, k = i % K;
        *reinterpret_cast<__half*>(sa + mn_offset(r, k)) = A[i];
    }
    for (int i = tid; i < N * K; i += 128) {
        const int r = i / K, k = i % K;
        *reinterpret_cast<__half*>(sb + sw128_offset(N, r, k)) = B[i];
    }
    const uint32_t mb = (uint32_t)__cvta_generic_to_shared(&mbar);
    if (tid == 0) { mbar_init(mb, 1); fence_mbar_init(); }
    if (warp == 0) tmem_alloc((uint32_t)__cvta_generic_to_shared(&tmem_base_slot), 256);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tbase = tmem_base_slot;
    const uint32_t sa_addr = (uint32_t)__cvta_generic_to_shared(sa), sb_addr = (uint32_t)__cvta_generic_to_shared(sb);
    const uint32_t idesc = idesc_f16(128, N) | (1u << 15);          // a_major = MN
    const uint32_t lbo = variant == 0 ? MBLK : KGRP, sbo = variant == 0 ? KGRP : MBLK;
    uint32_t parity = 0;
    bool ok = true;
    long long t0 = 0;
    for (int rep = 0; rep < reps; ++rep) {
        if (tid == 0) {
            if (rep == 1) t0 = clock64();
            for (int kk = 0; kk < K / 16; ++kk) {
                const uint64_t ad = mn_desc(sa_addr + kk * 2 * KGRP, lbo, sbo);
                const int atom = kk / 4, s = kk % 4;
                const uint64_t bd = sw128_desc(sb_addr + atom * N * 128 + s * 32);
                mma_f16(tbase, ad, bd, idesc, kk > 0 ? 1u : 0u);
            }
            mma_commit(mb);
        }
        ok = mbar_wait(mb, parity) && ok;
        parity ^= 1;
        if (!ok) break;
    }
    if (tid == 0) cycles[0] = clock64() - t0;
    tc_fence_after();
    if (ok) {
        for (int c0 = 0; c0 < N; c0 += 32) {
            float r[32];
            tmem_ld32(tbase + ((uint32_t)(warp * 32) << 16) + c0, r);
            tmem_ld_wait();
            for (int j = 0; j < 32; ++j) if (c0 + j < N) D[(size_t)tid * N + c0 + j] = r[j];
        }
    } else if (tid == 0) {
        status[0] = 1;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_free(tbase, 256);
}

template <int N, int K>
static int run(int variant, int reps) {
    std::vector<__half> hA(128 * K), hB(N * K);
    std::vector<float> fA(128 * K), fB(N * K);
    srand(777 + N + K);
    for (size_t i = 0; i < hA.size(); ++i) { fA[i] = (float)((rand() % 17) - 8) / 8.0f; hA[i] = __float2half(fA[i]); }
    for (size_t i = 0; i < hB.size(); ++i) { fB[i] = (float)((rand() % 13) - 6) / 4.0f; hB[i] = __float2half(fB[i]); }
    __half *dA, *dB; float* dD; int* dS; long long* dC;
    cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dB, hB.size() * 2); cudaMalloc(&dD, 128 * N * 4); cudaMalloc(&dS, 8); cudaMalloc(&dC, 8);
    cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
    cudaMemset(dS, 0, 8); cudaMemset(dD, 0, 128 * N * 4);
    const int smem = (128 + N) * K * 2 + 1024;
    cudaFuncSetAttribute(probe_mn_kernel<N, K>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    probe_mn_kernel<N, K><<<1, 128, smem>>>(dA, dB, dD, dS, variant, reps, dC);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("MN variant %d N=%d K=%d: CUDA error %s\n", variant, N, K, cudaGetErrorString(e)); return 1; }
    int st; long long cyc; std::vector<float> hD(128 * N);
    cudaMemcpy(&st, dS, 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(&cyc, dC, 8, cudaMemcpyDeviceToHost);
    cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost);
    double maxerr = 0; int nbad = 0;
    for (int m = 0; m < 128; ++m)
        for (int n = 0; n < N; ++n) {
            double s = 0;
            for (int k = 0; k < K; ++k) s += (double)fA[m * K + k] * fB[n * K + k];
            const double d = fabs(s - hD[m * N + n]);
            if (d > maxerr) maxerr = d;
            if (d > 1e-3) ++nbad;
        }
    printf("MN-major A, variant %d (LBO=%d SBO=%d) N=%d K=%d: status=%d bad=%d max|err|=%g cycles/rep=%.1f (%d MMAs)\n", variant,
           variant == 0 ? MBLK : KGRP, variant == 0 ? KGRP : MBLK, N, K, st, nbad, maxerr, reps > 1 ? (double)cyc / (reps - 1) : 0.0,
           K / 16);
    return (st != 0 || maxerr > 1e-3) ? 1 : 0;
}

int main() {
    const int b0 = run<64, 64>(0, 1);
    const int b1 = run<64, 64>(1, 1);
    if (!b0) run<64, 64>(0, 200);
    if (!b1) run<64, 64>(1, 200);
    printf((b0 && b1) ? "PROBE FAILED\n" : "PROBE OK (variant %d)\n", b0 ? 1 : 0);
    return b0 && b1;
}
