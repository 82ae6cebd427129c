// Probe for the tcgen05 building blocks in pyfft_b200/csrc/tc_core.cuh: one CTA computes
// D[128][N] = A[128][K] * B[N][K]^T (fp16 in, fp32 accumulate in TMEM) from K-major SW128
// shared-memory tiles and the host checks every element.  Also times a burst of MMAs.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/tc_probe tools/tc_probe.cu
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../pyfft_b200/csrc/tc_core.cuh"

using namespace pyfft::tc;

template <int N, int K>
__global__ void __launch_bounds__(128) probe_kernel(const __half* A, const __half* B, float* D, int* status, int reps,
                                                   long long* cycles) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) unsigned long long mbar;
    __shared__ uint32_t tmem_base_slot;
    unsigned char* sa = smem + ((1024u - ((uint32_t)__cvta_generic_to_shared(smem) & 1023u)) & 1023u);   // A: 128 rows x K
    unsigned char* sb = sa + 128 * K * 2;         // B: N rows x K
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 128 * K; i += 128) {
        const int r = i / K, k = i % K;
        *reinterpret_cast<__half*>(sa + sw128_offset(128, r, k)) = A[i];
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
    uint32_t parity = 0;
    bool ok = true;
    long long t0 = 0, t1 = 0;
    for (int rep = 0; rep < reps; ++rep) {
        if (tid == 0) {
            if (rep == 1) t0 = clock64();
            for (int kk = 0; kk < K / 16; ++kk) {
                const int atom = kk / 4, s = kk % 4;
                const uint64_t ad = sw128_desc(sa_addr + atom * 128 * 128 + s * 32);
                const uint64_t bd = sw128_desc(sb_addr + atom * N * 128 + s * 32);
                mma_f16(tbase, ad, bd, idesc_f16(128, N), kk > 0 ? 1u : 0u);
            }
            mma_commit(mb);
        }
        ok = mbar_wait(mb, parity) && ok;
        parity ^= 1;
        if (!ok) break;
    }
    if (tid == 0) { t1 = clock64(); cycles[0] = t1 - t0; }
    tc_fence_after();
    if (ok) {
        for (int c0 = 0; c0 < N; c0 += 32) {
            float r[32];
            tmem_ld32(tbase + ((uint32_t)(warp * 32) << 16) + c0, r);
            tmem_ld_wait();
            for (int j = 0; j < 32; ++j) if (c0 + j < N) D[(size_t)tid * N + c0 + j] = r[j];
        }
        // unaligned column offsets: 32 columns from column 34, 2 columns from column 66
        if (N >= 72) {
            float r[32], a, b;
            tmem_ld32(tbase + ((uint32_t)(warp * 32) << 16) + 34, r);
            tmem_ld2(tbase + ((uint32_t)(warp * 32) << 16) + 66, a, b);
            tmem_ld_wait();
            bool same = (a == D[(size_t)tid * N + 66]) && (b == D[(size_t)tid * N + 67]);
            for (int j = 0; j < 32; ++j) same = same && (r[j] == D[(size_t)tid * N + 34 + j]);
            if (!same) atomicAdd(&status[1], 1);
        }
    } else if (tid == 0) {
        status[0] = 1;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_free(tbase, 256);
}


// A operand from tensor memory: thread m writes row m of A, two halves (k = 2c, 2c+1) per 32-bit column
template <int N, int K>
__global__ void __launch_bounds__(128) probe_ts_kernel(const __half* A, const __half* B, float* D, int* status, int reps,
                                                      long long* cycles) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) unsigned long long mbar;
    __shared__ uint32_t tmem_base_slot;
    unsigned char* sb = smem + ((1024u - ((uint32_t)__cvta_generic_to_shared(smem) & 1023u)) & 1023u);
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < N * K; i += 128) {
        const int r = i / K, k = i % K;
        *reinterpret_cast<__half*>(sb + sw128_offset(N, r, k)) = B[i];
    }
    const uint32_t mb = (uint32_t)__cvta_generic_to_shared(&mbar);
    if (tid == 0) { mbar_init(mb, 1); fence_mbar_init(); }
    if (warp == 0) tmem_alloc((uint32_t)__cvta_generic_to_shared(&tmem_base_slot), 512);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tbase = tmem_base_slot;
    const uint32_t a_col = 256;                         // A lives at columns 256 .. 256 + K/2
    // row tid of A -> TMEM lane tid
    for (int c0 = 0; c0 < K / 2; c0 += 8) {
        uint32_t v[8];
        for (int j = 0; j < 8; ++j) {
            const __half2 h = __halves2half2(A[tid * K + 2 * (c0 + j)], A[tid * K + 2 * (c0 + j) + 1]);
            v[j] = *reinterpret_cast<const uint32_t*>(&h);
        }
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                     ::"r"(tbase + ((uint32_t)(warp * 32) << 16) + a_col + c0), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]),
                       "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t sb_addr = (uint32_t)__cvta_generic_to_shared(sb);
    bool ok = true;
    uint32_t parity = 0;
    long long t0 = 0;
    for (int rep = 0; rep < reps; ++rep) {
        if (tid == 0) {
            if (rep == 1) t0 = clock64();
            for (int kk = 0; kk < K / 16; ++kk) {
                const int atom = kk / 4, s = kk % 4;
                const uint64_t bd = sw128_desc(sb_addr + atom * N * 128 + s * 32);
                const uint32_t acc = kk > 0 ? 1u : 0u;
                asm volatile(
                    "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                    "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                    ::"r"(tbase), "r"(tbase + a_col + kk * 8), "l"(bd), "r"(idesc_f16(128, N)), "r"(acc)
                    : "memory");
            }
            mma_commit(mb);
        }
        ok = mbar_wait(mb, parity) && ok;
        parity ^= 1;
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
    if (warp == 0) tmem_free(tbase, 512);
}

template <int N, int K>
static int run_ts(int reps) {
    std::vector<__half> hA(128 * K), hB(N * K);
    std::vector<float> fA(128 * K), fB(N * K);
    srand(4321 + N + K);
    for (size_t i = 0; i < hA.size(); ++i) { fA[i] = (float)((rand() % 17) - 8) / 8.0f; hA[i] = __float2half(fA[i]); }
    for (size_t i = 0; i < hB.size(); ++i) { fB[i] = (float)((rand() % 13) - 6) / 4.0f; hB[i] = __float2half(fB[i]); }
    __half *dA, *dB; float* dD; int* dS; long long* dC;
    cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dB, hB.size() * 2); cudaMalloc(&dD, 128 * N * 4); cudaMalloc(&dS, 8); cudaMalloc(&dC, 8);
    cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
    cudaMemset(dS, 0, 8); cudaMemset(dD, 0, 128 * N * 4);
    const int smem = N * K * 2 + 1024;
    cudaFuncSetAttribute(probe_ts_kernel<N, K>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    probe_ts_kernel<N, K><<<1, 128, smem>>>(dA, dB, dD, dS, reps, dC);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("TS N=%d K=%d: CUDA error %s\n", N, K, cudaGetErrorString(e)); return 1; }
    int st; std::vector<float> hD(128 * N);
    cudaMemcpy(&st, dS, 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost);
    double maxerr = 0;
    for (int m = 0; m < 128; ++m)
        for (int n = 0; n < N; ++n) {
            double s = 0;
            for (int k = 0; k < K; ++k) s += (double)fA[m * K + k] * fB[n * K + k];
            const double d = fabs(s - hD[m * N + n]);
            if (d > maxerr) maxerr = d;
        }
    long long cyc; cudaMemcpy(&cyc, dC, 8, cudaMemcpyDeviceToHost);
    printf("TS (A from TMEM) N=%d K=%d: status=%d max|err|=%g D[0][0]=%g D[127][%d]=%g cycles/rep=%.1f (%d MMAs)\n", N, K, st, maxerr,
           hD[0], N - 1, hD[127 * N + N - 1], reps > 1 ? (double)cyc / (reps - 1) : 0.0, K / 16);
    return (st != 0 || maxerr > 1e-3) ? 1 : 0;
}

template <int N, int K>
static int run(int reps) {
    std::vector<__half> hA(128 * K), hB(N * K);
    std::vector<float> fA(128 * K), fB(N * K);
    srand(1234 + N + K);
    for (size_t i = 0; i < hA.size(); ++i) { fA[i] = (float)((rand() % 17) - 8) / 8.0f; hA[i] = __float2half(fA[i]); }
    for (size_t i = 0; i < hB.size(); ++i) { fB[i] = (float)((rand() % 13) - 6) / 4.0f; hB[i] = __float2half(fB[i]); }
    __half *dA, *dB;
    float* dD;
    int* dS;
    long long* dC;
    cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dB, hB.size() * 2); cudaMalloc(&dD, 128 * N * 4);
    cudaMalloc(&dS, 8); cudaMalloc(&dC, 8);
    cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
    cudaMemset(dS, 0, 8); cudaMemset(dD, 0, 128 * N * 4);
    const int smem = (128 + N) * K * 2 + 1024;
    cudaFuncSetAttribute(probe_kernel<N, K>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    probe_kernel<N, K><<<1, 128, smem>>>(dA, dB, dD, dS, reps, dC);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("N=%d K=%d: CUDA error %s\n", N, K, cudaGetErrorString(e)); return 1; }
    int st, st2; long long cyc;
    std::vector<float> hD(128 * N);
    cudaMemcpy(&st, dS, 4, cudaMemcpyDeviceToHost); cudaMemcpy(&st2, dS + 1, 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(&cyc, dC, 8, cudaMemcpyDeviceToHost);
    cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost);
    double maxerr = 0;
    for (int m = 0; m < 128; ++m)
        for (int n = 0; n < N; ++n) {
            double s = 0;
            for (int k = 0; k < K; ++k) s += (double)fA[m * K + k] * fB[n * K + k];
            const double d = fabs(s - hD[m * N + n]);
            if (d > maxerr) maxerr = d;
        }
    printf("N=%d K=%d: status=%d unaligned_ld_mismatch=%d max|err|=%g  D[0][0]=%g D[127][%d]=%g  cycles/rep=%.1f (reps=%d, %d MMAs per rep)\n", N,
           K, st, st2, maxerr, hD[0], N - 1, hD[127 * N + N - 1], reps > 1 ? (double)cyc / (reps - 1) : 0.0, reps, K / 16);
    return (st != 0 || maxerr > 1e-3) ? 1 : 0;
}

int main() {
    int bad = 0;
    bad += run<64, 64>(1);
    bad += run<64, 128>(1);
    bad += run<208, 128>(1);
    bad += run<256, 128>(1);
    bad += run<64, 64>(200);
    bad += run<208, 128>(200);
    bad += run<256, 128>(200);
    bad += run<144, 128>(200);
    bad += run<128, 64>(200);
    bad += run_ts<80, 128>(1);
    bad += run_ts<80, 128>(200);
    bad += run_ts<144, 128>(200);
    bad += run_ts<256, 128>(200);
    bad += run<80, 128>(200);
    printf(bad ? "PROBE FAILED\n" : "PROBE OK\n");
    return bad;
}
