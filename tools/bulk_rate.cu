// cp.async.bulk (global -> shared, mbarrier completion) issue cost and throughput on one SM of a B200, for the
// TMA ring of welch_tc_kernel: one thread issues NCOPY copies of SIZE bytes into a ring of shared memory.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/bulk_rate tools/bulk_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

__device__ __forceinline__ void mbar_init(uint32_t mbar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint32_t mbar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(mbar), "r"(parity) : "memory");
    return ok != 0;
}

template <int SIZE, int PER_BAR>
__global__ void __launch_bounds__(128) k(const unsigned char* src, int64_t src_stride, int offset, int ncopy, long long* out) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) unsigned long long bar[64];
    const uint32_t b0 = (uint32_t)__cvta_generic_to_shared(&bar[0]);
    const uint32_t s0 = (uint32_t)__cvta_generic_to_shared(smem);
    if (threadIdx.x == 0) {
        for (int i = 0; i < 64; ++i) mbar_init(b0 + 8 * i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned char* p = src + (size_t)blockIdx.x * src_stride * ncopy + offset;
        const long long t0 = clock64();
        int nb = 0;
        for (int i = 0; i < ncopy; i += PER_BAR, ++nb) {
            const uint32_t bar_i = b0 + 8 * (nb & 63);
            if (nb >= 64) { while (!mbar_try(bar_i, ((nb >> 6) - 1) & 1)) {} }      // ring of 64 groups in flight
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_i), "r"(SIZE * PER_BAR) : "memory");
            for (int q = 0; q < PER_BAR; ++q)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(s0 + ((nb & 7) * PER_BAR + q) * ((SIZE + 127) / 128 * 128)), "l"(p + (size_t)(i + q) * src_stride), "r"(SIZE), "r"(bar_i) : "memory");
        }
        const long long t_issue = clock64() - t0;
        for (int g = (nb > 64 ? nb - 64 : 0); g < nb; ++g) while (!mbar_try(b0 + 8 * (g & 63), (g >> 6) & 1)) {}
        out[2 * blockIdx.x] = t_issue;
        out[2 * blockIdx.x + 1] = clock64() - t0;
    }
}

template <int SIZE, int PER_BAR>
void run(const unsigned char* d, int64_t stride, int offset, int ctas, const char* what) {
    long long* o; cudaMalloc(&o, 16 * ctas);
    const int ncopy = 1024;
    const int smem = 8 * PER_BAR * ((SIZE + 127) / 128 * 128);
    cudaFuncSetAttribute(k<SIZE, PER_BAR>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    k<SIZE, PER_BAR><<<ctas, 128, smem>>>(d, stride, offset, ncopy, o);
    k<SIZE, PER_BAR><<<ctas, 128, smem>>>(d, stride, offset, ncopy, o);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[2 * 148]; cudaMemcpy(h, o, 16 * ctas, cudaMemcpyDeviceToHost);
    double iss = 0, tot = 0; for (int i = 0; i < ctas; ++i) { iss += h[2 * i]; tot += h[2 * i + 1]; }
    iss /= ctas; tot /= ctas;
    printf("%-34s size %5d x%d per barrier, %3d CTAs: issue %6.1f clk/copy, total %6.1f clk/copy = %5.1f B/clk/SM (%s)\n", what, SIZE, PER_BAR,
           ctas, iss / ncopy, tot / ncopy, SIZE * (double)ncopy / tot, cudaGetErrorString(e));
    cudaFree(o);
}

int main() {
    unsigned char* d; const size_t bytes = (size_t)3 << 30;           // 148 CTAs x 1024 copies x 16 KB = 2.3 GB
    cudaMalloc(&d, bytes); cudaMemset(d, 1, bytes);
    for (int ctas : {1, 148}) {
        run<2048, 1>(d, 2048, 0, ctas, "aligned 128, contiguous");
        run<2064, 1>(d, 2048, 16, ctas, "2064 B, start +16");
        run<2064, 2>(d, 2048, 16, ctas, "2064 B, start +16, 2 per barrier");
        run<4096, 1>(d, 4096, 0, ctas, "aligned 128, contiguous");
        run<8192, 1>(d, 8192, 0, ctas, "aligned 128, contiguous");
        run<8208, 1>(d, 8192, 48, ctas, "8208 B, start +48");
        run<16384, 1>(d, 16384, 0, ctas, "aligned 128, contiguous");
    }
    return 0;
}
