// Tensor-memory read throughput: W warps (W = 1, 4, 8) of one CTA read their lane quarter with
// tcgen05.ld.32x32b.xN back to back; prints bytes per clock for the CTA.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/tmem_bw tools/tmem_bw.cu
#include <cstdio>
#include "../pyfft_b200/csrc/tc_core.cuh"
using namespace pyfft::tc;

template <int X>
__global__ void __launch_bounds__(256) tmem_bw_kernel(int nwarps, int iters, long long* cycles, float* sink) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) tmem_alloc((uint32_t)__cvta_generic_to_shared(&slot), 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tl = slot + ((uint32_t)((warp & 3) * 32) << 16);
    float acc = 0.f;
    __syncthreads();
    const long long t0 = clock64();
    if (warp < nwarps) {
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (X == 32) { float r[32]; tmem_ld32(tl + ((it * 4 + u) & 7) * 32, r); tmem_ld_wait(); acc += r[0] + r[31]; }
                if (X == 8) {
                    float a[8], b[8], c[8], d[8];
                    const uint32_t c0 = tl + ((it * 4 + u) & 7) * 32;
                    tmem_ld8(c0, a); tmem_ld8(c0 + 8, b); tmem_ld8(c0 + 16, c); tmem_ld8(c0 + 24, d);
                    tmem_ld_wait();
                    acc += a[0] + b[7] + c[3] + d[5];
                }
            }
        }
    }
    const long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0) cycles[0] = t1 - t0;
    if (acc == 123.456f) sink[threadIdx.x] = acc;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_free(slot, 512);
}

int main() {
    long long* dC; float* dS;
    cudaMalloc(&dC, 8); cudaMalloc(&dS, 1024);
    const int iters = 2000;
    for (int x : {32, 8})
        for (int w : {1, 2, 4, 8}) {
            if (x == 32) tmem_bw_kernel<32><<<1, 256>>>(w, iters, dC, dS);
            else tmem_bw_kernel<8><<<1, 256>>>(w, iters, dC, dS);
            cudaError_t e = cudaDeviceSynchronize();
            long long c; cudaMemcpy(&c, dC, 8, cudaMemcpyDeviceToHost);
            const double bytes = (double)w * iters * 4 * 32 * 32 * 4;
            printf("x%-2d loads, %d warps: %lld clk, %.1f B/clk per CTA, %.1f clk per 4 KB warp-load group (%s)\n", x, w, c, bytes / c,
                   (double)c / iters, cudaGetErrorString(e));
        }
    return 0;
}
