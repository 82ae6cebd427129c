// Pipe-rate microbenchmarks for the kernel design (run once on the B200):
// scalar FADD/FMUL/FFMA vs packed f32x2, DFMA, LDS.64, STS.64, SHFL.
// Prints warp-instructions per clock per SM.   nvcc -arch=sm_100a -O3 microbench.cu
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 4096
#define NCH 8   // independent chains per thread

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, float a, float b, long long* cyc) {
    extern __shared__ float2 sm[];
    float r[NCH * 2];
    double d[NCH];
#pragma unroll
    for (int i = 0; i < NCH * 2; ++i) r[i] = threadIdx.x * 0.001f + i;
#pragma unroll
    for (int i = 0; i < NCH; ++i) d[i] = threadIdx.x * 0.001 + i;
    sm[threadIdx.x] = make_float2(a, b);
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < NCH; ++i) {
            if (MODE == 0) { r[2 * i] = r[2 * i] + a; r[2 * i + 1] = r[2 * i + 1] + b; }               // 2 FADD
            if (MODE == 1) { r[2 * i] = r[2 * i] * a; r[2 * i + 1] = r[2 * i + 1] * b; }               // 2 FMUL
            if (MODE == 2) { r[2 * i] = fmaf(r[2 * i], a, b); r[2 * i + 1] = fmaf(r[2 * i + 1], b, a); } // 2 FFMA (2 regs + const)
            if (MODE == 3) { r[2 * i] = fmaf(r[2 * i], r[2 * i + 1], a); r[2 * i + 1] = fmaf(r[2 * i + 1], r[2 * i], b); } // FFMA 3 distinct regs
            if (MODE == 4) { float2 v = make_float2(r[2 * i], r[2 * i + 1]);
                             v = __ffma2_rn(v, make_float2(a, b), make_float2(b, a)); r[2 * i] = v.x; r[2 * i + 1] = v.y; }
            if (MODE == 5) { float2 v = make_float2(r[2 * i], r[2 * i + 1]);
                             v = __fadd2_rn(v, make_float2(a, b)); r[2 * i] = v.x; r[2 * i + 1] = v.y; }
            if (MODE == 6) { d[i] = fma(d[i], (double)a, (double)b); }
            if (MODE == 7) { float2 v = sm[(threadIdx.x + i * 32 + it) & 255]; r[2 * i] += v.x; r[2 * i + 1] += v.y; } // LDS.64 + 2 FADD
            if (MODE == 8) { sm[(threadIdx.x + i * 32) & 255] = make_float2(r[2 * i], r[2 * i + 1]); r[2*i] += a; }   // STS.64 + FADD
            if (MODE == 9) { r[2 * i] = __shfl_xor_sync(0xffffffffu, r[2 * i], 16) + a; }                             // SHFL + FADD
            if (MODE == 10) { float2 v = make_float2(r[2 * i], r[2 * i + 1]); float2 w = make_float2(r[(2*i+2)%(2*NCH)], r[(2*i+3)%(2*NCH)]);
                             v = __ffma2_rn(v, w, make_float2(a, b)); r[2 * i] = v.x; r[2 * i + 1] = v.y; }            // FFMA2 two reg pairs
        }
    }
    long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < NCH * 2; ++i) s += r[i];
#pragma unroll
    for (int i = 0; i < NCH; ++i) s += (float)d[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name, double instr_per_chain, int ctas_per_sm) {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int grid = sms * ctas_per_sm;
    float* out; long long* cyc;
    cudaMalloc(&out, grid * 256 * sizeof(float)); cudaMalloc(&cyc, grid * sizeof(long long));
    k<MODE><<<grid, 256, 2048>>>(out, 1.0001f, 0.9999f, cyc);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<MODE><<<grid, 256, 2048>>>(out, 1.0001f, 0.9999f, cyc);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long h[4096]; cudaMemcpy(h, cyc, grid * sizeof(long long), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < grid; ++i) avg += h[i]; avg /= grid;
    double warp_instr_per_cta = 8.0 * ITERS * NCH * instr_per_chain;     // 8 warps
    printf("%-28s ctas/sm=%d  cyc/cta=%9.0f  warp-instr/clk/SM=%6.2f  (time %.3f ms, err=%s)\n", name, ctas_per_sm, avg,
           warp_instr_per_cta * ctas_per_sm / avg, ms, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out); cudaFree(cyc);
}

int main() {
    for (int c = 1; c <= 4; c *= 2) {
        run<0>("FADD x2", 2, c); run<1>("FMUL x2", 2, c); run<2>("FFMA x2 (reg,const,const)", 2, c);
        run<3>("FFMA x2 (3 regs)", 2, c); run<4>("FFMA2 (pair,const)", 1, c); run<10>("FFMA2 (pair,pair,const)", 1, c);
        run<5>("FADD2", 1, c); run<6>("DFMA", 1, c); run<7>("LDS.64 (+2 FADD)", 1, c); run<8>("STS.64 (+FADD)", 1, c);
        run<9>("SHFL (+FADD)", 1, c);
    }
    return 0;
}
