// Shared host/device helpers: status + last-error plumbing, the record map
// gather, twiddle-table cache.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

#include "../../include/pyfft_b200.h"
#include "fft_core.cuh"

namespace pyfft {

// ---- errors -------------------------------------------------------------- //
std::string& last_error_ref();
int fail(int code, const char* fmt, ...);

#define PYFFT_CUDA_OK(expr)                                                            \
    do {                                                                               \
        cudaError_t _e = (expr);                                                       \
        if (_e != cudaSuccess)                                                         \
            return ::pyfft::fail(PYFFT_ECUDA, "%s failed: %s (%s:%d)", #expr,          \
                                 cudaGetErrorString(_e), __FILE__, __LINE__);          \
    } while (0)

#define PYFFT_LAUNCH_OK(name)                                                          \
    do {                                                                               \
        cudaError_t _e = cudaGetLastError();                                           \
        if (_e != cudaSuccess)                                                         \
            return ::pyfft::fail(PYFFT_ECUDA, "launch of %s failed: %s", name,         \
                                 cudaGetErrorString(_e));                              \
    } while (0)

// ---- record map ----------------------------------------------------------- //
struct RecMap {
    int64_t nsig, i0, reflect, src_base, src_len;
};

// analysed index j -> element offset behind the channel pointer
// (reference: reflect-pad fft_analysis.py:196-205 then slice [i0:i1] :353-354)
__host__ __device__ __forceinline__ int64_t src_index(const RecMap& m, int64_t j) {
    return (j < m.reflect ? m.reflect - j : j - m.reflect + m.i0) - m.src_base;
}

int check_map(const pyfft_record_map* map, RecMap* out);
// validates that analysed indices [j_lo, j_hi) are inside the record and behind the pointers
int check_range(const RecMap& m, int64_t j_lo, int64_t j_hi);
int check_seg(const pyfft_segmentation* seg, int min_log2, int max_log2, int* log2l);

inline int ilog2(int64_t v) {
    int l = 0;
    while ((int64_t(1) << l) < v) ++l;
    return l;
}

// ---- twiddles ------------------------------------------------------------- //
// exp(-2 pi i k / L), k = 0..L-1, computed in long double on the host, cached
// per (device, L, dtype).  First use of an L allocates and copies
// synchronously -- do that (one warm-up call) outside CUDA graph capture.
const void* twiddle_table(int dtype, int64_t L, int* status);
// exp(-2 pi i k step / M), k = 0..count-1 (two-level tables of the long transforms)
const void* twiddle_table_ex(int dtype, int64_t M, int64_t step, int64_t count, int* status);
void free_twiddles();

int device_sm_count();

// ---- system-scope accesses of the NVLink peer exchanges (peer.cu, welch_tc.cu) ---- //
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ double ld_relaxed_sys(const double* p) {
    double v;
    asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}

// shared between welch.cu and welch_tc.cu
int welch_reduce_launch_f32(const float* partial, int nparts, int per_pair, int nch, double* acc, cudaStream_t st);
int welch_raw_sums_launch(const double* dev_sums, const double* est, int n, double count, double* stats,
                          cudaStream_t st);

}  // namespace pyfft
