// Transforms of ARBITRARY length (the reference never zero-pads: nfft = nwins = whatever Navr and the
// overlap give, fft_analysis.py:185-189, 219-221, 2411-2418; hilbert.py:38-64 takes any n) by Bluestein's
// chirp-z identity on top of the power-of-two kernels of bigfft.cu:
//     X[k] = c[k] * sum_n (x[n] c[n]) conj(c[k - n]),      c[m] = exp(-i pi m^2 / n)
// i.e. one circular convolution of length M = 2^p >= 2n - 1: pack (gather / detrend / window / chirp, zero
// padding) -> forward transform -> multiply by the transformed chirp filter -> inverse transform -> unpack
// (chirp, bin selection, scaling).  The inverse DFT is the conjugate of the forward one of the conjugate.
// Arithmetic is float64 whatever the input type: Bluestein's sums cancel heavily and the fp32 contract
// (1e-5 of the spectral peak) is not safe in single precision; inputs and outputs keep the caller's type.
// Reference lines replaced: the np.fft.fft calls of fft_pwelch (:375-388) and fftanal.fft_win (:2152-2176)
// for window lengths that are not powers of two, complex-valued inputs of fft_pwelch (:123-127, :422-428),
// np.fft.fft / ifft of hilbert.py:44-66.
#include "common.cuh"

namespace pyfft {

typedef cpx<double> zc;

// c[m] = exp(-i pi m^2 / n): the angle is reduced exactly in integers (m^2 mod 2n) before sincospi
__global__ void __launch_bounds__(256) any_chirp_kernel(zc* __restrict__ c, int64_t n) {
    const int64_t m = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (m >= n) return;
    const unsigned long long r = ((unsigned long long)m * (unsigned long long)m) % (unsigned long long)(2 * n);
    double s, co;
    sincospi(-(double)r / (double)n, &s, &co);
    c[m] = mk<double>(co, s);
}

// chirp filter b[m] = conj(c[|m|]), |m| < n, wrapped into length M (zero elsewhere)
__global__ void __launch_bounds__(256) any_filter_kernel(zc* __restrict__ b, const zc* __restrict__ c, int64_t n, int64_t M) {
    const int64_t m = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (m >= M) return;
    zc v = mk<double>(0.0, 0.0);
    if (m < n) v = cconj(c[m]);
    else if (M - m < n) v = cconj(c[M - m]);
    b[m] = v;
}

// rows of a contiguous array: a[r][m] = (inverse ? conj : id)(in[r][m]) c[m], zero for m >= n
template <typename In, bool CPLX>
__global__ void __launch_bounds__(256) any_pack_rows_kernel(const In* __restrict__ in, int64_t row_stride, int64_t n,
                                                            int64_t M, int inverse, const zc* __restrict__ c,
                                                            zc* __restrict__ a) {
    const int64_t r = blockIdx.y;
    const In* src = in + r * row_stride * (CPLX ? 2 : 1);
    zc* dst = a + r * M;
    for (int64_t m = (int64_t)blockIdx.x * 256 + threadIdx.x; m < M; m += (int64_t)gridDim.x * 256) {
        zc v = mk<double>(0.0, 0.0);
        if (m < n) {
            v = CPLX ? mk<double>((double)src[2 * m], (double)src[2 * m + 1]) : mk<double>((double)src[m], 0.0);
            if (inverse) v = cconj(v);
            v = cmul(v, c[m]);
        }
        dst[m] = v;
    }
}

// rows of laplace_1d (laplace.py:85-113): row r = u[m] * exp(sigma_r (n-1-m)) / sum_k exp(sigma_r k), then chirp
template <typename In>
__global__ void __launch_bounds__(256) any_pack_laplace_kernel(const In* __restrict__ u, int64_t n, int64_t M,
                                                               const double* __restrict__ sigma, const zc* __restrict__ c,
                                                               zc* __restrict__ a) {
    const int64_t r = blockIdx.y;
    const double sg = sigma[r];
    const double S = sg != 0.0 ? expm1(sg * (double)n) / expm1(sg) : (double)n;     // geometric series
    zc* dst = a + r * M;
    for (int64_t m = (int64_t)blockIdx.x * 256 + threadIdx.x; m < M; m += (int64_t)gridDim.x * 256) {
        zc v = mk<double>(0.0, 0.0);
        if (m < n) v = cmul(mk<double>((double)u[m] * (exp(sg * (double)(n - 1 - m)) / S), 0.0), c[m]);
        dst[m] = v;
    }
}

// segments of channel records (record map, trend, optional per-segment mean, window), one CTA per row (c, k)
template <typename In, bool CPLX>
__global__ void __launch_bounds__(256) any_pack_segments_kernel(const In* __restrict__ x, int64_t ch_stride, RecMap map,
                                                                int64_t hop, int64_t seg_first, int64_t seg_count,
                                                                int64_t n, int64_t M, const double* __restrict__ window,
                                                                const double* __restrict__ coef, int per_segment,
                                                                const zc* __restrict__ c, zc* __restrict__ a,
                                                                double* __restrict__ pseg) {
    const int64_t k = blockIdx.x, ch = blockIdx.y;
    const In* src = x + ch * ch_stride;
    const int64_t j0 = (seg_first + k) * hop;
    zc* dst = a + (ch * seg_count + k) * M;
    // trend: real input (a + b j) from coef[2c], coef[2c+1]; complex input: means at coef[4c], coef[4c+2]
    const double ta = coef ? (CPLX ? coef[4 * ch] : coef[2 * ch]) : 0.0;
    const double tb = coef ? (CPLX ? 0.0 : coef[2 * ch + 1]) : 0.0;
    const double ti = (coef && CPLX) ? coef[4 * ch + 2] : 0.0;
    __shared__ double red[3][8];
    double mr = 0.0, mi = 0.0;
    if (per_segment) {                                      // detrendwin=True: the segment's own mean on top
        double sr = 0.0, si = 0.0;
        for (int64_t m = threadIdx.x; m < n; m += 256) {
            const int64_t e = src_index(map, j0 + m);
            sr += (CPLX ? (double)src[2 * e] : (double)src[e]) - (ta + tb * (double)(j0 + m));
            if (CPLX) si += (double)src[2 * e + 1] - ti;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { sr += __shfl_xor_sync(0xffffffffu, sr, o); si += __shfl_xor_sync(0xffffffffu, si, o); }
        if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = sr; red[1][threadIdx.x >> 5] = si; }
        __syncthreads();
        for (int w = 0; w < 8; ++w) { mr += red[0][w]; mi += red[1][w]; }
        mr /= (double)n; mi /= (double)n;
        __syncthreads();
    }
    double g = 0.0;
    for (int64_t m = threadIdx.x; m < M; m += 256) {
        zc v = mk<double>(0.0, 0.0);
        if (m < n) {
            const int64_t e = src_index(map, j0 + m);
            const double w = window[m];
            double vr = (CPLX ? (double)src[2 * e] : (double)src[e]) - (ta + tb * (double)(j0 + m)) - mr;
            double vi = CPLX ? (double)src[2 * e + 1] - ti - mi : 0.0;
            vr *= w; vi *= w;
            const double gg = vr * vr + vi * vi;
            g += (m == 0 || m == n - 1) ? 0.5 * gg : gg;    // trapezoid rule core of fft_analysis.py:2174
            v = cmul(mk<double>(vr, vi), c[m]);
        }
        dst[m] = v;
    }
    if (pseg) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) g += __shfl_xor_sync(0xffffffffu, g, o);
        if ((threadIdx.x & 31) == 0) red[2][threadIdx.x >> 5] = g;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = 0.0;
            for (int w = 0; w < 8; ++w) t += red[2][w];
            pseg[ch * seg_count + k] = t;
        }
    }
}

// a[r][m] *= bhat[m]
__global__ void __launch_bounds__(256) any_multiply_kernel(zc* __restrict__ a, const zc* __restrict__ bhat, int64_t M, int64_t total) {
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256)
        a[i] = cmul(a[i], bhat[i & (M - 1)]);
}

// out[r][b] = scale_b * (inverse ? conj : id)(a[r][k] c[k]),  k = (bin_first + b) mod n
template <typename Out>
__global__ void __launch_bounds__(256) any_unpack_kernel(const zc* __restrict__ a, const zc* __restrict__ c, int64_t n, int64_t M,
                                                         int inverse, double scale, int64_t bin_first, int64_t nbins,
                                                         const double* __restrict__ bin_scale, int power_out,
                                                         Out* __restrict__ out) {
    const int64_t r = blockIdx.y;
    const zc* src = a + r * M;
    for (int64_t b = (int64_t)blockIdx.x * 256 + threadIdx.x; b < nbins; b += (int64_t)gridDim.x * 256) {
        int64_t k = (bin_first + b) % n;
        if (k < 0) k += n;
        zc v = cmul(src[k], c[k]);
        if (inverse) v = cconj(v);
        const double sc = scale * (bin_scale ? bin_scale[b] : 1.0);
        if (power_out) out[r * nbins + b] = (Out)((v.x * v.x + v.y * v.y) * sc);
        else {
            out[2 * (r * nbins + b)] = (Out)(v.x * sc);
            out[2 * (r * nbins + b) + 1] = (Out)(v.y * sc);
        }
    }
}

static int64_t any_pow2(int64_t n) {
    int64_t M = 16;
    while (M < 2 * n - 1) M *= 2;
    return M;
}

struct AnyWs {
    zc *chirp, *bhat, *work;
    void* big;
    int64_t big_bytes, M;
};

static int any_carve(int64_t n, int64_t rows, void* ws, int64_t wsb, AnyWs* o) {
    if (n < 1 || rows < 1) return fail(PYFFT_EINVAL, "anyfft: n and rows must be positive");
    if (n > (int64_t(1) << 26)) return fail(PYFFT_EUNSUPPORTED, "anyfft: n = %lld exceeds 2^26", (long long)n);
    const int64_t M = any_pow2(n);
    const int64_t cb = ((n * 16 + 255) / 256) * 256, bb = M * 16, wb = rows * M * 16;
    const int64_t big = pyfft_bigfft_workspace_bytes(PYFFT_F64, M, (int32_t)rows);
    if (!ws || wsb < cb + bb + wb + big)
        return fail(PYFFT_EINVAL, "anyfft: workspace of %lld bytes needed (pyfft_anyfft_workspace_bytes)",
                    (long long)(cb + bb + wb + big));
    char* p = static_cast<char*>(ws);
    o->chirp = reinterpret_cast<zc*>(p);
    o->bhat = reinterpret_cast<zc*>(p + cb);
    o->work = reinterpret_cast<zc*>(p + cb + bb);
    o->big = p + cb + bb + wb;
    o->big_bytes = big;
    o->M = M;
    return PYFFT_OK;
}

// chirp, transformed filter; then (after the caller packed `work`): convolve in place
static int any_prepare(int64_t n, const AnyWs& w, cudaStream_t st) {
    any_chirp_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(w.chirp, n);
    PYFFT_LAUNCH_OK("any_chirp_kernel");
    any_filter_kernel<<<(unsigned)((w.M + 255) / 256), 256, 0, st>>>(w.bhat, w.chirp, n, w.M);
    PYFFT_LAUNCH_OK("any_filter_kernel");
    return pyfft_fft_c2c(PYFFT_F64, w.bhat, w.bhat, w.M, 1, 0, w.big, w.big_bytes, st);
}

static int any_convolve(int64_t rows, const AnyWs& w, cudaStream_t st) {
    int rc = pyfft_fft_c2c(PYFFT_F64, w.work, w.work, w.M, (int32_t)rows, 0, w.big, w.big_bytes, st);
    if (rc) return rc;
    const int64_t total = rows * w.M;
    int64_t blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    any_multiply_kernel<<<(unsigned)blocks, 256, 0, st>>>(w.work, w.bhat, w.M, total);
    PYFFT_LAUNCH_OK("any_multiply_kernel");
    return pyfft_fft_c2c(PYFFT_F64, w.work, w.work, w.M, (int32_t)rows, 1, w.big, w.big_bytes, st);   // includes 1/M
}

template <typename Out>
static int any_unpack(const AnyWs& w, int64_t n, int64_t rows, int inverse, double scale, int64_t bin_first, int64_t nbins,
                      const double* bin_scale, int power_out, void* out, cudaStream_t st) {
    int64_t gx = (nbins + 255) / 256;
    if (gx > 64) gx = 64;
    any_unpack_kernel<Out><<<dim3((unsigned)gx, (unsigned)rows), 256, 0, st>>>(w.work, w.chirp, n, w.M, inverse, scale, bin_first,
                                                                           nbins, bin_scale, power_out, static_cast<Out*>(out));
    PYFFT_LAUNCH_OK("any_unpack_kernel");
    return PYFFT_OK;
}

}  // namespace pyfft

using namespace pyfft;

extern "C" {

int64_t pyfft_anyfft_workspace_bytes(int64_t n, int64_t rows) {
    if (n < 1 || rows < 1) return fail(PYFFT_EINVAL, "anyfft_workspace_bytes: bad arguments");
    if (n > (int64_t(1) << 26)) return fail(PYFFT_EUNSUPPORTED, "anyfft: n = %lld exceeds 2^26", (long long)n);
    const int64_t M = any_pow2(n);
    return ((n * 16 + 255) / 256) * 256 + M * 16 + rows * M * 16 + pyfft_bigfft_workspace_bytes(PYFFT_F64, M, (int32_t)rows);
}

int pyfft_fft_any(int dtype, int32_t complex_in, const void* in, int64_t row_stride, int64_t n, int32_t rows,
                  int32_t inverse, void* out, void* workspace, int64_t workspace_bytes, void* stream) {
    if (!in || !out) return fail(PYFFT_EINVAL, "fft_any: NULL pointer");
    if (rows > 65535) return fail(PYFFT_EINVAL, "fft_any: at most 65535 rows per call");
    if (dtype != PYFFT_F32 && dtype != PYFFT_F64) return fail(PYFFT_EINVAL, "unknown dtype %d", dtype);
    AnyWs w;
    int rc = any_carve(n, rows, workspace, workspace_bytes, &w);
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if ((rc = any_prepare(n, w, st))) return rc;
    int64_t gx = (w.M + 255) / 256;
    if (gx > 64) gx = 64;
    const dim3 grid((unsigned)gx, (unsigned)rows);
    if (dtype == PYFFT_F32) {
        if (complex_in) any_pack_rows_kernel<float, true><<<grid, 256, 0, st>>>(static_cast<const float*>(in), row_stride, n, w.M, inverse, w.chirp, w.work);
        else any_pack_rows_kernel<float, false><<<grid, 256, 0, st>>>(static_cast<const float*>(in), row_stride, n, w.M, inverse, w.chirp, w.work);
    } else {
        if (complex_in) any_pack_rows_kernel<double, true><<<grid, 256, 0, st>>>(static_cast<const double*>(in), row_stride, n, w.M, inverse, w.chirp, w.work);
        else any_pack_rows_kernel<double, false><<<grid, 256, 0, st>>>(static_cast<const double*>(in), row_stride, n, w.M, inverse, w.chirp, w.work);
    }
    PYFFT_LAUNCH_OK("any_pack_rows_kernel");
    if ((rc = any_convolve(rows, w, st))) return rc;
    const double scale = inverse ? 1.0 / (double)n : 1.0;
    if (dtype == PYFFT_F32) return any_unpack<float>(w, n, rows, inverse, scale, 0, n, nullptr, 0, out, st);
    return any_unpack<double>(w, n, rows, inverse, scale, 0, n, nullptr, 0, out, st);
}

int pyfft_laplace_1d(int dtype, const void* u, int64_t n, const double* sigma, int32_t nsigma, void* out,
                     void* workspace, int64_t workspace_bytes, void* stream) {
    if (!u || !sigma || !out) return fail(PYFFT_EINVAL, "laplace_1d: NULL pointer");
    if (nsigma > 65535) return fail(PYFFT_EINVAL, "laplace_1d: at most 65535 rows per call");
    if (dtype != PYFFT_F32 && dtype != PYFFT_F64) return fail(PYFFT_EINVAL, "unknown dtype %d", dtype);
    AnyWs w;
    int rc = any_carve(n, nsigma, workspace, workspace_bytes, &w);
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if ((rc = any_prepare(n, w, st))) return rc;
    int64_t gx = (w.M + 255) / 256;
    if (gx > 64) gx = 64;
    const dim3 grid((unsigned)gx, (unsigned)nsigma);
    if (dtype == PYFFT_F32) any_pack_laplace_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(u), n, w.M, sigma, w.chirp, w.work);
    else any_pack_laplace_kernel<double><<<grid, 256, 0, st>>>(static_cast<const double*>(u), n, w.M, sigma, w.chirp, w.work);
    PYFFT_LAUNCH_OK("any_pack_laplace_kernel");
    if ((rc = any_convolve(nsigma, w, st))) return rc;
    const int64_t nb = n / 2 + 1;                                   // numpy.fft.rfft
    if (dtype == PYFFT_F32) return any_unpack<float>(w, n, nsigma, 0, 1.0, 0, nb, nullptr, 0, out, st);
    return any_unpack<double>(w, n, nsigma, 0, 1.0, 0, nb, nullptr, 0, out, st);
}

int pyfft_stft_any(int dtype, int32_t complex_in, const void* x, int64_t ch_stride, int32_t nch,
                   const pyfft_record_map* map, const pyfft_segmentation* seg, const double* window,
                   const double* coef, int32_t detrend_per_segment, int32_t bin_first, int32_t nbins,
                   const double* bin_scale, int32_t power_out, void* out, double* pseg, void* workspace,
                   int64_t workspace_bytes, void* stream) {
    RecMap m;
    int rc = check_map(map, &m);
    if (rc) return rc;
    if (!seg || seg->nfft < 2 || seg->hop <= 0 || seg->seg_first < 0 || seg->seg_count < 0)
        return fail(PYFFT_EINVAL, "stft_any: bad segmentation");
    if (!x || !window || !out || nch <= 0) return fail(PYFFT_EINVAL, "stft_any: NULL pointer or nch <= 0");
    if (nbins <= 0 || nbins > seg->nfft) return fail(PYFFT_EINVAL, "stft_any: nbins must be in 1..nfft");
    if (dtype != PYFFT_F32 && dtype != PYFFT_F64) return fail(PYFFT_EINVAL, "unknown dtype %d", dtype);
    if (seg->seg_count == 0) return PYFFT_OK;
    const int64_t rows = (int64_t)nch * seg->seg_count;
    if (rows > 65535 || nch > 65535) return fail(PYFFT_EINVAL, "stft_any: at most 65535 rows (channels x segments) per call");
    const int64_t j_lo = seg->seg_first * (int64_t)seg->hop;
    const int64_t j_hi = (seg->seg_first + seg->seg_count - 1) * (int64_t)seg->hop + seg->nfft;
    if ((rc = check_range(m, j_lo, j_hi))) return rc;
    const int64_t n = seg->nfft;
    AnyWs w;
    if ((rc = any_carve(n, rows, workspace, workspace_bytes, &w))) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if ((rc = any_prepare(n, w, st))) return rc;
    const dim3 grid((unsigned)seg->seg_count, (unsigned)nch);
#define PYFFT_ANY_PACK(IN, CP)                                                                                      \
    any_pack_segments_kernel<IN, CP><<<grid, 256, 0, st>>>(static_cast<const IN*>(x), ch_stride, m, seg->hop,      \
                                                           seg->seg_first, seg->seg_count, n, w.M, window, coef,   \
                                                           detrend_per_segment, w.chirp, w.work, pseg)
    if (dtype == PYFFT_F32) { if (complex_in) PYFFT_ANY_PACK(float, true); else PYFFT_ANY_PACK(float, false); }
    else { if (complex_in) PYFFT_ANY_PACK(double, true); else PYFFT_ANY_PACK(double, false); }
#undef PYFFT_ANY_PACK
    PYFFT_LAUNCH_OK("any_pack_segments_kernel");
    if ((rc = any_convolve(rows, w, st))) return rc;
    if (dtype == PYFFT_F32) return any_unpack<float>(w, n, rows, 0, 1.0, bin_first, nbins, bin_scale, power_out, out, st);
    return any_unpack<double>(w, n, rows, 0, 1.0, bin_first, nbins, bin_scale, power_out, out, st);
}

}  // extern "C"
