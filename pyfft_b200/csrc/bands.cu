// Post-processing of spectra on the device: peak picking with exclusion
// windows, band integration, circular (phase) averaging over segments.
// Reference lines replaced: getNpeaks' selection loop fft_analysis.py:985-1027, integratespectra's
// trapezoid integrals and power-weighted frequency :890-903, :931-935, mean_angle :1334-1377 (with
// varphi :1300-1330).  All are batched over channels / bands / columns; arithmetic is float64.
#include "common.cuh"

namespace pyfft {

constexpr int PEAK_MAXF = 16384;

// One CTA per column of L[F][ncol].  Replicates the reference's loop LITERALLY, including that it
// compresses the arrays after every removal (indices below are ranks among the surviving bins):
//   imax = argmax(L);  h = irem // 2
//   if imax - h >= 0 and imax + h < n:   remove ranks [imax - h, imax + h)
//   elif imax + h < n:                   remove ranks [0, imax + h)
//   elif imax - h >= 0:                  remove everything   (iff[-(imax + h):] with imax + h >= n)
//   else:                                remove nothing
// idx[col][p] = ORIGINAL bin of peak p, or -1 once no bin survives (numpy raises there).
__global__ void __launch_bounds__(256) find_peaks_kernel(const double* __restrict__ L, int F, int ncol, int npeaks, int irem,
                                                         int32_t* __restrict__ idx) {
    __shared__ unsigned char alive[PEAK_MAXF];
    __shared__ double bv[8];
    __shared__ int bi[8], cnt[256], s_imax, s_rank, s_n;
    const int col = blockIdx.x, t = threadIdx.x;
    const int per = (F + 255) / 256, lo = t * per, hi = min(F, lo + per);
    for (int f = t; f < F; f += 256) alive[f] = 1;
    __syncthreads();
    const int h = irem / 2;
    for (int p = 0; p < npeaks; ++p) {
        // argmax over the survivors (first occurrence on ties, like numpy)
        double best = -1.0e308;
        int where = -1, mine = 0;
        for (int f = lo; f < hi; ++f)
            if (alive[f]) {
                ++mine;
                const double v = L[(size_t)f * ncol + col];
                if (v > best || where < 0) { if (v > best || where < 0) { best = v; where = f; } }
            }
        cnt[t] = mine;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int ow = __shfl_xor_sync(0xffffffffu, where, o);
            if (ow >= 0 && (where < 0 || ov > best || (ov == best && ow < where))) { best = ov; where = ow; }
        }
        if ((t & 31) == 0) { bv[t >> 5] = best; bi[t >> 5] = where; }
        __syncthreads();
        if (t == 0) {
            double b = bv[0];
            int w = bi[0];
            for (int k = 1; k < 8; ++k)
                if (bi[k] >= 0 && (w < 0 || bv[k] > b || (bv[k] == b && bi[k] < w))) { b = bv[k]; w = bi[k]; }
            int n = 0, r = 0;
            for (int k = 0; k < 256; ++k) {
                if (w >= 0 && k * per + per <= w) r += cnt[k];
                n += cnt[k];
            }
            s_imax = w; s_rank = r; s_n = n;
            idx[(size_t)col * npeaks + p] = w;
        }
        __syncthreads();
        const int imax = s_imax, n = s_n;
        if (imax < 0) { __syncthreads(); continue; }
        // rank of imax among the survivors: whole chunks before it (s_rank) + survivors of its own chunk before it
        int r = s_rank;
        {
            const int owner = imax / per;
            for (int f = owner * per; f < imax; ++f) r += alive[f] ? 1 : 0;
        }
        int r0, r1;                                             // ranks [r0, r1) are removed
        if (r - h >= 0 && r + h < n) { r0 = r - h; r1 = r + h; }
        else if (r + h < n) { r0 = 0; r1 = r + h; }
        else if (r - h >= 0) { r0 = 0; r1 = n; }
        else { r0 = 0; r1 = 0; }
        // exclusive prefix of the chunk counts -> rank of every survivor
        int base = 0;
        for (int k = 0; k < t; ++k) base += cnt[k];
        __syncthreads();
        int rr = base;
        for (int f = lo; f < hi; ++f)
            if (alive[f]) {
                if (rr >= r0 && rr < r1) alive[f] = 0;
                ++rr;
            }
        __syncthreads();
    }
}

// One CTA per (band, column): trapezoid integrals over the bins with lo <= freq <= hi of Pxy (complex), Pxx,
// Pyy (complex128 arrays [F][ncol]; Pxx may have one column shared by all), and the |Pxy|-weighted mean
// frequency with unit spacing (np.trapz default) -- out[band][col][8] = Re/Im Pxy_i, Re/Im Pxx_i, Re/Im Pyy_i,
// fweighted, number of bins.
__global__ void __launch_bounds__(256) band_integrate_kernel(const double* __restrict__ freq, int F, const double* __restrict__ Pxy,
                                                             const double* __restrict__ Pxx, int pxx_cols,
                                                             const double* __restrict__ Pyy, int ncol,
                                                             const double* __restrict__ bands, double* __restrict__ out) {
    const int band = blockIdx.x, col = blockIdx.y, t = threadIdx.x;
    const double lo = bands[2 * band], hi = bands[2 * band + 1];
    double s[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    const int xc = pxx_cols == 1 ? 0 : col;
    for (int f = t; f < F; f += 256) {
        const double fr = freq[f];
        if (!(fr >= lo && fr <= hi)) continue;
        s[8] += 1.0;
        const bool nxt = f + 1 < F && freq[f + 1] >= lo && freq[f + 1] <= hi;
        const bool prv = f > 0 && freq[f - 1] >= lo && freq[f - 1] <= hi;
        // trapezoid weights: half the spacing to each selected neighbour
        const double w = (nxt ? 0.5 * (freq[f + 1] - fr) : 0.0) + (prv ? 0.5 * (fr - freq[f - 1]) : 0.0);
        const double wu = (nxt ? 0.5 : 0.0) + (prv ? 0.5 : 0.0);       // unit spacing
        const size_t i = 2 * ((size_t)f * ncol + col), ix = 2 * ((size_t)f * pxx_cols + xc);
        s[0] += w * Pxy[i]; s[1] += w * Pxy[i + 1];
        s[2] += w * Pxx[ix]; s[3] += w * Pxx[ix + 1];
        s[4] += w * Pyy[i]; s[5] += w * Pyy[i + 1];
        const double a = sqrt(Pxy[i] * Pxy[i] + Pxy[i + 1] * Pxy[i + 1]);
        s[6] += wu * fr * a; s[7] += wu * a;
    }
    __shared__ double red[9][8];
#pragma unroll
    for (int q = 0; q < 9; ++q) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s[q] += __shfl_xor_sync(0xffffffffu, s[q], o);
        if ((t & 31) == 0) red[q][t >> 5] = s[q];
    }
    __syncthreads();
    if (t < 9) {
        double v = 0.0;
        for (int w = 0; w < 8; ++w) v += red[t][w];
        red[t][0] = v;
    }
    __syncthreads();
    if (t == 0) {
        double* o = out + ((size_t)band * ncol + col) * 8;
        for (int q = 0; q < 6; ++q) o[q] = red[q][0];
        o[6] = red[6][0] / red[7][0];
        o[7] = red[8][0];
    }
}

// Circular mean over the rows of phi[K][ncol] (one thread per column, coalesced across columns):
// out[col][4] = nanmean cos, nanmean sin, nanvar cos + sum vphi / K^2 + (sum sqrt(vsyst) / K)^2, same for sin
template <typename T>
__global__ void __launch_bounds__(256) mean_angle_kernel(const T* __restrict__ phi, const T* __restrict__ vphi,
                                                         const T* __restrict__ vsyst, int64_t K, int64_t ncol,
                                                         double* __restrict__ out) {
    const int64_t col = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (col >= ncol) return;
    double sc = 0, ss = 0, scc = 0, sss = 0, sv = 0, sy = 0, n = 0;
    for (int64_t k = 0; k < K; ++k) {
        const double p = (double)phi[k * ncol + col];
        if (vphi) { const double v = (double)vphi[k * ncol + col]; if (v == v) sv += v; }
        if (vsyst) { const double v = sqrt((double)vsyst[k * ncol + col]); if (v == v) sy += v; }
        if (p != p) continue;                                  // nanmean / nanvar skip NaN
        double s, c;
        sincos(p, &s, &c);
        sc += c; ss += s; scc += c * c; sss += s * s; n += 1.0;
    }
    const double mc = sc / n, ms = ss / n, kk = (double)K;
    double* o = out + col * 4;
    o[0] = mc;
    o[1] = ms;
    o[2] = (scc / n - mc * mc) + sv / (kk * kk) + (sy / kk) * (sy / kk);
    o[3] = (sss / n - ms * ms) + sv / (kk * kk) + (sy / kk) * (sy / kk);
}

// ------------------------------------------------------------------------- //
// the two steps either side of the path
// ------------------------------------------------------------------------- //
// Doppler.cog (Doppler.py:43-61): power-weighted mean frequency of every window's spectrum,
// out[row] = sum_b P[row][b] wnum[b] / sum_b P[row][b] wden[b] (0 when no bin is selected), one warp per row.
// P = |FFT|^2 of the windows as written by pyfft_stft (power_out, natural bin order); the host puts the signed
// frequencies (and the band selection of fmin / fmax) into wnum / wden.
template <typename T>
__global__ void __launch_bounds__(256) spectral_centroid_kernel(const T* __restrict__ P, int64_t rows, int nb,
                                                                const double* __restrict__ wnum,
                                                                const double* __restrict__ wden, double* __restrict__ out) {
    const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const T* p = P + row * nb;
    double num = 0.0, den = 0.0, any = 0.0;
    for (int b = lane; b < nb; b += 32) {
        const double v = (double)p[b];
        num += v * wnum[b];
        den += v * wden[b];
        any += wden[b];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        num += __shfl_xor_sync(0xffffffffu, num, o);
        den += __shfl_xor_sync(0xffffffffu, den, o);
        any += __shfl_xor_sync(0xffffffffu, any, o);
    }
    if (lane == 0) out[row] = any > 0.0 ? num / den : 0.0;
}

// filters.upsample (filters.py:20-34; fftanal.resample, fft_analysis.py:2363-2373): piecewise-linear interpolation
// of records sampled at fs_in onto the grid t_i = i / fs_out, i < m, written the way numpy.interp evaluates it
// (slope (y1 - y0) / (x1 - x0) times (x - x0), plus y0, on the abscissae x_j = j / fs_in).
template <typename T>
__global__ void __launch_bounds__(256) resample_linear_kernel(const T* __restrict__ in, int64_t in_stride, int64_t n,
                                                              double fs_in, double fs_out, int64_t m,
                                                              T* __restrict__ out, int64_t out_stride) {
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= m) return;
    const T* p = in + (int64_t)blockIdx.y * in_stride;
    const double step = 1.0 / fs_out;
    const double x = 0.0 + (double)i * step;                 // numpy.arange: start + i * step
    int64_t j = (int64_t)floor(x * fs_in);
    if (j > n - 2) j = n - 2;
    if (j < 0) j = 0;
    // the abscissae are j / fs_in: settle j so that x_j <= x < x_{j+1} in the arithmetic numpy sees
    while (j > 0 && (double)j / fs_in > x) --j;
    while (j < n - 2 && (double)(j + 1) / fs_in <= x) ++j;
    double r;
    if (n == 1) {
        r = (double)p[0];
    } else {
        const double x0 = (double)j / fs_in, x1 = (double)(j + 1) / fs_in;
        const double y0 = (double)p[j], y1 = (double)p[j + 1];
        if (x >= x1 && j == n - 2) r = y1;                      // at (or past) the last sample: numpy.interp clamps
        else r = (y1 - y0) / (x1 - x0) * (x - x0) + y0;
    }
    out[(int64_t)blockIdx.y * out_stride + i] = (T)r;
}

}  // namespace pyfft

using namespace pyfft;

extern "C" {

int pyfft_find_peaks(const double* L, int32_t nfreq, int32_t ncol, int32_t npeaks, int32_t irem, int32_t* idx, void* stream) {
    if (!L || !idx || nfreq <= 0 || ncol <= 0 || npeaks <= 0 || irem < 0) return fail(PYFFT_EINVAL, "find_peaks: bad arguments");
    if (nfreq > PEAK_MAXF) return fail(PYFFT_EUNSUPPORTED, "find_peaks: at most %d bins", PEAK_MAXF);
    find_peaks_kernel<<<ncol, 256, 0, static_cast<cudaStream_t>(stream)>>>(L, nfreq, ncol, npeaks, irem, idx);
    PYFFT_LAUNCH_OK("find_peaks_kernel");
    return PYFFT_OK;
}

int pyfft_band_integrate(const double* freq, int32_t nfreq, const double* Pxy, const double* Pxx, int32_t pxx_cols,
                         const double* Pyy, int32_t ncol, const double* bands, int32_t nbands, double* out, void* stream) {
    if (!freq || !Pxy || !Pxx || !Pyy || !bands || !out || nfreq <= 0 || ncol <= 0 || nbands <= 0 ||
        (pxx_cols != 1 && pxx_cols != ncol))
        return fail(PYFFT_EINVAL, "band_integrate: bad arguments");
    band_integrate_kernel<<<dim3(nbands, ncol), 256, 0, static_cast<cudaStream_t>(stream)>>>(freq, nfreq, Pxy, Pxx, pxx_cols, Pyy,
                                                                                          ncol, bands, out);
    PYFFT_LAUNCH_OK("band_integrate_kernel");
    return PYFFT_OK;
}

int pyfft_mean_angle(int dtype, const void* phi, const void* vphi, const void* vsyst, int64_t K, int64_t ncol, double* out,
                     void* stream) {
    if (!phi || !out || K <= 0 || ncol <= 0) return fail(PYFFT_EINVAL, "mean_angle: bad arguments");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned g = (unsigned)((ncol + 255) / 256);
    if (dtype == PYFFT_F32)
        mean_angle_kernel<float><<<g, 256, 0, st>>>(static_cast<const float*>(phi), static_cast<const float*>(vphi),
                                                    static_cast<const float*>(vsyst), K, ncol, out);
    else if (dtype == PYFFT_F64)
        mean_angle_kernel<double><<<g, 256, 0, st>>>(static_cast<const double*>(phi), static_cast<const double*>(vphi),
                                                     static_cast<const double*>(vsyst), K, ncol, out);
    else return fail(PYFFT_EINVAL, "unknown dtype %d", dtype);
    PYFFT_LAUNCH_OK("mean_angle_kernel");
    return PYFFT_OK;
}

int pyfft_spectral_centroid(int dtype, const void* P, int64_t rows, int32_t nbins, const double* wnum, const double* wden,
                            double* out, void* stream) {
    if (!P || !wnum || !wden || !out || rows <= 0 || nbins <= 0) return fail(PYFFT_EINVAL, "spectral_centroid: bad arguments");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned g = (unsigned)((rows + 7) / 8);
    if (dtype == PYFFT_F32)
        spectral_centroid_kernel<float><<<g, 256, 0, st>>>(static_cast<const float*>(P), rows, nbins, wnum, wden, out);
    else if (dtype == PYFFT_F64)
        spectral_centroid_kernel<double><<<g, 256, 0, st>>>(static_cast<const double*>(P), rows, nbins, wnum, wden, out);
    else return fail(PYFFT_EINVAL, "unknown dtype %d", dtype);
    PYFFT_LAUNCH_OK("spectral_centroid_kernel");
    return PYFFT_OK;
}

int pyfft_resample_linear(int dtype, const void* in, int64_t in_stride, int32_t nch, int64_t n, double fs_in, double fs_out,
                          int64_t m, void* out, int64_t out_stride, void* stream) {
    if (!in || !out || nch <= 0 || n <= 0 || m <= 0 || !(fs_in > 0.0) || !(fs_out > 0.0))
        return fail(PYFFT_EINVAL, "resample_linear: bad arguments");
    if ((double)(m - 1) / fs_out > (double)(n - 1) / fs_in * (1.0 + 1e-12))
        return fail(PYFFT_EINVAL, "resample_linear: the output grid runs past the last input sample");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const dim3 g((unsigned)((m + 255) / 256), (unsigned)nch);
    if (dtype == PYFFT_F32)
        resample_linear_kernel<float><<<g, 256, 0, st>>>(static_cast<const float*>(in), in_stride, n, fs_in, fs_out, m,
                                                         static_cast<float*>(out), out_stride);
    else if (dtype == PYFFT_F64)
        resample_linear_kernel<double><<<g, 256, 0, st>>>(static_cast<const double*>(in), in_stride, n, fs_in, fs_out, m,
                                                          static_cast<double*>(out), out_stride);
    else return fail(PYFFT_EINVAL, "unknown dtype %d", dtype);
    PYFFT_LAUNCH_OK("resample_linear_kernel");
    return PYFFT_OK;
}

}  // extern "C"
