/*
 * pyfft_b200 -- C ABI of the B200-native accelerator for PYFFT's segmented
 * spectral-estimation hot path (Welch / STFT / correlation / Hilbert).
 *
 * The reference (gmweir/PYFFT) has NO FFI, plugin or operator interface: the hot
 * path sits behind plain Python callables and bottoms out in numpy.fft call
 * sites inside per-segment Python loops.  Each entry point below therefore
 * replaces one of those loops *together with* the numpy.fft calls it feeds; the
 * "replaces" note on each function cites the reference lines (paths relative to
 * the reference tree).  INTEGRATION.md shows the ctypes binding a maintainer of
 * the reference would add.
 *
 * Conventions
 *  - every function returns 0 on success or a negative pyfft_status code;
 *    pyfft_last_error() returns a thread-local human readable message.
 *  - all data pointers are DEVICE pointers owned by the caller; the library
 *    allocates nothing persistent except cached twiddle tables (freed by
 *    pyfft_shutdown()).  Inputs are const and never written.
 *  - every call is asynchronous on `stream` (a cudaStream_t passed as void*);
 *    the library never synchronises the device.
 *  - no CPU fallback exists: unsupported shapes return PYFFT_EUNSUPPORTED.
 */
#ifndef PYFFT_B200_H
#define PYFFT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum pyfft_status {
    PYFFT_OK = 0,
    PYFFT_EINVAL = -1,       /* bad argument (maps to ValueError)              */
    PYFFT_EUNSUPPORTED = -2, /* valid but not implemented (NotImplementedError) */
    PYFFT_ECUDA = -3         /* CUDA runtime error (RuntimeError)               */
} pyfft_status;

typedef enum pyfft_dtype { PYFFT_F32 = 0, PYFFT_F64 = 1 } pyfft_dtype;

/* detrend modes of fftanal._detrend_func (fft_analysis.py:2538-2549) */
typedef enum pyfft_detrend { PYFFT_DETREND_NONE = 0, PYFFT_DETREND_MEAN = 1,
                             PYFFT_DETREND_LINEAR = 2 } pyfft_detrend;

/*
 * How the "analysed record" x_in[j], j = 0 .. nsig-1 (fft_analysis.py:353-354)
 * maps onto the caller's sample arrays.  The reference materialises
 * reflect-padded copies (fft_analysis.py:196-205); here the map is evaluated in
 * the gather, no copy is made:
 *     g(j) = (j < reflect) ? (reflect - j) : (j - reflect + i0)      global sample index
 *     element read = ptr[g(j) - src_base]
 * src_base lets one rank hold only its slice of a long record (multi-GPU).
 */
typedef struct pyfft_record_map {
    int64_t nsig;      /* length of the analysed record                          */
    int64_t i0;        /* first analysed sample (fft_analysis.py:143)            */
    int64_t reflect;   /* nwins-1 when the reference reflect-pads, else 0        */
    int64_t src_base;  /* global sample index of element 0 of the given pointers */
    int64_t src_len;   /* elements per channel reachable behind the pointers     */
} pyfft_record_map;

/* Segmentation (fft_analysis.py:189-221, 359-364): segment k covers analysed
 * indices [k*hop, k*hop + nfft). */
typedef struct pyfft_segmentation {
    int32_t nfft;       /* = nwins; power of two, 16 .. 8192                     */
    int32_t hop;        /* nwins - noverlap                                      */
    int64_t seg_first;  /* first segment handled by this call (shard offset)     */
    int64_t seg_count;  /* number of segments handled by this call               */
} pyfft_segmentation;

const char* pyfft_version(void);
const char* pyfft_last_error(void);
/* frees cached twiddle tables on every device */
int pyfft_shutdown(void);
/* number of SMs of the current device (grid sizing / reporting) */
int pyfft_sm_count(void);

/* ------------------------------------------------------------------------- *
 * K0  detrend statistics
 * replaces: pybaseutils detrend_mean / detrend_linear applied to the whole
 *           analysed record (fft_analysis.py:356-357, 2147-2149).
 * For each of the nch channels (channel c at ptr + c*ch_stride) ADDS
 *     stats[2*c+0] += sum_j x(j)      stats[2*c+1] += sum_j j*x(j)
 * over analysed indices j in [j_begin, j_end).  stats is double[2*nch] and must
 * be zeroed by the caller (it is a running sum so shards can be all-reduced).
 * ------------------------------------------------------------------------- */
int pyfft_detrend_stats(int dtype, const void* ptr, int64_t ch_stride, int32_t nch,
                        const pyfft_record_map* map, int64_t j_begin, int64_t j_end,
                        double* stats, void* stream);

/* The same sums for x and the nch y channels in ONE launch: stats = double[2*(nch+1)], x first
 * (the layout pyfft_welch_accumulate_mean and pyfft_welch_mean_correct use). */
int pyfft_detrend_stats_xy(int dtype, const void* x, const void* y, int64_t y_stride, int32_t nch,
                           const pyfft_record_map* map, int64_t j_begin, int64_t j_end,
                           double* stats, void* stream);

/* Turns the (all-reduced) stats into per-channel trend coefficients
 *     coef[2*c+0] = a, coef[2*c+1] = b   with   detrended(j) = x(j) - (a + b*j)
 * for mode MEAN (b = 0) or LINEAR (least-squares line over j = 0..nsig-1). */
int pyfft_detrend_coef(int mode, const double* stats, int32_t nch, int64_t nsig,
                       double* coef, void* stream);

/* ------------------------------------------------------------------------- *
 * K1  fused Welch auto/cross accumulation
 * replaces: the hot loop of fft_pwelch (fft_analysis.py:359-393) -- segment
 *           gather, window multiply, np.fft.fft per segment, X conj(X),
 *           Y conj(Y), Y conj(X) -- and the segment sum of :444-446, without
 *           materialising any per-segment spectrum.
 * For each pair p = 0..nch-1 the REAL signals u = x, v = y[p] (y channel p at
 * y + p*y_stride) are packed as u + i v into ONE complex FFT per segment and
 * split by Hermitian symmetry in registers.  Running sums are ADDED to
 *     acc[p][0][f] += sum_k |U_k[f]|^2          acc[p][1][f] += sum_k |V_k[f]|^2
 *     acc[p][2][f] += Re sum_k V_k[f] conj(U_k[f])   acc[p][3][f] += Im ...
 * for f = 0 .. nfft/2 (nfft/2+1 bins, unscaled), acc = double[nch][4][nfft/2+1].
 * coef_x (double[2]) / coef_y (double[2*nch]) are the trend coefficients from
 * pyfft_detrend_coef or NULL.  window = nfft values of `dtype`.
 * workspace: pyfft_welch_workspace_bytes() bytes of scratch.
 * ------------------------------------------------------------------------- */
int64_t pyfft_welch_workspace_bytes(int dtype, const pyfft_segmentation* seg, int32_t nch);

int pyfft_welch_accumulate(int dtype, const void* x, const void* y, int64_t y_stride,
                           int32_t nch, const pyfft_record_map* map,
                           const pyfft_segmentation* seg, const void* window,
                           const double* coef_x, const double* coef_y,
                           double* acc, void* workspace, int64_t workspace_bytes,
                           void* stream);

/* ------------------------------------------------------------------------- *
 * K1, mean detrend fused (one pass over HBM instead of K0 + K1)
 * replaces: the same lines as pyfft_welch_accumulate plus the global mean removal
 *           of fft_analysis.py:356-357, without a separate statistics pass.
 * est = double[2*(nch+1)] (a, b) pairs, x first then the y channels, as written by
 * pyfft_detrend_coef from ANY cheap estimate of the means (b ignored).  The kernel
 * accumulates the spectra of (sample - est) into acc and, by linearity of the FFT,
 * everything needed to move them to the exact global mean afterwards:
 *     lin   = double[nch][4][nfft/16]   += sum_k U_k, V_k (re, im) for bins f < nfft/16
 *     stats = double[2*(nch+1)]         stats[2c] += raw sum of the samples of analysed
 *             indices [seg_first*hop, (seg_first+seg_count)*hop)  (same layout as
 *             pyfft_detrend_stats, so shards and the uncovered tail add into one buffer)
 * pyfft_welch_mean_correct then applies, with delta_c = stats[2c]/nana - est[2c] and
 * W = FFT(window)[0 .. nfft/16) (wlow, double[nfft/16][2]):
 *     sum|U|^2 += -2 dx Re(W* sumU) + K dx^2 |W|^2         (same for V with dy)
 *     sumVU*   += -dx W* sumV - dy W (sumU)* + K dx dy |W|^2
 * K = number of segments accumulated in THIS acc (apply per shard, before the
 * all-reduce of acc; stats must already hold the whole record's sums).  Bins
 * f >= nfft/16 are left alone: the caller must check that |W[f]|^2 is negligible
 * there (true for every cosine-sum window; the Python host checks and otherwise
 * uses the two-pass route).
 * ------------------------------------------------------------------------- */
int pyfft_welch_accumulate_mean(int dtype, const void* x, const void* y, int64_t y_stride,
                                int32_t nch, const pyfft_record_map* map,
                                const pyfft_segmentation* seg, const void* window,
                                const double* est, double* acc, double* lin, double* stats,
                                void* workspace, int64_t workspace_bytes, void* stream);

int pyfft_welch_mean_correct(double* acc, const double* lin, int32_t nfft, int32_t nch,
                             int64_t K, const double* est, const double* stats, int64_t nana,
                             const double* wlow, void* stream);

/* Sharded records with ONE collective: every rank moves its sums from its
 * own provisional means to the common reference 0 (acc and lin become the sums of the raw
 * windowed spectra for bins f < nfft/16; K = segments in THIS acc), the ranks all-reduce
 * acc | lin | stats together, and pyfft_welch_finalize_mean is called with est = 0 and the
 * global K.  Replaces the pair (all-reduce stats, pyfft_welch_mean_correct per rank). */
int pyfft_welch_rereference(double* acc, double* lin, int32_t nfft, int32_t nch, int64_t K,
                            const double* est, const double* wlow, void* stream);

/* ------------------------------------------------------------------------- *
 * pyfft_welch_accumulate without any allocation.
 * No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments, meaning and result as pyfft_welch_accumulate, minus the workspace: the per-CTA partial
 * sums AND the twiddle table (computed on the device by every call) live in one file-scope
 * `static __device__` array of the library (16 MiB + 128 KiB; every GPU that loads the library holds it
 * until exit).  The call performs no cudaMalloc / cudaFree and uses no allocating library routine.
 * Limit: calls that would need more than pyfft_welch_static_scratch_bytes() = 16 MiB of partial sums
 * (pyfft_welch_workspace_bytes(dtype, seg, nch) tells) return PYFFT_EUNSUPPORTED with a message; config 2
 * (fp32, nfft 4096, one y channel: 10.9 MB on 148 SMs) fits, fp64 at nfft >= 4096 does not.
 * All calls share the array and it keeps the previous call's contents (each call overwrites what it
 * reads): calls that overlap on the GPU -- different streams, or different host threads -- race; issue
 * them on one stream.
 * ------------------------------------------------------------------------- */
int64_t pyfft_welch_static_scratch_bytes(void);

int pyfft_welch_accumulate_static(int dtype, const void* x, const void* y, int64_t y_stride, int32_t nch,
                                  const pyfft_record_map* map, const pyfft_segmentation* seg,
                                  const void* window, const double* coef_x, const double* coef_y,
                                  double* acc, void* stream);

/* ------------------------------------------------------------------------- *
 * K1-TC, the same accumulation on the tcgen05 tensor cores (float32 input,
 * nfft = 4096, hop 512, 1024, 2048 or 4096)
 * replaces: the same reference lines as pyfft_welch_accumulate_mean.
 * Every segment's 4096-point DFT is computed as two real DFT-matrix GEMMs
 * (64 x 64 decomposition, fp16 operands split hi + lo for the matrices, fp32
 * accumulation in tensor memory); the detrended, windowed samples are rounded to
 * fp16 ONCE after a power-of-two prescale, so every segment carries an independent
 * relative error of ~2^-11/sqrt(nfft) that averages down with the number of
 * segments.  The Python host therefore selects this path only for long averages
 * (Navr >= 16384, where it meets the 1e-5-of-peak bar) and the CUDA-core kernel
 * otherwise.  acc, lin, stats: exactly as pyfft_welch_accumulate_mean (lin holds
 * bins 0 .. 7 only, the rest of each row is zero); finish with
 * pyfft_welch_mean_correct and pyfft_welch_finalize.
 *   pyfft_welch_tc_prepare: est (pairs, as above; the mean is rounded to float so
 *     that the kernel's subtraction is the one the correction assumes) and
 *     prescale = double[16 (nch+1)]: [0, nch] powers of two bringing the rms of each
 *     channel to [1/8, 1/4), from 256 runs of 256 samples spread evenly over the analysed
 *     indices [j_begin, j_end); [nch+1, 2 nch+1] the repetition measure of each channel,
 *     min over l = 1..8 of sum (x[j + l hop] - x[j])^2 / (2 sum (x - mean)^2) over the same
 *     runs (1 when the span is shorter than 16 hops): a record that repeats exactly after a
 *     few hops rounds to fp16 identically segment after segment and the errors do not
 *     average out -- the host takes the tensor-core path only above 1e-4;
 *     [2 (nch+1), 3 nch+2] the stationarity of each channel, (sum_b v_b)^2 / (256 sum_b v_b^2)
 *     over the variances v_b of the 256 runs (1 = stationary; the effective number of segments
 *     is that fraction of Navr) -- the host wants >= 1/2; the rest is scratch.
 *   status_flag (device int32, zeroed by the caller): bit 0 is set if a tensor-core
 *     completion barrier timed out, bit 1 if a partial sum is not finite (an fp16
 *     operand overflowed: the record is far louder somewhere than the sampled
 *     statistics said); the result is then invalid and the host reruns on
 *     pyfft_welch_accumulate_mean.
 * ------------------------------------------------------------------------- */
int64_t pyfft_welch_tc_workspace_bytes(const pyfft_segmentation* seg, int32_t nch);

int pyfft_welch_tc_prepare(const void* x, const void* y, int64_t y_stride, int32_t nch,
                           const pyfft_record_map* map, int64_t j_begin, int64_t j_end, int64_t hop,
                           double* est, double* prescale, void* stream);

int pyfft_welch_accumulate_tc(const void* x, const void* y, int64_t y_stride, int32_t nch,
                              const pyfft_record_map* map, const pyfft_segmentation* seg,
                              const void* window, const double* est, const double* prescale,
                              double* acc, double* lin, double* stats, void* workspace,
                              int64_t workspace_bytes, int32_t* status_flag, void* stream);

/* The whole fused tensor-core step of a resident record in THREE launches (+ one 168 (nch+1)-byte fill):
 *   tc_pre_kernel   = pyfft_welch_tc_prepare over the segments' own span + the reflect-padded head + the raw
 *                     sums of the analysed samples outside that span (tail_a = [a0, a1), tail_b = [b0, b1):
 *                     what pyfft_detrend_stats_xy adds for the exact global mean, fft_analysis.py:356-357),
 *   welch_tc_kernel = the main kernel of pyfft_welch_accumulate_tc,
 *   tc_post_kernel  = sum of the per-CTA partials -> [sharded: rereference + exchange over NVLink peer memory]
 *                     -> state and the finalised spectra of pyfft_welch_finalize_mean, 16 bins per CTA.
 * Same results as prepare + accumulate_tc + detrend_stats_xy (+ rereference + peer_allreduce) + finalize_mean;
 * the step is launch-bound around the main kernel, so the boundaries between the small kernels are what it saves.
 *   aux   = double[21 (nch+1)], written: est (2 per channel) | prescale (16 per channel, as pyfft_welch_tc_prepare)
 *           | tail sums (2 per channel) | raw sums of the segments' span (1 per channel).
 *   state = acc [nch][4][2049] | lin [nch][4][256] | stats [2 (nch+1)], WRITTEN (not added to): the sums of this call,
 *           summed over the ranks when sharded (then about the reference 0: mean correction with est = 0).
 *   wlow  = the window's transform, bins 0..7 at least, (re, im) pairs (double).
 *   K_total / nana: segments and analysed samples of the WHOLE record (all ranks); seg = this rank's segments (>= 1).
 *   world > 1: peer_base = device array of `world` addresses of the ranks' peer buffers (each
 *           pyfft_welch_tc_step_peer_bytes(nch, world) bytes, zero-filled once, peer-mapped by the host:
 *           torch.distributed._symmetric_memory); every rank must make the same sequence of calls.
 *   status_flag bits: 1 pipeline barrier time-out, 2 non-finite partial sum, 4 a peer did not arrive (the outputs
 *           are then NaN).
 * No reference counterpart for the exchange (the reference is single-process): it is the sum over shards of the
 * per-segment sums of fft_analysis.py:444-446. */
int64_t pyfft_welch_tc_step_peer_bytes(int32_t nch, int32_t world);

int pyfft_welch_tc_step(const void* x, const void* y, int64_t y_stride, int32_t nch,
                        const pyfft_record_map* map, const pyfft_segmentation* seg, const void* window,
                        const double* wlow, int64_t tail_a0, int64_t tail_a1, int64_t tail_b0, int64_t tail_b1,
                        int64_t K_total, int64_t nana, double scale, int32_t onesided, double* aux, double* state,
                        double* Pxx, double* Pyy, double* Pxy, double* Cxy, double* Cxy2, double* phi,
                        void* workspace, int64_t workspace_bytes, int32_t* status_flag,
                        const int64_t* peer_base, int32_t rank, int32_t world, int32_t deferred, void* stream);

/* deferred != 0 (world > 1 only; a stream of steps over successive records): the exchange of a step is collected by
 * the NEXT deferred step (or by pyfft_welch_tc_step_flush), a whole main kernel after it was pushed, so that no rank
 * ever waits for a slower one inside a step.  The outputs and `state` written by a deferred call are those of the
 * PREVIOUS deferred call (nothing is written by the first one); _flush writes those of the last one.  Do not mix
 * deferred and plain calls on one peer buffer without a flush in between. */
int pyfft_welch_tc_step_flush(int32_t nch, const double* wlow, int64_t K_total, int64_t nana, double scale,
                              int32_t onesided, double* state, double* Pxx, double* Pyy, double* Pxy,
                              double* Cxy, double* Cxy2, double* phi, int32_t* status_flag,
                              const int64_t* peer_base, int32_t rank, int32_t world, void* stream);

/* ------------------------------------------------------------------------- *
 * the steps either side of the path (csrc/bands.cu)
 * pyfft_spectral_centroid replaces the reduction of Doppler.cog (Doppler.py:43-61) and of the per-window loop of
 *   cogspec (:72-81): out[row] = sum_b P[row][b] wnum[b] / sum_b P[row][b] wden[b] (0 when wden is all zero) over
 *   the window spectra P = |FFT|^2 [rows][nbins] as written by pyfft_stft with power_out (natural bin order); the
 *   host puts the signed frequencies and the fmin / fmax selection into wnum / wden (double[nbins]).
 * pyfft_resample_linear replaces filters.upsample (filters.py:20-34) behind fftanal.resample
 *   (fft_analysis.py:2363-2373): records sampled at fs_in interpolated linearly onto t_i = i / fs_out, i < m
 *   (evaluated like numpy.interp; pybaseutils.utils.interp, which the reference calls, is not in the image:
 *   parity unpinned for this entry).
 * ------------------------------------------------------------------------- */
int pyfft_spectral_centroid(int dtype, const void* P, int64_t rows, int32_t nbins, const double* wnum,
                            const double* wden, double* out, void* stream);

int pyfft_resample_linear(int dtype, const void* in, int64_t in_stride, int32_t nch, int64_t n, double fs_in,
                          double fs_out, int64_t m, void* out, int64_t out_stride, void* stream);

/* ------------------------------------------------------------------------- *
 * per-segment statistics as fused reductions (csrc/segstats.cu)
 * replaces: the reductions a caller makes over the per-segment spectra the reference materialises
 *           (Pxx_seg, Pyy_seg, Pxy_seg, phixy_seg: fft_analysis.py:397-460, :1955-1970): np.var over the
 *           segments -- the "real" variance estimates the reference comments out in favour of the stub
 *           (P/sqrt(Navr))^2 (:449-458, :503-505) -- and mean_angle(phixy_seg) (:1334-1377).
 * One pass over the records (the structure of pyfft_welch_accumulate: detrend coefficients, window, two real
 * channels per complex FFT), nothing per-segment is written.
 * out = double[nch][12][nfft/2 + 1], sums over the segments of, per bin of the half spectrum,
 *   0..3  Pxx, Pyy, Re Pxy, Im Pxy   (Pxy = Y conj(X), unscaled products of the windowed detrended segments)
 *   4..7  Pxx^2, Pyy^2, (Re Pxy)^2, (Im Pxy)^2
 *   8..11 cos phi, sin phi, cos^2 phi, sin^2 phi,  phi = angle(Pxy) of the segment (angle(0) = 0).
 * nfft = 64 .. 8192 (power of two); workspace from pyfft_segment_moments_workspace_bytes.
 * ------------------------------------------------------------------------- */
int64_t pyfft_segment_moments_workspace_bytes(int dtype, const pyfft_segmentation* seg, int32_t nch);

int pyfft_segment_moments(int dtype, const void* x, const void* y, int64_t y_stride, int32_t nch,
                          const pyfft_record_map* map, const pyfft_segmentation* seg, const void* window,
                          const double* coef_x, const double* coef_y, double* out, void* workspace,
                          int64_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------- *
 * Welch finalize
 * replaces: one-/two-sided selection and doubling (fft_analysis.py:402-428),
 *           scaling by 1/S1^2/ENBW (:432-440), the mean over segments
 *           (:444-446), Cxy_Cxy2 (:1662-1680) and phi_xy (:520).
 * acc as written by pyfft_welch_accumulate (after any all-reduce), K = total
 * number of segments summed, scale = 1/(S1^2 * ENBW) (the reference's constants).
 * onesided != 0: F = nnyq bins [0, nnyq) with x2 on bins 1..F-2 (and F-1 when
 *                nfft is odd -- never for the power-of-two sizes here);
 * onesided == 0: F = nfft bins in fftshift order, negative frequencies filled
 *                from Hermitian symmetry of the real inputs.
 * Outputs (device, complex128 = double[2]): Pxx[F], Pyy[F][nch], Pxy[F][nch],
 * Cxy[F][nch], Cxy2[F][nch]; phi[F][nch] double.  Any output may be NULL.
 * ------------------------------------------------------------------------- */
int pyfft_welch_finalize(const double* acc, int32_t nfft, int32_t nch, int64_t K,
                         double scale, int32_t onesided,
                         double* Pxx, double* Pyy, double* Pxy,
                         double* Cxy, double* Cxy2, double* phi, void* stream);

/* pyfft_welch_mean_correct + pyfft_welch_finalize in one launch for the unsharded case: the
 * correction is applied while acc is read (acc is left as accumulated); arguments as in the two
 * functions, K = number of segments in acc. */
int pyfft_welch_finalize_mean(const double* acc, const double* lin, int32_t nfft, int32_t nch,
                              int64_t K, const double* est, const double* stats, int64_t nana,
                              const double* wlow, double scale, int32_t onesided,
                              double* Pxx, double* Pyy, double* Pxy, double* Cxy, double* Cxy2,
                              double* phi, void* stream);

/* Dynamic range of the averaged auto-spectra, the host's data-aware check of the tensor-core path
 * (no reference line: the reference computes in float64 throughout).  The fp16 operands of
 * pyfft_welch_accumulate_tc leave a white error floor of ~1e-7 of the MEAN bin power in every bin; the
 * coherence (fft_analysis.py:1662-1680) of a bin whose own power is not far above that floor cannot meet
 * the 1e-4 contract.  out[2c] = min over bins, out[2c+1] = mean over bins of Re Pxx (c = 0) / Re Pyy[:, c-1]
 * (c = 1..nch), F = nfreq bins as written by pyfft_welch_finalize; the caller reruns on the CUDA-core
 * kernel when min < threshold * mean (pyfft_b200.fft_analysis.TC_DYNAMIC_RANGE / Navr). */
int pyfft_welch_dynamic_range(const double* Pxx, const double* Pyy, int32_t nfreq, int32_t nch,
                              int32_t onesided, double* out, void* stream);

/* ------------------------------------------------------------------------- *
 * K3a  correlation epilogue
 * replaces: fft_analysis.py:544-597 (and fftanal.crosscorr :1840-1878): the
 *           irfft / ifft of the averaged spectra, sqrt(nfft) scaling and
 *           fftshift.  spec = complex128 [F][ncol] (row-major, as produced by
 *           pyfft_welch_finalize), out = [nfft][ncol]: double when onesided
 *           (irfft), complex128 when two-sided (ifft(ifftshift)).
 * halve_interior != 0 applies the x0.5 on bins 1..F-2 the reference uses for
 * Rxx/Ryy/Rxy (not for iCxy).  Output is sqrt(nfft)*inverse, fftshift-ed.
 * ------------------------------------------------------------------------- */
int pyfft_welch_correlation(const double* spec, int32_t nfft, int32_t ncol,
                            int32_t onesided, int32_t halve_interior,
                            double* out, void* stream);

/* ------------------------------------------------------------------------- *
 * K2  short-time Fourier transform (per-segment spectra written out)
 * replaces: fftanal.fft_win (fft_analysis.py:2126-2203) -- the loop behind
 *           Xstft/Ystft and spectrogram.stft -- and spectrogram.specgram's loop
 *           (spectrogram.py:91-112).
 * Channel c (at x + c*ch_stride, real or interleaved complex samples) and
 * segment k produce nbins values
 *     out[c][k][b] = bin_scale[b] * FFT(window * detrended segment)[bin_first + b (mod nfft)]
 * bin_scale (double[nbins]) carries the reference's sqrt(2)/S1/sqrt(ENBW)
 * factors (or NULL = 1).  power_out != 0 writes |.|^2 * bin_scale[b] as a real
 * array instead (specgram).  Output type follows `dtype` (complex64/128, or float/double
 * for power_out), layout [c][k][b].  Real input packs two consecutive segments into one
 * complex transform.  complex_in: coef holds the real/imaginary means at coef[4c], coef[4c+2]
 * (mean detrend only).
 * pseg (optional, double[nch][seg_count]) receives
 *     sum_n g_n - (g_0 + g_{L-1})/2,  g = |window*x|^2     (trapezoid rule core of :2174)
 * detrend_per_segment != 0 removes the segment's own mean (detrendwin=True,
 * :2169-2172) instead of using coef.
 * ------------------------------------------------------------------------- */
int pyfft_stft(int dtype, int32_t complex_in, const void* x, int64_t ch_stride, int32_t nch,
               const pyfft_record_map* map, const pyfft_segmentation* seg,
               const void* window, const double* coef, int32_t detrend_per_segment,
               int32_t bin_first, int32_t nbins, const double* bin_scale,
               int32_t power_out, void* out, double* pseg, void* stream);

/* ------------------------------------------------------------------------- *
 * Spectrum post-processing on the device, float64, batched over columns.
 * pyfft_find_peaks
 *   replaces: the selection loop of getNpeaks (fft_analysis.py:985-1027): npeaks times argmax of
 *   L[nfreq][ncol] (column-wise; first occurrence on ties) followed by the removal of irem//2 surviving
 *   bins on either side -- in the reference's COMPRESSED indexing, with its literal edge rules (:1010-1017).
 *   idx[ncol][npeaks] receives the original bin of every peak, -1 once nothing survives.  nfreq <= 16384.
 * pyfft_band_integrate
 *   replaces: integratespectra's integrals (:890-903) and power-weighted frequency (:931-935): for every
 *   band [lo, hi] (bands = double[nbands][2]) and column, trapezoid integrals over the bins with
 *   lo <= freq <= hi of Pxy, Pxx, Pyy (complex128 [nfreq][ncol]; Pxx [nfreq][pxx_cols], pxx_cols 1 or ncol).
 *   out = double[nbands][ncol][8]: Re/Im Pxy_i, Re/Im Pxx_i, Re/Im Pyy_i, fweighted, bins in the band.
 * pyfft_mean_angle
 *   replaces: mean_angle (:1334-1377) along axis 0 of phi[K][ncol] (`dtype` real; vphi, vsyst optional,
 *   same shape): out = double[ncol][4] = nanmean cos, nanmean sin and their variances incl. the
 *   propagated terms; the caller finishes with varphi (:1300-1330).
 * ------------------------------------------------------------------------- */
int pyfft_find_peaks(const double* L, int32_t nfreq, int32_t ncol, int32_t npeaks, int32_t irem,
                     int32_t* idx, void* stream);

int pyfft_band_integrate(const double* freq, int32_t nfreq, const double* Pxy, const double* Pxx,
                         int32_t pxx_cols, const double* Pyy, int32_t ncol, const double* bands,
                         int32_t nbands, double* out, void* stream);

int pyfft_mean_angle(int dtype, const void* phi, const void* vphi, const void* vsyst, int64_t K,
                     int64_t ncol, double* out, void* stream);

/* ------------------------------------------------------------------------- *
 * Transforms of ARBITRARY length (Bluestein's chirp-z on the power-of-two kernels, float64 arithmetic)
 * replaces: np.fft.fft in fft_pwelch (fft_analysis.py:375-388) and fftanal.fft_win (:2152-2176) when
 *           nwins is not a power of two -- the reference never zero-pads (:219-221) and derives nwins
 *           from Navr (:185-189, :2411-2418), default Navr = 8 -- for complex-valued inputs of
 *           fft_pwelch (:123-127, :422-428), and np.fft.fft / ifft of hilbert.py:44-66 for any n.
 * pyfft_stft_any: the contract of pyfft_stft for any seg->nfft >= 2 (real or interleaved complex
 *   samples of `dtype`; window = double[nfft]; coef as pyfft_stft; out of `dtype`), at most 65535 rows
 *   (nch x seg_count) per call; workspace = pyfft_anyfft_workspace_bytes(nfft, nch * seg_count).
 * pyfft_fft_any: numpy.fft.fft / ifft (inverse != 0, includes 1/n) of `rows` rows of length n at
 *   in + r * row_stride (elements; complex elements when complex_in), out complex [rows][n] of `dtype`.
 * ------------------------------------------------------------------------- */
int64_t pyfft_anyfft_workspace_bytes(int64_t n, int64_t rows);

/* laplace.laplace_1d (laplace.py:85-113): out[r][k] = rfft over m of
 * u[m] exp(sigma_r (n-1-m)) / sum_j exp(sigma_r j), k = 0 .. n/2, for every damping sigma_r -- the
 * exponential weights are applied while the rows are packed; any n; out complex [nsigma][n/2+1] of
 * `dtype`; workspace = pyfft_anyfft_workspace_bytes(n, nsigma); nsigma <= 65535 per call. */
int pyfft_laplace_1d(int dtype, const void* u, int64_t n, const double* sigma, int32_t nsigma,
                     void* out, void* workspace, int64_t workspace_bytes, void* stream);

int pyfft_fft_any(int dtype, int32_t complex_in, const void* in, int64_t row_stride, int64_t n,
                  int32_t rows, int32_t inverse, void* out, void* workspace, int64_t workspace_bytes,
                  void* stream);

int pyfft_stft_any(int dtype, int32_t complex_in, const void* x, int64_t ch_stride, int32_t nch,
                   const pyfft_record_map* map, const pyfft_segmentation* seg, const double* window,
                   const double* coef, int32_t detrend_per_segment, int32_t bin_first, int32_t nbins,
                   const double* bin_scale, int32_t power_out, void* out, double* pseg,
                   void* workspace, int64_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------- *
 * Spectral products of materialised STFTs (the non-fused path; used for
 * complex inputs and for the fftanal class path that keeps *_seg arrays)
 * replaces: fftanal.Pstft + averagewins means (fft_analysis.py:1944-1990) and
 *           fft_pwelch :391-393/:444-446 for complex input.
 * X[K][nb], Y[nch][K][nb] complex of `dtype`; conj_x != 0 gives Y conj(X)
 * (fft_pwelch convention), else X conj(Y) (fftanal.Pstft convention).
 * Writes double sums (not means): sxx[nb], syy[nch][nb], sxy[nch][nb][2].
 * ------------------------------------------------------------------------- */
int pyfft_stft_products(int dtype, const void* X, const void* Y, int32_t nch, int64_t K,
                        int32_t nb, int32_t conj_x, double* sxx, double* syy, double* sxy,
                        void* stream);

/* ------------------------------------------------------------------------- *
 * K3b  long transforms: analytic signal and FFT cross-correlation
 * replaces: hilbert.hilbert / hilbert_1d (hilbert.py:22-112) and ccf.ccf /
 *           fftcorr / cross_correlation_fft (ccf.py:66-77, 239-303).
 * n must be a power of two for hilbert (2^4 .. 2^27); xcorr zero-pads to the
 * next power of two >= 2n-1 internally.
 * pyfft_hilbert: in = batch real records [batch][n]; analytic (complex, may be
 *   NULL) and/or envelope (|z|, real, may be NULL) of `dtype`.
 * pyfft_xcorr: for each pair p: out[p][0 .. 2n-2] = correlate(x_p - mean, y_p - mean, 'full')
 *   * norm_p (numpy.correlate lag order, ccf.py:69-74).  normalise = 1: means removed and
 *   norm_p = 1/(n*std_x*std_y) (ccf.ccf, ccf.py:74-76); normalise = 2: means removed,
 *   norm_p = 1; normalise = 0: raw correlation, no mean removal (fftcorr /
 *   cross_correlation_fft, ccf.py:239-303).  x_p = x + p*x_stride (x_stride may be 0 to
 *   reuse one reference channel); out is real of `dtype`, [pairs][2n-1].
 *   workspace: pyfft_bigfft_workspace_bytes(dtype, L, pairs + (x_stride ? pairs : 1)) with
 *   L = the power of two >= 2n-1 (at least 16).
 * pyfft_hilbert: with a workspace of pyfft_bigfft_workspace_bytes(dtype, n, (batch + 1) / 2) bytes
 *   and batch >= 2, two records share one complex transform (half the work; the Hilbert
 *   transform is linear).  Without it the records are transformed one by one, in place in
 *   `analytic`; pyfft_bigfft_workspace_bytes(dtype, n, batch) is then needed only when analytic
 *   is NULL.  pyfft_xcorr pairs the records of y the same way when x_stride = 0.
 * ------------------------------------------------------------------------- */
int64_t pyfft_bigfft_workspace_bytes(int dtype, int64_t nfft, int32_t batch);

int pyfft_hilbert(int dtype, const void* in, int64_t n, int32_t batch,
                  void* analytic, void* envelope, void* workspace,
                  int64_t workspace_bytes, void* stream);

int pyfft_xcorr(int dtype, const void* x, int64_t x_stride, const void* y, int64_t y_stride,
                int64_t n, int32_t pairs, int32_t normalise, void* out,
                void* workspace, int64_t workspace_bytes, void* stream);

/* Spectral derivative of `batch` real records on one uniform grid
 * replaces: sig = real(ifft(wavenumber * fft(sig))), fft_analysis.py:1528-1553 (fft_deriv), with
 * wavenumber = i sin(k dx)/dx / L (modified != 0) or i k / L, k = 2 pi fftfreq(n, d = dx/L), L = n dx.
 * n a power of two (2^4 .. 2^27); in/out real [batch][n] of `dtype`; two records share one complex
 * transform; workspace: pyfft_bigfft_workspace_bytes(dtype, n, (batch + 1) / 2). */
int pyfft_spectral_derivative(int dtype, const void* in, int64_t n, int32_t batch, int32_t modified, double dx,
                              void* out, void* workspace, int64_t workspace_bytes, void* stream);

/* Batched complex FFT of power-of-two length 2^4 .. 2^27 (forward: e^{-2 pi i nk/N}
 * unnormalised; inverse: 1/N), the numpy.fft.fft/ifft contract used at every
 * reference call site (fft_analysis.py:2096-2116).  in/out complex of `dtype`,
 * [batch][n]; in may equal out. */
int pyfft_fft_c2c(int dtype, const void* in, void* out, int64_t n, int32_t batch,
                  int32_t inverse, void* workspace, int64_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------- *
 * K4  multi-channel cross-spectral matrix
 * replaces: the per-reference-channel loop over fft_pwelch in
 *           HeatPulse_Funcs.py:576-583 (G[f][i][j] = sum_k X_j conj(X_i)).
 * x = [nch][..] real channels; ADDS unscaled sums into
 * G = double[nfft/2+1][nch][nch][2].
 * ------------------------------------------------------------------------- */
int64_t pyfft_csm_workspace_bytes(int dtype, const pyfft_segmentation* seg, int32_t nch);

int pyfft_csm_accumulate(int dtype, const void* x, int64_t ch_stride, int32_t nch,
                         const pyfft_record_map* map, const pyfft_segmentation* seg,
                         const void* window, const double* coef, double* G,
                         void* workspace, int64_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------- *
 * The exchange step of the sharded path over NVLink peer memory
 * replaces: the single all-reduce of the packed partial sums acc | lin | stats
 * (the sum over shards of the segment sums of fft_analysis.py:444-446).
 * Every rank owns a buffer of pyfft_peer_buffer_bytes(n, world) bytes, zero-filled
 * once, with the SAME layout on every GPU (double staging[2][n] | uint32 flags[world])
 * and peer-mapped by the caller; peer_base is a DEVICE array of `world` 64-bit
 * addresses: where rank r's buffer is mapped in THIS process.  One call per step on
 * every rank with the same epoch = 1, 2, 3, ...: state[0..n) becomes the sum over
 * ranks, formed in rank order (bit-identical on all ranks).  epoch = 0: the epoch is
 * kept on the device in ticket[1] and advanced by the kernel itself, so that the launch
 * can be replayed from a CUDA graph (all ranks must then use this mode from the first
 * call on).  ticket: two zeroed uint32 in device memory; status_flag: device int, set to 1 if a peer did not show
 * up within the (bounded) wait -- the result is then undefined and the caller must
 * fall back to another collective.
 * ------------------------------------------------------------------------- */
int64_t pyfft_peer_buffer_bytes(int64_t n, int32_t world);
int pyfft_peer_allreduce(double* state, int64_t n, const int64_t* peer_base, int32_t rank, int32_t world,
                         uint32_t epoch, uint32_t* ticket, int32_t* status_flag, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PYFFT_B200_H */
