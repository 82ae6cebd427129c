// Entry points declared in include/pyfft_b200.h whose kernels are not written
// yet.  They fail loudly (PYFFT_EUNSUPPORTED); there is no CPU fallback.
#include "common.cuh"

using namespace pyfft;

extern "C" {

int64_t pyfft_bigfft_workspace_bytes(int, int64_t, int32_t) {
    return fail(PYFFT_EUNSUPPORTED, "long transforms (K3b) are not implemented yet");
}
int pyfft_hilbert(int, const void*, int64_t, int32_t, void*, void*, void*, int64_t, void*) {
    return fail(PYFFT_EUNSUPPORTED, "pyfft_hilbert (K3b) is not implemented yet");
}
int pyfft_xcorr(int, const void*, int64_t, const void*, int64_t, int64_t, int32_t, int32_t, void*, void*,
                int64_t, void*) {
    return fail(PYFFT_EUNSUPPORTED, "pyfft_xcorr (K3b) is not implemented yet");
}
int pyfft_fft_c2c(int, const void*, void*, int64_t, int32_t, int32_t, void*, int64_t, void*) {
    return fail(PYFFT_EUNSUPPORTED, "pyfft_fft_c2c (K3b) is not implemented yet");
}
int64_t pyfft_csm_workspace_bytes(int, const pyfft_segmentation*, int32_t) {
    return fail(PYFFT_EUNSUPPORTED, "cross-spectral matrix (K4) is not implemented yet");
}
int pyfft_csm_accumulate(int, const void*, int64_t, int32_t, const pyfft_record_map*, const pyfft_segmentation*,
                         const void*, const double*, double*, void*, int64_t, void*) {
    return fail(PYFFT_EUNSUPPORTED, "pyfft_csm_accumulate (K4) is not implemented yet");
}
}  // extern "C"
