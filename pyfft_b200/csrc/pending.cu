// Entry points declared in include/pyfft_b200.h whose kernels are not written
// yet.  They fail loudly (PYFFT_EUNSUPPORTED); there is no CPU fallback.
#include "common.cuh"

using namespace pyfft;

extern "C" {

int64_t pyfft_csm_workspace_bytes(int, const pyfft_segmentation*, int32_t) {
    return fail(PYFFT_EUNSUPPORTED, "cross-spectral matrix (K4) is not implemented yet");
}
int pyfft_csm_accumulate(int, const void*, int64_t, int32_t, const pyfft_record_map*, const pyfft_segmentation*,
                         const void*, const double*, double*, void*, int64_t, void*) {
    return fail(PYFFT_EUNSUPPORTED, "pyfft_csm_accumulate (K4) is not implemented yet");
}
}  // extern "C"
