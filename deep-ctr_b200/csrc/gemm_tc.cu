// Tensor-core GEMM entry point (tcgen05.mma kind::tf32 with 3xTF32 error compensation).
// Round-1 state: the kernel is not written yet; the entry point reports the shape as "not covered" so
// callers stay on the fp32 SIMT path of gemm.cu.  No silent fallback happens inside this library.
#include "common.cuh"

extern "C" int dctr_gemm_tc(int, int, int, int, const float*, int, const float*, int, float*, int, int, int,
                            const float*, const float*, int, const float*, int, void*) {
    dctr::set_error("gemm_tc: no shape is covered by the tcgen05 path yet");
    return DCTR_EINVAL;
}
