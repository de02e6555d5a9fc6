// Shared helpers for the sm_100a kernels of libdctr.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/dctr.h"

namespace dctr {

constexpr int kNumSMs = 148;  // B200

// ---- per-thread error slot + launch counter (api.cu) ----
void set_error(const char* fmt, ...);
void count_launch(int n = 1);
int check_launch(const char* what);

#define DCTR_REQUIRE(cond, ...)            \
    do {                                   \
        if (!(cond)) {                     \
            dctr::set_error(__VA_ARGS__);  \
            return DCTR_EINVAL;            \
        }                                  \
    } while (0)

static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }
static inline int ceil_div(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

// ---- device helpers ----
__device__ __forceinline__ float sigmoidf_(float x) {
    // stable both sides; expf (not __expf) to stay within 1e-6 of the fp64 oracle
    if (x >= 0.f) { float e = expf(-x); return 1.f / (1.f + e); }
    float e = expf(x);
    return e / (1.f + e);
}

__device__ __forceinline__ float act_fwd(int act, float x) {
    switch (act) {
        case DCTR_ACT_TANH: return tanhf(x);
        case DCTR_ACT_RELU: return fmaxf(x, 0.f);
        case DCTR_ACT_SIGMOID: return sigmoidf_(x);
        default: return x;
    }
}
// derivative expressed through the activation OUTPUT t = act(x)
__device__ __forceinline__ float act_bwd_from_out(int act, float t) {
    switch (act) {
        case DCTR_ACT_TANH: return 1.f - t * t;
        case DCTR_ACT_RELU: return t > 0.f ? 1.f : 0.f;
        case DCTR_ACT_SIGMOID: return t * (1.f - t);
        default: return 1.f;
    }
}
// stable sigmoid cross entropy with logits (TF: max(z,0) - z*y + log1p(exp(-|z|)))
__device__ __forceinline__ float xent_logits(float z, float y) {
    return fmaxf(z, 0.f) - z * y + log1pf(expf(-fabsf(z)));
}

__device__ __forceinline__ float4 ldg_f4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace dctr
