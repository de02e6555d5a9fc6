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

// ---- programmatic dependent launch ---------------------------------------------------------------------------------
// Every kernel of the library starts with pdl_sync(): griddepcontrol.wait blocks until the preceding kernel of the stream
// has completed and flushed (so ordering is exactly what plain stream order gives), launch_dependents then lets the
// NEXT kernel's CTAs be scheduled onto idle SMs while this one runs.  Launching through launch_k() with the
// programmatic-serialization attribute hides the ~1-2 us kernel-to-kernel launch gap of a 20-launch step; a kernel
// launched this way MUST call pdl_sync() before touching global memory.  DCTR_PDL=0 falls back to plain launches.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_sync() { pdl_wait(); pdl_launch_dependents(); }
int pdl_mode();          // 0 = off, 1 = every kernel, 2 (default) = every kernel but the tensor-core GEMM
int32_t* id_status();    // device words {count of out-of-range ids, one offending id} of the calling thread, or nullptr (api.cu)

template <class... KArgs, class... Args>
static inline void launch_k_mode(bool pdl, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);       // errors are picked up by check_launch()
}
template <class K, class... Args>
static inline void launch_k(K kernel, dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    launch_k_mode(pdl_mode() != 0, kernel, grid, block, smem, st, args...);
}
// A CTA that owns a whole SM (192 KB of shared memory, all of TMEM): parked early behind its predecessor it would keep
// that SM from the kernels of the other streams (weight-gradient GEMMs, the sort), which costs more than the launch gap.
template <class K, class... Args>
static inline void launch_k_heavy(K kernel, dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    launch_k_mode(pdl_mode() == 1, kernel, grid, block, smem, st, args...);
}
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
// "Issue" loads: volatile asm keeps them in program order relative to each other.  ptxas otherwise sinks every
// load to its first use and serialises a batch of independent gathers; callers issue a batch of these and then
// consume the LAST-issued value first, which forces the whole batch to be in flight before any use.
__device__ __forceinline__ float4 ldg_f4_issue(const float* p) {
    float4 v;
    asm volatile("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

// Row ids come from data files: an id outside [0, X) (a forgotten -1 pad, a test-set feature above the training maximum) is
// counted in the caller's status words (dctr_set_id_status) and READ as row 0, so it can never address memory outside the
// table -- TensorFlow raises InvalidArgument for these; the host side raises after the step.
__device__ __forceinline__ int checked_id(int id, int X, int32_t* __restrict__ status) {
    if ((unsigned)id >= (unsigned)X) {
        if (status) { atomicAdd(status, 1); status[1] = id; }
        return 0;
    }
    return id;
}

// ids / weights of 8 consecutive fields [f0, f0+8) of one sample (row-major [N, F] arrays, pbase = n*F); slots past F
// get id 0 / weight 0.  `vec`: F % 4 == 0 and 16-byte aligned arrays, so the 8 values are two 128-bit loads.
__device__ __forceinline__ void load_fields8(const int32_t* __restrict__ ids, const float* __restrict__ wts, int64_t pbase,
                                             int f0, int F, bool vec, int (&id)[8], float (&w)[8], int X, int32_t* __restrict__ status) {
    if (vec && f0 + 8 <= F) {
        const int4 a = __ldg(reinterpret_cast<const int4*>(ids + pbase + f0)), b = __ldg(reinterpret_cast<const int4*>(ids + pbase + f0) + 1);
        id[0] = a.x; id[1] = a.y; id[2] = a.z; id[3] = a.w; id[4] = b.x; id[5] = b.y; id[6] = b.z; id[7] = b.w;
        if (wts) {
            const float4 c = __ldg(reinterpret_cast<const float4*>(wts + pbase + f0)), e = __ldg(reinterpret_cast<const float4*>(wts + pbase + f0) + 1);
            w[0] = c.x; w[1] = c.y; w[2] = c.z; w[3] = c.w; w[4] = e.x; w[5] = e.y; w[6] = e.z; w[7] = e.w;
        } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) w[k] = 1.f;
        }
    } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const bool ok = f0 + k < F;
            id[k] = ok ? __ldg(ids + pbase + f0 + k) : 0;
            w[k] = ok ? (wts ? __ldg(wts + pbase + f0 + k) : 1.f) : 0.f;
        }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) id[k] = checked_id(id[k], X, status);
}

__device__ __forceinline__ float4 ld_f4_issue(const float* p, bool pred) {      // plain (coherent) load, predicated
    float4 v;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %5, 0;\n\tmov.f32 %0, 0f00000000;\n\tmov.f32 %1, 0f00000000;\n\t"
                 "mov.f32 %2, 0f00000000;\n\tmov.f32 %3, 0f00000000;\n\t@p ld.global.v4.f32 {%0,%1,%2,%3}, [%4];\n\t}"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p), "r"((int)pred));
    return v;
}
__device__ __forceinline__ float4 ldg_f4_issue(const float* p, bool pred) {
    float4 v;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %5, 0;\n\tmov.f32 %0, 0f00000000;\n\tmov.f32 %1, 0f00000000;\n\t"
                 "mov.f32 %2, 0f00000000;\n\tmov.f32 %3, 0f00000000;\n\t@p ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];\n\t}"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p), "r"((int)pred));
    return v;
}
__device__ __forceinline__ float ldg_f_issue(const float* p, bool pred) {
    float v;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.s32 p, %2, 0;\n\tmov.f32 %0, 0f00000000;\n\t@p ld.global.nc.f32 %0, [%1];\n\t}"
                 : "=f"(v) : "l"(p), "r"((int)pred));
    return v;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace dctr
