// K1fm: fused gather + FM forward + sigmoid cross-entropy + dlogit (FM.py:55-64, 33-38).
// A 4-lane group owns one sample; lane q reads float4 q of each of the F rows (F independent 128-bit loads per
// lane) and keeps s = sum_f x_f*row_f and sum_f (x_f*row_f)^2 for its four components in registers, so the
// table rows are read ONCE and nothing is staged (the reference does three embedding_lookup_sparse passes and
// materialises V^2).  Two shuffles fold the four lanes into the logit.
#include "common.cuh"

namespace dctr {

constexpr int kFmThreads = 256;

__global__ void __launch_bounds__(kFmThreads)
fm_fwd_kernel(const int32_t* __restrict__ ids, const float* __restrict__ wts, const float* __restrict__ table,
              int ld, int N, int F, int d, const float* __restrict__ bias, const float* __restrict__ y,
              float gscale, float* __restrict__ logit, float* __restrict__ pred, float* __restrict__ loss_n,
              float* __restrict__ g, float* __restrict__ s, int lds, int vec, int X, int32_t* __restrict__ status) {
    pdl_sync();
    const int q = threadIdx.x & 3;
    const float b = bias ? __ldg(bias) : 0.f;
    const int64_t groups = ((int64_t)gridDim.x * kFmThreads) >> 2;
    // the trip count is uniform per warp (8 samples per warp and round) so the shuffles below stay converged
    for (int64_t n0 = (((int64_t)blockIdx.x * kFmThreads + threadIdx.x) >> 5) * 8; n0 < N; n0 += groups) {
        const int64_t n = n0 + ((threadIdx.x & 31) >> 2);
        const bool on = n < N;
        float part = 0.f;
        for (int c0 = q * 4; c0 < ld; c0 += 16) {
            float4 sv = make_float4(0.f, 0.f, 0.f, 0.f), qv = sv;
            if (on) {
                const int64_t pbase = n * F;
                for (int f0 = 0; f0 < F; f0 += 8) {
                    int id[8]; float w[8]; float4 v[8];
                    load_fields8(ids, wts, pbase, f0, F, vec, id, w, X, status);
#pragma unroll
                    for (int k = 0; k < 8; ++k) v[k] = ldg_f4_issue(table + (int64_t)id[k] * ld + c0);
#pragma unroll
                    for (int k = 7; k >= 0; --k) {     // slots past F carry weight 0; last-issued load first (common.cuh)
                        const float x = v[k].x * w[k], yv = v[k].y * w[k], z = v[k].z * w[k], u = v[k].w * w[k];
                        sv.x += x; sv.y += yv; sv.z += z; sv.w += u;
                        qv.x = fmaf(x, x, qv.x); qv.y = fmaf(yv, yv, qv.y); qv.z = fmaf(z, z, qv.z); qv.w = fmaf(u, u, qv.w);
                    }
                }
                if (s) *reinterpret_cast<float4*>(s + n * (int64_t)lds + c0) = sv;     // pad columns of the table are 0
                part += c0 == 0 ? sv.x : 0.5f * (sv.x * sv.x - qv.x);
                part += 0.5f * ((sv.y * sv.y - qv.y) + (sv.z * sv.z - qv.z) + (sv.w * sv.w - qv.w));
            }
        }
        if (on && s) for (int c = ld + q * 4; c < lds; c += 16) *reinterpret_cast<float4*>(s + n * (int64_t)lds + c) = make_float4(0.f, 0.f, 0.f, 0.f);
        part += __shfl_xor_sync(0xffffffffu, part, 1);
        part += __shfl_xor_sync(0xffffffffu, part, 2);
        if (on && q == 0) {
            const float z = part + b;
            const float p = sigmoidf_(z);
            if (logit) logit[n] = z;
            if (pred) pred[n] = p;
            if (y) {
                const float yn = y[n];
                if (loss_n) loss_n[n] = xent_logits(z, yn);
                if (g) g[n] = (p - yn) * gscale;
            }
        }
    }
}

}  // namespace dctr

using namespace dctr;

extern "C" int dctr_fm_fwd(const int32_t* ids, const float* wts, const float* table, int64_t X, int ld,
                           int N, int F, int d, const float* bias, const float* y, float gscale,
                           float* logit, float* pred, float* loss_n, float* g, float* s, int lds, void* stream) {
    DCTR_REQUIRE(ids && table, "fm_fwd: null pointer");
    DCTR_REQUIRE(N >= 0 && F > 0 && d > 0 && d <= ld && (ld & 3) == 0 && X > 0 && X < ((int64_t)1 << 31), "fm_fwd: bad shape N=%d F=%d d=%d ld=%d", N, F, d, ld);
    DCTR_REQUIRE(!s || lds >= d, "fm_fwd: lds=%d < d=%d", lds, d);
    if (N == 0) return DCTR_OK;
    DCTR_REQUIRE(((uintptr_t)table & 15) == 0 && (!s || (((uintptr_t)s & 15) == 0 && (lds & 3) == 0 && lds >= ld)),
                 "fm_fwd: table / s must be 16-byte aligned, lds a multiple of 4 and >= ld");
    const int vec = (F & 3) == 0 && ((uintptr_t)ids & 15) == 0 && (!wts || ((uintptr_t)wts & 15) == 0);
    int blocks = ceil_div(N, kFmThreads / 4);
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    launch_k(fm_fwd_kernel, blocks, kFmThreads, 0, as_stream(stream), ids, wts, table, ld, N, F, d, bias, y, gscale,
                                                                 logit, pred, loss_n, g, s, lds, vec, (int)X, id_status());
    return check_launch("fm_fwd_kernel");
}
