// K1fm: fused gather + FM forward + sigmoid cross-entropy + dlogit (FM.py:55-64, 33-38).
// One warp per sample; the field block is staged in shared memory exactly like K1 (gather.cu);
// s_c = sum_f x_f*row_f[c] and sum_f (x_f*row_f[c])^2 are folded per component by lane c, so the table
// rows are read ONCE (the reference does three embedding_lookup_sparse passes and materialises V^2).
#include "common.cuh"

namespace dctr {

constexpr int kFmWarps = 8;

__global__ void __launch_bounds__(kFmWarps * 32)
fm_fwd_kernel(const int32_t* __restrict__ ids, const float* __restrict__ wts, const float* __restrict__ table,
              int ld, int N, int F, int d, const float* __restrict__ bias, const float* __restrict__ y,
              float gscale, float* __restrict__ logit, float* __restrict__ pred, float* __restrict__ loss_n,
              float* __restrict__ g, float* __restrict__ s, int lds) {
    extern __shared__ __align__(16) float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* e = smem + warp * F * ld;
    const int Q = ld >> 2;
    const float b = bias ? __ldg(bias) : 0.f;

    for (int64_t n = (int64_t)blockIdx.x * kFmWarps + warp; n < N; n += (int64_t)gridDim.x * kFmWarps) {
        for (int t = lane; t < F * Q; t += 32) {
            const int f = t / Q;
            const int q = t - f * Q;
            const int64_t p = n * F + f;
            const int64_t id = ids[p];
            const float w = wts ? wts[p] : 1.f;
            float4 v = ldg_f4(table + id * ld + q * 4);
            v.x *= w; v.y *= w; v.z *= w; v.w *= w;
            *reinterpret_cast<float4*>(e + f * ld + q * 4) = v;
        }
        __syncwarp();
        float part = 0.f;
        for (int c = lane; c < d; c += 32) {
            float sc = 0.f, qc = 0.f;
            for (int f = 0; f < F; ++f) {
                const float v = e[f * ld + c];
                sc += v;
                qc = fmaf(v, v, qc);
            }
            if (s) s[n * (int64_t)lds + c] = sc;
            part += (c == 0) ? sc : 0.5f * (sc * sc - qc);
        }
        if (s) for (int c = d + lane; c < lds; c += 32) s[n * (int64_t)lds + c] = 0.f;
        const float z = warp_sum(part) + b;
        if (lane == 0) {
            const float p = sigmoidf_(z);
            if (logit) logit[n] = z;
            if (pred) pred[n] = p;
            if (y) {
                const float yn = y[n];
                if (loss_n) loss_n[n] = xent_logits(z, yn);
                if (g) g[n] = (p - yn) * gscale;
            }
        }
        __syncwarp();
    }
}

}  // namespace dctr

using namespace dctr;

extern "C" int dctr_fm_fwd(const int32_t* ids, const float* wts, const float* table, int64_t X, int ld,
                           int N, int F, int d, const float* bias, const float* y, float gscale,
                           float* logit, float* pred, float* loss_n, float* g, float* s, int lds, void* stream) {
    DCTR_REQUIRE(ids && table, "fm_fwd: null pointer");
    DCTR_REQUIRE(N >= 0 && F > 0 && d > 0 && d <= ld && (ld & 3) == 0 && X > 0, "fm_fwd: bad shape N=%d F=%d d=%d ld=%d", N, F, d, ld);
    DCTR_REQUIRE(!s || lds >= d, "fm_fwd: lds=%d < d=%d", lds, d);
    if (N == 0) return DCTR_OK;
    size_t smem = (size_t)kFmWarps * F * ld * sizeof(float);
    DCTR_REQUIRE(smem <= 200 * 1024, "fm_fwd: field block too large for shared memory");
    if (smem > 48 * 1024)
        cudaFuncSetAttribute(fm_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int blocks = ceil_div(N, kFmWarps);
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    fm_fwd_kernel<<<blocks, kFmWarps * 32, smem, as_stream(stream)>>>(ids, wts, table, ld, N, F, d, bias, y, gscale,
                                                                      logit, pred, loss_n, g, s, lds);
    return check_launch("fm_fwd_kernel");
}
