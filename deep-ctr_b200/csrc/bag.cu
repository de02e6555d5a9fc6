// Embedding-bag front end: the first layer of the reference's SNN models is a SUM over the rows of the
// active features followed by a sigmoid -- x_dense = 1 / (1 + exp(-(sum_f ww0[feat_f] + bb0)))  (get_fi_h1_y,
// SNN_DAE.py:234-260; SNN_RBM.py:238-262) -- with 200-wide rows (800 B, 6.25 sectors... 25 full 32-byte sectors).
// HBM-bound like K1: F row reads of ld*4 bytes per sample, one [H] write.  A warp owns a sample; lane l holds the float4
// at columns 128*cb + 4*l of each row, so a row is read as whole 512-byte warp requests; the id / weight loads of 8 fields
// are issued together and then the 8 row loads (volatile asm keeps them batched), accumulated in field order.
#include "common.cuh"

namespace dctr {

constexpr int kBagWarps = 8;

__global__ void __launch_bounds__(kBagWarps * 32)
bag_fwd_kernel(const int32_t* __restrict__ ids, const float* __restrict__ wts, const float* __restrict__ table, int ld,
               int N, int F, int H, const float* __restrict__ b0, int act, float* __restrict__ out, int ldo, bool vec,
               int X, int32_t* __restrict__ status) {
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int n = blockIdx.x * kBagWarps + (threadIdx.x >> 5);
    if (n >= N) return;
    const int64_t pbase = (int64_t)n * F;
    for (int c = 4 * lane; c < ldo; c += 128) {
        const bool on = c < ld;                     // columns the table rows have (ld % 4 == 0)
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        if (b0) {
            if (c < H) acc.x = __ldg(b0 + c);
            if (c + 1 < H) acc.y = __ldg(b0 + c + 1);
            if (c + 2 < H) acc.z = __ldg(b0 + c + 2);
            if (c + 3 < H) acc.w = __ldg(b0 + c + 3);
        }
        for (int f0 = 0; f0 < F; f0 += 8) {
            int id[8];
            float w[8];
            load_fields8(ids, wts, pbase, f0, F, vec, id, w, X, status);
            float4 r[8];
#pragma unroll
            for (int k = 7; k >= 0; --k) r[k] = ldg_f4_issue(table + (int64_t)id[k] * ld + c, on && f0 + k < F);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                acc.x = fmaf(w[k], r[k].x, acc.x); acc.y = fmaf(w[k], r[k].y, acc.y);
                acc.z = fmaf(w[k], r[k].z, acc.z); acc.w = fmaf(w[k], r[k].w, acc.w);
            }
        }
        float4 o;
        o.x = c < H ? act_fwd(act, acc.x) : 0.f;
        o.y = c + 1 < H ? act_fwd(act, acc.y) : 0.f;
        o.z = c + 2 < H ? act_fwd(act, acc.z) : 0.f;
        o.w = c + 3 < H ? act_fwd(act, acc.w) : 0.f;
        *reinterpret_cast<float4*>(out + (int64_t)n * ldo + c) = o;
    }
}

}  // namespace dctr

using namespace dctr;

extern "C" int dctr_bag_fwd(const int32_t* ids, const float* wts, const float* table, int64_t X, int ld, int N, int F, int H,
                            const float* b0, int act, float* out, int ldo, void* stream) {
    DCTR_REQUIRE(N >= 0 && F > 0 && H > 0 && H <= ld && (ld & 3) == 0 && ldo >= H && (ldo & 3) == 0 && X > 0, "bag_fwd: bad shape");
    if (N == 0) return DCTR_OK;
    DCTR_REQUIRE(ids && table && out, "bag_fwd: null pointer");
    DCTR_REQUIRE((((uintptr_t)table | (uintptr_t)out) & 15) == 0, "bag_fwd: table / out must be 16-byte aligned");
    const bool vec = (F & 3) == 0 && ((uintptr_t)ids & 15) == 0 && (!wts || ((uintptr_t)wts & 15) == 0);
    launch_k(bag_fwd_kernel, ceil_div(N, kBagWarps), kBagWarps * 32, 0, as_stream(stream), ids, wts, table, ld, N, F, H, b0, act,
             out, ldo, vec, (int)X, id_status());
    return check_launch("bag_fwd_kernel");
}
