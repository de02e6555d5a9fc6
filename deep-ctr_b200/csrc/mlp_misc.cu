// K5 / K7 and the small fixed-order reductions around the MLP: output unit + sigmoid cross-entropy,
// its backward, weighted column sums (bias and last-layer gradients), loss sum, dense SGD.
#include "common.cuh"

namespace dctr {

// one warp per sample: logit = <a[n,:], w> + b
__global__ void __launch_bounds__(256)
out_xent_fwd_kernel(const float* __restrict__ a, int lda, int N, int H, const float* __restrict__ w,
                    const float* __restrict__ b, const float* __restrict__ y, float gscale,
                    float* __restrict__ logit, float* __restrict__ pred, float* __restrict__ loss_n,
                    float* __restrict__ dlogit) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const float bv = b ? __ldg(b) : 0.f;
    for (int64_t n = warp; n < N; n += nwarps) {
        const float* row = a + n * lda;
        float acc = 0.f;
        for (int h = lane; h < H; h += 32) acc = fmaf(row[h], __ldg(w + h), acc);
        const float z = warp_sum(acc) + bv;
        if (lane == 0) {
            const float p = sigmoidf_(z);
            if (logit) logit[n] = z;
            if (pred) pred[n] = p;
            if (y) {
                const float yn = y[n];
                if (loss_n) loss_n[n] = xent_logits(z, yn);
                if (dlogit) dlogit[n] = (p - yn) * gscale;
            }
        }
    }
}

__global__ void __launch_bounds__(256)
out_bwd_kernel(const float* __restrict__ dlogit, const float* __restrict__ w, const float* __restrict__ a, int lda,
               int N, int H, int act, const float* __restrict__ mask, int ldmask, float* __restrict__ da, int ldda) {
    const int64_t total = (int64_t)N * H;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int n = (int)(t / H), h = (int)(t - (int64_t)n * H);
        const float mk = mask ? __ldg(mask + (int64_t)n * ldmask + h) : 1.f;
        float v = 0.f;
        if (mk != 0.f) {
            const float tt = a[(int64_t)n * lda + h] / mk;
            v = dlogit[n] * __ldg(w + h) * mk * act_bwd_from_out(act, tt);
        }
        da[(int64_t)n * ldda + h] = v;
    }
}

// stage 1: CTA (bx, by) reduces rows [by*rows_per, ...) of columns [bx*32, +32) -> partial[by][h]
__global__ void __launch_bounds__(256)
col_wsum_stage1(const float* __restrict__ A, int lda, int N, int H, const float* __restrict__ wv, int rows_per,
                float* __restrict__ partial) {
    __shared__ float sm[8][33];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int h = blockIdx.x * 32 + lane;
    const int r0 = blockIdx.y * rows_per;
    const int r1 = min(N, r0 + rows_per);
    float acc = 0.f;
    if (h < H)
        for (int n = r0 + warp; n < r1; n += 8) acc = fmaf(wv ? __ldg(wv + n) : 1.f, A[(int64_t)n * lda + h], acc);
    sm[warp][lane] = acc;
    __syncthreads();
    if (warp == 0 && h < H) {
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) s += sm[k][lane];
        partial[(int64_t)blockIdx.y * H + h] = s;
    }
}
__global__ void col_wsum_stage2(const float* __restrict__ partial, int S, int H, float* __restrict__ out) {
    const int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= H) return;
    float s = 0.f;
    for (int k = 0; k < S; ++k) s += partial[(int64_t)k * H + h];
    out[h] = s;
}

__global__ void __launch_bounds__(1024) vec_sum_kernel(const float* __restrict__ v, int N, float scale, float* __restrict__ out) {
    __shared__ float sm[1024];
    float acc = 0.f;
    for (int i = threadIdx.x; i < N; i += 1024) acc += v[i];
    sm[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 512; o > 0; o >>= 1) {
        if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[0] = sm[0] * scale;
}

__global__ void dense_sgd_kernel(float* __restrict__ p, const float* __restrict__ g, int64_t n, float lr, float decay) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        p[i] = fmaf(p[i], decay, -lr * g[i]);
}

static int col_chunks(int N, int* rows_per) {
    int rp = (N + 127) / 128;
    if (rp < 256) rp = 256;
    *rows_per = rp;
    return (N + rp - 1) / rp;
}

}  // namespace dctr

using namespace dctr;

extern "C" int dctr_out_xent_fwd(const float* a, int lda, int N, int H, const float* w, const float* b,
                                 const float* y, float gscale, float* logit, float* pred, float* loss_n,
                                 float* dlogit, void* stream) {
    DCTR_REQUIRE(a && w && N >= 0 && H > 0 && lda >= H, "out_xent_fwd: bad arguments");
    if (N == 0) return DCTR_OK;
    int blocks = ceil_div(N, 8);
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    out_xent_fwd_kernel<<<blocks, 256, 0, as_stream(stream)>>>(a, lda, N, H, w, b, y, gscale, logit, pred, loss_n, dlogit);
    return check_launch("out_xent_fwd_kernel");
}

extern "C" int dctr_out_bwd(const float* dlogit, const float* w, const float* a, int lda, int N, int H, int act,
                            const float* mask, int ldmask, float* da, int ldda, void* stream) {
    DCTR_REQUIRE(dlogit && w && a && da && N >= 0 && H > 0, "out_bwd: bad arguments");
    if (N == 0) return DCTR_OK;
    int blocks = ceil_div((int64_t)N * H, 256);
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    out_bwd_kernel<<<blocks, 256, 0, as_stream(stream)>>>(dlogit, w, a, lda, N, H, act, mask, ldmask, da, ldda);
    return check_launch("out_bwd_kernel");
}

extern "C" size_t dctr_col_wsum_workspace(int N, int H) {
    int rp;
    const int S = col_chunks(N > 0 ? N : 1, &rp);
    return (size_t)S * (size_t)(H > 0 ? H : 1) * sizeof(float);
}

extern "C" int dctr_col_wsum(const float* A, int lda, int N, int H, const float* wv, float* out, void* partial,
                             size_t partial_bytes, void* stream) {
    DCTR_REQUIRE(A && out && partial && N > 0 && H > 0, "col_wsum: bad arguments");
    if (partial_bytes < dctr_col_wsum_workspace(N, H)) {
        set_error("col_wsum: workspace %zu < %zu", partial_bytes, dctr_col_wsum_workspace(N, H));
        return DCTR_EWORKSPACE;
    }
    int rp;
    const int S = col_chunks(N, &rp);
    dim3 grid(ceil_div(H, 32), S);
    col_wsum_stage1<<<grid, 256, 0, as_stream(stream)>>>(A, lda, N, H, wv, rp, static_cast<float*>(partial));
    int rc = check_launch("col_wsum_stage1");
    if (rc) return rc;
    col_wsum_stage2<<<ceil_div(H, 128), 128, 0, as_stream(stream)>>>(static_cast<const float*>(partial), S, H, out);
    return check_launch("col_wsum_stage2");
}

extern "C" int dctr_vec_sum(const float* v, int N, float scale, float* out, void* stream) {
    DCTR_REQUIRE(v && out && N >= 0, "vec_sum: bad arguments");
    vec_sum_kernel<<<1, 1024, 0, as_stream(stream)>>>(v, N, scale, out);
    return check_launch("vec_sum_kernel");
}

extern "C" int dctr_dense_sgd(float* param, const float* grad, int64_t n, float lr, float decay, void* stream) {
    DCTR_REQUIRE(param && grad && n >= 0, "dense_sgd: bad arguments");
    if (n == 0) return DCTR_OK;
    int blocks = ceil_div(n, 256);
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    dense_sgd_kernel<<<blocks, 256, 0, as_stream(stream)>>>(param, grad, n, lr, decay);
    return check_launch("dense_sgd_kernel");
}
