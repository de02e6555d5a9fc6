// K5 / K7 and the small fixed-order reductions around the MLP: output unit + sigmoid cross-entropy,
// its backward, weighted column sums (bias and last-layer gradients), loss sum, dense SGD.
#include "common.cuh"

namespace dctr {

// ---- single-launch fixed-order reductions ------------------------------------------------------------------
// Every CTA owns a block of rows and writes one partial row; the CTA that takes the last ticket adds the partial rows in
// block order (fixed order => bit-reproducible, no float atomics) and resets the ticket, so the workspace stays
// reusable across launches and CUDA-graph replays.  The workspace must be zero-filled once before first use.
constexpr int kRedThreads = 256;
constexpr int kRedMaxBlocks = kNumSMs;

__device__ __forceinline__ bool take_last_ticket(unsigned* ticket, unsigned nblocks) {
    __shared__ unsigned s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(ticket, 1u);
        s_last = (t == nblocks - 1) ? 1u : 0u;
        if (s_last) *ticket = 0u;
    }
    __syncthreads();
    if (s_last) __threadfence();
    return s_last != 0;
}

// Last-CTA fold of partial[S][W] columns [0, Hn): column h is summed over the S partial rows in a fixed order -- by up to
// 32 k-slices in parallel (each slice in row order, the slices then in slice order), so narrow reductions (a loss, a bias
// column) do not serialise hundreds of dependent L2 round trips on one thread.  Called by every thread of the CTA.
template <class Dst>
__device__ __forceinline__ void ordered_reduce(const float* __restrict__ partial, int S, int W, int Hn, Dst dst) {
    __shared__ float s_part[kRedThreads];
    __syncthreads();                                        // s_part may still be read by a previous call
    int KS = 1;
    while (KS < 32 && KS * 2 * Hn <= kRedThreads) KS *= 2;
    if (KS == 1) {
        for (int h = threadIdx.x; h < Hn; h += kRedThreads) {
            float s = 0.f;
#pragma unroll 16
            for (int k = 0; k < S; ++k) s += __ldcg(partial + (int64_t)k * W + h);
            dst(h, s);
        }
        return;
    }
    const int slice = (S + KS - 1) / KS;
    const int h = threadIdx.x % Hn, ks = threadIdx.x / Hn;
    float s = 0.f;
    if (ks < KS) {
        const int k1 = min(S, (ks + 1) * slice);
#pragma unroll 16
        for (int k = ks * slice; k < k1; ++k) s += __ldcg(partial + (int64_t)k * W + h);
    }
    s_part[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x < Hn) {
        float t = 0.f;
        for (int q = 0; q < KS; ++q) t += s_part[q * Hn + threadIdx.x];
        dst(threadIdx.x, t);
    }
}

// one warp per sample: logit = <a[n,:], w> + b
__global__ void __launch_bounds__(256)
out_xent_fwd_kernel(const float* __restrict__ a, int lda, int N, int H, const float* __restrict__ w,
                    const float* __restrict__ b, const float* __restrict__ y, float gscale,
                    float* __restrict__ logit, float* __restrict__ pred, float* __restrict__ loss_n,
                    float* __restrict__ dlogit, float loss_scale, float* __restrict__ loss_sum,
                    float* __restrict__ partial, unsigned* __restrict__ ticket) {
    pdl_sync();
    __shared__ float s_loss[8];
    float my_loss = 0.f;
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
                const float ln = xent_logits(z, yn);
                my_loss += ln;
                if (loss_n) loss_n[n] = ln;
                if (dlogit) dlogit[n] = (p - yn) * gscale;
            }
        }
    }
    if (!loss_sum) return;
    // loss = scale * sum_n loss_n (FM.py:36 / FNN.py:57): warp partials in warp order, CTA partials in CTA order
    if (lane == 0) s_loss[threadIdx.x >> 5] = my_loss;
    __syncthreads();
    if (threadIdx.x == 0) {
        float s = 0.f;
        for (int k = 0; k < 8; ++k) s += s_loss[k];
        partial[blockIdx.x] = s;
    }
    if (!take_last_ticket(ticket, gridDim.x)) return;
    ordered_reduce(partial, (int)gridDim.x, 1, 1, [&](int, float s) { loss_sum[0] = s * loss_scale; });
}

// Output unit of a TRAIN step in one pass over the last activation (FNN.py:54-63 + its autodiff): logit, pred, loss, dlogit,
// d(activation), d(w_out), d(b_out).  A warp owns rows n = w, w + W, ...: the row stays in registers between the GEMV
// and the backward; per-lane partial sums of dlogit * a over the warp's rows, then warps in order (shared memory), then
// CTAs in order (last-ticket CTA): fixed order, no float atomics.  HV = ceil(H / 32) register columns per lane.
template <int HV>
__global__ void __launch_bounds__(256)
out_unit_train_kernel(const float* __restrict__ a, int lda, int N, int H, const float* __restrict__ w, const float* __restrict__ b,
                      const float* __restrict__ y, float gscale, int act, const float* __restrict__ mask, int ldmask,
                      float* __restrict__ logit, float* __restrict__ pred, float* __restrict__ loss_n, float* __restrict__ dlogit,
                      float loss_scale, float* __restrict__ loss_sum, float* __restrict__ da, int ldda,
                      float* __restrict__ gw, float* __restrict__ gb, float* __restrict__ partial, unsigned* __restrict__ ticket) {
    pdl_sync();
    constexpr int WC = HV * 32 + 2;                       // gw columns | gb | loss
    extern __shared__ float s_red[];                      // [8][WC]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t gwarp = (int64_t)blockIdx.x * 8 + warp, nwarps = (int64_t)gridDim.x * 8;
    const float bv = b ? __ldg(b) : 0.f;
    float wreg[HV], gacc[HV];
#pragma unroll
    for (int k = 0; k < HV; ++k) { const int h = lane + 32 * k; wreg[k] = h < H ? __ldg(w + h) : 0.f; gacc[k] = 0.f; }
    float gbacc = 0.f, lacc = 0.f;
    for (int64_t n = gwarp; n < N; n += nwarps) {
        const float* row = a + n * lda;
        float av[HV], acc = 0.f;
#pragma unroll
        for (int k = 0; k < HV; ++k) { const int h = lane + 32 * k; av[k] = h < H ? row[h] : 0.f; acc = fmaf(av[k], wreg[k], acc); }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        const float z = acc + bv, p = sigmoidf_(z), yn = y[n];
        const float ln = xent_logits(z, yn), dl = (p - yn) * gscale;
        if (lane == 0) {
            if (logit) logit[n] = z;
            if (pred) pred[n] = p;
            if (loss_n) loss_n[n] = ln;
            if (dlogit) dlogit[n] = dl;
        }
        lacc += ln;
        gbacc += dl;
        float* drow = da + n * ldda;
#pragma unroll
        for (int k = 0; k < HV; ++k) {
            const int h = lane + 32 * k;
            if (h < H) {
                gacc[k] = fmaf(dl, av[k], gacc[k]);
                const float mk = mask ? __ldg(mask + n * ldmask + h) : 1.f;
                drow[h] = mk != 0.f ? dl * wreg[k] * mk * act_bwd_from_out(act, mask ? av[k] / mk : av[k]) : 0.f;
            }
        }
    }
    float* mine = s_red + warp * WC;
#pragma unroll
    for (int k = 0; k < HV; ++k) mine[lane + 32 * k] = gacc[k];
    if (lane == 0) { mine[HV * 32] = gbacc; mine[HV * 32 + 1] = lacc; }
    __syncthreads();
    const int W = H + 2;
    for (int c = threadIdx.x; c < W; c += 256) {
        const int src = c < H ? c : HV * 32 + (c - H);
        float s = 0.f;
#pragma unroll
        for (int q = 0; q < 8; ++q) s += s_red[q * WC + src];
        partial[(int64_t)blockIdx.x * W + c] = s;
    }
    if (!take_last_ticket(ticket, gridDim.x)) return;
    ordered_reduce(partial, (int)gridDim.x, W, W, [&](int c, float s) {
        if (c < H) gw[c] = s;
        else if (c == H) gb[0] = s;
        else if (loss_sum) loss_sum[0] = s * loss_scale;
    });
}

__global__ void __launch_bounds__(256)
out_bwd_kernel(const float* __restrict__ dlogit, const float* __restrict__ w, const float* __restrict__ a, int lda,
               int N, int H, int act, const float* __restrict__ mask, int ldmask, float* __restrict__ da, int ldda) {
    pdl_sync();
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

// ---- column reductions over an [N, H] row-major matrix ------------------------------------------------------
// Grid = (column tiles, row chunks).  A CTA is TX column groups (V consecutive columns each: one LDG.128 when V = 4) by
// TY = 256 / TX row lanes; every thread walks its chunk's rows with stride TY, four independent loads in flight, the
// row lanes are then folded by a fixed shared-memory tree and the chunk's partial row goes to the workspace.  The last
// CTA of each COLUMN TILE (one ticket per tile) adds the chunk partials in chunk order: fixed order, no float atomics.
struct RedShape {
    int TX, V, col_tiles, chunks, rows_per;
};

template <int V>
__device__ __forceinline__ void ld_cols(const float* p, bool ok, float (&v)[V]) {
    if constexpr (V == 4) {
        const float4 t = ok ? *reinterpret_cast<const float4*>(p) : make_float4(0.f, 0.f, 0.f, 0.f);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else {
        v[0] = ok ? *p : 0.f;
    }
}

// fold acc[V] over the TY row lanes (fixed tree); the result is valid in the ty == 0 threads
template <int V>
__device__ __forceinline__ void fold_row_lanes(float (&acc)[V], int TX, float* s_acc) {
    const int TY = kRedThreads / TX, ty = threadIdx.x / TX;
#pragma unroll
    for (int c = 0; c < V; ++c) s_acc[c * kRedThreads + threadIdx.x] = acc[c];
    __syncthreads();
    for (int o = TY >> 1; o > 0; o >>= 1) {
        if (ty < o) {
#pragma unroll
            for (int c = 0; c < V; ++c) s_acc[c * kRedThreads + threadIdx.x] += s_acc[c * kRedThreads + threadIdx.x + o * TX];
        }
        __syncthreads();
    }
#pragma unroll
    for (int c = 0; c < V; ++c) acc[c] = s_acc[c * kRedThreads + threadIdx.x];
}

// out[h] = sum_n (wv ? wv[n] : 1) * A[n,h]
template <int V>
__global__ void __launch_bounds__(kRedThreads)
col_wsum_kernel(const float* __restrict__ A, int lda, int N, int H, const float* __restrict__ wv, RedShape sh,
                float* __restrict__ partial, unsigned* __restrict__ tickets, float* __restrict__ out, int sgd, float lr, float decay) {
    pdl_sync();
    __shared__ float s_acc[V * kRedThreads];
    const int TX = sh.TX, TY = kRedThreads / TX, tx = threadIdx.x % TX, ty = threadIdx.x / TX;
    const int c0 = blockIdx.x * TX * V, h0 = c0 + tx * V;
    const int r0 = blockIdx.y * sh.rows_per, r1 = min(N, r0 + sh.rows_per);
    const bool col_ok = h0 < H;
    float acc[V] = {};
    for (int n = r0 + ty; n < r1; n += 4 * TY) {
        float v[4][V], w[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int nn = n + u * TY;
            const bool ok = col_ok && nn < r1;
            ld_cols<V>(A + (int64_t)nn * lda + h0, ok, v[u]);
            w[u] = (wv && ok) ? __ldg(wv + nn) : 1.f;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int c = 0; c < V; ++c) acc[c] = fmaf(w[u], v[u][c], acc[c]);
    }
    fold_row_lanes<V>(acc, TX, s_acc);
    if (ty == 0 && col_ok) {
#pragma unroll
        for (int c = 0; c < V; ++c)
            if (h0 + c < H) partial[(int64_t)blockIdx.y * H + h0 + c] = acc[c];
    }
    if (!take_last_ticket(tickets + blockIdx.x, gridDim.y)) return;
    // sgd: out is the parameter vector itself (a bias): out = out * decay - lr * (column sum)
    ordered_reduce(partial + c0, (int)gridDim.y, H, min(TX * V, H - c0),
                   [&](int h, float s) { out[c0 + h] = sgd ? fmaf(out[c0 + h], decay, -lr * s) : s; });
}

// backward of the output unit in one launch (FNN.py:54-63 / FNN_wnzh.py:169-174 differentiated):
//   da[n,h] = dlogit[n]*w[h] * mask * act'(a/mask);  gw[h] = sum_n dlogit[n]*a[n,h];  gb = sum_n dlogit[n]
template <int V>
__global__ void __launch_bounds__(kRedThreads)
out_bwd_fused_kernel(const float* __restrict__ dlogit, const float* __restrict__ w, const float* __restrict__ a, int lda,
                     int N, int H, int act, const float* __restrict__ mask, int ldmask, float* __restrict__ da, int ldda,
                     RedShape sh, float* __restrict__ partial, unsigned* __restrict__ tickets,
                     float* __restrict__ gw, float* __restrict__ gb) {
    pdl_sync();
    __shared__ float s_acc[V * kRedThreads];
    const int TX = sh.TX, TY = kRedThreads / TX, tx = threadIdx.x % TX, ty = threadIdx.x / TX;
    const int c0 = blockIdx.x * TX * V, h0 = c0 + tx * V;
    const int r0 = blockIdx.y * sh.rows_per, r1 = min(N, r0 + sh.rows_per);
    const int HP = H + 1;
    const bool col_ok = h0 < H;
    float wh[V];
#pragma unroll
    for (int c = 0; c < V; ++c) wh[c] = (h0 + c < H) ? __ldg(w + h0 + c) : 0.f;
    float acc[V] = {};
    float dsum = 0.f;                                       // sum of dlogit over this thread's rows (used by tx == 0 of tile 0)
    for (int n = r0 + ty; n < r1; n += 4 * TY) {
        float av[4][V], mk[4][V], dl[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int nn = n + u * TY;
            const bool ok = col_ok && nn < r1;
            ld_cols<V>(a + (int64_t)nn * lda + h0, ok, av[u]);
            if (mask) ld_cols<V>(mask + (int64_t)nn * ldmask + h0, ok, mk[u]);
            dl[u] = ok ? __ldg(dlogit + nn) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int nn = n + u * TY;
            if (!(col_ok && nn < r1)) continue;
            dsum += dl[u];
            float o[V];
#pragma unroll
            for (int c = 0; c < V; ++c) {
                acc[c] = fmaf(dl[u], av[u][c], acc[c]);
                const float m = mask ? mk[u][c] : 1.f;
                o[c] = m != 0.f ? dl[u] * wh[c] * m * act_bwd_from_out(act, mask ? av[u][c] / m : av[u][c]) : 0.f;
            }
            float* dst = da + (int64_t)nn * ldda + h0;
            if (V == 4 && h0 + 3 < H) {
                *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
            } else {
#pragma unroll
                for (int c = 0; c < V; ++c)
                    if (h0 + c < H) dst[c] = o[c];
            }
        }
    }
    fold_row_lanes<V>(acc, TX, s_acc);
    if (ty == 0 && col_ok) {
#pragma unroll
        for (int c = 0; c < V; ++c)
            if (h0 + c < H) partial[(int64_t)blockIdx.y * HP + h0 + c] = acc[c];
    }
    if (blockIdx.x == 0) {                                  // bias gradient: the dlogit column, folded the same way
        __syncthreads();
        float d1[1] = {tx == 0 ? dsum : 0.f};
        fold_row_lanes<1>(d1, TX, s_acc);
        if (threadIdx.x == 0) partial[(int64_t)blockIdx.y * HP + H] = d1[0];
    }
    if (!take_last_ticket(tickets + blockIdx.x, gridDim.y)) return;
    ordered_reduce(partial + c0, (int)gridDim.y, HP, min(TX * V, H - c0), [&](int h, float s) { gw[c0 + h] = s; });
    if (blockIdx.x == 0) ordered_reduce(partial + H, (int)gridDim.y, HP, 1, [&](int, float s) { gb[0] = s; });
}

__global__ void __launch_bounds__(1024) vec_sum_kernel(const float* __restrict__ v, int N, float scale, float* __restrict__ out) {
    pdl_sync();
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
    pdl_sync();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        p[i] = fmaf(p[i], decay, -lr * g[i]);
}

// ---- Adam, TF form (tf_util.py:18-21): dense over whatever buffer it is given ------------------------------------------
// state = [t, lr_t, beta1^t, beta2^t]: advanced ON THE DEVICE once per step, so a captured CUDA graph replays the
// right bias correction without a host-side scalar:  lr_t = lr * sqrt(1 - beta2^t) / (1 - beta1^t)
__global__ void adam_tick_kernel(float* __restrict__ state, float lr, float beta1, float beta2) {
    pdl_sync();
    const float t = state[0] + 1.f;
    const float p1 = (t == 1.f ? 1.f : state[2]) * beta1, p2 = (t == 1.f ? 1.f : state[3]) * beta2;
    state[0] = t;
    state[1] = lr * sqrtf(1.f - p2) / (1.f - p1);
    state[2] = p1;
    state[3] = p2;
}

__device__ __forceinline__ float adam1(float& p, float g, float& m, float& v, float lr_t, float b1, float b2, float eps, float l2) {
    g = fmaf(l2, p, g);
    m = fmaf(b1, m, (1.f - b1) * g);
    v = fmaf(b2, v, (1.f - b2) * g * g);
    p -= lr_t * m / (sqrtf(v) + eps);
    return p;
}

template <bool ZERO>
__global__ void dense_adam_kernel(float4* __restrict__ p, float4* __restrict__ g, float4* __restrict__ m, float4* __restrict__ v,
                                  int64_t n4, const float* __restrict__ state, float b1, float b2, float eps, float l2) {
    pdl_sync();
    const float lr_t = __ldg(state + 1);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        float4 pp = p[i], gg = g[i], mm = m[i], vv = v[i];
        adam1(pp.x, gg.x, mm.x, vv.x, lr_t, b1, b2, eps, l2);
        adam1(pp.y, gg.y, mm.y, vv.y, lr_t, b1, b2, eps, l2);
        adam1(pp.z, gg.z, mm.z, vv.z, lr_t, b1, b2, eps, l2);
        adam1(pp.w, gg.w, mm.w, vv.w, lr_t, b1, b2, eps, l2);
        p[i] = pp; m[i] = mm; v[i] = vv;
        if (ZERO) g[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
}

// ---- FTRL-proximal, TensorFlow's ApplyFtrl form (tf_util.py:22-24 FtrlOptimizer(lr): lr_power -0.5, accumulators start at
// 0.1, l1 = l2 = 0).  Dense like Adam: an element whose gradient is 0 keeps accum / linear and is RE-DERIVED from linear, so
// a weight that never saw a gradient goes to 0 on the first step (linear starts at 0) -- TF's behaviour, kept.
__device__ __forceinline__ void ftrl1(float& p, float g, float& n, float& z, float lr_inv, float lr_power, float l1, float l2,
                                      float gl2) {
    g = fmaf(gl2, p, g);
    const float n_new = fmaf(g, g, n);
    float s_new, ds;
    if (lr_power == -0.5f) {      // sqrt(n + g^2) - sqrt(n) = g^2 / (sqrt(n + g^2) + sqrt(n)): no cancellation for small g
        const float s_old = sqrtf(n);
        s_new = sqrtf(n_new);
        ds = s_new + s_old > 0.f ? g * g / (s_new + s_old) : 0.f;
    } else {
        s_new = powf(n_new, -lr_power);
        ds = s_new - powf(n, -lr_power);
    }
    z += g - ds * lr_inv * p;
    const float quad = fmaf(s_new, lr_inv, 2.f * l2);
    const float sgn = z > 0.f ? 1.f : (z < 0.f ? -1.f : 0.f);
    p = fabsf(z) > l1 ? (l1 * sgn - z) / quad : 0.f;
    n = n_new;
}

template <bool ZERO>
__global__ void dense_ftrl_kernel(float4* __restrict__ p, float4* __restrict__ g, float4* __restrict__ n, float4* __restrict__ z,
                                  int64_t n4, float lr_inv, float lr_power, float l1, float l2, float gl2) {
    pdl_sync();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        float4 pp = p[i], gg = g[i], nn = n[i], zz = z[i];
        ftrl1(pp.x, gg.x, nn.x, zz.x, lr_inv, lr_power, l1, l2, gl2);
        ftrl1(pp.y, gg.y, nn.y, zz.y, lr_inv, lr_power, l1, l2, gl2);
        ftrl1(pp.z, gg.z, nn.z, zz.z, lr_inv, lr_power, l1, l2, gl2);
        ftrl1(pp.w, gg.w, nn.w, zz.w, lr_inv, lr_power, l1, l2, gl2);
        p[i] = pp; n[i] = nn; z[i] = zz;
        if (ZERO) g[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
}

// Shape of a column reduction: V = 4 when every row start is 16-byte aligned, TX = the narrowest power of two covering
// the columns (<= 32), enough row chunks for ~4 CTAs per SM but at least 4 rows per row lane and chunk.
constexpr int kRedHeader = 1024;                       // bytes: one ticket per column tile
constexpr int kRedMaxChunks = kNumSMs * 4;

static RedShape red_shape(int N, int H, bool vec) {
    RedShape s;
    s.V = vec ? 4 : 1;
    const int groups = (H + s.V - 1) / s.V;
    s.TX = 1;
    while (s.TX < 32 && s.TX < groups) s.TX *= 2;
    s.col_tiles = (groups + s.TX - 1) / s.TX;
    const int TY = kRedThreads / s.TX;
    int chunks = (kRedMaxChunks + s.col_tiles - 1) / s.col_tiles;
    const int max_chunks = (N + 4 * TY - 1) / (4 * TY);
    if (chunks > max_chunks) chunks = max_chunks;
    if (chunks < 1) chunks = 1;
    s.rows_per = ((N + chunks - 1) / chunks + TY - 1) / TY * TY;
    s.chunks = (N + s.rows_per - 1) / s.rows_per;
    return s;
}

static bool vec_ok(const void* p, int ld) { return (((uintptr_t)p & 15) == 0) && (ld & 3) == 0; }

}  // namespace dctr

using namespace dctr;

extern "C" size_t dctr_reduce_workspace(int N, int H) {
    if (N < 1) N = 1;
    if (H < 1) H = 1;
    (void)N;
    const size_t rows = (size_t)(kRedMaxChunks > kNumSMs * 8 ? kRedMaxChunks : kNumSMs * 8);   // also out_xent_fwd's CTA partials
    return kRedHeader + rows * (size_t)(H + 1) * sizeof(float);
}

extern "C" int dctr_out_xent_fwd(const float* a, int lda, int N, int H, const float* w, const float* b,
                                 const float* y, float gscale, float* logit, float* pred, float* loss_n,
                                 float* dlogit, float loss_scale, float* loss_sum, void* workspace,
                                 size_t workspace_bytes, void* stream) {
    DCTR_REQUIRE(a && w && N >= 0 && H > 0 && lda >= H, "out_xent_fwd: bad arguments");
    DCTR_REQUIRE(!loss_sum || (y && workspace && workspace_bytes >= dctr_reduce_workspace(N, 1)),
                 "out_xent_fwd: the loss sum needs labels and a dctr_reduce_workspace(N, 1) workspace");
    if (N == 0) return DCTR_OK;
    int blocks = ceil_div(N, 8);
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    unsigned* ticket = static_cast<unsigned*>(workspace);
    float* partial = workspace ? reinterpret_cast<float*>(static_cast<char*>(workspace) + kRedHeader) : nullptr;
    launch_k(out_xent_fwd_kernel, blocks, 256, 0, as_stream(stream), a, lda, N, H, w, b, y, gscale, logit, pred, loss_n, dlogit,
                                                                loss_scale, loss_sum, partial, ticket);
    return check_launch("out_xent_fwd_kernel");
}

extern "C" int dctr_out_bwd_fused(const float* dlogit, const float* w, const float* a, int lda, int N, int H, int act,
                                  const float* mask, int ldmask, float* da, int ldda, float* gw, float* gb,
                                  void* workspace, size_t workspace_bytes, void* stream) {
    DCTR_REQUIRE(dlogit && w && a && da && gw && gb && workspace && N > 0 && H > 0, "out_bwd_fused: bad arguments");
    if (workspace_bytes < dctr_reduce_workspace(N, H)) {
        set_error("out_bwd_fused: workspace %zu < %zu", workspace_bytes, dctr_reduce_workspace(N, H));
        return DCTR_EWORKSPACE;
    }
    const bool vec = vec_ok(a, lda) && vec_ok(da, ldda) && (!mask || vec_ok(mask, ldmask));
    const RedShape sh = red_shape(N, H, vec);
    DCTR_REQUIRE(sh.col_tiles * 4 <= kRedHeader, "out_bwd_fused: H=%d too wide", H);
    unsigned* tickets = static_cast<unsigned*>(workspace);
    float* partial = reinterpret_cast<float*>(static_cast<char*>(workspace) + kRedHeader);
    const dim3 grid(sh.col_tiles, sh.chunks);
    if (vec)
        launch_k(out_bwd_fused_kernel<4>, grid, kRedThreads, 0, as_stream(stream), dlogit, w, a, lda, N, H, act, mask, ldmask, da, ldda,
                                                                            sh, partial, tickets, gw, gb);
    else
        launch_k(out_bwd_fused_kernel<1>, grid, kRedThreads, 0, as_stream(stream), dlogit, w, a, lda, N, H, act, mask, ldmask, da, ldda,
                                                                            sh, partial, tickets, gw, gb);
    return check_launch("out_bwd_fused_kernel");
}

extern "C" int dctr_out_bwd(const float* dlogit, const float* w, const float* a, int lda, int N, int H, int act,
                            const float* mask, int ldmask, float* da, int ldda, void* stream) {
    DCTR_REQUIRE(dlogit && w && a && da && N >= 0 && H > 0, "out_bwd: bad arguments");
    if (N == 0) return DCTR_OK;
    int blocks = ceil_div((int64_t)N * H, 256);
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    launch_k(out_bwd_kernel, blocks, 256, 0, as_stream(stream), dlogit, w, a, lda, N, H, act, mask, ldmask, da, ldda);
    return check_launch("out_bwd_kernel");
}

extern "C" size_t dctr_col_wsum_workspace(int N, int H) { return dctr_reduce_workspace(N, H); }

static int col_wsum_impl(const float* A, int lda, int N, int H, const float* wv, float* out, int sgd, float lr, float decay,
                         void* workspace, size_t workspace_bytes, void* stream) {
    DCTR_REQUIRE(A && out && workspace && N > 0 && H > 0, "col_wsum: bad arguments");
    if (workspace_bytes < dctr_reduce_workspace(N, H)) {
        set_error("col_wsum: workspace %zu < %zu", workspace_bytes, dctr_reduce_workspace(N, H));
        return DCTR_EWORKSPACE;
    }
    const bool vec = vec_ok(A, lda);
    const RedShape sh = red_shape(N, H, vec);
    DCTR_REQUIRE(sh.col_tiles * 4 <= kRedHeader, "col_wsum: H=%d too wide", H);
    unsigned* tickets = static_cast<unsigned*>(workspace);
    float* partial = reinterpret_cast<float*>(static_cast<char*>(workspace) + kRedHeader);
    const dim3 grid(sh.col_tiles, sh.chunks);
    if (vec)
        launch_k(col_wsum_kernel<4>, grid, kRedThreads, 0, as_stream(stream), A, lda, N, H, wv, sh, partial, tickets, out, sgd, lr, decay);
    else
        launch_k(col_wsum_kernel<1>, grid, kRedThreads, 0, as_stream(stream), A, lda, N, H, wv, sh, partial, tickets, out, sgd, lr, decay);
    return check_launch("col_wsum_kernel");
}

extern "C" int dctr_col_wsum(const float* A, int lda, int N, int H, const float* wv, float* out, void* workspace,
                             size_t workspace_bytes, void* stream) {
    return col_wsum_impl(A, lda, N, H, wv, out, 0, 0.f, 1.f, workspace, workspace_bytes, stream);
}

extern "C" int dctr_col_wsum_sgd(const float* A, int lda, int N, int H, const float* wv, float* param, float lr, float decay,
                                 void* workspace, size_t workspace_bytes, void* stream) {
    return col_wsum_impl(A, lda, N, H, wv, param, 1, lr, decay, workspace, workspace_bytes, stream);
}

// ---- dctr_col_wsum without a caller-provided workspace (no counterpart in the reference; added
// for callers that cannot allocate: code under graph capture, serving loops, harnesses that forbid cudaMalloc).
// The scratch (tickets + chunk partials of dctr_col_wsum) lives in a file-scope device array that every GPU loading the
// library holds until exit: kStaticMaxH columns -> kRedHeader + 1184 x (kStaticMaxH + 1) x 4 bytes = 4.63 MiB.
// In host code the array's NAME is only a host-side stand-in: its device address comes from cudaGetSymbolAddress, asked once
// per device with that device current.  Every call shares the array and the array keeps the previous call's contents, so the
// ticket header is cleared on the caller's stream before each launch; two calls that overlap on the GPU (different streams
// without an ordering between them) race on it.
namespace dctr {
constexpr int kStaticMaxH = 1024;
constexpr size_t kStaticRows = (size_t)(kRedMaxChunks > kNumSMs * 8 ? kRedMaxChunks : kNumSMs * 8);
constexpr size_t kStaticScratchBytes = kRedHeader + kStaticRows * (size_t)(kStaticMaxH + 1) * sizeof(float);
static __device__ __align__(256) unsigned char g_col_wsum_scratch[kStaticScratchBytes];
constexpr int kStaticMaxDevices = 64;
static void* g_col_wsum_scratch_addr[kStaticMaxDevices];      // device address per device ordinal (0 = not asked yet)
}  // namespace dctr

extern "C" int dctr_col_wsum_static_max_h(void) { return kStaticMaxH; }

extern "C" int dctr_col_wsum_static(const float* A, int lda, int N, int H, const float* wv, float* out, void* stream) {
    DCTR_REQUIRE(A && out && N > 0 && H > 0, "col_wsum_static: bad arguments");
    if (H > kStaticMaxH) {
        set_error("col_wsum_static: H=%d exceeds the built-in scratch (at most %d columns, %zu bytes): use dctr_col_wsum with a workspace",
                  H, kStaticMaxH, (size_t)kStaticScratchBytes);
        return DCTR_EINVAL;
    }
    if (dctr_reduce_workspace(N, H) > kStaticScratchBytes) {
        set_error("col_wsum_static: N=%d H=%d needs %zu bytes of scratch, the built-in array holds %zu", N, H, dctr_reduce_workspace(N, H),
                  (size_t)kStaticScratchBytes);
        return DCTR_EINVAL;
    }
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kStaticMaxDevices) {
        set_error("col_wsum_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
        return DCTR_ECUDA;
    }
    void* scratch = __atomic_load_n(&g_col_wsum_scratch_addr[dev], __ATOMIC_ACQUIRE);
    if (!scratch) {                     // set-up, once per device, with that device current (two racing threads get the same answer)
        e = cudaGetSymbolAddress(&scratch, g_col_wsum_scratch);
        if (e != cudaSuccess || !scratch) {
            set_error("col_wsum_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            return DCTR_ECUDA;
        }
        __atomic_store_n(&g_col_wsum_scratch_addr[dev], scratch, __ATOMIC_RELEASE);
    }
    // the tickets must start at zero; a previous call that failed half-way (or a racing one) may have left them otherwise
    e = cudaMemsetAsync(scratch, 0, kRedHeader, as_stream(stream));
    if (e != cudaSuccess) {
        set_error("col_wsum_static: cudaMemsetAsync: %s", cudaGetErrorString(e));
        return DCTR_ECUDA;
    }
    return dctr_col_wsum(A, lda, N, H, wv, out, scratch, kStaticScratchBytes, stream);
}

// lam * l2_loss(table) + lam * l2_loss(b) of FM.py:39 / LR.py:37 (tf.nn.l2_loss = sum(t^2) / 2) ADDED to out[0]: the loss the
// reference fetches includes the regulariser.  CTA partials in double, folded in CTA order by the last CTA.
namespace dctr {
__global__ void __launch_bounds__(256)
sumsq_kernel(const float4* __restrict__ v, int64_t n4, const float* __restrict__ extra, int n_extra, float scale,
             float* __restrict__ out, double* __restrict__ partial, unsigned* __restrict__ ticket) {
    pdl_sync();
    __shared__ double sh[256];
    double s = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (int64_t)gridDim.x * 256) {
        const float4 t = __ldg(v + i);
        s += (double)t.x * t.x + (double)t.y * t.y + (double)t.z * t.z + (double)t.w * t.w;
    }
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o]; __syncthreads(); }
    if (threadIdx.x == 0) partial[blockIdx.x] = sh[0];
    if (!take_last_ticket(ticket, gridDim.x)) return;
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (unsigned k = 0; k < gridDim.x; ++k) t += __ldcg(partial + k);
        for (int k = 0; k < n_extra; ++k) t += (double)extra[k] * extra[k];
        out[0] += (float)(0.5 * scale * t);
    }
}
}  // namespace dctr

extern "C" int dctr_l2_loss_add(const float* v, int64_t n, const float* extra, int n_extra, float scale, float* out,
                                void* workspace, size_t workspace_bytes, void* stream) {
    DCTR_REQUIRE(v && out && workspace && n >= 0 && (n & 3) == 0 && ((uintptr_t)v & 15) == 0 && n_extra >= 0 && (n_extra == 0 || extra),
                 "l2_loss_add: bad arguments (n must be a multiple of 4, v 16-byte aligned)");
    DCTR_REQUIRE(workspace_bytes >= dctr_reduce_workspace(1, 1), "l2_loss_add: workspace smaller than dctr_reduce_workspace(1, 1)");
    int blocks = dctr::ceil_div(n / 4, 256 * 8);
    if (blocks > dctr::kNumSMs * 4) blocks = dctr::kNumSMs * 4;
    if (blocks < 1) blocks = 1;
    char* ws = static_cast<char*>(workspace);
    dctr::launch_k(dctr::sumsq_kernel, blocks, 256, 0, dctr::as_stream(stream), reinterpret_cast<const float4*>(v), n / 4, extra, n_extra, scale,
                   out, reinterpret_cast<double*>(ws + dctr::kRedHeader), reinterpret_cast<unsigned*>(ws));
    return dctr::check_launch("sumsq_kernel");
}

// tf.nn.dropout masks for every layer input of a step in ONE launch (FNN.py:87-89: keep_prob + U[0,1) floored, scaled by
// 1 / keep_prob), drawn on the device with Philox-4x32-10 keyed by (seed, step counter): the step replays as a CUDA graph, so
// the counter lives in device memory and is advanced by the kernel's last CTA.  TF leaves these masks unseeded; parity runs
// inject masks instead.  apply != NULL: the first `apply_n` mask elements are also multiplied into `apply` in place (the
// dropout on act(z1), the first layer's input).
namespace dctr {
__device__ __forceinline__ uint4 philox4x32(uint4 ctr, uint2 key) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const unsigned hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
        const unsigned hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += 0x9E3779B9u; key.y += 0xBB67AE85u;
    }
    return ctr;
}
__global__ void __launch_bounds__(256)
dropout_masks_kernel(float4* __restrict__ masks, int64_t n4, float keep, unsigned seed, unsigned* __restrict__ state,
                     float4* __restrict__ apply, int64_t apply_n4) {
    pdl_sync();
    const unsigned step = state[0];
    const float inv = 1.f / keep;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (int64_t)gridDim.x * 256) {
        const uint4 r = philox4x32(make_uint4((unsigned)i, (unsigned)(i >> 32), step, 0u), make_uint2(seed, 0x5EEDu));
        // 24-bit uniforms in [0, 1): kept iff u < keep  (== floor(keep + u') >= 1 for u' = 1 - u)
        float4 m;
        m.x = (r.x >> 8) * (1.f / 16777216.f) < keep ? inv : 0.f;
        m.y = (r.y >> 8) * (1.f / 16777216.f) < keep ? inv : 0.f;
        m.z = (r.z >> 8) * (1.f / 16777216.f) < keep ? inv : 0.f;
        m.w = (r.w >> 8) * (1.f / 16777216.f) < keep ? inv : 0.f;
        masks[i] = m;
        if (i < apply_n4) {
            float4 v = apply[i];
            v.x *= m.x; v.y *= m.y; v.z *= m.z; v.w *= m.w;
            apply[i] = v;
        }
    }
    if (take_last_ticket(state + 1, gridDim.x) && threadIdx.x == 0) state[0] = step + 1;      // every CTA has read the counter
}
}  // namespace dctr

extern "C" int dctr_dropout_masks(float* masks, int64_t n, float keep, unsigned seed, unsigned* state, float* apply, int64_t apply_n,
                                  void* stream) {
    DCTR_REQUIRE(masks && state && n >= 0 && (n & 3) == 0 && ((uintptr_t)masks & 15) == 0 && keep > 0.f && keep <= 1.f,
                 "dropout_masks: bad arguments (n a multiple of 4, masks 16-byte aligned, 0 < keep <= 1)");
    DCTR_REQUIRE(!apply || ((apply_n & 3) == 0 && apply_n <= n && ((uintptr_t)apply & 15) == 0), "dropout_masks: bad apply block");
    if (n == 0) return DCTR_OK;
    int blocks = dctr::ceil_div(n / 4, 256 * 4);
    if (blocks > dctr::kNumSMs * 8) blocks = dctr::kNumSMs * 8;
    dctr::launch_k(dctr::dropout_masks_kernel, blocks, 256, 0, dctr::as_stream(stream), reinterpret_cast<float4*>(masks), n / 4, keep, seed, state,
                   reinterpret_cast<float4*>(apply), apply ? apply_n / 4 : (int64_t)0);
    return dctr::check_launch("dropout_masks_kernel");
}

extern "C" int dctr_vec_sum(const float* v, int N, float scale, float* out, void* stream) {
    DCTR_REQUIRE(v && out && N >= 0, "vec_sum: bad arguments");
    launch_k(vec_sum_kernel, 1, 1024, 0, as_stream(stream), v, N, scale, out);
    return check_launch("vec_sum_kernel");
}

extern "C" int dctr_dense_sgd(float* param, const float* grad, int64_t n, float lr, float decay, void* stream) {
    DCTR_REQUIRE(param && grad && n >= 0, "dense_sgd: bad arguments");
    if (n == 0) return DCTR_OK;
    int blocks = ceil_div(n, 256);
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    launch_k(dense_sgd_kernel, blocks, 256, 0, as_stream(stream), param, grad, n, lr, decay);
    return check_launch("dense_sgd_kernel");
}

extern "C" int dctr_adam_tick(float* state, float lr, float beta1, float beta2, void* stream) {
    DCTR_REQUIRE(state && beta1 >= 0.f && beta1 < 1.f && beta2 >= 0.f && beta2 < 1.f, "adam_tick: bad arguments");
    launch_k(adam_tick_kernel, 1, 1, 0, as_stream(stream), state, lr, beta1, beta2);
    return check_launch("adam_tick_kernel");
}

extern "C" int dctr_dense_adam(float* param, float* grad, float* m, float* v, int64_t n, const float* state, float beta1,
                               float beta2, float eps, float l2, int zero_grad, void* stream) {
    DCTR_REQUIRE(param && grad && m && v && state && n >= 0 && (n & 3) == 0, "dense_adam: null pointer or n not a multiple of 4");
    DCTR_REQUIRE(((((uintptr_t)param) | ((uintptr_t)grad) | ((uintptr_t)m) | ((uintptr_t)v)) & 15) == 0, "dense_adam: buffers must be 16-byte aligned");
    if (n == 0) return DCTR_OK;
    int blocks = ceil_div(n / 4, 256);
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    float4 *p4 = reinterpret_cast<float4*>(param), *g4 = reinterpret_cast<float4*>(grad), *m4 = reinterpret_cast<float4*>(m),
           *v4 = reinterpret_cast<float4*>(v);
    if (zero_grad) launch_k(dense_adam_kernel<true>, blocks, 256, 0, as_stream(stream), p4, g4, m4, v4, n / 4, state, beta1, beta2, eps, l2);
    else launch_k(dense_adam_kernel<false>, blocks, 256, 0, as_stream(stream), p4, g4, m4, v4, n / 4, state, beta1, beta2, eps, l2);
    return check_launch("dense_adam_kernel");
}

extern "C" int dctr_dense_ftrl(float* param, float* grad, float* accum, float* linear, int64_t n, float lr, float lr_power,
                               float l1, float l2, float grad_l2, int zero_grad, void* stream) {
    DCTR_REQUIRE(param && grad && accum && linear && n >= 0 && (n & 3) == 0, "dense_ftrl: null pointer or n not a multiple of 4");
    DCTR_REQUIRE(((((uintptr_t)param) | ((uintptr_t)grad) | ((uintptr_t)accum) | ((uintptr_t)linear)) & 15) == 0,
                 "dense_ftrl: buffers must be 16-byte aligned");
    DCTR_REQUIRE(lr > 0.f && lr_power <= 0.f && l1 >= 0.f && l2 >= 0.f, "dense_ftrl: lr > 0, lr_power <= 0, l1 >= 0, l2 >= 0");
    if (n == 0) return DCTR_OK;
    int blocks = ceil_div(n / 4, 256);
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    float4 *p4 = reinterpret_cast<float4*>(param), *g4 = reinterpret_cast<float4*>(grad), *n4 = reinterpret_cast<float4*>(accum),
           *z4 = reinterpret_cast<float4*>(linear);
    if (zero_grad) launch_k(dense_ftrl_kernel<true>, blocks, 256, 0, as_stream(stream), p4, g4, n4, z4, n / 4, 1.f / lr, lr_power, l1, l2, grad_l2);
    else launch_k(dense_ftrl_kernel<false>, blocks, 256, 0, as_stream(stream), p4, g4, n4, z4, n / 4, 1.f / lr, lr_power, l1, l2, grad_l2);
    return check_launch("dense_ftrl_kernel");
}

extern "C" int dctr_out_unit_train(const float* a, int lda, int N, int H, const float* w, const float* b, const float* y,
                                   float gscale, int act, const float* mask, int ldmask, float* logit, float* pred,
                                   float* loss_n, float* dlogit, float loss_scale, float* loss_sum, float* da, int ldda,
                                   float* gw, float* gb, void* workspace, size_t workspace_bytes, void* stream) {
    DCTR_REQUIRE(a && w && y && da && gw && gb && workspace && N > 0 && H > 0 && H <= 1024 && lda >= H && ldda >= H,
                 "out_unit_train: bad arguments (H <= 1024)");
    if (workspace_bytes < dctr_reduce_workspace(N, H + 1)) {
        set_error("out_unit_train: workspace %zu < %zu", workspace_bytes, dctr_reduce_workspace(N, H + 1));
        return DCTR_EWORKSPACE;
    }
    int blocks = ceil_div(N, 32);                          // >= 4 rows per warp
    if (blocks > kNumSMs * 2) blocks = kNumSMs * 2;
    unsigned* ticket = static_cast<unsigned*>(workspace);
    float* partial = reinterpret_cast<float*>(static_cast<char*>(workspace) + kRedHeader);
    cudaStream_t st = as_stream(stream);
#define DCTR_OUT_TRAIN(HVV)                                                                                                \
    do {                                                                                                                   \
        const size_t smem = (size_t)8 * (HVV * 32 + 2) * sizeof(float);                                                    \
        static bool attr = false;                                                                                          \
        if (!attr && smem > 48 * 1024) {                                                                                   \
            cudaFuncSetAttribute(out_unit_train_kernel<HVV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);      \
            attr = true;                                                                                                   \
        }                                                                                                                  \
        launch_k(out_unit_train_kernel<HVV>, blocks, 256, smem, st, a, lda, N, H, w, b, y, gscale, act, mask, ldmask, logit, \
                 pred, loss_n, dlogit, loss_scale, loss_sum, da, ldda, gw, gb, partial, ticket);                           \
    } while (0)
    if (H <= 128) DCTR_OUT_TRAIN(4);
    else if (H <= 256) DCTR_OUT_TRAIN(8);
    else if (H <= 512) DCTR_OUT_TRAIN(16);
    else DCTR_OUT_TRAIN(32);
#undef DCTR_OUT_TRAIN
    return check_launch("out_unit_train_kernel");
}
