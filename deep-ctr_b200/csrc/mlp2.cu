// Fused train step of a TWO-hidden-layer MLP for small batches: the network of the Theano script (177-300-100-1,
// FNN_wnzh.py:133-183) and of FNN.py at iPinYou sizes.  At batch 4096 such a step is ~2 GFLOP and 15 MB: as a chain of
// per-layer GEMM launches it is launch- and latency-bound (19 launches, every GEMM 13-29 us for 0.2-0.4 GFLOP; the ncu capture
// profiles/r02_ncu_fnn_gemm_tc_small.txt shows the tensor pipe 7-14 % active), and the layers are far too small for tensor-core
// tiles to pay (north_star: tensor cores only where the GEMMs are real).  Here
//
//   mlp2_train_kernel     one CTA per 32 samples walks the whole network forward AND backward with the tile's activations in
//                         shared memory: x -> a1 -> a2 -> output unit (logit, pred, loss, dlogit, d w_out, d b_out) -> dz2 ->
//                         dz1 -> dz0 (= Theano's gx).  fp32 FFMA with a 4-sample x 4*NQ-column register tile per thread:
//                         warp = 4 samples, lane = columns {lane + 32 i}.  Activation tiles are kept TRANSPOSED ([column][sample],
//                         row stride 36) so that the 4 samples of a warp are one broadcast 128-bit read per k and the epilogue's
//                         128-bit stores are conflict-free; the weights stream through a 6-stage cp.async ring in slabs of 8
//                         k-rows (each slab is read from L2 once per CTA).  The backward GEMMs read TRANSPOSED copies of the
//                         weights (mlp2_transpose_kernel, a side-stream launch per step) so every phase is the same row-major
//                         [k][column] stream.  Leaves x (masked), a1, dz1, dz2 in global memory for the weight gradients.
//   mlp2_wgrad_kernel     gW1 = x^T dz1, gW2 = a1^T dz2 and both bias gradients as ONE launch of 8-CTA clusters, one 64x64 output tile per
//                         cluster; the 8 CTAs split the batch, park their partial tiles in shared memory and add them in rank
//                         order through distributed shared memory (fixed order, no float atomics, no global round trip), then
//                         store the gradient or apply plain SGD in place.
//
// What bounds the first kernel: shared-memory -> register bandwidth (128 B/clk/SM).  A thread tile of R samples x C columns
// needs (R + C) / (R C) operand floats per FFMA; with 32 samples x 300 columns per 256 threads R = 4, C = 12 is what fits:
// 0.33 floats per FFMA = 171 B/clk at the full 128 FFMA/clk, i.e. ~75 % of the fp32 peak at best.  (The first version used one
// sample per lane -- 1 float per FFMA -- and ran at 9 % of the peak: profiles/r02_bench_fnn_ipinyou_mlp2_v1.json.)
// Results are deterministic (fixed reduction orders) and equal to the per-layer path up to fp32 summation order.
#include "common.cuh"

namespace dctr {

constexpr int kM2Threads = 256;
constexpr int kM2Warps = 8;
constexpr int kM2Rows = 32;                 // samples per CTA: warp w owns samples 4w .. 4w+3
constexpr int kM2KC = 8;                    // k-depth of a weight slab
constexpr int kM2MaxW = 384;                // widest layer / input: 3 groups of 128 columns
constexpr int kM2StageFloats = kM2MaxW * kM2KC;      // 12 KB
constexpr int kM2Stages = 6;
constexpr int kM2LdT = 36;                  // row stride of a transposed activation tile [column][32 samples + 4]

struct Mlp2Args {
    int N, D0, H1, H2;
    int ld0, ld1, ldw1, ldw2, ldd;          // row strides: x / a1 (global), W1 / W2, dz1 / dz2
    int act0, act1, act2;
    float* x; const float* m0; int ldm0;    // x = act0(z) [N, ld0]; m0 != NULL: x *= m0 in place (dropout on the first layer's input)
    const float* W1; const float* b1; const float* W2; const float* b2; float* w3; float* b3;
    const float* W1t; const float* W2t;     // W1^T [H1, ld0], W2^T [H2, ld1]
    const float* m1; int ldm1; const float* m2; int ldm2;        // masks on a1 / a2 (ld 0 = one broadcast row), NULL = none
    const float* y; float gscale, loss_scale;
    float* a1;
    float* dz2; float* dz1; float* dz0;
    float* logit; float* pred; float* loss_n; float* dlogit; float* loss_sum;
    float* gw3; float* gb3;                 // written unless sgd_out != 0
    int sgd_out; float lr, decay_out;       // sgd_out: the last CTA applies w3 = w3 * decay_out - lr * gw3 (b3 likewise) itself
    float* partial; unsigned* ticket;       // [tiles][H2 + 2] CTA partials of (gw3 | gb3 | loss), ticket word
};

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__host__ __device__ inline int m2_kpad(int w) { return (w + kM2KC - 1) / kM2KC * kM2KC; }      // tile rows: whole slabs

// ---- streaming GEMM  acc[sample u][column] += sum_k At[k][4 warp + u] * B[k][column],  B row-major [K, ldb] -------------------
struct Slab { const float* B; int ldb, K, nchunks; };

__device__ __forceinline__ Slab make_slab(const float* B, int ldb, int K) {
    Slab s; s.B = B; s.ldb = ldb; s.K = K; s.nchunks = (K + kM2KC - 1) / kM2KC; return s;
}
__device__ __forceinline__ void slab_issue(const Slab& s, int chunk, float* stage) {
    if (chunk < s.nchunks) {
        // rows k0 .. k0+7 of B: kM2KC * ldb contiguous floats; rows past K are zero-filled (the A tile is zero there as well)
        const int k0 = chunk * kM2KC;
        const int rows = min(kM2KC, s.K - k0);
        const int n4 = rows * s.ldb / 4, tot4 = kM2KC * s.ldb / 4;
        const float4* src = reinterpret_cast<const float4*>(s.B + (size_t)k0 * s.ldb);
        for (int i = threadIdx.x; i < n4; i += kM2Threads) cp_async16(stage + 4 * i, src + i);
        for (int i = n4 + threadIdx.x; i < tot4; i += kM2Threads) *reinterpret_cast<float4*>(stage + 4 * i) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    cp_async_commit();            // an empty group keeps the wait_group arithmetic uniform
}
__device__ __forceinline__ void slab_prologue(const Slab& s, float* ring) {
#pragma unroll
    for (int c = 0; c < kM2Stages - 1; ++c) slab_issue(s, c, ring + c * kM2StageFloats);
}

// NQ groups of 128 columns; thread columns: lane + 32 i + 128 j  (i < 4, j < NQ)
template <int NQ>
__device__ __forceinline__ void slab_mainloop(const Slab& s, float* ring, const float* __restrict__ at, float (&acc)[NQ][4][4]) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int j = 0; j < NQ; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int u = 0; u < 4; ++u) acc[j][i][u] = 0.f;
    const float* arow = at + 4 * warp;
    for (int c = 0; c < s.nchunks; ++c) {
        cp_async_wait<kM2Stages - 2>();
        __syncthreads();                                   // slab c has landed for everyone; slab c-1 is no longer read
        slab_issue(s, c + kM2Stages - 1, ring + ((c + kM2Stages - 1) % kM2Stages) * kM2StageFloats);
        const float* st = ring + (c % kM2Stages) * kM2StageFloats + lane;
        // unrolled by 2 only: the fully unrolled body (8 k x 61 instructions = 7.8 KB) did not fit the instruction cache and the
        // kernel spent 2.9 cycles per issued instruction waiting for fetches (profiles/r02_ncu_mlp2_v2.txt)
#pragma unroll 2
        for (int k = 0; k < kM2KC; ++k) {
            const float4 a = *reinterpret_cast<const float4*>(arow + (c * kM2KC + k) * kM2LdT);      // the warp's 4 samples: broadcast
            const float* b = st + k * s.ldb;
#pragma unroll
            for (int j = 0; j < NQ; ++j)
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float w = b[32 * i + 128 * j];
                    acc[j][i][0] = fmaf(a.x, w, acc[j][i][0]); acc[j][i][1] = fmaf(a.y, w, acc[j][i][1]);
                    acc[j][i][2] = fmaf(a.z, w, acc[j][i][2]); acc[j][i][3] = fmaf(a.w, w, acc[j][i][3]);
                }
        }
    }
    cp_async_wait<0>();
    __syncthreads();                                       // the ring may be refilled, the A tile may be overwritten
}

// epilogue: f(u, column, value) for every column < nout of this thread; the 4 samples of a column leave as one 128-bit store
// into the transposed tile `tt` (which may be the tile f reads its own old values from)
template <int NQ, class F>
__device__ __forceinline__ void tile_store(float (&acc)[NQ][4][4], float* __restrict__ tt, int nout, F f) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
    for (int j = 0; j < NQ; ++j)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int c = lane + 32 * i + 128 * j;
            if (c < nout) {
                float4 v;
                v.x = f(0, c, acc[j][i][0]); v.y = f(1, c, acc[j][i][1]); v.z = f(2, c, acc[j][i][2]); v.w = f(3, c, acc[j][i][3]);
                *reinterpret_cast<float4*>(tt + c * kM2LdT + 4 * warp) = v;
            }
        }
}

template <class F>
__device__ __forceinline__ void gemm_phase(const Slab& s, float* ring, const float* __restrict__ at, float* __restrict__ tt, int nout, F f) {
    if (nout <= 128) { float acc[1][4][4]; slab_mainloop<1>(s, ring, at, acc); tile_store<1>(acc, tt, nout, f); }
    else if (nout <= 256) { float acc[2][4][4]; slab_mainloop<2>(s, ring, at, acc); tile_store<2>(acc, tt, nout, f); }
    else { float acc[3][4][4]; slab_mainloop<3>(s, ring, at, acc); tile_store<3>(acc, tt, nout, f); }
}

// transposed tile [column][sample] -> global rows [sample][ldg] (w4 float4 per row, columns >= width written as 0), coalesced
__device__ __forceinline__ void tile_to_global(const float* __restrict__ tt, int width, float* __restrict__ g, int ldg, int w4, int rows) {
    for (int i = threadIdx.x; i < rows * w4; i += kM2Threads) {
        const int r = i / w4, q = i - r * w4, c = 4 * q;
        float4 v;
        v.x = c < width ? tt[c * kM2LdT + r] : 0.f;
        v.y = c + 1 < width ? tt[(c + 1) * kM2LdT + r] : 0.f;
        v.z = c + 2 < width ? tt[(c + 2) * kM2LdT + r] : 0.f;
        v.w = c + 3 < width ? tt[(c + 3) * kM2LdT + r] : 0.f;
        *reinterpret_cast<float4*>(g + (size_t)r * ldg + c) = v;
    }
}

// derivative of `act` at a masked activation value v = m * act(z): m * act'(act(z)), 0 where the unit was dropped.
// Both helpers are deliberately NOT inlined: the epilogues call them for 48 register values at 12 call sites, and the inlined
// tanhf / expf expansions made the kernel 670 KB of code (instruction-fetch bound, profiles/r02_ncu_mlp2_v2.txt).
__device__ __noinline__ float dact_masked(int act, float v, float m) { return m != 0.f ? m * act_bwd_from_out(act, v / m) : 0.f; }
__device__ __noinline__ float m2_act(int act, float x) { return act_fwd(act, x); }

__global__ void __launch_bounds__(kM2Threads, 1)
mlp2_train_kernel(const Mlp2Args a) {
    pdl_sync();
    extern __shared__ __align__(16) float m2smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int k0p = m2_kpad(a.D0), k1p = m2_kpad(a.H1), k2p = m2_kpad(a.H2);
    float* xt = m2smem;                              // [k0p][36]   x^T, later dz0^T
    float* a1t = xt + k0p * kM2LdT;                  // [k1p][36]   a1^T, later dz1^T
    float* a2t = a1t + k1p * kM2LdT;                 // [k2p][36]   a2^T, later dz2^T
    float* ring = a2t + k2p * kM2LdT;                // [kM2Stages][kM2StageFloats]
    float* dl = ring + kM2Stages * kM2StageFloats;   // [32] dlogit of the tile
    const int64_t r0 = (int64_t)blockIdx.x * kM2Rows;
    const int rows = (int)min((int64_t)kM2Rows, (int64_t)a.N - r0);
    // global sample of this thread's register row u (idle rows of the last tile alias a valid one: they are never stored)
    int64_t grow[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) grow[u] = 4 * warp + u < rows ? r0 + 4 * warp + u : r0;

    Slab s1 = make_slab(a.W1, a.ldw1, a.D0);
    slab_prologue(s1, ring);                         // the first weight slabs travel while the x tile is loaded

    // ---- tiles start as zeros (pad rows of whole slabs), then x^T: lane = sample, a warp per group of 4 columns ----
    for (int i = threadIdx.x; i < (k0p + k1p + k2p) * kM2LdT / 4; i += kM2Threads) *reinterpret_cast<float4*>(m2smem + 4 * i) = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncthreads();
    for (int q = warp; q < (a.ld0 >> 2); q += kM2Warps) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (lane < rows) {
            float* px = a.x + (size_t)(r0 + lane) * a.ld0 + 4 * q;
            v = *reinterpret_cast<const float4*>(px);
            if (a.m0) {
                const float4 m = *reinterpret_cast<const float4*>(a.m0 + (size_t)(a.ldm0 ? r0 + lane : 0) * a.ldm0 + 4 * q);
                v.x *= m.x; v.y *= m.y; v.z *= m.z; v.w *= m.w;
                *reinterpret_cast<float4*>(px) = v;                  // the weight gradient reads the masked input
            }
        }
        const int c = 4 * q;
        if (c < a.D0) xt[c * kM2LdT + lane] = v.x;
        if (c + 1 < a.D0) xt[(c + 1) * kM2LdT + lane] = v.y;
        if (c + 2 < a.D0) xt[(c + 2) * kM2LdT + lane] = v.z;
        if (c + 3 < a.D0) xt[(c + 3) * kM2LdT + lane] = v.w;
    }
    __syncthreads();

    auto mask_of = [&](const float* m, int ldm, int u, int c) { return m ? __ldg(m + (ldm ? grow[u] : 0) * ldm + c) : 1.f; };

    // ---- layer 1: a1 = m1 * act1(x W1 + b1) ----
    gemm_phase(s1, ring, xt, a1t, a.H1, [&](int u, int c, float v) { return m2_act(a.act1, v + __ldg(a.b1 + c)) * mask_of(a.m1, a.ldm1, u, c); });
    Slab s2 = make_slab(a.W2, a.ldw2, a.H1);
    slab_prologue(s2, ring);
    __syncthreads();
    tile_to_global(a1t, a.H1, a.a1 + (size_t)r0 * a.ld1, a.ld1, a.ld1 >> 2, rows);

    // ---- layer 2: a2 = m2 * act2(a1 W2 + b2) ----
    gemm_phase(s2, ring, a1t, a2t, a.H2, [&](int u, int c, float v) { return m2_act(a.act2, v + __ldg(a.b2 + c)) * mask_of(a.m2, a.ldm2, u, c); });
    Slab s3 = make_slab(a.W2t, a.ld1, a.H2);         // dz2 * W2^T comes next: B = W2^T [H2, ld1]
    slab_prologue(s3, ring);
    __syncthreads();

    // ---- output unit: logit, pred, loss, dlogit (one warp, lane = sample), CTA partials of d w_out | d b_out | loss ----
    float* part = a.partial + (size_t)blockIdx.x * (a.H2 + 2);
    if (warp == 0) {
        float z = 0.f;
        for (int c = 0; c < a.H2; ++c) z = fmaf(a2t[c * kM2LdT + lane], a.w3[c], z);
        z += a.b3[0];
        const float p = sigmoidf_(z);
        float ln = 0.f, d = 0.f;
        if (lane < rows) {
            const int64_t n = r0 + lane;
            const float yn = a.y[n];
            ln = xent_logits(z, yn);
            d = (p - yn) * a.gscale;
            if (a.logit) a.logit[n] = z;
            if (a.pred) a.pred[n] = p;
            if (a.loss_n) a.loss_n[n] = ln;
            if (a.dlogit) a.dlogit[n] = d;
        }
        dl[lane] = d;
        float ls = ln, ds = d;                       // fixed-order butterfly over the 32 samples
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { ls += __shfl_xor_sync(0xffffffffu, ls, o); ds += __shfl_xor_sync(0xffffffffu, ds, o); }
        if (lane == 0) { part[a.H2] = ds; part[a.H2 + 1] = ls; }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < a.H2; c += kM2Threads) {       // d w_out[c] = sum_n dlogit[n] a2[n][c], samples in order
        float g = 0.f;
#pragma unroll 8
        for (int r = 0; r < kM2Rows; ++r) g = fmaf(dl[r], a2t[c * kM2LdT + r], g);
        part[c] = g;
    }
    __syncthreads();
    // ---- dz2 = dlogit * w_out * m2 * act2'(.)  in place of a2 (same thread <-> element mapping as the GEMM epilogues) ----
    {
        const float4 d4 = *reinterpret_cast<const float4*>(dl + 4 * warp);
        const float dd[4] = {d4.x, d4.y, d4.z, d4.w};
        for (int c = lane; c < a.H2; c += 32) {
            float4 v = *reinterpret_cast<const float4*>(a2t + c * kM2LdT + 4 * warp);
            const float w = a.w3[c];
            float* pv = &v.x;
#pragma unroll
            for (int u = 0; u < 4; ++u) pv[u] = dd[u] * w * dact_masked(a.act2, pv[u], mask_of(a.m2, a.ldm2, u, c));
            *reinterpret_cast<float4*>(a2t + c * kM2LdT + 4 * warp) = v;
        }
    }
    __syncthreads();
    tile_to_global(a2t, a.H2, a.dz2 + (size_t)r0 * a.ldd, a.ldd, (a.H2 + 3) >> 2, rows);

    // ---- dz1 = (dz2 W2^T) * m1 * act1'(.)  in place of a1 ----
    gemm_phase(s3, ring, a2t, a1t, a.H1,
               [&](int u, int c, float v) { return v * dact_masked(a.act1, a1t[c * kM2LdT + 4 * warp + u], mask_of(a.m1, a.ldm1, u, c)); });
    Slab s4 = make_slab(a.W1t, a.ld0, a.H1);         // B = W1^T [H1, ld0]
    slab_prologue(s4, ring);
    __syncthreads();
    tile_to_global(a1t, a.H1, a.dz1 + (size_t)r0 * a.ldd, a.ldd, (a.H1 + 3) >> 2, rows);

    // ---- dz0 = (dz1 W1^T) * m0 * act0'(.)  in place of x: the gradient the sparse update folds (Theano's gx) ----
    gemm_phase(s4, ring, a1t, xt, a.D0,
               [&](int u, int c, float v) { return v * dact_masked(a.act0, xt[c * kM2LdT + 4 * warp + u], mask_of(a.m0, a.ldm0, u, c)); });
    __syncthreads();
    tile_to_global(xt, a.D0, a.dz0 + (size_t)r0 * a.ld0, a.ld0, a.ld0 >> 2, rows);

    // ---- last CTA: d w_out, d b_out and the loss from the CTA partials, in CTA order ----
    __shared__ unsigned s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(a.ticket, 1u);
        s_last = (t == gridDim.x - 1) ? 1u : 0u;
        if (s_last) *a.ticket = 0u;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const int W = a.H2 + 2;
    for (int c = threadIdx.x; c < W; c += kM2Threads) {
        float s = 0.f;
#pragma unroll 8
        for (unsigned k = 0; k < gridDim.x; ++k) s += __ldcg(a.partial + (size_t)k * W + c);
        if (c == a.H2 + 1) { if (a.loss_sum) a.loss_sum[0] = s * a.loss_scale; }
        else if (a.sgd_out) {
            float* p = c < a.H2 ? a.w3 + c : a.b3;      // every CTA has finished reading w_out / b_out: the last ticket is taken
            *p = fmaf(*p, a.decay_out, -a.lr * s);
        } else {
            if (c < a.H2) a.gw3[c] = s; else a.gb3[0] = s;
        }
    }
}

// out[j][i] = in[i][j] for i < I, j < J; pad columns of out (i in [I, ldo)) are written as 0.  32x32 tiles through shared memory.
__global__ void __launch_bounds__(256)
mlp2_transpose_kernel(const float* __restrict__ W1, int ldw1, int D0, int H1, float* __restrict__ W1t, int ld0,
                      const float* __restrict__ W2, int ldw2, int H2, float* __restrict__ W2t, int ld1, int tiles1) {
    pdl_sync();
    __shared__ float t[32][33];
    const bool second = (int)blockIdx.x >= tiles1;
    const float* in = second ? W2 : W1;
    float* out = second ? W2t : W1t;
    const int I = second ? H1 : D0, J = second ? H2 : H1, ldi = second ? ldw2 : ldw1, ldo = second ? ld1 : ld0;
    const int b = second ? blockIdx.x - tiles1 : blockIdx.x;
    const int tj = (J + 31) / 32;
    const int i0 = (b / tj) * 32, j0 = (b % tj) * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int r = ty; r < 32; r += 8) t[r][tx] = (i0 + r < I && j0 + tx < J) ? in[(size_t)(i0 + r) * ldi + j0 + tx] : 0.f;
    __syncthreads();
    for (int r = ty; r < 32; r += 8)
        if (j0 + r < J && i0 + tx < ldo) out[(size_t)(j0 + r) * ldo + i0 + tx] = t[tx][r];
}

// ---- weight gradients of both layers (+ bias gradients) as one launch of 8-CTA clusters ------------------------------------------
// Work item = (layer, 64x64 output tile) per CLUSTER (25 tiles x 8 CTAs for 177-300-100: several CTAs per SM, a 128x64 tile with
// one 4-warp CTA per SM left every scheduler with a single warp, 47 % issue activity); the 8 CTAs of the cluster split the batch (k) dimension.  A[k][i] = layer
// input (x or a1, [N, lda]), B[k][j] = dz of the layer.  CTA: 128 threads, thread (ty, tx) owns outputs i0 + 4 ty .. +3 (ty < 16) and
// columns j0 + 4 tx .. +3 and j0 + 32 + 4 tx .. +3 (tx < 8): per k three 128-bit operand reads for 32 FFMA; k in slabs of 16 samples
// through a 3-stage cp.async ring.  The split-K reduction never touches global memory: every CTA parks its partial tile in its own
// shared memory, and after a cluster barrier CTA r adds rows 8 r .. 8 r + 7 of the 8 partials in rank order through distributed
// shared memory (fixed order: deterministic, no float atomics) and stores the gradient or applies plain SGD in place.  Rank 0 also
// folds the bias gradient (column sums of dz) of the tiles with i0 == 0.
constexpr int kWgTI = 64, kWgTJ = 64, kWgK = 16, kWgThreads = 128, kWgStages = 3, kWgSplit = 8;
constexpr int kWgRing = kWgStages * kWgK * (kWgTI + kWgTJ);       // floats: 9216
constexpr int kWgPart = kWgTI * kWgTJ + kWgTJ;                    // floats: the partial tile + its bias row
constexpr int kWgSmem = (kWgRing > kWgPart ? kWgRing : kWgPart) * 4;

struct WgLayer { const float* A; int lda; const float* B; int ldb; int K, I, J; float* W; int ldw; float* bvec; float* gW; float* gb; float decay; int tiles_j, tile0; };
struct WgArgs {
    WgLayer L[2];
    int rows_per_split;
    int sgd; float lr;
};

__global__ void __launch_bounds__(kWgThreads, 5)
mlp2_wgrad_kernel(const WgArgs a) {
    pdl_sync();
    extern __shared__ __align__(16) float wsm[];
    float (*As)[kWgK][kWgTI] = reinterpret_cast<float (*)[kWgK][kWgTI]>(wsm);
    float (*Bs)[kWgK][kWgTJ] = reinterpret_cast<float (*)[kWgK][kWgTJ]>(wsm + kWgStages * kWgK * kWgTI);
    unsigned split, nsplit;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(split));
    asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(nsplit));
    const int tile = blockIdx.x / kWgSplit;
    const WgLayer& L = tile >= a.L[1].tile0 ? a.L[1] : a.L[0];
    const int t = tile - L.tile0;
    const int ti_ = t / L.tiles_j, tj_ = t - ti_ * L.tiles_j;
    const int i0 = ti_ * kWgTI, j0 = tj_ * kWgTJ;
    const int k_lo = (int)split * a.rows_per_split, k_hi = min(L.K, k_lo + a.rows_per_split);
    const int tid = threadIdx.x, ty = tid >> 3, tx = tid & 7;
    float acc[4][8];
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v < 8; ++v) acc[u][v] = 0.f;
    float bsum[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    const bool do_bias = (i0 == 0) && ty == 0;
    const int nslabs = k_hi > k_lo ? (k_hi - k_lo + kWgK - 1) / kWgK : 0;
    // slab: 16 rows x 64 floats of A and of B (256 float4 each): 2 + 2 pieces per thread, zero outside the matrix
    auto load = [&](int s, int buf) {
        if (s < nslabs) {
            const int kb = k_lo + s * kWgK;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int idx = tid + kWgThreads * h, r = idx >> 4, q = (idx & 15) * 4;
                float* d = &As[buf][r][q];
                if (kb + r < k_hi && i0 + q + 3 < L.lda) cp_async16(d, L.A + (size_t)(kb + r) * L.lda + i0 + q);
                else *reinterpret_cast<float4*>(d) = make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int idx = tid + kWgThreads * h, r = idx >> 4, q = (idx & 15) * 4;
                float* d = &Bs[buf][r][q];
                if (kb + r < k_hi && j0 + q + 3 < L.ldb) cp_async16(d, L.B + (size_t)(kb + r) * L.ldb + j0 + q);
                else *reinterpret_cast<float4*>(d) = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
        cp_async_commit();
    };
    load(0, 0);
    load(1, 1);
    for (int s = 0; s < nslabs; ++s) {
        cp_async_wait<kWgStages - 2>();
        __syncthreads();
        load(s + 2, (s + 2) % kWgStages);
        const int buf = s % kWgStages;
#pragma unroll 4
        for (int k = 0; k < kWgK; ++k) {
            const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][k][4 * ty]);
            const float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][k][4 * tx]), b1 = *reinterpret_cast<const float4*>(&Bs[buf][k][32 + 4 * tx]);
            const float aa[4] = {a0.x, a0.y, a0.z, a0.w}, bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int v = 0; v < 8; ++v) acc[u][v] = fmaf(aa[u], bb[v], acc[u][v]);
            if (do_bias) {
#pragma unroll
                for (int v = 0; v < 8; ++v) bsum[v] += bb[v];
            }
        }
    }
    cp_async_wait<0>();
    __syncthreads();                 // the operand ring is dead: the partial tile takes its place
    float* part = wsm;               // [64][64] | bias [64]
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        float* pr = part + (4 * ty + u) * kWgTJ;
        *reinterpret_cast<float4*>(pr + 4 * tx) = make_float4(acc[u][0], acc[u][1], acc[u][2], acc[u][3]);
        *reinterpret_cast<float4*>(pr + 32 + 4 * tx) = make_float4(acc[u][4], acc[u][5], acc[u][6], acc[u][7]);
    }
    if (do_bias) {
        float* pb = part + kWgTI * kWgTJ;
        *reinterpret_cast<float4*>(pb + 4 * tx) = make_float4(bsum[0], bsum[1], bsum[2], bsum[3]);
        *reinterpret_cast<float4*>(pb + 32 + 4 * tx) = make_float4(bsum[4], bsum[5], bsum[6], bsum[7]);
    }
    // cluster barrier (release / acquire): every partial tile is visible to the whole cluster
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    // ---- CTA r: rows 8 r .. 8 r + 7 of the tile (128 float4, one per thread) over the partials of ranks 0 .. 7, in rank order ----
    const unsigned my = (unsigned)__cvta_generic_to_shared(part);
    const int rpc = kWgTI / kWgSplit;            // 8 rows per CTA
    for (int e = tid; e < (rpc + (split == 0 ? 1 : 0)) * (kWgTJ / 4); e += kWgThreads) {
        const int rr = e / (kWgTJ / 4), q = (e - rr * (kWgTJ / 4)) * 4;
        const bool bias_row = rr == rpc;
        const int r = bias_row ? kWgTI : (int)split * rpc + rr;
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (unsigned s = 0; s < kWgSplit; ++s) {
            unsigned remote;
            asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(my + (unsigned)(r * kWgTJ + q) * 4u), "r"(s));
            float4 p;
            asm volatile("ld.shared::cluster.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(p.x), "=f"(p.y), "=f"(p.z), "=f"(p.w) : "r"(remote));
            g.x += p.x; g.y += p.y; g.z += p.z; g.w += p.w;
        }
        const float gg[4] = {g.x, g.y, g.z, g.w};
        if (!bias_row) {
            const int i = i0 + r;
            if (i < L.I) {
#pragma unroll
                for (int v = 0; v < 4; ++v) {
                    const int j = j0 + q + v;
                    if (j < L.J) {
                        if (a.sgd) { float* w = L.W + (size_t)i * L.ldw + j; *w = fmaf(*w, L.decay, -a.lr * gg[v]); }
                        else L.gW[(size_t)i * L.ldw + j] = gg[v];
                    }
                }
            }
        } else if (i0 == 0) {
#pragma unroll
            for (int v = 0; v < 4; ++v) {
                const int j = j0 + q + v;
                if (j < L.J) {
                    if (a.sgd) L.bvec[j] = fmaf(L.bvec[j], L.decay, -a.lr * gg[v]);
                    else L.gb[j] = gg[v];
                }
            }
        }
    }
    // nobody leaves while its shared memory may still be read by a peer
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

static size_t mlp2_smem(int D0, int H1, int H2) {
    return (size_t)((m2_kpad(D0) + m2_kpad(H1) + m2_kpad(H2)) * kM2LdT + kM2Stages * kM2StageFloats + kM2Rows) * sizeof(float);
}
static int wg_tiles(int I, int J) { return ceil_div(I, kWgTI) * ceil_div(J, kWgTJ); }

// workspace layout (dctr_mlp2_workspace): ticket | CTA partials of the output unit
struct Mlp2Ws { unsigned* ticket; float* partial; };
static Mlp2Ws mlp2_ws(void* workspace) {
    Mlp2Ws w;
    char* ws = static_cast<char*>(workspace);
    w.ticket = reinterpret_cast<unsigned*>(ws);
    w.partial = reinterpret_cast<float*>(ws + 256);
    return w;
}

}  // namespace dctr

using namespace dctr;

extern "C" int dctr_mlp2_supported(int D0, int H1, int H2) {
    return D0 > 0 && H1 > 0 && H2 > 0 && D0 <= kM2MaxW && H1 <= kM2MaxW && H2 <= kM2MaxW && mlp2_smem(D0, H1, H2) <= 200 * 1024;
}

extern "C" size_t dctr_mlp2_workspace(int N, int D0, int H1, int H2) {
    if (N <= 0) return 1024;
    (void)D0; (void)H1;
    return 1024 + (size_t)ceil_div(N, kM2Rows) * (H2 + 2) * 4;
}

extern "C" int dctr_mlp2_transpose(int D0, int H1, int H2, const float* W1, int ldw1, const float* W2, int ldw2, float* W1t, int ld0,
                                   float* W2t, int ld1, void* stream) {
    DCTR_REQUIRE(dctr_mlp2_supported(D0, H1, H2) && W1 && W2 && W1t && W2t && ldw1 >= H1 && ldw2 >= H2 && ld0 >= D0 && ld1 >= H1,
                 "mlp2_transpose: bad arguments");
    const int t1 = ceil_div(ld0, 32) * ceil_div(H1, 32), t2 = ceil_div(ld1, 32) * ceil_div(H2, 32);
    launch_k(mlp2_transpose_kernel, t1 + t2, 256, 0, as_stream(stream), W1, ldw1, D0, H1, W1t, ld0, W2, ldw2, H2, W2t, ld1, t1);
    return check_launch("mlp2_transpose_kernel");
}

extern "C" int dctr_mlp2_train(int N, int D0, int H1, int H2, float* x, int ld0, const float* m0, int ldm0, int act0,
                               const float* W1, int ldw1, const float* b1, int act1, const float* m1, int ldm1,
                               const float* W2, int ldw2, const float* b2, int act2, const float* m2, int ldm2,
                               const float* W1t, const float* W2t, float* w3, float* b3, const float* y, float gscale, float loss_scale,
                               float* a1, int ld1, float* dz1, float* dz2, int ldd, float* dz0,
                               float* logit, float* pred, float* loss_n, float* dlogit, float* loss_sum,
                               float* gw3, float* gb3, int sgd, float lr, float decay_out,
                               void* workspace, size_t workspace_bytes, void* stream) {
    DCTR_REQUIRE(N >= 0 && dctr_mlp2_supported(D0, H1, H2), "mlp2_train: unsupported shape D0=%d H1=%d H2=%d (each <= %d)", D0, H1, H2, kM2MaxW);
    if (N == 0) return DCTR_OK;
    DCTR_REQUIRE(x && W1 && b1 && W2 && b2 && W1t && W2t && w3 && b3 && y && a1 && dz1 && dz2 && dz0 && workspace, "mlp2_train: null pointer");
    DCTR_REQUIRE(sgd || (gw3 && gb3), "mlp2_train: gradient outputs missing");
    DCTR_REQUIRE((ld0 & 3) == 0 && (ld1 & 3) == 0 && (ldw1 & 3) == 0 && (ldw2 & 3) == 0 && (ldd & 3) == 0 && ld0 >= D0 && ld1 >= H1 &&
                 ldw1 >= H1 && ldw2 >= H2 && ldd >= H1 && ldd >= H2 && (!m0 || (ldm0 & 3) == 0) &&
                 ldw1 <= kM2MaxW && ldw2 <= kM2MaxW && ld0 <= kM2MaxW && ld1 <= kM2MaxW,
                 "mlp2_train: leading dimensions must be multiples of 4, cover the widths and not exceed %d", kM2MaxW);
    DCTR_REQUIRE((((uintptr_t)x | (uintptr_t)W1 | (uintptr_t)W2 | (uintptr_t)W1t | (uintptr_t)W2t | (uintptr_t)a1 | (uintptr_t)dz1 |
                   (uintptr_t)dz2 | (uintptr_t)dz0 | (uintptr_t)workspace | (uintptr_t)m0) & 15) == 0, "mlp2_train: buffers must be 16-byte aligned");
    if (workspace_bytes < dctr_mlp2_workspace(N, D0, H1, H2)) {
        set_error("mlp2_train: workspace %zu < %zu bytes", workspace_bytes, dctr_mlp2_workspace(N, D0, H1, H2));
        return DCTR_EWORKSPACE;
    }
    const Mlp2Ws w = mlp2_ws(workspace);
    Mlp2Args a{};
    a.N = N; a.D0 = D0; a.H1 = H1; a.H2 = H2; a.ld0 = ld0; a.ld1 = ld1; a.ldw1 = ldw1; a.ldw2 = ldw2; a.ldd = ldd;
    a.act0 = act0; a.act1 = act1; a.act2 = act2;
    a.x = x; a.m0 = m0; a.ldm0 = ldm0;
    a.W1 = W1; a.b1 = b1; a.W2 = W2; a.b2 = b2; a.w3 = w3; a.b3 = b3; a.W1t = W1t; a.W2t = W2t;
    a.m1 = m1; a.ldm1 = ldm1; a.m2 = m2; a.ldm2 = ldm2;
    a.y = y; a.gscale = gscale; a.loss_scale = loss_scale;
    a.a1 = a1; a.dz2 = dz2; a.dz1 = dz1; a.dz0 = dz0;
    a.logit = logit; a.pred = pred; a.loss_n = loss_n; a.dlogit = dlogit; a.loss_sum = loss_sum;
    a.gw3 = gw3; a.gb3 = gb3; a.sgd_out = sgd; a.lr = lr; a.decay_out = decay_out;
    a.partial = w.partial; a.ticket = w.ticket;
    const size_t smem = mlp2_smem(D0, H1, H2);
    static bool attr = false;
    if (!attr) { cudaFuncSetAttribute(mlp2_train_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); attr = true; }
    launch_k(mlp2_train_kernel, ceil_div(N, kM2Rows), kM2Threads, smem, as_stream(stream), a);
    return check_launch("mlp2_train_kernel");
}

extern "C" int dctr_mlp2_wgrad(int N, int D0, int H1, int H2, const float* x, int ld0, const float* a1, int ld1,
                               const float* dz1, const float* dz2, int ldd, float* W1, int ldw1, float* b1, float* W2, int ldw2,
                               float* b2, float* gW1, float* gb1, float* gW2, float* gb2, int sgd, float lr, float decay,
                               void* workspace, size_t workspace_bytes, void* stream) {
    DCTR_REQUIRE(N >= 0 && dctr_mlp2_supported(D0, H1, H2), "mlp2_wgrad: unsupported shape");
    if (N == 0) return DCTR_OK;
    DCTR_REQUIRE(x && a1 && dz1 && dz2 && workspace && (sgd ? (W1 && b1 && W2 && b2) : (gW1 && gb1 && gW2 && gb2)), "mlp2_wgrad: null pointer");
    DCTR_REQUIRE((ld0 & 3) == 0 && (ld1 & 3) == 0 && (ldd & 3) == 0 && (((uintptr_t)x | (uintptr_t)a1 | (uintptr_t)dz1 | (uintptr_t)dz2 | (uintptr_t)workspace) & 15) == 0,
                 "mlp2_wgrad: buffers must be 16-byte aligned, leading dimensions multiples of 4");
    (void)workspace_bytes;
    WgArgs w{};
    w.L[0] = WgLayer{x, ld0, dz1, ldd, N, D0, H1, W1, ldw1, b1, gW1, gb1, decay, ceil_div(H1, kWgTJ), 0};
    w.L[1] = WgLayer{a1, ld1, dz2, ldd, N, H1, H2, W2, ldw2, b2, gW2, gb2, decay, ceil_div(H2, kWgTJ), wg_tiles(D0, H1)};
    w.rows_per_split = ceil_div(ceil_div(N, kWgSplit), kWgK) * kWgK;
    w.sgd = sgd; w.lr = lr;
    const int nt = wg_tiles(D0, H1) + wg_tiles(H1, H2);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(nt * kWgSplit);
    cfg.blockDim = dim3(kWgThreads);
    cfg.dynamicSmemBytes = kWgSmem;
    cfg.stream = as_stream(stream);
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kWgSplit; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_mode() != 0 ? 2 : 1;
    cudaLaunchKernelEx(&cfg, mlp2_wgrad_kernel, w);
    return check_launch("mlp2_wgrad_kernel");
}
