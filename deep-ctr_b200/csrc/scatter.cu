// K6b: deterministic segmented fold + SGD into the touched embedding rows only.
// Input is the row-sorted (row, position) list from sort.cu.  Equal rows form a segment whose elements
// are in (sample, field) order.  A 16-lane group owns one sorted index; the group at a segment head folds
// the segment, lane c owning component c, so a row is read and written exactly once and the summation
// order is fixed.  Segments longer than DCTR_LONG_SEG (hot rows: Zipf heads, the 13 Criteo numeric rows
// that every sample touches) are cut into chunks of DCTR_SEG_CHUNK, folded chunk-parallel into affine
// partials (A, B) and then composed in chunk order -- still deterministic.
//   SUM (TF semantics, FNN.py:46-49):          row = row*decay - lr * sum_k grad_k
//   SEQ_DECAY (Theano, FNN_wnzh.py:300-306):   for k in order: row = row*decay - lr*grad_k
#include "common.cuh"

namespace dctr {

constexpr int kGroup = 16;          // lanes per sorted element
constexpr int kFoldThreads = 256;   // 16 groups per CTA

struct DenseSrc {
    const float* g; int64_t ld_n; int ld_f; const float* wts;
    __device__ __forceinline__ float operator()(int p, int n, int f, int c, float) const {
        const float v = g[n * ld_n + (int64_t)f * ld_f + c];
        return wts ? wts[p] * v : v;
    }
};
struct FmSrc {
    const float* g; const float* s; int lds; const float* wts;
    __device__ __forceinline__ float operator()(int p, int n, int, int c, float old) const {
        const float x = wts ? wts[p] : 1.f;
        const float gn = g[n];
        return c == 0 ? gn * x : gn * (x * s[(int64_t)n * lds + c] - x * x * old);
    }
};

// header of the workspace: [0] = number of long segments, [1] = number of chunk partials reserved
struct LongSeg { int32_t start, len, chunk_off, pad; };
// chunk length grows with the segment so that a segment never has more than 256 partials to compose
__device__ __forceinline__ int chunk_len_of(int len) {
    const int c = (len + 255) >> 8;
    return c < DCTR_SEG_CHUNK ? DCTR_SEG_CHUNK : c;
}

__device__ __forceinline__ int64_t segment_end(const int32_t* __restrict__ rows, int64_t i, int64_t M, int32_t row) {
    // gallop then bisect: first index > i whose row differs
    int64_t step = 1;
    while (i + step < M && rows[i + step] == row) step <<= 1;
    int64_t lo = i + (step >> 1), hi = (i + step < M) ? i + step : M;   // rows[lo]==row, rows[hi]!=row or hi==M
    while (hi - lo > 1) {
        const int64_t mid = (lo + hi) >> 1;
        if (rows[mid] == row) lo = mid; else hi = mid;
    }
    return hi;
}

template <bool SEQ, class Src>
__global__ void __launch_bounds__(kFoldThreads)
fold_short_kernel(Src src, const int32_t* __restrict__ rows, const int32_t* __restrict__ pos, int64_t M, int F, int d,
                  float lr, float decay, float* __restrict__ table, int ld, int32_t* __restrict__ hdr, LongSeg* __restrict__ longs,
                  int32_t* __restrict__ chunk_seg) {
    const int64_t i = ((int64_t)blockIdx.x * kFoldThreads + threadIdx.x) / kGroup;
    const int gl = threadIdx.x & (kGroup - 1);
    if (i >= M) return;
    const int32_t row = rows[i];
    if (i > 0 && rows[i - 1] == row) return;   // not a segment head
    const int64_t end = segment_end(rows, i, M, row);
    const int len = (int)(end - i);
    if (len > DCTR_LONG_SEG) {
        const int nchunks = (len + chunk_len_of(len) - 1) / chunk_len_of(len);
        int slot = 0, coff = 0;
        if (gl == 0) {
            slot = atomicAdd(&hdr[0], 1);
            coff = atomicAdd(&hdr[1], nchunks);
            LongSeg ls; ls.start = (int32_t)i; ls.len = len; ls.chunk_off = coff; ls.pad = 0;
            longs[slot] = ls;   // slot order is arbitrary; every segment is folded independently
        }
        const unsigned gmask = 0xffffu << (threadIdx.x & 16);
        slot = __shfl_sync(gmask, slot, 0, kGroup);
        coff = __shfl_sync(gmask, coff, 0, kGroup);
        for (int j = gl; j < nchunks; j += kGroup) chunk_seg[coff + j] = slot;   // chunk -> owning long segment
        return;
    }
    float* trow = table + (int64_t)row * ld;
    for (int c = gl; c < d; c += kGroup) {
        const float old = trow[c];
        float w = SEQ ? old : 0.f;
#pragma unroll 4
        for (int64_t k = i; k < end; ++k) {
            const int p = pos[k];
            const int n = p / F;
            const float gk = src(p, n, p - n * F, c, old);
            w = SEQ ? fmaf(w, decay, -lr * gk) : w + gk;
        }
        trow[c] = SEQ ? w : fmaf(old, decay, -lr * w);
    }
}

// one 16-lane group per chunk of a long segment: partial[chunk][c] = B, partial[chunk][ld + c] = A
template <bool SEQ, class Src>
__global__ void __launch_bounds__(kFoldThreads)
fold_chunk_kernel(Src src, const int32_t* __restrict__ pos, int F, int d, float lr, float decay,
                  const float* __restrict__ table, const int32_t* __restrict__ rows, int ld,
                  const int32_t* __restrict__ hdr, const LongSeg* __restrict__ longs,
                  const int32_t* __restrict__ chunk_seg, float* __restrict__ partial) {
    const int total_chunks = hdr[1];
    const int gl = threadIdx.x & (kGroup - 1);
    const int groups_total = gridDim.x * (kFoldThreads / kGroup);
    for (int g = blockIdx.x * (kFoldThreads / kGroup) + threadIdx.x / kGroup; g < total_chunks; g += groups_total) {
        const LongSeg ls = longs[chunk_seg[g]];
        const int j = g - ls.chunk_off;
        const float* trow = table + (int64_t)rows[ls.start] * ld;
        const int cl = chunk_len_of(ls.len);
        const int64_t k0 = (int64_t)ls.start + (int64_t)j * cl;
        const int64_t k1 = (k0 + cl < (int64_t)ls.start + ls.len) ? k0 + cl : (int64_t)ls.start + ls.len;
        float* out = partial + (int64_t)g * 2 * ld;
        for (int c = gl; c < d; c += kGroup) {
            const float old = trow[c];
            float a = 1.f, b = 0.f;
#pragma unroll 4
            for (int64_t k = k0; k < k1; ++k) {
                const int p = pos[k];
                const int n = p / F;
                const float gk = src(p, n, p - n * F, c, old);
                if (SEQ) { b = fmaf(b, decay, -lr * gk); a *= decay; } else { b += gk; }
            }
            out[c] = b;
            out[ld + c] = a;
        }
    }
}

// one group per long segment: compose the chunk partials in chunk order and write the row
template <bool SEQ>
__global__ void __launch_bounds__(kFoldThreads)
fold_combine_kernel(int d, float lr, float decay, float* __restrict__ table, const int32_t* __restrict__ rows, int ld,
                    const int32_t* __restrict__ hdr, const LongSeg* __restrict__ longs, const float* __restrict__ partial) {
    const int nlong = hdr[0];
    const int gl = threadIdx.x & (kGroup - 1);
    const int groups_total = gridDim.x * (kFoldThreads / kGroup);
    for (int sidx = blockIdx.x * (kFoldThreads / kGroup) + threadIdx.x / kGroup; sidx < nlong; sidx += groups_total) {
        const LongSeg ls = longs[sidx];
        const int nchunks = (ls.len + chunk_len_of(ls.len) - 1) / chunk_len_of(ls.len);
        float* trow = table + (int64_t)rows[ls.start] * ld;
        for (int c = gl; c < d; c += kGroup) {
            const float old = trow[c];
            float w = SEQ ? old : 0.f;
#pragma unroll 8
            for (int j = 0; j < nchunks; ++j) {
                const float* pj = partial + (int64_t)(ls.chunk_off + j) * 2 * ld;
                w = SEQ ? fmaf(w, pj[ld + c], pj[c]) : w + pj[c];
            }
            trow[c] = SEQ ? w : fmaf(old, decay, -lr * w);
        }
    }
}

__global__ void decay_sweep_kernel(float4* __restrict__ table, int64_t X, int q, float decay,
                                   const int32_t* __restrict__ rows, int64_t M) {
    const int64_t total = X * q;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int32_t r = (int32_t)(t / q);
        // skip rows the batch touched (their decay was applied by the fold): lower_bound in the sorted list
        int64_t lo = 0, hi = M;
        while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (rows[mid] < r) lo = mid + 1; else hi = mid; }
        if (lo < M && rows[lo] == r) continue;
        float4 v = table[t];
        v.x *= decay; v.y *= decay; v.z *= decay; v.w *= decay;
        table[t] = v;
    }
}

static size_t longs_capacity(int64_t M) { return (size_t)(M / DCTR_LONG_SEG + 2); }
static size_t chunks_capacity(int64_t M) { return (size_t)(M / DCTR_SEG_CHUNK + M / DCTR_LONG_SEG + 4); }

template <class Src>
static int run_fold(int mode, Src src, const int32_t* rows, const int32_t* pos, int64_t M, int F, int d, float lr,
                    float decay, float* table, int ld, void* workspace, size_t workspace_bytes, cudaStream_t st) {
    if (workspace_bytes < dctr_scatter_workspace(M, ld)) {
        set_error("scatter_sgd: workspace %zu < %zu bytes", workspace_bytes, dctr_scatter_workspace(M, ld));
        return DCTR_EWORKSPACE;
    }
    int32_t* hdr = static_cast<int32_t*>(workspace);
    LongSeg* longs = reinterpret_cast<LongSeg*>(hdr + 4);
    int32_t* chunk_seg = reinterpret_cast<int32_t*>(longs + longs_capacity(M));
    float* partial = reinterpret_cast<float*>(chunk_seg + ((chunks_capacity(M) + 3) & ~(size_t)3));
    cudaMemsetAsync(hdr, 0, 16, st);
    const int blocks = ceil_div(M * kGroup, kFoldThreads);
    const int pgrid = kNumSMs * 4;
    const bool seq = mode == DCTR_SCATTER_SEQ_DECAY;
    int rc;
    if (seq) fold_short_kernel<true, Src><<<blocks, kFoldThreads, 0, st>>>(src, rows, pos, M, F, d, lr, decay, table, ld, hdr, longs, chunk_seg);
    else fold_short_kernel<false, Src><<<blocks, kFoldThreads, 0, st>>>(src, rows, pos, M, F, d, lr, decay, table, ld, hdr, longs, chunk_seg);
    if ((rc = check_launch("fold_short_kernel"))) return rc;
    if (M > DCTR_LONG_SEG) {
        if (seq) fold_chunk_kernel<true, Src><<<pgrid, kFoldThreads, 0, st>>>(src, pos, F, d, lr, decay, table, rows, ld, hdr, longs, chunk_seg, partial);
        else fold_chunk_kernel<false, Src><<<pgrid, kFoldThreads, 0, st>>>(src, pos, F, d, lr, decay, table, rows, ld, hdr, longs, chunk_seg, partial);
        if ((rc = check_launch("fold_chunk_kernel"))) return rc;
        if (seq) fold_combine_kernel<true><<<kNumSMs, kFoldThreads, 0, st>>>(d, lr, decay, table, rows, ld, hdr, longs, partial);
        else fold_combine_kernel<false><<<kNumSMs, kFoldThreads, 0, st>>>(d, lr, decay, table, rows, ld, hdr, longs, partial);
        if ((rc = check_launch("fold_combine_kernel"))) return rc;
    }
    return DCTR_OK;
}

}  // namespace dctr

using namespace dctr;

extern "C" size_t dctr_scatter_workspace(int64_t M, int ld) {
    if (M < 0) M = 0;
    return 16 + longs_capacity(M) * sizeof(LongSeg) + ((chunks_capacity(M) + 3) & ~(size_t)3) * sizeof(int32_t) +
           chunks_capacity(M) * 2 * (size_t)ld * sizeof(float);
}

extern "C" int dctr_scatter_sgd(int mode, const int32_t* sorted_rows, const int32_t* sorted_pos, int64_t M, int F,
                                int d, const float* g, int64_t g_ld_n, int g_ld_f, const float* wts, float lr,
                                float decay, float* table, int64_t X, int ld, void* workspace,
                                size_t workspace_bytes, void* stream) {
    DCTR_REQUIRE(mode == DCTR_SCATTER_SUM || mode == DCTR_SCATTER_SEQ_DECAY, "scatter_sgd: bad mode %d", mode);
    DCTR_REQUIRE(M >= 0 && F > 0 && d > 0 && d <= ld && (ld & 3) == 0 && X > 0, "scatter_sgd: bad shape");
    if (M == 0) return DCTR_OK;
    DCTR_REQUIRE(sorted_rows && sorted_pos && g && table && workspace, "scatter_sgd: null pointer");
    DenseSrc src{g, g_ld_n, g_ld_f, wts};
    return run_fold(mode, src, sorted_rows, sorted_pos, M, F, d, lr, decay, table, ld, workspace, workspace_bytes, as_stream(stream));
}

extern "C" int dctr_scatter_sgd_fm(const int32_t* sorted_rows, const int32_t* sorted_pos, int64_t M, int F, int d,
                                   const float* g, const float* s, int lds, const float* wts, float lr, float decay,
                                   float* table, int64_t X, int ld, void* workspace, size_t workspace_bytes,
                                   void* stream) {
    DCTR_REQUIRE(M >= 0 && F > 0 && d > 0 && d <= ld && (ld & 3) == 0 && X > 0 && lds >= d, "scatter_sgd_fm: bad shape");
    if (M == 0) return DCTR_OK;
    DCTR_REQUIRE(sorted_rows && sorted_pos && g && s && table && workspace, "scatter_sgd_fm: null pointer");
    FmSrc src{g, s, lds, wts};
    return run_fold(DCTR_SCATTER_SUM, src, sorted_rows, sorted_pos, M, F, d, lr, decay, table, ld, workspace, workspace_bytes, as_stream(stream));
}

extern "C" int dctr_decay_sweep(float* table, int64_t X, int ld, float decay, const int32_t* sorted_rows, int64_t M,
                                void* stream) {
    DCTR_REQUIRE(table && X > 0 && (ld & 3) == 0 && M >= 0 && (M == 0 || sorted_rows), "decay_sweep: bad arguments");
    const int64_t total = X * (ld / 4);
    int blocks = ceil_div(total, 256);
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    decay_sweep_kernel<<<blocks, 256, 0, as_stream(stream)>>>(reinterpret_cast<float4*>(table), X, ld / 4, decay, sorted_rows, M);
    return check_launch("decay_sweep_kernel");
}
