// K6b: deterministic segmented fold + SGD into the touched embedding rows only.
//
// Input is the row-sorted (row, position) list from sort.cu: equal rows form a segment whose elements are in
// (sample, field) order.  The fold is FLAT over the sorted list -- the work per launch does not depend on how
// the list splits into segments (Zipf heads and the Criteo numeric rows that every sample touches cost the
// same per element as singletons):
//
//   level 0  a 4-lane group owns DCTR_FOLD_CHUNK = 4 consecutive sorted elements; lane q owns one float4 of the
//            row.  All gradient gathers (random 44-48 B reads) and the old rows of the segments that END in
//            the chunk are issued up front (8 independent 128-bit loads per lane), then the chunk is folded
//            serially in list order.  (Chunks of 8 need 105-116 registers = 2 CTAs per SM; chunks of 4 fit 64 registers =
//            4 CTAs per SM and measure 10-15 % faster on lists far larger than L2, equal at step sizes.)  A segment that starts and ends inside the chunk is written straight to
//            the table (one read-modify-write per touched row).
//   level 1  pieces that cross chunk borders are parked in shared memory as partials (P, Q) and composed in
//            group order by the group that holds the segment head (CTA tile = 64 groups = 256 elements).
//   level 2  pieces that cross tile borders leave the CTA as one carry-in and one carry-out partial per tile;
//            fold_stitch_kernel walks them in tile order and writes the row.
//
// Every composition happens in list order with a fixed association, so two runs are bit-identical and no float
// atomics are used.  Partials are per-mode:
//   SUM  (TF, FNN.py:46-49 densified gradient + SGD)      Q = sum g                    row = row*decay - lr*Q
//   SEQ  (Theano, FNN_wnzh.py:300-306, per occurrence)    w -> w*decay - lr*g composed as w -> P*w + Q
//   FM   (FM.py:55-64 differentiated)          Q = sum gn*x*s, P = sum gn*x^2  row = row*decay - lr*(Q - P*row)
#include "common.cuh"

namespace dctr {

constexpr int kC = DCTR_FOLD_CHUNK;            // sorted elements per group
constexpr int kG = 4;                          // lanes per group
constexpr int kFoldThreads = 256;
constexpr int kGroups = kFoldThreads / kG;     // 64 groups per CTA
constexpr int kTile = kGroups * kC;            // 256 sorted elements per CTA
constexpr int kSlab = 4 * kG;                  // 16 table columns per launch
static_assert(kTile == DCTR_FOLD_TILE && (kC == 4 || kC == 8), "dctr.h out of sync");

enum { kSum = 0, kSeq = 1, kFm = 2 };
enum { kHasF = 1, kFClosed = 2, kHasL = 4 };
enum { kInValid = 1, kInClosed = 2 };

struct FoldArgs {
    const int32_t* rows; const int32_t* pos; int64_t M;
    int F; uint64_t fmul;          // n = p / F  ==  (p * fmul) >> 40 when fmul != 0 (F < 512, p < 2^31)
    int d, ld, col0;
    int X;                         // rows >= X (the sort's sentinel for out-of-range ids) are folded but never read or written
    int skip_row;                  // segments of this row are folded but never written (-1: none) -- the pad row of absent fields
    float lr, decay;
    // decay as 1 - eps_d with eps_d and log(decay) rounded from DOUBLE: the Theano decay 1 - 2*lambda*lr/N (FNN_wnzh.py:305) is
    // 1 - 4.9e-8 at the headline configuration, which float32 can only represent as 1 - 5.96e-8 -- applied once per occurrence
    // that is a 2e-3 relative error on a row 70 000 samples hit (tests/test_baseline_sizes_gpu.py)
    float eps_d, log_decay;
    float* table;                  // rows are READ from here ...
    float* dst;                    // ... and the folded result is written here (== table except for the gradient-only fold)
    // gradient source: dense g[n*ld_n + f*ld_f + c] (* wts[p]); FM: g[n], s[n*ld_f + c], x = wts[p]
    const float* g; const float* s; int64_t ld_n; int ld_f; const float* wts;
    // per-tile carries
    float* in_q; float* out_q; float* in_p; float* out_p; int32_t* in_flag; int32_t* out_row;
};

// P of a partial: SUM unused, FM sum of g x^2, SEQ the NUMBER of occurrences m folded so far (the map is w -> decay^m w + Q;
// decay^m = exp(m log decay) is formed where it is applied, a running float product of decays would lose the increment)
template <int MODE> __device__ __forceinline__ float p_identity() { return 0.f; }
__device__ __forceinline__ float seq_pow(float m, float log_decay) { return log_decay == 0.f ? 1.f : expf(m * log_decay); }
// x * decay as x - eps_d * x: one rounding of the exact product
__device__ __forceinline__ float mul_decay(float x, float eps_d) { return fmaf(-eps_d, x, x); }

template <int MODE>
__device__ __forceinline__ void fold_elem(float& P, float4& Q, const float4& e, float v, float lr, float eps_d) {
    if (MODE == kSeq) {
        P += 1.f;
        Q.x = fmaf(-lr, e.x, mul_decay(Q.x, eps_d)); Q.y = fmaf(-lr, e.y, mul_decay(Q.y, eps_d));
        Q.z = fmaf(-lr, e.z, mul_decay(Q.z, eps_d)); Q.w = fmaf(-lr, e.w, mul_decay(Q.w, eps_d));
    } else {
        if (MODE == kFm) P += v;
        Q.x += e.x; Q.y += e.y; Q.z += e.z; Q.w += e.w;
    }
}
// (P, Q) followed in list order by (P2, Q2)
template <int MODE>
__device__ __forceinline__ void fold_combine(float& P, float4& Q, float P2, const float4& Q2, float log_decay) {
    if (MODE == kSeq) {
        const float f = seq_pow(P2, log_decay);
        Q.x = fmaf(Q.x, f, Q2.x); Q.y = fmaf(Q.y, f, Q2.y); Q.z = fmaf(Q.z, f, Q2.z); Q.w = fmaf(Q.w, f, Q2.w);
        P += P2;
    } else {
        if (MODE == kFm) P += P2;
        Q.x += Q2.x; Q.y += Q2.y; Q.z += Q2.z; Q.w += Q2.w;
    }
}
template <int MODE>
__device__ __forceinline__ float4 fold_final(const float4& old, float P, const float4& Q, int c0, float lr, float eps_d, float log_decay) {
    float4 r;
    if (MODE == kSeq) {
        const float f = seq_pow(P, log_decay);
        r.x = fmaf(f, old.x, Q.x); r.y = fmaf(f, old.y, Q.y); r.z = fmaf(f, old.z, Q.z); r.w = fmaf(f, old.w, Q.w);
    } else if (MODE == kFm) {
        const float gx = c0 == 0 ? Q.x : fmaf(-P, old.x, Q.x);     // component 0 is the linear weight: grad = gn*x
        r.x = fmaf(-lr, gx, mul_decay(old.x, eps_d));
        r.y = fmaf(-lr, fmaf(-P, old.y, Q.y), mul_decay(old.y, eps_d));
        r.z = fmaf(-lr, fmaf(-P, old.z, Q.z), mul_decay(old.z, eps_d));
        r.w = fmaf(-lr, fmaf(-P, old.w, Q.w), mul_decay(old.w, eps_d));
    } else {
        r.x = fmaf(-lr, Q.x, mul_decay(old.x, eps_d)); r.y = fmaf(-lr, Q.y, mul_decay(old.y, eps_d));
        r.z = fmaf(-lr, Q.z, mul_decay(old.z, eps_d)); r.w = fmaf(-lr, Q.w, mul_decay(old.w, eps_d));
    }
    return r;
}

__device__ __forceinline__ float4 mask_cols(float4 v, int c0, int d) {
    if (c0 + 3 >= d) {
        if (c0 >= d) v.x = 0.f;
        if (c0 + 1 >= d) v.y = 0.f;
        if (c0 + 2 >= d) v.z = 0.f;
        v.w = 0.f;
    }
    return v;
}

template <int MODE, bool ALIGNED>
__global__ void __launch_bounds__(kFoldThreads, kC == 8 ? 2 : 4)
fold_tile_kernel(const FoldArgs a) {
    pdl_sync();
    __shared__ float4 sQF[kFoldThreads], sQL[kFoldThreads];
    __shared__ float sPF[kFoldThreads], sPL[kFoldThreads];
    __shared__ int sFlag[kGroups];
    const int tid = threadIdx.x, q = tid & (kG - 1), grp = tid >> 2;
    const int c0 = a.col0 + 4 * q;
    const bool lane_on = c0 < a.ld;
    const int64_t tile = blockIdx.x;
    const int64_t k0 = (tile * kGroups + grp) * kC;

    // ---- the chunk's (row, position) pairs and its two neighbours ----
    int r[kC], p[kC];
    if (k0 + kC <= a.M) {
#pragma unroll
        for (int h = 0; h < kC / 4; ++h) {
            const int4 rr = __ldg(reinterpret_cast<const int4*>(a.rows + k0) + h), pp = __ldg(reinterpret_cast<const int4*>(a.pos + k0) + h);
            r[4 * h] = rr.x; r[4 * h + 1] = rr.y; r[4 * h + 2] = rr.z; r[4 * h + 3] = rr.w;
            p[4 * h] = pp.x; p[4 * h + 1] = pp.y; p[4 * h + 2] = pp.z; p[4 * h + 3] = pp.w;
        }
    } else {
#pragma unroll
        for (int j = 0; j < kC; ++j) {
            const bool ok = k0 + j < a.M;
            r[j] = ok ? __ldg(a.rows + k0 + j) : -1;
            p[j] = ok ? __ldg(a.pos + k0 + j) : 0;
        }
    }
#pragma unroll
    for (int j = 0; j < kC; ++j) if (r[j] >= a.X) r[j] = -1;       // sentinel tail of the sorted list == past the end
    int prev = (k0 > 0 && k0 <= a.M) ? __ldg(a.rows + k0 - 1) : -1;
    int next = (k0 + kC < a.M) ? __ldg(a.rows + k0 + kC) : -1;
    if (prev >= a.X) prev = -1;
    if (next >= a.X) next = -1;
    unsigned head = 0, tail = 0, valid = 0;
#pragma unroll
    for (int j = 0; j < kC; ++j) {
        const bool ok = r[j] >= 0;
        valid |= (unsigned)ok << j;
        head |= (unsigned)(ok && r[j] != (j ? r[j - 1] : prev)) << j;
        tail |= (unsigned)(ok && r[j] != (j < kC - 1 ? r[j + 1] : next)) << j;
    }

    // ---- issue every gather of the chunk before folding: 8 gradient reads + the old rows of the closing segments ----
    // (issued in reverse list order with "issue" loads, so the fold -- which starts with element 0, the last one
    // issued -- cannot begin before all 16 gathers are in flight; see common.cuh)
    float4 e[kC], old[kC];
    float v[kC], w[kC], gn[kC];
#pragma unroll
    for (int j = kC - 1; j >= 0; --j)
        old[j] = ld_f4_issue(a.table + (int64_t)(r[j] < 0 ? 0 : r[j]) * a.ld + c0,
                             lane_on && ((tail >> j & 1) || (j == kC - 1 && (valid >> j & 1))));
#pragma unroll
    for (int j = kC - 1; j >= 0; --j) {
        const bool on = lane_on && (valid >> j & 1);
        const uint32_t pj = (uint32_t)p[j];
        const uint32_t n = a.fmul ? (uint32_t)(((uint64_t)pj * a.fmul) >> 40) : pj / (uint32_t)a.F;
        w[j] = a.wts ? ldg_f_issue(a.wts + pj, on) : 1.f;
        if (MODE == kFm) {
            gn[j] = ldg_f_issue(a.g + n, on);
            e[j] = ldg_f4_issue(a.s + (int64_t)n * a.ld_f + c0, on);
        } else {
            const uint32_t f = pj - n * (uint32_t)a.F;
            const float* src = a.g + (int64_t)n * a.ld_n + (int64_t)f * a.ld_f + c0;
            if (ALIGNED) {
                e[j] = ldg_f4_issue(src, on);
            } else {
                e[j].w = ldg_f_issue(src + 3, on && c0 + 3 < a.d);
                e[j].z = ldg_f_issue(src + 2, on && c0 + 2 < a.d);
                e[j].y = ldg_f_issue(src + 1, on && c0 + 1 < a.d);
                e[j].x = ldg_f_issue(src, on && c0 < a.d);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < kC; ++j) {
        if (MODE == kFm) {
            const float gx = gn[j] * w[j];
            float4 s4 = e[j];
            if (c0 == 0) s4.x = 1.f;
            e[j] = mask_cols(make_float4(gx * s4.x, gx * s4.y, gx * s4.z, gx * s4.w), c0, a.d);
            v[j] = gx * w[j];
        } else {
            const float4 g4 = mask_cols(e[j], c0, a.d);
            e[j] = make_float4(g4.x * w[j], g4.y * w[j], g4.z * w[j], g4.w * w[j]);
            v[j] = 0.f;
        }
    }

    // ---- serial fold of the chunk in list order ----
    float P = p_identity<MODE>(), PF = 0.f, PL = 0.f;
    float4 Q = make_float4(0.f, 0.f, 0.f, 0.f), QF = Q, QL = Q, oldL = Q;
    bool open_left = (valid & 1) && !(head & 1);
    int flag = 0, rowL = -1;
#pragma unroll
    for (int j = 0; j < kC; ++j) {
        if (valid >> j & 1) {
            if (head >> j & 1) { P = p_identity<MODE>(); Q = make_float4(0.f, 0.f, 0.f, 0.f); open_left = false; }
            fold_elem<MODE>(P, Q, e[j], v[j], a.lr, a.eps_d);
            if (tail >> j & 1) {
                if (!open_left) {
                    if (lane_on && r[j] != a.skip_row)
                        *reinterpret_cast<float4*>(a.dst + (int64_t)r[j] * a.ld + c0) = fold_final<MODE>(old[j], P, Q, c0, a.lr, a.eps_d, a.log_decay);
                } else {
                    PF = P; QF = Q; flag |= kHasF | kFClosed;
                }
            } else if (j == kC - 1) {
                if (open_left) { PF = P; QF = Q; flag |= kHasF; }
                else { PL = P; QL = Q; flag |= kHasL; oldL = old[j]; rowL = r[j]; }
            }
        }
    }
    sPF[tid] = PF; sQF[tid] = QF; sPL[tid] = PL; sQL[tid] = QL;
    if (q == 0) sFlag[grp] = flag;
    const int tile_has_head = __syncthreads_or(head != 0);

    // ---- level 1: compose the pieces that cross chunk borders, in group order ----
    if (flag & kHasL) {
        P = PL; Q = QL;
        bool closed = false;
        for (int h = grp + 1; h < kGroups; ++h) {
            const int fh = sFlag[h];
            fold_combine<MODE>(P, Q, sPF[h * kG + q], sQF[h * kG + q], a.log_decay);
            if (fh & kFClosed) { closed = true; break; }
        }
        if (closed) {
            if (lane_on && rowL != a.skip_row)
                *reinterpret_cast<float4*>(a.dst + (int64_t)rowL * a.ld + c0) = fold_final<MODE>(oldL, P, Q, c0, a.lr, a.eps_d, a.log_decay);
        } else {      // the segment runs past this tile: carry-out
            reinterpret_cast<float4*>(a.out_q)[tile * kG + q] = Q;
            if (q == 0) { a.out_p[tile] = P; a.out_row[tile] = rowL; }
        }
    }
    if (grp == 0) {   // carry-in: the tile's first piece if it continues a segment of an earlier tile
        int fl = 0;
        if (flag & kHasF) {
            P = PF; Q = QF;
            bool closed = flag & kFClosed;
            for (int h = 1; h < kGroups && !closed; ++h) {
                const int fh = sFlag[h];
                if (!(fh & kHasF)) break;     // only past the end of the list
                fold_combine<MODE>(P, Q, sPF[h * kG + q], sQF[h * kG + q], a.log_decay);
                closed = fh & kFClosed;
            }
            reinterpret_cast<float4*>(a.in_q)[tile * kG + q] = Q;
            if (q == 0) a.in_p[tile] = P;
            fl = kInValid | (closed ? kInClosed : 0);
        }
        if (q == 0) a.in_flag[tile] = fl;
    }
    if (grp == kGroups - 1 && q == 0) {
        // a carry-out exists iff the tile's last element does not close its segment and the segment's head is in this tile
        const bool right_open = (flag & kHasL) || ((flag & kHasF) && !(flag & kFClosed));
        if (!(right_open && tile_has_head)) a.out_row[tile] = -1;
    }
}

// level 2: one WARP per tile with a carry-out; walks the carry-ins of the following tiles in tile order.  A hot row (a small
// field's value that 10 % of the samples share) spans hundreds of tiles and its walk is a chain of dependent round trips, so
// the warp fetches 64 carry-ins per trip: lane = 4 * i + q loads the 8 consecutive tiles u + 8 i + j for float4 column q,
// composes them locally, and the 8 group results are then composed in order through shuffles (the 8 lanes of a column keep
// identical accumulators).  The association is fixed by the tile index alone, so results are reproducible run to run.
// (The first version walked 4 tiles per trip with one 4-lane group: 286 us of the fold's 0.78 ms at 16.7 M slots.)
constexpr int kSt = 8;           // consecutive carry-ins per lane and trip

template <int MODE>
__global__ void __launch_bounds__(kFoldThreads)
fold_stitch_kernel(const FoldArgs a, int tiles) {
    pdl_sync();
    const int64_t t = ((int64_t)blockIdx.x * kFoldThreads + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31, q = lane & (kG - 1), i0 = lane >> 2;
    if (t >= tiles) return;
    const int row = a.out_row[t];
    if (row < 0) return;                               // warp-uniform
    const int c0 = a.col0 + 4 * q;
    const bool lane_on = c0 < a.ld;
    float4 old = make_float4(0.f, 0.f, 0.f, 0.f);
    if (lane_on) old = *reinterpret_cast<const float4*>(a.table + (int64_t)row * a.ld + c0);
    float P = a.out_p[t];
    float4 Q = reinterpret_cast<const float4*>(a.out_q)[t * kG + q];
    int64_t u = t + 1;
    bool closed = false;
    while (!closed && u < tiles) {
        // lane (i0, q) fetches the kSt CONSECUTIVE carry-ins u + kSt * i0 + j together and composes them locally in order ...
        int fl[kSt]; float pp[kSt]; float4 qq[kSt];
#pragma unroll
        for (int j = 0; j < kSt; ++j) {
            const int64_t w = u + kSt * i0 + j < tiles ? u + kSt * i0 + j : tiles - 1;
            fl[j] = a.in_flag[w]; pp[j] = a.in_p[w]; qq[j] = reinterpret_cast<const float4*>(a.in_q)[w * kG + q];
        }
        float gP = p_identity<MODE>();
        float4 gQ = make_float4(0.f, 0.f, 0.f, 0.f);
        int gclosed = 0;
#pragma unroll
        for (int j = 0; j < kSt; ++j) {
            if (!gclosed && u + kSt * i0 + j < tiles) {
                fold_combine<MODE>(gP, gQ, pp[j], qq[j], a.log_decay);
                gclosed = fl[j] & kInClosed;
            }
        }
        // ... then the 8 group results are composed in order through shuffles (a group past the end of the list is the identity)
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int src = 4 * i + q;
            const int c = __shfl_sync(0xffffffffu, gclosed, src);
            const float p2 = __shfl_sync(0xffffffffu, gP, src);
            float4 q2;
            q2.x = __shfl_sync(0xffffffffu, gQ.x, src); q2.y = __shfl_sync(0xffffffffu, gQ.y, src);
            q2.z = __shfl_sync(0xffffffffu, gQ.z, src); q2.w = __shfl_sync(0xffffffffu, gQ.w, src);
            if (!closed) {
                fold_combine<MODE>(P, Q, p2, q2, a.log_decay);
                closed = c != 0;
            }
        }
        u += 8 * kSt;
    }
    if (i0 == 0 && lane_on && row != a.skip_row)
        *reinterpret_cast<float4*>(a.dst + (int64_t)row * a.ld + c0) = fold_final<MODE>(old, P, Q, c0, a.lr, a.eps_d, a.log_decay);
}

__global__ void decay_sweep_kernel(float4* __restrict__ table, int64_t X, int q, float decay,
                                   const int32_t* __restrict__ rows, int64_t M) {
    pdl_sync();
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

static int64_t num_tiles(int64_t M) { return (M + kTile - 1) / kTile; }

template <int MODE>
static int launch_fold(FoldArgs a, bool aligned, cudaStream_t st) {
    const int tiles = (int)num_tiles(a.M);
    for (int col0 = 0; col0 < a.ld; col0 += kSlab) {
        a.col0 = col0;
        if (aligned) launch_k(fold_tile_kernel<MODE, true>, tiles, kFoldThreads, 0, st, a);
        else launch_k(fold_tile_kernel<MODE, false>, tiles, kFoldThreads, 0, st, a);
        int rc = check_launch("fold_tile_kernel");
        if (rc) return rc;
        if (tiles > 1) {
            launch_k(fold_stitch_kernel<MODE>, ceil_div((int64_t)tiles * 32, kFoldThreads), kFoldThreads, 0, st, a, tiles);
            if ((rc = check_launch("fold_stitch_kernel"))) return rc;
        }
    }
    return DCTR_OK;
}

static int run_fold(int mode, FoldArgs a, bool aligned, void* workspace, size_t workspace_bytes, cudaStream_t st) {
    if (workspace_bytes < dctr_scatter_workspace(a.M, a.ld)) {
        set_error("scatter_sgd: workspace %zu < %zu bytes", workspace_bytes, dctr_scatter_workspace(a.M, a.ld));
        return DCTR_EWORKSPACE;
    }
    const int64_t tiles = num_tiles(a.M);
    float* w = static_cast<float*>(workspace);
    a.in_q = w;
    a.out_q = w + tiles * kSlab;
    a.in_p = w + 2 * tiles * kSlab;
    a.out_p = a.in_p + tiles;
    a.in_flag = reinterpret_cast<int32_t*>(a.out_p + tiles);
    a.out_row = a.in_flag + tiles;
    a.fmul = (a.F < 512) ? (((uint64_t)1 << 40) / (uint64_t)a.F + 1) : 0;
    if (mode == kSeq) return launch_fold<kSeq>(a, aligned, st);
    if (mode == kFm) return launch_fold<kFm>(a, aligned, st);
    return launch_fold<kSum>(a, aligned, st);
}

}  // namespace dctr

using namespace dctr;

extern "C" size_t dctr_scatter_workspace(int64_t M, int ld) {
    (void)ld;
    if (M < 0) M = 0;
    const int64_t tiles = num_tiles(M) + 1;
    return (size_t)tiles * (2 * kSlab + 4) * sizeof(float) + 64;
}

static bool aligned16(const void* p) { return ((uintptr_t)p & 15) == 0; }
#include <cmath>
static void set_decay(dctr::FoldArgs& a, double decay) {
    a.decay = (float)decay;
    a.eps_d = (float)(1.0 - decay);
    a.log_decay = decay > 0.0 ? (float)std::log(decay) : 0.f;
}

static int scatter_sgd_impl(int mode, const int32_t* sorted_rows, const int32_t* sorted_pos, int64_t M, int F,
                            int d, const float* g, int64_t g_ld_n, int g_ld_f, const float* wts, int32_t skip_row, float lr,
                            double decay, float* table, int64_t X, int ld, void* workspace,
                            size_t workspace_bytes, void* stream) {
    DCTR_REQUIRE(mode == DCTR_SCATTER_SUM || mode == DCTR_SCATTER_SEQ_DECAY, "scatter_sgd: bad mode %d", mode);
    DCTR_REQUIRE(M >= 0 && M < ((int64_t)1 << 31) && F > 0 && d > 0 && d <= ld && (ld & 3) == 0 && X > 0 && X < ((int64_t)1 << 31), "scatter_sgd: bad shape");
    if (M == 0) return DCTR_OK;
    DCTR_REQUIRE(sorted_rows && sorted_pos && g && table && workspace, "scatter_sgd: null pointer");
    DCTR_REQUIRE(aligned16(sorted_rows) && aligned16(sorted_pos) && aligned16(table) && aligned16(workspace),
                 "scatter_sgd: rows / pos / table / workspace must be 16-byte aligned");
    FoldArgs a{};
    a.skip_row = -1;
    a.X = (int)X; a.rows = sorted_rows; a.pos = sorted_pos; a.M = M; a.F = F; a.d = d; a.ld = ld; a.lr = lr; set_decay(a, decay); a.table = table; a.dst = table;
    a.g = g; a.s = nullptr; a.ld_n = g_ld_n; a.ld_f = g_ld_f; a.wts = wts; a.skip_row = skip_row;
    // 128-bit gradient loads need whole float4s inside the source row: per-field blocks of >= ld floats, or -- g_ld_f = 0, the
    // embedding bag: every field of a sample shares the sample's row -- rows of >= ld floats
    const bool aligned = aligned16(g) && (g_ld_n & 3) == 0 && (g_ld_f & 3) == 0 && (g_ld_f >= ld || (g_ld_f == 0 && g_ld_n >= ld));
    return run_fold(mode == DCTR_SCATTER_SEQ_DECAY ? kSeq : kSum, a, aligned, workspace, workspace_bytes, as_stream(stream));
}

extern "C" int dctr_scatter_sgd(int mode, const int32_t* sorted_rows, const int32_t* sorted_pos, int64_t M, int F,
                                int d, const float* g, int64_t g_ld_n, int g_ld_f, const float* wts, float lr,
                                double decay, float* table, int64_t X, int ld, void* workspace,
                                size_t workspace_bytes, void* stream) {
    return scatter_sgd_impl(mode, sorted_rows, sorted_pos, M, F, d, g, g_ld_n, g_ld_f, wts, -1, lr, decay, table, X, ld, workspace,
                            workspace_bytes, stream);
}

extern "C" int dctr_scatter_sgd_masked(int mode, const int32_t* sorted_rows, const int32_t* sorted_pos, int64_t M, int F,
                                       int d, const float* g, int64_t g_ld_n, int g_ld_f, int32_t skip_row, float lr,
                                       double decay, float* table, int64_t X, int ld, void* workspace,
                                       size_t workspace_bytes, void* stream) {
    return scatter_sgd_impl(mode, sorted_rows, sorted_pos, M, F, d, g, g_ld_n, g_ld_f, nullptr, skip_row, lr, decay, table, X, ld,
                            workspace, workspace_bytes, stream);
}

extern "C" int dctr_scatter_sgd_fm(const int32_t* sorted_rows, const int32_t* sorted_pos, int64_t M, int F, int d,
                                   const float* g, const float* s, int lds, const float* wts, float lr, double decay,
                                   float* table, int64_t X, int ld, void* workspace, size_t workspace_bytes,
                                   void* stream) {
    DCTR_REQUIRE(M >= 0 && M < ((int64_t)1 << 31) && F > 0 && d > 0 && d <= ld && (ld & 3) == 0 && X > 0 && lds >= ld && (lds & 3) == 0,
                 "scatter_sgd_fm: bad shape (lds must be a multiple of 4 and >= ld)");
    if (M == 0) return DCTR_OK;
    DCTR_REQUIRE(sorted_rows && sorted_pos && g && s && table && workspace, "scatter_sgd_fm: null pointer");
    DCTR_REQUIRE(aligned16(sorted_rows) && aligned16(sorted_pos) && aligned16(table) && aligned16(workspace) && aligned16(s),
                 "scatter_sgd_fm: rows / pos / table / s / workspace must be 16-byte aligned");
    FoldArgs a{};
    a.skip_row = -1;
    a.X = (int)X; a.rows = sorted_rows; a.pos = sorted_pos; a.M = M; a.F = F; a.d = d; a.ld = ld; a.lr = lr; set_decay(a, decay); a.table = table; a.dst = table;
    a.g = g; a.s = s; a.ld_n = 0; a.ld_f = lds; a.wts = wts;
    return run_fold(kFm, a, true, workspace, workspace_bytes, as_stream(stream));
}

extern "C" int dctr_scatter_grad_fm(const int32_t* sorted_rows, const int32_t* sorted_pos, int64_t M, int F, int d,
                                    const float* g, const float* s, int lds, const float* wts, const float* table, int64_t X,
                                    int ld, float* grad_out, void* workspace, size_t workspace_bytes, void* stream) {
    DCTR_REQUIRE(M >= 0 && M < ((int64_t)1 << 31) && F > 0 && d > 0 && d <= ld && (ld & 3) == 0 && X > 0 && lds >= ld && (lds & 3) == 0,
                 "scatter_grad_fm: bad shape (lds must be a multiple of 4 and >= ld)");
    if (M == 0) return DCTR_OK;
    DCTR_REQUIRE(sorted_rows && sorted_pos && g && s && table && grad_out && workspace, "scatter_grad_fm: null pointer");
    DCTR_REQUIRE(aligned16(sorted_rows) && aligned16(sorted_pos) && aligned16(table) && aligned16(grad_out) && aligned16(workspace) && aligned16(s),
                 "scatter_grad_fm: rows / pos / table / grad_out / s / workspace must be 16-byte aligned");
    FoldArgs a{};
    a.skip_row = -1;
    // the FM fold ends in  old*decay - lr*(Q - P*old):  decay = 0, lr = -1 leaves exactly the row's gradient
    a.X = (int)X; a.rows = sorted_rows; a.pos = sorted_pos; a.M = M; a.F = F; a.d = d; a.ld = ld; a.lr = -1.f; set_decay(a, 0.0);
    a.table = const_cast<float*>(table); a.dst = grad_out;
    a.g = g; a.s = s; a.ld_n = 0; a.ld_f = lds; a.wts = wts;
    return run_fold(kFm, a, true, workspace, workspace_bytes, as_stream(stream));
}

extern "C" int dctr_decay_sweep(float* table, int64_t X, int ld, float decay, const int32_t* sorted_rows, int64_t M,
                                void* stream) {
    DCTR_REQUIRE(table && X > 0 && (ld & 3) == 0 && M >= 0 && (M == 0 || sorted_rows), "decay_sweep: bad arguments");
    const int64_t total = X * (ld / 4);
    int blocks = ceil_div(total, 256);
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    launch_k(decay_sweep_kernel, blocks, 256, 0, as_stream(stream), reinterpret_cast<float4*>(table), X, ld / 4, decay, sorted_rows, M);
    return check_launch("decay_sweep_kernel");
}
