// Eval metrics on the device (baseline.py:246-259 watch_train, FNN_wnzh.py:187-221): ROC AUC with tied
// predictions counted by average rank (what sklearn.roc_auc_score computes), mean log-loss, RMSE -- without copying the
// predictions of a whole evaluation set to the host.
//
//   AUC = ( sum over positives of avg_rank  -  n_pos (n_pos + 1) / 2 ) / ( n_pos * n_neg )
// The predictions are sorted by their bit pattern (non-negative floats order like their int32 bits) with the library's own
// radix sort; the tie group of a positive is found by two binary searches in the sorted keys; avg_rank = (lo + hi + 1) / 2
// with lo / hi the 0-based bounds of the group, so 2 * sum(avg_rank) is an INTEGER and is accumulated with 64-bit integer
// atomics: exact, hence independent of the order the atomics land in.  Log-loss and squared error are summed in double
// per CTA and the CTA partials are added in CTA order by the last CTA (ticket), like every other reduction here.
#include "common.cuh"

namespace dctr {

constexpr int kMetThreads = 256;

struct MetricArgs {
    const float* pred; const float* y; int64_t N;
    int32_t* keys; const int32_t* skeys; const int32_t* spos;
    double* partial;                 // [blocks][2]
    unsigned long long* acc;         // [0] = 2 * sum of positive average ranks, [1] = n_pos
    unsigned* ticket;
    double* out;                     // [4] = auc (-1 when one class is absent), mean log-loss, rmse, n_pos
};

__global__ void __launch_bounds__(kMetThreads) metric_keys_kernel(const MetricArgs a) {
    pdl_sync();
    for (int64_t i = (int64_t)blockIdx.x * kMetThreads + threadIdx.x; i < a.N; i += (int64_t)gridDim.x * kMetThreads) {
        const float p = fminf(fmaxf(a.pred[i], 0.f), 1.f);      // a probability: its bits are a non-negative int32, monotone in p
        a.keys[i] = __float_as_int(p);
    }
}

__global__ void __launch_bounds__(kMetThreads) metric_reduce_kernel(const MetricArgs a) {
    pdl_sync();
    __shared__ double s_ll[kMetThreads], s_se[kMetThreads];
    __shared__ unsigned s_last;
    double ll = 0.0, se = 0.0;
    unsigned long long ranks2 = 0, npos = 0;
    for (int64_t i = (int64_t)blockIdx.x * kMetThreads + threadIdx.x; i < a.N; i += (int64_t)gridDim.x * kMetThreads) {
        const int32_t k = a.skeys[i];
        const int32_t src = a.spos[i];
        const float yv = a.y[src];
        const double p = (double)a.pred[src];
        const double pc = fmin(fmax(p, 1e-15), 1.0 - 1e-15);    // sklearn.metrics.log_loss clips the same way
        ll -= yv > 0.5f ? log(pc) : log1p(-pc);
        se += (p - (double)yv) * (p - (double)yv);
        if (yv > 0.5f) {
            int64_t lo = 0, hi = i;                              // first index with key == k lies in [0, i]
            while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (a.skeys[mid] < k) lo = mid + 1; else hi = mid; }
            int64_t l2 = i + 1, h2 = a.N;                        // first index with key > k lies in [i + 1, N]
            while (l2 < h2) { const int64_t mid = (l2 + h2) >> 1; if (a.skeys[mid] <= k) l2 = mid + 1; else h2 = mid; }
            ranks2 += (unsigned long long)(lo + l2 + 1);         // 2 * average 1-based rank of the tie group [lo, l2)
            npos += 1;
        }
    }
    // integer partials: warp-reduce, then one atomic per warp
    for (int o = 16; o > 0; o >>= 1) {
        ranks2 += __shfl_down_sync(0xffffffffu, ranks2, o);
        npos += __shfl_down_sync(0xffffffffu, npos, o);
    }
    if ((threadIdx.x & 31) == 0 && npos) { atomicAdd(a.acc, ranks2); atomicAdd(a.acc + 1, npos); }
    s_ll[threadIdx.x] = ll; s_se[threadIdx.x] = se;
    __syncthreads();
    for (int o = kMetThreads / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) { s_ll[threadIdx.x] += s_ll[threadIdx.x + o]; s_se[threadIdx.x] += s_se[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { a.partial[2 * blockIdx.x] = s_ll[0]; a.partial[2 * blockIdx.x + 1] = s_se[0]; }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(a.ticket, 1u);
        s_last = (t == gridDim.x - 1) ? 1u : 0u;
    }
    __syncthreads();
    if (!s_last || threadIdx.x != 0) return;
    __threadfence();
    double L = 0.0, S = 0.0;
    for (unsigned b = 0; b < gridDim.x; ++b) { L += __ldcg(a.partial + 2 * b); S += __ldcg(a.partial + 2 * b + 1); }
    const unsigned long long r2 = __ldcg(a.acc), np = __ldcg(a.acc + 1);
    const double n_pos = (double)np, n_neg = (double)a.N - n_pos;
    a.out[0] = (np == 0 || n_neg == 0.0) ? -1.0 : (0.5 * (double)r2 - 0.5 * n_pos * (n_pos + 1.0)) / (n_pos * n_neg);
    a.out[1] = L / (double)a.N;
    a.out[2] = sqrt(S / (double)a.N);
    a.out[3] = n_pos;
}

static int met_blocks(int64_t N) {
    int b = ceil_div(N, kMetThreads * 4);
    if (b > kNumSMs * 4) b = kNumSMs * 4;
    return b < 1 ? 1 : b;
}
static size_t r16(size_t v) { return (v + 15) / 16 * 16; }

}  // namespace dctr

using namespace dctr;

extern "C" size_t dctr_eval_metrics_workspace(int64_t N) {
    if (N < 1) N = 1;
    return 3 * r16((size_t)N * 4) + r16(dctr_sort_rows_workspace(N)) + r16((size_t)met_blocks(N) * 16) + 64;
}

extern "C" int dctr_eval_metrics(const float* pred, const float* y, int64_t N, double* out, void* workspace,
                                 size_t workspace_bytes, void* stream) {
    DCTR_REQUIRE(pred && y && out && workspace && N > 0 && N < ((int64_t)1 << 30), "eval_metrics: bad arguments");
    DCTR_REQUIRE(((uintptr_t)workspace & 15) == 0 && ((uintptr_t)out & 7) == 0, "eval_metrics: workspace / out alignment");
    if (workspace_bytes < dctr_eval_metrics_workspace(N)) {
        set_error("eval_metrics: workspace %zu < %zu bytes", workspace_bytes, dctr_eval_metrics_workspace(N));
        return DCTR_EWORKSPACE;
    }
    cudaStream_t st = as_stream(stream);
    char* w = static_cast<char*>(workspace);
    MetricArgs a{};
    a.pred = pred; a.y = y; a.N = N; a.out = out;
    a.keys = reinterpret_cast<int32_t*>(w); w += r16((size_t)N * 4);
    int32_t* skeys = reinterpret_cast<int32_t*>(w); w += r16((size_t)N * 4);
    int32_t* spos = reinterpret_cast<int32_t*>(w); w += r16((size_t)N * 4);
    void* sort_ws = w; const size_t sort_bytes = dctr_sort_rows_workspace(N); w += r16(sort_bytes);
    a.partial = reinterpret_cast<double*>(w); w += r16((size_t)met_blocks(N) * 16);
    a.acc = reinterpret_cast<unsigned long long*>(w);
    a.ticket = reinterpret_cast<unsigned*>(w + 16);
    a.skeys = skeys; a.spos = spos;
    cudaMemsetAsync(a.acc, 0, 32, st);
    launch_k(metric_keys_kernel, met_blocks(N), kMetThreads, 0, st, a);
    int rc = check_launch("metric_keys_kernel");
    if (rc) return rc;
    // bit patterns of [0, 1] are < 0x3F800001 < 2^30: four 8-bit passes
    if ((rc = dctr_sort_rows(a.keys, N, (int64_t)0x3F800001, skeys, spos, sort_ws, sort_bytes, stream))) return rc;
    launch_k(metric_reduce_kernel, met_blocks(N), kMetThreads, 0, st, a);
    return check_launch("metric_reduce_kernel");
}
