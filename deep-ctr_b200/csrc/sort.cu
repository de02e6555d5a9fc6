// K6a: stable LSD radix sort of (row id, position) pairs -- the deterministic ordering step of the
// sparse update.  Hand-written (no CUB): per pass  histogram -> exclusive scan -> stable scatter.
// Stability inside a tile comes from ranking 32-element chunks with __match_any_sync and scanning the
// per-chunk digit counts in chunk order, so equal rows keep their (sample, field) order: that is what
// lets SEQ_DECAY reproduce the sample-major loop of FNN_wnzh.py:300-306 and makes SUM bit-reproducible.
#include "common.cuh"

namespace dctr {

constexpr int kSortThreads = 256;
constexpr int kSortItems = 8;                       // chunks per warp
constexpr int kSortTile = kSortThreads * kSortItems;  // 2048 elements per CTA
constexpr int kSortChunks = kSortTile / 32;         // 64
constexpr int kRadixBits = 8;
constexpr int kRadix = 1 << kRadixBits;

__device__ __forceinline__ int digit_of(int32_t key, int shift) { return (key >> shift) & (kRadix - 1); }

__global__ void __launch_bounds__(kSortThreads)
sort_hist_kernel(const int32_t* __restrict__ keys, int64_t M, int shift, int num_tiles, int32_t* __restrict__ hist) {
    __shared__ int32_t h[kRadix];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * kSortTile;
#pragma unroll
    for (int j = 0; j < kSortItems; ++j) {
        const int64_t i = base + j * kSortThreads + threadIdx.x;
        if (i < M) atomicAdd(&h[digit_of(keys[i], shift)], 1);
    }
    __syncthreads();
    hist[(int64_t)threadIdx.x * num_tiles + blockIdx.x] = h[threadIdx.x];
}

// one CTA per digit: exclusive scan of that digit's per-tile counts (in place, tile order) + the digit total.
// The scatter kernel turns the 256 totals into digit base offsets itself, so the scan is parallel over digits.
__global__ void __launch_bounds__(1024) sort_scan_kernel(int32_t* __restrict__ hist, int num_tiles, int32_t* __restrict__ totals) {
    __shared__ int32_t warp_tot[32];
    __shared__ int32_t carry_s;
    int32_t* row = hist + (int64_t)blockIdx.x * num_tiles;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < num_tiles; base += 1024) {
        const int i = base + tid;
        const int32_t v = i < num_tiles ? row[i] : 0;
        int32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int32_t u = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += u;
        }
        if (lane == 31) warp_tot[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            const int32_t w = warp_tot[lane];
            int32_t wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int32_t u = __shfl_up_sync(0xffffffffu, wi, o);
                if (lane >= o) wi += u;
            }
            warp_tot[lane] = wi - w;
        }
        __syncthreads();
        const int32_t carry = carry_s;
        if (i < num_tiles) row[i] = carry + warp_tot[warp] + incl - v;
        __syncthreads();
        if (tid == 1023) carry_s = carry + warp_tot[31] + incl;
        __syncthreads();
    }
    if (tid == 0) totals[blockIdx.x] = carry_s;
}

// pos_in == nullptr: payload is the element index itself (first pass)
__global__ void __launch_bounds__(kSortThreads)
sort_scatter_kernel(const int32_t* __restrict__ keys_in, const int32_t* __restrict__ pos_in, int64_t M, int shift,
                    int num_tiles, const int32_t* __restrict__ hist_scanned, const int32_t* __restrict__ totals,
                    int32_t* __restrict__ keys_out, int32_t* __restrict__ pos_out) {
    __shared__ unsigned short cnt[kSortChunks][kRadix];  // 32 KB: per-chunk digit counts -> exclusive prefix
    __shared__ int32_t gbase[kRadix];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < kSortChunks * kRadix / 2; i += kSortThreads) reinterpret_cast<uint32_t*>(&cnt[0][0])[i] = 0u;
    {   // digit base = exclusive scan of the 256 digit totals (one value per thread)
        __shared__ int32_t wsum[8];
        const int32_t t = totals[tid];
        int32_t incl = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int32_t u = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += u;
        }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        int32_t off = 0;
        for (int w = 0; w < warp; ++w) off += wsum[w];
        gbase[tid] = off + incl - t + hist_scanned[(int64_t)tid * num_tiles + blockIdx.x];
    }
    __syncthreads();

    const int64_t base = (int64_t)blockIdx.x * kSortTile;
    int32_t key[kSortItems];
    int rank[kSortItems];
#pragma unroll
    for (int j = 0; j < kSortItems; ++j) {
        const int chunk = warp * kSortItems + j;
        const int64_t i = base + chunk * 32 + lane;
        const bool valid = i < M;
        key[j] = valid ? keys_in[i] : 0;
        const int dg = valid ? digit_of(key[j], shift) : kRadix + lane;  // invalid lanes: private groups
        const unsigned peers = __match_any_sync(0xffffffffu, dg);
        rank[j] = __popc(peers & ((1u << lane) - 1u));
        if (valid && rank[j] == 0) cnt[chunk][dg] = (unsigned short)__popc(peers);
    }
    __syncthreads();
    {   // thread `tid` owns digit `tid`: exclusive scan over the chunks, in chunk order
        unsigned run = 0;
#pragma unroll 8
        for (int c = 0; c < kSortChunks; ++c) {
            const unsigned v = cnt[c][tid];
            cnt[c][tid] = (unsigned short)run;
            run += v;
        }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < kSortItems; ++j) {
        const int chunk = warp * kSortItems + j;
        const int64_t i = base + chunk * 32 + lane;
        if (i < M) {
            const int dg = digit_of(key[j], shift);
            const int64_t dst = (int64_t)gbase[dg] + cnt[chunk][dg] + rank[j];
            keys_out[dst] = key[j];
            pos_out[dst] = pos_in ? pos_in[i] : (int32_t)i;
        }
    }
}

static int num_passes(int64_t X) {
    int bits = 1;
    while (bits < 31 && ((int64_t)1 << bits) < X) ++bits;
    return (bits + kRadixBits - 1) / kRadixBits;
}

}  // namespace dctr

using namespace dctr;

extern "C" size_t dctr_sort_rows_workspace(int64_t M) {
    if (M <= 0) return 16;
    const int64_t tiles = (M + kSortTile - 1) / kSortTile;
    const int64_t m4 = (M + 3) / 4 * 4;
    return (size_t)(2 * m4 + (int64_t)kRadix * tiles + kRadix) * sizeof(int32_t);
}

extern "C" int dctr_sort_rows(const int32_t* ids, int64_t M, int64_t X, int32_t* sorted_rows, int32_t* sorted_pos,
                              void* workspace, size_t workspace_bytes, void* stream) {
    DCTR_REQUIRE(M >= 0 && M < ((int64_t)1 << 31) && X > 0 && X <= ((int64_t)1 << 31), "sort_rows: bad M/X");
    if (M == 0) return DCTR_OK;
    DCTR_REQUIRE(ids && sorted_rows && sorted_pos && workspace, "sort_rows: null pointer");
    if (workspace_bytes < dctr_sort_rows_workspace(M)) {
        set_error("sort_rows: workspace %zu < %zu bytes", workspace_bytes, dctr_sort_rows_workspace(M));
        return DCTR_EWORKSPACE;
    }
    const int tiles = (int)((M + kSortTile - 1) / kSortTile);
    const int64_t m4 = (M + 3) / 4 * 4;
    int32_t* alt_keys = static_cast<int32_t*>(workspace);
    int32_t* alt_pos = alt_keys + m4;
    int32_t* hist = alt_pos + m4;
    int32_t* totals = hist + (int64_t)kRadix * tiles;
    cudaStream_t st = as_stream(stream);
    const int passes = num_passes(X);
    const int32_t* kin = ids;
    const int32_t* pin = nullptr;
    for (int p = 0; p < passes; ++p) {
        const bool to_out = ((passes - p) & 1) == 1;
        int32_t* kout = to_out ? sorted_rows : alt_keys;
        int32_t* pout = to_out ? sorted_pos : alt_pos;
        const int shift = p * kRadixBits;
        sort_hist_kernel<<<tiles, kSortThreads, 0, st>>>(kin, M, shift, tiles, hist);
        int rc = check_launch("sort_hist_kernel");
        if (rc) return rc;
        sort_scan_kernel<<<kRadix, 1024, 0, st>>>(hist, tiles, totals);
        rc = check_launch("sort_scan_kernel");
        if (rc) return rc;
        sort_scatter_kernel<<<tiles, kSortThreads, 0, st>>>(kin, pin, M, shift, tiles, hist, totals, kout, pout);
        rc = check_launch("sort_scatter_kernel");
        if (rc) return rc;
        kin = kout;
        pin = pout;
    }
    return DCTR_OK;
}
