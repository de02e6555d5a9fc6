// K6a: stable LSD radix sort of (row id, position) pairs -- the deterministic ordering step of the
// sparse update.  Hand-written (no CUB): one histogram kernel for all digits, then ONE kernel per 8-bit digit
// (tile ranking + decoupled look-back + scatter).  Stability inside a tile comes from ranking 32-key rows with
// __match_any_sync on top of per-warp running digit counters, across tiles from the look-back running in tile
// order, so equal rows keep their (sample, field) order: that is what lets SEQ_DECAY reproduce the sample-major
// loop of FNN_wnzh.py:300-306 and makes SUM bit-reproducible.
#include "common.cuh"

namespace dctr {

constexpr int kSortThreads = 256;
constexpr int kSortItems = 8;                       // chunks per warp
constexpr int kSortTile = kSortThreads * kSortItems;  // 2048 elements per CTA
constexpr int kRadixBits = 8;
constexpr int kRadix = 1 << kRadixBits;

__device__ __forceinline__ int digit_of(int32_t key, int shift) { return (key >> shift) & (kRadix - 1); }

constexpr int kMaxPasses = 4;
enum : unsigned { kFlagAgg = 1u, kFlagPrefix = 2u };       // tile status = (count << 2) | flag

// digit histograms of ALL passes in one read of the keys (they do not depend on the order of the list)
__global__ void __launch_bounds__(kSortThreads)
sort_hist_kernel(const int32_t* __restrict__ keys, int64_t M, int passes, unsigned* __restrict__ ghist) {
    pdl_sync();
    __shared__ unsigned h[kMaxPasses][kRadix];
    for (int p = 0; p < passes; ++p) h[p][threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * kSortTile;
#pragma unroll
    for (int j = 0; j < kSortItems; ++j) {
        const int64_t i = base + j * kSortThreads + threadIdx.x;
        if (i < M) {
            const int32_t k = keys[i];
            for (int p = 0; p < passes; ++p) atomicAdd(&h[p][digit_of(k, p * kRadixBits)], 1u);
        }
    }
    __syncthreads();
    for (int p = 0; p < passes; ++p) {
        const unsigned v = h[p][threadIdx.x];
        if (v) atomicAdd(&ghist[p * kRadix + threadIdx.x], v);      // integer atomics: the totals are order-independent
    }
}

// One pass of the LSD sort in ONE kernel ("onesweep"): a CTA takes the next tile (ticket), ranks its 2048 keys stably,
// publishes its per-digit counts, obtains the counts of all earlier tiles by decoupled look-back (thread d walks digit
// d back to the nearest tile whose inclusive prefix is published), and scatters.  Tiles are handed out in ticket order,
// so a CTA only ever waits on tiles that are already running: no deadlock, and the output does not depend on timing.
// pos_in == nullptr: payload is the element index itself (first pass)
__global__ void __launch_bounds__(kSortThreads)
sort_pass_kernel(const int32_t* __restrict__ keys_in, const int32_t* __restrict__ pos_in, int64_t M, int shift,
                 const unsigned* __restrict__ ghist, unsigned* __restrict__ ticket, volatile unsigned* __restrict__ status,
                 int32_t* __restrict__ keys_out, int32_t* __restrict__ pos_out) {
    pdl_sync();
    __shared__ unsigned cnt[kSortThreads / 32][kRadix];   // per-warp digit counts -> exclusive prefix over the warps
    __shared__ unsigned gbase[kRadix];
    __shared__ unsigned lbase[kRadix];
    __shared__ unsigned wsum[kSortThreads / 32], lsum[kSortThreads / 32];
    __shared__ int32_t skey[kSortTile], spos[kSortTile];
    __shared__ unsigned s_tile;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
#pragma unroll
    for (int w = 0; w < kSortThreads / 32; ++w) cnt[w][tid] = 0;
    __syncthreads();
    const unsigned tile = s_tile;
    const int64_t base = (int64_t)tile * kSortTile;

    // ---- stable ranks inside the tile: warp w owns rows w*8 .. w*8+7 of 32 consecutive keys ----
    int32_t key[kSortItems];
    unsigned rank[kSortItems];
#pragma unroll
    for (int j = 0; j < kSortItems; ++j) {
        const int64_t i = base + (warp * kSortItems + j) * 32 + lane;
        const bool valid = i < M;
        key[j] = valid ? keys_in[i] : 0;
        const int dg = digit_of(key[j], shift);
        // lanes holding the same digit: eight ballots (one per digit bit) instead of __match_any_sync, whose cost grows
        // with the number of distinct values in the warp
        unsigned peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
        for (int bit = 0; bit < kRadixBits; ++bit) {
            const unsigned m = __ballot_sync(0xffffffffu, (dg >> bit) & 1);
            peers &= ((dg >> bit) & 1) ? m : ~m;
        }
        const int leader = __ffs(peers) - 1;
        unsigned before = 0;
        if (valid && lane == leader) { before = cnt[warp][dg]; cnt[warp][dg] = before + __popc(peers); }
        before = __shfl_sync(0xffffffffu, before, valid ? leader : lane);
        rank[j] = before + __popc(peers & ((1u << lane) - 1u));
        __syncwarp();
    }
    __syncthreads();

    // ---- thread d owns digit d: prefix over the warps, publish, look back ----
    unsigned total = 0;
#pragma unroll
    for (int w = 0; w < kSortThreads / 32; ++w) { const unsigned v = cnt[w][tid]; cnt[w][tid] = total; total += v; }
    volatile unsigned* mine = status + (int64_t)tile * kRadix + tid;
    unsigned excl = 0;
    if (tile == 0) {
        *mine = (total << 2) | kFlagPrefix;
    } else {
        *mine = (total << 2) | kFlagAgg;
        for (int64_t t = (int64_t)tile - 1; t >= 0; --t) {
            unsigned v;
            do { v = status[t * kRadix + tid]; } while ((v & 3u) == 0u);
            excl += v >> 2;
            if (v & kFlagPrefix) break;
        }
        *mine = ((excl + total) << 2) | kFlagPrefix;
    }
    // ---- exclusive scans over the 256 digits: of the global totals (digit base) and of this tile's totals (local base) ----
    unsigned lbase_d;
    {
        const unsigned t = ghist[tid];
        unsigned incl = t, lincl = total;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned u = __shfl_up_sync(0xffffffffu, incl, o), lu = __shfl_up_sync(0xffffffffu, lincl, o);
            if (lane >= o) { incl += u; lincl += lu; }
        }
        if (lane == 31) { wsum[warp] = incl; lsum[warp] = lincl; }
        __syncthreads();
        unsigned off = 0, loff = 0;
        for (int w = 0; w < warp; ++w) { off += wsum[w]; loff += lsum[w]; }
        lbase_d = loff + lincl - total;
        lbase[tid] = lbase_d;
        // destination of the tile's digit-d run, minus its offset inside the tile: dst = gbase[d] + (index in sorted tile)
        gbase[tid] = off + incl - t + excl - lbase_d;
    }
    __syncthreads();
    // ---- reorder the tile in shared memory (stable), then write out runs: consecutive threads -> consecutive addresses ----
#pragma unroll
    for (int j = 0; j < kSortItems; ++j) {
        const int64_t i = base + (warp * kSortItems + j) * 32 + lane;
        if (i < M) {
            const int dg = digit_of(key[j], shift);
            const unsigned lp = lbase[dg] + cnt[warp][dg] + rank[j];
            skey[lp] = key[j];
            spos[lp] = pos_in ? pos_in[i] : (int32_t)i;
        }
    }
    __syncthreads();
    const int nvalid = (int)min((int64_t)kSortTile, M - base);
#pragma unroll
    for (int j = 0; j < kSortItems; ++j) {
        const int t = j * kSortThreads + tid;
        if (t < nvalid) {
            const int32_t k = skey[t];
            const unsigned dst = gbase[digit_of(k, shift)] + (unsigned)t;      // modulo 2^32: gbase may have wrapped below 0
            keys_out[dst] = k;
            pos_out[dst] = spos[t];
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

static size_t sort_header_words(int64_t tiles) { return (size_t)kMaxPasses * kRadix + 16 + (size_t)kMaxPasses * tiles * kRadix; }

extern "C" size_t dctr_sort_rows_workspace(int64_t M) {
    if (M <= 0) return 16;
    const int64_t tiles = (M + kSortTile - 1) / kSortTile;
    const int64_t m4 = (M + 3) / 4 * 4;
    return (size_t)(2 * m4) * sizeof(int32_t) + sort_header_words(tiles) * sizeof(unsigned);
}

extern "C" int dctr_sort_rows(const int32_t* ids, int64_t M, int64_t X, int32_t* sorted_rows, int32_t* sorted_pos,
                              void* workspace, size_t workspace_bytes, void* stream) {
    DCTR_REQUIRE(M >= 0 && M < ((int64_t)1 << 30) && X > 0 && X <= ((int64_t)1 << 31), "sort_rows: bad M/X");
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
    unsigned* ghist = reinterpret_cast<unsigned*>(alt_pos + m4);       // [passes][256] digit totals
    unsigned* tickets = ghist + kMaxPasses * kRadix;                    // [passes] tile tickets (padded to 16 words)
    unsigned* status = tickets + 16;                                    // [passes][tiles][256] look-back state
    cudaStream_t st = as_stream(stream);
    const int passes = num_passes(X);
    cudaMemsetAsync(ghist, 0, (size_t)(kMaxPasses * kRadix + 16 + (size_t)passes * tiles * kRadix) * sizeof(unsigned), st);
    launch_k(sort_hist_kernel, tiles, kSortThreads, 0, st, ids, M, passes, ghist);
    int rc = check_launch("sort_hist_kernel");
    if (rc) return rc;
    const int32_t* kin = ids;
    const int32_t* pin = nullptr;
    for (int p = 0; p < passes; ++p) {
        const bool to_out = ((passes - p) & 1) == 1;
        int32_t* kout = to_out ? sorted_rows : alt_keys;
        int32_t* pout = to_out ? sorted_pos : alt_pos;
        launch_k(sort_pass_kernel, tiles, kSortThreads, 0, st, kin, pin, M, p * kRadixBits, ghist + p * kRadix, tickets + p,
                                                         status + (size_t)p * tiles * kRadix, kout, pout);
        if ((rc = check_launch("sort_pass_kernel"))) return rc;
        kin = kout;
        pin = pout;
    }
    return DCTR_OK;
}
