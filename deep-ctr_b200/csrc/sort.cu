// K6a: stable LSD radix sort of (row id, position) pairs -- the deterministic ordering step of the
// sparse update.  Hand-written (no CUB): one histogram kernel for all digits, then ONE kernel per 8-bit digit
// (tile ranking + decoupled look-back + scatter).  Stability inside a tile comes from ranking 32-key rows with
// __match_any_sync on top of per-warp running digit counters, across tiles from the look-back running in tile
// order, so equal rows keep their (sample, field) order: that is what lets SEQ_DECAY reproduce the sample-major
// loop of FNN_wnzh.py:300-306 and makes SUM bit-reproducible.
#include <cstdlib>
#include "common.cuh"

namespace dctr {

constexpr int kSortThreads = 256;
constexpr int kSortItems = 8;                       // 32-key rows per warp: 2048 keys per CTA (16 rows = 4096 keys on long lists)
constexpr int kSortTile = kSortThreads * kSortItems;  // histogram tile
constexpr int kMaxPasses = 4;
constexpr int kMaxRadix = 1024;
enum : unsigned { kFlagAgg = 1u, kFlagPrefix = 2u };       // tile status = (count << 2) | flag

// Digit width RB: 8 bits (256 bins, static shared memory) or 10 bits (1024 bins, 56 KB of dynamic shared memory).  The
// wide digit is chosen when it saves a pass: 20-bit keys sort in 2 passes instead of 3, 25..30-bit keys in 3 instead of 4.
template <int RB>
__device__ __forceinline__ int digit_of(int32_t key, int shift) { return (key >> shift) & ((1 << RB) - 1); }
// keys outside [0, X) (malformed input) become the sentinel X: they sort behind every table row and the fold never touches
// them; the key width is that of X + 1, so the sentinel is representable
__device__ __forceinline__ int32_t sane_key(int32_t k, int32_t X) { return (unsigned)k < (unsigned)X ? k : X; }

// digit histograms of ALL passes in one read of the keys (they do not depend on the order of the list)
template <int RB>
__global__ void __launch_bounds__(kSortThreads)
sort_hist_kernel(const int32_t* __restrict__ keys, int64_t M, int passes, unsigned* __restrict__ ghist, int32_t X,
                 int32_t* __restrict__ id_stat) {
    pdl_sync();
    constexpr int R = 1 << RB;
    extern __shared__ unsigned h[];                  // [passes][R]
    for (int i = threadIdx.x; i < passes * R; i += kSortThreads) h[i] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * kSortTile;
#pragma unroll
    for (int j = 0; j < kSortItems; ++j) {
        const int64_t i = base + j * kSortThreads + threadIdx.x;
        if (i < M) {
            const int32_t raw = keys[i], k = sane_key(raw, X);
            if (k != raw && id_stat) { atomicAdd(id_stat, 1); id_stat[1] = raw; }
            for (int p = 0; p < passes; ++p) atomicAdd(&h[p * R + digit_of<RB>(k, p * RB)], 1u);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * R; i += kSortThreads) {
        const unsigned v = h[i];
        if (v) atomicAdd(&ghist[i], v);              // integer atomics: the totals are order-independent
    }
}

// One pass of the LSD sort in ONE kernel ("onesweep"): a CTA takes the next tile (ticket), ranks its 2048 keys stably,
// publishes its per-digit counts, obtains the counts of all earlier tiles by decoupled look-back (every thread walks its
// R/256 digits back to the nearest tile whose inclusive prefix is published), and scatters.  Tiles are handed out in ticket
// order, so a CTA only ever waits on tiles that are already running: no deadlock, and the output does not depend on timing.
// pos_in == nullptr: payload is the element index itself (first pass)
template <int RB, int ITEMS>
__global__ void __launch_bounds__(kSortThreads)
sort_pass_kernel(const int32_t* __restrict__ keys_in, const int32_t* __restrict__ pos_in, int64_t M, int shift,
                 const unsigned* __restrict__ ghist, unsigned* __restrict__ ticket, volatile unsigned* __restrict__ status,
                 int32_t* __restrict__ keys_out, int32_t* __restrict__ pos_out, int32_t X, int first) {
    pdl_sync();
    constexpr int R = 1 << RB;
    constexpr int DPT = R / kSortThreads;                 // digits owned by a thread: tid*DPT .. tid*DPT + DPT-1
    constexpr int NW = kSortThreads / 32;
    extern __shared__ unsigned sort_smem[];
    unsigned* cnt = sort_smem;                            // [NW][R] per-warp digit counts -> exclusive prefix over the warps
    unsigned* gbase = cnt + NW * R;                       // [R]
    unsigned* lbase = gbase + R;                          // [R]
    constexpr int TILE = kSortThreads * ITEMS;
    int32_t* skey = reinterpret_cast<int32_t*>(lbase + R);   // [TILE]
    int32_t* spos = skey + TILE;                          // [TILE]
    __shared__ unsigned wsum[NW], lsum[NW];
    __shared__ unsigned s_tile;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    for (int i = tid; i < NW * R; i += kSortThreads) cnt[i] = 0;
    __syncthreads();
    const unsigned tile = s_tile;
    const int64_t base = (int64_t)tile * TILE;

    // ---- stable ranks inside the tile: warp w owns rows w*ITEMS .. w*ITEMS+ITEMS-1 of 32 consecutive keys ----
    // keys AND payloads are fetched up front (the payload is only needed by the reorder step, but loading it there exposed a
    // full memory latency per tile)
    int32_t key[ITEMS], pay[ITEMS];
    unsigned rank[ITEMS];
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
        const int64_t i = base + (warp * ITEMS + j) * 32 + lane;
        const bool valid = i < M;
        key[j] = valid ? keys_in[i] : 0;
        pay[j] = valid ? (pos_in ? pos_in[i] : (int32_t)i) : 0;
    }
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
        const int64_t i = base + (warp * ITEMS + j) * 32 + lane;
        const bool valid = i < M;
        if (first) key[j] = sane_key(key[j], X);
        const int dg = digit_of<RB>(key[j], shift);
        // lanes holding the same digit: one ballot per digit bit instead of __match_any_sync, whose cost grows with the
        // number of distinct values in the warp
        unsigned peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
        for (int bit = 0; bit < RB; ++bit) {
            const unsigned m = __ballot_sync(0xffffffffu, (dg >> bit) & 1);
            peers &= ((dg >> bit) & 1) ? m : ~m;
        }
        const int leader = __ffs(peers) - 1;
        unsigned before = 0;
        if (valid && lane == leader) { before = cnt[warp * R + dg]; cnt[warp * R + dg] = before + __popc(peers); }
        before = __shfl_sync(0xffffffffu, before, valid ? leader : lane);
        rank[j] = before + __popc(peers & ((1u << lane) - 1u));
        __syncwarp();
    }
    __syncthreads();

    // ---- per owned digit: prefix over the warps, publish, look back ----
    unsigned total[DPT], excl[DPT];
#pragma unroll
    for (int q = 0; q < DPT; ++q) {
        const int d = tid * DPT + q;
        unsigned t = 0;
#pragma unroll
        for (int w = 0; w < NW; ++w) { const unsigned v = cnt[w * R + d]; cnt[w * R + d] = t; t += v; }
        total[q] = t;
        excl[q] = 0;
    }
    volatile unsigned* mine = status + (int64_t)tile * R + tid * DPT;
    if (tile == 0) {
#pragma unroll
        for (int q = 0; q < DPT; ++q) mine[q] = (total[q] << 2) | kFlagPrefix;
    } else {
#pragma unroll
        for (int q = 0; q < DPT; ++q) mine[q] = (total[q] << 2) | kFlagAgg;
        unsigned open = (1u << DPT) - 1u;                 // digits whose walk has not met an inclusive prefix yet
        for (int64_t t = (int64_t)tile - 1; t >= 0 && open; --t) {
            const volatile unsigned* theirs = status + t * R + tid * DPT;
            unsigned v[DPT];
            bool ready;
            do {                                           // the DPT words of one tile are read together
                ready = true;
#pragma unroll
                for (int q = 0; q < DPT; ++q) { v[q] = theirs[q]; ready = ready && ((v[q] & 3u) != 0u || !((open >> q) & 1u)); }
            } while (!ready);
#pragma unroll
            for (int q = 0; q < DPT; ++q) {
                if ((open >> q) & 1u) {
                    excl[q] += v[q] >> 2;
                    if (v[q] & kFlagPrefix) open &= ~(1u << q);
                }
            }
        }
#pragma unroll
        for (int q = 0; q < DPT; ++q) mine[q] = ((excl[q] + total[q]) << 2) | kFlagPrefix;
    }
    // ---- exclusive scans over the R digits: of the global totals (digit base) and of this tile's totals (local base) ----
    {
        unsigned gt[DPT], gsum = 0, lsum_t = 0;
#pragma unroll
        for (int q = 0; q < DPT; ++q) { gt[q] = ghist[tid * DPT + q]; gsum += gt[q]; lsum_t += total[q]; }
        unsigned incl = gsum, lincl = lsum_t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned u = __shfl_up_sync(0xffffffffu, incl, o), lu = __shfl_up_sync(0xffffffffu, lincl, o);
            if (lane >= o) { incl += u; lincl += lu; }
        }
        if (lane == 31) { wsum[warp] = incl; lsum[warp] = lincl; }
        __syncthreads();
        unsigned off = 0, loff = 0;
        for (int w = 0; w < warp; ++w) { off += wsum[w]; loff += lsum[w]; }
        unsigned g_run = off + incl - gsum, l_run = loff + lincl - lsum_t;       // exclusive prefixes of this thread's first digit
#pragma unroll
        for (int q = 0; q < DPT; ++q) {
            const int d = tid * DPT + q;
            lbase[d] = l_run;
            // destination of the tile's digit-d run, minus its offset inside the tile: dst = gbase[d] + (index in sorted tile)
            gbase[d] = g_run + excl[q] - l_run;
            g_run += gt[q];
            l_run += total[q];
        }
    }
    __syncthreads();
    // ---- reorder the tile in shared memory (stable), then write out runs: consecutive threads -> consecutive addresses ----
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
        const int64_t i = base + (warp * ITEMS + j) * 32 + lane;
        if (i < M) {
            const int dg = digit_of<RB>(key[j], shift);
            const unsigned lp = lbase[dg] + cnt[warp * R + dg] + rank[j];
            skey[lp] = key[j];
            spos[lp] = pay[j];            // first pass: the caller's payload, or the element index
        }
    }
    __syncthreads();
    const int nvalid = (int)min((int64_t)TILE, M - base);
#pragma unroll
    for (int j = 0; j < ITEMS; ++j) {
        const int t = j * kSortThreads + tid;
        if (t < nvalid) {
            const int32_t k = skey[t];
            const unsigned dst = gbase[digit_of<RB>(k, shift)] + (unsigned)t;      // modulo 2^32: gbase may have wrapped below 0
            keys_out[dst] = k;
            pos_out[dst] = spos[t];
        }
    }
}

static int key_bits(int64_t X) {
    int bits = 1;
    while (bits < 31 && ((int64_t)1 << bits) < X) ++bits;
    return bits;
}
// digit width: 10 bits when that saves a pass over 8-bit digits AND the list is short.  A short list is latency-bound (a
// pass costs ~8 us whatever it does: M = 65536, 20-bit keys: 36.8 -> 28.6 us); on a long one the wide pass loses more to
// its 56 KB of shared memory, 10 ballots per key and 4 KB of look-back state per tile than the saved pass returns
// (M = 16.7 M, 25-bit keys: 0.90 -> 1.19 ms), and at M = 1 M the two are equal.
// DCTR_SORT_BITS = 8 / 9 / 10 forces the digit width (experiments: 9-bit digits sort 25-27-bit row ids in 3 passes instead of 4)
static int forced_bits() {
    static const int v = [] {
        const char* e = getenv("DCTR_SORT_BITS");
        const int b = e ? atoi(e) : 0;
        return (b == 8 || b == 9 || b == 10) ? b : 0;
    }();
    return v;
}
static int radix_bits(int64_t X, int64_t M) {
    if (forced_bits() && (forced_bits() < 10 || M <= ((int64_t)1 << 18))) return forced_bits();
    const int bits = key_bits(X + 1);
    if (M <= ((int64_t)1 << 18) && (bits + 9) / 10 < (bits + 7) / 8) return 10;
    // long lists: 9-bit digits when they save a pass (25..27-bit row ids sort in 3 passes instead of 4: 1.00 -> 0.86 ms at
    // 16.7 M keys, profiles/r02_sort_experiments.txt); 512 bins still fit the per-warp counters in shared memory
    return (bits + 8) / 9 < (bits + 7) / 8 ? 9 : 8;
}
static int num_passes(int64_t X, int64_t M) {
    const int rb = radix_bits(X, M);
    return (key_bits(X + 1) + rb - 1) / rb;
}
template <int RB, int ITEMS>
static constexpr size_t pass_smem() { return ((size_t)(kSortThreads / 32 + 2) * (1 << RB) + 2 * kSortThreads * ITEMS) * sizeof(unsigned); }

// keys per CTA of the pass kernel (DCTR_SORT_ITEMS = 8 | 16 rows of 32 keys per warp: experiment knob)
static int pass_items(int64_t M) {
    static const int forced = [] { const char* e = getenv("DCTR_SORT_ITEMS"); const int v = e ? atoi(e) : 0; return (v == 8 || v == 16) ? v : 0; }();
    if (forced) return forced;
    (void)M;
    return 8;          // 4096-key tiles (128-146 registers, 1-2 CTAs per SM) measured slower: 0.957 vs 0.767 ms at 16.7 M keys
}

template <int RB, int ITEMS>
static int run_sort(const int32_t* ids, const int32_t* payload, int64_t M, int32_t X, int passes, int32_t* sorted_rows, int32_t* sorted_pos,
                    int32_t* alt_keys, int32_t* alt_pos, unsigned* ghist, unsigned* tickets, unsigned* status, cudaStream_t st) {
    constexpr int R = 1 << RB;
    const int htiles = (int)((M + kSortTile - 1) / kSortTile);
    const int tiles = (int)((M + kSortThreads * ITEMS - 1) / (kSortThreads * ITEMS));
    static bool attr_set = false;
    if (!attr_set) {       // above the 48 KB default for wide digits / large tiles
        cudaFuncSetAttribute(sort_pass_kernel<RB, ITEMS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pass_smem<RB, ITEMS>());
        attr_set = true;
    }
    launch_k(sort_hist_kernel<RB>, htiles, kSortThreads, (size_t)passes * R * sizeof(unsigned), st, ids, M, passes, ghist, X, id_status());
    int rc = check_launch("sort_hist_kernel");
    if (rc) return rc;
    const int32_t* kin = ids;
    const int32_t* pin = payload;
    for (int p = 0; p < passes; ++p) {
        const bool to_out = ((passes - p) & 1) == 1;
        int32_t* kout = to_out ? sorted_rows : alt_keys;
        int32_t* pout = to_out ? sorted_pos : alt_pos;
        launch_k(sort_pass_kernel<RB, ITEMS>, tiles, kSortThreads, pass_smem<RB, ITEMS>(), st, kin, pin, M, p * RB, ghist + p * R, tickets + p,
                 status + (size_t)p * tiles * R, kout, pout, X, (int)(p == 0));
        if ((rc = check_launch("sort_pass_kernel"))) return rc;
        kin = kout;
        pin = pout;
    }
    return DCTR_OK;
}

}  // namespace dctr

using namespace dctr;

// ghist [kMaxPasses][kMaxRadix] | tickets [16] | status [passes][tiles][R]: 4 passes x 256 bins, or (short lists only)
// 3 passes x 1024 bins
static int64_t pass_tiles(int64_t M) { const int64_t t = (int64_t)kSortThreads * pass_items(M); return (M + t - 1) / t; }
static size_t sort_header_words(int64_t M) {
    // look-back words per tile: up to 3 passes x 1024 bins on short lists, 4 passes x 512 bins otherwise
    const size_t per_tile = M <= ((int64_t)1 << 18) ? (size_t)3 * kMaxRadix : (size_t)kMaxPasses * 512;
    return (size_t)kMaxPasses * kMaxRadix + 16 + per_tile * (size_t)pass_tiles(M);
}

extern "C" size_t dctr_sort_rows_workspace(int64_t M) {
    if (M <= 0) return 16;
    const int64_t m4 = (M + 3) / 4 * 4;
    return (size_t)(2 * m4) * sizeof(int32_t) + sort_header_words(M) * sizeof(unsigned);
}

extern "C" int dctr_sort_rows(const int32_t* ids, int64_t M, int64_t X, int32_t* sorted_rows, int32_t* sorted_pos,
                              void* workspace, size_t workspace_bytes, void* stream) {
    return dctr_sort_pairs(ids, nullptr, M, X, sorted_rows, sorted_pos, workspace, workspace_bytes, stream);
}

extern "C" int dctr_sort_pairs(const int32_t* ids, const int32_t* payload, int64_t M, int64_t X, int32_t* sorted_rows,
                               int32_t* sorted_pos, void* workspace, size_t workspace_bytes, void* stream) {
    DCTR_REQUIRE(M >= 0 && M < ((int64_t)1 << 30) && X > 0 && X < ((int64_t)1 << 31) - 1, "sort_rows: bad M/X");
    if (M == 0) return DCTR_OK;
    DCTR_REQUIRE(ids && sorted_rows && sorted_pos && workspace, "sort_rows: null pointer");
    if (workspace_bytes < dctr_sort_rows_workspace(M)) {
        set_error("sort_rows: workspace %zu < %zu bytes", workspace_bytes, dctr_sort_rows_workspace(M));
        return DCTR_EWORKSPACE;
    }
    const int64_t tiles = pass_tiles(M);
    const int64_t m4 = (M + 3) / 4 * 4;
    int32_t* alt_keys = static_cast<int32_t*>(workspace);
    int32_t* alt_pos = alt_keys + m4;
    unsigned* ghist = reinterpret_cast<unsigned*>(alt_pos + m4);       // [passes][R] digit totals
    unsigned* tickets = ghist + kMaxPasses * kMaxRadix;                 // [passes] tile tickets (padded to 16 words)
    unsigned* status = tickets + 16;                                    // [passes][tiles][R] look-back state
    cudaStream_t st = as_stream(stream);
    const int rb = radix_bits(X, M), passes = num_passes(X, M), R = 1 << rb;
    cudaMemsetAsync(ghist, 0, (size_t)(kMaxPasses * kMaxRadix + 16 + (size_t)passes * tiles * R) * sizeof(unsigned), st);
#define DCTR_RUN_SORT(RB_, IT_) run_sort<RB_, IT_>(ids, payload, M, (int32_t)X, passes, sorted_rows, sorted_pos, alt_keys, alt_pos, ghist, tickets, status, st)
    const bool big = pass_items(M) == 16;
    if (rb == 10) return DCTR_RUN_SORT(10, 8);
    if (rb == 9) return big ? DCTR_RUN_SORT(9, 16) : DCTR_RUN_SORT(9, 8);
    return big ? DCTR_RUN_SORT(8, 16) : DCTR_RUN_SORT(8, 8);
#undef DCTR_RUN_SORT
}
