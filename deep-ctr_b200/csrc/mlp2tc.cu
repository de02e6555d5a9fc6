// Fused train step of a TWO-hidden-layer MLP on the tensor cores: forward, output unit and the whole data-gradient chain of
// the Theano script's network (177-300-100-1, FNN_wnzh.py:133-183) and of FNN.py at iPinYou sizes as ONE kernel, one CTA per
// 128 samples.  The per-layer tcgen05 chain (gemm_tc.cu) spends 15-28 us per GEMM on ~6 k-blocks of work (launch, pipeline
// fill, an epilogue through L2 and back: profiles/r02_ncu_fnn_gemm_tc_small.txt); the FFMA fusion (mlp2.cu) removed the launches
// but lost to the tensor cores.  Here the layers are chained INSIDE the kernel and stay on tcgen05:
//
//   phase F1   z1 = x W1            A = x chunks of 32 columns, B = W1 [k][n]  (MN-major tiles)
//   phase F2   z2 = a1 W2           A = a1 chunks, written by the F1 epilogue straight into the A ring;  B = W2 [k][n]
//   out unit   logit = a2 w3 + b3 -> pred, loss, dlogit, CTA partials of d w3 | d b3 | loss; dz2 = dlogit w3 m2 act2'(a2)
//   phase B2   da1 = dz2 W2^T       A = dz2 chunks, B = W2 read K-major ([n = row of W2][k])
//   phase B1   dz0 = dz1 W1^T       A = dz1 chunks = da1 m1 act1'(a1) from the B2 epilogue, B = W1 read K-major
//
//   Accumulators live in tensor memory with lane = sample, an A operand is a K-major SWIZZLE_128B tile (row = sample, 128 bytes
//   = 32 columns), and global memory wants whole rows per warp instruction.  An epilogue warp therefore (1) reads 32 columns
//   of its 32 rows with tcgen05.ld (lane = row), (2) parks them in the A slot it is about to fill -- the swizzle makes the
//   row-per-lane 128-bit stores conflict-free --, (3) reads them back "row-major" (lane = (row % 4, 16-byte column), 8 steps of
//   4 rows), applies bias / activation / mask / the stored activation's derivative there (elementwise work commutes with the
//   transposition, and every global access is now a coalesced 128-bit one), (4) stores the result to global memory (a1, dz2,
//   dz1 are what the weight gradients read, dz0 what the sparse update folds) and (5) writes it back into the same slot split
//   into tf32 hi / lo planes.  No activation makes a round trip through L2 between layers.
//
//   3xTF32 as in gemm_tc.cu (acc += A_lo B_hi + A_hi B_lo + A_hi B_hi).  The weights arrive PRE-PACKED (mlp2tc_pack_kernel, a
//   side-stream launch per step: W only changes in the update): split into tf32 hi / lo and laid out tile by tile in exactly
//   the swizzled shared-memory image the MMAs read, in the order the kernel consumes them -- a B stage is ONE contiguous bulk
//   copy (cp.async.bulk, <= 40 KB).  One accumulator per output (K <= 320: the truncating accumulation stays below 1e-6).
//
//   CTA = 14 warps: warps 0-11 workers (row quadrant = warp % 4; the three warps of a quadrant take the 32-column chunks in turn),
//   warp 12 allocates TMEM and issues the MMAs (all lanes run the loop, one elected lane issues: tc_common.cuh), one lane of
//   warp 13 streams the weight tiles through a 3-stage ring.
//   Rings: A 3 slots x (hi 16 KB | lo 16 KB), B 3 stages x (hi 20 KB | lo 20 KB: up to 160 output columns x 32 k).
//   TMEM: forward z1 [0, 320) z2 [320, 448); backward da1 [0, 320) dz0 [320, 512).
#include <stdlib.h>
#include "tc_common.cuh"

namespace dctr {
namespace m2tc {
using namespace tc;

constexpr int kRows = 128;                 // rows of an MMA tile (UMMA M)
// Samples per CTA.  The epilogues, not the MMAs, bound this kernel (~30 instructions per activation element against ~45 ns per
// MMA), and 4096 / 128 = 32 CTAs leave 116 SMs idle: a CTA therefore takes only 16 * 4 = 64 samples -- rows 0-15 of each
// 32-lane TMEM quadrant, so that every worker warp (which can only reach the TMEM lanes of quadrant warp % 4) keeps a
// share -- and issues the same M = 128 MMAs over half-empty tiles (rows 16-31 of a quadrant stay zero).
#ifndef DCTR_M2TC_QROWS
#define DCTR_M2TC_QROWS 16
#endif
constexpr int kQRows = DCTR_M2TC_QROWS;     // valid rows per quadrant: 16 or 32
constexpr int kTileRows = 4 * kQRows;
static_assert(kQRows == 8 || kQRows == 16 || kQRows == 32, "rows per quadrant");
constexpr int kWorkers = 384;             // 12 warps = 3 groups of 4 row quadrants: three worker warps per scheduler hide each other's latencies
constexpr int kGroups = kWorkers / 128;   // == kASlots: inside a phase a group always fills the same slot, so a writer is never more than
                                           // one mbarrier phase ahead of the MMAs (a parity wait cannot tell phases two apart)
constexpr int kThreads = kWorkers + 64;
constexpr int kMmaWarp = kWorkers / 32, kTmaWarp = kMmaWarp + 1;
constexpr int kASlots = 3, kBStages = 3;
static_assert(kGroups == kASlots, "chunk -> group and chunk -> slot must stay in step");
constexpr int kATile = kRows * 128;            // [128 rows x 32 fp32]
constexpr int kASlot = 2 * kATile;             // hi | lo
constexpr int kMaxHalf = 160;                  // output columns per MMA group
constexpr int kBStage = 2 * kMaxHalf * 128;    // hi | lo
constexpr int kScratch = 9728;                 // output unit (partial dots 512, column partials 512, warp sums 32 floats), constants (1216), dlogit (128)
constexpr int kSmem = kASlots * kASlot + kBStages * kBStage + kScratch + 256 + 1024;
constexpr int kMaxD0 = 192, kMaxH1 = 320, kMaxH2 = 128;
#ifndef DCTR_M2TC_POLL_NS
#define DCTR_M2TC_POLL_NS 50
#endif
constexpr int kPollNs = DCTR_M2TC_POLL_NS;      // back-off of the MMA issuer's polls (try_wait returns after a few dozen ns: measured 1 poll / 60 ns)
constexpr int kPollTmaNs = 200;                 // the TMA lane runs three stages ahead

// one GEMM phase: acc[128, nh * hw] (+)= A[128, 32 * nchunks] * B
struct Phase {
    int nchunks;        // k-blocks of 32
    int nh, hw;         // column groups ("halves") and their width in TMEM columns / B-tile rows (multiple of 32)
    int nmma[2];        // UMMA N of each group (multiple of 16)
    int acc_col;        // first TMEM column of the accumulator
    int kmajor;         // B tile layout: 1 = [n][32 k] SWIZZLE_128B, 0 = [32 k][32 n] atoms (SWIZZLE_128B_BASE32B)
};

struct Args {
    int N, D0, H1, H2;
    int ld0, ld1, ldd;
    int act0, act1, act2;
    float* x; const float* m0; int ldm0;
    const float* b1; const float* b2; float* w3; float* b3;
    const float* m1; int ldm1; const float* m2; int ldm2;
    const float* y; float gscale, loss_scale;
    float* a1; float* dz2; float* dz1; float* dz0;
    float* logit; float* pred; float* loss_n; float* dlogit; float* loss_sum;
    float* gw3; float* gb3;
    int sgd_out; float lr, decay_out;
    float* partial; unsigned* ticket;       // [tiles][H2 + 2] CTA partials of (d w3 | d b3 | loss), ticket word
    const uint8_t* packed;                  // weight tiles in consumption order (mlp2tc_pack_kernel)
    unsigned long long* timing;             // diagnostic (DCTR_M2TC_TIMING=1): %globaltimer stamps of CTA 0, else NULL
    Phase ph[4];
};

// tanh / sigmoid through ex2.approx + rcp.approx (5 instructions, ABSOLUTE error < 3e-7 -- small |x| included --, exact saturation), inlined so that the
// independent chains of a chunk overlap (a non-inlined tanhf per element made the first version's epilogues a serial chain)
__device__ __forceinline__ float ex2_approx(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float rcp_approx(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float fast_tanh(float x) { return fmaf(-2.f, rcp_approx(ex2_approx(x * 2.8853900817779268f) + 1.f), 1.f); }
__device__ __forceinline__ float fast_sigmoid(float x) { return rcp_approx(1.f + ex2_approx(x * -1.4426950408889634f)); }
__device__ __forceinline__ float4 act4(int act, float4 v) {
    if (act == DCTR_ACT_TANH) return make_float4(fast_tanh(v.x), fast_tanh(v.y), fast_tanh(v.z), fast_tanh(v.w));
    if (act == DCTR_ACT_RELU) return make_float4(fmaxf(v.x, 0.f), fmaxf(v.y, 0.f), fmaxf(v.z, 0.f), fmaxf(v.w, 0.f));
    if (act == DCTR_ACT_SIGMOID) return make_float4(fast_sigmoid(v.x), fast_sigmoid(v.y), fast_sigmoid(v.z), fast_sigmoid(v.w));
    return v;
}
// a = m * act(z) (the stored, masked activation) -> m * act'(z), 0 where the unit was dropped; masked = false: m = 1
__device__ __forceinline__ float4 dact4(int act, bool masked, float4 a, float4 m) {
    if (masked) { a.x *= rcp_approx(m.x); a.y *= rcp_approx(m.y); a.z *= rcp_approx(m.z); a.w *= rcp_approx(m.w); }   // 0 * inf = NaN: selected away below
    float4 g = make_float4(1.f, 1.f, 1.f, 1.f);
    if (act == DCTR_ACT_TANH) g = make_float4(fmaf(-a.x, a.x, 1.f), fmaf(-a.y, a.y, 1.f), fmaf(-a.z, a.z, 1.f), fmaf(-a.w, a.w, 1.f));
    else if (act == DCTR_ACT_RELU) g = make_float4(a.x > 0.f ? 1.f : 0.f, a.y > 0.f ? 1.f : 0.f, a.z > 0.f ? 1.f : 0.f, a.w > 0.f ? 1.f : 0.f);
    else if (act == DCTR_ACT_SIGMOID) g = make_float4(a.x * (1.f - a.x), a.y * (1.f - a.y), a.z * (1.f - a.z), a.w * (1.f - a.w));
    if (masked) {
        g.x = m.x != 0.f ? g.x * m.x : 0.f; g.y = m.y != 0.f ? g.y * m.y : 0.f;
        g.z = m.z != 0.f ? g.z * m.z : 0.f; g.w = m.w != 0.f ? g.w * m.w : 0.f;
    }
    return g;
}
__device__ __forceinline__ float4 mul4(float4 a, float4 b) { return make_float4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
__device__ __forceinline__ float4 add4(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
// columns col .. col+3: zero those >= H, everything if !ok
__device__ __forceinline__ float4 clip4(float4 v, int col, int H, bool ok) {
    if (!ok || col >= H) v.x = 0.f;
    if (!ok || col + 1 >= H) v.y = 0.f;
    if (!ok || col + 2 >= H) v.z = 0.f;
    if (!ok || col + 3 >= H) v.w = 0.f;
    return v;
}

// ---- the two views of a warp's [32 rows x 32 columns] chunk inside a K-major SWIZZLE_128B tile -------------------------------------
// Two accumulators per sample out of the idle half of the tile (kQRows == 16).  The tensor core TRUNCATES when it adds an MMA's
// products to the accumulator -- a bias that grows with the number of inexact additions (one accumulator, K = 300: ~3e-6 relative,
// and it compounds over the four chained GEMMs).  Tensor memory has no room for a second accumulator, but rows 16-31 of every
// quadrant are idle: the 16-byte column j of sample row r is stored in row r + 16 * ((j >> 1) & 1), i.e. the k-steps (8 columns)
// of a chunk alternate between the sample's two rows and every other position of both rows stays zero.  Adding an exact zero is
// exact, so each row sees half of the inexact additions; the epilogue adds TMEM lanes r and r + 16 (one shuffle per value).
__device__ __forceinline__ int split_row(int j) { return kQRows == 16 ? 16 * ((j >> 1) & 1) : 0; }
// row-per-lane (TMEM order): lane = row, 16-byte column i
__device__ __forceinline__ uint32_t off_lane(int quad, int lane, int i) {
    const int R = quad * 32 + lane + split_row(i);
    return (uint32_t)((R >> 3) * 1024 + (R & 7) * 128 + ((i ^ (R & 7)) << 4));
}
// row-major (global-memory order): step t < 8, lane -> row 4 t + (lane >> 3), 16-byte column lane & 7
__device__ __forceinline__ int rm_row(int t, int lane) { return 4 * t + (lane >> 3); }
__device__ __forceinline__ uint32_t off_rm(int quad, int t, int lane) {
    const int j = lane & 7, R = quad * 32 + rm_row(t, lane) + split_row(j);
    return (uint32_t)((R >> 3) * 1024 + (R & 7) * 128 + ((j ^ (R & 7)) << 4));
}
// the two partial accumulators of a sample live in TMEM lanes r and r + 16: after this, lanes < 16 (and their mirrors) hold the sum
template <int NV>
__device__ __forceinline__ void join_split(float (&v)[NV]) {
    if (kQRows != 16) return;
#pragma unroll
    for (int j = 0; j < NV; ++j) v[j] += __shfl_xor_sync(0xffffffffu, v[j], 16);
}
__device__ __forceinline__ void park_rows(uint8_t* tile, int quad, int lane, const float (&v)[32]) {
    if (lane >= kQRows) return;                // rows past the quadrant's share stay zero
#pragma unroll
    for (int i = 0; i < 8; ++i)
        *reinterpret_cast<float4*>(tile + off_lane(quad, lane, i)) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
}
// value -> tf32 hi / lo planes of the slot at the same swizzled position
__device__ __forceinline__ void put_split(uint8_t* slot, uint32_t off, float4 v) {
    float4 h, l;
    split_tf32(v.x, h.x, l.x); split_tf32(v.y, h.y, l.y); split_tf32(v.z, h.z, l.z); split_tf32(v.w, h.w, l.w);
    *reinterpret_cast<float4*>(slot + off) = h;
    *reinterpret_cast<float4*>(slot + kATile + off) = l;
}
// the slot is complete for this thread: make the generic-proxy stores visible to the tensor core and signal the MMA warp
__device__ __forceinline__ void publish(uint64_t* a_full_slot) {
    fence_proxy_async();
    tc_fence_before();                // orders this thread's earlier tcgen05.ld before whatever the MMA warp issues next
    mbar_arrive(a_full_slot);
}
// mask of columns col .. col+3 of a sample: [N, ld] (ldm = ld), one broadcast row (ldm = 0: staged in shared memory) or none
__device__ __forceinline__ float4 mask4(const float* m, int ldm, const float* srow, int64_t row, int col, bool ok) {
    if (!m) return make_float4(1.f, 1.f, 1.f, 1.f);
    if (!ldm) return *reinterpret_cast<const float4*>(srow + col);
    return ok ? __ldg(reinterpret_cast<const float4*>(m + row * ldm + col)) : make_float4(0.f, 0.f, 0.f, 0.f);
}

// t[j] of lane l -> lane l returns the sum over the 32 lanes of t[l]: 31 shuffles, fixed association (reproducible)
__device__ __forceinline__ float warp_colsum32(float (&t)[32], int lane) {
#pragma unroll
    for (int s = 0; s < 5; ++s) {
        const int o = 16 >> s, n = 16 >> s;          // n results stay in t[0..n) after this stage
        const bool up = (lane & o) != 0;
#pragma unroll
        for (int i = 0; i < n; ++i) {
            const float send = up ? t[i] : t[i + n];
            const float keep = up ? t[i + n] : t[i];
            t[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
    }
    return t[0];
}
__device__ __forceinline__ void worker_bar() { asm volatile("bar.sync 1, %0;" ::"n"(kWorkers) : "memory"); }


// ---- compact epilogue loops ---------------------------------------------------------------------------------------------------------
// The kernel runs ONCE per SM and every epilogue instruction only a handful of times: straight-line code is paid for in instruction
// fetches from L2, not in issue slots (the fully unrolled first version of these epilogues was 11 600 instructions = 186 KB and spent
// ~2 us per chunk and warp, most of it waiting for code).  Hence loops over the 8 row-major steps, unrolled by 2, and ONE shared
// routine for the two backward epilogues.
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float (&v)[8]) {
    uint32_t* r = reinterpret_cast<uint32_t*>(v);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, bool ok) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem)), "l"(gmem), "r"(ok ? 16 : 0) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
// global rows row0 + rm_row(t), columns col .. col+3 -> plane (same swizzled positions this thread reads back); zero fill outside
__device__ __forceinline__ void rm_fetch(uint8_t* plane, const float* base, int ld, int64_t row0, int col, int N, int lim, int quad, int lane) {
#pragma unroll 4
    for (int t = 0; t < kQRows / 4; ++t) {
        const int64_t grow = row0 + rm_row(t, lane);
        const bool ok = grow < N && col < lim;
        cp_async16(plane + off_rm(quad, t, lane), ok ? base + grow * ld + col : base, ok);
    }
}
// acc = the parked accumulators (hi plane of `tile`), aux = the stored activation (lo plane): out = acc * m * act'(aux / m), written to
// global memory and -- to_slot -- back into the tile as tf32 hi / lo.  Shared by the B2 (dz1) and B1 (dz0) epilogues.
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, float4 v) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __noinline__ void rm_backward(uint32_t tile, int quad, int lane, int64_t row0, int col, int N, int act, const float* m, int ldm,
                                         uint32_t srow, int mlim, int H, float* out, int ldo, int olim, int to_slot) {
#pragma unroll 2
    for (int t = 0; t < kQRows / 4; ++t) {
        const uint32_t off = off_rm(quad, t, lane);
        const int64_t grow = row0 + rm_row(t, lane);
        const bool ok = grow < N;
        float4 u = lds128(tile + off);
        const float4 aux = lds128(tile + kATile + off);
        float4 mk = make_float4(1.f, 1.f, 1.f, 1.f);
        if (m) mk = ldm ? ((ok && col < mlim) ? __ldg(reinterpret_cast<const float4*>(m + grow * ldm + col)) : make_float4(0.f, 0.f, 0.f, 0.f))
                        : lds128(srow + 4 * col);
        u = clip4(mul4(u, dact4(act, m != nullptr, aux, mk)), col, H, ok);
        if (ok && col < olim) *reinterpret_cast<float4*>(out + grow * ldo + col) = u;
        if (to_slot) {
            float4 h, l;
            split_tf32(u.x, h.x, l.x); split_tf32(u.y, h.y, l.y); split_tf32(u.z, h.z, l.z); split_tf32(u.w, h.w, l.w);
            sts128(tile + off, h);
            sts128(tile + kATile + off, l);
        }
    }
}

__global__ void __launch_bounds__(kThreads, 1)
mlp2tc_train_kernel(const Args a) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);     // pointer arithmetic keeps the shared address space (LDS / STS, not generic LD / ST)
    uint8_t* aring = smem;
    uint8_t* bring = smem + kASlots * kASlot;
    float* scratch = reinterpret_cast<float*>(bring + kBStages * kBStage);
    uint64_t* a_full = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(scratch) + kScratch);
    uint64_t* a_empty = a_full + kASlots;
    uint64_t* b_full = a_empty + kASlots;
    uint64_t* b_empty = b_full + kBStages;
    uint64_t* acc_full = b_empty + kBStages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 4);
    __shared__ unsigned s_last;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (a.timing && tid == 0) atomicMin(a.timing + 14, (unsigned long long)global_ns());      // first CTA to start
    if (tid == 0) {
        for (int s = 0; s < kASlots; ++s) { mbar_init(a_full + s, 128); mbar_init(a_empty + s, 1); }
        for (int s = 0; s < kBStages; ++s) { mbar_init(b_full + s, 1); mbar_init(b_empty + s, 1); }
        for (int p = 0; p < 4; ++p) mbar_init(acc_full + p, 1);
        fence_mbar_init();
    }
    if (warp == kMmaWarp) tmem_alloc(tmem_slot, 512);
    if (kQRows < 32) {                          // the unused rows of every A tile are zeros for the whole kernel
        for (int i = tid; i < kASlots * kASlot / 16; i += kThreads) reinterpret_cast<float4*>(aring)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        fence_proxy_async();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_sync();

    if (warp < kMmaWarp) {
        // =================================================== workers ===================================================
        const int quad = warp & 3, grp = warp >> 2, j4 = (lane & 7) * 4;
        const int64_t row0 = (int64_t)blockIdx.x * kTileRows + quad * kQRows;     // first sample of this warp's quadrant
        const int r = quad * 32 + lane;                                     // row layout: this lane's row of the tile
        const int64_t row = row0 + lane;
        const bool row_ok = lane < kQRows && row < a.N;
        const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
        float* sdot = scratch;                  // [3][128] partial logits of the warp groups (+ 128 unused)
        float* colpart = scratch + 512;         // [4][128] per-quadrant column sums of dlogit * a2
        float* wsum = scratch + 1024;           // [4][2]   per-quadrant sums of dlogit, loss
        // constants of the step, zero-filled to the widest chunk: every later access is a shared-memory broadcast instead of a
        // dependent global load in the middle of an epilogue
        float* sb1 = scratch + 1056;            // [320]
        float* sb2 = sb1 + kMaxH1;              // [128]
        float* sw3 = sb2 + kMaxH2;              // [128]
        float* sm0 = sw3 + kMaxH2;              // [192] broadcast mask rows (ldm = 0)
        float* sm1 = sm0 + kMaxD0;              // [320]
        float* sm2 = sm1 + kMaxH1;              // [128]
        float* dl = sm2 + kMaxH2;               // [128] dlogit of the tile's samples (for the row-major pass)
        for (int i = tid; i < kMaxH1; i += kWorkers) {
            sb1[i] = i < a.H1 ? a.b1[i] : 0.f;
            sm1[i] = (a.m1 && !a.ldm1 && i < a.H1) ? a.m1[i] : 0.f;
            if (i < kMaxH2) {
                sb2[i] = i < a.H2 ? a.b2[i] : 0.f;
                sw3[i] = i < a.H2 ? a.w3[i] : 0.f;
                sm2[i] = (a.m2 && !a.ldm2 && i < a.H2) ? a.m2[i] : 0.f;
            }
            if (i < kMaxD0) sm0[i] = (a.m0 && !a.ldm0 && i < a.D0) ? a.m0[i] : 0.f;
        }
        const float yn = row_ok ? a.y[row] : 0.f;
        const float b3v = a.b3[0];
        const int nc0 = a.ph[0].nchunks, nc1 = a.ph[1].nchunks, nc2 = a.ph[2].nchunks;
        const int lim1 = (a.H1 + 3) & ~3, lim2 = (a.H2 + 3) & ~3;
        const bool stamp = a.timing && blockIdx.x == 0 && tid == 0;
#define M2TC_STAMP(i) do { if (stamp) a.timing[i] = global_ns(); } while (0)
        M2TC_STAMP(0);
        worker_bar();
        float v[32];

        // ---- x chunks -> A ring, coalesced (dropout on the input: x *= m0 in place, the weight gradient reads the masked input) ----
        for (int c = grp; c < nc0; c += kGroups) {
            const int slot = c % kASlots, col = 32 * c + j4;
            uint8_t* sl = aring + slot * kASlot;
            mbar_wait(a_empty + slot, ((c / kASlots) & 1) ^ 1);           // the MMAs that read this slot have completed
            rm_fetch(sl, a.x, a.ld0, row0, col, a.N, a.ld0, quad, lane);
            cp_async_wait_all();
#pragma unroll 2
            for (int t = 0; t < kQRows / 4; ++t) {
                const uint32_t off = off_rm(quad, t, lane);
                const int64_t grow = row0 + rm_row(t, lane);
                const bool ok = grow < a.N;
                float4 u = *reinterpret_cast<const float4*>(sl + off);
                if (a.m0) {
                    u = mul4(u, mask4(a.m0, a.ldm0, sm0, grow, col, ok && col < a.ld0));
                    if (ok && col < a.ld0) *reinterpret_cast<float4*>(a.x + grow * a.ld0 + col) = u;
                }
                put_split(sl, off, clip4(u, col, a.D0, ok));      // x pad columns / rows past N: zeros
            }
            publish(a_full + slot);
        }
        int gbase = nc0;
        M2TC_STAMP(1);

        // ---- F1 epilogue: a1 = m1 * act1(z1 + b1) -> global + A ring of F2 ----
        mbar_wait(acc_full + 0, 0);
        tc_fence_after();
        M2TC_STAMP(2);
        for (int c = grp; c < nc1; c += kGroups) {
            const int g = gbase + c, slot = g % kASlots, col = 32 * c + j4;
            uint8_t* sl = aring + slot * kASlot;
            tmem_ld32(tlane + a.ph[0].acc_col + 32 * c, v);
            join_split(v);
            mbar_wait(a_empty + slot, ((g / kASlots) & 1) ^ 1);
            park_rows(sl, quad, lane, v);
            __syncwarp();
            const float4 bias = *reinterpret_cast<const float4*>(sb1 + col);
#pragma unroll 2
            for (int t = 0; t < kQRows / 4; ++t) {
                const uint32_t off = off_rm(quad, t, lane);
                const int64_t grow = row0 + rm_row(t, lane);
                const bool ok = grow < a.N;
                float4 u = act4(a.act1, add4(*reinterpret_cast<const float4*>(sl + off), bias));
                if (a.m1) u = mul4(u, mask4(a.m1, a.ldm1, sm1, grow, col, ok && col < lim1));
                u = clip4(u, col, a.H1, ok);
                if (ok && col < a.ld1) *reinterpret_cast<float4*>(a.a1 + grow * a.ld1 + col) = u;
                put_split(sl, off, u);
            }
            publish(a_full + slot);
        }
        gbase += nc1;
        M2TC_STAMP(3);

        // ---- F2 epilogue, pass 1 (row layout, 8 columns at a time): this sample's partial logit over the group's chunks ----
        mbar_wait(acc_full + 1, 0);
        tc_fence_after();
        M2TC_STAMP(4);
        {
            float dotp = 0.f;
            for (int c = grp; c < nc2; c += kGroups) {
#pragma unroll 1
                for (int q = 0; q < 4; ++q) {
                    const int c0 = 32 * c + 8 * q;
                    float z8[8];
                    tmem_ld8(tlane + a.ph[1].acc_col + c0, z8);
                    join_split(z8);
                    float4 u0 = act4(a.act2, add4(make_float4(z8[0], z8[1], z8[2], z8[3]), *reinterpret_cast<const float4*>(sb2 + c0)));
                    float4 u1 = act4(a.act2, add4(make_float4(z8[4], z8[5], z8[6], z8[7]), *reinterpret_cast<const float4*>(sb2 + c0 + 4)));
                    if (a.m2) {
                        u0 = mul4(u0, mask4(a.m2, a.ldm2, sm2, row, c0, row_ok && c0 < lim2));
                        u1 = mul4(u1, mask4(a.m2, a.ldm2, sm2, row, c0 + 4, row_ok && c0 + 4 < lim2));
                    }
                    u0 = clip4(u0, c0, a.H2, row_ok);
                    u1 = clip4(u1, c0 + 4, a.H2, row_ok);
                    const float4 w0 = *reinterpret_cast<const float4*>(sw3 + c0), w1 = *reinterpret_cast<const float4*>(sw3 + c0 + 4);
                    dotp = fmaf(u0.x, w0.x, dotp); dotp = fmaf(u0.y, w0.y, dotp); dotp = fmaf(u0.z, w0.z, dotp); dotp = fmaf(u0.w, w0.w, dotp);
                    dotp = fmaf(u1.x, w1.x, dotp); dotp = fmaf(u1.y, w1.y, dotp); dotp = fmaf(u1.z, w1.z, dotp); dotp = fmaf(u1.w, w1.w, dotp);
                }
            }
            sdot[grp * kRows + r] = dotp;
        }
        worker_bar();
        // ---- output unit (row layout: a lane owns its sample) ----
        const float z = ((sdot[r] + sdot[kRows + r]) + sdot[2 * kRows + r]) + b3v;
        const float p = sigmoidf_(z);
        float ln = 0.f, d = 0.f;
        if (row_ok) {
            ln = xent_logits(z, yn);
            d = (p - yn) * a.gscale;
            if (grp == 0) {
                if (a.logit) a.logit[row] = z;
                if (a.pred) a.pred[row] = p;
                if (a.loss_n) a.loss_n[row] = ln;
                if (a.dlogit) a.dlogit[row] = d;
            }
        }
        if (grp == 0) {
            const float ds = warp_sum(d), ls = warp_sum(ln);
            if (lane == 0) { wsum[2 * quad] = ds; wsum[2 * quad + 1] = ls; }
        }
        // dlogit per row for the row-major pass (every group computed the same value)
        if (grp == 0) dl[r] = d;
        worker_bar();
        // ---- pass 2: d w3 partials (column sums of dlogit * a2) and dz2 = dlogit * w3 * m2 * act2'(a2) -> global + A ring of B2 ----
        for (int c = grp; c < nc2; c += kGroups) {
            const int g = gbase + c, slot = g % kASlots, col = 32 * c + j4;
            uint8_t* sl = aring + slot * kASlot;
            tmem_ld32(tlane + a.ph[1].acc_col + 32 * c, v);
            join_split(v);
            mbar_wait(a_empty + slot, ((g / kASlots) & 1) ^ 1);
            park_rows(sl, quad, lane, v);
            __syncwarp();
            const float4 bias = *reinterpret_cast<const float4*>(sb2 + col), wq = *reinterpret_cast<const float4*>(sw3 + col);
            float4 cs = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 2
            for (int t = 0; t < kQRows / 4; ++t) {
                const uint32_t off = off_rm(quad, t, lane);
                const int R = rm_row(t, lane);
                const int64_t grow = row0 + R;
                const bool ok = grow < a.N;
                const float4 mk = mask4(a.m2, a.ldm2, sm2, grow, col, ok && col < lim2);
                float4 a2 = act4(a.act2, add4(*reinterpret_cast<const float4*>(sl + off), bias));      // bit-identical to pass 1
                if (a.m2) a2 = mul4(a2, mk);
                a2 = clip4(a2, col, a.H2, ok);
                const float dr = dl[quad * 32 + R];
                cs.x = fmaf(dr, a2.x, cs.x); cs.y = fmaf(dr, a2.y, cs.y); cs.z = fmaf(dr, a2.z, cs.z); cs.w = fmaf(dr, a2.w, cs.w);
                float4 u = mul4(dact4(a.act2, a.m2 != nullptr, a2, mk), make_float4(dr * wq.x, dr * wq.y, dr * wq.z, dr * wq.w));
                u = clip4(u, col, a.H2, ok);
                if (ok && col < lim2) *reinterpret_cast<float4*>(a.dz2 + grow * a.ldd + col) = u;
                put_split(sl, off, u);
            }
            publish(a_full + slot);
            // rows 4 t + (lane >> 3): fold the four row groups in a fixed order
            cs.x += __shfl_xor_sync(0xffffffffu, cs.x, 8); cs.y += __shfl_xor_sync(0xffffffffu, cs.y, 8);
            cs.z += __shfl_xor_sync(0xffffffffu, cs.z, 8); cs.w += __shfl_xor_sync(0xffffffffu, cs.w, 8);
            cs.x += __shfl_xor_sync(0xffffffffu, cs.x, 16); cs.y += __shfl_xor_sync(0xffffffffu, cs.y, 16);
            cs.z += __shfl_xor_sync(0xffffffffu, cs.z, 16); cs.w += __shfl_xor_sync(0xffffffffu, cs.w, 16);
            if (lane < 8) *reinterpret_cast<float4*>(colpart + quad * kRows + col) = cs;
        }
        gbase += nc2;
        worker_bar();
        {
            float* part = a.partial + (size_t)blockIdx.x * (a.H2 + 2);
            if (tid < a.H2) part[tid] = ((colpart[tid] + colpart[kRows + tid]) + colpart[2 * kRows + tid]) + colpart[3 * kRows + tid];
            else if (tid == a.H2) part[tid] = ((wsum[0] + wsum[2]) + wsum[4]) + wsum[6];
            else if (tid == a.H2 + 1) part[tid] = ((wsum[1] + wsum[3]) + wsum[5]) + wsum[7];
        }
        M2TC_STAMP(5);

        // ---- B2 epilogue: dz1 = da1 * m1 * act1'(a1) -> global + A ring of B1 ----
        mbar_wait(acc_full + 2, 0);
        tc_fence_after();
        M2TC_STAMP(6);
        for (int c = grp; c < nc1; c += kGroups) {
            const int g = gbase + c, slot = g % kASlots, col = 32 * c + j4;
            uint8_t* sl = aring + slot * kASlot;
            mbar_wait(a_empty + slot, ((g / kASlots) & 1) ^ 1);
            rm_fetch(sl + kATile, a.a1, a.ld1, row0, col, a.N, a.ld1, quad, lane);      // this thread's own stores of the F1 epilogue (same mapping)
            tmem_ld32(tlane + a.ph[2].acc_col + 32 * c, v);
            join_split(v);
            park_rows(sl, quad, lane, v);
            cp_async_wait_all();
            __syncwarp();
            rm_backward(smem_u32(sl), quad, lane, row0, col, a.N, a.act1, a.m1, a.ldm1, smem_u32(sm1), lim1, a.H1, a.dz1, a.ldd, lim1, 1);
            publish(a_full + slot);
        }
        M2TC_STAMP(7);

        // ---- B1 epilogue: dz0 = (dz1 W1^T) * m0 * act0'(x): the gradient the sparse update folds (Theano's gx) ----
        mbar_wait(acc_full + 3, 0);
        tc_fence_after();
        M2TC_STAMP(8);
        // every MMA has completed: the A ring is free, one staging tile (hi | lo planes) per warp group
        uint8_t* stg = aring + grp * kASlot;
        for (int c = grp; c < nc0; c += kGroups) {
            const int col = 32 * c + j4;
            rm_fetch(stg + kATile, a.x, a.ld0, row0, col, a.N, a.ld0, quad, lane);      // masked in place by this thread in the first loop
            tmem_ld32(tlane + a.ph[3].acc_col + 32 * c, v);
            join_split(v);
            park_rows(stg, quad, lane, v);
            cp_async_wait_all();
            __syncwarp();
            rm_backward(smem_u32(stg), quad, lane, row0, col, a.N, a.act0, a.m0, a.ldm0, smem_u32(sm0), a.ld0, a.D0, a.dz0, a.ld0, a.ld0, 0);
            __syncwarp();                           // the next chunk's park_rows overwrites what other lanes have just read
        }
        M2TC_STAMP(9);
#undef M2TC_STAMP
    } else if (warp == kTmaWarp && lane == 0) {
        // =================================================== weight tiles ===================================================
        int it = 0;
        const uint8_t* src = a.packed;
        for (int p = 0; p < 4; ++p) {
            const Phase ph = a.ph[p];
            const uint32_t bytes = 2u * ph.hw * 128u;          // hi tile | lo tile
            for (int c = 0; c < ph.nchunks; ++c)
                for (int h = 0; h < ph.nh; ++h, ++it, src += bytes) {
                    const int s = it % kBStages;
                    mbar_wait<kPollTmaNs>(b_empty + s, ((it / kBStages) & 1) ^ 1);
                    mbar_expect_tx(b_full + s, bytes);
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 ::"r"(smem_u32(bring + s * kBStage)), "l"(src), "r"(bytes), "r"(smem_u32(b_full + s)) : "memory");
                }
        }
    } else if (warp == kMmaWarp) {
        // =================================================== MMA issuer ===================================================
        // All 32 lanes run the loops and the waits; one elected lane issues (see elect_one()).
        // instruction descriptor: D = F32 (1 << 4), A = B = TF32 (2 << 7, 2 << 10), A K-major (bit 15 = 0), b_major bit 16,
        // N >> 3 at bit 17, M >> 4 at bit 24
        const bool tstamp = a.timing && blockIdx.x == 0;
        int it = 0, g = 0;
        for (int p = 0; p < 4; ++p) {
            const int nchunks = a.ph[p].nchunks, nh = a.ph[p].nh, hw = a.ph[p].hw, kmajor = a.ph[p].kmajor, acc_col = a.ph[p].acc_col;
            const uint32_t idesc0 = (1u << 4) | (2u << 7) | (2u << 10) | ((kmajor ? 0u : 1u) << 16) | ((uint32_t)(kRows >> 4) << 24);
            for (int c = 0; c < nchunks; ++c, ++g) {
                const int slot = g % kASlots;
                mbar_wait<kPollNs>(a_full + slot, (g / kASlots) & 1);
                tc_fence_after();
                const uint32_t sa = smem_u32(aring + slot * kASlot);
                const uint64_t a_hi = make_desc(sa), a_lo = make_desc(sa + kATile);
                for (int h = 0; h < nh; ++h, ++it) {
                    const int s = it % kBStages;
                    mbar_wait<kPollNs>(b_full + s, (it / kBStages) & 1);
                    tc_fence_after();
                    const uint32_t sb = smem_u32(bring + s * kBStage);
                    const uint32_t idesc = idesc0 | ((uint32_t)(a.ph[p].nmma[h] >> 3) << 17);
                    const uint32_t dcol = tmem_base + acc_col + h * hw;
                    // K-major B: +32 bytes per k-step inside the 128-byte rows; MN-major: +1024 bytes (8 k-rows = 2 atoms)
                    const uint64_t b_hi = kmajor ? make_desc(sb) : make_desc_mn(sb), b_lo = kmajor ? make_desc(sb + hw * 128) : make_desc_mn(sb + hw * 128);
                    const uint64_t bstep = kmajor ? 2 : (1024 >> 4);
                    if (elect_one()) {
#pragma unroll
                        for (int k4 = 0; k4 < 4; ++k4) {
                            mma_tf32(dcol, a_lo + 2 * k4, b_hi + bstep * k4, idesc, (c | k4) != 0);
                            mma_tf32(dcol, a_hi + 2 * k4, b_lo + bstep * k4, idesc, 1);
                            mma_tf32(dcol, a_hi + 2 * k4, b_hi + bstep * k4, idesc, 1);
                        }
                        mma_commit(b_empty + s);
                    }
                    __syncwarp();
                }
                if (elect_one()) mma_commit(a_empty + slot);
                __syncwarp();
            }
            if (elect_one()) {
                mma_commit(acc_full + p);
                if (tstamp) a.timing[10 + p] = global_ns();        // issue time of the phase's last MMA
            }
            __syncwarp();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) tmem_dealloc(tmem_base, 512);

    // ---- last CTA: d w3, d b3 and the loss from the CTA partials, in CTA order ----
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const unsigned t = atomicAdd(a.ticket, 1u);
        s_last = (t == gridDim.x - 1) ? 1u : 0u;
        if (s_last) *a.ticket = 0u;
    }
    __syncthreads();
    if (a.timing && tid == 0) atomicMax(a.timing + 15, (unsigned long long)global_ns());      // last CTA to finish its tile
    if (!s_last) return;
    __threadfence();
    const int W = a.H2 + 2;
    for (int c = tid; c < W; c += kThreads) {
        float s = 0.f;
        for (unsigned k = 0; k < gridDim.x; ++k) s += __ldcg(a.partial + (size_t)k * W + c);
        if (c == a.H2 + 1) { if (a.loss_sum) a.loss_sum[0] = s * a.loss_scale; }
        else if (a.sgd_out) {
            float* p = c < a.H2 ? a.w3 + c : a.b3;      // every CTA has finished reading w3 / b3: the last ticket is taken
            *p = fmaf(*p, a.decay_out, -a.lr * s);
        } else {
            if (c < a.H2) a.gw3[c] = s; else a.gb3[0] = s;
        }
    }
}

// One thread per element of every weight tile: phase p, stage (k-block c, column group h), tile-local (n, k).
//   K-major tiles (B2, B1: B[n][k] = W[n * ld + k])   byte offset = SWIZZLE_128B of row n: (n/8) 1024 + (n%8) 128 + ((k/4 ^ n%8) << 4) + (k%4) 4
//   MN-major tiles (F1, F2: B[k][n] = W[k * ld + n])  32-column boxes of 4096 B, row k at k 128, 32-byte chunks XOR k%4
//                                                     (SWIZZLE_128B_BASE32B: Swizzle<2,5,2> on the byte address)
// Elements outside the matrix are zeros.
struct PackPhase { const float* W; int ld, nK, nN, kmajor, nchunks, nh, hw; long long first, byte0; };
struct PackArgs { PackPhase ph[4]; long long total; uint8_t* out; };

__global__ void __launch_bounds__(256)
mlp2tc_pack_kernel(const PackArgs a) {
    pdl_sync();
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.total; i += (long long)gridDim.x * blockDim.x) {
        int p = 0;
        while (p < 3 && i >= a.ph[p + 1].first) ++p;
        const PackPhase ph = a.ph[p];
        const int per_stage = ph.hw * 32;
        const long long j = i - ph.first;
        const int stage = (int)(j / per_stage), e = (int)(j - (long long)stage * per_stage);
        const int c = stage / ph.nh, h = stage - c * ph.nh;
        int n, k;
        uint32_t off;
        if (ph.kmajor) {
            n = e >> 5; k = e & 31;
            off = (uint32_t)((n >> 3) * 1024 + (n & 7) * 128 + ((((k >> 2) ^ (n & 7))) << 4) + (k & 3) * 4);
        } else {
            k = e / ph.hw; n = e - k * ph.hw;
            const int box = n >> 5, nn = n & 31;
            off = (uint32_t)(box * 4096 + k * 128 + ((((nn >> 3) ^ (k & 3))) << 5) + (nn & 7) * 4);
        }
        const int gn = h * ph.hw + n, gk = 32 * c + k;
        float w = 0.f;
        if (gn < ph.nN && gk < ph.nK) w = ph.kmajor ? ph.W[(size_t)gn * ph.ld + gk] : ph.W[(size_t)gk * ph.ld + gn];
        float hi, lo;
        split_tf32(w, hi, lo);
        uint8_t* tile = a.out + ph.byte0 + (size_t)stage * (2 * ph.hw * 128);
        *reinterpret_cast<float*>(tile + off) = hi;
        *reinterpret_cast<float*>(tile + ph.hw * 128 + off) = lo;
    }
}

static int r32(int x) { return (x + 31) / 32 * 32; }
static void plan_phase(Phase& p, int K, int Nc, int acc_col, int kmajor) {
    p.nchunks = (K + 31) / 32;
    const int npad = r32(Nc);
    p.nh = npad > kMaxHalf ? 2 : 1;
    p.hw = p.nh == 2 ? r32((npad + 1) / 2) : npad;
    if (p.nh == 2 && Nc <= p.hw) p.nh = 1;
    for (int h = 0; h < 2; ++h) {
        int valid = Nc - h * p.hw;
        if (valid > p.hw) valid = p.hw;
        if (valid < 16) valid = 16;
        p.nmma[h] = (valid + 15) / 16 * 16;
    }
    p.acc_col = acc_col;
    p.kmajor = kmajor;
}




}  // namespace m2tc
}  // namespace dctr

using namespace dctr;

extern "C" int dctr_mlp2tc_supported(int D0, int H1, int H2) {
    return D0 >= 1 && H1 >= 1 && H2 >= 1 && D0 <= m2tc::kMaxD0 && H1 <= m2tc::kMaxH1 && H2 <= m2tc::kMaxH2;
}

extern "C" size_t dctr_mlp2tc_workspace(int N, int H2) {
    if (N < 0) N = 0;
    const size_t tiles = (size_t)(N + m2tc::kTileRows - 1) / m2tc::kTileRows;
    return 256 + tiles * (size_t)(H2 + 2) * sizeof(float);
}

static void plan_all(m2tc::Phase (&ph)[4], int D0, int H1, int H2) {
    m2tc::plan_phase(ph[0], D0, H1, 0, 0);
    m2tc::plan_phase(ph[1], H1, H2, ph[0].nh * ph[0].hw, 0);
    m2tc::plan_phase(ph[2], H2, H1, 0, 1);
    m2tc::plan_phase(ph[3], H1, D0, ph[2].nh * ph[2].hw, 1);
}

extern "C" size_t dctr_mlp2tc_packed_bytes(int D0, int H1, int H2) {
    if (!dctr_mlp2tc_supported(D0, H1, H2)) return 0;
    m2tc::Phase ph[4];
    plan_all(ph, D0, H1, H2);
    size_t b = 0;
    for (int p = 0; p < 4; ++p) b += (size_t)ph[p].nchunks * ph[p].nh * 2 * ph[p].hw * 128;
    return b;
}

extern "C" int dctr_mlp2tc_pack(int D0, int H1, int H2, const float* W1, int ldw1, const float* W2, int ldw2, void* packed,
                                size_t packed_bytes, void* stream) {
    DCTR_REQUIRE(dctr_mlp2tc_supported(D0, H1, H2), "mlp2tc_pack: unsupported widths %d-%d-%d", D0, H1, H2);
    DCTR_REQUIRE(W1 && W2 && packed, "mlp2tc_pack: null pointer");
    DCTR_REQUIRE(ldw1 >= H1 && ldw2 >= H2, "mlp2tc_pack: bad leading dimensions");
    DCTR_REQUIRE(((uintptr_t)packed & 127) == 0, "mlp2tc_pack: the packed buffer must be 128-byte aligned");
    DCTR_REQUIRE(packed_bytes >= dctr_mlp2tc_packed_bytes(D0, H1, H2), "mlp2tc_pack: packed buffer %zu < %zu bytes", packed_bytes,
                 dctr_mlp2tc_packed_bytes(D0, H1, H2));
    m2tc::Phase ph[4];
    plan_all(ph, D0, H1, H2);
    m2tc::PackArgs a{};
    // F1 / F2 read W as [k][n] (n contiguous), B2 / B1 read the SAME matrices as [n][k] (k contiguous)
    const float* Ws[4] = {W1, W2, W2, W1};
    const int lds[4] = {ldw1, ldw2, ldw2, ldw1}, nN[4] = {H1, H2, H1, D0}, nK[4] = {D0, H1, H2, H1};
    long long first = 0, byte0 = 0;
    for (int p = 0; p < 4; ++p) {
        m2tc::PackPhase& q = a.ph[p];
        q.W = Ws[p]; q.ld = lds[p]; q.nK = nK[p]; q.nN = nN[p]; q.kmajor = ph[p].kmajor; q.nchunks = ph[p].nchunks; q.nh = ph[p].nh; q.hw = ph[p].hw;
        q.first = first; q.byte0 = byte0;
        first += (long long)q.nchunks * q.nh * q.hw * 32;
        byte0 += (long long)q.nchunks * q.nh * 2 * q.hw * 128;
    }
    a.total = first;
    a.out = static_cast<uint8_t*>(packed);
    launch_k(m2tc::mlp2tc_pack_kernel, ceil_div(a.total, 256 * 2), 256, 0, as_stream(stream), a);
    return check_launch("mlp2tc_pack_kernel");
}

extern "C" int dctr_mlp2tc_train(int N, int D0, int H1, int H2, float* x, int ld0, const float* m0, int ldm0, int act0,
                                 const void* packed, const float* b1, int act1, const float* m1, int ldm1,
                                 const float* b2, int act2, const float* m2, int ldm2,
                                 float* w3, float* b3, const float* y, float gscale, float loss_scale,
                                 float* a1, int ld1, float* dz1, float* dz2, int ldd, float* dz0,
                                 float* logit, float* pred, float* loss_n, float* dlogit, float* loss_sum,
                                 float* gw3, float* gb3, int sgd, float lr, float decay_out,
                                 void* workspace, size_t workspace_bytes, void* stream) {
    using namespace m2tc;
    DCTR_REQUIRE(N >= 0, "mlp2tc_train: negative batch");
    DCTR_REQUIRE(dctr_mlp2tc_supported(D0, H1, H2), "mlp2tc_train: unsupported widths %d-%d-%d (limits %d-%d-%d)", D0, H1, H2, kMaxD0, kMaxH1, kMaxH2);
    if (N == 0) return DCTR_OK;
    DCTR_REQUIRE(x && packed && b1 && b2 && w3 && b3 && y && a1 && dz1 && dz2 && dz0 && workspace, "mlp2tc_train: null pointer");
    DCTR_REQUIRE(sgd || (gw3 && gb3), "mlp2tc_train: gw3 / gb3 are required unless sgd");
    DCTR_REQUIRE((ld0 & 3) == 0 && (ld1 & 3) == 0 && (ldd & 3) == 0 && ld0 >= D0 && ld1 >= H1 && ldd >= ((H1 + 3) & ~3) && ldd >= ((H2 + 3) & ~3) &&
                 (!m0 || (ldm0 & 3) == 0) && (!m1 || (ldm1 & 3) == 0) && (!m2 || (ldm2 & 3) == 0) &&
                 (!m0 || !ldm0 || ldm0 >= ld0) && (!m1 || !ldm1 || ldm1 >= ((H1 + 3) & ~3)) && (!m2 || !ldm2 || ldm2 >= ((H2 + 3) & ~3)),
                 "mlp2tc_train: leading dimensions must be multiples of 4 and cover the widths");
    DCTR_REQUIRE((((uintptr_t)x | (uintptr_t)packed | (uintptr_t)a1 | (uintptr_t)dz1 | (uintptr_t)dz2 | (uintptr_t)dz0 | (uintptr_t)workspace |
                   (uintptr_t)m0 | (uintptr_t)m1 | (uintptr_t)m2) & 15) == 0, "mlp2tc_train: buffers must be 16-byte aligned");
    DCTR_REQUIRE(workspace_bytes >= dctr_mlp2tc_workspace(N, H2), "mlp2tc_train: workspace %zu < %zu bytes", workspace_bytes, dctr_mlp2tc_workspace(N, H2));
    Args a{};
    a.N = N; a.D0 = D0; a.H1 = H1; a.H2 = H2; a.ld0 = ld0; a.ld1 = ld1; a.ldd = ldd;
    a.act0 = act0; a.act1 = act1; a.act2 = act2;
    a.x = x; a.m0 = m0; a.ldm0 = ldm0; a.b1 = b1; a.b2 = b2; a.w3 = w3; a.b3 = b3;
    a.m1 = m1; a.ldm1 = ldm1; a.m2 = m2; a.ldm2 = ldm2; a.y = y; a.gscale = gscale; a.loss_scale = loss_scale;
    a.a1 = a1; a.dz2 = dz2; a.dz1 = dz1; a.dz0 = dz0;
    a.logit = logit; a.pred = pred; a.loss_n = loss_n; a.dlogit = dlogit; a.loss_sum = loss_sum;
    a.gw3 = gw3; a.gb3 = gb3; a.sgd_out = sgd; a.lr = lr; a.decay_out = decay_out;
    a.ticket = static_cast<unsigned*>(workspace);
    a.partial = reinterpret_cast<float*>(static_cast<uint8_t*>(workspace) + 256);
    plan_all(a.ph, D0, H1, H2);
    DCTR_REQUIRE(a.ph[1].acc_col + a.ph[1].nh * a.ph[1].hw <= 512 && a.ph[3].acc_col + a.ph[3].nh * a.ph[3].hw <= 512 && a.ph[1].nh == 1,
                 "mlp2tc_train: accumulators do not fit tensor memory");
    a.packed = static_cast<const uint8_t*>(packed);
    static unsigned long long* timing = nullptr;
    static const bool want_timing = getenv("DCTR_M2TC_TIMING") != nullptr;
    if (want_timing && !timing) cudaMalloc(&timing, 32 * sizeof(unsigned long long));      // diagnostic runs only
    a.timing = want_timing ? timing : nullptr;
    if (a.timing) { cudaMemsetAsync(timing + 14, 0xFF, 8, as_stream(stream)); cudaMemsetAsync(timing + 15, 0, 8, as_stream(stream)); }
    cudaError_t e = cudaFuncSetAttribute(mlp2tc_train_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
    if (e != cudaSuccess) { set_error("mlp2tc_train: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return DCTR_ECUDA; }
    // programmatic dependent launch: barrier set-up, the TMEM allocation and the zero fill of the A ring overlap the tail of the
    // kernel before (64 CTAs leave the other SMs to the concurrent streams, unlike the per-layer GEMM's full waves)
    launch_k(mlp2tc_train_kernel, dim3((N + kTileRows - 1) / kTileRows), kThreads, kSmem, as_stream(stream), a);
    const int rc = check_launch("mlp2tc_train_kernel");
    if (want_timing && timing && rc == DCTR_OK && cudaStreamSynchronize(as_stream(stream)) == cudaSuccess) {
        // workers (thread 0 of CTA 0): start | x chunks written | z1 complete | a1 chunks written | z2 complete | dz2 chunks written |
        // da1 complete | dz1 chunks written | dz0 complete | end; MMA warp: last MMA of F1 / F2 / B2 / B1 issued.  ns since the start.
        unsigned long long t[32];
        cudaMemcpy(t, timing, sizeof(t), cudaMemcpyDeviceToHost);
        fprintf(stderr, "mlp2tc timing (ns, CTA 0):");
        for (int i = 0; i < 14; ++i) fprintf(stderr, " %lld", (long long)(t[i] - t[0]));
        fprintf(stderr, " | first CTA start %lld, last CTA end %lld\n", (long long)(t[14] - t[0]), (long long)(t[15] - t[0]));
    }
    return rc;
}

