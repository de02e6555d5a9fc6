// K4-tc: tensor-core GEMM for the MLP -- tcgen05.mma kind::tf32 with 3xTF32 error compensation, TMA-staged operands,
// TMEM accumulators.  Same contract as dctr_gemm (gemm.cu): NN forward / NT dgrad / TN wgrad (split-K slabs), fused epilogues.
//
//   C = A*B with A = A_hi + A_lo, B = B_hi + B_lo (hi = rna-tf32(x), lo = rna-tf32(x - hi)):
//       acc += A_lo*B_hi + A_hi*B_lo + A_hi*B_hi        (fp32 accumulate in TMEM; the dropped A_lo*B_lo term is ~2^-22)
//
//   CTA = 10 warps, one 128 x BN output tile, BK = 32 fp32 (one 128-byte row), NS-stage mbarrier pipeline:
//     warp 9     one lane issues the TMA loads (cp.async.bulk.tensor, out-of-bounds zero fill) of the raw fp32 operand
//                boxes.  B goes straight into its UMMA layout: K-contiguous as one [BN x 32] SWIZZLE_128B box, MN-contiguous
//                as BN/32 boxes of [32 k x 32 n] (SWIZZLE_128B_ATOM_32B, the only MN-major fp32 UMMA layout).  A is only
//                ever read by the converter warps, so it lands in whatever layout they read without bank conflicts;
//     warps 0-7  split every landed stage: B raw -> hi in place and lo into a twin buffer (fence.proxy.async); A raw ->
//                hi / lo written to TENSOR MEMORY (tcgen05.st, lane = row, column = k), so the MMAs take A from TMEM and
//                only B crosses the shared-memory port a second time -- the port, not the tensor pipe, bounds this
//                kernel.  After the last k-block the same warps run the epilogue: tcgen05.ld (lane =
//                output row) -> bias / act / mask or dgrad -> 128-bit stores;
//     warp 8     allocates TMEM; one lane issues 4 k-steps x 3 tcgen05.mma (A in TMEM, B by descriptor) per stage;
//                tcgen05.commit frees the shared-memory stage for the TMA lane and the TMEM A slot for the converters,
//                and finally signals the epilogue.
#include "tc_common.cuh"

namespace dctr {
namespace tc {

constexpr int kBM = 128;
constexpr int kBKf = 32;                 // fp32 per k-block = 128 bytes
constexpr int kConverters = 256;          // 8 warps (16 measured 4% slower: the split is not what bounds the kernel)
constexpr int kThreads = kConverters + 64;
constexpr int kMmaWarp = kConverters / 32, kTmaWarp = kMmaWarp + 1;



struct TcParams {
    int M, N, K;
    const float* A; int lda;
    const float* B; int ldb;
    float* C; int ldc;
    int epilogue, act;
    const float* bias;
    const float* aux; int ldaux;
    const float* mask; int ldmask;
    int k_per_split;   // multiple of kBKf
    int direct_epilogue;   // small outputs (L2-resident): store straight from the accumulator registers, lane = row
};

// Epilogue of one 32-column chunk of one output row, specialised at compile time on (epilogue, activation): the
// runtime dispatch happens once per chunk, not once per element (a per-element switch inlined 128 times made the
// kernel 41k instructions long and the epilogue instruction-cache bound: 14 us -> measured with %globaltimer).
template <int EPI, int ACT>
__device__ __forceinline__ float tc_epi1(const TcParams& p, int m, int n, float acc, float auxv) {
    if (EPI == DCTR_EPI_FWD) {
        float v = act_fwd(ACT, acc + (p.bias ? __ldg(p.bias + n) : 0.f));
        if (p.mask) v *= __ldg(p.mask + (int64_t)m * p.ldmask + n);
        return v;
    } else if (EPI == DCTR_EPI_DGRAD) {
        float mk = 1.f, t = auxv;
        if (p.mask) {
            mk = __ldg(p.mask + (int64_t)m * p.ldmask + n);
            if (mk == 0.f) return 0.f;
            t *= __frcp_rn(mk);
        }
        return acc * mk * act_bwd_from_out(ACT, t);
    }
    return acc;
}
// One 32 x 32 chunk: the accumulators arrive with lane = row (TMEM), global memory wants lane = column.  The chunk goes
// through a padded [32][33] shared-memory tile (conflict-free both ways); afterwards every warp instruction loads aux /
// mask / bias and stores C along ONE row: 128 contiguous bytes instead of 32 rows x 4 bytes (the row-per-lane form cost
// ~13 us per tile in the dgrad epilogue, more than a K = 300 main loop).
// the dgrad epilogue's second operand (the stored activation) for rows mbase..mbase+31 of column n: 32 independent loads,
// issued before the accumulators are even complete so their DRAM latency hides behind the main loop's tail
__device__ __forceinline__ void tc_prefetch_aux(const TcParams& p, int mbase, int n, float (&ax)[32]) {
    const bool on = p.epilogue == DCTR_EPI_DGRAD && n < p.N;
    const float* src = p.aux + (int64_t)mbase * p.ldaux + n;
#pragma unroll
    for (int r = 0; r < 32; ++r) ax[r] = (on && mbase + r < p.M) ? __ldg(src + (int64_t)r * p.ldaux) : 0.f;
}
template <int EPI, int ACT>
__device__ __forceinline__ void tc_store_chunk(const TcParams& p, float* __restrict__ C, float* tile, int mbase, int nbase, int lane,
                                               const float (&v)[32], const float (&ax)[32]) {
#pragma unroll
    for (int j = 0; j < 32; ++j) tile[lane * 33 + j] = v[j];
    __syncwarp();
    const int n = nbase + lane;
    if (n < p.N) {
        if (EPI == DCTR_EPI_DGRAD) {        // fully unrolled: ax[] must stay in registers
#pragma unroll
            for (int r = 0; r < 32; ++r)
                if (mbase + r < p.M) C[(int64_t)(mbase + r) * p.ldc + n] = tc_epi1<EPI, ACT>(p, mbase + r, n, tile[r * 33 + lane], ax[r]);
        } else {                            // 32 inlined activations would only bloat the instruction stream
            const int rows = min(32, p.M - mbase);
#pragma unroll 8
            for (int r = 0; r < rows; ++r)
                C[(int64_t)(mbase + r) * p.ldc + n] = tc_epi1<EPI, ACT>(p, mbase + r, n, tile[r * 33 + lane], 0.f);
        }
    }
    __syncwarp();
}
// Small problems (C, aux and mask stay in L2, a single wave of CTAs): the transposition costs more than the scattered
// 16-byte accesses it avoids, so every lane stores its own row directly (measured on the N = 4096 step: 8 us per step).
template <int EPI, int ACT>
__device__ __forceinline__ void tc_store_rows(const TcParams& p, float* __restrict__ C, int m, int nbase, const float (&v)[32]) {
    if (m >= p.M) return;
    float* dst = C + (int64_t)m * p.ldc + nbase;
    const int nlast = p.N - 1;
    // dgrad: this row's 32 stored activations, fetched up front as 128-bit loads where the row allows it (a scalar load
    // per element costs 32 wavefronts per warp instruction here, lanes being rows)
    float ax[32];
    if (EPI == DCTR_EPI_DGRAD) {
        const float* arow = p.aux + (int64_t)m * p.ldaux;
        const bool vec = ((p.ldaux & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.aux) & 15) == 0);
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
            const int n = nbase + j;
            if (vec && n + 3 <= nlast) {
                const float4 t = __ldg(reinterpret_cast<const float4*>(arow + n));
                ax[j] = t.x; ax[j + 1] = t.y; ax[j + 2] = t.z; ax[j + 3] = t.w;
            } else {
#pragma unroll
                for (int c = 0; c < 4; ++c) ax[j + c] = __ldg(arow + min(n + c, nlast));
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
        const int n = nbase + j;
        float o[4];      // columns past N reuse the last column's bias / mask / aux (loads stay in range) and are not stored
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int nn = min(n + c, nlast);
            o[c] = tc_epi1<EPI, ACT>(p, m, nn, v[j + c], EPI == DCTR_EPI_DGRAD ? ax[j + c] : 0.f);
        }
        if (n + 3 < p.N) *reinterpret_cast<float4*>(dst + j) = make_float4(o[0], o[1], o[2], o[3]);
        else {
#pragma unroll
            for (int c = 0; c < 3; ++c)
                if (n + c < p.N) dst[j + c] = o[c];
        }
    }
}
__device__ __forceinline__ void tc_store_rows_dispatch(const TcParams& p, float* C, int m, int nbase, const float (&v)[32]) {
    switch (p.epilogue == DCTR_EPI_NONE ? 0 : p.epilogue * 4 + p.act) {
        case DCTR_EPI_FWD * 4 + DCTR_ACT_NONE: tc_store_rows<DCTR_EPI_FWD, DCTR_ACT_NONE>(p, C, m, nbase, v); break;
        case DCTR_EPI_FWD * 4 + DCTR_ACT_TANH: tc_store_rows<DCTR_EPI_FWD, DCTR_ACT_TANH>(p, C, m, nbase, v); break;
        case DCTR_EPI_FWD * 4 + DCTR_ACT_RELU: tc_store_rows<DCTR_EPI_FWD, DCTR_ACT_RELU>(p, C, m, nbase, v); break;
        case DCTR_EPI_FWD * 4 + DCTR_ACT_SIGMOID: tc_store_rows<DCTR_EPI_FWD, DCTR_ACT_SIGMOID>(p, C, m, nbase, v); break;
        case DCTR_EPI_DGRAD * 4 + DCTR_ACT_NONE: tc_store_rows<DCTR_EPI_DGRAD, DCTR_ACT_NONE>(p, C, m, nbase, v); break;
        case DCTR_EPI_DGRAD * 4 + DCTR_ACT_TANH: tc_store_rows<DCTR_EPI_DGRAD, DCTR_ACT_TANH>(p, C, m, nbase, v); break;
        case DCTR_EPI_DGRAD * 4 + DCTR_ACT_RELU: tc_store_rows<DCTR_EPI_DGRAD, DCTR_ACT_RELU>(p, C, m, nbase, v); break;
        case DCTR_EPI_DGRAD * 4 + DCTR_ACT_SIGMOID: tc_store_rows<DCTR_EPI_DGRAD, DCTR_ACT_SIGMOID>(p, C, m, nbase, v); break;
        default: tc_store_rows<DCTR_EPI_NONE, DCTR_ACT_NONE>(p, C, m, nbase, v); break;
    }
}
__device__ __forceinline__ void tc_store_dispatch(const TcParams& p, float* C, float* tile, int mbase, int nbase, int lane, const float (&v)[32],
                                                  const float (&ax)[32]) {
    switch (p.epilogue == DCTR_EPI_NONE ? 0 : p.epilogue * 4 + p.act) {
        case DCTR_EPI_FWD * 4 + DCTR_ACT_NONE: tc_store_chunk<DCTR_EPI_FWD, DCTR_ACT_NONE>(p, C, tile, mbase, nbase, lane, v, ax); break;
        case DCTR_EPI_FWD * 4 + DCTR_ACT_TANH: tc_store_chunk<DCTR_EPI_FWD, DCTR_ACT_TANH>(p, C, tile, mbase, nbase, lane, v, ax); break;
        case DCTR_EPI_FWD * 4 + DCTR_ACT_RELU: tc_store_chunk<DCTR_EPI_FWD, DCTR_ACT_RELU>(p, C, tile, mbase, nbase, lane, v, ax); break;
        case DCTR_EPI_FWD * 4 + DCTR_ACT_SIGMOID: tc_store_chunk<DCTR_EPI_FWD, DCTR_ACT_SIGMOID>(p, C, tile, mbase, nbase, lane, v, ax); break;
        case DCTR_EPI_DGRAD * 4 + DCTR_ACT_NONE: tc_store_chunk<DCTR_EPI_DGRAD, DCTR_ACT_NONE>(p, C, tile, mbase, nbase, lane, v, ax); break;
        case DCTR_EPI_DGRAD * 4 + DCTR_ACT_TANH: tc_store_chunk<DCTR_EPI_DGRAD, DCTR_ACT_TANH>(p, C, tile, mbase, nbase, lane, v, ax); break;
        case DCTR_EPI_DGRAD * 4 + DCTR_ACT_RELU: tc_store_chunk<DCTR_EPI_DGRAD, DCTR_ACT_RELU>(p, C, tile, mbase, nbase, lane, v, ax); break;
        case DCTR_EPI_DGRAD * 4 + DCTR_ACT_SIGMOID: tc_store_chunk<DCTR_EPI_DGRAD, DCTR_ACT_SIGMOID>(p, C, tile, mbase, nbase, lane, v, ax); break;
        default: tc_store_chunk<DCTR_EPI_NONE, DCTR_ACT_NONE>(p, C, tile, mbase, nbase, lane, v, ax); break;
    }
}

template <int BN, int NS>
struct TcSmem {
    static constexpr int kA = kBM * 128;           // bytes of one [128 x 32] fp32 tile
    static constexpr int kB = BN * 128;
    static constexpr int kStage = kA + 2 * kB;      // A raw | B raw->hi | B lo
    static constexpr int kBytes = NS * kStage + 1024 /* alignment slack */ + 256 /* barriers */;
};

// TMEM columns: kHiAcc accumulators for hi*hi + one for the cross terms, then a ring of A slots (hi 32 | lo 32 columns)
constexpr int kHiAcc = 2;
template <int BN>
struct TcTmem {
    static constexpr int kAcc = (kHiAcc + 1) * BN;
    static constexpr int kSlots = BN <= 64 ? 4 : 2;
    static constexpr int kCols = 512;
    static_assert(kAcc + kSlots * 64 <= kCols, "TMEM budget");
};

// raw fp32 -> hi in place, lo into the twin buffer (same swizzled layout): 16-byte vectors, conflict-free
template <int BYTES>
__device__ __forceinline__ void split_in_place(uint8_t* hi, uint8_t* lo, int t) {
    constexpr int NV = BYTES / 16 / kConverters;
    float4 x[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) x[i] = *reinterpret_cast<const float4*>(hi + (t + i * kConverters) * 16);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        float4 h, l;
        split_tf32(x[i].x, h.x, l.x); split_tf32(x[i].y, h.y, l.y);
        split_tf32(x[i].z, h.z, l.z); split_tf32(x[i].w, h.w, l.w);
        *reinterpret_cast<float4*>(hi + (t + i * kConverters) * 16) = h;
        *reinterpret_cast<float4*>(lo + (t + i * kConverters) * 16) = l;
    }
}

// A tile -> tensor memory.  Thread = (row quadrant q = warp % 4, lane = row in the quadrant, k part h = warp / 4): reads its
// kAVals raw values from shared memory, splits them, writes hi and lo to its TMEM lane.
//   AK  (K-contiguous A, one SWIZZLE_128B box): row r at r*128 B, 16-byte chunk c at position c ^ (r % 8)
//   !AK (MN-contiguous A, four plain [32 k x 32 m] boxes): element (k, m) of quadrant q at q*4096 + k*128 + m*4
constexpr int kAVals = kBKf / (kConverters / 128);      // 16 with 8 converter warps
template <bool AK>
__device__ __forceinline__ void a_to_tmem(const uint8_t* a_raw, uint32_t slot_addr, int warp, int lane) {
    const int q = warp & 3, h = warp >> 2;
    float x[kAVals];
    if (AK) {
        const int r = q * 32 + lane;
        const uint8_t* row = a_raw + (r >> 3) * 1024 + (r & 7) * 128;
#pragma unroll
        for (int c = 0; c < kAVals / 4; ++c) {
            const float4 v = *reinterpret_cast<const float4*>(row + ((((kAVals / 4) * h + c) ^ (r & 7)) << 4));
            x[4 * c] = v.x; x[4 * c + 1] = v.y; x[4 * c + 2] = v.z; x[4 * c + 3] = v.w;
        }
    } else {
        const uint8_t* col = a_raw + q * 4096 + lane * 4 + (kAVals * h) * 128;
#pragma unroll
        for (int j = 0; j < kAVals; ++j) x[j] = *reinterpret_cast<const float*>(col + j * 128);
    }
    const uint32_t ta = slot_addr + ((uint32_t)(q * 32) << 16) + (uint32_t)(kAVals * h);
#pragma unroll
    for (int g = 0; g < kAVals / 8; ++g) {
        float hi[8], lo[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) split_tf32(x[8 * g + j], hi[j], lo[j]);
        tmem_st8(ta + 8 * g, hi);
        tmem_st8(ta + 32 + 8 * g, lo);
    }
}

template <int BN, int NS, bool AK, bool BK>
__global__ void __launch_bounds__(kThreads, 1)
gemm_tc_kernel(const TcParams p, const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB) {
    using S = TcSmem<BN, NS>;
    using T = TcTmem<BN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);     // pointer arithmetic keeps the shared address space (LDS / STS, not generic LD / ST)
    uint64_t* raw_full = reinterpret_cast<uint64_t*>(smem + NS * S::kStage);
    uint64_t* conv_full = raw_full + NS;
    uint64_t* empty = conv_full + NS;
    uint64_t* a_empty = empty + NS;
    uint64_t* accum_full = a_empty + T::kSlots;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_full + 1);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m0 = blockIdx.y * kBM, n0 = blockIdx.x * BN;
    const int kbeg = blockIdx.z * p.k_per_split;
    const int kend = min(p.K, kbeg + p.k_per_split);
    const int nkb = kend > kbeg ? (kend - kbeg + kBKf - 1) / kBKf : 0;
    float* C = p.C + (int64_t)blockIdx.z * p.M * p.ldc;

    if (tid == 0) {
        for (int s = 0; s < NS; ++s) { mbar_init(raw_full + s, 1); mbar_init(conv_full + s, kConverters); mbar_init(empty + s, 1); }
        for (int t = 0; t < T::kSlots; ++t) mbar_init(a_empty + t, 1);
        mbar_init(accum_full, 1);
        fence_mbar_init();
    }
    if (warp == kMmaWarp) tmem_alloc(tmem_slot, T::kCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tmem_a0 = tmem_base + T::kAcc;
    // barriers initialised, tensor memory allocated: only now wait for the kernel that produced the operands (programmatic
    // dependent launch: everything above overlapped its tail)
    pdl_sync();

    if (warp < kMmaWarp) {
        // ===== converters =====
        for (int kb = 0; kb < nkb; ++kb) {
            const int s = kb % NS, t = kb % T::kSlots;
            mbar_wait(raw_full + s, (kb / NS) & 1);
            uint8_t* st = smem + s * S::kStage;
            split_in_place<S::kB>(st + S::kA, st + S::kA + S::kB, tid);
            fence_proxy_async();          // generic-proxy stores -> visible to the tensor core's async proxy
            mbar_wait(a_empty + t, ((kb / T::kSlots) & 1) ^ 1);      // the MMAs that read this TMEM slot have completed
            tc_fence_after();
            a_to_tmem<AK>(st, tmem_a0 + t * 64, warp, lane);
            tmem_st_wait();
            tc_fence_before();
            mbar_arrive(conv_full + s);
        }
        // ===== epilogue: TMEM lane = output row, global memory lane = output column =====
        // a warp reaches the TMEM lanes 32*(warp%4)..+31 only: warps w, w+4, ... share a row quadrant and split the columns
        const int quad = warp & 3;
        const int mbase = m0 + quad * 32;
        constexpr int kCStep = (kConverters / 128) * 32;
        float ax[32];
        if (!p.direct_epilogue) tc_prefetch_aux(p, mbase, n0 + (warp >> 2) * 32 + lane, ax);
        if (nkb > 0) {
            mbar_wait(accum_full, 0);
            tc_fence_after();
        }
        // every pipeline stage is drained by now (accum_full = all MMAs complete): its first bytes become the transpose tiles
        float* tile = reinterpret_cast<float*>(smem) + warp * (32 * 33);
#pragma unroll 1
        for (int c = (warp >> 2) * 32; c < BN; c += kCStep) {
            if (n0 + c >= p.N) break;                   // uniform across the warp
            float v[32];
            if (nkb > 0) {
                const uint32_t ta = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)c;
                tmem_ld32(ta, v);
#pragma unroll
                for (int a = 1; a <= kHiAcc; ++a) {
                    float u[32];
                    tmem_ld32(ta + a * BN, u);
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] += u[j];
                }
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = 0.f;
            }
            if (p.direct_epilogue) {
                tc_store_rows_dispatch(p, C, mbase + lane, n0 + c, v);
                continue;
            }
            if (mbase < p.M) tc_store_dispatch(p, C, tile, mbase, n0 + c, lane, v, ax);
            if (c + kCStep < BN) tc_prefetch_aux(p, mbase, n0 + c + kCStep + lane, ax);
        }
    } else if (warp == kTmaWarp) {
        // ===== TMA producer ===== (all lanes run the loop, one elected lane issues -- see the MMA issuer below)
        if (elect_one()) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&mapA)) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&mapB)) : "memory");
        }
        __syncwarp();
        for (int kb = 0; kb < nkb; ++kb) {
            const int s = kb % NS;
            mbar_wait<64>(empty + s, ((kb / NS) & 1) ^ 1);
            uint8_t* st = smem + s * S::kStage;
            const int k0 = kbeg + kb * kBKf;
            if (elect_one()) {
                mbar_expect_tx(raw_full + s, S::kA + S::kB);
                if (AK) tma_load_2d(st, &mapA, k0, m0, raw_full + s);
                else {
#pragma unroll
                    for (int i = 0; i < kBM / 32; ++i) tma_load_2d(st + i * 4096, &mapA, m0 + 32 * i, k0, raw_full + s);
                }
                uint8_t* sb = st + S::kA;
                if (BK) tma_load_2d(sb, &mapB, k0, n0, raw_full + s);
                else {
#pragma unroll
                    for (int i = 0; i < BN / 32; ++i) tma_load_2d(sb + i * 4096, &mapB, n0 + 32 * i, k0, raw_full + s);
                }
            }
            __syncwarp();
        }
    } else if (warp == kMmaWarp) {
        // ===== MMA issuer =====
        // All 32 lanes run the loop and the waits, one elected lane issues: inside `if (lane == 0)` ptxas cannot prove the
        // descriptors warp-uniform and wraps EVERY tcgen05.mma in an ELECT / R2UR.BROADCAST / BRA.U.ANY loop (~70 ns per MMA,
        // measured in mlp2tc.cu -- more than the 34 ns a 128 x 128 x 8 tf32 MMA takes).
        // instruction descriptor (cute::UMMA::InstrDescriptor): D = F32 (1 << 4), A = B = TF32 (2 << 7, 2 << 10),
        // a_major bit 15 = 0 (A comes from TMEM: always K-major) / b_major bit 16 (0 = K-major, 1 = MN-major),
        // N >> 3 at bit 17, M >> 4 at bit 24
        constexpr uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((BK ? 0u : 1u) << 16) |
                                   ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(kBM >> 4) << 24);
        for (int kb = 0; kb < nkb; ++kb) {
            const int s = kb % NS, ts = kb % T::kSlots;
            mbar_wait<32>(conv_full + s, (kb / NS) & 1);
            tc_fence_after();
            const uint32_t sb = smem_u32(smem + s * S::kStage) + S::kA;
            const uint32_t ta = tmem_a0 + ts * 64;
            const uint64_t b_hi0 = BK ? make_desc(sb) : make_desc_mn(sb), b_lo0 = BK ? make_desc(sb + S::kB) : make_desc_mn(sb + S::kB);
            constexpr uint64_t bstep = BK ? 2 : (1024 >> 4);     // B: +32 B inside a K-major row, +1024 B (8 k-rows = 2 atoms) MN-major
            if (elect_one()) {
#pragma unroll
                for (int k4 = 0; k4 < kBKf / 8; ++k4) {           // UMMA_K = 8 tf32: 8 TMEM columns of A
                    const uint32_t a_hi = ta + k4 * 8, a_lo = ta + 32 + k4 * 8;
                    const uint64_t b_hi = b_hi0 + bstep * k4, b_lo = b_lo0 + bstep * k4;
                    // Several accumulators shorten every addition chain (the tensor core truncates when it accumulates): the
                    // hi*hi products of k-step t go to accumulator t % kHiAcc, the two small cross terms to the last one;
                    // the epilogue adds them up in fp32.
                    const int t = kb * (kBKf / 8) + k4;
                    mma_tf32_ts(tmem_base + kHiAcc * BN, a_lo, b_hi, idesc, t != 0);
                    mma_tf32_ts(tmem_base + kHiAcc * BN, a_hi, b_lo, idesc, 1);
                    mma_tf32_ts(tmem_base + (t % kHiAcc) * BN, a_hi, b_hi, idesc, t >= kHiAcc);
                }
                mma_commit(empty + s);        // frees the shared-memory stage when these MMAs have read it ...
                mma_commit(a_empty + ts);     // ... and the TMEM A slot
            }
            __syncwarp();
        }
        if (nkb > 0 && elect_one()) mma_commit(accum_full);
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) tmem_dealloc(tmem_base, T::kCols);
}


template <int BN, int NS>
static int launch_tc(int layout, const TcParams& p, dim3 grid, cudaStream_t st) {
    constexpr int bytes = TcSmem<BN, NS>::kBytes;
    const bool ak = layout != DCTR_GEMM_TN, bk = layout == DCTR_GEMM_NT;
    CUtensorMap mapA, mapB;
    int rc = make_map(&mapA, p.A, p.lda, p.M, p.K, ak, kBM, /*plain_mn=*/true);    // A is read by the converter warps only
    if (rc) return rc;
    if ((rc = make_map(&mapB, p.B, p.ldb, p.N, p.K, bk, BN))) return rc;
    cudaError_t e = cudaSuccess;
#define DCTR_TC_LAUNCH(AKV, BKV)                                                                                      \
    do {                                                                                                              \
        e = cudaFuncSetAttribute(gemm_tc_kernel<BN, NS, AKV, BKV>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes); \
        if (e == cudaSuccess) launch_k_heavy(gemm_tc_kernel<BN, NS, AKV, BKV>, grid, kThreads, bytes, st, p, mapA, mapB);          \
    } while (0)
    if (layout == DCTR_GEMM_NN) DCTR_TC_LAUNCH(true, false);
    else if (layout == DCTR_GEMM_NT) DCTR_TC_LAUNCH(true, true);
    else DCTR_TC_LAUNCH(false, false);
#undef DCTR_TC_LAUNCH
    if (e != cudaSuccess) {
        set_error("gemm_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return DCTR_ECUDA;
    }
    return check_launch("gemm_tc_kernel");
}

}  // namespace tc
}  // namespace dctr

using namespace dctr;

extern "C" int dctr_gemm_tc(int layout, int M, int N, int K, const float* A, int lda, const float* B, int ldb,
                            float* C, int ldc, int epilogue, int act, const float* bias, const float* aux, int ldaux,
                            const float* mask, int ldmask, int splits, void* stream) {
    DCTR_REQUIRE(layout >= 0 && layout <= 2, "gemm_tc: bad layout %d", layout);
    DCTR_REQUIRE(M >= 0 && N >= 0 && K >= 0, "gemm_tc: negative shape");
    if (M == 0 || N == 0) return DCTR_OK;
    DCTR_REQUIRE(A && B && C, "gemm_tc: null pointer");
    DCTR_REQUIRE((lda & 3) == 0 && (ldb & 3) == 0 && (ldc & 3) == 0, "gemm_tc: leading dimensions must be multiples of 4");
    DCTR_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0 && ((uintptr_t)C & 15) == 0, "gemm_tc: operands must be 16-byte aligned");
    DCTR_REQUIRE(epilogue >= 0 && epilogue <= 2 && (epilogue != DCTR_EPI_DGRAD || aux), "gemm_tc: bad epilogue");
    if (splits < 1) splits = 1;
    DCTR_REQUIRE(splits == 1 || epilogue == DCTR_EPI_NONE, "gemm_tc: split-K needs EPI_NONE");
    tc::TcParams p{M, N, K, A, lda, B, ldb, C, ldc, epilogue, act, bias, aux, ldaux, mask, ldmask, 0, 0};
    int kper = (K + splits - 1) / splits;
    kper = (kper + tc::kBKf - 1) / tc::kBKf * tc::kBKf;
    if (kper == 0) kper = tc::kBKf;
    p.k_per_split = kper;
    // outputs (+ the dgrad operand) of up to 16 MB stay in the 126 MB L2 next to everything else of a small step
    p.direct_epilogue = (int64_t)splits * M * N * 4 <= ((int64_t)16 << 20);
    cudaStream_t st = as_stream(stream);
    const int mt = ceil_div(M, tc::kBM);
    DCTR_REQUIRE(mt < 65536, "gemm_tc: M too large");
    if (K == 0) { p.K = 0; }
    // narrow N tiles when a 128-wide tiling would leave most SMs idle
    if (N <= 64 || mt * ceil_div(N, 128) * splits < kNumSMs / 2) {
        dim3 grid(ceil_div(N, 64), mt, splits);
        return tc::launch_tc<64, 6>(layout, p, grid, st);
    }
    dim3 grid(ceil_div(N, 128), mt, splits);
    return tc::launch_tc<128, 4>(layout, p, grid, st);
}
