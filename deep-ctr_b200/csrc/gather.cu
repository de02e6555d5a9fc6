// K1 / K2 / K3: per-field embedding gather + concat with the inner- / outer-product layer fused in,
// and the backward of that layer into a per-(sample, field) gradient buffer.
//
// Rows are fetched with 128-bit loads (row stride ld is a multiple of 4 floats, so a row is ld/4 LDG.128), scaled
// by the field weight and staged in shared memory; the concat / IP / OP outputs are produced from the staged
// block and written with warp-coalesced 128-bit stores.  HBM-bound: no tensor cores here by design.
#include <string.h>
#include "common.cuh"

namespace dctr {

constexpr int kGatherWarps = 8;

// stage e[f][0:ld) = w_f * table[id_f][0:ld) for one sample into this warp's smem block
__device__ __forceinline__ void stage_field_block(const int32_t* __restrict__ ids, const float* __restrict__ wts,
                                                  const float* __restrict__ table, int ld, int F, int64_t n,
                                                  int lane, float* e, int X, int32_t* __restrict__ status) {
    const int Q = ld >> 2;
    const int T = F * Q;
    for (int t = lane; t < T; t += 32) {
        const int f = t / Q;
        const int q = t - f * Q;
        const int64_t p = n * F + f;
        const int64_t id = checked_id(ids[p], X, status);
        const float w = wts ? wts[p] : 1.f;
        float4 v = ldg_f4(table + id * ld + q * 4);
        v.x *= w; v.y *= w; v.z *= w; v.w *= w;
        *reinterpret_cast<float4*>(e + f * ld + q * 4) = v;
    }
    __syncwarp();
}

// Forward: a warp owns 8 consecutive samples at a time.  A 4-lane group gathers one sample -- lane q reads float4
// q of each of the F rows (F independent 128-bit loads in flight per lane) -- and scatters the scaled values
// into the warp's shared-memory image of the 8 output rows, already in concat order.  The inner- / outer-product
// columns are then formed from that image by all 32 lanes, and the image (8 * ldz contiguous floats, also
// contiguous in global memory) leaves with coalesced 128-bit stores, the input activation applied on the way.
constexpr int kSamplesPerWarp = 8;

template <int CROSS>
__global__ void __launch_bounds__(kGatherWarps * 32)
gather_concat_kernel(const int32_t* __restrict__ ids, const float* __restrict__ wts,
                     const float* __restrict__ table, int ld, int N, int F, int d, int bias_first,
                     const float* __restrict__ bias, int act0, float* __restrict__ out, int ldz, int warps, int vec,
                     int X, int32_t* __restrict__ status) {
    pdl_sync();
    extern __shared__ __align__(16) float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = lane & 3, g = lane >> 2;
    const int Fd = F * d;
    const int P = F * (F - 1) / 2;
    const int crossw = CROSS == DCTR_CROSS_IP ? P : (CROSS == DCTR_CROSS_OP ? d * d : 0);
    const int D0 = Fd + crossw + 1;
    const int eoff = bias_first ? 1 : 0;
    const int Q = ld >> 2, ldz4 = ldz >> 2;
    float* zs = smem + (size_t)warp * kSamplesPerWarp * ldz;
    // pair table (i<<8 | j), i-outer / j-inner: the order of FNN_IP_L3.py:91-94
    unsigned short* pairs = reinterpret_cast<unsigned short*>(smem + (size_t)warps * kSamplesPerWarp * ldz);
    if (CROSS == DCTR_CROSS_IP) {
        for (int i = threadIdx.x; i < F - 1; i += blockDim.x) {
            int base = i * F - i * (i + 1) / 2 - i - 1;
            for (int j = i + 1; j < F; ++j) pairs[base + j] = (unsigned short)((i << 8) | j);
        }
        __syncthreads();
    }
    const float bval = bias ? __ldg(bias) : 0.f;

    for (int64_t n0 = ((int64_t)blockIdx.x * warps + warp) * kSamplesPerWarp; n0 < N;
         n0 += (int64_t)gridDim.x * warps * kSamplesPerWarp) {
        const int cnt = (int)min((int64_t)kSamplesPerWarp, (int64_t)N - n0);
        float* zr = zs + g * ldz;
        if (g < cnt) {
            const int64_t pbase = (n0 + g) * F;
            for (int qq = q; qq < Q; qq += 4) {
                const int c0 = qq * 4;
                for (int f0 = 0; f0 < F; f0 += 8) {
                    int id[8]; float w[8]; float4 v[8];
                    load_fields8(ids, wts, pbase, f0, F, vec, id, w, X, status);
#pragma unroll
                    for (int k = 0; k < 8; ++k) v[k] = ldg_f4_issue(table + (int64_t)id[k] * ld + c0);
#pragma unroll
                    for (int k = 7; k >= 0; --k) {         // last-issued load first (common.cuh)
                        if (f0 + k < F) {
                            float* dst = zr + eoff + (f0 + k) * d + c0;
                            if (c0 < d) dst[0] = v[k].x * w[k];
                            if (c0 + 1 < d) dst[1] = v[k].y * w[k];
                            if (c0 + 2 < d) dst[2] = v[k].z * w[k];
                            if (c0 + 3 < d) dst[3] = v[k].w * w[k];
                        }
                    }
                }
            }
            if (q == 3) {
                zr[bias_first ? 0 : Fd + crossw] = bval;
                for (int c = D0; c < ldz; ++c) zr[c] = 0.f;
            }
        }
        __syncwarp();
        if (CROSS == DCTR_CROSS_IP) {
            for (int s = 0; s < cnt; ++s) {
                const float* e = zs + s * ldz;
                for (int t = lane; t < P; t += 32) {
                    const int ij = pairs[t];
                    const float* ei = e + (ij >> 8) * d;
                    const float* ej = e + (ij & 255) * d;
                    float acc = 0.f;
                    for (int c = 0; c < d; ++c) acc = fmaf(ei[c], ej[c], acc);
                    zs[s * ldz + Fd + t] = acc;
                }
            }
            __syncwarp();
        } else if (CROSS == DCTR_CROSS_OP) {
            // q[a,b] = sum_j prefix_{<j}[a] * e_j[b]   (== sum_{i<j} e_i[a] e_j[b], FNN_OP_L3.py:93-97)
            for (int s = 0; s < cnt; ++s) {
                const float* e = zs + s * ldz;
                for (int t = lane; t < crossw; t += 32) {
                    const int a = t / d, b = t - a * d;
                    float acc = 0.f, pre = 0.f;
                    for (int j = 0; j < F; ++j) {
                        acc = fmaf(pre, e[j * d + b], acc);
                        pre += e[j * d + a];
                    }
                    zs[s * ldz + Fd + t] = acc;
                }
            }
            __syncwarp();
        }
        // the image is cnt * ldz contiguous floats both here and in global memory
        float4* dst = reinterpret_cast<float4*>(out + n0 * (int64_t)ldz);
        const float4* src = reinterpret_cast<const float4*>(zs);
        const int total4 = cnt * ldz4;
        if (act0 == DCTR_ACT_NONE) {
            for (int i = lane; i < total4; i += 32) dst[i] = src[i];
        } else {
            int col4 = lane;                   // float4 column inside the row, tracked incrementally
            while (col4 >= ldz4) col4 -= ldz4;
            for (int i = lane; i < total4; i += 32) {
                float4 v = src[i];
                const int c = col4 * 4;
                v.x = c < D0 ? act_fwd(act0, v.x) : 0.f;
                v.y = c + 1 < D0 ? act_fwd(act0, v.y) : 0.f;
                v.z = c + 2 < D0 ? act_fwd(act0, v.z) : 0.f;
                v.w = c + 3 < D0 ? act_fwd(act0, v.w) : 0.f;
                dst[i] = v;
                col4 += 32;
                while (col4 >= ldz4) col4 -= ldz4;
            }
        }
        __syncwarp();
    }
}

// first element of the output row of slot `slot` (see cross_bwd_kernel); ldg is applied by the caller through the row pointer
struct OutRow {
    float* p;
    __device__ __forceinline__ float& operator[](int64_t i) const { return p[i]; }
};
__device__ __forceinline__ OutRow out_row_impl(float* ge, int ldg, const int32_t* dest, float* const* peers, int64_t C, int rank, int64_t slot) {
    if (!dest) return OutRow{ge + slot * ldg};
    const int64_t dd = dest[slot];
    if (!peers) return OutRow{ge + dd * ldg};
    const int64_t o = dd / C;
    return OutRow{peers[o] + ((int64_t)rank * C + (dd - o * C)) * ldg};
}
#define out_row(ge, dest, peers, C, rank, slot) out_row_impl(ge, ldg, dest, peers, C, rank, slot)

template <int CROSS>
__global__ void __launch_bounds__(kGatherWarps * 32)
cross_bwd_kernel(const int32_t* __restrict__ ids, const float* __restrict__ wts, const float* __restrict__ table,
                 int ld, int N, int F, int d, const float* __restrict__ dz, int ldz,
                 float* __restrict__ ge, int ldg, int X, int32_t* __restrict__ status, const int32_t* __restrict__ dest,
                 float* const* __restrict__ peers, int64_t C, int rank) {
    // dest != NULL: slot (n, f) writes row dest[n * F + f] of ge instead of row n * F + f (the bucket order of the row-sharded exchange);
    // peers != NULL: bucket o = dest / C lives on rank o -- the row goes to peers[o] at row rank * C + dest % C (a store over NVLink)
    pdl_sync();
    extern __shared__ __align__(16) float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int Fd = F * d;
    const int P = F * (F - 1) / 2;
    const int crossw = CROSS == DCTR_CROSS_IP ? P : (CROSS == DCTR_CROSS_OP ? d * d : 0);
    const int per_warp = (F * ld * (CROSS == DCTR_CROSS_OP ? 2 : 1) + crossw + (CROSS == DCTR_CROSS_OP ? d : 0) + 3) & ~3;  // keep float4 alignment
    float* e = smem + warp * per_warp;
    float* pre = e + F * ld;                                        // OP only: prefix_{<f}[a]
    float* dc = e + F * ld * (CROSS == DCTR_CROSS_OP ? 2 : 1);      // dL/dcross
    float* tot = dc + crossw;                                       // OP only: sum_f e_f

    // CROSS_NONE leaves the warp's shared-memory block unused: with a destination index it stages the sample's F rows there
    const bool vec_rows = dest != nullptr && ldg <= ld && (ldg & 3) == 0;
    for (int64_t n = (int64_t)blockIdx.x * kGatherWarps + warp; n < N; n += (int64_t)gridDim.x * kGatherWarps) {
        const float* dzr = dz + n * (int64_t)ldz;
        if (CROSS != DCTR_CROSS_NONE) {
            for (int t = lane; t < crossw; t += 32) dc[t] = dzr[Fd + t];
            stage_field_block(ids, wts, table, ld, F, n, lane, e, X, status);   // ends with __syncwarp
        }
        if (CROSS == DCTR_CROSS_OP) {
            for (int c = lane; c < d; c += 32) {
                float run = 0.f;
                for (int j = 0; j < F; ++j) { pre[j * ld + c] = run; run += e[j * ld + c]; }
                tot[c] = run;
            }
            __syncwarp();
        }
        for (int t = lane; t < Fd; t += 32) {
            const int f = t / d;
            const int c = t - f * d;
            float acc = dzr[t];
            if (CROSS == DCTR_CROSS_IP) {
                // d e_f[c] += sum_{j != f} dp[pi(min,max)] * e_j[c]
                for (int j = 0; j < f; ++j) acc = fmaf(dc[j * F - j * (j + 1) / 2 + f - j - 1], e[j * ld + c], acc);
                const int base = f * F - f * (f + 1) / 2 - f - 1;
                for (int j = f + 1; j < F; ++j) acc = fmaf(dc[base + j], e[j * ld + c], acc);
            } else if (CROSS == DCTR_CROSS_OP) {
                // as the j of a pair: sum_a dq[a,c] * prefix_{<f}[a]; as the i: sum_b dq[c,b] * suffix_{>f}[b]
                for (int a = 0; a < d; ++a) {
                    const float suf = tot[a] - pre[f * ld + a] - e[f * ld + a];
                    acc = fmaf(dc[a * d + c], pre[f * ld + a], acc);
                    acc = fmaf(dc[c * d + a], suf, acc);
                }
            }
            const float w = wts ? wts[n * F + f] : 1.f;
            if (CROSS == DCTR_CROSS_NONE && vec_rows) e[f * ldg + c] = w * acc;       // staged: the rows leave as 128-bit stores below
            else out_row(ge, dest, peers, C, rank, n * F + f)[(int64_t)0 * ldg + c] = w * acc;
        }
        if (CROSS == DCTR_CROSS_NONE && vec_rows) {
            // rows written in a caller-given order (bucket order, possibly into peer memory over NVLink): whole rows as 16-byte
            // stores instead of one 4-byte store per lane and element
            for (int t = lane; t < F * (ldg - d); t += 32) { const int f = t / (ldg - d); e[f * ldg + d + (t - f * (ldg - d))] = 0.f; }
            __syncwarp();
            const int q = ldg >> 2;
            for (int t = lane; t < F * q; t += 32) {
                const int f = t / q, j = t - f * q;
                reinterpret_cast<float4*>(out_row(ge, dest, peers, C, rank, n * F + f).p)[j] = reinterpret_cast<const float4*>(e + f * ldg)[j];
            }
        } else if (ldg > d) {
            const int padw = ldg - d;
            for (int t = lane; t < F * padw; t += 32) {
                const int f = t / padw;
                out_row(ge, dest, peers, C, rank, n * F + f)[d + (t - f * padw)] = 0.f;
            }
        }
        __syncwarp();
    }
}

// K0: slot of a feature = its field (feats_to_layer_one_array, FNN_wnzh.py:91-95).  One warp per sample: lanes fill the
// row with (pad_row, 0), then lane j places feature j; when several features of one sample share a field the highest j
// wins (__match_any_sync picks it), which is the reference loop's "later feature overwrites".
// upd_rows / upd_pos (optional, [N, J]): the UPDATE list of the sample -- the reference's scatter loops over every feature
// of the sample, the overwritten ones included, and gives each the gradient of its field's slot (FNN_wnzh.py:300-306):
// entry j = (feature id, n * F + field); absent / unknown features become (pad_row, n * F), which the fold skips.
__global__ void __launch_bounds__(256)
arrange_fields_kernel(const int32_t* __restrict__ feats, const int32_t* __restrict__ feat_field, int N, int J, int F,
                      int32_t X, int32_t pad_row, int32_t* __restrict__ ids, float* __restrict__ wts,
                      int32_t* __restrict__ upd_rows, int32_t* __restrict__ upd_pos, int32_t* __restrict__ status) {
    pdl_sync();
    const int n = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (n >= N) return;
    int32_t* irow = ids + (int64_t)n * F;
    float* wrow = wts + (int64_t)n * F;
    for (int f = lane; f < F; f += 32) { irow[f] = pad_row; wrow[f] = 0.f; }
    const int32_t* frow = feats + (int64_t)n * J;
    for (int j0 = 0; j0 < J; j0 += 32) {          // chunks in list order, so a later chunk overwrites an earlier one
        __syncwarp();
        const int j = j0 + lane;
        const int32_t feat = j < J ? frow[j] : -1;
        int slot = -1 - lane;                     // invalid lanes: private keys
        if (feat >= 0) {
            if (feat < X) {
                const int s = __ldg(feat_field + feat);
                if (s >= 0 && s < F) slot = s;
            } else if (status) {                  // a feature id the FM model does not have (KeyError in the reference)
                atomicAdd(status, 1);
                status[1] = feat;
            }
        }
        const unsigned peers = __match_any_sync(0xffffffffu, slot);
        if (slot >= 0 && (31 - __clz(peers)) == lane) { irow[slot] = feat; wrow[slot] = 1.f; }
        if (upd_rows && j < J) {
            upd_rows[(int64_t)n * J + j] = slot >= 0 ? feat : pad_row;
            upd_pos[(int64_t)n * J + j] = n * F + (slot >= 0 ? slot : 0);
        }
    }
}

// Row gather out[i, :] = table[ids[i], :] -- the owner-side lookup of the row-sharded table and the
// permutation of gradient rows into exchange order.  One thread per 128-bit piece of a row.
__global__ void __launch_bounds__(256)
gather_rows_kernel(const int32_t* __restrict__ ids, const float4* __restrict__ table, int q, int64_t M, float4* __restrict__ out,
                   int X, int32_t* __restrict__ status) {
    pdl_sync();
    const int64_t total = M * q;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = t / q;
        const int c = (int)(t - i * q);
        out[t] = __ldg(table + (int64_t)checked_id(__ldg(ids + i), X, c == 0 ? status : nullptr) * q + c);
    }
}

static int gather_grid(int N) {
    int blocks = ceil_div(N, kGatherWarps);
    int cap = kNumSMs * 8;  // 8 resident CTAs of 8 warps = 64 warps / SM
    return blocks < cap ? (blocks > 0 ? blocks : 1) : cap;
}

}  // namespace dctr

using namespace dctr;

extern "C" int dctr_gather_concat_fwd(const int32_t* ids, const float* wts, const float* table, int64_t X, int ld,
                                      int N, int F, int d, int cross, int bias_first, const float* bias, int act0,
                                      float* out, int ldz, void* stream) {
    DCTR_REQUIRE(ids && table && out, "gather_concat_fwd: null pointer");
    DCTR_REQUIRE(N >= 0 && F > 0 && F < 256 && d > 0 && d <= ld && (ld & 3) == 0, "gather_concat_fwd: bad shape N=%d F=%d d=%d ld=%d", N, F, d, ld);
    DCTR_REQUIRE(cross >= 0 && cross <= 2 && !(cross && bias_first), "gather_concat_fwd: bad cross/bias_first");
    const int crossw = cross == DCTR_CROSS_IP ? F * (F - 1) / 2 : (cross == DCTR_CROSS_OP ? d * d : 0);
    DCTR_REQUIRE(ldz >= F * d + crossw + 1, "gather_concat_fwd: ldz=%d < D0=%d", ldz, F * d + crossw + 1);
    DCTR_REQUIRE(X > 0 && X < ((int64_t)1 << 31) && X * (int64_t)ld < ((int64_t)1 << 40), "gather_concat_fwd: bad X");
    if (N == 0) return DCTR_OK;
    DCTR_REQUIRE((ldz & 3) == 0 && ((uintptr_t)out & 15) == 0 && ((uintptr_t)table & 15) == 0,
                 "gather_concat_fwd: out / table must be 16-byte aligned and ldz a multiple of 4");
    // shared memory: one image of 8 output rows per warp; fewer warps per CTA when the rows are wide
    const size_t per_warp = (size_t)kSamplesPerWarp * ldz * sizeof(float);
    const size_t extra = cross == DCTR_CROSS_IP ? (size_t)crossw * 2 + 16 : 0;
    int warps = kGatherWarps;
    while (warps > 1 && warps * per_warp + extra > 96 * 1024) warps >>= 1;
    const size_t smem = warps * per_warp + extra;
    DCTR_REQUIRE(smem <= 200 * 1024, "gather_concat_fwd: output row too wide for shared memory (ldz=%d)", ldz);
    int blocks = ceil_div(N, warps * kSamplesPerWarp);
    const int cap = kNumSMs * (smem > 48 * 1024 ? 2 : 4);
    if (blocks > cap) blocks = cap;
    dim3 grid(blocks), block(warps * 32);
    cudaStream_t st = as_stream(stream);
    const int vec = (F & 3) == 0 && ((uintptr_t)ids & 15) == 0 && (!wts || ((uintptr_t)wts & 15) == 0);
#define LAUNCH_G(C)                                                                                          \
    do {                                                                                                     \
        if (smem > 48 * 1024)                                                                                \
            cudaFuncSetAttribute(gather_concat_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        launch_k(gather_concat_kernel<C>, grid, block, smem, st, ids, wts, table, ld, N, F, d, bias_first, bias,   \
                                                           act0, out, ldz, warps, vec, (int)X, id_status()); \
    } while (0)
    if (cross == DCTR_CROSS_IP) LAUNCH_G(DCTR_CROSS_IP);
    else if (cross == DCTR_CROSS_OP) LAUNCH_G(DCTR_CROSS_OP);
    else LAUNCH_G(DCTR_CROSS_NONE);
#undef LAUNCH_G
    return check_launch("gather_concat_kernel");
}

static int cross_bwd_impl(const int32_t* ids, const float* wts, const float* table, int64_t X, int ld,
                          int N, int F, int d, int cross, const float* dz, int ldz,
                          float* ge, int ldg, const int32_t* dest, float* const* peers, int64_t C, int rank, void* stream) {
    DCTR_REQUIRE(ids && table && dz && ge, "cross_bwd: null pointer");
    DCTR_REQUIRE(N >= 0 && F > 0 && F < 256 && d > 0 && d <= ld && (ld & 3) == 0 && ldg >= d, "cross_bwd: bad shape");
    DCTR_REQUIRE(cross >= 0 && cross <= 2 && X > 0, "cross_bwd: bad cross");
    if (N == 0) return DCTR_OK;
    const int crossw = cross == DCTR_CROSS_IP ? F * (F - 1) / 2 : (cross == DCTR_CROSS_OP ? d * d : 0);
    const int per_warp = (F * ld * (cross == DCTR_CROSS_OP ? 2 : 1) + crossw + (cross == DCTR_CROSS_OP ? d : 0) + 3) & ~3;
    size_t smem = (size_t)kGatherWarps * per_warp * sizeof(float);
    DCTR_REQUIRE(smem <= 200 * 1024, "cross_bwd: field block too large for shared memory");
    dim3 grid(gather_grid(N)), block(kGatherWarps * 32);
    cudaStream_t st = as_stream(stream);
#define LAUNCH_B(CR)                                                                                          \
    do {                                                                                                     \
        if (smem > 48 * 1024)                                                                                \
            cudaFuncSetAttribute(cross_bwd_kernel<CR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
        launch_k(cross_bwd_kernel<CR>, grid, block, smem, st, ids, wts, table, ld, N, F, d, dz, ldz, ge, ldg, (int)X, id_status(), dest, peers, C, rank); \
    } while (0)
    if (cross == DCTR_CROSS_IP) LAUNCH_B(DCTR_CROSS_IP);
    else if (cross == DCTR_CROSS_OP) LAUNCH_B(DCTR_CROSS_OP);
    else LAUNCH_B(DCTR_CROSS_NONE);
#undef LAUNCH_B
    return check_launch("cross_bwd_kernel");
}

extern "C" int dctr_cross_bwd(const int32_t* ids, const float* wts, const float* table, int64_t X, int ld,
                              int N, int F, int d, int cross, const float* dz, int ldz,
                              float* ge, int ldg, void* stream) {
    return cross_bwd_impl(ids, wts, table, X, ld, N, F, d, cross, dz, ldz, ge, ldg, nullptr, nullptr, 1, 0, stream);
}

extern "C" int dctr_cross_bwd_to(const int32_t* ids, const float* wts, const float* table, int64_t X, int ld,
                                 int N, int F, int d, int cross, const float* dz, int ldz,
                                 float* ge, int ldg, const int32_t* dest, void* stream) {
    DCTR_REQUIRE(dest, "cross_bwd_to: null destination index");
    return cross_bwd_impl(ids, wts, table, X, ld, N, F, d, cross, dz, ldz, ge, ldg, dest, nullptr, 1, 0, stream);
}

extern "C" int dctr_cross_bwd_push(const int32_t* ids, const float* wts, const float* table, int64_t X, int ld,
                                   int N, int F, int d, int cross, const float* dz, int ldz, int ldg, const int32_t* dest,
                                   float* const* peers, int64_t C, int rank, void* stream) {
    DCTR_REQUIRE(dest && peers && C >= 1 && rank >= 0, "cross_bwd_push: null destination index / peer table");
    return cross_bwd_impl(ids, wts, table, X, ld, N, F, d, cross, dz, ldz, const_cast<float*>(dz), ldg, dest, peers, C, rank, stream);
}

extern "C" int dctr_arrange_fields(const int32_t* feats, const int32_t* feat_field, int N, int J, int F, int32_t pad_row,
                                   int32_t* ids, float* wts, void* stream) {
    return dctr_arrange_update_list(feats, feat_field, N, J, F, (int64_t)1 << 30, pad_row, ids, wts, nullptr, nullptr, stream);
}

extern "C" int dctr_arrange_update_list(const int32_t* feats, const int32_t* feat_field, int N, int J, int F, int64_t X,
                                        int32_t pad_row, int32_t* ids, float* wts, int32_t* upd_rows, int32_t* upd_pos,
                                        void* stream) {
    DCTR_REQUIRE(N >= 0 && J > 0 && F > 0 && X > 0 && X <= ((int64_t)1 << 30) && (int64_t)N * F < ((int64_t)1 << 31),
                 "arrange_fields: bad shape N=%d J=%d F=%d", N, J, F);
    if (N == 0) return DCTR_OK;
    DCTR_REQUIRE(feats && feat_field && ids && wts && (!upd_rows == !upd_pos), "arrange_fields: null pointer");
    launch_k(arrange_fields_kernel, ceil_div((int64_t)N * 32, 256), 256, 0, as_stream(stream), feats, feat_field, N, J, F, (int32_t)X,
             pad_row, ids, wts, upd_rows, upd_pos, id_status());
    return check_launch("arrange_fields_kernel");
}

// ---- fixed-capacity exchange plan of the row-sharded table ---------------------------------------------------------------------------
// The slots of a batch, already ordered by owner rank (perm / owner_sorted from the stable sort on the owner key), go into `world`
// buckets of C entries each: slot perm[i] of owner o is entry i - first(o).  Entries past a bucket's fill keep the sentinel -1
// (send_ids) / 0 (src_of_dest); a bucket that would overflow drops the slot and counts it in status[0] (status[1] = the largest
// fill seen) -- the caller sizes C from a calibration step and treats status[0] != 0 as an error.
namespace dctr {
__global__ void __launch_bounds__(256)
shard_pack_kernel(const int32_t* __restrict__ ids, const int32_t* __restrict__ perm, const int32_t* __restrict__ owner_sorted, int64_t M,
                  int world, int64_t C, int32_t* __restrict__ send_ids, int32_t* __restrict__ src_of_dest, int32_t* __restrict__ inv,
                  int32_t* __restrict__ status, int rank, int32_t* const* __restrict__ peers) {
    pdl_sync();
    __shared__ int64_t s_first[65];
    if ((int)threadIdx.x <= world) {            // first position of every owner in the sorted list (lower bound), [world] = M
        const int o = threadIdx.x;
        int64_t lo = 0, hi = M;
        while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (owner_sorted[mid] < o) lo = mid + 1; else hi = mid; }
        s_first[o] = lo;
    }
    __syncthreads();
    if (blockIdx.x == 0 && (int)threadIdx.x < world) atomicMax(status + 1, (int)(s_first[threadIdx.x + 1] - s_first[threadIdx.x]));
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < M; i += (int64_t)gridDim.x * blockDim.x) {
        const int o = owner_sorted[i], slot = perm[i];
        const int64_t k = i - s_first[o];
        if (k < C) {
            const int64_t dest = (int64_t)o * C + k;
            send_ids[dest] = ids[slot] / world;
            src_of_dest[dest] = slot;
            inv[slot] = (int32_t)dest;
        } else {
            inv[slot] = (int32_t)((int64_t)o * C);      // a valid position; the step is reported as failed
            atomicAdd(status, 1);
        }
    }
    if (!peers) return;
    // delivery of the id buckets over peer memory: bucket o -- its unused tail (-1) included -- is block `rank` of rank o's
    // request list, walked in destination order so that the remote stores are contiguous
    const int64_t total = (int64_t)world * C;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int o = (int)(t / C);
        const int64_t k = t - (int64_t)o * C, i = s_first[o] + k;
        peers[o][(int64_t)rank * C + k] = i < s_first[o + 1] ? ids[perm[i]] / world : -1;
    }
}
}  // namespace dctr

static int shard_pack_impl(const int32_t* ids, const int32_t* perm, const int32_t* owner_sorted, int64_t M, int world, int64_t C,
                           int32_t* send_ids, int32_t* src_of_dest, int32_t* inv, int32_t* status, int rank, int32_t* const* peers, void* stream) {
    DCTR_REQUIRE(M >= 0 && world >= 1 && world <= 64 && C >= 1 && (int64_t)world * C < ((int64_t)1 << 31), "shard_pack: bad shape");
    DCTR_REQUIRE(send_ids && src_of_dest && status && (M == 0 || (ids && perm && owner_sorted && inv)), "shard_pack: null pointer");
    cudaStream_t st = as_stream(stream);
    cudaError_t e = cudaMemsetAsync(send_ids, 0xFF, (size_t)world * C * sizeof(int32_t), st);
    if (e == cudaSuccess) e = cudaMemsetAsync(src_of_dest, 0, (size_t)world * C * sizeof(int32_t), st);
    if (e != cudaSuccess) { set_error("shard_pack: cudaMemsetAsync: %s", cudaGetErrorString(e)); return DCTR_ECUDA; }
    if (M == 0 && !peers) return DCTR_OK;
    int blocks = ceil_div(M > (int64_t)world * C ? M : (int64_t)world * C, 256);
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    launch_k(shard_pack_kernel, blocks, 256, 0, st, ids, perm, owner_sorted, M, world, C, send_ids, src_of_dest, inv, status, rank, peers);
    return check_launch("shard_pack_kernel");
}

extern "C" int dctr_shard_pack(const int32_t* ids, const int32_t* perm, const int32_t* owner_sorted, int64_t M, int world, int64_t C,
                               int32_t* send_ids, int32_t* src_of_dest, int32_t* inv, int32_t* status, void* stream) {
    return shard_pack_impl(ids, perm, owner_sorted, M, world, C, send_ids, src_of_dest, inv, status, 0, nullptr, stream);
}

extern "C" int dctr_shard_pack_push(const int32_t* ids, const int32_t* perm, const int32_t* owner_sorted, int64_t M, int world, int64_t C,
                                    int32_t* send_ids, int32_t* src_of_dest, int32_t* inv, int32_t* status, int rank, int32_t* const* peers,
                                    void* stream) {
    DCTR_REQUIRE(peers && rank >= 0 && rank < world, "shard_pack_push: null peer table / bad rank");
    return shard_pack_impl(ids, perm, owner_sorted, M, world, C, send_ids, src_of_dest, inv, status, rank, peers, stream);
}

// ---- owner-side lookup that delivers over peer memory: entry k of bucket s (the ids rank s asked this rank for) goes to row
// rank * C + k of rank s's row buffer (peers[s]); unused entries (-1) are not transmitted.  One kernel instead of gather + all-to-all.
namespace dctr {
__global__ void __launch_bounds__(256)
gather_rows_push_kernel(const int32_t* __restrict__ ids, const float4* __restrict__ table, int q, int64_t C, int world, int rank,
                        float* const* __restrict__ peers, int X) {
    pdl_sync();
    const int64_t total = (int64_t)world * C * q;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = t / q;
        const int c = (int)(t - i * q);
        const int id = __ldg(ids + i);
        if ((unsigned)id >= (unsigned)X) continue;
        const int64_t s = i / C, k = i - s * C;
        reinterpret_cast<float4*>(peers[s])[((int64_t)rank * C + k) * q + c] = __ldg(table + (int64_t)id * q + c);
    }
}
}  // namespace dctr

extern "C" int dctr_gather_rows_push(const int32_t* ids, const float* table, int64_t X, int ld, int64_t C, int world, int rank,
                                     float* const* peers, void* stream) {
    DCTR_REQUIRE(C >= 1 && world >= 1 && rank >= 0 && rank < world && X > 0 && ld > 0 && (ld & 3) == 0, "gather_rows_push: bad shape");
    DCTR_REQUIRE(ids && table && peers && ((uintptr_t)table & 15) == 0, "gather_rows_push: null or unaligned pointer");
    int blocks = ceil_div((int64_t)world * C * (ld / 4), 256);
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    launch_k(gather_rows_push_kernel, blocks, 256, 0, as_stream(stream), ids, reinterpret_cast<const float4*>(table), ld / 4, C, world, rank,
             peers, (int)X);
    return check_launch("gather_rows_push_kernel");
}

// ---- per-slot gradients of a model WITHOUT a cross layer, delivered in BUCKET order: entry i of the owner buckets (slot
// src_of_dest[i], valid where send_ids[i] >= 0) is wts * dz[n, f*d .. f*d+d) and goes to row rank * C + i % C of rank i / C's gradient
// buffer.  Walking the destination order makes every remote store stream contiguous rows (the sample-order kernel, dctr_cross_bwd_push,
// scatters 48-byte rows over the owners: 360 GB/s at 8 GPUs); the scattered side is the local, cheap one.
namespace dctr {
__global__ void __launch_bounds__(256)
slot_grads_push_kernel(const int32_t* __restrict__ send_ids, const int32_t* __restrict__ src_of_dest, const float* __restrict__ wts,
                       const float* __restrict__ dz, int64_t ldz, int F, int d, int q4, int64_t C, int world, int rank,
                       float* const* __restrict__ peers) {
    pdl_sync();
    const int64_t total = (int64_t)world * C * q4;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = t / q4;
        const int c = 4 * (int)(t - i * q4);
        if (__ldg(send_ids + i) < 0) continue;
        const int slot = __ldg(src_of_dest + i);
        const int n = slot / F, f = slot - n * F;
        const float w = wts ? __ldg(wts + slot) : 1.f;
        const float* src = dz + (int64_t)n * ldz + (int64_t)f * d + c;
        float4 v;
        v.x = c < d ? w * __ldg(src) : 0.f;
        v.y = c + 1 < d ? w * __ldg(src + 1) : 0.f;
        v.z = c + 2 < d ? w * __ldg(src + 2) : 0.f;
        v.w = c + 3 < d ? w * __ldg(src + 3) : 0.f;
        const int64_t o = i / C, k = i - o * C;
        reinterpret_cast<float4*>(peers[o])[((int64_t)rank * C + k) * q4 + (c >> 2)] = v;
    }
}
}  // namespace dctr

extern "C" int dctr_slot_grads_push(const int32_t* send_ids, const int32_t* src_of_dest, const float* wts, const float* dz, int64_t ldz,
                                    int F, int d, int ldg, int64_t C, int world, int rank, float* const* peers, void* stream) {
    DCTR_REQUIRE(send_ids && src_of_dest && dz && peers, "slot_grads_push: null pointer");
    DCTR_REQUIRE(F > 0 && d > 0 && ldg >= d && (ldg & 3) == 0 && C >= 1 && world >= 1 && rank >= 0 && rank < world, "slot_grads_push: bad shape");
    int blocks = ceil_div((int64_t)world * C * (ldg / 4), 256);
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    launch_k(slot_grads_push_kernel, blocks, 256, 0, as_stream(stream), send_ids, src_of_dest, wts, dz, ldz, F, d, ldg / 4, C, world, rank, peers);
    return check_launch("slot_grads_push_kernel");
}

// ---- device buffers other processes of the node can map (CUDA IPC): the exchange buffers of the row-sharded step ------------------
extern "C" int dctr_ipc_alloc(size_t bytes, void** ptr, unsigned char* handle64) {
    DCTR_REQUIRE(ptr && handle64 && bytes > 0, "ipc_alloc: bad arguments");
    cudaError_t e = cudaMalloc(ptr, bytes);
    if (e == cudaSuccess) e = cudaMemset(*ptr, 0, bytes);
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, *ptr);
    if (e != cudaSuccess) { set_error("ipc_alloc: %s", cudaGetErrorString(e)); return DCTR_ECUDA; }
    static_assert(sizeof(h) == 64, "CUDA IPC handle size");
    memcpy(handle64, &h, 64);
    return DCTR_OK;
}
extern "C" int dctr_ipc_open(const unsigned char* handle64, void** ptr) {
    DCTR_REQUIRE(ptr && handle64, "ipc_open: bad arguments");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    cudaError_t e = cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { set_error("ipc_open: %s", cudaGetErrorString(e)); (void)cudaGetLastError(); return DCTR_ECUDA; }
    return DCTR_OK;
}
extern "C" int dctr_ipc_close(void* ptr) {
    cudaError_t e = cudaIpcCloseMemHandle(ptr);
    if (e != cudaSuccess) { set_error("ipc_close: %s", cudaGetErrorString(e)); return DCTR_ECUDA; }
    return DCTR_OK;
}
extern "C" int dctr_ipc_free(void* ptr) {
    cudaError_t e = cudaFree(ptr);
    if (e != cudaSuccess) { set_error("ipc_free: %s", cudaGetErrorString(e)); return DCTR_ECUDA; }
    return DCTR_OK;
}

extern "C" int dctr_gather_rows(const int32_t* ids, const float* table, int64_t X, int ld, int64_t M, float* out, void* stream) {
    DCTR_REQUIRE(M >= 0 && X > 0 && ld > 0 && (ld & 3) == 0, "gather_rows: bad shape M=%lld ld=%d", (long long)M, ld);
    if (M == 0) return DCTR_OK;
    DCTR_REQUIRE(ids && table && out && ((uintptr_t)table & 15) == 0 && ((uintptr_t)out & 15) == 0, "gather_rows: null or unaligned pointer");
    int blocks = ceil_div(M * (ld / 4), 256);
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    launch_k(gather_rows_kernel, blocks, 256, 0, as_stream(stream), ids, reinterpret_cast<const float4*>(table), ld / 4, M,
                                                              reinterpret_cast<float4*>(out), (int)X, id_status());
    return check_launch("gather_rows_kernel");
}
