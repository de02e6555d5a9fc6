// K4: fp32 SIMT GEMM with the MLP epilogues fused (bias + activation + dropout mask forward,
// mask * act' backward).  This is the PARITY path (fp32 FFMA, 1e-5 vs the fp64 oracle); the tensor-core
// path (tcgen05 kind::tf32, 3xTF32 split) lives in gemm_tc.cu and is checked against this one.
//   CTA = 256 threads, BMxBNx16 tile, (BM/16)x(BN/16) micro-tile per thread, register-prefetched
//   double-buffered shared memory, 128-bit global and shared accesses, zero-filled edges.
#include "common.cuh"

namespace dctr {

constexpr int kBK = 16;
constexpr int kGemmThreads = 256;

struct GemmParams {
    int M, N, K;
    const float* A; int lda;
    const float* B; int ldb;
    float* C; int ldc;
    int epilogue, act;
    const float* bias;
    const float* aux; int ldaux;
    const float* mask; int ldmask;
    int k_per_split;   // multiple of kBK
};

template <int BT, bool KCONTIG>
struct TileLoader {
    static constexpr int NV = BT * kBK / 4 / kGemmThreads;
    static constexpr int S = BT + 4;
    float4 r[NV];
    // operand element (mn, k): KCONTIG -> base[mn*ld + k], else base[k*ld + mn]
    __device__ __forceinline__ void load(const float* __restrict__ base, int ld, int mn0, int k0, int MN, int Kend) {
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int idx = threadIdx.x + v * kGemmThreads;
            float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
            if (KCONTIG) {
                const int mn = mn0 + (idx >> 2), k = k0 + (idx & 3) * 4;
                if (mn < MN) {
                    const float* src = base + (int64_t)mn * ld + k;
                    if (k + 3 < Kend) x = __ldg(reinterpret_cast<const float4*>(src));
                    else {
                        if (k < Kend) x.x = __ldg(src);
                        if (k + 1 < Kend) x.y = __ldg(src + 1);
                        if (k + 2 < Kend) x.z = __ldg(src + 2);
                    }
                }
            } else {
                const int k = k0 + idx / (BT / 4), mn = mn0 + (idx % (BT / 4)) * 4;
                if (k < Kend) {
                    const float* src = base + (int64_t)k * ld + mn;
                    if (mn + 3 < MN) x = __ldg(reinterpret_cast<const float4*>(src));
                    else {
                        if (mn < MN) x.x = __ldg(src);
                        if (mn + 1 < MN) x.y = __ldg(src + 1);
                        if (mn + 2 < MN) x.z = __ldg(src + 2);
                    }
                }
            }
            r[v] = x;
        }
    }
    __device__ __forceinline__ void store(float* __restrict__ sm) const {   // sm: [kBK][S]
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const int idx = threadIdx.x + v * kGemmThreads;
            if (KCONTIG) {
                const int mn = idx >> 2, k = (idx & 3) * 4;
                sm[(k + 0) * S + mn] = r[v].x;
                sm[(k + 1) * S + mn] = r[v].y;
                sm[(k + 2) * S + mn] = r[v].z;
                sm[(k + 3) * S + mn] = r[v].w;
            } else {
                const int k = idx / (BT / 4), mn = (idx % (BT / 4)) * 4;
                *reinterpret_cast<float4*>(sm + k * S + mn) = r[v];
            }
        }
    }
};

__device__ __forceinline__ float epilogue_apply(const GemmParams& p, int m, int n, float acc) {
    if (p.epilogue == DCTR_EPI_FWD) {
        float v = act_fwd(p.act, acc + (p.bias ? __ldg(p.bias + n) : 0.f));
        if (p.mask) v *= __ldg(p.mask + (int64_t)m * p.ldmask + n);
        return v;
    } else if (p.epilogue == DCTR_EPI_DGRAD) {
        const float mk = p.mask ? __ldg(p.mask + (int64_t)m * p.ldmask + n) : 1.f;
        if (mk == 0.f) return 0.f;
        const float t = __ldg(p.aux + (int64_t)m * p.ldaux + n) / mk;
        return acc * mk * act_bwd_from_out(p.act, t);
    }
    return acc;
}

template <int BM, int BN, bool AK, bool BKC>
__global__ void __launch_bounds__(kGemmThreads)
gemm_kernel(GemmParams p) {
    pdl_sync();
    constexpr int TM = BM / 16, TN = BN / 16;
    constexpr int SA = BM + 4, SB = BN + 4;
    __shared__ __align__(16) float As[2][kBK * SA];
    __shared__ __align__(16) float Bs[2][kBK * SB];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int kbeg = blockIdx.z * p.k_per_split;
    const int kend = min(p.K, kbeg + p.k_per_split);
    float* C = p.C + (int64_t)blockIdx.z * p.M * p.ldc;

    TileLoader<BM, AK> la;
    TileLoader<BN, BKC> lb;
    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

    la.load(p.A, p.lda, m0, kbeg, p.M, kend);
    lb.load(p.B, p.ldb, n0, kbeg, p.N, kend);
    la.store(As[0]);
    lb.store(Bs[0]);
    __syncthreads();
    int buf = 0;
    for (int k0 = kbeg; k0 < kend; k0 += kBK) {
        const bool more = k0 + kBK < kend;
        if (more) {
            la.load(p.A, p.lda, m0, k0 + kBK, p.M, kend);
            lb.load(p.B, p.ldb, n0, k0 + kBK, p.N, kend);
        }
        const float* as = As[buf];
        const float* bs = Bs[buf];
#pragma unroll
        for (int kk = 0; kk < kBK; ++kk) {
            float a[TM], b[TN];
            *reinterpret_cast<float4*>(a) = *reinterpret_cast<const float4*>(as + kk * SA + ty * 4);
            if (TM == 8) *reinterpret_cast<float4*>(a + 4) = *reinterpret_cast<const float4*>(as + kk * SA + BM / 2 + ty * 4);
            *reinterpret_cast<float4*>(b) = *reinterpret_cast<const float4*>(bs + kk * SB + tx * 4);
            if (TN == 8) *reinterpret_cast<float4*>(b + 4) = *reinterpret_cast<const float4*>(bs + kk * SB + BN / 2 + tx * 4);
#pragma unroll
            for (int i = 0; i < TM; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        if (more) {
            la.store(As[buf ^ 1]);
            lb.store(Bs[buf ^ 1]);
            __syncthreads();
            buf ^= 1;
        }
    }
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        const int m = m0 + (i < 4 ? ty * 4 + i : BM / 2 + ty * 4 + (i - 4));
        if (m >= p.M) continue;
#pragma unroll
        for (int jh = 0; jh < TN / 4; ++jh) {
            const int n = n0 + (jh == 0 ? tx * 4 : BN / 2 + tx * 4);
            float v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = (n + j < p.N) ? epilogue_apply(p, m, n + j, acc[i][jh * 4 + j]) : 0.f;
            float* dst = C + (int64_t)m * p.ldc + n;
            if (n + 3 < p.N) *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
            else
                for (int j = 0; j < 4; ++j) if (n + j < p.N) dst[j] = v[j];
        }
    }
}

// SGD: out is the weight matrix itself, out = out * decay - lr * (sum of the slabs) -- the gradient never goes to memory
template <bool SGD>
__global__ void splitk_reduce_kernel(const float* __restrict__ part, int splits, int M, int N, int ldc,
                                     float* __restrict__ out, int ldo, float lr, float decay) {
    pdl_sync();
    const int64_t total = (int64_t)M * N;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int m = (int)(t / N), n = (int)(t - (int64_t)m * N);
        float acc = 0.f;
        for (int z = 0; z < splits; ++z) acc += part[((int64_t)z * M + m) * ldc + n];
        float* o = out + (int64_t)m * ldo + n;
        *o = SGD ? fmaf(*o, decay, -lr * acc) : acc;
    }
}

template <int BM, int BN>
static void launch_gemm(int layout, const GemmParams& p, dim3 grid, cudaStream_t st) {
    if (layout == DCTR_GEMM_NN) launch_k(gemm_kernel<BM, BN, true, false>, grid, kGemmThreads, 0, st, p);
    else if (layout == DCTR_GEMM_NT) launch_k(gemm_kernel<BM, BN, true, true>, grid, kGemmThreads, 0, st, p);
    else launch_k(gemm_kernel<BM, BN, false, false>, grid, kGemmThreads, 0, st, p);
}

}  // namespace dctr

using namespace dctr;

extern "C" int dctr_gemm(int layout, int M, int N, int K, const float* A, int lda, const float* B, int ldb,
                         float* C, int ldc, int epilogue, int act, const float* bias, const float* aux, int ldaux,
                         const float* mask, int ldmask, int splits, void* stream) {
    DCTR_REQUIRE(layout >= 0 && layout <= 2, "gemm: bad layout %d", layout);
    DCTR_REQUIRE(M >= 0 && N >= 0 && K >= 0, "gemm: negative shape");
    if (M == 0 || N == 0) return DCTR_OK;
    DCTR_REQUIRE(A && B && C, "gemm: null pointer");
    DCTR_REQUIRE((lda & 3) == 0 && (ldb & 3) == 0 && (ldc & 3) == 0, "gemm: leading dimensions must be multiples of 4 (lda=%d ldb=%d ldc=%d)", lda, ldb, ldc);
    DCTR_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0 && ((uintptr_t)C & 15) == 0, "gemm: operands must be 16-byte aligned");
    DCTR_REQUIRE(epilogue >= 0 && epilogue <= 2 && (epilogue != DCTR_EPI_DGRAD || aux), "gemm: bad epilogue");
    if (splits < 1) splits = 1;
    DCTR_REQUIRE(splits == 1 || epilogue == DCTR_EPI_NONE, "gemm: split-K needs EPI_NONE");
    GemmParams p{M, N, K, A, lda, B, ldb, C, ldc, epilogue, act, bias, aux, ldaux, mask, ldmask, 0};
    int kper = (K + splits - 1) / splits;
    kper = (kper + kBK - 1) / kBK * kBK;
    if (kper == 0) kper = kBK;
    p.k_per_split = kper;
    cudaStream_t st = as_stream(stream);
    const int big = ceil_div(M, 128) * ceil_div(N, 128) * splits;
    if (big >= kNumSMs) {
        dim3 grid(ceil_div(N, 128), ceil_div(M, 128), splits);
        DCTR_REQUIRE(grid.y < 65536, "gemm: M too large");
        launch_gemm<128, 128>(layout, p, grid, st);
    } else {
        dim3 grid(ceil_div(N, 64), ceil_div(M, 64), splits);
        DCTR_REQUIRE(grid.y < 65536, "gemm: M too large");
        launch_gemm<64, 64>(layout, p, grid, st);
    }
    return check_launch("gemm_kernel");
}

extern "C" int dctr_splitk_reduce(const float* partials, int splits, int M, int N, int ldc, float* out, int ldo,
                                  void* stream) {
    DCTR_REQUIRE(partials && out && splits >= 1 && M >= 0 && N >= 0, "splitk_reduce: bad arguments");
    if (M == 0 || N == 0) return DCTR_OK;
    int blocks = ceil_div((int64_t)M * N, 256);
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    launch_k(splitk_reduce_kernel<false>, blocks, 256, 0, as_stream(stream), partials, splits, M, N, ldc, out, ldo, 0.f, 1.f);
    return check_launch("splitk_reduce_kernel");
}

extern "C" int dctr_splitk_reduce_sgd(const float* partials, int splits, int M, int N, int ldc, float* W, int ldw, float lr,
                                      float decay, void* stream) {
    DCTR_REQUIRE(partials && W && splits >= 1 && M >= 0 && N >= 0, "splitk_reduce_sgd: bad arguments");
    if (M == 0 || N == 0) return DCTR_OK;
    int blocks = ceil_div((int64_t)M * N, 256);
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    launch_k(splitk_reduce_kernel<true>, blocks, 256, 0, as_stream(stream), partials, splits, M, N, ldc, W, ldw, lr, decay);
    return check_launch("splitk_reduce_kernel");
}
