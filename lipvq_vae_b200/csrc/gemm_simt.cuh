// gemm_simt.cuh -- fp32 CUDA-core tile GEMM with fused epilogues, used for the small MLP layers
// of the tokenizer (v5:54-68) and their backward.  Exact fp32 FMA arithmetic (the 1e-5 parity
// tier); strides are generic so one kernel serves X*W^T, dY*W and dY^T*X (split over rows).
#pragma once
#include <cstdlib>
#include <cstring>

#include "common.cuh"
#include "bwd_umma.cuh"

namespace lipvq {

struct GemmArgs {
    const float *A; int64_t sam, sak;   // A(m,k) = A[m*sam + k*sak]
    const float *B; int64_t sbk, sbn;   // B(k,n) = B[k*sbk + n*sbn]
    float *C; int64_t ldc;              // C[m*ldc + n]
    int64_t M; int N; int64_t K;
    float alpha;                        // scale on the accumulator
    const float *alpha_dev;             // device scalar multiplying alpha (nullable)
    const float *bias;                  // [N] or null
    float *pre; int64_t ldpre;          // pre-activation output (nullable)
    const float *dpre; int64_t lddpre;  // multiply by DACT'(dpre[m,n]) (nullable)
    const float *aux1, *aux2; int64_t ldaux; float aux_alpha;  // += aux_alpha*(aux1-aux2)
    const float *aux_scale_dev;         // device scalar multiplying aux_alpha (nullable)
    int b_ones_col;                     // >=0: B(k, b_ones_col) == 1 (bias-gradient column)
    int64_t k_chunk;                    // split-K: rows per blockIdx.z (0 = no split)
};

constexpr int GBM = 64, GBN = 64, GBK = 16, GLD = 68, GNT = 256;

// A_KC: A is contiguous along k (sak==1); else contiguous along m (sam==1).
// B_KC: B is contiguous along k (sbk==1); else contiguous along n (sbn==1).
template <int ACT, int DACT, bool A_KC, bool B_KC>
__global__ void __launch_bounds__(GNT) gemm_simt_kernel(const GemmArgs g) {
    __shared__ __align__(16) float As[GBK][GLD];
    __shared__ __align__(16) float Bs[GBK][GLD];
    const int tid = threadIdx.x;
    const int ty = tid >> 4, tx = tid & 15;
    const int64_t m0 = (int64_t)blockIdx.x * GBM;
    const int n0 = blockIdx.y * GBN;
    int64_t kbeg = 0, kend = g.K;
    if (g.k_chunk > 0) {
        kbeg = (int64_t)blockIdx.z * g.k_chunk;
        kend = kbeg + g.k_chunk < g.K ? kbeg + g.k_chunk : g.K;
    }
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    float ra[4], rb[4];
    auto fetch = [&](int64_t k0) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int e = tid + q * GNT;
            int m, k;
            if (A_KC) { k = e & (GBK - 1); m = e >> 4; } else { m = e & (GBM - 1); k = e >> 6; }
            const int64_t gm = m0 + m, gk = k0 + k;
            ra[q] = (gm < g.M && gk < kend) ? __ldg(g.A + gm * g.sam + gk * g.sak) : 0.f;
            int n, kb;
            if (B_KC) { kb = e & (GBK - 1); n = e >> 4; } else { n = e & (GBN - 1); kb = e >> 6; }
            const int gn = n0 + n;
            const int64_t gkb = k0 + kb;
            float v = 0.f;
            if (gn < g.N && gkb < kend)
                v = (gn == g.b_ones_col) ? 1.f : __ldg(g.B + gkb * g.sbk + (int64_t)gn * g.sbn);
            rb[q] = v;
        }
    };
    auto stash = [&]() {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int e = tid + q * GNT;
            int m, k;
            if (A_KC) { k = e & (GBK - 1); m = e >> 4; } else { m = e & (GBM - 1); k = e >> 6; }
            As[k][m] = ra[q];
            int n, kb;
            if (B_KC) { kb = e & (GBK - 1); n = e >> 4; } else { n = e & (GBN - 1); kb = e >> 6; }
            Bs[kb][n] = rb[q];
        }
    };

    if (kbeg < kend) fetch(kbeg);
    for (int64_t k0 = kbeg; k0 < kend; k0 += GBK) {
        stash();
        __syncthreads();
        if (k0 + GBK < kend) fetch(k0 + GBK);
#pragma unroll
        for (int k = 0; k < GBK; ++k) {
            const float4 a = *reinterpret_cast<const float4 *>(&As[k][ty * 4]);
            const float4 b = *reinterpret_cast<const float4 *>(&Bs[k][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w};
            const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }

    float *C = g.C;
    const float aux_alpha = g.aux_alpha * ((g.aux1 && g.aux_scale_dev) ? __ldg(g.aux_scale_dev) : 1.f);
    const float alpha = g.alpha * (g.alpha_dev ? __ldg(g.alpha_dev) : 1.f);
    if (g.k_chunk > 0) C += (int64_t)blockIdx.z * g.M * g.ldc;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int64_t m = m0 + ty * 4 + i;
        if (m >= g.M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + tx * 4 + j;
            if (n >= g.N) continue;
            float v = alpha * acc[i][j];
            if (g.bias) v += __ldg(g.bias + n);
            if (g.pre) g.pre[m * g.ldpre + n] = v;
            v = apply_act<ACT>(v);
            if (DACT == ACT_GELU) v *= gelu_grad_f(__ldg(g.dpre + m * g.lddpre + n));
            if (DACT == ACT_RELU) v = __ldg(g.dpre + m * g.lddpre + n) > 0.f ? v : 0.f;
            if (g.aux1) v += aux_alpha * (__ldg(g.aux1 + m * g.ldaux + n) - __ldg(g.aux2 + m * g.ldaux + n));
            C[m * g.ldc + n] = v;
        }
    }
}

// ---------------------------------------------------------------- 3xTF32 variant for the big backward GEMMs
// Same tiling, operand staging and fused epilogue as gemm_simt_kernel, but the 64x64x16 tile product runs on the
// tensor cores as three tf32 mma.sync passes (a_hi*b_hi + a_lo*b_hi + a_hi*b_lo, hi = cvt.rna.tf32, lo = exact
// remainder): fp32-grade products.  The mma accumulators are folded into plain fp32 (round-to-nearest) sums after
// every 16-deep slab, so a reduction over 10^5..10^6 rows (weight gradients) does not inherit the tensor core's
// accumulation rounding.  8 warps = 2 (m) x 4 (n); a warp owns 32x16 = 2x2 m16n8 tiles.
constexpr int TLD = 72;   // slab row pitch: lane (g, t) reads word t*72 + g -> 32 distinct banks
__device__ __forceinline__ void split_tf32_rna(float v, uint32_t &hi, uint32_t &lo) {
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi) : "f"(v));
    lo = __float_as_uint(v - __uint_as_float(hi));
}
__device__ __forceinline__ void mma_tf32_16x8x8(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <int ACT, int DACT, bool A_KC, bool B_KC>
__global__ void __launch_bounds__(GNT) gemm_tc3_kernel(const GemmArgs g) {
    __shared__ __align__(16) float As[GBK][TLD];
    __shared__ __align__(16) float Bs[GBK][TLD];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int wm = warp >> 2, wn = warp & 3, gq = lane >> 2, tq = lane & 3;
    const int64_t m0 = (int64_t)blockIdx.x * GBM;
    const int n0 = blockIdx.y * GBN;
    int64_t kbeg = 0, kend = g.K;
    if (g.k_chunk > 0) {
        kbeg = (int64_t)blockIdx.z * g.k_chunk;
        kend = kbeg + g.k_chunk < g.K ? kbeg + g.k_chunk : g.K;
    }
    float acc[2][2][4];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[i][j][q] = 0.f;

    float ra[4], rb[4];
    auto fetch = [&](int64_t k0) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int e = tid + q * GNT;
            int m, k;
            if (A_KC) { k = e & (GBK - 1); m = e >> 4; } else { m = e & (GBM - 1); k = e >> 6; }
            const int64_t gm = m0 + m, gk = k0 + k;
            ra[q] = (gm < g.M && gk < kend) ? __ldg(g.A + gm * g.sam + gk * g.sak) : 0.f;
            int n, kb;
            if (B_KC) { kb = e & (GBK - 1); n = e >> 4; } else { n = e & (GBN - 1); kb = e >> 6; }
            const int gn = n0 + n;
            const int64_t gkb = k0 + kb;
            float v = 0.f;
            if (gn < g.N && gkb < kend)
                v = (gn == g.b_ones_col) ? 1.f : __ldg(g.B + gkb * g.sbk + (int64_t)gn * g.sbn);
            rb[q] = v;
        }
    };
    auto stash = [&]() {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int e = tid + q * GNT;
            int m, k;
            if (A_KC) { k = e & (GBK - 1); m = e >> 4; } else { m = e & (GBM - 1); k = e >> 6; }
            As[k][m] = ra[q];
            int n, kb;
            if (B_KC) { kb = e & (GBK - 1); n = e >> 4; } else { n = e & (GBN - 1); kb = e >> 6; }
            Bs[kb][n] = rb[q];
        }
    };

    if (kbeg < kend) fetch(kbeg);
    for (int64_t k0 = kbeg; k0 < kend; k0 += GBK) {
        stash();
        __syncthreads();
        if (k0 + GBK < kend) fetch(k0 + GBK);
        float slab[2][2][4];
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int q = 0; q < 4; ++q) slab[i][j][q] = 0.f;
#pragma unroll
        for (int ks = 0; ks < GBK; ks += 8) {
            uint32_t ah[2][4], al[2][4], bh[2][2], bl[2][2];
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const int m = wm * 32 + i * 16 + gq;
                split_tf32_rna(As[ks + tq][m], ah[i][0], al[i][0]);
                split_tf32_rna(As[ks + tq][m + 8], ah[i][1], al[i][1]);
                split_tf32_rna(As[ks + tq + 4][m], ah[i][2], al[i][2]);
                split_tf32_rna(As[ks + tq + 4][m + 8], ah[i][3], al[i][3]);
            }
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int n = wn * 16 + j * 8 + gq;
                split_tf32_rna(Bs[ks + tq][n], bh[j][0], bl[j][0]);
                split_tf32_rna(Bs[ks + tq + 4][n], bh[j][1], bl[j][1]);
            }
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    mma_tf32_16x8x8(slab[i][j], al[i], bh[j]);
                    mma_tf32_16x8x8(slab[i][j], ah[i], bl[j]);
                    mma_tf32_16x8x8(slab[i][j], ah[i], bh[j]);
                }
        }
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[i][j][q] += slab[i][j][q];
        __syncthreads();
    }

    float *C = g.C;
    const float aux_alpha = g.aux_alpha * ((g.aux1 && g.aux_scale_dev) ? __ldg(g.aux_scale_dev) : 1.f);
    const float alpha = g.alpha * (g.alpha_dev ? __ldg(g.alpha_dev) : 1.f);
    if (g.k_chunk > 0) C += (int64_t)blockIdx.z * g.M * g.ldc;
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int64_t m = m0 + wm * 32 + i * 16 + gq + (q >> 1) * 8;
                const int n = n0 + wn * 16 + j * 8 + 2 * tq + (q & 1);
                if (m >= g.M || n >= g.N) continue;
                float v = alpha * acc[i][j][q];
                if (g.bias) v += __ldg(g.bias + n);
                if (g.pre) g.pre[m * g.ldpre + n] = v;
                v = apply_act<ACT>(v);
                if (DACT == ACT_GELU) v *= gelu_grad_f(__ldg(g.dpre + m * g.lddpre + n));
                if (DACT == ACT_RELU) v = __ldg(g.dpre + m * g.lddpre + n) > 0.f ? v : 0.f;
            if (DACT == ACT_RELU) v = __ldg(g.dpre + m * g.lddpre + n) > 0.f ? v : 0.f;
                if (g.aux1) v += aux_alpha * (__ldg(g.aux1 + m * g.ldaux + n) - __ldg(g.aux2 + m * g.ldaux + n));
                C[m * g.ldc + n] = v;
            }
}

inline GemmArgs gemm_args_zero() {
    GemmArgs g;
    g.A = nullptr; g.sam = g.sak = 0; g.B = nullptr; g.sbk = g.sbn = 0; g.C = nullptr; g.ldc = 0;
    g.M = 0; g.N = 0; g.K = 0; g.alpha = 1.f; g.bias = nullptr; g.pre = nullptr; g.ldpre = 0;
    g.dpre = nullptr; g.lddpre = 0; g.aux1 = g.aux2 = nullptr; g.ldaux = 0; g.aux_alpha = 0.f;
    g.b_ones_col = -1; g.k_chunk = 0; g.aux_scale_dev = nullptr; g.alpha_dev = nullptr;
    return g;
}

template <int ACT, int DACT, bool A_KC, bool B_KC>
inline int gemm_launch(const GemmArgs &g, int splits, cudaStream_t st) {
    if (g.M <= 0 || g.N <= 0) return LIPVQ_OK;
    dim3 grid((unsigned)((g.M + GBM - 1) / GBM), (unsigned)((g.N + GBN - 1) / GBN), (unsigned)splits);
    // big products (the N-row backward GEMMs) go to the 3xTF32 tensor-core variant; LIPVQ_GEMM=simt forces FFMA
    static const bool simt = [] { const char *e = getenv("LIPVQ_GEMM"); return e && !strcmp(e, "simt"); }();
    if (!simt && (double)g.M * g.N * (double)g.K >= 1.0e7) {
        ProfScope prof("gemm_tc3_kernel", st);
        gemm_tc3_kernel<ACT, DACT, A_KC, B_KC><<<grid, GNT, 0, st>>>(g);
        LIPVQ_LAUNCH_OK();
        return LIPVQ_OK;
    }
    ProfScope prof("gemm_simt_kernel", st);
    gemm_simt_kernel<ACT, DACT, A_KC, B_KC><<<grid, GNT, 0, st>>>(g);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

// Y[M,N] = act(X[M,K] W[N,K]^T + b);  optional pre-activation output
template <int ACT>
inline int linear_fwd(const float *X, const float *W, const float *b, float *pre, float *Y,
                      int64_t M, int N, int K, cudaStream_t st) {
    GemmArgs g = gemm_args_zero();
    g.A = X; g.sam = K; g.sak = 1;
    g.B = W; g.sbk = 1; g.sbn = K;
    g.C = Y; g.ldc = N; g.M = M; g.N = N; g.K = K;
    g.bias = b; g.pre = pre; g.ldpre = N;
    return gemm_launch<ACT, ACT_NONE, true, true>(g, 1, st);
}

__device__ __forceinline__ void cp_async16(void *dst, const void *src, bool valid) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst);
    const int sz = valid ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(sz) : "memory");
}

// ---------------------------------------------------------------- input gradient on the tensor cores
// dX[M, KK] = (dY[M, Nn] W[Nn, KK]) (.) gelu'(dpre[M, KK])   (KK = 64 or 128 output columns, 3xTF32 mma.sync).
// Persistent CTAs: W stays in shared memory for the CTA's lifetime, dY streams through a 3-stage cp.async pipeline
// of [128 rows x 32 k] slabs that runs across row-tile boundaries; 8 warps = (8/WN) x WN with WN = KK/32, a warp
// owns (128*WN/8) rows x 32 columns.  Needs Nn % 4 == 0 and 16-byte aligned operands.
constexpr int DG_BM = 128, DG_BK = 32, DG_ST = 3, DG_ALD = DG_BK + 4;
template <int KK, int DACT>
__global__ void __launch_bounds__(GNT, 1) dgrad_tc3_kernel(const float *__restrict__ dY, const float *__restrict__ W,
                                                           const float *__restrict__ dpre, float *__restrict__ dX,
                                                           int64_t M, int Nn, const float *__restrict__ scale_dev) {
    constexpr int WN = KK / 32, WM = 8 / WN, MF = DG_BM / WM / 16, BLD = KK + 8;
    extern __shared__ __align__(16) float dg_smem[];
    float *Bs = dg_smem;                                             // [Nn_pad][BLD]
    const int n_slabs = (Nn + DG_BK - 1) / DG_BK;
    const float out_scale = scale_dev ? __ldg(scale_dev) : 1.f;
    float(*As)[DG_BM][DG_ALD] = reinterpret_cast<float(*)[DG_BM][DG_ALD]>(dg_smem + (size_t)n_slabs * DG_BK * BLD);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int wm = warp / WN, wn = warp % WN, gq = lane >> 2, tq = lane & 3;
    for (int i = tid; i < n_slabs * DG_BK * (KK / 4); i += GNT) {
        const int k = i / (KK / 4), c = (i - k * (KK / 4)) << 2;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (k < Nn) v = __ldg(reinterpret_cast<const float4 *>(W + (size_t)k * KK + c));
        *reinterpret_cast<float4 *>(Bs + (size_t)k * BLD + c) = v;
    }
    const int64_t n_tiles = (M + DG_BM - 1) / DG_BM;
    const int64_t my_tiles = blockIdx.x < n_tiles ? (n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int64_t total = my_tiles * n_slabs;
    auto issue = [&](int64_t s) {
        if (s < total) {
            const int64_t tile = blockIdx.x + (s / n_slabs) * gridDim.x;
            const int slab = (int)(s % n_slabs), stage = (int)(s % DG_ST);
            const int64_t r0 = tile * DG_BM;
#pragma unroll
            for (int q = 0; q < 4; ++q) {                           // 128 rows x 8 float4 = 1024 copies
                const int e = tid + q * GNT, r = e >> 3, c = (e & 7) << 2;
                const int k = slab * DG_BK + c;
                const bool ok = r0 + r < M && k < Nn;
                cp_async16(&As[stage][r][c], ok ? dY + (r0 + r) * Nn + k : dY, ok);
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    for (int s = 0; s < DG_ST - 1; ++s) issue(s);
    float acc[MF][4][4];
    for (int64_t s = 0; s < total; ++s) {
        const int slab = (int)(s % n_slabs), stage = (int)(s % DG_ST);
        if (slab == 0) {
#pragma unroll
            for (int i = 0; i < MF; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j)
#pragma unroll
                    for (int q = 0; q < 4; ++q) acc[i][j][q] = 0.f;
        }
        asm volatile("cp.async.wait_group %0;" ::"n"(DG_ST - 2) : "memory");
        __syncthreads();
        issue(s + DG_ST - 1);
        const float *Bk = Bs + (size_t)slab * DG_BK * BLD;
#pragma unroll
        for (int ks = 0; ks < DG_BK; ks += 8) {
            uint32_t bh[4][2], bl[4][2];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int n = wn * 32 + j * 8 + gq;
                split_tf32_rna(Bk[(ks + tq) * BLD + n], bh[j][0], bl[j][0]);
                split_tf32_rna(Bk[(ks + tq + 4) * BLD + n], bh[j][1], bl[j][1]);
            }
#pragma unroll
            for (int i = 0; i < MF; ++i) {
                const int m = wm * (MF * 16) + i * 16 + gq;
                uint32_t ah[4], al[4];
                split_tf32_rna(As[stage][m][ks + tq], ah[0], al[0]);
                split_tf32_rna(As[stage][m + 8][ks + tq], ah[1], al[1]);
                split_tf32_rna(As[stage][m][ks + tq + 4], ah[2], al[2]);
                split_tf32_rna(As[stage][m + 8][ks + tq + 4], ah[3], al[3]);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    mma_tf32_16x8x8(acc[i][j], al, bh[j]);
                    mma_tf32_16x8x8(acc[i][j], ah, bl[j]);
                    mma_tf32_16x8x8(acc[i][j], ah, bh[j]);
                }
            }
        }
        if (slab == n_slabs - 1) {
            const int64_t tile = blockIdx.x + (s / n_slabs) * gridDim.x;
#pragma unroll
            for (int i = 0; i < MF; ++i)
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int64_t m = tile * DG_BM + wm * (MF * 16) + i * 16 + gq + h * 8;
                    if (m >= M) continue;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int n = wn * 32 + j * 8 + 2 * tq;
                        float2 v = make_float2(out_scale * acc[i][j][2 * h], out_scale * acc[i][j][2 * h + 1]);
                        if (DACT == ACT_GELU) {
                            const float2 a = __ldg(reinterpret_cast<const float2 *>(dpre + m * KK + n));
                            v.x *= gelu_grad_f(a.x);
                            v.y *= gelu_grad_f(a.y);
                        }
                        if (DACT == ACT_RELU) {
                            const float2 a = __ldg(reinterpret_cast<const float2 *>(dpre + m * KK + n));
                            v.x = a.x > 0.f ? v.x : 0.f;
                            v.y = a.y > 0.f ? v.y : 0.f;
                        }
                        *reinterpret_cast<float2 *>(dX + m * KK + n) = v;
                    }
                }
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

template <int KK, int DACT>
inline int dgrad_tc3_launch(const float *dY, const float *W, const float *dpre, float *dX, int64_t M, int Nn,
                            const float *scale_dev, cudaStream_t st) {
    const int n_slabs = (Nn + DG_BK - 1) / DG_BK;
    const size_t smem = ((size_t)n_slabs * DG_BK * (KK + 8) + (size_t)DG_ST * DG_BM * DG_ALD) * sizeof(float);
    if (smem > 220 * 1024) return LIPVQ_ESHAPE;
    LIPVQ_CUDA_OK(cudaFuncSetAttribute(dgrad_tc3_kernel<KK, DACT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem));
    const int64_t tiles = (M + DG_BM - 1) / DG_BM;
    const int per_sm = smem <= 110 * 1024 ? 2 : 1;
    const int grid = (int)(tiles < per_sm * kNumSMs ? tiles : per_sm * kNumSMs);
    ProfScope prof("dgrad_tc3_kernel", st);
    dgrad_tc3_kernel<KK, DACT><<<grid, GNT, smem, st>>>(dY, W, dpre, dX, M, Nn, scale_dev);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

// dX[M,K] = (dY[M,N] W[N,K]) (.) DACT'(dpre[M,K])  (+ aux_alpha*(aux1-aux2))
template <int DACT>
inline int linear_bwd_input(const float *dY, const float *W, const float *dpre, float *dX,
                            int64_t M, int N, int K, const float *aux1, const float *aux2,
                            float aux_alpha, const float *aux_scale_dev, cudaStream_t st,
                            const float *out_scale_dev = nullptr) {
    static const bool simt = [] { const char *e = getenv("LIPVQ_GEMM"); return e && !strcmp(e, "simt"); }();
    static const bool legacy = [] { const char *e = getenv("LIPVQ_GEMM"); return e && !strcmp(e, "mma_sync"); }();
    if (!simt && !legacy && !aux1 && bwd5::dgrad_umma_ok(M, N, K) && aligned16(dY) && aligned16(W) && aligned16(dX) &&
        (DACT == ACT_NONE || aligned16(dpre)))
        return bwd5::dgrad_umma_launch<DACT>(dY, W, dpre, dX, M, N, K, out_scale_dev, st);
    if (!simt && !aux1 && (K == 64 || K == 128) && (N & 3) == 0 && aligned16(dY) && aligned16(W) && aligned16(dX) &&
        (DACT == ACT_NONE || aligned16(dpre)) && (double)M * N * K >= 1.0e7) {
        const size_t smem = ((size_t)((N + DG_BK - 1) / DG_BK) * DG_BK * (K + 8) + (size_t)DG_ST * DG_BM * DG_ALD) * 4;
        if (smem <= 220 * 1024)
            return K == 64 ? dgrad_tc3_launch<64, DACT>(dY, W, dpre, dX, M, N, out_scale_dev, st)
                           : dgrad_tc3_launch<128, DACT>(dY, W, dpre, dX, M, N, out_scale_dev, st);
    }
    GemmArgs g = gemm_args_zero();
    g.A = dY; g.sam = N; g.sak = 1;
    g.B = W; g.sbk = K; g.sbn = 1;
    g.C = dX; g.ldc = K; g.M = M; g.N = K; g.K = N;
    g.dpre = dpre; g.lddpre = K;
    g.aux1 = aux1; g.aux2 = aux2; g.ldaux = K; g.aux_alpha = aux_alpha; g.aux_scale_dev = aux_scale_dev;
    g.alpha_dev = out_scale_dev;
    return gemm_launch<ACT_NONE, DACT, true, false>(g, 1, st);
}

// Sum the split-K partials in a FIXED order (deterministic) and separate weight / bias gradients.
// part: [splits][Nout][Kin+1]; dW[Nout,Kin]; db[Nout] (nullable).  A block owns 64 consecutive elements; its four
// 64-thread slices each sum every 4th partial (coalesced 256-byte reads, 4 loads in flight per thread) and meet in
// shared memory in slice order -- a one-thread-per-element loop over 148 partials of 107 KB was 30 us of dependent loads.
constexpr int SKR_E = 64, SKR_S = 4;
__global__ void __launch_bounds__(SKR_E * SKR_S) splitk_reduce_kernel(const float *__restrict__ part, int splits, int Nout,
                                                                      int Kin, float *__restrict__ dW, float *__restrict__ db,
                                                                      const float *__restrict__ scale_dev = nullptr) {
    __shared__ float sm[SKR_S][SKR_E];
    const float out_scale = scale_dev ? __ldg(scale_dev) : 1.f;
    const int ld = Kin + 1;
    const int total = Nout * ld;
    const int e = threadIdx.x & (SKR_E - 1), sl = threadIdx.x / SKR_E;
    for (int i0 = blockIdx.x * SKR_E; i0 < total; i0 += gridDim.x * SKR_E) {
        const int i = i0 + e;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        if (i < total) {
            int z = sl;
            for (; z + 3 * SKR_S < splits; z += 4 * SKR_S) {
                a0 += part[(size_t)z * total + i];
                a1 += part[(size_t)(z + SKR_S) * total + i];
                a2 += part[(size_t)(z + 2 * SKR_S) * total + i];
                a3 += part[(size_t)(z + 3 * SKR_S) * total + i];
            }
            for (; z < splits; z += SKR_S) a0 += part[(size_t)z * total + i];
        }
        sm[sl][e] = (a0 + a1) + (a2 + a3);
        __syncthreads();
        if (sl == 0 && i < total) {
            const float v = (sm[0][e] + sm[1][e]) + (sm[2][e] + sm[3][e]);
            const int m = i / ld, n = i - m * ld;
            if (n < Kin) dW[(size_t)m * Kin + n] = out_scale * v;
            else if (db) db[m] = out_scale * v;
        }
        __syncthreads();
    }
}
inline void splitk_reduce_launch(const float *part, int splits, int Nout, int Kin, float *dW, float *db,
                                 const float *scale_dev, cudaStream_t st) {
    const int total = Nout * (Kin + 1);
    splitk_reduce_kernel<<<(total + SKR_E - 1) / SKR_E, SKR_E * SKR_S, 0, st>>>(part, splits, Nout, Kin, dW, db, scale_dev);
}

inline int splits_for_rows(int64_t rows, int Nout, int Kin) {
    const int tiles = ((Nout + GBM - 1) / GBM) * ((Kin + 1 + GBN - 1) / GBN);
    int64_t s = (4 * kNumSMs + tiles - 1) / tiles;
    const int64_t max_s = (rows + 255) / 256;   // at least 256 rows per split
    if (s > max_s) s = max_s;
    if (s < 1) s = 1;
    return (int)s;
}
inline size_t splitk_ws_floats(int64_t rows, int Nout, int Kin) {
    size_t s = (size_t)splits_for_rows(rows, Nout, Kin);
    if (bwd5::wgrad_umma_ok(rows, Nout, Kin) && s < (size_t)kNumSMs) s = kNumSMs;   // one partial per CTA
    return s * Nout * (Kin + 1);
}

// ---------------------------------------------------------------- weight gradient on the tensor cores
// part[z][Nout][Kin+1] = dY[rows_z, Nout]^T [X[rows_z, Kin] | 1]  for the row range of split z (3xTF32 mma.sync).
// The reduction axis is the ROW axis (10^5..10^6 long): both operands stream through a 4-stage cp.async pipeline
// of 32-row slabs (16-byte copies, zero-filled past the edges), a CTA owns a 64 x 64 tile of dW for its row range;
// the bias column (column sums of dY) is accumulated by 64 threads of the first column tile from the same slabs.
// Needs Nout % 4 == 0, Kin % 4 == 0 and 16-byte aligned operands (row pitches are then 16-byte multiples).
constexpr int WG_BK = 32, WG_ST = 4;
__global__ void __launch_bounds__(GNT) wgrad_tc3_kernel(const float *__restrict__ dY, const float *__restrict__ X,
                                                        int64_t rows, int Nout, int Kin, int64_t k_chunk,
                                                        float *__restrict__ part) {
    extern __shared__ __align__(16) float wg_smem[];
    float(*As)[WG_BK][TLD] = reinterpret_cast<float(*)[WG_BK][TLD]>(wg_smem);
    float(*Bs)[WG_BK][TLD] = reinterpret_cast<float(*)[WG_BK][TLD]>(wg_smem + WG_ST * WG_BK * TLD);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int wm = warp >> 2, wn = warp & 3, gq = lane >> 2, tq = lane & 3;
    const int m0 = blockIdx.x * GBM, n0 = blockIdx.y * GBN;
    const int64_t kbeg = (int64_t)blockIdx.z * k_chunk;
    const int64_t kend = kbeg + k_chunk < rows ? kbeg + k_chunk : rows;
    const int n_slabs = kbeg < kend ? (int)((kend - kbeg + WG_BK - 1) / WG_BK) : 0;
    // copy assignment: 32 rows x 16 float4 per operand per slab = 512 copies each, 2 + 2 per thread
    auto issue = [&](int slab, int stage) {
        const int64_t r0 = kbeg + (int64_t)slab * WG_BK;
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int e = tid + q * GNT, r = e >> 4, c = (e & 15) << 2;
            const int64_t row = r0 + r;
            const bool rok = slab < n_slabs && row < kend;
            const bool aok = rok && m0 + c < Nout, bok = rok && n0 + c < Kin;
            cp_async16(&As[stage][r][c], aok ? dY + row * Nout + m0 + c : dY, aok);
            cp_async16(&Bs[stage][r][c], bok ? X + row * Kin + n0 + c : X, bok);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    float acc[2][2][4];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[i][j][q] = 0.f;
    float bsum = 0.f;
    for (int s = 0; s < WG_ST - 1; ++s) issue(s, s);
    for (int slab = 0; slab < n_slabs; ++slab) {
        const int stage = slab % WG_ST;
        asm volatile("cp.async.wait_group %0;" ::"n"(WG_ST - 2) : "memory");
        __syncthreads();                                   // slab landed; everyone left the stage refilled below
        issue(slab + WG_ST - 1, (slab + WG_ST - 1) % WG_ST);
        float part_acc[2][2][4];
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int q = 0; q < 4; ++q) part_acc[i][j][q] = 0.f;
#pragma unroll
        for (int ks = 0; ks < WG_BK; ks += 8) {
            uint32_t ah[2][4], al[2][4], bh[2][2], bl[2][2];
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const int m = wm * 32 + i * 16 + gq;
                split_tf32_rna(As[stage][ks + tq][m], ah[i][0], al[i][0]);
                split_tf32_rna(As[stage][ks + tq][m + 8], ah[i][1], al[i][1]);
                split_tf32_rna(As[stage][ks + tq + 4][m], ah[i][2], al[i][2]);
                split_tf32_rna(As[stage][ks + tq + 4][m + 8], ah[i][3], al[i][3]);
            }
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int n = wn * 16 + j * 8 + gq;
                split_tf32_rna(Bs[stage][ks + tq][n], bh[j][0], bl[j][0]);
                split_tf32_rna(Bs[stage][ks + tq + 4][n], bh[j][1], bl[j][1]);
            }
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    mma_tf32_16x8x8(part_acc[i][j], al[i], bh[j]);
                    mma_tf32_16x8x8(part_acc[i][j], ah[i], bl[j]);
                    mma_tf32_16x8x8(part_acc[i][j], ah[i], bh[j]);
                }
        }
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[i][j][q] += part_acc[i][j][q];
        if (blockIdx.y == 0 && tid < GBM) {                // bias gradient: column sums of dY
            float t = 0.f;
#pragma unroll
            for (int k = 0; k < WG_BK; ++k) t += As[stage][k][tid];
            bsum += t;
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    const int ld = Kin + 1;
    float *C = part + (size_t)blockIdx.z * Nout * ld;
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j)
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int m = m0 + wm * 32 + i * 16 + gq + (q >> 1) * 8;
                const int n = n0 + wn * 16 + j * 8 + 2 * tq + (q & 1);
                if (m < Nout && n < Kin) C[(size_t)m * ld + n] = acc[i][j][q];
            }
    if (blockIdx.y == 0 && tid < GBM && m0 + tid < Nout) C[(size_t)(m0 + tid) * ld + Kin] = bsum;
}

// ---------------------------------------------------------------- weight gradient of a NARROW-input layer (Kin <= 16)
// The first encoder layer has Kin = feature_dim = 12 (v5:54): its weight gradient is 2 * Nout * Kin = 1.5 kflop per row
// against 304 bytes -- a streaming reduction, not a GEMM.  On the tensor-core kernel the per-slab machinery costs more
// than the data (measured 137 us for 200 MB); here a thread owns 4 rows of dW (4 x Kin accumulators), a "row lane" of
// Nout / 4 threads walks every RL-th row of the block's row range with plain fp32 FMAs (exact, no split products:
// 48 FMAs for 4 vector loads), and the row lanes are summed in a fixed order: part[z][Nout][Kin+1], bias column
// included.  KG = ceil(Kin / 4) float4 groups of X per row.
// dY and X rows of a block are contiguous in memory: they stream through a WN_NS-deep ring of WN_SR-row stages filled
// by 16-byte asynchronous copies (zero-filled past the block's last row), so the bytes in flight do not depend on
// registers.
constexpr int WN_NT = 256, WN_SR = 64, WN_NS = 4;
template <int KG>
__global__ void __launch_bounds__(WN_NT, 2) wgrad_narrow_kernel(const float *__restrict__ dY, const float *__restrict__ X,
                                                                 int64_t rows, int Nout, int Kin, int64_t k_chunk,
                                                                 float *__restrict__ part) {
    extern __shared__ __align__(16) float wn_sm[];
    const int tid = threadIdx.x;
    const int per_lane = Nout >> 2, RL = WN_NT / per_lane;
    const int lane_row = tid / per_lane, j0 = (tid - lane_row * per_lane) << 2;
    const bool active = lane_row < RL;
    const int64_t r_beg = (int64_t)blockIdx.x * k_chunk, r_end = r_beg + k_chunk < rows ? r_beg + k_chunk : rows;
    const int n_rows = r_beg < r_end ? (int)(r_end - r_beg) : 0, n_st = (n_rows + WN_SR - 1) / WN_SR;
    const int stage_floats = WN_SR * (Nout + Kin);
    float acc[4][4 * KG], bs[4];
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        bs[a] = 0.f;
#pragma unroll
        for (int b = 0; b < 4 * KG; ++b) acc[a][b] = 0.f;
    }
    auto issue = [&](int st_i) {
        if (st_i < n_st) {
            float *sy = wn_sm + (size_t)(st_i % WN_NS) * stage_floats, *sx = sy + WN_SR * Nout;
            const int live = n_rows - st_i * WN_SR < WN_SR ? n_rows - st_i * WN_SR : WN_SR;   // rows of this stage
            const float *gy = dY + (r_beg + (int64_t)st_i * WN_SR) * Nout, *gx = X + (r_beg + (int64_t)st_i * WN_SR) * Kin;
            const int ny = WN_SR * Nout / 4, ly = live * Nout / 4, nx = WN_SR * Kin / 4, lx = live * Kin / 4;
            for (int i = tid; i < ny; i += WN_NT) cp_async16(sy + 4 * i, i < ly ? gy + 4 * i : dY, i < ly);
            for (int i = tid; i < nx; i += WN_NT) cp_async16(sx + 4 * i, i < lx ? gx + 4 * i : X, i < lx);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    for (int s0 = 0; s0 < WN_NS - 1; ++s0) issue(s0);
    for (int st_i = 0; st_i < n_st; ++st_i) {
        issue(st_i + WN_NS - 1);   // refills the buffer consumed in the previous iteration (barrier below)
        asm volatile("cp.async.wait_group %0;" ::"n"(WN_NS - 1) : "memory");
        __syncthreads();
        if (active) {
            const float *sy = wn_sm + (size_t)(st_i % WN_NS) * stage_floats, *sx = sy + WN_SR * Nout;
#pragma unroll 2
            for (int rr = lane_row; rr < WN_SR; rr += RL) {
                const float4 g = *reinterpret_cast<const float4 *>(sy + rr * Nout + j0);
                const float gv[4] = {g.x, g.y, g.z, g.w};
                float4 x[KG];
#pragma unroll
                for (int q = 0; q < KG; ++q) x[q] = *reinterpret_cast<const float4 *>(sx + rr * Kin + 4 * q);
#pragma unroll
                for (int a = 0; a < 4; ++a) {
                    bs[a] += gv[a];
#pragma unroll
                    for (int q = 0; q < KG; ++q) {
                        acc[a][4 * q] = fmaf(gv[a], x[q].x, acc[a][4 * q]);
                        acc[a][4 * q + 1] = fmaf(gv[a], x[q].y, acc[a][4 * q + 1]);
                        acc[a][4 * q + 2] = fmaf(gv[a], x[q].z, acc[a][4 * q + 2]);
                        acc[a][4 * q + 3] = fmaf(gv[a], x[q].w, acc[a][4 * q + 3]);
                    }
                }
            }
        }
        __syncthreads();
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    // row lanes -> one partial, summed in lane order (round-robin over the shared scratch in groups of WN_G lanes)
    const int ld = Kin + 1, tile = Nout * ld;
    float *C = part + (size_t)blockIdx.x * tile;
    constexpr int WN_G = 4;   // row lanes staged per round: WN_G * tile floats of shared memory
    for (int l0 = 0; l0 < RL; l0 += WN_G) {
        if (active && lane_row >= l0 && lane_row < l0 + WN_G) {
            float *dst = wn_sm + (size_t)(lane_row - l0) * tile;
#pragma unroll
            for (int a = 0; a < 4; ++a) {
#pragma unroll
                for (int b = 0; b < 4 * KG; ++b)
                    if (b < Kin) dst[(j0 + a) * ld + b] = acc[a][b];
                dst[(j0 + a) * ld + Kin] = bs[a];
            }
        }
        __syncthreads();
        const int nl = RL - l0 < WN_G ? RL - l0 : WN_G;
        for (int i = tid; i < tile; i += WN_NT) {
            float v = l0 ? C[i] : 0.f;
            for (int l = 0; l < nl; ++l) v += wn_sm[(size_t)l * tile + i];
            C[i] = v;
        }
        __syncthreads();
    }
}
inline int wgrad_narrow_smem(int Nout, int Kin) {
    const int ring = WN_NS * WN_SR * (Nout + Kin) * (int)sizeof(float), red = 4 * Nout * (Kin + 1) * (int)sizeof(float);
    return ring > red ? ring : red;
}
inline bool wgrad_narrow_ok(int64_t rows, int Nout, int Kin) {
    return Kin >= 4 && Kin <= 16 && (Kin & 3) == 0 && (Nout & 3) == 0 && Nout >= 4 && Nout <= 4 * WN_NT && rows >= 2048 &&
           wgrad_narrow_smem(Nout, Kin) <= 100 * 1024;
}

// dW[Nout,Kin] = dY[rows,Nout]^T X[rows,Kin];  db[Nout] = column sums of dY.  part = scratch.
inline int linear_bwd_weight(const float *dY, const float *X, float *dW, float *db, int64_t rows,
                             int Nout, int Kin, float *part, cudaStream_t st, const float *out_scale_dev = nullptr) {
    const int splits = splits_for_rows(rows, Nout, Kin);
    static const bool simt = [] { const char *e = getenv("LIPVQ_GEMM"); return e && !strcmp(e, "simt"); }();
    static const bool legacy = [] { const char *e = getenv("LIPVQ_GEMM"); return e && !strcmp(e, "mma_sync"); }();
    static const bool no_narrow = [] { const char *e = getenv("LIPVQ_WGRAD_NARROW"); return e && !strcmp(e, "0"); }();
    if (!simt && !legacy && !no_narrow && wgrad_narrow_ok(rows, Nout, Kin) && aligned16(dY) && aligned16(X)) {
        const int64_t k_chunk = (rows + splits - 1) / splits;
        const int smem = wgrad_narrow_smem(Nout, Kin);
        {
            ProfScope prof("wgrad_narrow_kernel", st);
            auto go = [&](auto kern) -> int {
                LIPVQ_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
                kern<<<splits, WN_NT, smem, st>>>(dY, X, rows, Nout, Kin, k_chunk, part);
                return LIPVQ_OK;
            };
            switch (Kin >> 2) {
            case 1: LIPVQ_TRY(go(wgrad_narrow_kernel<1>)); break;
            case 2: LIPVQ_TRY(go(wgrad_narrow_kernel<2>)); break;
            case 3: LIPVQ_TRY(go(wgrad_narrow_kernel<3>)); break;
            default: LIPVQ_TRY(go(wgrad_narrow_kernel<4>)); break;
            }
            LIPVQ_LAUNCH_OK();
        }
        const int total = Nout * (Kin + 1);
        splitk_reduce_launch(part, splits, Nout, Kin, dW, db, out_scale_dev, st);
        LIPVQ_LAUNCH_OK();
        return LIPVQ_OK;
    }
    if (!simt && !legacy && bwd5::wgrad_umma_ok(rows, Nout, Kin) && aligned16(dY) && aligned16(X)) {
        int parts = 0;
        LIPVQ_TRY(bwd5::wgrad_umma_launch(dY, X, rows, Nout, Kin, part, &parts, st));
        const int total = Nout * (Kin + 1);
        splitk_reduce_launch(part, parts, Nout, Kin, dW, db, out_scale_dev, st);
        LIPVQ_LAUNCH_OK();
        return LIPVQ_OK;
    }
    if (!simt && (Nout & 3) == 0 && (Kin & 3) == 0 && aligned16(dY) && aligned16(X) &&
        (double)rows * Nout * Kin >= 1.0e7) {
        int64_t k_chunk = (rows + splits - 1) / splits;
        k_chunk = (k_chunk + WG_BK - 1) / WG_BK * WG_BK;
        const int smem = 2 * WG_ST * WG_BK * TLD * (int)sizeof(float);
        LIPVQ_CUDA_OK(cudaFuncSetAttribute(wgrad_tc3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        dim3 grid((unsigned)((Nout + GBM - 1) / GBM), (unsigned)((Kin + GBN - 1) / GBN), (unsigned)splits);
        {
            ProfScope prof("wgrad_tc3_kernel", st);
            wgrad_tc3_kernel<<<grid, GNT, smem, st>>>(dY, X, rows, Nout, Kin, k_chunk, part);
            LIPVQ_LAUNCH_OK();
        }
        const int total = Nout * (Kin + 1);
        splitk_reduce_launch(part, splits, Nout, Kin, dW, db, out_scale_dev, st);
        LIPVQ_LAUNCH_OK();
        return LIPVQ_OK;
    }
    GemmArgs g = gemm_args_zero();
    g.A = dY; g.sam = 1; g.sak = Nout;
    g.B = X; g.sbk = Kin; g.sbn = 1;
    g.C = part; g.ldc = Kin + 1; g.M = Nout; g.N = Kin + 1; g.K = rows;
    g.b_ones_col = Kin;
    g.k_chunk = (rows + splits - 1) / splits;
    g.k_chunk = (g.k_chunk + GBK - 1) / GBK * GBK;
    LIPVQ_TRY((gemm_launch<ACT_NONE, ACT_NONE, false, false>(g, splits, st)));
    const int total = Nout * (Kin + 1);
    splitk_reduce_launch(part, splits, Nout, Kin, dW, db, out_scale_dev, st);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

}  // namespace lipvq
