// vq_tensor.cuh -- nearest-codeword search on the 5th-gen tensor cores (LIPVQ_SEARCH_TENSOR).
//
// Replaces the [N,K,D] broadcast of LFQQuantizer.forward (v5:41-46) by
//   pass 0  vq_prep_*      : codebook -> mean mu, bf16 operand  Eb[k] = [-2(e_k-mu) | en_hi en_mid en_lo | 0..]
//                            (en = ||e_k-mu||^2 split in three bf16 so the GEMM itself adds it exactly)
//   pass 1  vq_convert_z   : z -> bf16 operand Zb[n] = [z_n-mu | 1 1 1 | 0..], zn[n] = ||z_n-mu||
//   pass 2  vq_tc_kernel   : S = Zb Eb^T  (= ||e'||^2 - 2 z'.e', the distance up to the row constant)
//                            as a TMA-fed tcgen05 GEMM with TMEM accumulators; the epilogue never writes
//                            S: it keeps, per row and column half, the 4 smallest chunk minima.
//   pass 3  vq_rescore     : every code of every chunk within tau(row) of the row minimum is rescored
//                            in fp32 direct form sum (z-e)^2 (the reference's arithmetic); first minimal
//                            index wins.  tau is a rigorous bound on the bf16 rounding of pass 2, so the
//                            true nearest code is always rescored; rows whose 4th chunk is still inside
//                            the window go to the exact fp32 scan (vq_exact.cuh).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"
#include "ptx_sm100.cuh"
#include "vq_exact.cuh"

namespace lipvq {
namespace tc {

constexpr int BM = 128;          // rows per CTA tile (UMMA M)
constexpr int BN = 256;          // codes per accumulator (UMMA N)
constexpr int KBLK = 64;         // bf16 elements per 128-byte swizzle row
constexpr int UK = 16;           // UMMA K for 16-bit inputs
constexpr int A_KB_BYTES = BM * 128;
constexpr int B_ST_BYTES = BN * 128;
constexpr int NT = 384;          // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warps 4..11 epilogue
constexpr int NCAND = 4;         // chunk minima kept per (row, column half)
constexpr float BIG = 3.0e38f;
constexpr float PAD_SCORE = 1.0e30f;   // en of the padding codes (never inside any window)

struct Geometry {
    int D, K, DK, DKp, NKB, NK16, Kp, nct;
    int64_t n, Np, n_tiles;
};
inline Geometry geometry(int64_t n, int D, int K) {
    Geometry g;
    g.D = D; g.K = K; g.n = n;
    g.DK = D + 3;
    g.DKp = (g.DK + KBLK - 1) / KBLK * KBLK;
    g.NKB = g.DKp / KBLK;
    g.NK16 = (g.DK + UK - 1) / UK;
    g.Kp = (K + BN - 1) / BN * BN;
    g.nct = g.Kp / BN;
    g.Np = (n + BM - 1) / BM * BM;
    g.n_tiles = g.Np / BM;
    return g;
}

// ------------------------------------------------------------------ pass 0: codebook operand
// mu[d] = mean_k E[k,d]   (one thread per column, fixed order)
__global__ void vq_prep_mean_kernel(const float *__restrict__ E, int K, int D, float *__restrict__ mu) {
    const int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= D) return;
    float s = 0.f;
    for (int k = 0; k < K; ++k) s += E[(size_t)k * D + d];
    mu[d] = s / (float)K;
}

__device__ __forceinline__ void split3_bf16(float v, __nv_bfloat16 &a, __nv_bfloat16 &b, __nv_bfloat16 &c) {
    a = __float2bfloat16_rn(v);
    const float r1 = v - __bfloat162float(a);
    b = __float2bfloat16_rn(r1);
    const float r2 = r1 - __bfloat162float(b);
    c = __float2bfloat16_rn(r2);
}

// one warp per (padded) code row; writes Eb row and enorm[k] = ||e'_k||
__global__ void vq_prep_rows_kernel(const float *__restrict__ E, const float *__restrict__ mu, int K, int Kp,
                                    int D, int DKp, __nv_bfloat16 *__restrict__ Eb, float *__restrict__ enorm) {
    const int k = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (k >= Kp) return;
    __nv_bfloat16 *row = Eb + (size_t)k * DKp;
    float s = 0.f;
    for (int d = lane; d < DKp; d += 32) {
        float v = 0.f;
        if (k < K && d < D) {
            v = E[(size_t)k * D + d] - mu[d];
            s = fmaf(v, v, s);
        }
        row[d] = __float2bfloat16_rn(-2.0f * v);
    }
    s = warp_sum(s);
    if (lane == 0) {
        const float en = (k < K) ? s : PAD_SCORE;
        __nv_bfloat16 a, b, c;
        split3_bf16(en, a, b, c);
        row[D] = a; row[D + 1] = b; row[D + 2] = c;
        if (k < K) enorm[k] = sqrtf(s);
    }
}

// emax = max_k enorm[k] (single block)
__global__ void vq_prep_emax_kernel(const float *__restrict__ enorm, int K, float *__restrict__ emax) {
    __shared__ float sm[32];
    float m = 0.f;
    for (int i = threadIdx.x; i < K; i += blockDim.x) m = fmaxf(m, enorm[i]);
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x < 32) {
        m = threadIdx.x < (blockDim.x >> 5) ? sm[threadIdx.x] : 0.f;
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (threadIdx.x == 0) *emax = m;
    }
}

// ------------------------------------------------------------------ pass 1: row operand
// one warp per (padded) row: Zb row = [bf16(z - mu) | 1 1 1 | 0...]; zn = ||z - mu||
__global__ void __launch_bounds__(256) vq_convert_z_kernel(const float *__restrict__ z, const float *__restrict__ mu,
                                                           int64_t n, int64_t Np, int D, int DKp,
                                                           __nv_bfloat16 *__restrict__ Zb, float *__restrict__ zn) {
    const int lane = threadIdx.x & 31;
    for (int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); r < Np; r += (int64_t)gridDim.x * 8) {
        __nv_bfloat16 *row = Zb + r * DKp;
        float s = 0.f;
        // two bf16 per lane per step -> 4-byte stores, 128 B per warp
        for (int d = lane * 2; d < DKp; d += 64) {
            float v0 = 0.f, v1 = 0.f;
            if (r < n) {
                if (d < D) v0 = z[r * D + d] - mu[d];
                else if (d < D + 3) v0 = 1.0f;
                if (d + 1 < D) v1 = z[r * D + d + 1] - mu[d + 1];
                else if (d + 1 < D + 3) v1 = 1.0f;
                if (d < D) s = fmaf(v0, v0, s);
                if (d + 1 < D) s = fmaf(v1, v1, s);
            }
            *reinterpret_cast<__nv_bfloat162 *>(row + d) = __floats2bfloat162_rn(v0, v1);
        }
        s = warp_sum(s);
        if (lane == 0 && r < n) zn[r] = sqrtf(s);
    }
}

}  // namespace tc
}  // namespace lipvq
