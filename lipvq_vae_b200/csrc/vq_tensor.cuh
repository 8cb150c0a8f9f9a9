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
#include <cuda_fp16.h>
#include <stdlib.h>

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
constexpr int NT = 640;          // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warps 4..19 epilogue
constexpr int NCW = 2;            // converter warps of the standalone search: warps 2 and 3 (extra warps would cut the
                                  // register budget of the 16 epilogue warps below what they need)
constexpr int NT_CONV = NT;
constexpr int NPART = 4;         // column parts of an accumulator (64 columns each), one epilogue warp per (quarter, part)
constexpr int NCAND = 4;         // chunk minima kept per (row, column half, code tile)
constexpr int LCAP = 16;         // candidate entries kept per (row, column half)
constexpr float BIG = 3.0e38f;

// GEMM operand type.  fp16 (default): 11-bit significand -> the rounding window tau is 8x narrower than with
// bf16 at the same tensor-core rate, so far fewer codes need the fp32 rescore.  Its narrow exponent range is
// guarded twice: an absolute (subnormal) term in tau, and a range flag (counters[3]) that sends every row to
// the exact scan when a value would overflow.  -DLIPVQ_VQ_BF16 restores bf16 operands (A/B measurements).
#ifdef LIPVQ_VQ_BF16
using op_t = __nv_bfloat16;
using op2_t = __nv_bfloat162;
constexpr float OP_U4 = 0.015625f;          // 4u, u = 2^-8
constexpr float OP_ABS = 0.f;               // no subnormal term (fp32 exponent range)
constexpr float OP_MAX = 3.0e38f;
__device__ __forceinline__ op_t f2op(float v) { return __float2bfloat16_rn(v); }
__device__ __forceinline__ float op2f(op_t v) { return __bfloat162float(v); }
__device__ __forceinline__ op2_t f2op2(float a, float b) { return __floats2bfloat162_rn(a, b); }
#else
using op_t = __half;
using op2_t = __half2;
constexpr float OP_U4 = 0.001953125f;       // 4u, u = 2^-11
constexpr float OP_ABS = 5.9604645e-8f * 1.01f;   // 2^-24: twice the largest subnormal rounding error
constexpr float OP_MAX = 60000.f;
__device__ __forceinline__ op_t f2op(float v) { return __float2half_rn(v); }
__device__ __forceinline__ float op2f(op_t v) { return __half2float(v); }
__device__ __forceinline__ op2_t f2op2(float a, float b) { return __floats2half2_rn(a, b); }
#endif
constexpr float PAD_SCORE = 1.0e30f;   // en of the padding codes (never inside any window)

struct Geometry {
    int D, K, DK, DKp, DKz, NKB, NK16, Kp, nct;   // DKp: Eb row pitch (whole K blocks); DKz: Zb row pitch (whole MMA K steps)
    int64_t n, Np, n_tiles;
};
inline Geometry geometry(int64_t n, int D, int K) {
    Geometry g;
    g.D = D; g.K = K; g.n = n;
    g.DK = D + 3;
    g.DKp = (g.DK + KBLK - 1) / KBLK * KBLK;
    g.NKB = g.DKp / KBLK;
    g.NK16 = (g.DK + UK - 1) / UK;
    g.DKz = g.NK16 * UK;   // the part of the last K block past DKz is out of bounds for the TMA loads (zero fill)
    g.Kp = (K + BN - 1) / BN * BN;
    g.nct = g.Kp / BN;
    g.Np = (n + BM - 1) / BM * BM;
    g.n_tiles = g.Np / BM;
    return g;
}

// ------------------------------------------------------------------ pass 0: codebook operand
// mu[d] = mean_k E[k,d]: block = 32 columns x 32 row-slices, combined in slice order (deterministic)
__global__ void __launch_bounds__(1024) vq_prep_mean_kernel(const float *__restrict__ E, int K, int D,
                                                            float *__restrict__ mu) {
    __shared__ float part[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int d = blockIdx.x * 32 + tx;
    float s = 0.f;
    if (d < D)
        for (int k = ty; k < K; k += 32) s += E[(size_t)k * D + d];
    part[ty][tx] = s;
    __syncthreads();
    if (ty == 0 && d < D) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < 32; ++i) t += part[i][tx];
        mu[d] = t / (float)K;
    }
}

__device__ __forceinline__ void split3_op(float v, op_t &a, op_t &b, op_t &c) {
    a = f2op(v);
    const float r1 = v - op2f(a);
    b = f2op(r1);
    const float r2 = r1 - op2f(b);
    c = f2op(r2);
}

// one warp per (padded) code row; writes Eb row and enorm[k] = ||e'_k||
__global__ void vq_prep_rows_kernel(const float *__restrict__ E, const float *__restrict__ mu, int K, int Kp,
                                    int D, int DKp, op_t *__restrict__ Eb, float *__restrict__ enorm,
                                    int *__restrict__ counters, float *__restrict__ emax) {
    const int k = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (k >= Kp) return;
    op_t *row = Eb + (size_t)k * DKp;
    float s = 0.f, amax = 0.f;
    for (int d = lane; d < DKp; d += 32) {
        float v = 0.f;
        if (k < K && d < D) {
            v = E[(size_t)k * D + d] - mu[d];
            s = fmaf(v, v, s);
            amax = fmaxf(amax, fabsf(v));
        }
        row[d] = f2op(-2.0f * v);
    }
    s = warp_sum(s);
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if (lane == 0) {
        if (k < K && !(2.f * amax <= OP_MAX && s <= 0.25f * OP_MAX)) atomicOr(counters + 3, 1);   // out of range
        const float en = (k < K) ? s : fminf(PAD_SCORE, OP_MAX);
        op_t a, b, c;
        split3_op(en, a, b, c);
        row[D] = a; row[D + 1] = b; row[D + 2] = c;
        if (k < K) {
            const float en_k = sqrtf(s);
            enorm[k] = en_k;
            // emax = max_k enorm[k]: non-negative floats order like their bit patterns (emax is zeroed with the counters)
            atomicMax(reinterpret_cast<int *>(emax), __float_as_int(en_k));
        }
    }
}

// ------------------------------------------------------------------ pass 1: row operand
// Zb row = [op(z - mu) | 1 1 1 | 0...]; zn = ||z - mu||.  One thread per 8 consecutive columns (32 B in, 16 B
// out, fully coalesced); the row norm is combined in shared memory in column order (deterministic).
// Rows never straddle a warp: a warp holds floor(32 / cpr) rows (cpr = 8-column chunks per row <= 32), the row norm
// is a fixed-order shuffle sum inside the row's lanes -- no block barrier in the streaming loop.
__global__ void __launch_bounds__(256) vq_convert_z_warp_kernel(const float *__restrict__ z, const float *__restrict__ mu,
                                                                int64_t n, int64_t Np, int D, int DKz,
                                                                op_t *__restrict__ Zb, float *__restrict__ zn,
                                                                int *__restrict__ counters) {
    const int cpr = DKz >> 3, rpw = 32 / cpr, rpb = 8 * rpw;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int rl = lane / cpr, ch = lane - rl * cpr;
    const bool active = rl < rpw, vec = (D & 3) == 0;
    const int d0 = ch * 8;
    // ncu (round 2): this pass was ISSUE bound (72 % issue-active, 273 instructions per 3 rows), not memory bound.  So: a
    // thread's 8 means live in registers (its columns never change), chunks that lie wholly inside the row take a
    // 3-instruction-per-value path (the "1 1 1 | 0" tail only exists in the last chunks), rows past n are written as
    // zeros on their own, and the row norm is a log-depth segmented tree instead of a cpr-step scan.
    float mu_r[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) mu_r[i] = (active && d0 + i < D) ? __ldg(mu + d0 + i) : 0.f;
    const bool full = active && d0 + 8 <= D && vec;      // whole chunk inside the row, 16-byte loads allowed
    int top = 1;
    while (top < cpr) top <<= 1;                         // first tree step: top / 2
    float amax = 0.f;
    bool bad = false;
    constexpr int U = 2;                                 // row sets per iteration: all loads issued before the first use
    const int64_t stride = (int64_t)gridDim.x * rpb;
    for (int64_t r0 = (int64_t)blockIdx.x * rpb; r0 < Np; r0 += U * stride) {
        float4 t[U][2];
        int64_t rr[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            rr[u] = r0 + u * stride + warp * rpw + rl;
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const int d = d0 + 4 * q;
                t[u][q] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (active && rr[u] < n && d < D) {
                    const float *p = z + rr[u] * D + d;
                    if (vec) t[u][q] = __ldg(reinterpret_cast<const float4 *>(p));
                    else {
                        t[u][q].x = __ldg(p);
                        if (d + 1 < D) t[u][q].y = __ldg(p + 1);
                        if (d + 2 < D) t[u][q].z = __ldg(p + 2);
                        if (d + 3 < D) t[u][q].w = __ldg(p + 3);
                    }
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t r = rr[u];
            if (r0 + u * stride >= Np) break;                 // warp-uniform
            float ss = 0.f;
            if (active && r < Np) {
                float v[8] = {t[u][0].x, t[u][0].y, t[u][0].z, t[u][0].w, t[u][1].x, t[u][1].y, t[u][1].z, t[u][1].w};
                if (r >= n) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) v[i] = 0.f;   // padding rows of the last tile: all-zero operand
                } else if (full) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        v[i] -= mu_r[i];
                        ss = fmaf(v[i], v[i], ss);
                        amax = fmaxf(amax, fabsf(v[i]));
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int d = d0 + i;
                        if (d < D) {
                            v[i] -= mu_r[i];
                            ss = fmaf(v[i], v[i], ss);
                            amax = fmaxf(amax, fabsf(v[i]));
                        } else {
                            v[i] = d < D + 3 ? 1.0f : 0.f;
                        }
                    }
                }
                __align__(16) op2_t o[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) o[i] = f2op2(v[2 * i], v[2 * i + 1]);
                *reinterpret_cast<uint4 *>(Zb + r * DKz + d0) = *reinterpret_cast<const uint4 *>(o);
            }
            bad |= !(ss <= 3.0e38f);                          // NaN / inf in the row (fmaxf above drops NaNs)
            // row norm: segmented tree over the row's cpr lanes (lanes ch < o add lane ch + o; fixed order)
            float tsum = ss;
            for (int o = top >> 1; o > 0; o >>= 1) {
                const float other = __shfl_down_sync(0xffffffffu, tsum, o);
                if (ch < o && ch + o < cpr) tsum += other;
            }
            if (active && ch == 0 && r < n) zn[r] = sqrtf(tsum);
        }
    }
    bad |= !(amax <= OP_MAX);
    if (__any_sync(0xffffffffu, bad) && lane == 0) atomicOr(counters + 3, 1);   // operand out of range -> exact search
}

// Rows wider than 32 chunks (D + 3 > 256): one warp per row, a lane takes chunks lane and lane + 32, the row norm is a
// full-warp shuffle tree.  (The block-wide kernel below, which this replaces for 33 <= cpr <= 64, synchronised the block
// once per 7 rows and ran at 2.4 TB/s.)
__global__ void __launch_bounds__(256) vq_convert_z_wide_kernel(const float *__restrict__ z, const float *__restrict__ mu,
                                                                int64_t n, int64_t Np, int D, int DKz,
                                                                op_t *__restrict__ Zb, float *__restrict__ zn,
                                                                int *__restrict__ counters) {
    const int cpr = DKz >> 3;                    // 33 .. 64
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool vec = (D & 3) == 0;
    float mu_r[2][8];
#pragma unroll
    for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int d = (lane + 32 * c) * 8 + i;
            mu_r[c][i] = d < D ? __ldg(mu + d) : 0.f;
        }
    float amax = 0.f;
    bool bad = false;
    const int64_t stride = (int64_t)gridDim.x * 8;
    for (int64_t r = (int64_t)blockIdx.x * 8 + warp; r < Np; r += stride) {
        float4 t[2][2];
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const int d = (lane + 32 * c) * 8 + 4 * q;
                t[c][q] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (r < n && d < D) {
                    const float *p = z + r * D + d;
                    if (vec) t[c][q] = __ldg(reinterpret_cast<const float4 *>(p));
                    else {
                        t[c][q].x = __ldg(p);
                        if (d + 1 < D) t[c][q].y = __ldg(p + 1);
                        if (d + 2 < D) t[c][q].z = __ldg(p + 2);
                        if (d + 3 < D) t[c][q].w = __ldg(p + 3);
                    }
                }
            }
        float ss = 0.f;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const int ch = lane + 32 * c, d0 = ch * 8;
            if (ch >= cpr) continue;
            float v[8] = {t[c][0].x, t[c][0].y, t[c][0].z, t[c][0].w, t[c][1].x, t[c][1].y, t[c][1].z, t[c][1].w};
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int d = d0 + i;
                if (r >= n) v[i] = 0.f;
                else if (d < D) {
                    v[i] -= mu_r[c][i];
                    ss = fmaf(v[i], v[i], ss);
                    amax = fmaxf(amax, fabsf(v[i]));
                } else {
                    v[i] = d < D + 3 ? 1.0f : 0.f;
                }
            }
            __align__(16) op2_t o[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) o[i] = f2op2(v[2 * i], v[2 * i + 1]);
            *reinterpret_cast<uint4 *>(Zb + r * DKz + d0) = *reinterpret_cast<const uint4 *>(o);
        }
        bad |= !(ss <= 3.0e38f);
        ss = warp_sum(ss);
        if (lane == 0 && r < n) zn[r] = sqrtf(ss);
    }
    bad |= !(amax <= OP_MAX);
    if (__any_sync(0xffffffffu, bad) && lane == 0) atomicOr(counters + 3, 1);   // operand out of range -> exact search
}

__global__ void __launch_bounds__(256) vq_convert_z_kernel(const float *__restrict__ z, const float *__restrict__ mu,
                                                           int64_t n, int64_t Np, int D, int DKp,
                                                           op_t *__restrict__ Zb, float *__restrict__ zn,
                                                           int *__restrict__ counters) {
    __shared__ float s_mu[5 * KBLK];
    __shared__ float s_ss[256];
    __shared__ int s_bad;
    const int cpr = DKp >> 3, rpb = 256 / cpr;            // 8-column chunks per row, rows per block iteration
    for (int i = threadIdx.x; i < DKp; i += 256) s_mu[i] = i < D ? mu[i] : 0.f;
    if (threadIdx.x == 0) s_bad = 0;
    __syncthreads();
    const int rl = threadIdx.x / cpr, ch = threadIdx.x - rl * cpr;
    const bool vec = (D & 3) == 0, pow2 = (cpr & (cpr - 1)) == 0 && cpr <= 32;
    bool bad = false;
    for (int64_t r0 = (int64_t)blockIdx.x * rpb; r0 < Np; r0 += (int64_t)gridDim.x * rpb) {
        const int64_t r = r0 + rl;
        float ss = 0.f;
        if (rl < rpb && r < Np) {
            float v[8];
            const int d0 = ch * 8;
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const int d = d0 + 4 * q;
                float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
                if (r < n && d < D) {
                    const float *p = z + r * D + d;
                    if (vec) t = __ldg(reinterpret_cast<const float4 *>(p));
                    else {
                        t.x = __ldg(p);
                        if (d + 1 < D) t.y = __ldg(p + 1);
                        if (d + 2 < D) t.z = __ldg(p + 2);
                        if (d + 3 < D) t.w = __ldg(p + 3);
                    }
                }
                v[4 * q] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int d = d0 + i;
                if (r < n && d < D) {
                    v[i] -= s_mu[d];
                    ss = fmaf(v[i], v[i], ss);
                    bad |= !(fabsf(v[i]) <= OP_MAX);
                } else {
                    v[i] = (r < n && d < D + 3) ? 1.0f : 0.f;
                }
            }
            __align__(16) op2_t o[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) o[i] = f2op2(v[2 * i], v[2 * i + 1]);
            *reinterpret_cast<uint4 *>(Zb + r * DKp + d0) = *reinterpret_cast<const uint4 *>(o);
        }
        if (pow2) {     // the row's chunks sit in one warp: fixed shuffle tree, no block barrier
            for (int o = cpr >> 1; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
            if (ch == 0 && r < n) zn[r] = sqrtf(ss);
        } else {
            s_ss[threadIdx.x] = ss;
            __syncthreads();
            if (threadIdx.x < rpb && r0 + threadIdx.x < n) {
                float t = 0.f;
                for (int c = 0; c < cpr; ++c) t += s_ss[threadIdx.x * cpr + c];
                zn[r0 + threadIdx.x] = sqrtf(t);
            }
            __syncthreads();
        }
    }
    if (bad) s_bad = 1;
    __syncthreads();
    if (threadIdx.x == 0 && s_bad) atomicOr(counters + 3, 1);   // operand out of range -> whole launch goes exact
}


// ------------------------------------------------------------------ pass 2: tcgen05 GEMM + fused chunk-min
// Shared memory (dynamic, 1024-B aligned): A[A_BUFS][NKB] 16 KB K-blocks | B[STAGES] 32 KB | barriers.
// All operand tiles are K-major with the 128-byte TMA/UMMA swizzle.
template <int NKB, int A_BUFS, int STAGES, bool PAIR = false>
struct SmemLayout {
    // PAIR (two CTAs, cta_group::2): a CTA holds HALF of every codebook stage, so the same ring bytes make 2 * STAGES
    // stages; extra barriers a_peer[A_BUFS], b_peer[BST] are used by the leader only
    static constexpr int BST = PAIR ? 2 * STAGES : STAGES;
    static constexpr int B_STAGE = PAIR ? B_ST_BYTES / 2 : B_ST_BYTES;
    static constexpr int A_BYTES = A_BUFS * NKB * A_KB_BYTES;
    static constexpr int B_BYTES = STAGES * B_ST_BYTES;
    static constexpr int BAR_OFF = A_BYTES + B_BYTES;
    static constexpr int NBARS = 2 * A_BUFS + 2 * BST + 4 + (PAIR ? A_BUFS + BST : 0);
    static constexpr int RES_OFF = BAR_OFF + NBARS * 8 + 16;        // row exchange of the in-kernel resolve: 3 x 128 words
    static constexpr int TOTAL = RES_OFF + 3 * BM * 4 + 1024;       // + alignment slack
};

constexpr float RESC_C1 = OP_U4 * 1.004f;       // 2*(2u+u^2), u = unit roundoff of the operand type
constexpr float RESC_C2 = 4.0e-5f;
constexpr int RESC_MAXD = 320;
// window half-width*2 for a row with ||z'|| = znr: relative rounding (C1), fp32 accumulation / split / payload
// bits (C2) and the absolute error of fp16 subnormals: sum_i 2^-25 (2|e_i| + |z_i|) <= 2^-25 sqrt(D) (2 emax + znr)
__device__ __forceinline__ float tau_of(float znr, float emax, float abs_c) {
    return 2.f * (RESC_C1 * znr * emax + RESC_C2 * (2.f * znr * emax + emax * emax)) + abs_c * (2.f * emax + znr) +
           1e-30f;
}

// Candidate record written by the epilogue for each (row, column part), laid out so that both the epilogue
// (thread = row) and the resolve pass (thread = row) touch it coalesced:
//   entry e of (tile, part, row r)  at cand[((tile*NPART + part)*LCAP + e)*BM + r] = {chunk-min bits, code0 | mask<<24}
//   meta of (tile, part, row r)     at cand_meta[(tile*NPART + part)*BM + r]       = entry count (LCAP+1 = overflow)
// One entry per 8-code CHUNK whose minimum was within tau(row) of the running row minimum when it was read
// (a superset of the final window: the running minimum only decreases); code0 = first code of the chunk and
// mask = which of its 8 elements were inside that window.  The common path costs the chunk minima only.
__device__ __noinline__ void cand_compact(uint2 *out, int &cnt, float T) {
    int k = 0;                  // nearly full: drop entries that fell out of the (shrinking) window
    const int n = cnt < LCAP ? cnt : LCAP;
    for (int e = 0; e < n; ++e) {
        const uint2 x = out[e * BM];
        if (__uint_as_float(x.x) <= T) out[(k++) * BM] = x;
    }
    cnt = k;
}
__device__ __forceinline__ float min8(const uint32_t *r) {
    const float a = ptx::fmin3(__uint_as_float(r[0]), __uint_as_float(r[1]), __uint_as_float(r[2]));
    const float b = ptx::fmin3(__uint_as_float(r[3]), __uint_as_float(r[4]), __uint_as_float(r[5]));
    return ptx::fmin3(a, b, fminf(__uint_as_float(r[6]), __uint_as_float(r[7])));
}

// predicated 8-byte global store (keeps the epilogue's record path free of divergent branches)
__device__ __forceinline__ void st_global_v2_if(uint2 *p, uint32_t a, uint32_t b, bool pred) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %3, 0;\n"
        "@p st.global.v2.b32 [%0], {%1, %2};\n"
        "}\n" ::"l"(p), "r"(a), "r"(b), "r"((uint32_t)pred)
        : "memory");
}

// order-preserving map float -> unsigned (atomicMin on the key == min on the float) and back
__device__ __forceinline__ unsigned fkey(float f) {
    const unsigned b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float fkey_inv(unsigned k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// CONV: the row operand is not read from a pre-converted fp16 image (tmapA) but built in shared memory by four
// converter warps straight from the fp32 rows z (thread = row; [fp16(z - mu) | 1 1 1 | 0..] in the TMA/UMMA 128-byte
// swizzle, zn = ||z - mu|| to global): the standalone search API needs no separate conversion pass over z.
// PAIR (opt-in, LIPVQ_VQ_PAIR=1): two CTAs of a cluster work as one cta_group::2 unit on TWO row tiles at a time
// (M = 256): each CTA holds its own 128 rows (A tile, TMEM accumulators, epilogue, decision) and HALF of every codebook
// stage; the leader CTA issues the MMAs for both, and both CTAs' TMA loads are counted on the leader's barriers.
// Built to test whether the codebook bytes an SM ingests bound the kernel at D = 208; they do not.  Measured with the
// epilogue's arithmetic switched off (tools/vq_shape_probe.py, N = 2^20, K = 1024): the MMA side alone takes
// ~570 + 185 cycles per MMA and accumulator in the one-CTA kernel (D = 253: 3 535 cycles for 16 MMAs) and the same in the
// pair (3 452) -- 185 cycles per 128 x 256 x 16 MMA is what this B200 sustains on streamed operands (69 % of the nominal
// 128; cuBLAS's measured burst peak is the same 69 % of nominal), whatever the ring depth, the issue loop or the B
// bytes per SM.  Parity-tested (tests/test_gpu_scale_parity.py, the pair-search test); not the default
// because it is no faster (269 vs 263 us in the forward).
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at the same shared-memory offset in CTA `rank` of the cluster
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar, uint32_t rank) {
    asm volatile(
        "{\n.reg .b32 ra;\n"
        "mapa.shared::cluster.u32 ra, %0, %1;\n"
        "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n}\n" ::"r"(bar), "r"(rank)
        : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {   // acquire at cluster scope
    asm volatile(
        "{\n.reg .pred p;\n"
        "WAITC_LOOP:\n"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAITC_DONE;\n"
        "bra WAITC_LOOP;\n"
        "WAITC_DONE:\n}\n" ::"r"(bar), "r"(parity)
        : "memory");
}
// TMA load whose completion is counted on the LEADER CTA's barrier (cta_group::2: the barrier may live in the peer CTA;
// clearing bit 24 of a shared::cluster address selects the even CTA of the pair), so the leader's MMA thread waits on
// ONE barrier for both halves and no relay thread sits in the operand path
__device__ __forceinline__ void tma_load_2d_pair(const CUtensorMap *m, uint32_t bar, uint32_t dst, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar & 0xFEFFFFFFu), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void umma_f16_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc),
        "r"(acc)
        : "memory");
}
// completion of all MMAs issued so far -> arrive on the barrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
                 "h"((uint16_t)3)
                 : "memory");
}

template <int NKB, int A_BUFS, int STAGES, bool CONV = false, bool PAIR = false>
__global__ void __launch_bounds__(CONV ? NT_CONV : NT, 1) vq_tc_kernel(const __grid_constant__ CUtensorMap tmapA,
                                                       const __grid_constant__ CUtensorMap tmapB,
                                                       int64_t n_rows, int n_tiles, int nct, int nk16,
                                                       const float *__restrict__ zn, const float *__restrict__ emax_p,
                                                       float abs_c, uint2 *__restrict__ cand,
                                                       int *__restrict__ cand_meta, int K,
                                                       int64_t *__restrict__ idx_out, int want_d2,
                                                       int *__restrict__ resc_list, int *__restrict__ exact_list,
                                                       int *__restrict__ counters,
                                                       const float *__restrict__ zsrc = nullptr,
                                                       const float *__restrict__ mu = nullptr, int D = 0,
                                                       float *__restrict__ zn_out = nullptr) {
    static_assert(!(CONV && PAIR), "the pair kernel reads the pre-converted row operand");
    using L = SmemLayout<NKB, A_BUFS, STAGES, PAIR>;
    constexpr int BST = L::BST, B_STAGE = L::B_STAGE;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = ptx::smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;      // (the same offset in both CTAs of a pair: same kernel, same layout)
    uint8_t *gen = smem_raw + (base - raw);
    const uint32_t sA = base, sB = base + L::A_BYTES, sBar = base + L::BAR_OFF;
    const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
    // barrier slots
    auto a_full = [&](int i) { return sBar + 8u * i; };
    auto a_empty = [&](int i) { return sBar + 8u * (A_BUFS + i); };
    auto b_full = [&](int i) { return sBar + 8u * (2 * A_BUFS + i); };
    auto b_empty = [&](int i) { return sBar + 8u * (2 * A_BUFS + BST + i); };
    auto t_full = [&](int i) { return sBar + 8u * (2 * A_BUFS + 2 * BST + i); };
    auto t_empty = [&](int i) { return sBar + 8u * (2 * A_BUFS + 2 * BST + 2 + i); };
    auto a_peer = [&](int i) { return sBar + 8u * (2 * A_BUFS + 2 * BST + 4 + i); };            // PAIR, leader: peer's A tile landed
    auto b_peer = [&](int i) { return sBar + 8u * (3 * A_BUFS + 2 * BST + 4 + i); };            // PAIR, leader: peer's B half landed
    volatile uint32_t *tmem_ptr = reinterpret_cast<volatile uint32_t *>(gen + L::BAR_OFF + L::NBARS * 8);
    // per-row exchange between the four column parts of a row (in-kernel resolve): ordered key of the row minimum,
    // number of listed elements inside the final window, lowest such code
    unsigned *s_min = reinterpret_cast<unsigned *>(gen + L::RES_OFF);
    int *s_nin = reinterpret_cast<int *>(s_min + BM), *s_code = s_nin + BM;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x < BM) { s_min[threadIdx.x] = 0xffffffffu; s_nin[threadIdx.x] = 0; s_code[threadIdx.x] = 0x7fffffff; }

    if (warp == 0 && lane == 0) {
        ptx::tma_prefetch_desc(&tmapA);
        ptx::tma_prefetch_desc(&tmapB);
        for (int i = 0; i < A_BUFS; ++i) { ptx::mbar_init(a_full(i), CONV ? NCW : 1); ptx::mbar_init(a_empty(i), 1); }
        for (int i = 0; i < BST; ++i) { ptx::mbar_init(b_full(i), 1); ptx::mbar_init(b_empty(i), 1); }
        // PAIR: the leader's t_empty collects the epilogue warps of BOTH CTAs (the MMAs write both CTAs' TMEM)
        for (int i = 0; i < 2; ++i) { ptx::mbar_init(t_full(i), 1); ptx::mbar_init(t_empty(i), (PAIR ? 2 : 1) * 4 * NPART); }
        if (PAIR) {
            for (int i = 0; i < A_BUFS; ++i) ptx::mbar_init(a_peer(i), 1);
            for (int i = 0; i < BST; ++i) ptx::mbar_init(b_peer(i), 1);
        }
        ptx::fence_barrier_init();
    }
    if (warp == 2) {
        if (PAIR) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(ptx::smem_u32((const void *)tmem_ptr)), "r"(512) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            ptx::tmem_alloc<512>(ptx::smem_u32((const void *)tmem_ptr));
        }
    }
    ptx::tc_fence_before();
    if (PAIR) cluster_sync_all(); else __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    // PAIR: the pair walks tile pairs (2 tp, 2 tp + 1); a CTA's own tile may lie past the end in the last pair (its rows
    // then fail row_ok; the row operand of the last real tile is loaded in its place)
    const int n_units = PAIR ? (n_tiles + 1) / 2 : n_tiles;
    const int unit0 = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x, unit_step = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    auto tile_of = [&](int u) { return PAIR ? 2 * u + (int)rank : u; };

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            int st = 0, sph = 0, ab = 0, aph = 0;
            auto load_a = [&](int u) {
                int tile = tile_of(u);
                if (tile >= n_tiles) tile = n_tiles - 1;      // PAIR, odd tile count: any valid rows (results are not used)
                ptx::mbar_wait(a_empty(ab), aph ^ 1);
                // PAIR: both CTAs' bytes are counted on the leader's barrier; only the leader arms it
                if (!PAIR || rank == 0) ptx::mbar_arrive_expect_tx(a_full(ab), (PAIR ? 2 : 1) * NKB * A_KB_BYTES);
#pragma unroll
                for (int kb = 0; kb < NKB; ++kb) {
                    if (PAIR) tma_load_2d_pair(&tmapA, a_full(ab), sA + (ab * NKB + kb) * A_KB_BYTES, kb * KBLK, tile * BM);
                    else ptx::tma_load_2d(&tmapA, a_full(ab), sA + (ab * NKB + kb) * A_KB_BYTES, kb * KBLK, tile * BM);
                }
                if (++ab == A_BUFS) { ab = 0; aph ^= 1; }
            };
            if (!CONV && A_BUFS > 1 && unit0 < n_units) load_a(unit0);
            for (int u = unit0; u < n_units; u += unit_step) {
                if (!CONV && A_BUFS == 1) load_a(u);
                for (int c = 0; c < nct; ++c) {
                    // double-buffered rows: the NEXT tile's rows are fetched after this tile's first accumulator has its
                    // codebook stages in flight.  Fetching them first (as this loop did) parks the producer on a_empty
                    // -- the previous tile's last MMAs -- with an EMPTY codebook ring, and every tile then starts with a
                    // TMA round trip nobody overlaps: the ~570 cycles per accumulator that tools/vq_shape_probe.py shows on
                    // top of the MMAs themselves.
                    if (!CONV && A_BUFS > 1 && c == (nct > 1 ? 1 : 0) && u + unit_step < n_units) load_a(u + unit_step);
                    for (int kb = 0; kb < NKB; ++kb) {
                        ptx::mbar_wait(b_empty(st), sph ^ 1);
                        if (PAIR) {   // this CTA's half of the 256 codes (tmapB then has 128-row boxes)
                            if (rank == 0) ptx::mbar_arrive_expect_tx(b_full(st), 2 * B_STAGE);
                            tma_load_2d_pair(&tmapB, b_full(st), sB + st * B_STAGE, kb * KBLK, c * BN + (int)rank * (BN / 2));
                        } else {
                            ptx::mbar_arrive_expect_tx(b_full(st), B_STAGE);
                            ptx::tma_load_2d(&tmapB, b_full(st), sB + st * B_STAGE, kb * KBLK, c * BN);
                        }
                        if (++st == BST) { st = 0; sph ^= 1; }
                    }
                }
            }
        }
    } else if (PAIR && warp == 1 && rank != 0) {
        // ===== PAIR, second CTA: the leader issues the MMAs for both =====
    } else if (warp == 1) {
        // ===== MMA issuer (one thread) =====
        if (lane == 0) {
            #ifdef LIPVQ_VQ_BF16
            constexpr uint32_t idesc = ptx::umma_idesc_bf16(PAIR ? 2 * BM : BM, BN);
#else
            constexpr uint32_t idesc = ptx::umma_idesc_f16(PAIR ? 2 * BM : BM, BN);
#endif
            // The issuing thread is a serial resource: measured (tools/mma_rate_probe.cu) an MMA costs max(50, N/2) cycles
            // when its operands are ready in uniform registers, but 160-235 when descriptor arithmetic, a counted loop
            // and register->uniform moves sit between two of them -- with 14 MMAs per accumulator at D=208 that, not
            // the tensor pipe (128 cycles per instruction), set the pace of this kernel.  So: descriptors are a constant
            // high word and a running low word (16-byte units; the 14-bit start field cannot overflow below 256 KB of
            // shared memory), the four K=16 slices of a K block are written out, and nothing else is computed here.
            const uint32_t desc_hi = (uint32_t)(1024 >> 4) | (1u << 14) | (2u << 29);      // SBO | version 1 | SWIZZLE_128B
            auto mk = [&](uint32_t lo) {
                uint64_t d;
                asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(desc_hi));
                return d;
            };
            const uint32_t a_lo0 = ((sA >> 4) & 0x3FFFu) | (1u << 16), b_lo0 = ((sB >> 4) & 0x3FFFu) | (1u << 16);
            constexpr uint32_t A_KB16 = A_KB_BYTES >> 4, B_ST16 = (uint32_t)B_STAGE >> 4, A_BUF16 = (uint32_t)NKB * A_KB16;
            auto mma = [&](uint32_t d, uint64_t ad, uint64_t bd, uint32_t acc_flag) {
                if (PAIR) umma_f16_pair(d, ad, bd, idesc, acc_flag);
                else ptx::umma_bf16(d, ad, bd, idesc, acc_flag);
            };
            auto commit = [&](uint32_t bar) {
                if (PAIR) umma_commit_pair(bar);
                else ptx::umma_commit(bar);
            };
            const int nk_last = nk16 - (NKB - 1) * (KBLK / UK);       // K=16 slices of the last K block (1..4)
            int st = 0, sph = 0, ab = 0, aph = 0, acc = 0, accph = 0;
            uint32_t b_lo = b_lo0, a_buf_lo = a_lo0;
            for (int u = unit0; u < n_units; u += unit_step) {
                if (PAIR) mbar_wait_cluster(a_full(ab), aph);
                else ptx::mbar_wait(a_full(ab), aph);
                ptx::tc_fence_after();
                for (int c = 0; c < nct; ++c) {
                    if (PAIR) mbar_wait_cluster(t_empty(acc), accph ^ 1);
                    else ptx::mbar_wait(t_empty(acc), accph ^ 1);
                    ptx::tc_fence_after();
                    const uint32_t d_tmem = tmem_base + acc * BN;
                    uint32_t a_lo = a_buf_lo;
#pragma unroll
                    for (int kb = 0; kb < NKB; ++kb) {
                        if (PAIR) mbar_wait_cluster(b_full(st), sph);
                        else ptx::mbar_wait(b_full(st), sph);
                        ptx::tc_fence_after();
                        if (kb < NKB - 1 || nk_last == 4) {
                            mma(d_tmem, mk(a_lo), mk(b_lo), kb ? 1u : 0u);
                            mma(d_tmem, mk(a_lo + 2), mk(b_lo + 2), 1u);
                            mma(d_tmem, mk(a_lo + 4), mk(b_lo + 4), 1u);
                            mma(d_tmem, mk(a_lo + 6), mk(b_lo + 6), 1u);
                        } else {
                            for (int k = 0; k < nk_last; ++k)   // +32 B per K=16 slice inside the swizzle row
                                mma(d_tmem, mk(a_lo + 2 * k), mk(b_lo + 2 * k), (kb | k) ? 1u : 0u);
                        }
                        commit(b_empty(st));
                        a_lo += A_KB16;
                        b_lo += B_ST16;
                        if (++st == BST) { st = 0; sph ^= 1; b_lo = b_lo0; }
                    }
                    commit(t_full(acc));
                    if (++acc == 2) { acc = 0; accph ^= 1; }
                }
                commit(a_empty(ab));
                a_buf_lo += A_BUF16;
                if (++ab == A_BUFS) { ab = 0; aph ^= 1; a_buf_lo = a_lo0; }
            }
        }
    } else if (CONV && (warp == 2 || warp == 3)) {
        // ===== converter: a warp takes row quads of the tile, one row at a time with its lanes ACROSS the row (coalesced
        // 512-byte requests; a lane-per-row layout cost 32 L1 wavefronts per request and made this role the limit of
        // the kernel); lane <-> 4 consecutive columns -> two half2 = 8 bytes of the swizzled K-major tile =====
        const int cw = warp - 2;
        const int Q = nk16 * 4;                         // 4-column groups per row (whole MMA K steps)
        const bool vec = (D & 3) == 0;
        int ab = 0, aph = 0;
        float row_ss = 0.f;
        bool row_bad_acc = false;
        auto load_quad = [&](const float *zr, bool ok, int q) -> float4 {
            float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
            const int d = 4 * q;
            if (ok && q < Q && d < D) {
                if (vec) t = __ldg(reinterpret_cast<const float4 *>(zr + d));
                else {
                    t.x = __ldg(zr + d);
                    if (d + 1 < D) t.y = __ldg(zr + d + 1);
                    if (d + 2 < D) t.z = __ldg(zr + d + 2);
                    if (d + 3 < D) t.w = __ldg(zr + d + 3);
                }
            }
            return t;
        };
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            ptx::mbar_wait(a_empty(ab), aph ^ 1);
            const uint32_t tile_s = sA + (uint32_t)(ab * NKB * A_KB_BYTES);
            for (int r0 = 8 * cw; r0 < BM; r0 += 8 * NCW) {        // 8 rows in flight; row groups dealt round-robin to the warps
                for (int p = 0; p * 32 < Q; ++p) {
                    const int q = p * 32 + lane;
                    float4 t[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int64_t row = (int64_t)tile * BM + r0 + u;
                        t[u] = load_quad(zsrc + row * D, row < n_rows, q);
                    }
                    float4 m4 = make_float4(0.f, 0.f, 0.f, 0.f);
                    const int d = 4 * q;
                    if (q < Q && d < D) {
                        if (vec) m4 = __ldg(reinterpret_cast<const float4 *>(mu + d));
                        else {
                            m4.x = __ldg(mu + d);
                            if (d + 1 < D) m4.y = __ldg(mu + d + 1);
                            if (d + 2 < D) m4.z = __ldg(mu + d + 2);
                            if (d + 3 < D) m4.w = __ldg(mu + d + 3);
                        }
                    }
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int r = r0 + u;
                        const bool ok = (int64_t)tile * BM + r < n_rows;
                        float v[4] = {t[u].x - m4.x, t[u].y - m4.y, t[u].z - m4.z, t[u].w - m4.w};
                        float ssq = 0.f;
                        bool bad = false;
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int dd = d + i;
                            if (ok && dd < D) {
                                ssq = fmaf(v[i], v[i], ssq);
                                bad |= !(fabsf(v[i]) <= OP_MAX);     // (also catches NaN)
                            } else {
                                v[i] = (ok && dd < D + 3) ? 1.0f : 0.f;
                            }
                        }
                        if (q < Q) {
                            const op2_t p0 = f2op2(v[0], v[1]), p1 = f2op2(v[2], v[3]);
                            const uint32_t dst = tile_s + (uint32_t)((q >> 4) * A_KB_BYTES) + (uint32_t)r * 128u +
                                                 (((uint32_t)((q & 15) >> 1) ^ (uint32_t)(r & 7)) << 4) + (uint32_t)(q & 1) * 8u;
                            asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(dst),
                                         "r"(*reinterpret_cast<const uint32_t *>(&p0)),
                                         "r"(*reinterpret_cast<const uint32_t *>(&p1))
                                         : "memory");
                        }
                        // row norm: this pass's lanes, fixed shuffle tree; passes accumulate in order (deterministic)
                        ssq = warp_sum(ssq);
                        const bool any_bad = __any_sync(0xffffffffu, bad);
                        if (lane == u) {          // lane u keeps row r0+u's running (ss, bad) across the passes
                            row_ss = (p == 0 ? 0.f : row_ss) + ssq;
                            row_bad_acc = (p == 0 ? false : row_bad_acc) || any_bad;
                        }
                    }
                }
                if (lane < 8) {
                    const int64_t row = (int64_t)tile * BM + r0 + lane;
                    if (row < n_rows) zn_out[row] = row_bad_acc ? -1.0f : sqrtf(row_ss);   // negative: exact scan for this row
                }
            }
            __threadfence();                 // zn is read by this CTA's epilogue warps (ld.cg) and by the rescore pass
            ptx::fence_proxy_async();        // the tile is read by the tensor core through the async proxy
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(a_full(ab));
            if (++ab == A_BUFS) { ab = 0; aph ^= 1; }
        }
    } else if (warp >= 4) {
        // ===== epilogue: 16 warps; warp%4 = TMEM lane quarter, (warp-4)/4 = 64-column part =====
        // Measured (tools/tmem_probe.cu, profiles/r02_tmem_probe.txt): 16 warps of tcgen05.ld.32x32b.x32 read TMEM at
        // 530 B/clk/SM = 250 cycles per 128x256 fp32 accumulator, also under a concurrent MMA stream -- far below the
        // MMA time of an accumulator (161 cycles per M=128 N=256 K=16 instruction with both operands in shared memory:
        // 805 cycles at D=64, 2254 at D=208).  What the epilogue must not exceed is its INSTRUCTION budget, so the
        // per-chunk work below is branch-free: with 32 rows per warp a divergent "record this chunk" body is taken by
        // some lane almost every time (each row has its own running minimum), and used to cost ~3000 issue cycles
        // per accumulator.
        const int quarter = warp & 3, part = (warp - 4) >> 2;
        const uint32_t lane_addr = ((uint32_t)(quarter * 32) << 16);
        constexpr int PW = BN / NPART;            // columns per part
        int acc = 0, accph = 0;
        const float emax = __ldg(emax_p);
        // set by the operand passes, which finished before this launch (CONV: a row whose own operand left the fp16
        // range is flagged by its converter thread through a negative zn and goes to the exact scan by itself)
        const bool range_bad = __ldg(counters + 3) != 0;
        for (int u = unit0; u < n_units; u += unit_step) {
            const int tile = tile_of(u);      // (PAIR: may be n_tiles in the last pair -- every row then fails row_ok)
            const int r_in = quarter * 32 + lane;
            const int64_t row = (int64_t)tile * BM + r_in;
            const bool row_ok = row < n_rows;
            // CONV: zn of this tile is written by this CTA's converter warps before they release the tile to the MMA
            // issuer, so it may only be read once the tile's FIRST accumulator has been signalled (below); reading it
            // here would race with the converter (and pick up the previous call's value from the workspace)
            float znr = (row_ok && !CONV) ? __ldg(zn + row) : 0.f;
            bool row_bad = false;
            float tau = tau_of(znr, emax, abs_c);
            uint2 *out = cand + ((size_t)tile * NPART + part) * LCAP * BM + r_in;
            float gb = BIG, m1 = BIG, m2 = BIG;
            uint32_t info1 = 0;
            int cnt = 0;
            bool ovf = false;
            for (int c = 0; c < nct; ++c) {
                ptx::mbar_wait(t_full(acc), accph);
                ptx::tc_fence_after();
                if (CONV && c == 0) {
                    znr = row_ok ? __ldcg(zn + row) : 0.f;
                    row_bad = !(znr >= 0.f);
                    if (row_bad) znr = 0.f;
                    tau = tau_of(znr, emax, abs_c);
                }
                const uint32_t col0 = tmem_base + acc * BN + part * PW;
                const int code0 = c * BN + part * PW;
#pragma unroll
                for (int b = 0; b < PW / 32; ++b) {
                    uint32_t r[32];
                    ptx::tmem_ld_32x32b_x32(lane_addr + col0 + b * 32, r);
                    ptx::tmem_ld_wait();
                    if (b == PW / 32 - 1) {   // accumulator part fully in registers: hand the TMEM stage back
                        ptx::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) {
                            if (PAIR && rank != 0) mbar_arrive_cluster(t_empty(acc), 0u);   // the leader issues the MMAs for both
                            else ptx::mbar_arrive(t_empty(acc));
                        }
                    }
                    float cm[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) cm[j] = min8(r + 8 * j);
                    gb = fminf(gb, ptx::fmin3(cm[0], cm[1], fminf(cm[2], cm[3])));
                    const float T = gb + tau;
                    const float Tm = T + fabsf(T) * 1.2e-7f + 1e-37f;     // > T: (v - Tm) is negative for every v <= T
                    // room for this batch's (at most 4) records, so the per-chunk stores need no capacity test
                    if (cnt > LCAP - 4) {                                  // rare
                        cand_compact(out, cnt, T);
                        if (cnt > LCAP - 4) { ovf = true; cnt = 0; }       // flat score landscape: the row takes the exact scan
                    }
#ifndef LIPVQ_EPI_CHUNKWISE
                    // Records are rare per ROW (a handful of new running minima per part) but not per WARP: with 32 rows in
                    // a warp some lane has one in ~98 % of the chunks, so code under `if (put)` per chunk runs for every
                    // chunk -- and the 8-element window mask (16 instructions) was 60 % of this epilogue.  Here the four
                    // chunk tests of a batch are collected first and the record body runs ONCE per batch, every lane on
                    // ITS chunk (values picked with selects); a second trip only for a lane with two chunks of one batch
                    // inside the window.  The two smallest chunk minima are tracked in the same body: a chunk that is a
                    // new minimum or lies inside the final window is always a record.
                    uint32_t pm = (cm[0] <= T ? 1u : 0u) | (cm[1] <= T ? 2u : 0u) | (cm[2] <= T ? 4u : 0u) | (cm[3] <= T ? 8u : 0u);
                    pm = row_ok ? pm : 0u;
                    while (pm) {
                        const int j = __ffs(pm) - 1;
                        pm &= pm - 1;
                        const bool j0 = j == 0, j1 = j == 1, j2 = j == 2;
                        const float cmj = j0 ? cm[0] : j1 ? cm[1] : j2 ? cm[2] : cm[3];
                        uint32_t m = 0;
#pragma unroll
                        for (int e = 7; e >= 0; --e) {
                            const uint32_t bits = j0 ? r[e] : j1 ? r[8 + e] : j2 ? r[16 + e] : r[24 + e];
                            m = __funnelshift_l(__float_as_uint(__uint_as_float(bits) - Tm), m, 1);
                        }
                        const uint32_t info = (uint32_t)(code0 + b * 32 + 8 * j) | (m << 24);
                        out[cnt * BM] = make_uint2(__float_as_uint(cmj), info);
                        ++cnt;
                        const bool lt = cmj < m1;
                        m2 = lt ? m1 : fminf(m2, cmj);
                        info1 = lt ? info : info1;
                        m1 = lt ? cmj : m1;
                    }
#else
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const bool put = (cm[j] <= T) && row_ok;
                        uint32_t info = (uint32_t)(code0 + b * 32 + 8 * j);
#ifdef LIPVQ_EPI_PRED
                        // branch-free variant (A/B: measured SLOWER, 4.4 k vs 3.1-3.5 k cycles per accumulator -- the ALU
                        // pipe, not divergence, is what the epilogue runs out of): mask always, predicated record
                        uint32_t m = 0;
#pragma unroll
                        for (int e = 7; e >= 0; --e)
                            m = __funnelshift_l(__float_as_uint(__uint_as_float(r[8 * j + e]) - Tm), m, 1);
                        info |= m << 24;
                        st_global_v2_if(out + cnt * BM, __float_as_uint(cm[j]), info, put);
                        cnt += put ? 1 : 0;
#else
                        // one record per chunk whose minimum is inside the running window: the window mask of its 8
                        // elements (sign bit of v - Tm shifted in, last element first so that element e lands in bit e)
                        if (put) {
                            uint32_t m = 0;
#pragma unroll
                            for (int e = 7; e >= 0; --e)
                                m = __funnelshift_l(__float_as_uint(__uint_as_float(r[8 * j + e]) - Tm), m, 1);
                            info |= m << 24;
                            out[cnt * BM] = make_uint2(__float_as_uint(cm[j]), info);
                            ++cnt;
                        }
#endif
                        // the two smallest chunk minima of this (row, part) stay in registers (a new part minimum is
                        // always inside the running window, so its record -- and mask -- exists): the tile-end decision
                        // below needs nothing from memory
                        const bool lt = cm[j] < m1;
                        m2 = lt ? m1 : fminf(m2, cm[j]);
                        info1 = lt ? info : info1;
                        m1 = lt ? cm[j] : m1;
                    }
#endif
                }
                if (++acc == 2) { acc = 0; accph ^= 1; }
            }
            if (ovf) cnt = LCAP + 1;
            if (row_ok) cand_meta[((size_t)tile * NPART + part) * BM + r_in] = cnt;
            // ---- resolve (was a separate kernel): the four parts of a row agree on the row minimum, count the listed
            // elements inside the FINAL window tau(row) and decide: exactly one -> it IS the nearest code, idx is
            // written here; several -> fp32 rescore list; list overflow / range flag -> exact scan list.
            asm volatile("bar.sync 1, 512;" ::: "memory");      // the previous tile's decision (and reset) is done
            if (row_ok) atomicMin(&s_min[r_in], fkey(m1));
            asm volatile("bar.sync 1, 512;" ::: "memory");
            if (row_ok) {
                // (the mask of the best chunk was taken against the window of ITS time, a superset of the final one: a
                //  row can only be sent to the rescore list needlessly, never decided wrongly)
                const float T = fkey_inv(s_min[r_in]) + tau;
                int nin = ovf ? (1 << 20) : 0, code = 0x7fffffff;
                if (m1 <= T) {
                    const int c0 = (int)(info1 & 0xffffffu);
                    uint32_t m = info1 >> 24;
                    if (c0 + 8 > K) m &= (1u << max(K - c0, 0)) - 1u;     // padding codes of the last chunk
                    nin += __popc(m);
                    if (m) code = c0 + __ffs(m) - 1;
                }
                if (m2 <= T) nin += 1 << 10;                              // a second chunk of this part: rescore decides
                if (nin) atomicAdd(&s_nin[r_in], nin);
                if (code != 0x7fffffff) atomicMin(&s_code[r_in], code);
            }
            asm volatile("bar.sync 1, 512;" ::: "memory");
            if (part == 0) {
                bool fast = false;
                if (row_ok) {
                    const int nin = s_nin[r_in];
                    if (range_bad || row_bad || nin >= (1 << 20) || nin == 0)   // 0 cannot happen for finite inputs
                        exact_list[atomicAdd(counters + 0, 1)] = (int)row;
                    else if (nin == 1 && !want_d2) { idx_out[row] = (int64_t)s_code[r_in]; fast = true; }
                    else resc_list[atomicAdd(counters + 4, 1)] = (int)row;
                }
                s_min[r_in] = 0xffffffffu; s_nin[r_in] = 0; s_code[r_in] = 0x7fffffff;
                const unsigned nf = __popc(__ballot_sync(0xffffffffu, fast));
                if (lane == 0 && nf) atomicAdd(counters + 2, (int)nf);
            }
        }
    }
    ptx::tc_fence_before();
    if (PAIR) cluster_sync_all(); else __syncthreads();      // (PAIR: the peer's barriers and shared memory stay alive until both are done)
    if (warp == 2) {
        if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
        else ptx::tmem_dealloc<512>(tmem_base);
    }
}


// ------------------------------------------------------------------ pass 3: resolve + fp32 direct-form rescore
// Rounding bound of pass 2 for row n and ANY code k (u = unit roundoff of the operand type, round-to-nearest on
// both operands, Cauchy-Schwarz; fp32 accumulation and the 3-way split of en are covered by C2, fp16
// subnormals by the absolute term):
//   |S_nk - (||e'_k||^2 - 2 z'_n.e'_k)| <= eps_n = tau_of(zn, emax)/2
// so every code whose exact score is minimal has S <= min_k S + 2*eps_n.
//
// counters: [0] rows sent to the exact scan, [1] codes rescored, [2] rows resolved without rescoring,
//           [3] operand-range flag, [4] rows on the rescore list.

// 3b. one WARP per listed row: the listed elements inside the window (<= 64) are rescored NG at a time in fp32
// direct form sum (z-e)^2 (the reference's arithmetic, v5:41-45); the first minimal index wins.
template <int NG, bool VEC>
__global__ void __launch_bounds__(256) vq_rescore_kernel(
    const float *__restrict__ z, const float *__restrict__ E, int D, int K, const uint2 *__restrict__ cand,
    const int *__restrict__ cand_meta, const float *__restrict__ zn, const float *__restrict__ emax_p, float abs_c,
    int64_t *__restrict__ idx_out, float *__restrict__ d2_out, const int *__restrict__ resc_list,
    int *__restrict__ counters, int *__restrict__ exact_list) {
    constexpr int LPG = 32 / NG;       // lanes per code
    __shared__ int s_codes[8][64];
    __shared__ int s_cnt[8];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, g = lane / LPG, l = lane % LPG;
    const float emax = __ldg(emax_p);
    const int n_list = counters[4];
    int n_resc = 0;
    // The per-row chain is list entry -> entry counts -> entries -> z / E rows: four dependent round trips for a few
    // hundred flops.  The first three of the NEXT listed row are issued before the current row's distances are
    // computed (registers), so a warp pays ~two round trips per row instead of four.
    const int li_step = gridDim.x * 8;
    auto fetch_meta = [&](int li, int64_t &row, int (&cnt)[2]) {
        row = li < n_list ? (int64_t)resc_list[li] : -1;
        cnt[0] = cnt[1] = 0;
        if (row >= 0) {
            const int64_t tile = row >> 7;
            const int r = (int)(row & 127);
#pragma unroll
            for (int h = 0; h < 2; ++h) cnt[h] = __ldg(cand_meta + (tile * NPART + 2 * h + (lane >> 4)) * BM + r);
        }
    };
    auto fetch_ent = [&](int64_t row, const int (&cnt)[2], uint2 (&ent)[2]) {
        ent[0] = ent[1] = make_uint2(0u, 0u);
        if (row >= 0) {
            const int64_t tile = row >> 7;
            const int r = (int)(row & 127), e = lane & 15;
#pragma unroll
            for (int h = 0; h < 2; ++h)
                if (e < cnt[h]) ent[h] = __ldg(cand + ((tile * NPART + 2 * h + (lane >> 4)) * LCAP + e) * BM + r);
        }
    };
    int64_t row_n;
    int cnt_n[2];
    uint2 ent_n[2];
    fetch_meta(blockIdx.x * 8 + wib, row_n, cnt_n);
    fetch_ent(row_n, cnt_n, ent_n);
    for (int li = blockIdx.x * 8 + wib; li < n_list; li += li_step) {
        const int64_t row = row_n;
        // lanes 0..15 / 16..31 hold the entries of parts (0,1) then (2,3)
        uint2 ent[2] = {ent_n[0], ent_n[1]};
        bool has_e[2] = {(lane & 15) < cnt_n[0], (lane & 15) < cnt_n[1]};
        fetch_meta(li + li_step, row_n, cnt_n);          // next row: in flight during this row's work
        float gb = BIG;
#pragma unroll
        for (int h = 0; h < 2; ++h)
            if (has_e[h]) gb = fminf(gb, __uint_as_float(ent[h].x));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) gb = fminf(gb, __shfl_xor_sync(0xffffffffu, gb, o));
        const float T = gb + tau_of(__ldg(zn + row), emax, abs_c);
        if (lane == 0) s_cnt[wib] = 0;
        __syncwarp();
#pragma unroll
        for (int h = 0; h < 2; ++h)
            if (has_e[h] && __uint_as_float(ent[h].x) <= T) {
                const int c0 = (int)(ent[h].y & 0xffffffu);
                uint32_t m = ent[h].y >> 24;
                if (c0 + 8 > K) m &= (1u << max(K - c0, 0)) - 1u;
                while (m) {
                    const int b = __ffs(m) - 1;
                    m &= m - 1;
                    const int pos = atomicAdd(&s_cnt[wib], 1);
                    if (pos < 64) s_codes[wib][pos] = c0 + b;
                }
            }
        __syncwarp();
        if (s_cnt[wib] > 64) {                   // too many listed elements for the warp list: the exact scan takes the row
            if (lane == 0) exact_list[atomicAdd(counters + 0, 1)] = (int)row;
            __syncwarp();
            fetch_ent(row_n, cnt_n, ent_n);
            continue;
        }
        const int nc = s_cnt[wib];
        fetch_ent(row_n, cnt_n, ent_n);
        float best = INFINITY;
        int bidx = 0x7fffffff;
        const float *zr = z + row * D;
        for (int p = 0; p * NG < nc; ++p) {
            const int e = p * NG + g;
            const bool has = e < nc;
            const int cj = has ? s_codes[wib][e] : 0;
            float s = 0.f;
            if (has) {
                const float *ek = E + (size_t)cj * D;
                if (VEC) {
                    const float4 *z4 = reinterpret_cast<const float4 *>(zr), *e4 = reinterpret_cast<const float4 *>(ek);
                    const int d4 = D >> 2;
                    for (int i0 = l; i0 < d4; i0 += 4 * LPG) {   // 4 independent load pairs in flight
                        float4 a[4], b[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int i = i0 + u * LPG;
                            a[u] = b[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                            if (i < d4) { a[u] = __ldg(z4 + i); b[u] = __ldg(e4 + i); }
                        }
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const float dx = a[u].x - b[u].x, dy = a[u].y - b[u].y, dz = a[u].z - b[u].z, dw = a[u].w - b[u].w;
                            s = fmaf(dx, dx, s); s = fmaf(dy, dy, s); s = fmaf(dz, dz, s); s = fmaf(dw, dw, s);
                        }
                    }
                } else {
                    for (int i = l; i < D; i += LPG) { const float df = __ldg(zr + i) - __ldg(ek + i); s = fmaf(df, df, s); }
                }
            }
#pragma unroll
            for (int o = LPG / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            float cs = has ? s : INFINITY;
            int ci = has ? cj : 0x7fffffff;
#pragma unroll
            for (int o = LPG; o < 32; o <<= 1) {
                const float os = __shfl_xor_sync(0xffffffffu, cs, o);
                const int oi = __shfl_xor_sync(0xffffffffu, ci, o);
                argmin_merge(cs, ci, os, oi);
            }
            argmin_merge(best, bidx, cs, ci);
            n_resc += min(NG, nc - p * NG);
        }
        if (lane == 0) {
            idx_out[row] = (int64_t)bidx;
            if (d2_out) d2_out[row] = best;
        }
        __syncwarp();
    }
    if (lane == 0 && n_resc) atomicAdd(counters + 1, n_resc);
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}
// bf16 [rows, cols] row-major, box = {64 cols, box_rows}, 128-byte swizzle
// pitch_cols (default = cols): row pitch of the array in elements when only the first `cols` columns are exposed -- the
// rest of a box is then out of bounds and zero-filled by the TMA unit instead of being fetched
inline int make_tmap(CUtensorMap *m, const void *ptr, uint64_t rows, uint64_t cols, uint32_t box_rows,
                     uint64_t pitch_cols = 0) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return LIPVQ_ECUDA;
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {(pitch_cols ? pitch_cols : cols) * 2};
    cuuint32_t box[2] = {(cuuint32_t)KBLK, box_rows};
    cuuint32_t estr[2] = {1, 1};
#ifdef LIPVQ_VQ_BF16
    const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
#else
    const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
#endif
    CUresult r = fn(m, dt, 2, const_cast<void *>(ptr), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? LIPVQ_OK : LIPVQ_ECUDA;
}

struct Workspace {
    float *mu, *enorm, *emax, *zn;
    uint2 *cand;
    int *cand_meta, *list, *resc_list, *counters;
    op_t *Eb, *Zb;
    size_t bytes;
};
inline Workspace carve(void *ws, const Geometry &g) {
    Workspace w;
    size_t off = 0;
    auto take = [&](size_t b) { off = align_up(off, 1024); size_t o = off; off += b; return (char *)ws + o; };
    w.mu = (float *)take(sizeof(float) * g.D);
    w.enorm = (float *)take(sizeof(float) * g.K);
    w.counters = (int *)take(sizeof(int) * 8);
    w.emax = (float *)(w.counters + 6);          // zeroed together with the counters, filled by atomicMax
    w.Eb = (op_t *)take((size_t)g.Kp * g.DKp * 2);
    w.Zb = (op_t *)take((size_t)g.Np * g.DKz * 2);
    w.zn = (float *)take(sizeof(float) * g.Np);
    w.cand = (uint2 *)take(sizeof(uint2) * g.Np * NPART * LCAP);
    w.cand_meta = (int *)take(sizeof(int) * g.Np * NPART);
    w.list = (int *)take(sizeof(int) * g.Np);
    w.resc_list = (int *)take(sizeof(int) * g.Np);
    w.bytes = align_up(off, 1024);
    return w;
}

// PAIR launch: clusters of two CTAs, an even grid; tb must have 128-row boxes
template <int NKB, int A_BUFS, int STAGES>
inline int launch_tc_pair(const CUtensorMap &ta, const CUtensorMap &tb_half, const Geometry &g, const Workspace &w,
                          int64_t *idx, bool want_d2, cudaStream_t st) {
    using L = SmemLayout<NKB, A_BUFS, STAGES, true>;
    auto kern = vq_tc_kernel<NKB, A_BUFS, STAGES, false, true>;
    LIPVQ_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
    const int64_t pairs = (g.n_tiles + 1) / 2;
    int grid = 2 * (int)(pairs < kNumSMs / 2 ? pairs : kNumSMs / 2);
    static const int env_grid = [] { const char *e = getenv("LIPVQ_VQ_GRID"); return e ? atoi(e) : 0; }();   // A/B probe
    if (env_grid > 1 && env_grid < grid) grid = env_grid & ~1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(NT);
    cfg.dynamicSmemBytes = L::TOTAL;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    ProfScope prof("vq_tc_kernel", st);
    LIPVQ_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, ta, tb_half, (int64_t)g.n, (int)g.n_tiles, g.nct, g.NK16, (const float *)w.zn,
                                     (const float *)w.emax, OP_ABS * sqrtf((float)g.D), w.cand, w.cand_meta, g.K, idx,
                                     want_d2 ? 1 : 0, w.resc_list, w.list, w.counters, (const float *)nullptr,
                                     (const float *)w.mu, g.D, w.zn));
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

template <int NKB, int A_BUFS, int STAGES, bool CONV = false>
inline int launch_tc(const CUtensorMap &ta, const CUtensorMap &tb, const Geometry &g, const Workspace &w,
                     int64_t *idx, bool want_d2, cudaStream_t st, const float *zsrc = nullptr) {
    using L = SmemLayout<NKB, A_BUFS, STAGES>;
    auto kern = vq_tc_kernel<NKB, A_BUFS, STAGES, CONV>;
    // (the attribute is per device and the call is cheap: set it on every launch, as set_smem in mlp_tc.cuh does)
    LIPVQ_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
    int grid = (int)(g.n_tiles < kNumSMs ? g.n_tiles : kNumSMs);
    static const int env_grid = [] { const char *e = getenv("LIPVQ_VQ_GRID"); return e ? atoi(e) : 0; }();   // A/B probe
    if (env_grid > 0 && env_grid < grid) grid = env_grid;
    ProfScope prof("vq_tc_kernel", st);
    kern<<<grid, CONV ? NT_CONV : NT, L::TOTAL, st>>>(ta, tb, g.n, (int)g.n_tiles, g.nct, g.NK16, w.zn, w.emax,
                                                      OP_ABS * sqrtf((float)g.D), w.cand, w.cand_meta, g.K, idx,
                                                      want_d2 ? 1 : 0, w.resc_list, w.list, w.counters, zsrc, w.mu, g.D,
                                                      w.zn);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

}  // namespace tc

inline bool vq_tensor_supported(int D, int K) { return D >= 1 && D + 3 <= 5 * tc::KBLK && K >= 1; }
inline size_t vq_tensor_ws_bytes(int64_t n, int D, int K) {
    if (!vq_tensor_supported(D, K)) return 256;
    return tc::carve(nullptr, tc::geometry(n > 0 ? n : 1, D, K)).bytes + 1024;
}

// pass 0: codebook operand (mu, Eb, enorm, emax) into the workspace; zeroes the counters
inline int vq_tensor_prep(const float *E, const tc::Geometry &g, const tc::Workspace &w, cudaStream_t st) {
    using namespace tc;
    ProfScope prof("vq_prep (2 kernels)", st);
    LIPVQ_CUDA_OK(cudaMemsetAsync(w.counters, 0, 8 * sizeof(int), st));
    vq_prep_mean_kernel<<<(g.D + 31) / 32, 1024, 0, st>>>(E, g.K, g.D, w.mu);
    LIPVQ_LAUNCH_OK();
    vq_prep_rows_kernel<<<(g.Kp + 7) / 8, 256, 0, st>>>(E, w.mu, g.K, g.Kp, g.D, g.DKp, w.Eb, w.enorm, w.counters,
                                                        w.emax);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

// passes 2-3 (+ exact tier) on operands already in the workspace (Zb, zn written by pass 1 or by the encoder)
// conv: the row operand is built inside the search kernel from the fp32 rows z (standalone API); otherwise it is
// read from the fp16 image w.Zb / w.zn that the fused encoder (or vq_convert_z) wrote
inline int vq_tensor_search(const float *z, const float *E, const tc::Geometry &g, const tc::Workspace &w,
                            int64_t *idx, float *d2, int32_t *stats, cudaStream_t st, bool conv = false) {
    using namespace tc;
    const int64_t n = g.n;
    const int D = g.D, K = g.K;
    CUtensorMap ta, tb;
    // only the DKz columns the MMA K steps read are exposed: the padding of the last 64-column block is zero-filled
    // by the TMA unit, not streamed from L2 (the kernel is bound by the bytes each SM ingests per accumulator)
    LIPVQ_TRY(make_tmap(&tb, w.Eb, (uint64_t)g.Kp, (uint64_t)g.DKz, BN, (uint64_t)g.DKp));
    if (conv) ta = tb;                    // (unused by the kernel in this mode)
    else LIPVQ_TRY(make_tmap(&ta, w.Zb, (uint64_t)g.Np, (uint64_t)g.DKz, BM));
    int rc = LIPVQ_ESHAPE;
    if (conv) switch (g.NKB) {
        case 1: rc = launch_tc<1, 2, 6, true>(ta, tb, g, w, idx, d2 != nullptr, st, z); break;
        case 2: rc = launch_tc<2, 2, 5, true>(ta, tb, g, w, idx, d2 != nullptr, st, z); break;
        case 3: rc = launch_tc<3, 2, 4, true>(ta, tb, g, w, idx, d2 != nullptr, st, z); break;
        case 4: rc = launch_tc<4, 2, 3, true>(ta, tb, g, w, idx, d2 != nullptr, st, z); break;
        case 5: rc = launch_tc<5, 1, 4, true>(ta, tb, g, w, idx, d2 != nullptr, st, z); break;
        default: break;
    }
    static const int pair_env = [] { const char *e = getenv("LIPVQ_VQ_PAIR"); return e ? atoi(e) : -1; }();   // A/B probe: 0 off, 1 always
    const bool pair = !conv && g.n_tiles >= 2 && pair_env == 1;   // opt-in: measured no faster, see vq_tc_kernel
    if (pair) {
        CUtensorMap tbh;
        LIPVQ_TRY(make_tmap(&tbh, w.Eb, (uint64_t)g.Kp, (uint64_t)g.DKz, BN / 2, (uint64_t)g.DKp));
        switch (g.NKB) {
            case 1: rc = launch_tc_pair<1, 2, 6>(ta, tbh, g, w, idx, d2 != nullptr, st); break;
            case 2: rc = launch_tc_pair<2, 2, 5>(ta, tbh, g, w, idx, d2 != nullptr, st); break;
            case 3: rc = launch_tc_pair<3, 2, 4>(ta, tbh, g, w, idx, d2 != nullptr, st); break;
            case 4: rc = launch_tc_pair<4, 2, 3>(ta, tbh, g, w, idx, d2 != nullptr, st); break;
            case 5: rc = launch_tc_pair<5, 1, 4>(ta, tbh, g, w, idx, d2 != nullptr, st); break;
            default: break;
        }
    } else if (conv) {
    } else switch (g.NKB) {
        case 1: rc = launch_tc<1, 2, 6>(ta, tb, g, w, idx, d2 != nullptr, st); break;
        case 2: rc = launch_tc<2, 2, 5>(ta, tb, g, w, idx, d2 != nullptr, st); break;
        case 3: rc = launch_tc<3, 2, 4>(ta, tb, g, w, idx, d2 != nullptr, st); break;
        case 4: {
            static const bool a1 = [] { const char *e = getenv("LIPVQ_VQ_RING"); return e && !strcmp(e, "a1b5"); }();   // A/B probe
            rc = a1 ? launch_tc<4, 1, 5>(ta, tb, g, w, idx, d2 != nullptr, st) : launch_tc<4, 2, 3>(ta, tb, g, w, idx, d2 != nullptr, st);
            break;
        }
        case 5: rc = launch_tc<5, 1, 4>(ta, tb, g, w, idx, d2 != nullptr, st); break;
        default: break;
    }
    LIPVQ_TRY(rc);
    const float abs_c = OP_ABS * sqrtf((float)D);
    {
        int64_t blocks = (n + 7) / 8;
        if (blocks > 8 * kNumSMs) blocks = 8 * kNumSMs;   // grid-stride over the device-side list length
        ProfScope prof("vq_rescore_kernel", st);
        if ((D & 3) == 0)
            vq_rescore_kernel<4, true><<<(unsigned)blocks, 256, 0, st>>>(z, E, D, K, w.cand, w.cand_meta, w.zn, w.emax,
                                                                        abs_c, idx, d2, w.resc_list, w.counters, w.list);
        else
            vq_rescore_kernel<4, false><<<(unsigned)blocks, 256, 0, st>>>(z, E, D, K, w.cand, w.cand_meta, w.zn, w.emax,
                                                                         abs_c, idx, d2, w.resc_list, w.counters, w.list);
        LIPVQ_LAUNCH_OK();
    }
    // tier 2: rows whose candidate list overflowed -> exact fp32 scan (count stays on the device)
    LIPVQ_TRY(vq_exact_launch(z, E, n, D, K, w.list, w.counters, idx, d2, st));
    if (stats) LIPVQ_CUDA_OK(cudaMemcpyAsync(stats, w.counters, 4 * sizeof(int), cudaMemcpyDeviceToDevice, st));
    return LIPVQ_OK;
}

inline int vq_tensor_launch(const float *z, const float *E, int64_t n, int D, int K, int64_t *idx, float *d2,
                            int32_t *stats, void *ws, cudaStream_t st) {
    using namespace tc;
    if (n > (int64_t)0x7fffffff - BM) return LIPVQ_ESHAPE;
    const Geometry g = geometry(n, D, K);
    const Workspace w = carve((void *)align_up((size_t)ws, 1024), g);
    LIPVQ_TRY(vq_tensor_prep(E, g, w, st));
    // LIPVQ_VQ_CONVERT=fused builds the row operand inside the search kernel (converter warps) instead of the separate
    // conversion pass below.  Measured at N=2^20, D=64, K=1024 (profiles/r02_*): pass 0.118 + search 0.33 ms against
    // 0.49 ms (thread-per-row converter, 4 extra warps) or 1.4 ms (coalesced converter on the 2 spare warps: latency
    // bound) for the fused kernel -- the register budget of the 16 epilogue warps leaves no room for enough converter
    // warps, so the separate pass stays the default.
    static const bool fused = [] { const char *e = getenv("LIPVQ_VQ_CONVERT"); return e && !strcmp(e, "fused"); }();
    if (fused) return vq_tensor_search(z, E, g, w, idx, d2, stats, st, true);
    {
        const int cpr = g.DKz >> 3;
        const int rpb = cpr <= 32 ? 8 * (32 / cpr) : 256 / cpr;
        int64_t blocks = (g.Np + rpb - 1) / rpb;
        if (blocks > 8 * kNumSMs) blocks = 8 * kNumSMs;
        ProfScope prof("vq_convert_z_kernel", st);
        if (cpr <= 32)
            vq_convert_z_warp_kernel<<<(unsigned)blocks, 256, 0, st>>>(z, w.mu, n, g.Np, D, g.DKz, w.Zb, w.zn, w.counters);
        else if (cpr <= 64)
            vq_convert_z_wide_kernel<<<(unsigned)(8 * kNumSMs), 256, 0, st>>>(z, w.mu, n, g.Np, D, g.DKz, w.Zb, w.zn, w.counters);
        else
            vq_convert_z_kernel<<<(unsigned)blocks, 256, 0, st>>>(z, w.mu, n, g.Np, D, g.DKz, w.Zb, w.zn, w.counters);
        LIPVQ_LAUNCH_OK();
    }
    return vq_tensor_search(z, E, g, w, idx, d2, stats, st);
}

}  // namespace lipvq
