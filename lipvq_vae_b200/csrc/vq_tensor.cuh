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
constexpr int NT = 384;          // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warps 4..11 epilogue
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
                                    int *__restrict__ counters) {
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
                                                           op_t *__restrict__ Zb, float *__restrict__ zn,
                                                           int *__restrict__ counters) {
    const int lane = threadIdx.x & 31;
    for (int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); r < Np; r += (int64_t)gridDim.x * 8) {
        op_t *row = Zb + r * DKp;
        float s = 0.f, amax = 0.f;
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
                amax = fmaxf(amax, fmaxf(fabsf(v0), fabsf(v1)));
            }
            *reinterpret_cast<op2_t *>(row + d) = f2op2(v0, v1);
        }
        s = warp_sum(s);
        if (!(amax <= OP_MAX)) atomicOr(counters + 3, 1);   // operand out of range -> whole launch goes exact
        if (lane == 0 && r < n) zn[r] = sqrtf(s);
    }
}


// ------------------------------------------------------------------ pass 2: tcgen05 GEMM + fused chunk-min
// Shared memory (dynamic, 1024-B aligned): A[A_BUFS][NKB] 16 KB K-blocks | B[STAGES] 32 KB | barriers.
// All operand tiles are K-major with the 128-byte TMA/UMMA swizzle.
template <int NKB, int A_BUFS, int STAGES>
struct SmemLayout {
    static constexpr int A_BYTES = A_BUFS * NKB * A_KB_BYTES;
    static constexpr int B_BYTES = STAGES * B_ST_BYTES;
    static constexpr int BAR_OFF = A_BYTES + B_BYTES;
    static constexpr int NBARS = 2 * A_BUFS + 2 * STAGES + 4;
    static constexpr int TOTAL = BAR_OFF + NBARS * 8 + 16 + 1024;   // + tmem ptr + alignment slack
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

// sorted insert of x into m0<=m1<=m2<=m3 (payload rides in the low mantissa bits): 7 FMNMX
__device__ __forceinline__ void insert4(float (&m)[NCAND], float x) {
    float t;
    t = fminf(m[0], x); x = fmaxf(m[0], x); m[0] = t;
    t = fminf(m[1], x); x = fmaxf(m[1], x); m[1] = t;
    t = fminf(m[2], x); x = fmaxf(m[2], x); m[2] = t;
    m[3] = fminf(m[3], x);
}
// Candidate record written by the epilogue for each (row, column half):
//   entries[LCAP] = {packed score bits, code}: every chunk minimum that was within tau(row) of the running
//                   row minimum when its code tile was finished (a superset of the final window);
//   meta = {count, m4min, smin}: m4min = smallest 4th-best chunk minimum of any code tile (if it is inside
//          the final window a 5th chunk may have been dropped -> exact scan); smin = smallest SECOND-best
//          element of any chunk (ELEM only: if it is outside the window every window element is listed).
// ELEM (CHUNK==4): the packed payload is the column inside the tile half (7 bits) -> entries name single codes.
// otherwise the payload is the chunk id and an entry names CHUNK consecutive codes.
template <int NKB, int A_BUFS, int STAGES, int CHUNK, bool ELEM>
__global__ void __launch_bounds__(NT, 1) vq_tc_kernel(const __grid_constant__ CUtensorMap tmapA,
                                                       const __grid_constant__ CUtensorMap tmapB,
                                                       int64_t n_rows, int n_tiles, int nct, int nk16,
                                                       const float *__restrict__ zn, const float *__restrict__ emax_p,
                                                       float abs_c, uint2 *__restrict__ cand,
                                                       float4 *__restrict__ cand_meta) {
    using L = SmemLayout<NKB, A_BUFS, STAGES>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = ptx::smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t *gen = smem_raw + (base - raw);
    const uint32_t sA = base, sB = base + L::A_BYTES, sBar = base + L::BAR_OFF;
    // barrier slots
    auto a_full = [&](int i) { return sBar + 8u * i; };
    auto a_empty = [&](int i) { return sBar + 8u * (A_BUFS + i); };
    auto b_full = [&](int i) { return sBar + 8u * (2 * A_BUFS + i); };
    auto b_empty = [&](int i) { return sBar + 8u * (2 * A_BUFS + STAGES + i); };
    auto t_full = [&](int i) { return sBar + 8u * (2 * A_BUFS + 2 * STAGES + i); };
    auto t_empty = [&](int i) { return sBar + 8u * (2 * A_BUFS + 2 * STAGES + 2 + i); };
    volatile uint32_t *tmem_ptr = reinterpret_cast<volatile uint32_t *>(gen + L::BAR_OFF + L::NBARS * 8);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        ptx::tma_prefetch_desc(&tmapA);
        ptx::tma_prefetch_desc(&tmapB);
        for (int i = 0; i < A_BUFS; ++i) { ptx::mbar_init(a_full(i), 1); ptx::mbar_init(a_empty(i), 1); }
        for (int i = 0; i < STAGES; ++i) { ptx::mbar_init(b_full(i), 1); ptx::mbar_init(b_empty(i), 1); }
        for (int i = 0; i < 2; ++i) { ptx::mbar_init(t_full(i), 1); ptx::mbar_init(t_empty(i), 8); }
        ptx::fence_barrier_init();
    }
    if (warp == 2) ptx::tmem_alloc<512>(ptx::smem_u32((const void *)tmem_ptr));
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            int st = 0, sph = 0, ab = 0, aph = 0;
            auto load_a = [&](int tile) {
                ptx::mbar_wait(a_empty(ab), aph ^ 1);
                ptx::mbar_arrive_expect_tx(a_full(ab), NKB * A_KB_BYTES);
#pragma unroll
                for (int kb = 0; kb < NKB; ++kb)
                    ptx::tma_load_2d(&tmapA, a_full(ab), sA + (ab * NKB + kb) * A_KB_BYTES, kb * KBLK, tile * BM);
                if (++ab == A_BUFS) { ab = 0; aph ^= 1; }
            };
            if (A_BUFS > 1 && (int)blockIdx.x < n_tiles) load_a(blockIdx.x);
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                // double-buffered rows: fetch the NEXT tile's rows before streaming this tile's codebook
                if (A_BUFS > 1) { if (tile + (int)gridDim.x < n_tiles) load_a(tile + gridDim.x); }
                else load_a(tile);
                for (int c = 0; c < nct; ++c)
                    for (int kb = 0; kb < NKB; ++kb) {
                        ptx::mbar_wait(b_empty(st), sph ^ 1);
                        ptx::mbar_arrive_expect_tx(b_full(st), B_ST_BYTES);
                        ptx::tma_load_2d(&tmapB, b_full(st), sB + st * B_ST_BYTES, kb * KBLK, c * BN);
                        if (++st == STAGES) { st = 0; sph ^= 1; }
                    }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (one thread) =====
        if (lane == 0) {
            #ifdef LIPVQ_VQ_BF16
            constexpr uint32_t idesc = ptx::umma_idesc_bf16(BM, BN);
#else
            constexpr uint32_t idesc = ptx::umma_idesc_f16(BM, BN);
#endif
            int st = 0, sph = 0, ab = 0, aph = 0, acc = 0, accph = 0;
            for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                ptx::mbar_wait(a_full(ab), aph);
                ptx::tc_fence_after();
                for (int c = 0; c < nct; ++c) {
                    ptx::mbar_wait(t_empty(acc), accph ^ 1);
                    ptx::tc_fence_after();
                    const uint32_t d_tmem = tmem_base + acc * BN;
                    for (int kb = 0; kb < NKB; ++kb) {
                        ptx::mbar_wait(b_full(st), sph);
                        ptx::tc_fence_after();
                        const uint64_t ad = ptx::umma_desc_sw128(sA + (ab * NKB + kb) * A_KB_BYTES);
                        const uint64_t bd = ptx::umma_desc_sw128(sB + st * B_ST_BYTES);
                        const int nk = min(KBLK / UK, nk16 - kb * (KBLK / UK));
                        for (int k = 0; k < nk; ++k)   // +32 B per K=16 slice inside the swizzle row
                            ptx::umma_bf16(d_tmem, ad + 2 * k, bd + 2 * k, idesc, (kb | k) ? 1u : 0u);
                        ptx::umma_commit(b_empty(st));
                        if (++st == STAGES) { st = 0; sph ^= 1; }
                    }
                    ptx::umma_commit(t_full(acc));
                    if (++acc == 2) { acc = 0; accph ^= 1; }
                }
                ptx::umma_commit(a_empty(ab));
                if (++ab == A_BUFS) { ab = 0; aph ^= 1; }
            }
        }
    } else if (warp >= 4) {
        // ===== epilogue: 8 warps; warp%4 = TMEM lane quarter, (warp-4)/4 = column half =====
        static_assert(!ELEM || CHUNK == 4, "element payload needs 4-wide chunks");
        const int quarter = warp & 3, half = (warp - 4) >> 2;
        const uint32_t lane_addr = ((uint32_t)(quarter * 32) << 16);
        int acc = 0, accph = 0;
        constexpr int CPB = 32 / CHUNK;               // chunks per 32-column batch
        constexpr uint32_t PMASK = ELEM ? 127u : 31u; // payload bits in the low mantissa
        const float emax = __ldg(emax_p);
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const int64_t row = (int64_t)tile * BM + quarter * 32 + lane;
            const bool row_ok = row < n_rows;
            const float znr = row_ok ? __ldg(zn + row) : 0.f;
            const float tau = tau_of(znr, emax, abs_c);
            uint2 *out = cand + (row * 2 + half) * LCAP;
            float gb = BIG, T = BIG, m4min = BIG, smin = BIG;
            int cnt = 0;
            for (int c = 0; c < nct; ++c) {
                ptx::mbar_wait(t_full(acc), accph);
                ptx::tc_fence_after();
                float m[NCAND] = {BIG, BIG, BIG, BIG};
                const uint32_t col0 = tmem_base + acc * BN + half * (BN / 2);
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    uint32_t r[32];
                    ptx::tmem_ld_32x32b_x32(lane_addr + col0 + b * 32, r);
                    ptx::tmem_ld_wait();
                    if (b == 3) {   // accumulator fully in registers: hand the TMEM stage back
                        ptx::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) ptx::mbar_arrive(t_empty(acc));
                    }
#pragma unroll
                    for (int j = 0; j < CPB; ++j) {
                        float v;
                        if (ELEM) {
                            // payload = column inside the tile half (b*32 + 4j + pos), constant per unrolled slot
                            const float e0 = __uint_as_float((r[4 * j + 0] & ~PMASK) | (uint32_t)(b * 32 + 4 * j + 0));
                            const float e1 = __uint_as_float((r[4 * j + 1] & ~PMASK) | (uint32_t)(b * 32 + 4 * j + 1));
                            const float e2 = __uint_as_float((r[4 * j + 2] & ~PMASK) | (uint32_t)(b * 32 + 4 * j + 2));
                            const float e3 = __uint_as_float((r[4 * j + 3] & ~PMASK) | (uint32_t)(b * 32 + 4 * j + 3));
                            const float lo1 = fminf(e0, e1), hi1 = fmaxf(e0, e1);
                            const float lo2 = fminf(e2, e3), hi2 = fmaxf(e2, e3);
                            v = fminf(lo1, lo2);
                            smin = ptx::fmin3(smin, fmaxf(lo1, lo2), fminf(hi1, hi2));   // chunk's 2nd smallest
                        } else if (CHUNK == 4) {
                            v = ptx::fmin3(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                           fminf(__uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3])));
                            v = __uint_as_float((__float_as_uint(v) & ~PMASK) | (uint32_t)(b * CPB + j));
                        } else {
                            const float v0 = ptx::fmin3(__uint_as_float(r[8 * j]), __uint_as_float(r[8 * j + 1]),
                                                        __uint_as_float(r[8 * j + 2]));
                            const float v1 = ptx::fmin3(__uint_as_float(r[8 * j + 3]), __uint_as_float(r[8 * j + 4]),
                                                        __uint_as_float(r[8 * j + 5]));
                            v = ptx::fmin3(v0, v1, fminf(__uint_as_float(r[8 * j + 6]), __uint_as_float(r[8 * j + 7])));
                            v = __uint_as_float((__float_as_uint(v) & ~PMASK) | (uint32_t)(b * CPB + j));
                        }
                        insert4(m, v);
                    }
                }
                if (++acc == 2) { acc = 0; accph ^= 1; }
                // keep the chunk minima of this code tile that are inside the window of the running row minimum
                m4min = fminf(m4min, m[NCAND - 1]);
                gb = fminf(gb, m[0]);
                T = gb + tau;
                const int code0 = c * BN + half * (BN / 2);
#pragma unroll
                for (int i = 0; i < NCAND; ++i) {
                    if (row_ok && m[i] <= T) {
                        const uint32_t bits = __float_as_uint(m[i]);
                        if (cnt == LCAP) {   // full: drop entries that fell out of the (shrinking) window
                            int k = 0;
                            for (int e = 0; e < LCAP; ++e) {
                                const uint2 x = out[e];
                                if (__uint_as_float(x.x) <= T) out[k++] = x;
                            }
                            cnt = k;
                        }
                        if (cnt < LCAP)
                            out[cnt] = make_uint2(bits, (uint32_t)(code0 + (int)(bits & PMASK) * (ELEM ? 1 : CHUNK)));
                        ++cnt;
                    }
                }
            }
            if (row_ok) cand_meta[row * 2 + half] = make_float4(__int_as_float(cnt), m4min, smin, 0.f);
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 2) ptx::tmem_dealloc<512>(tmem_base);
}


// ------------------------------------------------------------------ pass 3: fp32 direct-form rescore
// Rounding bound of pass 2 for row n and ANY code k (u = 2^-8, bf16 round-to-nearest on both operands,
// Cauchy-Schwarz; fp32 accumulation, the 3-way split of en and the 5 payload bits are covered by C2):
//   |S_nk - (||e'_k||^2 - 2 z'_n.e'_k)| <= eps_n = C1*zn*emax + C2*(2*zn*emax + emax^2)
// so every code whose exact score is minimal has S <= min_k S + 2*eps_n.

template <int CHUNK, bool ELEM, bool VEC>
__global__ void __launch_bounds__(256) vq_rescore_kernel(
    const float *__restrict__ z, const float *__restrict__ E, int64_t n_rows, int D, int K,
    const uint2 *__restrict__ cand, const float4 *__restrict__ cand_meta, const float *__restrict__ zn,
    const float *__restrict__ emax_p, int64_t *__restrict__ idx_out, float *__restrict__ d2_out,
    int *__restrict__ list, int *__restrict__ counters, float abs_c) {
    constexpr int NG = CHUNK;          // codes evaluated side by side
    constexpr int LPG = 32 / NG;       // lanes per code
    const int lane = threadIdx.x & 31, g = lane / LPG, l = lane % LPG;
    const float emax = __ldg(emax_p);
    const bool range_bad = counters[3] != 0;   // set by the operand passes, which finished before this launch
    int n_resc = 0, n_fast = 0;
    for (int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); row < n_rows; row += (int64_t)gridDim.x * 8) {
        const float4 m0 = __ldg(cand_meta + row * 2), m1 = __ldg(cand_meta + row * 2 + 1);
        const int cnt0 = __float_as_int(m0.x), cnt1 = __float_as_int(m1.x);
        const int n0 = min(cnt0, LCAP), n1 = min(cnt1, LCAP);
        float v = BIG;
        int code = 0x7fffffff;
        if (lane < n0 + n1) {
            const uint2 e = __ldg(cand + (row * 2 + (lane < n0 ? 0 : 1)) * LCAP + (lane < n0 ? lane : lane - n0));
            v = __uint_as_float(e.x);
            code = (int)e.y;
        }
        float gb = v;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) gb = fminf(gb, __shfl_xor_sync(0xffffffffu, gb, o));
        const float znr = __ldg(zn + row);
        const float T = gb + tau_of(znr, emax, abs_c);
        if (range_bad || cnt0 > LCAP || cnt1 > LCAP || m0.y <= T || m1.y <= T) {
            // the candidate record may be incomplete -> exact fp32 scan of the whole codebook
            if (lane == 0) list[atomicAdd(counters + 0, 1)] = (int)row;
            continue;
        }
        const bool in = (v <= T) && (code < K);
        unsigned mask = __ballot_sync(0xffffffffu, in);
        const int nc = __popc(mask);
        const bool elem_ok = ELEM && !(m0.z <= T || m1.z <= T);   // no chunk holds two window elements
        if (elem_ok && nc == 1 && d2_out == nullptr) {
            // a single element inside the rounding window IS the nearest code: nothing to rescore
            const int w = __shfl_sync(0xffffffffu, code, __ffs(mask) - 1);
            if (lane == 0) idx_out[row] = (int64_t)w;
            ++n_fast;
            continue;
        }
        float best = INFINITY;
        int bidx = 0x7fffffff;
        const float *zr = z + row * D;
        const int passes = elem_ok ? (nc + NG - 1) / NG : nc;
        for (int p = 0; p < passes; ++p) {
            int cj;
            bool has;
            if (elem_ok) {
                const int e = p * NG + g;
                has = e < nc;
                const int src = __fns(mask, 0, (has ? e : 0) + 1);
                cj = __shfl_sync(0xffffffffu, code, src);
            } else {
                const int src = __fns(mask, 0, p + 1);
                int c0 = __shfl_sync(0xffffffffu, code, src);
                if (ELEM) c0 &= ~(CHUNK - 1);
                cj = c0 + g;
                has = cj < K;
            }
            float s = 0.f;
            if (has) {
                const float *e = E + (size_t)cj * D;
                if (VEC) {
                    // issue 4 independent (z,e) load pairs before using them: the loop is latency-bound
                    const float4 *z4 = reinterpret_cast<const float4 *>(zr), *e4 = reinterpret_cast<const float4 *>(e);
                    const int d4 = D >> 2;
                    for (int i0 = l; i0 < d4; i0 += 4 * LPG) {
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
                    for (int i0 = l; i0 < D; i0 += 4 * LPG) {
                        float a[4], b[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int i = i0 + u * LPG;
                            a[u] = b[u] = 0.f;
                            if (i < D) { a[u] = __ldg(zr + i); b[u] = __ldg(e + i); }
                        }
#pragma unroll
                        for (int u = 0; u < 4; ++u) { const float df = a[u] - b[u]; s = fmaf(df, df, s); }
                    }
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
            n_resc += elem_ok ? min(NG, nc - p * NG) : NG;
        }
        if (lane == 0) {
            idx_out[row] = (int64_t)bidx;
            if (d2_out) d2_out[row] = best;
        }
    }
    if (lane == 0) {
        if (n_resc) atomicAdd(counters + 1, n_resc);
        if (n_fast) atomicAdd(counters + 2, n_fast);
    }
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
inline int make_tmap(CUtensorMap *m, const void *ptr, uint64_t rows, uint64_t cols, uint32_t box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return LIPVQ_ECUDA;
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * 2};
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
    float4 *cand_meta;
    int *list, *counters;
    op_t *Eb, *Zb;
    size_t bytes;
};
inline Workspace carve(void *ws, const Geometry &g) {
    Workspace w;
    size_t off = 0;
    auto take = [&](size_t b) { off = align_up(off, 1024); size_t o = off; off += b; return (char *)ws + o; };
    w.mu = (float *)take(sizeof(float) * g.D);
    w.enorm = (float *)take(sizeof(float) * g.K);
    w.emax = (float *)take(sizeof(float) * 4);
    w.counters = (int *)take(sizeof(int) * 4);
    w.Eb = (op_t *)take((size_t)g.Kp * g.DKp * 2);
    w.Zb = (op_t *)take((size_t)g.Np * g.DKp * 2);
    w.zn = (float *)take(sizeof(float) * g.Np);
    w.cand = (uint2 *)take(sizeof(uint2) * g.Np * 2 * LCAP);
    w.cand_meta = (float4 *)take(sizeof(float4) * g.Np * 2);
    w.list = (int *)take(sizeof(int) * g.Np);
    w.bytes = align_up(off, 1024);
    return w;
}

template <int NKB, int A_BUFS, int STAGES, int CHUNK, bool ELEM>
inline int launch_tc(const CUtensorMap &ta, const CUtensorMap &tb, const Geometry &g, const Workspace &w,
                     cudaStream_t st) {
    using L = SmemLayout<NKB, A_BUFS, STAGES>;
    auto kern = vq_tc_kernel<NKB, A_BUFS, STAGES, CHUNK, ELEM>;
    static bool attr_done = false;
    if (!attr_done) {
        LIPVQ_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
        attr_done = true;
    }
    const int grid = (int)(g.n_tiles < kNumSMs ? g.n_tiles : kNumSMs);
    kern<<<grid, NT, L::TOTAL, st>>>(ta, tb, g.n, (int)g.n_tiles, g.nct, g.NK16, w.zn, w.emax,
                                     OP_ABS * sqrtf((float)g.D), w.cand, w.cand_meta);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

}  // namespace tc

inline bool vq_tensor_supported(int D, int K) { return D >= 1 && D + 3 <= 5 * tc::KBLK && K >= 1; }
inline size_t vq_tensor_ws_bytes(int64_t n, int D, int K) {
    if (!vq_tensor_supported(D, K)) return 256;
    return tc::carve(nullptr, tc::geometry(n > 0 ? n : 1, D, K)).bytes + 1024;
}

inline int vq_tensor_launch(const float *z, const float *E, int64_t n, int D, int K, int64_t *idx, float *d2,
                            int32_t *stats, void *ws, cudaStream_t st) {
    using namespace tc;
    if (n > (int64_t)0x7fffffff - BM) return LIPVQ_ESHAPE;
    const Geometry g = geometry(n, D, K);
    void *ws_al = (void *)align_up((size_t)ws, 1024);
    const Workspace w = carve(ws_al, g);
    LIPVQ_CUDA_OK(cudaMemsetAsync(w.counters, 0, 4 * sizeof(int), st));
    vq_prep_mean_kernel<<<(D + 31) / 32, 1024, 0, st>>>(E, K, D, w.mu);
    LIPVQ_LAUNCH_OK();
    vq_prep_rows_kernel<<<(g.Kp + 7) / 8, 256, 0, st>>>(E, w.mu, K, g.Kp, D, g.DKp, w.Eb, w.enorm, w.counters);
    LIPVQ_LAUNCH_OK();
    vq_prep_emax_kernel<<<1, 256, 0, st>>>(w.enorm, K, w.emax);
    LIPVQ_LAUNCH_OK();
    {
        int64_t blocks = (g.Np + 7) / 8;
        if (blocks > 16 * kNumSMs) blocks = 16 * kNumSMs;
        vq_convert_z_kernel<<<(unsigned)blocks, 256, 0, st>>>(z, w.mu, n, g.Np, D, g.DKp, w.Zb, w.zn, w.counters);
        LIPVQ_LAUNCH_OK();
    }
    CUtensorMap ta, tb;
    LIPVQ_TRY(make_tmap(&ta, w.Zb, (uint64_t)g.Np, (uint64_t)g.DKp, BM));
    LIPVQ_TRY(make_tmap(&tb, w.Eb, (uint64_t)g.Kp, (uint64_t)g.DKp, BN));
    // variant: small D has little epilogue slack per MMA -> 8-wide chunks without element payload
    // (LIPVQ_TC_SMALL_D_ELEM=1 forces the element payload there too, for measurements)
    static const int elem_env = [] { const char *e = getenv("LIPVQ_TC_ELEM"); return e ? atoi(e) : -1; }();
    const bool elem = elem_env >= 0 ? elem_env != 0 : D > 96;
    const bool c8 = !elem && D <= 96;     // chunk-level records: 8-wide chunks for small D, 4-wide otherwise
    int rc = LIPVQ_ESHAPE;
#define LIPVQ_TC(NKB_, AB_, ST_)                                                                          \
    rc = elem ? launch_tc<NKB_, AB_, ST_, 4, true>(ta, tb, g, w, st)                                      \
              : (c8 ? launch_tc<NKB_, AB_, ST_, 8, false>(ta, tb, g, w, st)                               \
                    : launch_tc<NKB_, AB_, ST_, 4, false>(ta, tb, g, w, st))
    switch (g.NKB) {
        case 1: LIPVQ_TC(1, 2, 6); break;
        case 2: LIPVQ_TC(2, 2, 5); break;
        case 3: LIPVQ_TC(3, 2, 4); break;
        case 4: LIPVQ_TC(4, 2, 3); break;
        case 5: LIPVQ_TC(5, 1, 4); break;
        default: break;
    }
#undef LIPVQ_TC
    LIPVQ_TRY(rc);
    {
        int64_t blocks = (n + 7) / 8;
        if (blocks > 16 * kNumSMs) blocks = 16 * kNumSMs;
        const bool vec = (D & 3) == 0;
#define LIPVQ_RESC(CH, EL, VE)                                                                                   \
    vq_rescore_kernel<CH, EL, VE><<<(unsigned)blocks, 256, 0, st>>>(z, E, n, D, K, w.cand, w.cand_meta, w.zn, w.emax, \
                                                                    idx, d2, w.list, w.counters, OP_ABS * sqrtf((float)D))
        if (elem)    { if (vec) LIPVQ_RESC(4, true, true);   else LIPVQ_RESC(4, true, false); }
        else if (c8) { if (vec) LIPVQ_RESC(8, false, true);  else LIPVQ_RESC(8, false, false); }
        else         { if (vec) LIPVQ_RESC(4, false, true);  else LIPVQ_RESC(4, false, false); }
#undef LIPVQ_RESC
        LIPVQ_LAUNCH_OK();
    }
    // tier 2: rows whose candidate list overflowed -> exact fp32 scan (count stays on the device)
    LIPVQ_TRY(vq_exact_launch(z, E, n, D, K, w.list, w.counters, idx, d2, st));
    if (stats) LIPVQ_CUDA_OK(cudaMemcpyAsync(stats, w.counters, 4 * sizeof(int), cudaMemcpyDeviceToDevice, st));
    return LIPVQ_OK;
}

}  // namespace lipvq
