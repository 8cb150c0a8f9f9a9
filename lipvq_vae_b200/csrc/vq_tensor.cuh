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

// sorted insert of x into m0<=m1<=m2<=m3 (payload rides in the low mantissa bits): 7 FMNMX
__device__ __forceinline__ void insert4(float (&m)[NCAND], float x) {
    float t;
    t = fminf(m[0], x); x = fmaxf(m[0], x); m[0] = t;
    t = fminf(m[1], x); x = fmaxf(m[1], x); m[1] = t;
    t = fminf(m[2], x); x = fmaxf(m[2], x); m[2] = t;
    m[3] = fminf(m[3], x);
}
// sorted insert of (v,id) into the row-level list starting at level L0 (earlier levels are known smaller)
template <int L0>
__device__ __forceinline__ void insert4_payload(float (&G)[NCAND], int (&Gid)[NCAND], float v, int id) {
#pragma unroll
    for (int i = L0; i < NCAND; ++i) {
        const bool lt = v < G[i];
        const float gv = G[i];
        const int gi = Gid[i];
        G[i] = lt ? v : gv;
        Gid[i] = lt ? id : gi;
        v = lt ? gv : v;
        id = lt ? gi : id;
    }
}

template <int NKB, int A_BUFS, int STAGES, int CHUNK>
__global__ void __launch_bounds__(NT, 1) vq_tc_kernel(const __grid_constant__ CUtensorMap tmapA,
                                                       const __grid_constant__ CUtensorMap tmapB,
                                                       int64_t n_rows, int n_tiles, int nct, int nk16,
                                                       float *__restrict__ cand_val, int *__restrict__ cand_code) {
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
            constexpr uint32_t idesc = ptx::umma_idesc_bf16(BM, BN);
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
        const int quarter = warp & 3, half = (warp - 4) >> 2;
        const uint32_t lane_addr = ((uint32_t)(quarter * 32) << 16);
        int acc = 0, accph = 0;
        constexpr int CPB = 32 / CHUNK;               // chunks per 32-column batch
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            float G[NCAND];
            int Gid[NCAND];
#pragma unroll
            for (int i = 0; i < NCAND; ++i) { G[i] = BIG; Gid[i] = 0; }
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
                        if (CHUNK == 4) {
                            v = ptx::fmin3(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                           fminf(__uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3])));
                        } else {
                            const float v0 = ptx::fmin3(__uint_as_float(r[8 * j]), __uint_as_float(r[8 * j + 1]),
                                                        __uint_as_float(r[8 * j + 2]));
                            const float v1 = ptx::fmin3(__uint_as_float(r[8 * j + 3]), __uint_as_float(r[8 * j + 4]),
                                                        __uint_as_float(r[8 * j + 5]));
                            v = ptx::fmin3(v0, v1, fminf(__uint_as_float(r[8 * j + 6]), __uint_as_float(r[8 * j + 7])));
                        }
                        // chunk id (< 32) rides in the 5 low mantissa bits
                        const float pk = __uint_as_float((__float_as_uint(v) & ~31u) | (uint32_t)(b * CPB + j));
                        insert4(m, pk);
                    }
                }
                if (++acc == 2) { acc = 0; accph ^= 1; }
                // merge this tile-half's 4 minima into the row-level list (code index of the chunk start)
                const int code0 = c * BN + half * (BN / 2);
                insert4_payload<0>(G, Gid, m[0], code0 + (int)(__float_as_uint(m[0]) & 31u) * CHUNK);
                insert4_payload<1>(G, Gid, m[1], code0 + (int)(__float_as_uint(m[1]) & 31u) * CHUNK);
                insert4_payload<2>(G, Gid, m[2], code0 + (int)(__float_as_uint(m[2]) & 31u) * CHUNK);
                insert4_payload<3>(G, Gid, m[3], code0 + (int)(__float_as_uint(m[3]) & 31u) * CHUNK);
            }
            const int64_t row = (int64_t)tile * BM + quarter * 32 + lane;
            if (row < n_rows) {
                const int64_t o = (row * 2 + half) * NCAND;
                *reinterpret_cast<float4 *>(cand_val + o) = make_float4(G[0], G[1], G[2], G[3]);
                *reinterpret_cast<int4 *>(cand_code + o) = make_int4(Gid[0], Gid[1], Gid[2], Gid[3]);
            }
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
constexpr float RESC_C1 = 0.015625f * 1.004f;   // 2*(2u+u^2), u = 2^-8
constexpr float RESC_C2 = 4.0e-5f;
constexpr int RESC_MAXD = 320;

template <int CHUNK>
__global__ void __launch_bounds__(256) vq_rescore_kernel(
    const float *__restrict__ z, const float *__restrict__ E, int64_t n_rows, int D, int K,
    const float *__restrict__ cand_val, const int *__restrict__ cand_code, const float *__restrict__ zn,
    const float *__restrict__ emax_p, int64_t *__restrict__ idx_out, float *__restrict__ d2_out,
    int *__restrict__ list, int *__restrict__ counters) {
    const int lane = threadIdx.x & 31;
    const float emax = *emax_p;
    int n_resc = 0;
    for (int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); row < n_rows; row += (int64_t)gridDim.x * 8) {
        float v = BIG;
        int code = 0;
        if (lane < 2 * NCAND) { v = cand_val[row * 2 * NCAND + lane]; code = cand_code[row * 2 * NCAND + lane]; }
        float gb = v;
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) gb = fminf(gb, __shfl_xor_sync(0xffffffffu, gb, o));
        gb = __shfl_sync(0xffffffffu, gb, 0);
        const float znr = zn[row];
        const float eps = RESC_C1 * znr * emax + RESC_C2 * (2.f * znr * emax + emax * emax);
        const float T = gb + 2.f * eps + 1e-30f;
        const bool in = (lane < 2 * NCAND) && (v <= T) && (v < 1.0e29f) && (code < K);
        const unsigned over = __ballot_sync(0xffffffffu, in && ((lane & (NCAND - 1)) == NCAND - 1));
        if (over) {   // a 4th chunk is inside the window: more may have been dropped -> exact scan
            if (lane == 0) list[atomicAdd(counters + 0, 1)] = (int)row;
            continue;
        }
        unsigned mask = __ballot_sync(0xffffffffu, in);
        float zr[RESC_MAXD / 32];
#pragma unroll
        for (int i = 0; i < RESC_MAXD / 32; ++i) {
            const int d = lane + 32 * i;
            zr[i] = d < D ? __ldg(z + row * D + d) : 0.f;
        }
        float best = INFINITY;
        int bidx = 0x7fffffff;
        while (mask) {
            const int l = __ffs(mask) - 1;
            mask &= mask - 1;
            const int c0 = __shfl_sync(0xffffffffu, code, l);
#pragma unroll 1
            for (int j = 0; j < CHUNK; ++j) {
                const int cj = c0 + j;
                if (cj >= K) break;
                const float *e = E + (size_t)cj * D;
                float s = 0.f;
#pragma unroll
                for (int i = 0; i < RESC_MAXD / 32; ++i) {
                    const int d = lane + 32 * i;
                    if (d < D) {
                        const float df = zr[i] - __ldg(e + d);
                        s = fmaf(df, df, s);
                    }
                }
                s = warp_sum(s);
                argmin_merge(best, bidx, s, cj);
                ++n_resc;
            }
        }
        if (lane == 0) {
            idx_out[row] = (int64_t)bidx;
            if (d2_out) d2_out[row] = best;
        }
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
inline int make_tmap(CUtensorMap *m, const void *ptr, uint64_t rows, uint64_t cols, uint32_t box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return LIPVQ_ECUDA;
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * 2};
    cuuint32_t box[2] = {(cuuint32_t)KBLK, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(ptr), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? LIPVQ_OK : LIPVQ_ECUDA;
}

struct Workspace {
    float *mu, *enorm, *emax, *zn, *cand_val;
    int *cand_code, *list, *counters;
    __nv_bfloat16 *Eb, *Zb;
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
    w.Eb = (__nv_bfloat16 *)take((size_t)g.Kp * g.DKp * 2);
    w.Zb = (__nv_bfloat16 *)take((size_t)g.Np * g.DKp * 2);
    w.zn = (float *)take(sizeof(float) * g.Np);
    w.cand_val = (float *)take(sizeof(float) * g.Np * 2 * NCAND);
    w.cand_code = (int *)take(sizeof(int) * g.Np * 2 * NCAND);
    w.list = (int *)take(sizeof(int) * g.Np);
    w.bytes = align_up(off, 1024);
    return w;
}

template <int NKB, int A_BUFS, int STAGES, int CHUNK>
inline int launch_tc(const CUtensorMap &ta, const CUtensorMap &tb, const Geometry &g, const Workspace &w,
                     cudaStream_t st) {
    using L = SmemLayout<NKB, A_BUFS, STAGES>;
    auto kern = vq_tc_kernel<NKB, A_BUFS, STAGES, CHUNK>;
    static bool attr_done = false;
    if (!attr_done) {
        LIPVQ_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
        attr_done = true;
    }
    const int grid = (int)(g.n_tiles < kNumSMs ? g.n_tiles : kNumSMs);
    kern<<<grid, NT, L::TOTAL, st>>>(ta, tb, g.n, (int)g.n_tiles, g.nct, g.NK16, w.cand_val, w.cand_code);
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
    vq_prep_mean_kernel<<<(D + 127) / 128, 128, 0, st>>>(E, K, D, w.mu);
    LIPVQ_LAUNCH_OK();
    vq_prep_rows_kernel<<<(g.Kp + 7) / 8, 256, 0, st>>>(E, w.mu, K, g.Kp, D, g.DKp, w.Eb, w.enorm);
    LIPVQ_LAUNCH_OK();
    vq_prep_emax_kernel<<<1, 256, 0, st>>>(w.enorm, K, w.emax);
    LIPVQ_LAUNCH_OK();
    {
        int64_t blocks = (g.Np + 7) / 8;
        if (blocks > 16 * kNumSMs) blocks = 16 * kNumSMs;
        vq_convert_z_kernel<<<(unsigned)blocks, 256, 0, st>>>(z, w.mu, n, g.Np, D, g.DKp, w.Zb, w.zn);
        LIPVQ_LAUNCH_OK();
    }
    CUtensorMap ta, tb;
    LIPVQ_TRY(make_tmap(&ta, w.Zb, (uint64_t)g.Np, (uint64_t)g.DKp, BM));
    LIPVQ_TRY(make_tmap(&tb, w.Eb, (uint64_t)g.Kp, (uint64_t)g.DKp, BN));
    const bool c8 = D <= 96;
    int rc = LIPVQ_ESHAPE;
    switch (g.NKB) {
        case 1: rc = c8 ? launch_tc<1, 2, 6, 8>(ta, tb, g, w, st) : launch_tc<1, 2, 6, 4>(ta, tb, g, w, st); break;
        case 2: rc = c8 ? launch_tc<2, 2, 5, 8>(ta, tb, g, w, st) : launch_tc<2, 2, 5, 4>(ta, tb, g, w, st); break;
        case 3: rc = launch_tc<3, 2, 4, 4>(ta, tb, g, w, st); break;
        case 4: rc = launch_tc<4, 2, 3, 4>(ta, tb, g, w, st); break;
        case 5: rc = launch_tc<5, 1, 4, 4>(ta, tb, g, w, st); break;
        default: break;
    }
    LIPVQ_TRY(rc);
    {
        int64_t blocks = (n + 7) / 8;
        if (blocks > 16 * kNumSMs) blocks = 16 * kNumSMs;
        if (c8)
            vq_rescore_kernel<8><<<(unsigned)blocks, 256, 0, st>>>(z, E, n, D, K, w.cand_val, w.cand_code, w.zn, w.emax,
                                                                  idx, d2, w.list, w.counters);
        else
            vq_rescore_kernel<4><<<(unsigned)blocks, 256, 0, st>>>(z, E, n, D, K, w.cand_val, w.cand_code, w.zn, w.emax,
                                                                  idx, d2, w.list, w.counters);
        LIPVQ_LAUNCH_OK();
    }
    // tier 2: rows whose candidate list overflowed -> exact fp32 scan (count stays on the device)
    LIPVQ_TRY(vq_exact_launch(z, E, n, D, K, w.list, w.counters, idx, d2, st));
    if (stats) LIPVQ_CUDA_OK(cudaMemcpyAsync(stats, w.counters, 4 * sizeof(int), cudaMemcpyDeviceToDevice, st));
    return LIPVQ_OK;
}

}  // namespace lipvq
