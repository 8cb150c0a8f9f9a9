// bwd_umma.cuh -- the N-row GEMMs of the tokenizer backward (autograd of v5:71-72, invoked icl.py:969) on the
// 5th-gen tensor cores: tcgen05.mma kind::tf32, accumulators in TMEM, 3xTF32 split products
//   A*B ~= A_lo*B_hi + A_hi*B_lo + A_hi*B_hi   (hi = tf32-rounded value, lo = exact fp32 remainder)
// so the gradients keep fp32 parity (1e-5) with the reference's autograd.
//
//   wgrad_umma_kernel : part[cta][Nout][Kin+1] = [ dY[rows_cta, Nout]^T X[rows_cta, Kin] | column sums of dY ]
//   dgrad_umma_kernel : dX = scale * (dY W) (.) act'(pre)                       (second half of this file)
//
// The reduction axis of a weight gradient is the ROW axis, so both operands are "MN-major" for the tensor core:
// a row-major [rows, cols] slab IS the canonical MN-major operand (rows of 32 fp32 = 128 B, swizzled in 32-byte
// chunks, see desc_mn_sw128), no transposition anywhere -- TMA drops 16-row x 32-column boxes of dY and X straight
// into operand position.  The tensor core reads only the upper 19 bits of an fp32 operand (measured: feeding raw
// values and hi = rna(v) give the same 2e-6 error), so the raw image serves as the hi operand and only the lo
// image (v - trunc_tf32(v)) is computed, by 8 converter warps, shared memory to shared memory.  Every CTA walks
// its own contiguous row range, keeps the whole [Nout, Kin] partial in TMEM and writes it once; a fixed-order
// reduction over the CTAs (splitk_reduce_kernel) makes the result deterministic.
#pragma once
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace lipvq {
namespace bwd5 {

constexpr int NLOAD = 512;          // 16 converter warps
constexpr int NTHR = NLOAD + 64;    // + 1 MMA-issuing warp + 1 TMA producer warp
constexpr int R = 16;               // rows per slab
constexpr int KG = R / 8;           // one kind::tf32 MMA consumes 8 rows (K = 8)
constexpr int NST_MAX = 12;         // slab ring depth: as many stages as fit (TMA runs that many slabs ahead)
constexpr int CH_DEFAULT = 16;      // slabs per accumulation chunk (256 rows), see below
constexpr int MAX_NOUT = 256, MAX_KIN = 128;

__device__ __forceinline__ void split_rna(float v, uint32_t &hi, uint32_t &lo) {
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi) : "f"(v));
    lo = __float_as_uint(v - __uint_as_float(hi));
}

// 256-bit global accesses (sm_100: LDG.E.ENL2.256 / STG.E.ENL2.256): a thread that walks its OWN row touches a
// different 128-byte line than its neighbours, so the L1 spends one wavefront per thread and instruction whatever
// the width -- 32 bytes per wavefront instead of 16 is what takes dgrad_umma_kernel off the L1 limit (ncu round 2:
// l1tex 69-97 % of peak with 128-bit accesses).
struct __align__(32) F8 { float v[8]; };
__device__ __forceinline__ F8 ldg256(const float *p) {
    F8 r;
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7])
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void stg256(float *p, const F8 &r) {
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(r.v[0]), "f"(r.v[1]), "f"(r.v[2]),
                 "f"(r.v[3]), "f"(r.v[4]), "f"(r.v[5]), "f"(r.v[6]), "f"(r.v[7]) : "memory");
}

// MN-major fp32/tf32 operands have ONE legal shared-memory layout on sm_100: SWIZZLE_128B_BASE32B (layout type 1),
// canonical form Swizzle<2,5,2> o ((T,8,m),(4,k)):((1,T,LBO),(.,SBO)), T = 4 fp32 -- an atom is 4 k-rows of 128
// bytes (32 consecutive M/N elements), the 32-byte chunk c of row r sits at chunk (c ^ (r & 3)) -- what TMA's
// CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B writes.  Here: the two 4-row atoms of one K = 8 MMA are 512 B apart (SBO),
// consecutive 32-column blocks one 16-row TMA box = 2 KB apart (LBO); the 8-row k groups inside a box are walked
// by the issuing thread through the start address.
__device__ __forceinline__ uint64_t desc_mn_sw128(uint32_t saddr) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((KG * 1024) >> 4) << 16) | ((uint64_t)(512 >> 4) << 32) |
           (1ull << 46) | (1ull << 61);
}
// descriptor = {low word: start >> 4 | LBO >> 4 << 16, high word: SBO >> 4 | version 1 << 14 | layout type << 29}; the
// issuing thread only ever adds 16-byte offsets to the low word (shared memory is < 256 KB: the 14-bit field never
// overflows), which keeps the per-MMA instruction count of that single thread small
__device__ __forceinline__ uint64_t mk64(uint32_t lo, uint32_t hi) {
    uint64_t d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(hi));
    return d;
}
// kind::tf32, fp32 accumulate, A and B both MN-major (bits 15 / 16)
__host__ __device__ constexpr uint32_t idesc_tf32_mn(int M, int N) {
    return ptx::umma_idesc_tf32(M, N) | (1u << 15) | (1u << 16);
}

// The operand with at most 128 columns plays A (M = 128, zero padded), the other one B (N <= 256): for
// Nout > 128 the product is computed transposed (A = X, B = dY) so one M tile always suffices.
struct WgGeom {
    int swap, colsP, colsQ, NI, nblkQ, imgP, imgQ, stage, nst, smem;
};
inline __host__ __device__ WgGeom wg_geom(int Nout, int Kin) {
    WgGeom g;
    g.swap = Nout > 128;
    g.colsP = g.swap ? Kin : Nout;
    g.colsQ = g.swap ? Nout : Kin;
    g.NI = (g.colsQ + 15) / 16 * 16;
    g.nblkQ = (g.NI + 31) / 32;
    g.imgP = 4 * R * 128;
    g.imgQ = g.nblkQ * R * 128;
    g.stage = 2 * (g.imgP + g.imgQ);
    g.nst = (227 * 1024 - 2048) / g.stage;
    if (g.nst > NST_MAX) g.nst = NST_MAX;
    g.smem = g.nst * g.stage + 1024 + 1024;
    return g;
}

// byte offset of 16-byte chunk (row r of the slab, column c) inside an image: 32-column blocks of R rows x 128 B
__device__ __forceinline__ int img_off(int r, int c) {
    const int j = (c & 31) >> 2;
    return ((c >> 5) * (R * 128)) + (r << 7) + (((j >> 1) ^ (r & 3)) << 5) + ((j & 1) << 4);
}
__device__ __forceinline__ void named_sync_loaders() { asm volatile("bar.sync 1, %0;" ::"n"(NLOAD) : "memory"); }

// Roles: warp 9 (one lane) issues the TMA box loads of a slab into a ring stage, up to NST slabs ahead; warps 0-7
// write the lo images of a landed stage and sum the bias gradient (column sums of dY) from the values they
// touch; warp 8 (one lane) issues the MMAs of a converted stage and frees it through tcgen05.commit.  The tensor
// core accumulates with truncation, a bias that grows with the number of accumulations (measured 7e-9 per row),
// so a TMEM accumulator only collects a chunk of CH slabs (256 rows); at a chunk end the converter warps add it
// (fp32, round-to-nearest) to a master accumulator in the other half of TMEM (tcgen05.ld / add / tcgen05.st).
template <int W>   // converter warps per group (a multiple of 4); NLOAD / 32 / W groups convert that many slabs concurrently
__global__ void __launch_bounds__(NTHR, 1)   // 18 warps = 5 on one scheduler: 96 registers per thread
wgrad_umma_kernel(const __grid_constant__ CUtensorMap tm_p, const __grid_constant__ CUtensorMap tm_q, int64_t rows,
                  int Nout, int Kin, int64_t rows_per_cta, float *__restrict__ part, int CH) {
    constexpr int NW = NLOAD / 32, groups = NW / W;
    static_assert(W % 4 == 0 && R % W == 0 && NW % W == 0, "a warp's rows must share (row & 3), the swizzle phase");
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = ptx::smem_u32(smem_raw);
    uint8_t *sm = smem_raw + (((raw + 1023u) & ~1023u) - raw);
    const uint32_t s0 = ptx::smem_u32(sm);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const WgGeom G = wg_geom(Nout, Kin);
    const int NST = G.nst, STAGE = G.stage, P_IMG = G.imgP, Q_IMG = G.imgQ, NI = G.NI;
    // stage layout: P_raw (= hi) | P_lo | Q_raw (= hi) | Q_lo
    uint8_t *bars = sm + NST * STAGE;
    const uint32_t bar_tma = ptx::smem_u32(bars), bar_full = bar_tma + 8 * NST_MAX, bar_free = bar_full + 8 * NST_MAX,
                   bar_chunk = bar_free + 8 * NST_MAX, bar_flush = bar_chunk + 8;
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(bars + 8 * (3 * NST_MAX + 2));

    // zero the ring once: blocks past the last real column block are never written again
    for (int i = tid; i < NST * STAGE / 16; i += NTHR) reinterpret_cast<uint4 *>(sm)[i] = make_uint4(0, 0, 0, 0);
    if (tid == 0) {
        for (int i = 0; i < NST; ++i) {
            ptx::mbar_init(bar_tma + 8 * i, 1);
            ptx::mbar_init(bar_full + 8 * i, W);
            ptx::mbar_init(bar_free + 8 * i, 1);
        }
        ptx::mbar_init(bar_chunk, 1);
        ptx::mbar_init(bar_flush, NLOAD / 32);
        ptx::fence_barrier_init();
    }
    if (warp == 0) ptx::tmem_alloc<512>(ptx::smem_u32((const void *)tmem_slot));
    ptx::fence_proxy_async();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    const int64_t row_beg = (int64_t)blockIdx.x * rows_per_cta;
    const int64_t row_end = row_beg + rows_per_cta < rows ? row_beg + rows_per_cta : rows;
    const int n_slabs = row_beg < row_end ? (int)((row_end - row_beg + R - 1) / R) : 0;
    const int ld = Kin + 1;
    float *C = part + (size_t)blockIdx.x * Nout * ld;
    const int boxes_p = (G.colsP + 31) >> 5, boxes_q = (G.colsQ + 31) >> 5;
    // chunk accumulators: columns 0.. (hi*hi), 128.. (lo*hi), 384.. (hi*lo) when NI <= 128, else one at 0; master at 256
    const bool chains3 = NI <= 128;
    const uint32_t acc1 = chains3 ? 128u : 0u, acc2 = chains3 ? 384u : 0u;

    if (warp == NLOAD / 32 + 1) {
        // ---------------------------------------------------------------- TMA producer (one lane)
        if (lane == 0) {
            ptx::tma_prefetch_desc(&tm_p);
            ptx::tma_prefetch_desc(&tm_q);
        }
        // lane b issues box b of a slab (P boxes first), lane 0 arms the stage barrier before anyone issues
        int stage = 0;
        uint32_t ph = 1;   // parity of the previous use of this stage
        for (int slab = 0; slab < n_slabs; ++slab) {
            const uint32_t bar = bar_tma + 8 * stage, p_raw = s0 + stage * STAGE, q_raw = p_raw + 2 * P_IMG;
            const int row = (int)row_beg + slab * R;
            if (lane == 0) {
                if (slab >= NST) ptx::mbar_wait(bar_free + 8 * stage, ph);
                ptx::mbar_arrive_expect_tx(bar, (uint32_t)(boxes_p + boxes_q) * (R * 128));
            }
            __syncwarp();
            if (lane < boxes_p) ptx::tma_load_2d(&tm_p, bar, p_raw + lane * (R * 128), lane * 32, row);
            else if (lane < boxes_p + boxes_q)
                ptx::tma_load_2d(&tm_q, bar, q_raw + (lane - boxes_p) * (R * 128), (lane - boxes_p) * 32, row);
            if (++stage == NST) { stage = 0; ph ^= 1u; }
        }
    } else if (warp == NLOAD / 32) {
        // ---------------------------------------------------------------- MMA issuer (one lane)
        if (lane == 0) {
            const uint32_t idesc = idesc_tf32_mn(128, NI);
            const uint32_t hi_w = (512u >> 4) | (1u << 14) | (1u << 29);
            const uint32_t lo0 = (s0 >> 4) | ((uint32_t)((KG * 1024) >> 4) << 16);
            const uint32_t o_al = (uint32_t)P_IMG >> 4, o_bh = (uint32_t)(2 * P_IMG) >> 4, o_bl = (uint32_t)(2 * P_IMG + Q_IMG) >> 4;
            const uint32_t st_step = (uint32_t)STAGE >> 4;
            int stage = 0, in_chunk = 0;
            uint32_t ph = 0, fl_ph = 0, lo_w = lo0;
            for (int slab = 0; slab < n_slabs; ++slab) {
                if (in_chunk == 0 && slab > 0) {   // chunk accumulators were drained
                    ptx::mbar_wait(bar_flush, fl_ph);
                    fl_ph ^= 1u;
                }
                ptx::mbar_wait(bar_full + 8 * stage, ph);
                ptx::tc_fence_after();
#pragma unroll
                for (int g = 0; g < KG; ++g) {
                    const uint32_t w = lo_w + 64u * g;
                    const uint64_t ah = mk64(w, hi_w), al = mk64(w + o_al, hi_w), bh = mk64(w + o_bh, hi_w),
                                   bl = mk64(w + o_bl, hi_w);
                    // the three split terms go to separate accumulators when they fit (NI <= 128): the small terms
                    // then do not suffer the truncation of being added to the large running sum
                    const uint32_t first = (in_chunk | g) ? 1u : 0u;
                    ptx::umma_tf32_ss(tmem + acc1, al, bh, idesc, first);
                    ptx::umma_tf32_ss(tmem + acc2, ah, bl, idesc, chains3 ? first : 1u);
                    ptx::umma_tf32_ss(tmem, ah, bh, idesc, chains3 ? first : 1u);
                }
                ptx::umma_commit(bar_free + 8 * stage);
                const bool last = in_chunk == CH - 1 || slab == n_slabs - 1;
                if (last) ptx::umma_commit(bar_chunk);
                in_chunk = last ? 0 : in_chunk + 1;
                lo_w += st_step;
                if (++stage == NST) { stage = 0; ph ^= 1u; lo_w = lo0; }
            }
        }
    } else {
        // ---------------------------------------------------------------- converters / flushers
        const int lq = warp & 3, half = warp >> 2;
        constexpr int NH = NW / 4;   // warps per TMEM lane quarter
        const uint32_t t_lane = tmem + ((uint32_t)(lq * 32) << 16);
        {   // master accumulator (columns 256..511) starts at zero
            uint32_t z[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) z[i] = 0u;
            for (int cb = half; cb < 16; cb += NH) ptx::tmem_st_32x32b_x16(t_lane + 256 + cb * 16, z);
            ptx::tmem_st_wait();
        }
        // Converter warps work in `groups` groups of W warps.  A group owns whole slabs (slab % groups), a warp of the
        // group the rows wi, wi + W, .. of the slab, a lane the 16-byte column chunks lane (P and Q) and lane + 32 (Q).
        // W is a multiple of 4, so all rows of a warp share the swizzle phase (row & 3): a lane's byte offsets inside
        // a stage are fixed for the whole launch and the per-slab code is loads, 2 ALU ops per value and stores.  All
        // loads of a slab are issued before the first store (the compiler cannot reorder them itself: raw and lo
        // images live in the same shared array), which is what hides the shared-memory latency with 4 warps per
        // scheduler.
        const int p4 = G.colsP >> 2, q4 = G.colsQ >> 2;
        const int gi = warp / W, wi = warp % W;
        constexpr int RW = R / W;
        const bool hasP = lane < p4, hasQ0 = lane < q4, hasQ1 = lane + 32 < q4;
        const int offP = img_off(wi, lane << 2), offQ1 = img_off(wi, (lane + 32) << 2);   // offQ0 == offP
        float4 bsum[2];   // column sums of dY (the P operand, or Q when the roles are swapped)
        bsum[0] = bsum[1] = make_float4(0.f, 0.f, 0.f, 0.f);
        auto lo4 = [](const float4 &v) {
            auto lo_of = [](float x) { return x - __uint_as_float(__float_as_uint(x) & 0xffffe000u); };
            return make_float4(lo_of(v.x), lo_of(v.y), lo_of(v.z), lo_of(v.w));
        };
        auto add4 = [](float4 &a, const float4 &v) { a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w; };
        // a parity wait is only meaningful one phase ahead: a warp's next slab must not be more than a ring away
        const int ncw = NST < groups ? NST : groups;
        const int n_chunks = (n_slabs + CH - 1) / CH;
        int cv_slab = 0, cv_stage = 0;
        uint32_t cv_ph = 0;
        for (int chunk = 0; chunk < n_chunks; ++chunk) {
            const int s_end = (chunk + 1) * CH < n_slabs ? (chunk + 1) * CH : n_slabs;
            for (int slab = chunk * CH + gi; slab < s_end && gi < ncw; slab += ncw) {
                while (cv_slab < slab) {   // advance (stage, phase) to this slab without dividing
                    ++cv_slab;
                    if (++cv_stage == NST) { cv_stage = 0; cv_ph ^= 1u; }
                }
                const int stage = cv_stage;
                ptx::mbar_wait(bar_tma + 8 * stage, cv_ph);
                uint8_t *sp = sm + stage * STAGE, *sq = sp + 2 * P_IMG;
                float4 vp[RW], vq0[RW], vq1[RW];
#pragma unroll
                for (int k = 0; k < RW; ++k) {
                    if (hasP) vp[k] = *reinterpret_cast<const float4 *>(sp + offP + k * (W * 128));
                    if (hasQ0) vq0[k] = *reinterpret_cast<const float4 *>(sq + offP + k * (W * 128));
                    if (hasQ1) vq1[k] = *reinterpret_cast<const float4 *>(sq + offQ1 + k * (W * 128));
                }
#pragma unroll
                for (int k = 0; k < RW; ++k) {
                    if (hasP) {
                        *reinterpret_cast<float4 *>(sp + P_IMG + offP + k * (W * 128)) = lo4(vp[k]);
                        if (!G.swap) add4(bsum[0], vp[k]);
                    }
                    if (hasQ0) {
                        *reinterpret_cast<float4 *>(sq + Q_IMG + offP + k * (W * 128)) = lo4(vq0[k]);
                        if (G.swap) add4(bsum[0], vq0[k]);
                    }
                    if (hasQ1) {
                        *reinterpret_cast<float4 *>(sq + Q_IMG + offQ1 + k * (W * 128)) = lo4(vq1[k]);
                        if (G.swap) add4(bsum[1], vq1[k]);
                    }
                }
                ptx::fence_proxy_async();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(bar_full + 8 * stage);
            }
            // chunk accumulator (columns 0..) += into the master (columns 256..), fp32 round-to-nearest
            ptx::mbar_wait(bar_chunk, chunk & 1);
            ptx::tc_fence_after();
            for (int cb = half; cb < NI / 16; cb += NH) {
                uint32_t a[16], m[16];
                ptx::tmem_ld_32x32b_x16(t_lane + cb * 16, a);
                ptx::tmem_ld_32x32b_x16(t_lane + 256 + cb * 16, m);
                ptx::tmem_ld_wait();
                if (chains3) {
                    uint32_t a1[16], a2[16];
                    ptx::tmem_ld_32x32b_x16(t_lane + acc1 + cb * 16, a1);
                    ptx::tmem_ld_32x32b_x16(t_lane + acc2 + cb * 16, a2);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        a[i] = __float_as_uint((__uint_as_float(a1[i]) + __uint_as_float(a2[i])) + __uint_as_float(a[i]));
                }
#pragma unroll
                for (int i = 0; i < 16; ++i) m[i] = __float_as_uint(__uint_as_float(m[i]) + __uint_as_float(a[i]));
                ptx::tmem_st_32x32b_x16(t_lane + 256 + cb * 16, m);
            }
            ptx::tmem_st_wait();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(bar_flush);
        }
        // every MMA has completed (the last flush waited for it) and every TMA load has landed: the ring is free
        // -> bias-gradient scratch [NW warps][Nout], summed in warp order
        named_sync_loaders();
        float *sdb = reinterpret_cast<float *>(sm);
        {
            const int d4 = Nout >> 2;
            if (lane < d4) *reinterpret_cast<float4 *>(sdb + warp * Nout + (lane << 2)) = bsum[0];
            if (lane + 32 < d4) *reinterpret_cast<float4 *>(sdb + warp * Nout + ((lane + 32) << 2)) = bsum[1];
        }
        named_sync_loaders();
        for (int c = tid; c < Nout; c += NLOAD) {
            float t = 0.f;
#pragma unroll
            for (int w = 0; w < NLOAD / 32; ++w) t += sdb[w * Nout + c];
            C[(size_t)c * ld + Kin] = t;
        }
        // master accumulator -> this CTA's partial (D[m][n]: m = P column, n = Q column)
        ptx::tc_fence_after();
        const int m = lq * 32 + lane;
        for (int cb = half; cb < NI / 16; cb += NH) {
            uint32_t v[16];
            ptx::tmem_ld_32x32b_x16(t_lane + 256 + cb * 16, v);
            ptx::tmem_ld_wait();
            if (m < G.colsP) {
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int n = cb * 16 + i;
                    if (n < G.colsQ) {
                        if (G.swap) C[(size_t)n * ld + m] = __uint_as_float(v[i]);
                        else C[(size_t)m * ld + n] = __uint_as_float(v[i]);
                    }
                }
            }
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc<512>(tmem);
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline EncodeTiledFn encode_tiled() {
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
// fp32 [rows, cols] row-major, box = 32 columns x box_rows rows, 32-byte-atom 128-byte swizzle, zero fill past the edges
inline int make_slab_tmap(CUtensorMap *m, const float *ptr, uint64_t rows, uint64_t cols, uint32_t box_rows) {
    EncodeTiledFn fn = encode_tiled();
    if (!fn) return LIPVQ_ECUDA;
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * sizeof(float)};
    cuuint32_t box[2] = {32, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(ptr), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? LIPVQ_OK : LIPVQ_ECUDA;
}

inline bool wgrad_umma_ok(int64_t rows, int Nout, int Kin) {
    return Nout >= 4 && Nout <= MAX_NOUT && Kin >= 4 && Kin <= MAX_KIN && (Nout & 3) == 0 && (Kin & 3) == 0 &&
           rows >= 2048 && rows < (1ll << 31) && wg_geom(Nout, Kin).nst >= 2;
}
inline int wgrad_umma_grid(int64_t rows, int64_t *rows_per_cta) {
    int64_t slabs = (rows + R - 1) / R;
    int64_t per = (slabs + kNumSMs - 1) / kNumSMs * R;
    *rows_per_cta = per;
    return (int)((rows + per - 1) / per);
}
// part: [grid][Nout][Kin+1]; returns the number of partials in *splits
inline int wgrad_umma_launch(const float *dY, const float *X, int64_t rows, int Nout, int Kin, float *part,
                             int *splits, cudaStream_t st) {
    int64_t per;
    const int grid = wgrad_umma_grid(rows, &per);
    const WgGeom G = wg_geom(Nout, Kin);
    const int smem = G.smem;
    CUtensorMap tm_dy, tm_x;
    LIPVQ_TRY(make_slab_tmap(&tm_dy, dY, (uint64_t)rows, (uint64_t)Nout, R));
    LIPVQ_TRY(make_slab_tmap(&tm_x, X, (uint64_t)rows, (uint64_t)Kin, R));
    static const int wpg = [] { const char *e = getenv("LIPVQ_WGRAD_WPG"); return e ? atoi(e) : 16; }();   // A/B probe
    ProfScope prof("wgrad_umma_kernel", st);
    // all 16 converter warps on one slab (one row each), 16 slabs (256 rows) per accumulation chunk (1e-6 accurate)
    auto go = [&](auto kern) -> int {
        LIPVQ_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        kern<<<grid, NTHR, smem, st>>>(G.swap ? tm_x : tm_dy, G.swap ? tm_dy : tm_x, rows, Nout, Kin, per, part, CH_DEFAULT);
        return LIPVQ_OK;
    };
    if (wpg == 8) LIPVQ_TRY(go(wgrad_umma_kernel<8>));
    else if (wpg == 4) LIPVQ_TRY(go(wgrad_umma_kernel<4>));
    else LIPVQ_TRY(go(wgrad_umma_kernel<16>));
    LIPVQ_LAUNCH_OK();
    *splits = grid;
    return LIPVQ_OK;
}


// =====================================================================================================================
//   dgrad_umma_kernel : dX[M, Kout] = scale * (dY[M, Nn] W[Nn, Kout]) (.) act'(dpre[M, Kout])      (autograd of a Linear
//   followed by GELU / ReLU, v5:54-59 and backbone.py:17-32)
//
// W in its natural row-major layout IS the MN-major B operand (rows = reduction axis): its tf32 hi / lo images stay
// resident in shared memory for the whole launch.  The A operand lives in TMEM (tcgen05.mma with A in tensor memory,
// lane = row): 8 loader warps read 64 B of their own row per 32-column block straight from global memory into
// registers (two blocks ahead), split into hi / lo and tcgen05.st them into a 4-slot ring of TMEM columns -- dY never
// touches shared memory.  One lane of warp 16 issues the 3 MMAs per K = 8 step into one of two accumulators; 8
// epilogue warps read the finished accumulator, apply the activation derivative and store dX while the next tile's
// MMAs run into the other accumulator.   TMEM columns: acc0 0 | acc1 128 | A ring 256 + 64 * slot (hi 32 | lo 32).
namespace dg {
constexpr int TM = 128;
constexpr int NLD = 256, NEP = 256, NT = NLD + NEP + 32;
constexpr int SLOTS = 4;
constexpr int LOOK = 3;             // A blocks a loader thread keeps in flight (registers)
__host__ __device__ constexpr int nn_pad(int Nn) { return (Nn + 7) / 8 * 8; }
__host__ __device__ constexpr int w_img_bytes(int Nn, int Kout) { return (Kout / 32) * nn_pad(Nn) * 128; }
__host__ __device__ constexpr int smem_bytes(int Nn, int Kout) { return 2 * w_img_bytes(Nn, Kout) + 256 + 1024; }
}  // namespace dg

template <int DACT, bool WIDE>
__global__ void __launch_bounds__(dg::NT, 1)
dgrad_umma_kernel(const float *__restrict__ dY, const float *__restrict__ W, const float *__restrict__ dpre,
                  float *__restrict__ dX, int64_t M, int Nn, int Kout, const float *__restrict__ scale_dev, int PF) {
    using namespace dg;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = ptx::smem_u32(smem_raw);
    uint8_t *sm = smem_raw + (((raw + 1023u) & ~1023u) - raw);
    const uint32_t s0 = ptx::smem_u32(sm);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int NnP = nn_pad(Nn), IMG = w_img_bytes(Nn, Kout), KB = (Nn + 31) >> 5;
    uint8_t *bars = sm + 2 * IMG;
    const uint32_t bar_afull = ptx::smem_u32(bars), bar_afree = bar_afull + 8 * SLOTS, bar_cfull = bar_afree + 8 * SLOTS,
                   bar_cfree = bar_cfull + 16;
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(bars + 8 * (2 * SLOTS + 4));

    // W -> hi | lo operand images: 32-column blocks of NnP rows x 128 B, 32-byte chunks swizzled by (row & 3)
    {
        const int k4 = Kout >> 2;
        for (int i = tid; i < NnP * k4; i += NT) {
            const int k = i / k4, c = (i - k * k4) << 2, j = (c & 31) >> 2;
            const float4 v = k < Nn ? __ldg(reinterpret_cast<const float4 *>(W + (size_t)k * Kout + c))
                                    : make_float4(0.f, 0.f, 0.f, 0.f);
            uint32_t h[4], l[4];
            split_rna(v.x, h[0], l[0]); split_rna(v.y, h[1], l[1]);
            split_rna(v.z, h[2], l[2]); split_rna(v.w, h[3], l[3]);
            const int off = (c >> 5) * (NnP * 128) + (k << 7) + (((j >> 1) ^ (k & 3)) << 5) + ((j & 1) << 4);
            *reinterpret_cast<uint4 *>(sm + off) = make_uint4(h[0], h[1], h[2], h[3]);
            *reinterpret_cast<uint4 *>(sm + IMG + off) = make_uint4(l[0], l[1], l[2], l[3]);
        }
    }
    if (tid == 0) {
        for (int i = 0; i < SLOTS; ++i) {
            ptx::mbar_init(bar_afull + 8 * i, NLD / 32);
            ptx::mbar_init(bar_afree + 8 * i, 1);
        }
        for (int i = 0; i < 2; ++i) {
            ptx::mbar_init(bar_cfull + 8 * i, 1);
            ptx::mbar_init(bar_cfree + 8 * i, NEP / 32);
        }
        ptx::fence_barrier_init();
    }
    if (warp == 0) ptx::tmem_alloc<512>(ptx::smem_u32((const void *)tmem_slot));
    ptx::fence_proxy_async();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    const int64_t n_tiles = (M + TM - 1) / TM;
    const int my_tiles = blockIdx.x < n_tiles ? (int)((n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x) : 0;

    if (warp == (NLD + NEP) / 32) {
        // ---------------------------------------------------------------- MMA issuer (one lane)
        if (lane == 0) {
            const uint32_t idesc = ptx::umma_idesc_tf32(TM, Kout) | (1u << 16);   // B is MN-major
            // B descriptor: 32-column blocks NnP*128 B apart (LBO), the two 4-row atoms of a K = 8 step 512 B apart (SBO)
            const uint32_t hi_w = (512u >> 4) | (1u << 14) | (1u << 29);
            const uint32_t lo_h = (s0 >> 4) | ((uint32_t)((NnP * 128) >> 4) << 16), lo_l = lo_h + ((uint32_t)IMG >> 4);
            const uint32_t kb_step = 256u, s_step = 64u;   // 32 rows / 8 rows of 128 B, in 16-byte units
            int slot = 0;
            uint32_t a_ph = 0;
            for (int t = 0; t < my_tiles; ++t) {
                const int buf = t & 1;
                if (t >= 2) {
                    ptx::mbar_wait(bar_cfree + 8 * buf, ((t >> 1) - 1) & 1);
                    ptx::tc_fence_after();
                }
                const uint32_t acc = tmem + buf * 128;
                int k_left = (Nn + 7) >> 3;      // K = 8 steps left in this tile
                uint32_t kw = 0;                 // (k0 * 128) >> 4
                for (int kb = 0; kb < KB; ++kb) {
                    ptx::mbar_wait(bar_afull + 8 * slot, a_ph);
                    ptx::tc_fence_after();
                    const int ks = k_left < 4 ? k_left : 4;
                    const uint32_t a0 = tmem + 256 + slot * 64;
#pragma unroll
                    for (int s = 0; s < 4; ++s)
                        if (s < ks) {
                            const uint64_t bh = mk64(lo_h + kw + s_step * s, hi_w), bl = mk64(lo_l + kw + s_step * s, hi_w);
                            ptx::umma_tf32_ts(acc, a0 + 32 + 8 * s, bh, idesc, (kb | s) ? 1u : 0u);
                            ptx::umma_tf32_ts(acc, a0 + 8 * s, bl, idesc, 1u);
                            ptx::umma_tf32_ts(acc, a0 + 8 * s, bh, idesc, 1u);
                        }
                    ptx::umma_commit(bar_afree + 8 * slot);
                    k_left -= 4;
                    kw += kb_step;
                    if (++slot == SLOTS) { slot = 0; a_ph ^= 1u; }
                }
                ptx::umma_commit(bar_cfull + 8 * buf);
            }
        }
    } else if (warp < NLD / 32) {
        // ---------------------------------------------------------------- A loaders: global -> registers -> TMEM
        const int r = tid & 127, h = tid >> 7;
        const uint32_t t_lane = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        const int total = my_tiles * KB;
        // WIDE (Nn % 8 == 0, 32-byte aligned rows): two 256-bit loads per row and block instead of four 128-bit ones
        auto load = [&](int b, float(&v)[16]) {
            const int t = b / KB, kb = b - t * KB;
            const int64_t row = ((int64_t)blockIdx.x + (int64_t)t * gridDim.x) * TM + r;
            const int c0 = kb * 32 + h * 16;
            const bool live = b < total && row < M;
            const float *src = dY + row * Nn + c0;
            if (WIDE) {
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    if (live && c0 + 8 * i < Nn) {
                        const F8 q = ldg256(src + 8 * i);
#pragma unroll
                        for (int j = 0; j < 8; ++j) v[8 * i + j] = q.v[j];
                    } else {
#pragma unroll
                        for (int j = 0; j < 8; ++j) v[8 * i + j] = 0.f;
                    }
                }
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float4 q = (live && c0 + 4 * i < Nn) ? __ldg(reinterpret_cast<const float4 *>(src) + i)
                                                               : make_float4(0.f, 0.f, 0.f, 0.f);
                    v[4 * i] = q.x; v[4 * i + 1] = q.y; v[4 * i + 2] = q.z; v[4 * i + 3] = q.w;
                }
            }
        };
        auto step = [&](int b, float(&v)[16]) {
            const int slot = b % SLOTS;
            if (b >= SLOTS) {
                ptx::mbar_wait(bar_afree + 8 * slot, ((b / SLOTS) - 1) & 1);
                ptx::tc_fence_after();
            }
            uint32_t hi[16], lo[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) split_rna(v[i], hi[i], lo[i]);
            ptx::tmem_st_32x32b_x16(t_lane + 256 + slot * 64 + h * 16, hi);
            ptx::tmem_st_32x32b_x16(t_lane + 256 + slot * 64 + 32 + h * 16, lo);
            load(b + LOOK, v);
            if (h == 0) {   // the 128-byte line of this row PF blocks further ahead -> L2
                const int bp = b + PF;
                const int tp = bp / KB, kp = bp - tp * KB;
                const int64_t prow = ((int64_t)blockIdx.x + (int64_t)tp * gridDim.x) * TM + r;
                if (bp < total && prow < M && kp * 32 < Nn) prefetch_l2(dY + prow * Nn + kp * 32);
            }
            ptx::tmem_st_wait();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(bar_afull + 8 * slot);
        };
        float v0[16], v1[16], v2[16];
        load(0, v0);
        load(1, v1);
        load(2, v2);
        for (int b = 0; b < total; b += 3) {
            step(b, v0);
            if (b + 1 < total) step(b + 1, v1);
            if (b + 2 < total) step(b + 2, v2);
        }
    } else {
        // ---------------------------------------------------------------- epilogue: acc -> act' -> dX
        const int e = tid - NLD, r = e & 127, h = e >> 7;
        const uint32_t t_lane = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        const float out_scale = scale_dev ? __ldg(scale_dev) : 1.f;
        for (int t = 0; t < my_tiles; ++t) {
            const int buf = t & 1;
            const int64_t row = ((int64_t)blockIdx.x + (int64_t)t * gridDim.x) * TM + r;
            // the saved pre-activations of this thread's row are requested BEFORE the accumulator is waited for
            // (and the next column block's before the current one is used): their latency hides behind the MMAs
            const bool live = row < M;
            const float *pre = dpre + row * Kout;
            float *dst = dX + row * Kout;
            float pv[16] = {};
            auto load_pre = [&](int cb) {
                if (DACT == ACT_NONE || !live || cb >= Kout / 16) return;
                if (WIDE) {
#pragma unroll
                    for (int i = 0; i < 2; ++i) {
                        const F8 q = ldg256(pre + cb * 16 + 8 * i);
#pragma unroll
                        for (int j = 0; j < 8; ++j) pv[8 * i + j] = q.v[j];
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float4 q = __ldg(reinterpret_cast<const float4 *>(pre + cb * 16) + i);
                        pv[4 * i] = q.x; pv[4 * i + 1] = q.y; pv[4 * i + 2] = q.z; pv[4 * i + 3] = q.w;
                    }
                }
            };
            load_pre(h);
            if (DACT != ACT_NONE && t + 1 < my_tiles) {   // next tile's half row -> L2 (no registers held)
                const int64_t nrow = row + (int64_t)gridDim.x * TM;
                if (nrow < M)
                    for (int c = h * (Kout >> 1); c < (h + 1) * (Kout >> 1); c += 32) prefetch_l2(dpre + nrow * Kout + c);
            }
            ptx::mbar_wait(bar_cfull + 8 * buf, (t >> 1) & 1);
            ptx::tc_fence_after();
            for (int cb = h; cb < Kout / 16; cb += 2) {
                uint32_t a[16];
                ptx::tmem_ld_32x32b_x16(t_lane + buf * 128 + cb * 16, a);
                ptx::tmem_ld_wait();
                float o[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    o[i] = out_scale * __uint_as_float(a[i]);
                    if (DACT == ACT_GELU) o[i] *= gelu_grad_fast_f(pv[i]);
                    if (DACT == ACT_RELU) o[i] = pv[i] > 0.f ? o[i] : 0.f;
                }
                load_pre(cb + 2);
                if (live) {
                    if (WIDE) {
#pragma unroll
                        for (int i = 0; i < 2; ++i) {
                            F8 q;
#pragma unroll
                            for (int j = 0; j < 8; ++j) q.v[j] = o[8 * i + j];
                            stg256(dst + cb * 16 + 8 * i, q);
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            reinterpret_cast<float4 *>(dst + cb * 16)[i] = make_float4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]);
                    }
                }
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(bar_cfree + 8 * buf);
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc<512>(tmem);
}

inline bool dgrad_umma_ok(int64_t M, int Nn, int Kout) {
    return (Kout == 64 || Kout == 128) && Nn >= 8 && (Nn & 3) == 0 && M >= 2048 &&
           dg::smem_bytes(Nn, Kout) <= 227 * 1024;
}
template <int DACT>
inline int dgrad_umma_launch(const float *dY, const float *W, const float *dpre, float *dX, int64_t M, int Nn, int Kout,
                             const float *scale_dev, cudaStream_t st) {
    const int smem = dg::smem_bytes(Nn, Kout);
    const int64_t tiles = (M + dg::TM - 1) / dg::TM;
    const int grid = (int)(tiles < kNumSMs ? tiles : kNumSMs);
    static const int pf = [] { const char *e = getenv("LIPVQ_DGRAD_PF"); return e ? atoi(e) : 10; }();   // L2 prefetch distance (blocks)
    static const bool narrow = [] { const char *e = getenv("LIPVQ_DGRAD"); return e && !strcmp(e, "narrow"); }();   // A/B
    auto al32 = [](const void *p) { return (reinterpret_cast<uintptr_t>(p) & 31) == 0; };
    const bool wide = (Nn & 7) == 0 && al32(dY) && al32(dX) && (DACT == ACT_NONE || al32(dpre)) &&
                      !narrow;
    ProfScope prof("dgrad_umma_kernel", st);
    if (wide) {
        LIPVQ_CUDA_OK(cudaFuncSetAttribute(dgrad_umma_kernel<DACT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        dgrad_umma_kernel<DACT, true><<<grid, dg::NT, smem, st>>>(dY, W, dpre, dX, M, Nn, Kout, scale_dev, pf);
    } else {
        LIPVQ_CUDA_OK(cudaFuncSetAttribute(dgrad_umma_kernel<DACT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        dgrad_umma_kernel<DACT, false><<<grid, dg::NT, smem, st>>>(dY, W, dpre, dX, M, Nn, Kout, scale_dev, pf);
    }
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

}  // namespace bwd5
}  // namespace lipvq
