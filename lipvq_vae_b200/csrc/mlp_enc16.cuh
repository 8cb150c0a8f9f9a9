// mlp_enc16.cuh -- the fused encoder (mlp_enc.cuh) with every GEMM as a 3-term fp16 split product on the
// kind::f16 tensor path (twice the tf32 rate, half the operand bytes):
//   v = hi + lo,  hi = fp16(v),  lo = fp16(v - hi)      (22 significand bits)
//   A*B ~= A_hi*B_hi + A_lo*B_hi + A_hi*B_lo            (fp32 accumulate in TMEM; dropped term ~2^-24)
// Operands are pre-scaled by 2^8 (exact) so the lo parts of O(1e-3..1e2) values stay normal in fp16; the
// epilogue folds the 2^-16 back into the bias FMA.  Values that would leave the fp16 range (|v| >= 255 for an
// activation or a weight) raise a device flag and the launch is redone by the tf32 kernel (gated launch).
//
// What the narrower operands buy besides the MMA rate:
//  * all three weights' hi/lo images fit in shared memory (152 KB at D=208): no per-tile Wn ring, no L2 traffic;
//  * a packed fp16 hi+lo pair of 16 activations takes exactly the 16 TMEM columns of the fp32 accumulator block
//    it was computed from, so every epilogue writes the next layer's A operand IN PLACE and the tile needs
//    128 (acc2/h2) + 64 (acc1/h1) + D16 (acc3) + 16 (x) columns with no aliasing between tiles' layers.
// TMEM columns: acc2/h2 [0,128) | acc1/h1 [128,192) | acc3 [192,192+D16) | x [400,416).
// Block cb of an A operand (K elements [16cb, 16cb+16)) = columns [16cb, 16cb+8) hi, [16cb+8, 16cb+16) lo.
#pragma once
#include <cuda_fp16.h>
#include "mlp_enc.cuh"

namespace lipvq {
namespace mlp {

constexpr float E16_SCALE = 256.0f;              // operand pre-scale (A and W)
constexpr float E16_DESCALE = 1.0f / 65536.0f;   // accumulator = 2^16 x true value
constexpr float E16_MAX = 255.0f;                // |v| * 256 must stay below 65504

constexpr uint32_t E16_R2 = 0, E16_R1 = 128, E16_R3 = 192, E16_X = 400;

struct Enc16Smem {
    static constexpr int W0 = 0, W0_IMG = H1 * 128;                 // 64 rows x one 128-byte K block (16 of 64 used)
    static constexpr int W1 = W0 + 2 * W0_IMG, W1_IMG = HH * 128;   // 128 rows x one K block (64 fp16)
    static constexpr int WN = W1 + 2 * W1_IMG, WN_IMG = ENC_DMAX * 128 * 2;   // D16 rows x two K blocks
    static constexpr int WBUF_W = 3072;                             // per warp: z_e slot (2 KB) + Zb slot (1 KB)
    static constexpr int WBUF = WN + 2 * WN_IMG;                    // 16 warps x 3 KB store staging
    static constexpr int BIAS = WBUF + 16 * WBUF_W;
    static constexpr int MU = BIAS + (H1 + HH + 256) * 4;
    static constexpr int ZNP = MU + 272 * 4;
    static constexpr int BAR = ZNP + 2 * 4 * TM * 4;                // (partial row norms: two tiles in flight)
    static constexpr int NBAR = 2 + 2 + 1 + ENC_NC + 1;             // a_ready[2] acc_ready[2] acc3 a2_ready[4] r3_free
    static constexpr int TOTAL = BAR + NBAR * 8 + 16 + 1024;
};

// two sigmoids with ONE reciprocal: 1/a = b * rcp(a b), 1/b = a * rcp(a b), a = 1 + 2^t.  t is clamped at 60 so
// the product cannot overflow (sigmoid(x) for x < -41.6 returns 2^-60 instead of a smaller positive number).
__device__ __forceinline__ void sigmoid_fast_pair(float &x0, float &x1) {
    const float a0 = 1.0f + ex2_approx(fminf(-1.4426950408889634f * x0, 60.0f));
    const float a1 = 1.0f + ex2_approx(fminf(-1.4426950408889634f * x1, 60.0f));
    const float r = rcp_approx(a0 * a1);
    x0 = r * a1;
    x1 = r * a0;
}

// two scaled values -> packed hi / lo halves (low 16 bits = first value = even k)
__device__ __forceinline__ void split_f16x2(float a, float b, uint32_t &hi, uint32_t &lo) {
    const __half2 h = __floats2half2_rn(a, b);
    const float2 f = __half22float2(h);
    const __half2 l = __floats2half2_rn(a - f.x, b - f.y);
    hi = *reinterpret_cast<const uint32_t *>(&h);
    lo = *reinterpret_cast<const uint32_t *>(&l);
}

// 16 activations of this thread's row -> one in-place A-operand block (8 columns hi | 8 columns lo)
__device__ __forceinline__ void pack_block16(const float (&v)[16], uint32_t (&o)[16], float &amax) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        amax = fmaxf(amax, fmaxf(fabsf(v[2 * i]), fabsf(v[2 * i + 1])));
        split_f16x2(v[2 * i] * E16_SCALE, v[2 * i + 1] * E16_SCALE, o[i], o[8 + i]);
    }
}

// W[n][k] (dense row-major fp32) -> scaled fp16 hi/lo images: K blocks of 64 (128 B per row), block kb at
// kb*n_pad*128, row n at n*128, 16-byte piece j at ((j ^ (n & 7)) * 16)  == SWIZZLE_128B, K-major, SBO = 1024.
__device__ __forceinline__ void stage_weight16(uint8_t *hi, uint8_t *lo, const float *__restrict__ W, int n_real,
                                               int k_real, int n_pad, int k_blocks, float &wmax) {
    const int total = n_pad * k_blocks * 8;
    for (int c = threadIdx.x; c < total; c += blockDim.x) {
        const int j = c & 7, n = (c >> 3) % n_pad, kb = (c >> 3) / n_pad;
        uint32_t h[4], l[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float v[2];
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                const int k = kb * 64 + j * 8 + 2 * i + t;
                v[t] = (n < n_real && k < k_real) ? __ldg(W + (size_t)n * k_real + k) : 0.f;
                wmax = fmaxf(wmax, fabsf(v[t]));
            }
            split_f16x2(v[0] * E16_SCALE, v[1] * E16_SCALE, h[i], l[i]);
        }
        const int off = kb * n_pad * 128 + n * 128 + ((j ^ (n & 7)) << 4);
        *reinterpret_cast<uint4 *>(hi + off) = make_uint4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<uint4 *>(lo + off) = make_uint4(l[0], l[1], l[2], l[3]);
    }
}

// D[tmem] (+)= A[tmem: lane = row, two fp16 k per 32-bit column] * B[smem]^T  (cute::SM100_MMA_F16BF16_TS)
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

// one thread: K steps [s0, s1) of acc[128, n_pad] (+)= A * W^T; A block s at t_a + 16 s (hi) / + 8 (lo)
__device__ __forceinline__ void issue_steps16(uint32_t t_acc, uint32_t t_a, uint32_t s_w_hi, uint32_t s_w_lo, int n_pad,
                                              int s0, int s1, bool fresh) {
    const uint32_t idesc = ptx::umma_idesc_f16(TM, n_pad);
    for (int s = s0; s < s1; ++s) {
        const uint32_t blk = (uint32_t)(s >> 2) * (uint32_t)n_pad * 128u;
        const uint64_t bh = ptx::umma_desc_sw128(s_w_hi + blk) + 2u * (uint32_t)(s & 3);
        const uint64_t bl = ptx::umma_desc_sw128(s_w_lo + blk) + 2u * (uint32_t)(s & 3);
        const uint32_t a_hi = t_a + 16u * s, a_lo = a_hi + 8u;
        umma_f16_ts(t_acc, a_hi, bh, idesc, (fresh && s == s0) ? 0u : 1u);
        umma_f16_ts(t_acc, a_lo, bh, idesc, 1u);
        umma_f16_ts(t_acc, a_hi, bl, idesc, 1u);
    }
}

// Epilogue over the 16-column blocks cb in [cb0, cb1) with cb % nsub == sub, for this thread's row:
//   v = acc * 2^-16 + bias -> [pre-activation rows to global] -> act -> [activation rows to global]
//   -> next layer's A block written over the accumulator columns just read.
template <int ACT, class After>
__device__ __forceinline__ void epilogue_blocks16(uint32_t t_acc, int cb0, int cb1, int sub, int nsub, int n_real,
                                                  const float *sbias, uint8_t *wbuf, float *__restrict__ pre,
                                                  float *__restrict__ post, int ld, int rows_valid, float &amax,
                                                  After after) {
    int cb = cb0 + ((sub - cb0) % nsub + nsub) % nsub;
    for (; cb < cb1; cb += nsub) {
        const int c0 = cb << 4;
        uint32_t r[16];
        ptx::tmem_ld_32x32b_x16(t_acc + c0, r);
        ptx::tmem_ld_wait();
        float v[16];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float4 bq = *reinterpret_cast<const float4 *>(sbias + c0 + 4 * i);
            v[4 * i] = fmaf(__uint_as_float(r[4 * i]), E16_DESCALE, bq.x);
            v[4 * i + 1] = fmaf(__uint_as_float(r[4 * i + 1]), E16_DESCALE, bq.y);
            v[4 * i + 2] = fmaf(__uint_as_float(r[4 * i + 2]), E16_DESCALE, bq.z);
            v[4 * i + 3] = fmaf(__uint_as_float(r[4 * i + 3]), E16_DESCALE, bq.w);
        }
        const int valid = min(16, n_real - c0);
        if (pre) warp_store16(wbuf, pre + c0, ld, v, rows_valid, valid, true);
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = apply_act<ACT>(v[i]);
        if (post) warp_store16(wbuf, post + c0, ld, v, rows_valid, valid, true);
        pack_block16(v, r, amax);
        ptx::tmem_st_32x32b_x16(t_acc + c0, r);
        after(cb);
    }
}

template <bool TMA, bool KEEP>   // KEEP: a1/h1/a2/h2 are written for backward (training forward)
__global__ void __launch_bounds__(ENC_THR, 1)
enc16_fused_kernel(const float *__restrict__ x, const float *__restrict__ W0, const float *__restrict__ b0,
                   const float *__restrict__ W1, const float *__restrict__ b1, const float *__restrict__ Wn,
                   const float *__restrict__ b2, int64_t n_rows, int F, int D, float *__restrict__ a1,
                   float *__restrict__ h1, float *__restrict__ a2, float *__restrict__ h2, float *__restrict__ z_e,
                   const float *__restrict__ mu, tc::op_t *__restrict__ Zb, float *__restrict__ zn, int DKp,
                   int *__restrict__ counters, int *__restrict__ range_flag,
                   const __grid_constant__ CUtensorMap tm_ze, const __grid_constant__ CUtensorMap tm_zb) {
    using L = Enc16Smem;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = ptx::smem_u32(smem_raw);
    uint8_t *sm = smem_raw + (((raw + 1023u) & ~1023u) - raw);
    const uint32_t s0 = ptx::smem_u32(sm);
    float *sb0 = reinterpret_cast<float *>(sm + L::BIAS), *sb1 = sb0 + H1, *sb2 = sb1 + HH;
    const uint32_t bars = s0 + L::BAR;
    auto bar_a = [&](int l) { return bars + 8u * l; };               // A operand of layer 0/1 ready (16 warp arrivals)
    auto bar_acc = [&](int l) { return bars + 8u * (2 + l); };       // accumulator of layer 0/1 ready
    const uint32_t bar_acc3 = bars + 8u * 4;                         // layer-2 accumulator ready
    // h2 blocks 2c, 2c+1 (K steps 2c, 2c+1 of layer 2) ready: 2 column owners x 4 lane quarters = 8 warp arrivals
    auto bar_a2 = [&](int c) { return bars + 8u * (5 + c); };
    const uint32_t bar_r3 = bars + 8u * (5 + ENC_NC);                // every warp has read acc3 of the previous tile
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(sm + L::BAR + L::NBAR * 8);

    const int D16 = pad16(D);
    float amax = 0.f;                                                // largest |operand value| this thread produced
    stage_weight16(sm + L::W0, sm + L::W0 + L::W0_IMG, W0, H1, F, H1, 1, amax);
    stage_weight16(sm + L::W1, sm + L::W1 + L::W1_IMG, W1, HH, H1, HH, 1, amax);
    stage_weight16(sm + L::WN, sm + L::WN + D16 * 256, Wn, D, HH, D16, 2, amax);
    for (int i = threadIdx.x; i < H1; i += ENC_THR) sb0[i] = b0[i];
    for (int i = threadIdx.x; i < HH; i += ENC_THR) sb1[i] = b1[i];
    for (int i = threadIdx.x; i < 256; i += ENC_THR) sb2[i] = i < D ? b2[i] : 0.f;
    float *smu = reinterpret_cast<float *>(sm + L::MU), *sznp = reinterpret_cast<float *>(sm + L::ZNP);
    for (int i = threadIdx.x; i < 272; i += ENC_THR) smu[i] = (Zb && i < D) ? mu[i] : 0.f;
    ptx::fence_proxy_async();
    if (threadIdx.x == 0) {
        for (int l = 0; l < 2; ++l) ptx::mbar_init(bar_a(l), ENC_EPI / 32);
        for (int l = 0; l < 2; ++l) ptx::mbar_init(bar_acc(l), 1);
        ptx::mbar_init(bar_acc3, 1);
        for (int c = 0; c < ENC_NC; ++c) ptx::mbar_init(bar_a2(c), 8);
        ptx::mbar_init(bar_r3, ENC_EPI / 32);
        ptx::fence_barrier_init();
    }
    if ((threadIdx.x >> 5) == 0) ptx::tmem_alloc<512>(ptx::smem_u32((const void *)tmem_slot));
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem0 = *tmem_slot;

    const int64_t n_tiles = (n_rows + TM - 1) / TM;
    const int my_tiles = (int)((n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x);   // blockIdx.x < n_tiles

    if (threadIdx.x >= ENC_EPI) {
        // ============================ control warp: one thread issues all MMAs
        if (threadIdx.x == ENC_EPI) {
            const uint32_t w0h = s0 + L::W0, w0l = w0h + L::W0_IMG, w1h = s0 + L::W1, w1l = w1h + L::W1_IMG;
            const uint32_t wnh = s0 + L::WN, wnl = wnh + (uint32_t)D16 * 256u;
            auto issue_layer0 = [&](int t) {   // x of tile t is staged -> acc1 (its columns are free: layer 1 of
                ptx::mbar_wait(bar_a(0), t & 1);   // tile t-1 completed before that tile's layer-1 epilogue began)
                ptx::tc_fence_after();
                issue_steps16(tmem0 + E16_R1, tmem0 + E16_X, w0h, w0l, H1, 0, 1, true);
                ptx::umma_commit(bar_acc(0));
            };
            // Order in the tensor pipe: L0(0) | L1(t) L0(t+1) L2(t) ...  The epilogue warps run
            // epi0(t) -> epi2(t-1) -> epi1(t): layer 1 of tile t executes under the previous tile's long sigmoid
            // epilogue, layer 0 of tile t+1 is done long before it is needed, and the tail of layer 2 (released per
            // pair of h2 blocks during epi1) drains under epi0(t+1).
            issue_layer0(0);
            for (int t = 0; t < my_tiles; ++t) {
                const uint32_t par = t & 1;
                ptx::mbar_wait(bar_a(1), par);
                ptx::tc_fence_after();
                issue_steps16(tmem0 + E16_R2, tmem0 + E16_R1, w1h, w1l, HH, 0, H1 / 16, true);
                ptx::umma_commit(bar_acc(1));
                if (t + 1 < my_tiles) issue_layer0(t + 1);      // (queued behind L1(t), which reads h1(t) from acc1's columns)
                if (t >= 1) ptx::mbar_wait(bar_r3, (t - 1) & 1);    // acc3 of tile t-1 has been read by every warp
                for (int c = 0; c < ENC_NC; ++c) {      // K steps 2c, 2c+1 of Wn against h2 blocks 2c, 2c+1
                    ptx::mbar_wait(bar_a2(c), par);
                    ptx::tc_fence_after();
                    issue_steps16(tmem0 + E16_R3, tmem0 + E16_R2, wnh, wnl, D16, 2 * c, 2 * c + 2, c == 0);
                }
                ptx::umma_commit(bar_acc3);
            }
        }
    } else {
        // ============================ input + epilogue warps
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, sub = warp >> 2, r_in = threadIdx.x & 127;
        const uint32_t tm = tmem0 + ((uint32_t)((warp & 3) * 32) << 16);
        uint8_t *wbuf = sm + L::WBUF + warp * L::WBUF_W;
        const bool vec_x = (F & 3) == 0, vec_d = (D & 3) == 0;
        float xv[16];
        auto fetch_x = [&](int64_t tile) {
            const int64_t row = tile * TM + r_in;
            const bool ok = sub == 0 && tile < n_tiles && row < n_rows;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                const int k0 = 4 * j;
                if (ok && k0 < F) {
                    const float *p = x + row * F + k0;
                    if (vec_x && k0 + 3 < F) v = __ldg(reinterpret_cast<const float4 *>(p));
                    else {
                        v.x = __ldg(p);
                        if (k0 + 1 < F) v.y = __ldg(p + 1);
                        if (k0 + 2 < F) v.z = __ldg(p + 2);
                        if (k0 + 3 < F) v.w = __ldg(p + 3);
                    }
                }
                xv[k0] = v.x; xv[k0 + 1] = v.y; xv[k0 + 2] = v.z; xv[k0 + 3] = v.w;
            }
        };
        auto stage_x = [&]() {
            if (sub == 0) {
                uint32_t o[16];
                pack_block16(xv, o, amax);
                ptx::tmem_st_32x32b_x16(tm + E16_X, o);
            }
            a_ready_arrive(bar_a(0));
        };
        fetch_x(blockIdx.x);
        int slot_i = 0;
        bool zbad = false;
        int64_t zn_row0 = 0;
        int zn_rows = -1;                               // pending row-norm combine of the previous tile
#ifdef ENC_PROFILE
        long long pc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, pt = clock64();
#endif
        stage_x();
        fetch_x(blockIdx.x + (int64_t)gridDim.x);
        // iteration it: epi0(it) | epi2(it-1) | epi1(it)   (tile index of local tile i = blockIdx.x + i * gridDim.x)
        for (int it = 0; it <= my_tiles; ++it) {
            const int64_t tile = blockIdx.x + (int64_t)it * gridDim.x;
            const uint32_t par = it & 1;
            const int64_t wrow0_cur = tile * TM + (r_in - lane);           // first row of this warp
            const int rows_valid_cur = (int)max((int64_t)0, min((int64_t)32, n_rows - wrow0_cur));
            ENC_TICK(0);
            if (it < my_tiles) {
                // ---- layer 0 epilogue: h1 (in place over acc1)
                ptx::mbar_wait(bar_acc(0), par);
                ptx::tc_fence_after();
                ENC_TICK(1);
                epilogue_blocks16<ACT_GELU_FAST>(tm + E16_R1, 0, H1 / 16, sub, 4, H1, sb0, wbuf,
                                                 (KEEP && a1) ? a1 + wrow0_cur * H1 : nullptr,
                                                 (KEEP && h1) ? h1 + wrow0_cur * H1 : nullptr, H1, rows_valid_cur, amax,
                                                 NoAfter());
                a_ready_arrive(bar_a(1));
                // ---- next tile's x -> its TMEM columns (layer 0 of this tile has completed)
                if (it + 1 < my_tiles) {
                    stage_x();
                    fetch_x(tile + 2 * (int64_t)gridDim.x);
                }
                ENC_TICK(2);
            }
            if (it >= 1) {
            // ---- layer 2 epilogue of the PREVIOUS tile: sigmoid and store z_e
            const int64_t wrow0 = wrow0_cur - (int64_t)gridDim.x * TM;
            const int rows_valid = (int)max((int64_t)0, min((int64_t)32, n_rows - wrow0));
            float *sznp_w = sznp + ((it - 1) & 1) * 4 * TM, *sznp_r = sznp + (it & 1) * 4 * TM;
            ptx::mbar_wait(bar_acc3, par ^ 1u);
            ptx::tc_fence_after();
            if (zn_rows >= 0) {   // tile it-2: every warp wrote its partial norms before layer 2 of tile it-1 could start
                if (sub == 0 && lane < zn_rows)
                    zn[zn_row0 + lane] = sqrtf(((sznp_r[r_in] + sznp_r[TM + r_in]) + sznp_r[2 * TM + r_in]) + sznp_r[3 * TM + r_in]);
                zn_rows = -1;
            }
            ENC_TICK(3);
            if constexpr (TMA) {
                // z_e (+ the search's row operand Zb = [fp16(z - mu) | 1 1 1], zn = ||z - mu||) leave through bulk
                // tensor stores: per 16-column block a thread writes its 64 bytes of z_e (64-byte-swizzle slot, 2 KB
                // per warp) and its 32 bytes of Zb (32-byte-swizzle slot, 1 KB); ONE proxy fence and ONE bulk group
                // per block, lane 0 issues both stores.  The slots are rewritten once that group has read them
                // (wait_group.read 0 after the next block's arithmetic).
                float ss0 = 0.f, ss1 = 0.f;
                const int nblk = Zb ? (D + 3 + 15) / 16 : D16 / 16;
                const int sw32 = (lane >> 2) & 1, sw64 = (lane >> 1) & 3;
                uint8_t *sl_ze = wbuf, *sl_zb = wbuf + 2048;
                const uint32_t wbuf_s = ptx::smem_u32(wbuf);
                for (int cb = sub; cb < nblk; cb += 4) {
                    const int c0 = cb << 4;
                    float v[16];
                    const bool has_ze = c0 < D;
                    if (c0 < D16) {
                        uint32_t r[16];
                        ptx::tmem_ld_32x32b_x16(tm + E16_R3 + c0, r);
                        ptx::tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const float4 bq = *reinterpret_cast<const float4 *>(sb2 + c0 + 4 * i);
                            v[4 * i] = fmaf(__uint_as_float(r[4 * i]), E16_DESCALE, bq.x);
                            v[4 * i + 1] = fmaf(__uint_as_float(r[4 * i + 1]), E16_DESCALE, bq.y);
                            v[4 * i + 2] = fmaf(__uint_as_float(r[4 * i + 2]), E16_DESCALE, bq.z);
                            v[4 * i + 3] = fmaf(__uint_as_float(r[4 * i + 3]), E16_DESCALE, bq.w);
                        }
#pragma unroll
                        for (int i = 0; i < 8; ++i) sigmoid_fast_pair(v[2 * i], v[2 * i + 1]);
                    }
                    (void)slot_i;
                    if (lane == 0) ptx::bulk_wait_read<0>();
                    __syncwarp();
                    if (has_ze) {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            *reinterpret_cast<float4 *>(sl_ze + lane * 64 + ((j ^ sw64) << 4)) =
                                make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                    }
                    if (Zb) {
                        if (c0 + 16 <= D) {           // full block: no per-element bounds
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const float4 mq = *reinterpret_cast<const float4 *>(smu + c0 + 4 * i);
                                v[4 * i] -= mq.x; v[4 * i + 1] -= mq.y; v[4 * i + 2] -= mq.z; v[4 * i + 3] -= mq.w;
                            }
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                ss0 = fmaf(v[2 * i], v[2 * i], ss0);
                                ss1 = fmaf(v[2 * i + 1], v[2 * i + 1], ss1);
                            }
                        } else {
#pragma unroll
                            for (int i = 0; i < 16; ++i) {
                                const int d = c0 + i;
                                if (d < D) {
                                    v[i] -= smu[d];
                                    ss0 = fmaf(v[i], v[i], ss0);
                                } else {
                                    v[i] = d < D + 3 ? 1.0f : 0.f;
                                }
                            }
                        }
                        uint32_t o[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const tc::op2_t pr = tc::f2op2(v[2 * i], v[2 * i + 1]);
                            o[i] = *reinterpret_cast<const uint32_t *>(&pr);
                        }
                        *reinterpret_cast<uint4 *>(sl_zb + lane * 32 + ((0 ^ sw32) << 4)) = make_uint4(o[0], o[1], o[2], o[3]);
                        *reinterpret_cast<uint4 *>(sl_zb + lane * 32 + ((1 ^ sw32) << 4)) = make_uint4(o[4], o[5], o[6], o[7]);
                    }
                    ptx::fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) {
#ifdef ENC_EXP_SKIP
                        if (!(ENC_EXP_SKIP & 1))
#endif
                        if (has_ze) ptx::tma_store_2d(&tm_ze, wbuf_s, c0, (int)wrow0);
#ifdef ENC_EXP_SKIP
                        if (!(ENC_EXP_SKIP & 2))
#endif
                        if (Zb) ptx::tma_store_2d(&tm_zb, wbuf_s + 2048u, c0, (int)wrow0);
                        ptx::bulk_commit();
                    }
                }
                const float ss = ss0 + ss1;
                if (Zb) {
                    zbad |= !(ss <= tc::OP_MAX * tc::OP_MAX);
                    // partial norms of the four column owners are combined (fixed order) after the NEXT tile's
                    // layer-1 accumulator wait, which every thread of this tile has passed its epilogue to reach
                    sznp_w[sub * TM + r_in] = ss;
                    zn_row0 = wrow0;
                    zn_rows = rows_valid;
                }
            } else {
                // per-thread store path (row pitch not 16-byte aligned, or bulk stores disabled)
                float ss = 0.f;
                const int nblk = Zb ? DKp / 16 : D16 / 16;
                for (int cb = sub; cb < nblk; cb += 4) {
                    const int c0 = cb << 4;
                    float v[16];
                    if (c0 < D16) {
                        uint32_t r[16];
                        ptx::tmem_ld_32x32b_x16(tm + E16_R3 + c0, r);
                        ptx::tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            v[i] = sigmoid_fast_f(fmaf(__uint_as_float(r[i]), E16_DESCALE, sb2[c0 + i]));
                        warp_store16(wbuf, z_e + wrow0 * D + c0, D, v, rows_valid, min(16, D - c0), vec_d);
                    }
                    if (Zb) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            const int d = c0 + i;
                            if (d < D) {
                                v[i] -= smu[d];
                                ss = fmaf(v[i], v[i], ss);
                                zbad |= !(fabsf(v[i]) <= tc::OP_MAX);
                            } else {
                                v[i] = d < D + 3 ? 1.0f : 0.f;
                            }
                        }
                        uint4 q0, q1;
                        {
                            __align__(16) tc::op2_t o[8];
#pragma unroll
                            for (int i = 0; i < 8; ++i) o[i] = tc::f2op2(v[2 * i], v[2 * i + 1]);
                            q0 = reinterpret_cast<const uint4 *>(o)[0];
                            q1 = reinterpret_cast<const uint4 *>(o)[1];
                        }
                        *reinterpret_cast<uint4 *>(wbuf + lane * 32 + ((0 ^ ((lane >> 2) & 1)) << 4)) = q0;
                        *reinterpret_cast<uint4 *>(wbuf + lane * 32 + ((1 ^ ((lane >> 2) & 1)) << 4)) = q1;
                        __syncwarp();
#pragma unroll
                        for (int it = 0; it < 2; ++it) {
                            const int f = it * 32 + lane, rw = f >> 1, j = f & 1;
                            const uint4 q = *reinterpret_cast<const uint4 *>(wbuf + rw * 32 + ((j ^ ((rw >> 2) & 1)) << 4));
                            if (rw < rows_valid)
                                *reinterpret_cast<uint4 *>(Zb + (size_t)(wrow0 + rw) * DKp + c0 + j * 8) = q;
                        }
                        __syncwarp();
                    }
                }
                if (Zb) {
                    // row norm: the four column owners of a row combine through shared memory in a fixed order
                    sznp_w[sub * TM + r_in] = ss;
                    asm volatile("bar.sync 1, 512;" ::: "memory");
                    if (sub == 0 && lane < rows_valid)
                        zn[wrow0 + lane] = sqrtf(((sznp_w[r_in] + sznp_w[TM + r_in]) + sznp_w[2 * TM + r_in]) + sznp_w[3 * TM + r_in]);
                    asm volatile("bar.sync 1, 512;" ::: "memory");
                }
            }
            // acc3 of that tile is free for layer 2 of this one
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(bar_r3);
            ENC_TICK(4);
            }
            if (it < my_tiles) {
                if (TMA && KEEP && it >= 1) {   // the bulk stores just issued must have read the staging slots
                    if (lane == 0) ptx::bulk_wait_read<0>();
                    __syncwarp();
                }
                // ---- layer 1 epilogue: h2 (in place over acc2), released to layer 2 per pair of blocks
                ptx::mbar_wait(bar_acc(1), par);
                ptx::tc_fence_after();
                ENC_TICK(5);
                epilogue_blocks16<ACT_GELU_FAST>(tm + E16_R2, 0, HH / 16, sub, 4, HH, sb1, wbuf,
                                                 (KEEP && a2) ? a2 + wrow0_cur * HH : nullptr,
                                                 (KEEP && h2) ? h2 + wrow0_cur * HH : nullptr, HH, rows_valid_cur, amax,
                                                 [&](int cb) { a_ready_arrive(bar_a2(cb >> 1)); });
                ENC_TICK(6);
            }
        }
        if (TMA && Zb) {                       // row norms of the last tile
            asm volatile("bar.sync 1, 512;" ::: "memory");
            const float *sl = sznp + ((my_tiles - 1) & 1) * 4 * TM;
            if (zn_rows >= 0 && sub == 0 && lane < zn_rows)
                zn[zn_row0 + lane] = sqrtf(((sl[r_in] + sl[TM + r_in]) + sl[2 * TM + r_in]) + sl[3 * TM + r_in]);
        }
        if (zbad && counters) atomicOr(counters + 3, 1);   // search operand out of range -> exact search
        if (TMA && lane == 0) ptx::bulk_wait<0>();   // every bulk store has landed before the CTA retires
#ifdef ENC_PROFILE
        if (blockIdx.x == 3 && (threadIdx.x & 31) == 0 && (threadIdx.x >> 5) < 8)
            printf("w%d tiles %d cyc/tile: loop %lld | wait0 %lld epi0 %lld | wait3 %lld epi2 %lld | wait1 %lld epi1 %lld\n",
                   threadIdx.x >> 5, my_tiles, pc[0] / my_tiles, pc[1] / my_tiles, pc[2] / my_tiles, pc[3] / my_tiles,
                   pc[4] / my_tiles, pc[5] / my_tiles, pc[6] / my_tiles);
#endif
    }
    // an operand left the fp16 range (or was NaN): the tf32 kernel redoes this launch
    if (!(amax < E16_MAX)) atomicOr(range_flag, 1);
    ptx::tc_fence_before();
    __syncthreads();
    if ((threadIdx.x >> 5) == 0) ptx::tmem_dealloc<512>(tmem0);
}

// Lipschitz normalisation (v5:6-12) + the tf32 Wn image of the fallback kernel (one 16-byte piece per lane) +
// reset of the range flag.  One warp per row of W[D,128]; rows [D, D16) of the image are zero.
__global__ void lipschitz_prep_img_kernel(const float *__restrict__ W, const float *__restrict__ ci, int D, int D16,
                                          float *__restrict__ scale, float *__restrict__ Wn, uint8_t *__restrict__ img,
                                          int *__restrict__ range_flag) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (blockIdx.x == 0 && threadIdx.x == 0) *range_flag = 0;
    if (row >= D16) return;
    float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row < D) {
        w = *reinterpret_cast<const float4 *>(W + (size_t)row * HH + 4 * lane);
        float s = (fabsf(w.x) + fabsf(w.y)) + (fabsf(w.z) + fabsf(w.w));
        s = warp_sum(s);
        const float sc = fminf(1.0f, softplus_f(ci[row]) / s);
        if (lane == 0) scale[row] = sc;
        w.x *= sc; w.y *= sc; w.z *= sc; w.w *= sc;
        *reinterpret_cast<float4 *>(Wn + (size_t)row * HH + 4 * lane) = w;
    }
    const int kb = lane >> 3, j = lane & 7;
    uint32_t h[4], l[4];
    split_tf32(w.x, h[0], l[0]); split_tf32(w.y, h[1], l[1]); split_tf32(w.z, h[2], l[2]); split_tf32(w.w, h[3], l[3]);
    uint8_t *base = img + (size_t)kb * D16 * 256;
    const int off = row * 128 + ((j ^ (row & 7)) << 4);
    *reinterpret_cast<uint4 *>(base + off) = make_uint4(h[0], h[1], h[2], h[3]);
    *reinterpret_cast<uint4 *>(base + D16 * 128 + off) = make_uint4(l[0], l[1], l[2], l[3]);
}

// scratch of the fused encoder: the fallback kernel's Wn image + the range flag
inline size_t enc_scratch_bytes(int D) { return enc_img_bytes(D) + 256; }
inline int *enc_range_flag(uint8_t *img, int D) { return reinterpret_cast<int *>(img + enc_img_bytes(D)); }

// Whole encoder: Lipschitz prep -> fp16-split fused kernel -> (gated) tf32 fused kernel.
// img: enc_scratch_bytes(D) of 1024-byte aligned scratch.  mu/Zb/zn/counters: optional search-operand fusion.
inline int encoder_fused16_launch(const lipvq_dims_t *d, const lipvq_params_t *p, const float *x, int64_t n, float *a1,
                                  float *h1, float *a2, float *h2, float *scale, float *Wn, float *z_e, uint8_t *img,
                                  const float *mu, tc::op_t *Zb, float *zn, int DKp, int *counters, cudaStream_t st) {
    const int D16 = pad16(d->D);
    int *flag = enc_range_flag(img, d->D);
    lipschitz_prep_img_kernel<<<(D16 + 7) / 8, 256, 0, st>>>(p->lat_W, p->lat_ci, d->D, D16, scale, Wn, img, flag);
    LIPVQ_LAUNCH_OK();
    const bool keep = a1 || h1 || a2 || h2;
    static const bool no_tma = [] { const char *e = getenv("LIPVQ_ENC_TMA"); return e && !strcmp(e, "0"); }();
    static const bool force_tf32 = [] { const char *e = getenv("LIPVQ_ENC"); return e && !strcmp(e, "tf32"); }();
    CUtensorMap tm_ze, tm_zb, tm_ze64;   // z_e boxes: 32 bytes x 32 rows (tf32 kernel), 64 bytes x 32 rows (this one)
    memset(&tm_ze64, 0, sizeof(tm_ze64));
    memset(&tm_ze, 0, sizeof(tm_ze));
    memset(&tm_zb, 0, sizeof(tm_zb));
    const int use_tma = !no_tma && (d->D & 3) == 0 && n < (int64_t)0x7fffffff - TM && (!Zb || (DKp & 7) == 0);
    if (use_tma) {
        LIPVQ_TRY(make_store_tmap(&tm_ze, z_e, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, (uint64_t)n, (uint64_t)d->D,
                                  (uint64_t)d->D));
        LIPVQ_TRY(make_store_tmap(&tm_ze64, z_e, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, (uint64_t)n, (uint64_t)d->D,
                                  (uint64_t)d->D, 64));
        if (Zb) {
#ifdef LIPVQ_VQ_BF16
            const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
#else
            const CUtensorMapDataType dt = CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
#endif
            const uint64_t np = (uint64_t)((n + TM - 1) / TM * TM);
            LIPVQ_TRY(make_store_tmap(&tm_zb, Zb, dt, 2, np, (uint64_t)DKp, (uint64_t)DKp));
        }
    }
    if (!force_tf32) {
        ProfScope prof("enc16_fused_kernel", st);
        auto launch = [&](auto kern) -> int {
            LIPVQ_TRY(set_smem(kern, Enc16Smem::TOTAL));
            kern<<<grid_for(n), ENC_THR, Enc16Smem::TOTAL, st>>>(x, p->enc0_w, p->enc0_b, p->enc2_w, p->enc2_b, Wn,
                                                                 p->lat_b, n, d->F, d->D, a1, h1, a2, h2, z_e, mu, Zb, zn,
                                                                 DKp, counters, flag, tm_ze64, tm_zb);
            return LIPVQ_OK;
        };
        if (use_tma) LIPVQ_TRY(keep ? launch(enc16_fused_kernel<true, true>) : launch(enc16_fused_kernel<true, false>));
        else LIPVQ_TRY(keep ? launch(enc16_fused_kernel<false, true>) : launch(enc16_fused_kernel<false, false>));
        LIPVQ_LAUNCH_OK();
    }
    {   // tf32 kernel: runs only when the flag is set (returns at once otherwise)
        ProfScope prof(force_tf32 ? "enc_fused_kernel (tf32)" : "enc_fused_kernel (tf32 fallback, gated)", st);
        const int *gate = force_tf32 ? nullptr : flag;
        auto launch = [&](auto kern) -> int {
            LIPVQ_TRY(set_smem(kern, EncSmem::TOTAL));
            kern<<<grid_for(n), ENC_THR, EncSmem::TOTAL, st>>>(x, p->enc0_w, p->enc0_b, p->enc2_w, p->enc2_b, img,
                                                               p->lat_b, n, d->F, d->D, a1, h1, a2, h2, z_e, mu, Zb, zn,
                                                               DKp, counters, gate, tm_ze, tm_zb);
            return LIPVQ_OK;
        };
        if (use_tma) LIPVQ_TRY(keep ? launch(enc_fused_kernel<true, true>) : launch(enc_fused_kernel<true, false>));
        else LIPVQ_TRY(keep ? launch(enc_fused_kernel<false, true>) : launch(enc_fused_kernel<false, false>));
        LIPVQ_LAUNCH_OK();
    }
    return LIPVQ_OK;
}

}  // namespace mlp
}  // namespace lipvq
