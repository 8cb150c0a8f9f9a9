// mlp_enc.cuh -- the whole encoder in ONE persistent kernel (v5:54-60, 6-12, 22-24, called v5:71-72):
//   x[N,F] -> h1 = GELU(x W0^T + b0) -> h2 = GELU(h1 W1^T + b1) -> z_e = sigmoid(h2 Wn^T + b)
// h1 and h2 never leave the SM.  Every GEMM is 3xTF32 on tcgen05 with the A operand in TMEM (mlp_tc.cuh).
//
// Shared memory cannot hold the hi/lo images of all three weights (288 KB at D=208), so W0 and W1 stay
// resident (80 KB) and the Lipschitz-normalised Wn is streamed from L2 per tile in K blocks of 32 (all D16
// rows, 52 KB) through a 2-stage ring of 1-D bulk copies (cp.async.bulk + mbarrier).  Chunking along K keeps
// every MMA at N = D16: measured, a tcgen05.mma with A in TMEM costs ~80 cycles however small N is, so
// N-chunks of 48 made layer 2 four times slower than K-chunks.
//
// Warps 0..15 (512 threads): staging + epilogues (warp%4 = TMEM lane quarter, warp/4 = 16-column block
// owner).  Warp 16: control -- one thread issues every tcgen05.mma and every bulk copy.
// TMEM columns: h1_hi [0,64) h1_lo [64,128) -> later h2_lo [0,128) | acc1 [128,192) -> acc2 -> h2_hi [128,256) |
//               acc3 [256,256+D16) | x_hi [464,480) x_lo [480,496).
// x of tile t+1 has its own columns, and acc1 / h1 reuse columns that are dead once tile t's layer-2 MMAs have
// completed, so layer 0 of tile t+1 is issued while the epilogue warps are still in tile t's last epilogue.
#pragma once
#include "mlp_tc.cuh"
#include "vq_tensor.cuh"

namespace lipvq {
namespace mlp {

constexpr int ENC_EPI = 512;             // epilogue threads
constexpr int ENC_THR = ENC_EPI + 32;    // + control warp
constexpr int ENC_NC = HH / 32;            // Wn is streamed per K block of 32 (all D16 rows): 4 chunks per tile
constexpr int ENC_DMAX = 208;
constexpr int ENC_NCMAX = 1;
constexpr int ENC_STAGE_BYTES = ENC_DMAX * 256;   // one K block: hi (D16 x 128 B) + lo

struct EncSmem {
    static constexpr int W0 = 0, W0_IMG = H1 * 128;
    static constexpr int W1 = W0 + 2 * W0_IMG, W1_IMG = HH * 128 * 2;
    static constexpr int RING = W1 + 2 * W1_IMG;
    static constexpr int WBUF = RING + 2 * ENC_STAGE_BYTES;        // 16 warps x 2 KB store-transpose buffers
    static constexpr int BIAS = WBUF + 16 * 2048;                  // b0[64] b1[128] b[256]
    static constexpr int MU = BIAS + (H1 + HH + 256) * 4;          // codebook mean mu[<=256+16] (search operand fusion)
    static constexpr int ZNP = MU + 272 * 4;                       // [4 subs][128 rows] partial row norms
    static constexpr int BAR = ZNP + 4 * TM * 4;                   // mbarriers (8 B each), then the TMEM slot
    static constexpr int NBAR = 3 + 2 + ENC_NCMAX + 4 + ENC_NC;    // a_ready[3] acc_ready[2] acc3[NC] w_full[2] w_empty[2] a2_ready[4]
    static constexpr int TOTAL = BAR + NBAR * 8 + 16 + 1024;
};

__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
// epilogue thread: my part of an A operand is in TMEM -> tell the control thread
// (one arrival per WARP: 512 threads arriving on one mbarrier serialise on its shared-memory word)
__device__ __forceinline__ void a_ready_arrive(uint32_t bar) {
    ptx::tmem_st_wait();
    ptx::tc_fence_before();
    __syncwarp();
    if ((threadIdx.x & 31) == 0) ptx::mbar_arrive(bar);
}

// Epilogue over the 16-column blocks cb in [cb0, cb1) with cb % nsub == sub, for this thread's row:
//   v = acc + bias -> [pre-activation rows to global] -> act -> [activation rows to global] -> [next A operand]
// pre/post: &dst[warp_row0 * ld] or nullptr; stores go through the warp's transpose buffer (warp_store16).
// after(cb) runs once block cb is complete (its operand stored).  (Prefetching the next block's TMEM load was
// measured slower: the extra live registers spill at the 96-register cap of a 17-warp CTA.)
struct NoAfter { __device__ __forceinline__ void operator()(int) const {} };
template <int ACT, class After>
__device__ __forceinline__ void epilogue_blocks(uint32_t t_acc, int cb0, int cb1, int sub, int nsub, int n_real,
                                                const float *sbias, bool write_next, uint32_t t_next_hi,
                                                uint32_t t_next_lo, uint8_t *wbuf, float *__restrict__ pre,
                                                float *__restrict__ post, int ld, int rows_valid, bool vec,
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
            v[4 * i] = __uint_as_float(r[4 * i]) + bq.x; v[4 * i + 1] = __uint_as_float(r[4 * i + 1]) + bq.y;
            v[4 * i + 2] = __uint_as_float(r[4 * i + 2]) + bq.z; v[4 * i + 3] = __uint_as_float(r[4 * i + 3]) + bq.w;
        }
        const int valid = min(16, n_real - c0);
        if (pre) warp_store16(wbuf, pre + c0, ld, v, rows_valid, valid, vec);
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = apply_act<ACT>(v[i]);
        if (post) warp_store16(wbuf, post + c0, ld, v, rows_valid, valid, vec);
        if (write_next) store_split16(t_next_hi + c0, t_next_lo + c0, v);
        after(cb);
    }
}

template <bool TMA, bool KEEP>   // KEEP: a1/h1/a2/h2 are written for backward (training forward)
__global__ void __launch_bounds__(ENC_THR, 1)
enc_fused_kernel(const float *__restrict__ x, const float *__restrict__ W0, const float *__restrict__ b0,
                 const float *__restrict__ W1, const float *__restrict__ b1, const uint8_t *__restrict__ wn_img,
                 const float *__restrict__ b2, int64_t n_rows, int F, int D, float *__restrict__ a1,
                 float *__restrict__ h1, float *__restrict__ a2, float *__restrict__ h2, float *__restrict__ z_e,
                 // optional fusion of the search's row-operand pass (vq_tensor.cuh pass 1): Zb/zn/flag outputs
                 const float *__restrict__ mu, tc::op_t *__restrict__ Zb, float *__restrict__ zn, int DKp,
                 int *__restrict__ counters,
                 // gate != nullptr: run only if *gate != 0 (fallback of the fp16-split kernel, mlp_enc16.cuh)
                 const int *__restrict__ gate,
                 // bulk tensor stores of z_e (fp32, box 8 x 32, 32-byte swizzle) and Zb (16-bit, box 16 x 32)
                 const __grid_constant__ CUtensorMap tm_ze, const __grid_constant__ CUtensorMap tm_zb) {
    using L = EncSmem;
    if (gate != nullptr && *gate == 0) return;                       // (uniform over the grid)
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = ptx::smem_u32(smem_raw);
    uint8_t *sm = smem_raw + (((raw + 1023u) & ~1023u) - raw);
    const uint32_t s0 = ptx::smem_u32(sm);
    float *sb0 = reinterpret_cast<float *>(sm + L::BIAS), *sb1 = sb0 + H1, *sb2 = sb1 + HH;
    const uint32_t bars = s0 + L::BAR;
    auto bar_a = [&](int l) { return bars + 8u * l; };               // A operand of layer l ready (16 warp arrivals)
    auto bar_acc = [&](int l) { return bars + 8u * (3 + l); };       // accumulator of layer 0/1 ready
    auto bar_acc3 = [&](int c) { return bars + 8u * (5 + c); };      // chunk c of layer 2 ready
    auto bar_wfull = [&](int s) { return bars + 8u * (5 + ENC_NCMAX + s); };
    auto bar_wempty = [&](int s) { return bars + 8u * (7 + ENC_NCMAX + s); };
    // h2 columns [32c, 32c+32) (K block c of layer 2) ready: 2 column owners x 4 lane quarters = 8 warp arrivals, so
    // the layer-2 MMAs of the first K blocks run while the layer-1 epilogue is still producing the later ones
    auto bar_a2 = [&](int c) { return bars + 8u * (9 + ENC_NCMAX + c); };
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(sm + L::BAR + L::NBAR * 8);

    const int D16 = pad16(D), NC = ENC_NC;
    stage_weight(sm + L::W0, sm + L::W0 + L::W0_IMG, W0, H1, F, H1, 1);
    stage_weight(sm + L::W1, sm + L::W1 + L::W1_IMG, W1, HH, H1, HH, 2);
    for (int i = threadIdx.x; i < H1; i += ENC_THR) sb0[i] = b0[i];
    for (int i = threadIdx.x; i < HH; i += ENC_THR) sb1[i] = b1[i];
    for (int i = threadIdx.x; i < 256; i += ENC_THR) sb2[i] = i < D ? b2[i] : 0.f;
    float *smu = reinterpret_cast<float *>(sm + L::MU), *sznp = reinterpret_cast<float *>(sm + L::ZNP);
    for (int i = threadIdx.x; i < 272; i += ENC_THR) smu[i] = (Zb && i < D) ? mu[i] : 0.f;
    ptx::fence_proxy_async();
    if (threadIdx.x == 0) {
        for (int l = 0; l < 2; ++l) ptx::mbar_init(bar_a(l), ENC_EPI / 32);
        ptx::mbar_init(bar_a(2), 1);   // (unused: layer 2 is released per K block, bar_a2)
        for (int l = 0; l < 2; ++l) ptx::mbar_init(bar_acc(l), 1);
        for (int c = 0; c < ENC_NCMAX; ++c) ptx::mbar_init(bar_acc3(c), 1);
        for (int s = 0; s < 2; ++s) { ptx::mbar_init(bar_wfull(s), 1); ptx::mbar_init(bar_wempty(s), 1); }
        for (int c = 0; c < ENC_NC; ++c) ptx::mbar_init(bar_a2(c), 8);
        ptx::fence_barrier_init();
    }
    if ((threadIdx.x >> 5) == 0) ptx::tmem_alloc<512>(ptx::smem_u32((const void *)tmem_slot));
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem0 = *tmem_slot;

    const int64_t n_tiles = (n_rows + TM - 1) / TM;
    const int my_tiles = (int)((n_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x);   // blockIdx.x < n_tiles
    const int k0_steps = (F + 7) / 8;

    if (threadIdx.x >= ENC_EPI) {
        // ============================ control warp: one thread issues all MMAs and weight copies
        if (threadIdx.x == ENC_EPI) {
            const int total = my_tiles * NC;
            int g_load = 0;
            auto load_next = [&]() {   // chunk g_load -> ring stage g_load & 1 (waits until that stage is free)
                const int s = g_load & 1, c = g_load % NC;
                if (g_load >= 2) ptx::mbar_wait(bar_wempty(s), ((g_load >> 1) - 1) & 1);
                const uint32_t bytes = (uint32_t)D16 * 256u;
                ptx::mbar_arrive_expect_tx(bar_wfull(s), bytes);
                bulk_g2s(s0 + L::RING + s * ENC_STAGE_BYTES, wn_img + (size_t)c * bytes, bytes, bar_wfull(s));
                ++g_load;
            };
            auto issue_layer0 = [&](int t) {   // x of tile t is staged -> acc1
                ptx::mbar_wait(bar_a(0), t & 1);
                ptx::tc_fence_after();
                issue_layer(tmem0 + 128, tmem0 + 464, tmem0 + 480, s0 + L::W0, s0 + L::W0 + L::W0_IMG, H1, k0_steps,
                            bar_acc(0));
            };
            for (int t = 0; t < my_tiles; ++t) {
                const uint32_t par = t & 1;
                while (g_load < total && g_load < t * NC + 2) load_next();
                if (t == 0) issue_layer0(0);
                ptx::mbar_wait(bar_a(1), par);
                ptx::tc_fence_after();
                issue_layer(tmem0 + 128, tmem0 + 0, tmem0 + 64, s0 + L::W1, s0 + L::W1 + L::W1_IMG, HH, H1 / 8,
                            bar_acc(1));
                const uint32_t idesc = ptx::umma_idesc_tf32(TM, D16);
                for (int c = 0; c < NC; ++c) {      // K block c of Wn against h2 columns [32c, 32c+32)
                    const int g = t * NC + c, s = g & 1;
                    ptx::mbar_wait(bar_a2(c), par);
                    ptx::mbar_wait(bar_wfull(s), (g >> 1) & 1);
                    ptx::tc_fence_after();
                    const uint32_t st = s0 + L::RING + s * ENC_STAGE_BYTES;
                    const uint64_t bh = ptx::umma_desc_sw128(st), bl = ptx::umma_desc_sw128(st + D16 * 128);
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) {
                        const uint32_t a_lo = tmem0 + 32 * c + 8 * ks, a_hi = a_lo + 128;
                        ptx::umma_tf32_ts(tmem0 + 256, a_hi, bh + 2 * ks, idesc, (c | ks) ? 1u : 0u);
                        ptx::umma_tf32_ts(tmem0 + 256, a_lo, bh + 2 * ks, idesc, 1u);
                        ptx::umma_tf32_ts(tmem0 + 256, a_hi, bl + 2 * ks, idesc, 1u);
                    }
                    ptx::umma_commit(bar_wempty(s));
                    if (c == NC - 1) ptx::umma_commit(bar_acc3(0));
                    while (g_load < total && g_load <= g + 1) load_next();
                }
                if (t + 1 < my_tiles) {
                    // h2 is dead once this tile's layer-2 MMAs have completed: layer 0 of the next tile (its x was
                    // staged before the epilogue warps started waiting for acc3) overlaps this tile's last epilogue
                    ptx::mbar_wait(bar_acc3(0), par);
                    issue_layer0(t + 1);
                }
            }
        }
    } else {
        // ============================ input + epilogue warps
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, sub = warp >> 2, r_in = threadIdx.x & 127;
        const uint32_t tm = tmem0 + ((uint32_t)((warp & 3) * 32) << 16);
        uint8_t *wbuf = sm + L::WBUF + warp * 2048;
        const bool vec_x = (F & 3) == 0, vec_d = (D & 3) == 0;
        // x: the sub-0 warp of each lane quarter loads its rows directly (16 floats per row, prefetched one tile
        // ahead).  x_hi / x_lo live in acc3 blocks 0 and 4, both owned by sub 0, so no other warp touches them.
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
        fetch_x(blockIdx.x);
        int t = 0, slot_i = 0;
        bool zbad = false;
        constexpr bool keeps_acts = KEEP;               // their stores reuse the staging slots synchronously
        int64_t zn_row0 = 0;
        int zn_rows = -1;                               // pending row-norm combine of the previous tile
#ifdef ENC_PROFILE
        long long pc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, pt = clock64();
#define ENC_TICK(i) do { const long long _n = clock64(); pc[i] += _n - pt; pt = _n; } while (0)
#else
#define ENC_TICK(i) do { } while (0)
#endif
        for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++t) {
            const uint32_t par = t & 1;
            const int64_t wrow0 = tile * TM + (r_in - lane);               // first row of this warp
            const int rows_valid = (int)max((int64_t)0, min((int64_t)32, n_rows - wrow0));
            if (TMA && keeps_acts) {   // the previous tile's bulk stores must have read the slots
                if (lane == 0) ptx::bulk_wait_read<0>();
                __syncwarp();
            }
            // ---- x -> A operand (hi/lo): only the first tile stages here, later tiles were staged during the
            //      previous tile (below, before its acc3 wait)
            if (t == 0) {
                if (sub == 0) store_split16(tm + 464, tm + 480, xv);
                a_ready_arrive(bar_a(0));
                fetch_x(tile + gridDim.x);
            }
            ENC_TICK(0);
            // ---- layer 0 epilogue: h1
            ptx::mbar_wait(bar_acc(0), par);
            ptx::tc_fence_after();
            ENC_TICK(1);
            epilogue_blocks<ACT_GELU_FAST>(tm + 128, 0, H1 / 16, sub, 4, H1, sb0, true, tm + 0, tm + 64, wbuf,
                                           (KEEP && a1) ? a1 + wrow0 * H1 : nullptr, (KEEP && h1) ? h1 + wrow0 * H1 : nullptr, H1,
                                           rows_valid, true, NoAfter());
            a_ready_arrive(bar_a(1));
            ENC_TICK(2);
            // ---- layer 1 epilogue: h2 (hi written in place over the accumulator columns just read)
            ptx::mbar_wait(bar_acc(1), par);
            ptx::tc_fence_after();
            if (zn_rows >= 0) {   // (every warp passed the previous tile's last epilogue before acc2 could complete)
                if (sub == 0 && lane < zn_rows)
                    zn[zn_row0 + lane] = sqrtf(((sznp[r_in] + sznp[TM + r_in]) + sznp[2 * TM + r_in]) + sznp[3 * TM + r_in]);
                zn_rows = -1;
            }
            ENC_TICK(3);
            epilogue_blocks<ACT_GELU_FAST>(tm + 128, 0, HH / 16, sub, 4, HH, sb1, true, tm + 128, tm + 0, wbuf,
                                           (KEEP && a2) ? a2 + wrow0 * HH : nullptr, (KEEP && h2) ? h2 + wrow0 * HH : nullptr, HH,
                                           rows_valid, true, [&](int cb) { a_ready_arrive(bar_a2(cb >> 1)); });
            // ---- next tile's x -> its own TMEM columns (layer 0 of tile t+1 runs during this tile's last epilogue)
            if (tile + gridDim.x < n_tiles) {
                if (sub == 0) store_split16(tm + 464, tm + 480, xv);
                a_ready_arrive(bar_a(0));
                fetch_x(tile + 2 * (int64_t)gridDim.x);
            }
            ENC_TICK(4);
            // ---- layer 2 epilogue: sigmoid and store z_e
            ptx::mbar_wait(bar_acc3(0), par);
            ptx::tc_fence_after();
            ENC_TICK(5);
            if constexpr (TMA) {
                // z_e (+ the search's row operand Zb = [fp16(z - mu) | 1 1 1], zn = ||z - mu||) leave through bulk
                // tensor stores: a thread only writes its 16 columns into the warp's staging slots (two 1 KB
                // slots in the 32-byte-swizzle pattern the tensor map expects); lane 0 issues the store.  A slot
                // is rewritten once the store issued two stores ago has read it (wait_group.read 1).
                float ss = 0.f;
                const int nblk = Zb ? (D + 3 + 15) / 16 : D16 / 16;
                const int sw = (lane >> 2) & 1;
                auto slot_take = [&]() -> uint8_t * {
                    if (lane == 0) ptx::bulk_wait_read<1>();
                    __syncwarp();
                    uint8_t *sl = wbuf + (slot_i & 1) * 1024;
                    ++slot_i;
                    return sl;
                };
                const uint32_t wbuf_s = ptx::smem_u32(wbuf);
                auto slot_send = [&](const CUtensorMap *tmap, uint8_t *sl, int c, int64_t row) {
                    ptx::fence_proxy_async();
                    __syncwarp();
#ifdef ENC_EXP_SKIP
                    if (tmap == &tm_ze ? (ENC_EXP_SKIP & 1) : (ENC_EXP_SKIP & 2)) return;
#endif
                    if (lane == 0) {
                        ptx::tma_store_2d(tmap, wbuf_s + (uint32_t)(sl - wbuf), c, (int)row);
                        ptx::bulk_commit();
                    }
                };
                for (int cb = sub; cb < nblk; cb += 4) {
                    const int c0 = cb << 4;
                    float v[16];
                    if (c0 < D16) {
                        uint32_t r[16];
                        ptx::tmem_ld_32x32b_x16(tm + 256 + c0, r);
                        ptx::tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const float4 bq = *reinterpret_cast<const float4 *>(sb2 + c0 + 4 * i);
                            v[4 * i] = __uint_as_float(r[4 * i]) + bq.x; v[4 * i + 1] = __uint_as_float(r[4 * i + 1]) + bq.y;
                            v[4 * i + 2] = __uint_as_float(r[4 * i + 2]) + bq.z; v[4 * i + 3] = __uint_as_float(r[4 * i + 3]) + bq.w;
                        }
#pragma unroll
                        for (int i = 0; i < 16; ++i) v[i] = sigmoid_fast_f(v[i]);
#pragma unroll
                        for (int hf = 0; hf < 2; ++hf) {
                            if (c0 + 8 * hf >= D) break;
                            uint8_t *sl = slot_take();
                            *reinterpret_cast<float4 *>(sl + lane * 32 + ((0 ^ sw) << 4)) =
                                make_float4(v[8 * hf], v[8 * hf + 1], v[8 * hf + 2], v[8 * hf + 3]);
                            *reinterpret_cast<float4 *>(sl + lane * 32 + ((1 ^ sw) << 4)) =
                                make_float4(v[8 * hf + 4], v[8 * hf + 5], v[8 * hf + 6], v[8 * hf + 7]);
                            slot_send(&tm_ze, sl, c0 + 8 * hf, wrow0);
                        }
                    }
                    if (Zb) {
                        if (c0 + 16 <= D) {           // full block: no per-element bounds
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const float4 mq = *reinterpret_cast<const float4 *>(smu + c0 + 4 * i);
                                v[4 * i] -= mq.x; v[4 * i + 1] -= mq.y; v[4 * i + 2] -= mq.z; v[4 * i + 3] -= mq.w;
                            }
#pragma unroll
                            for (int i = 0; i < 16; ++i) ss = fmaf(v[i], v[i], ss);
                        } else {
#pragma unroll
                            for (int i = 0; i < 16; ++i) {
                                const int d = c0 + i;
                                if (d < D) {
                                    v[i] -= smu[d];
                                    ss = fmaf(v[i], v[i], ss);
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
                        uint8_t *sl = slot_take();
                        *reinterpret_cast<uint4 *>(sl + lane * 32 + ((0 ^ sw) << 4)) = make_uint4(o[0], o[1], o[2], o[3]);
                        *reinterpret_cast<uint4 *>(sl + lane * 32 + ((1 ^ sw) << 4)) = make_uint4(o[4], o[5], o[6], o[7]);
                        slot_send(&tm_zb, sl, c0, wrow0);
                    }
                }
                if (Zb) {
                    // any |z - mu| beyond the 16-bit operand range (or a NaN) makes the partial norm fail this test
                    zbad |= !(ss <= tc::OP_MAX * tc::OP_MAX);
                    // the four column owners of a row leave their partial norms in shared memory; they are combined
                    // (fixed order) after the NEXT tile's layer-1 accumulator wait, which every thread of this tile
                    // has passed its epilogue to reach -- no extra barrier
                    sznp[sub * TM + r_in] = ss;
                    zn_row0 = wrow0;
                    zn_rows = rows_valid;
                }
            } else if (Zb == nullptr) {
                epilogue_blocks<ACT_SIGMOID_FAST>(tm + 256, 0, D16 / 16, sub, 4, D, sb2, false, 0u, 0u, wbuf, nullptr,
                                                  z_e + wrow0 * D, D, rows_valid, vec_d, NoAfter());
            } else {
                // z_e plus the search's row operand: Zb = [fp16(z - mu) | 1 1 1 | 0...], zn = ||z - mu||
                float ss = 0.f;
                for (int cb = sub; cb < DKp / 16; cb += 4) {
                    const int c0 = cb << 4;
                    float v[16];
                    if (c0 < D16) {
                        uint32_t r[16];
                        ptx::tmem_ld_32x32b_x16(tm + 256 + c0, r);
                        ptx::tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 16; ++i) v[i] = sigmoid_fast_f(__uint_as_float(r[i]) + sb2[c0 + i]);
                        warp_store16(wbuf, z_e + wrow0 * D + c0, D, v, rows_valid, min(16, D - c0), vec_d);
                    }
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
                    // 32 bytes of fp16 per row, transposed through the warp buffer: 16 rows x 32 B per store
                    uint4 h0, h1v;
                    {
                        __align__(16) tc::op2_t o[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) o[i] = tc::f2op2(v[2 * i], v[2 * i + 1]);
                        h0 = reinterpret_cast<const uint4 *>(o)[0];
                        h1v = reinterpret_cast<const uint4 *>(o)[1];
                    }
                    *reinterpret_cast<uint4 *>(wbuf + lane * 32 + ((0 ^ ((lane >> 2) & 1)) << 4)) = h0;
                    *reinterpret_cast<uint4 *>(wbuf + lane * 32 + ((1 ^ ((lane >> 2) & 1)) << 4)) = h1v;
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
                // row norm: the four column owners of a row combine through shared memory in a fixed order
                sznp[sub * TM + r_in] = ss;
                asm volatile("bar.sync 1, 512;" ::: "memory");
                if (sub == 0 && lane < rows_valid)
                    zn[wrow0 + lane] = sqrtf(((sznp[r_in] + sznp[TM + r_in]) + sznp[2 * TM + r_in]) + sznp[3 * TM + r_in]);
            }
            ENC_TICK(6);
        }
        if (TMA && Zb) {                       // row norms of the last tile
            asm volatile("bar.sync 1, 512;" ::: "memory");
            if (zn_rows >= 0 && sub == 0 && lane < zn_rows)
                zn[zn_row0 + lane] = sqrtf(((sznp[r_in] + sznp[TM + r_in]) + sznp[2 * TM + r_in]) + sznp[3 * TM + r_in]);
        }
        if (zbad && counters) atomicOr(counters + 3, 1);   // operand out of fp16 range -> the search goes exact for this launch
        if (TMA && lane == 0) ptx::bulk_wait<0>();   // every bulk store has landed before the CTA retires
#ifdef ENC_PROFILE
        if (blockIdx.x == 3 && (threadIdx.x & 31) == 0 && (threadIdx.x >> 5) < 8)
            printf("w%d tiles %d cyc/tile: xstage %lld | wait0 %lld epi0 %lld | wait1 %lld epi1 %lld | wait2 %lld epi2 %lld\n",
                   threadIdx.x >> 5, t, pc[0] / t, pc[1] / t, pc[2] / t, pc[3] / t, pc[4] / t, pc[5] / t, pc[6] / t);
#endif
    }
    ptx::tc_fence_before();
    __syncthreads();
    if ((threadIdx.x >> 5) == 0) ptx::tmem_dealloc<512>(tmem0);
}

inline size_t enc_img_bytes(int D) { return (size_t)ENC_NC * pad16(D) * 256; }

// row-major [rows, cols] tensor of `esize`-byte elements, box = {32 bytes of columns, 32 rows}, 32-byte swizzle
inline int make_store_tmap(CUtensorMap *m, const void *ptr, CUtensorMapDataType dt, int esize, uint64_t rows,
                           uint64_t cols, uint64_t pitch_elems, int box_bytes = 32) {
    tc::EncodeTiledFn fn = tc::encode_fn();
    if (!fn) return LIPVQ_ECUDA;
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {pitch_elems * (uint64_t)esize};
    cuuint32_t box[2] = {(cuuint32_t)(box_bytes / esize), 32};   // swizzle span = box row bytes (32 or 64)
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(m, dt, 2, const_cast<void *>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    box_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? LIPVQ_OK : LIPVQ_ECUDA;
}

}  // namespace mlp
}  // namespace lipvq
