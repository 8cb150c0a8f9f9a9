// mlp_tc.cuh -- the tokenizer's MLP chains on the 5th-gen tensor cores (north_star item 4).
//
//   enc12 : x[N,F] -> h1 = GELU(x W0^T + b0) -> h2 = GELU(h1 W1^T + b1)            (v5:54-59, 71)
//   lat   : h2[N,128] -> z_e = sigmoid(h2 Wn^T + b),  Wn = W * lipschitz scale     (v5:6-12, 22-24)
//   (the decoder runs once per CODE on the CUDA cores: decoder_code_table_kernel, misc_kernels.cuh)
//
// fp32 parity (1e-5) rules out one tf32/bf16 pass, so every GEMM runs as 3xTF32:
//   A*B ~= A_hi*B_hi + A_lo*B_hi + A_hi*B_lo   (hi = tf32-rounded value, lo = exact remainder)
// accumulated in fp32 in TMEM (error ~2^-21).  Activations never leave the SM between layers: the
// epilogue reads the accumulator with tcgen05.ld, applies bias + exact-erf GELU / sigmoid, splits into
// hi/lo and writes both back to TMEM with tcgen05.st as the NEXT layer's A operand (tcgen05.mma with A in
// TMEM, B in shared memory).  Weights are split once per CTA into hi/lo and kept in shared memory in the
// K-major 128-byte-swizzle layout UMMA expects.  One persistent CTA per SM, 128 rows per tile.
#pragma once
#include "common.cuh"
#include "ptx_sm100.cuh"

namespace lipvq {
namespace mlp {

constexpr int TM = 128;       // rows per tile (UMMA M)
constexpr int NTHR = 256;     // 8 warps: warp%4 = TMEM lane quarter, warp/4 = which 16-column blocks
constexpr int H1 = 64;        // encoder.0 / decoder.0 width (fixed in the reference, v5:55,63)
constexpr int HH = 128;       // hidden_dim supported by the tensor path
constexpr int STG_BYTES = TM * 128;   // one [128 rows x 32 fp32] staging block

__device__ __forceinline__ void split_tf32(float v, uint32_t &hi, uint32_t &lo) {
    const uint32_t h = (__float_as_uint(v) + 0x1000u) & 0xffffe000u;   // round to nearest tf32
    hi = h;
    lo = __float_as_uint(v - __uint_as_float(h));                       // exact in fp32
}

// W[n][k] (n < n_real, k < k_real, dense row-major) -> hi/lo operand images in shared memory.
// Image layout: K blocks of 32 fp32 (128 B per row); block kb at kb*n_pad*128; row n at n*128;
// 16-byte chunk j of a row at ((j ^ (n & 7)) * 16)  == TMA/UMMA SWIZZLE_128B, K-major, SBO = 1024.
__device__ __forceinline__ void stage_weight(uint8_t *hi, uint8_t *lo, const float *__restrict__ W, int n_real,
                                             int k_real, int n_pad, int k_blocks) {
    const int total = n_pad * k_blocks * 8;
    for (int c = threadIdx.x; c < total; c += blockDim.x) {
        const int j = c & 7, n = (c >> 3) % n_pad, kb = (c >> 3) / n_pad;
        uint32_t h[4], l[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int k = kb * 32 + j * 4 + i;
            const float v = (n < n_real && k < k_real) ? __ldg(W + (size_t)n * k_real + k) : 0.f;
            split_tf32(v, h[i], l[i]);
        }
        const int off = kb * n_pad * 128 + n * 128 + ((j ^ (n & 7)) << 4);
        *reinterpret_cast<uint4 *>(hi + off) = make_uint4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<uint4 *>(lo + off) = make_uint4(l[0], l[1], l[2], l[3]);
    }
}

// one thread: acc[128, n_pad] = A[128, 8*k_steps] * W^T as 3xTF32, A in TMEM (hi/lo column ranges)
__device__ __forceinline__ void issue_layer(uint32_t t_acc, uint32_t t_a_hi, uint32_t t_a_lo, uint32_t s_w_hi,
                                            uint32_t s_w_lo, int n_pad, int k_steps, uint32_t bar) {
    const uint32_t idesc = ptx::umma_idesc_tf32(TM, n_pad);
    for (int s = 0; s < k_steps; ++s) {
        const uint32_t blk = (uint32_t)(s >> 2) * (uint32_t)n_pad * 128u;
        const uint64_t bh = ptx::umma_desc_sw128(s_w_hi + blk) + 2u * (uint32_t)(s & 3);
        const uint64_t bl = ptx::umma_desc_sw128(s_w_lo + blk) + 2u * (uint32_t)(s & 3);
        ptx::umma_tf32_ts(t_acc, t_a_hi + 8u * s, bh, idesc, s ? 1u : 0u);
        ptx::umma_tf32_ts(t_acc, t_a_lo + 8u * s, bh, idesc, 1u);
        ptx::umma_tf32_ts(t_acc, t_a_hi + 8u * s, bl, idesc, 1u);
    }
    ptx::umma_commit(bar);
}

// all threads: A operand is complete in TMEM -> thread 0 issues the layer -> everyone waits for it
__device__ __forceinline__ void run_layer(uint32_t t_acc, uint32_t t_a_hi, uint32_t t_a_lo, uint32_t s_w_hi,
                                          uint32_t s_w_lo, int n_pad, int k_steps, uint32_t bar, uint32_t &phase) {
    ptx::tmem_st_wait();
    ptx::tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) {
        ptx::tc_fence_after();
        issue_layer(t_acc, t_a_hi, t_a_lo, s_w_hi, s_w_lo, n_pad, k_steps, bar);
    }
    ptx::mbar_wait(bar, phase);
    phase ^= 1u;
    ptx::tc_fence_after();
}

// 16 fp32 of this thread's row -> hi/lo columns of the next A operand
__device__ __forceinline__ void store_split16(uint32_t t_hi, uint32_t t_lo, const float (&v)[16]) {
    uint32_t h[16], l[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) split_tf32(v[i], h[i], l[i]);
    ptx::tmem_st_32x32b_x16(t_hi, h);
    ptx::tmem_st_32x32b_x16(t_lo, l);
}

// 16 consecutive floats of one row -> global (vector stores when the row pitch allows)
__device__ __forceinline__ void store_row16(float *__restrict__ dst, const float (&v)[16], int valid, bool vec) {
    if (vec && (valid & 3) == 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (4 * i < valid)
                reinterpret_cast<float4 *>(dst)[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    } else {
#pragma unroll
        for (int i = 0; i < 16; ++i)
            if (i < valid) dst[i] = v[i];
    }
}

// All 32 lanes of a warp: lane l holds v[16] = columns [c0, c0+16) of row (warp_row0 + l).  The block is
// transposed through a warp-private 2 KB shared buffer so each store instruction covers 8 rows x 64
// contiguous bytes (8 memory wavefronts instead of 32 with one row per lane).
// base = &dst[warp_row0 * ld + c0]; rows >= rows_valid are not written.
__device__ __forceinline__ void warp_store16(uint8_t *wbuf, float *__restrict__ base, int ld, const float (&v)[16],
                                             int rows_valid, int valid_cols, bool vec) {
    const int lane = threadIdx.x & 31;
    if (vec && valid_cols == 16) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
            *reinterpret_cast<float4 *>(wbuf + lane * 64 + ((j ^ ((lane >> 1) & 3)) << 4)) =
                make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        __syncwarp();
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int f = it * 32 + lane, row = f >> 2, j = f & 3;
            const float4 q = *reinterpret_cast<const float4 *>(wbuf + row * 64 + ((j ^ ((row >> 1) & 3)) << 4));
            if (row < rows_valid) *reinterpret_cast<float4 *>(base + (size_t)row * ld + j * 4) = q;
        }
        __syncwarp();
    } else if (lane < rows_valid) {
#pragma unroll
        for (int i = 0; i < 16; ++i)
            if (i < valid_cols) base[(size_t)lane * ld + i] = v[i];
    }
}

// Epilogue of one layer for this thread's row (TMEM lane): for each 16-column block owned by this warp
//   v = acc + bias -> [pre-activation to global] -> act -> fn(c0, v) -> [next A operand hi/lo in TMEM]
// `half` (warp / 4) selects the 16-column blocks cb with cb % 2 == half.
template <int ACT, class Fn>
__device__ __forceinline__ void epilogue(uint32_t t_acc, int n_pad, int n_real, const float *sbias,
                                         bool write_next, uint32_t t_next_hi, uint32_t t_next_lo,
                                         float *__restrict__ a_out, int ld_out, bool vec, int half, Fn fn) {
    const int nb = n_pad >> 4;
    for (int cb = half; cb < nb; cb += 2) {
        const int c0 = cb << 4;
        uint32_t r[16];
        ptx::tmem_ld_32x32b_x16(t_acc + c0, r);
        ptx::tmem_ld_wait();
        float v[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]) + sbias[c0 + i];
        const int valid = min(16, n_real - c0);
        if (a_out) store_row16(a_out + c0, v, valid, vec);
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = apply_act<ACT>(v[i]);
        fn(c0, valid, v);
        if (write_next) store_split16(t_next_hi + c0, t_next_lo + c0, v);
    }
}

// Cooperative load of columns [32*kb, 32*kb+32) of the tile's TM rows into a swizzled staging block,
// then every thread moves 16 columns of ITS row (TMEM lane) into the A operand as hi/lo.
//   rowptr(r): pointer to row r's first element, or nullptr for rows past the end (-> zeros)
//   on_chunk(r, k0, v): called by the loading thread with the 4 values it fetched (fused side work)
// k_cols = columns allocated to the operand (multiple of 16).
// Ends with the staging block free for reuse only after the caller's next __syncthreads().
template <class RowPtr, class OnChunk>
__device__ __forceinline__ void stage_block_to_tmem(uint8_t *stage, int kb, int k_real, bool vec, RowPtr rowptr,
                                                    OnChunk on_chunk, uint32_t t_hi, uint32_t t_lo, int k_cols,
                                                    int half) {
    for (int c = threadIdx.x; c < TM * 8; c += NTHR) {
        const int r = c >> 3, j = c & 7, k0 = kb * 32 + j * 4;
        const float *p = rowptr(r);
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (p != nullptr && k0 < k_real) {
            if (vec && k0 + 3 < k_real) {
                v = __ldg(reinterpret_cast<const float4 *>(p + k0));
            } else {
                v.x = __ldg(p + k0);
                if (k0 + 1 < k_real) v.y = __ldg(p + k0 + 1);
                if (k0 + 2 < k_real) v.z = __ldg(p + k0 + 2);
                if (k0 + 3 < k_real) v.w = __ldg(p + k0 + 3);
            }
            on_chunk(r, k0, v);
        }
        *reinterpret_cast<float4 *>(stage + r * 128 + ((j ^ (r & 7)) << 4)) = v;
    }
    __syncthreads();
    if (kb * 32 + half * 16 >= k_cols) return;   // warp-uniform: this 16-column block is past the operand
    const int r = threadIdx.x & 127;   // TMEM lane of this thread: 32*(warp%4) + lane
    float v[16];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int j = half * 4 + i;
        const float4 q = *reinterpret_cast<const float4 *>(stage + r * 128 + ((j ^ (r & 7)) << 4));
        v[4 * i] = q.x; v[4 * i + 1] = q.y; v[4 * i + 2] = q.z; v[4 * i + 3] = q.w;
    }
    store_split16(t_hi + kb * 32 + half * 16, t_lo + kb * 32 + half * 16, v);
}

struct NoChunk { __device__ __forceinline__ void operator()(int, int, const float4 &) const {} };
struct NoFn { __device__ __forceinline__ void operator()(int, int, float (&)[16]) const {} };

// common prologue/epilogue of the three kernels: barrier + TMEM allocation
struct Ctx {
    uint32_t tmem, bar, phase, lane_addr;
    int half;
};
__device__ __forceinline__ Ctx ctx_begin(uint8_t *bar_mem, volatile uint32_t *tmem_slot) {
    Ctx c;
    c.bar = ptx::smem_u32(bar_mem);
    if (threadIdx.x == 0) {
        ptx::mbar_init(c.bar, 1);
        ptx::fence_barrier_init();
    }
    if ((threadIdx.x >> 5) == 0) ptx::tmem_alloc<512>(ptx::smem_u32((const void *)tmem_slot));
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const int warp = threadIdx.x >> 5;
    c.lane_addr = ((uint32_t)((warp & 3) * 32) << 16);
    c.tmem = *tmem_slot + c.lane_addr;     // every TMEM address below carries this warp's lane quarter
    c.phase = 0;
    c.half = warp >> 2;
    return c;
}
__device__ __forceinline__ void ctx_end(const Ctx &c) {
    ptx::tc_fence_before();
    __syncthreads();
    if ((threadIdx.x >> 5) == 0) ptx::tmem_dealloc<512>(c.tmem - c.lane_addr);
}

__host__ __device__ constexpr int pad16(int v) { return (v + 15) / 16 * 16; }
__host__ __device__ constexpr int blocks32(int v) { return (v + 31) / 32; }

// ------------------------------------------------------------------ enc12: x -> h1 -> h2
// smem: W0 hi|lo (64 x 1 block) | W1 hi|lo (128 x 2 blocks) | stage | b0 b1 | barrier, tmem slot
// TMEM columns: x_hi 0 | x_lo 16 | acc1 32 | h1_hi 96 | h1_lo 160 | acc2 224
struct Enc12Smem {
    static constexpr int W0 = 0, W0_IMG = H1 * 128;
    static constexpr int W1 = W0 + 2 * W0_IMG, W1_IMG = HH * 128 * 2;
    static constexpr int STG = W1 + 2 * W1_IMG;
    static constexpr int BIAS = STG + STG_BYTES;
    static constexpr int BAR = BIAS + (H1 + HH) * 4;
    static constexpr int TOTAL = BAR + 16 + 1024;
};

__global__ void __launch_bounds__(NTHR, 1)
enc12_kernel(const float *__restrict__ x, const float *__restrict__ W0, const float *__restrict__ b0,
             const float *__restrict__ W1, const float *__restrict__ b1, int64_t n_rows, int F,
             float *__restrict__ a1, float *__restrict__ h1, float *__restrict__ a2, float *__restrict__ h2) {
    using L = Enc12Smem;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = ptx::smem_u32(smem_raw);
    uint8_t *sm = smem_raw + (((raw + 1023u) & ~1023u) - raw);
    const uint32_t s0 = ptx::smem_u32(sm);
    float *sb0 = reinterpret_cast<float *>(sm + L::BIAS), *sb1 = sb0 + H1;

    stage_weight(sm + L::W0, sm + L::W0 + L::W0_IMG, W0, H1, F, H1, 1);
    stage_weight(sm + L::W1, sm + L::W1 + L::W1_IMG, W1, HH, H1, HH, 2);
    for (int i = threadIdx.x; i < H1; i += NTHR) sb0[i] = b0[i];
    for (int i = threadIdx.x; i < HH; i += NTHR) sb1[i] = b1[i];
    ptx::fence_proxy_async();   // generic-proxy smem writes -> visible to the tensor core (async proxy)
    Ctx c = ctx_begin(sm + L::BAR, reinterpret_cast<volatile uint32_t *>(sm + L::BAR + 8));

    const int64_t n_tiles = (n_rows + TM - 1) / TM;
    const bool vec_x = (F & 3) == 0;
    const int r_in = threadIdx.x & 127;
    const int k0_steps = (F + 7) / 8;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t row0 = tile * TM, row = row0 + r_in;
        const bool row_ok = row < n_rows;
        stage_block_to_tmem(
            sm + L::STG, 0, F, vec_x,
            [&](int r) { return row0 + r < n_rows ? x + (row0 + r) * F : (const float *)nullptr; }, NoChunk(),
            c.tmem + 0, c.tmem + 16, 16, c.half);
        run_layer(c.tmem + 32, c.tmem + 0, c.tmem + 16, s0 + L::W0, s0 + L::W0 + L::W0_IMG, H1, k0_steps, c.bar,
                  c.phase);
        {
            float *h1r = (h1 && row_ok) ? h1 + row * H1 : nullptr;
            epilogue<ACT_GELU>(c.tmem + 32, H1, H1, sb0, true, c.tmem + 96, c.tmem + 160,
                               (a1 && row_ok) ? a1 + row * H1 : nullptr, H1, true, c.half,
                               [&](int c0, int valid, float(&v)[16]) { if (h1r) store_row16(h1r + c0, v, valid, true); });
        }
        run_layer(c.tmem + 224, c.tmem + 96, c.tmem + 160, s0 + L::W1, s0 + L::W1 + L::W1_IMG, HH, H1 / 8, c.bar,
                  c.phase);
        {
            float *h2r = row_ok ? h2 + row * HH : nullptr;
            epilogue<ACT_GELU>(c.tmem + 224, HH, HH, sb1, false, 0u, 0u, (a2 && row_ok) ? a2 + row * HH : nullptr, HH,
                               true, c.half,
                               [&](int c0, int valid, float(&v)[16]) { if (h2r) store_row16(h2r + c0, v, valid, true); });
        }
    }
    ctx_end(c);
}

// ------------------------------------------------------------------ lat: h2 -> z_e = sigmoid(h2 Wn^T + b)
// smem: Wn hi|lo (D16 x 4 blocks) | stage | bias | barrier.   TMEM: h2_hi 0 | h2_lo 128 | acc 256
__host__ __device__ inline int lat_smem_bytes(int D16) { return 2 * D16 * 512 + STG_BYTES + D16 * 4 + 16 + 1024; }

__global__ void __launch_bounds__(NTHR, 1)
lat_kernel(const float *__restrict__ h2, const float *__restrict__ Wn, const float *__restrict__ b, int64_t n_rows,
           int D, float *__restrict__ z_e) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = ptx::smem_u32(smem_raw);
    uint8_t *sm = smem_raw + (((raw + 1023u) & ~1023u) - raw);
    const uint32_t s0 = ptx::smem_u32(sm);
    const int D16 = pad16(D), img = D16 * 512;
    uint8_t *stg = sm + 2 * img;
    float *sb = reinterpret_cast<float *>(stg + STG_BYTES);
    uint8_t *bar = reinterpret_cast<uint8_t *>(sb + D16);

    stage_weight(sm, sm + img, Wn, D, HH, D16, 4);
    for (int i = threadIdx.x; i < D16; i += NTHR) sb[i] = i < D ? b[i] : 0.f;
    ptx::fence_proxy_async();
    Ctx c = ctx_begin(bar, reinterpret_cast<volatile uint32_t *>(bar + 8));

    const int64_t n_tiles = (n_rows + TM - 1) / TM;
    const bool vec_d = (D & 3) == 0;
    const int r_in = threadIdx.x & 127;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t row0 = tile * TM, row = row0 + r_in;
        const bool row_ok = row < n_rows;
        for (int kb = 0; kb < 4; ++kb) {
            stage_block_to_tmem(
                stg, kb, HH, true,
                [&](int r) { return row0 + r < n_rows ? h2 + (row0 + r) * HH : (const float *)nullptr; }, NoChunk(),
                c.tmem + 0, c.tmem + 128, HH, c.half);
            __syncthreads();   // single staging block: everyone has read it before it is refilled
        }
        run_layer(c.tmem + 256, c.tmem + 0, c.tmem + 128, s0, s0 + img, D16, HH / 8, c.bar, c.phase);
        float *zr = row_ok ? z_e + row * D : nullptr;
        epilogue<ACT_SIGMOID>(c.tmem + 256, D16, D, sb, false, 0u, 0u, nullptr, D, vec_d, c.half,
                              [&](int c0, int valid, float(&v)[16]) { if (zr) store_row16(zr + c0, v, valid, vec_d); });
    }
    ctx_end(c);
}

// ------------------------------------------------------------------ host side
inline bool supported(const lipvq_dims_t *d) {
    return d->H1 == H1 && d->H == HH && d->F >= 1 && d->F <= 16 && d->D >= 1 && d->D <= 208;
}

template <class Kern>
inline int set_smem(Kern kern, int bytes) {
    LIPVQ_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    return LIPVQ_OK;
}
inline int grid_for(int64_t n) {
    const int64_t t = (n + TM - 1) / TM;
    return (int)(t < kNumSMs ? t : kNumSMs);
}

// encoder trunk + Lipschitz layer; Wn = normalised weight (lipschitz_prep_kernel); a1/h1/a2 nullable
inline int encoder_launch(const lipvq_dims_t *d, const lipvq_params_t *p, const float *x, int64_t n, float *a1,
                          float *h1, float *a2, float *h2, const float *Wn, float *z_e, cudaStream_t st) {
    LIPVQ_TRY(set_smem(enc12_kernel, Enc12Smem::TOTAL));
    enc12_kernel<<<grid_for(n), NTHR, Enc12Smem::TOTAL, st>>>(x, p->enc0_w, p->enc0_b, p->enc2_w, p->enc2_b, n, d->F,
                                                              a1, h1, a2, h2);
    LIPVQ_LAUNCH_OK();
    const int lb = lat_smem_bytes(pad16(d->D));
    LIPVQ_TRY(set_smem(lat_kernel, lb));
    lat_kernel<<<grid_for(n), NTHR, lb, st>>>(h2, Wn, p->lat_b, n, d->D, z_e);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

}  // namespace mlp
}  // namespace lipvq
