// mlp_dec2.cuh -- decoder after the codeword table trick.
//
// decoder.0 acts on z_q = E[idx] (v5:47, 75), so its pre-activation is a row of the K x 64 table
//   T = E * Wd0^T   (computed once per forward by dec_kernel's table mode, 3xTF32)
// and the per-row work shrinks to  d1 = GELU(T[idx] + bd0)  ->  decoder.2 (64 -> 128, tcgen05 3xTF32, A in
// TMEM)  ->  GELU  ->  to_output (128 -> F <= 16) on the CUDA cores straight from the epilogue registers
// (v5:62-68, 75-76) -> x_recon and sum (x_recon - x)^2 (v5:79).
//
// One persistent CTA per SM holds ONE copy of the weights and runs TWO independent 128-row tiles at a time:
// two "slots" of 8 warps, each with its own 256 TMEM columns, mbarrier and named barrier, so one slot's
// GELU epilogue overlaps the other slot's MMA and global-memory latency.
#pragma once
#include "mlp_tc.cuh"

namespace lipvq {
namespace mlp {

constexpr int D2_THR = 512;          // 2 slots x 8 warps
constexpr int D2_FP = 16;            // to_output columns padded to 16

struct Dec2Smem {
    static constexpr int W1 = 0, W1_IMG = HH * 128 * 2;              // decoder.2 hi | lo images
    static constexpr int WO = W1 + 2 * W1_IMG;                       // to_output^T [128][16] fp32
    static constexpr int BIAS = WO + HH * D2_FP * 4;                 // bd0[64] bd1[128] bo[16]
    static constexpr int COMB = BIAS + (H1 + HH + D2_FP) * 4;        // [2 slots][128 rows][16] partial outputs
    static constexpr int RED = COMB + 2 * TM * D2_FP * 4;
    static constexpr int BAR = RED + 16 * 8;
    static constexpr int TOTAL = BAR + 32 + 1024;
};

__device__ __forceinline__ void slot_sync(int slot) {
    asm volatile("bar.sync %0, 256;" ::"r"(slot + 1) : "memory");
}

__global__ void __launch_bounds__(D2_THR, 1)
dec2_kernel(const int64_t *__restrict__ idx, const float *__restrict__ T, const float *__restrict__ bd0,
            const float *__restrict__ Wd1, const float *__restrict__ bd1, const float *__restrict__ Wo,
            const float *__restrict__ bo, const float *__restrict__ x, int64_t n_rows, int K, int F,
            float *__restrict__ c1, float *__restrict__ d1, float *__restrict__ c2, float *__restrict__ d2,
            float *__restrict__ x_recon, double *__restrict__ part /* [gridDim.x] sse_r partials */) {
    using L = Dec2Smem;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = ptx::smem_u32(smem_raw);
    uint8_t *sm = smem_raw + (((raw + 1023u) & ~1023u) - raw);
    const uint32_t s0 = ptx::smem_u32(sm);
    float *sWo = reinterpret_cast<float *>(sm + L::WO);
    float *sb0 = reinterpret_cast<float *>(sm + L::BIAS), *sb1 = sb0 + H1, *sbo = sb1 + HH;
    double *sred = reinterpret_cast<double *>(sm + L::RED);
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(sm + L::BAR + 16);

    stage_weight(sm + L::W1, sm + L::W1 + L::W1_IMG, Wd1, HH, H1, HH, 2);
    for (int i = threadIdx.x; i < HH * D2_FP; i += D2_THR) {
        const int cc = i / D2_FP, f = i % D2_FP;
        sWo[i] = f < F ? Wo[(size_t)f * HH + cc] : 0.f;
    }
    for (int i = threadIdx.x; i < H1; i += D2_THR) sb0[i] = bd0[i];
    for (int i = threadIdx.x; i < HH; i += D2_THR) sb1[i] = bd1[i];
    for (int i = threadIdx.x; i < D2_FP; i += D2_THR) sbo[i] = i < F ? bo[i] : 0.f;
    ptx::fence_proxy_async();
    if (threadIdx.x == 0) {
        ptx::mbar_init(ptx::smem_u32(sm + L::BAR), 1);
        ptx::mbar_init(ptx::smem_u32(sm + L::BAR + 8), 1);
        ptx::fence_barrier_init();
    }
    if ((threadIdx.x >> 5) == 0) ptx::tmem_alloc<512>(ptx::smem_u32((const void *)tmem_slot));
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();

    const int slot = threadIdx.x >> 8, ltid = threadIdx.x & 255;
    const int warp = ltid >> 5, half = warp >> 2, r_in = ltid & 127;
    const uint32_t tm = *tmem_slot + ((uint32_t)((warp & 3) * 32) << 16) + 256u * slot;
    const uint32_t t_d1hi = tm, t_d1lo = tm + 64, t_acc = tm + 128;
    const uint32_t bar = ptx::smem_u32(sm + L::BAR + 8 * slot);
    float *comb = reinterpret_cast<float *>(sm + L::COMB) + (size_t)slot * TM * D2_FP + r_in * D2_FP;
    uint32_t phase = 0;
    const int64_t n_tiles = (n_rows + TM - 1) / TM;
    const bool vec_f = (F & 3) == 0;
    double sse_r = 0.0;

    float4 q[8];
    auto fetch = [&](int64_t tile) {
        const int64_t row = tile * TM + r_in;
        int64_t k = (tile < n_tiles && row < n_rows) ? idx[row] : 0;
        if (k < 0 || k >= K) k = 0;
        const float4 *src = reinterpret_cast<const float4 *>(T + k * H1 + half * 32);
#pragma unroll
        for (int i = 0; i < 8; ++i) q[i] = __ldg(src + i);
    };
    int64_t tile = (int64_t)blockIdx.x * 2 + slot;
    const int64_t tstep = (int64_t)gridDim.x * 2;
    if (tile < n_tiles) fetch(tile);
    for (; tile < n_tiles; tile += tstep) {
        const int64_t row = tile * TM + r_in;
        const bool row_ok = row < n_rows;
        // ---- layer 0 from the table: d1 = GELU(T[idx] + bd0) -> A operand (hi/lo) in TMEM
#pragma unroll
        for (int g = 0; g < 2; ++g) {
            const int c0 = half * 32 + g * 16;
            float v[16];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float4 t = q[g * 4 + i];
                v[4 * i] = t.x + sb0[c0 + 4 * i]; v[4 * i + 1] = t.y + sb0[c0 + 4 * i + 1];
                v[4 * i + 2] = t.z + sb0[c0 + 4 * i + 2]; v[4 * i + 3] = t.w + sb0[c0 + 4 * i + 3];
            }
            if (c1 && row_ok) store_row16(c1 + row * H1 + c0, v, 16, true);
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = gelu_f(v[i]);
            if (d1 && row_ok) store_row16(d1 + row * H1 + c0, v, 16, true);
            store_split16(t_d1hi + c0, t_d1lo + c0, v);
        }
        ptx::tmem_st_wait();
        ptx::tc_fence_before();
        slot_sync(slot);
        if (ltid == 0) {
            ptx::tc_fence_after();
            issue_layer(t_acc, t_d1hi, t_d1lo, s0 + L::W1, s0 + L::W1 + L::W1_IMG, HH, H1 / 8, bar);
        }
        fetch(tile + tstep);                        // next tile's table rows fly while the MMA runs
        ptx::mbar_wait(bar, phase);
        phase ^= 1u;
        ptx::tc_fence_after();
        // ---- layer 1 epilogue + to_output on the CUDA cores
        float out[D2_FP];
#pragma unroll
        for (int f = 0; f < D2_FP; ++f) out[f] = 0.f;
        for (int cb = half; cb < HH / 16; cb += 2) {
            const int c0 = cb << 4;
            uint32_t rr[16];
            ptx::tmem_ld_32x32b_x16(t_acc + c0, rr);
            ptx::tmem_ld_wait();
            float v[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(rr[i]) + sb1[c0 + i];
            if (c2 && row_ok) store_row16(c2 + row * HH + c0, v, 16, true);
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = gelu_f(v[i]);
            if (d2 && row_ok) store_row16(d2 + row * HH + c0, v, 16, true);
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float4 *w = reinterpret_cast<const float4 *>(sWo + (c0 + i) * D2_FP);
#pragma unroll
                for (int f4 = 0; f4 < D2_FP / 4; ++f4) {
                    const float4 ww = w[f4];
                    out[4 * f4] = fmaf(v[i], ww.x, out[4 * f4]);
                    out[4 * f4 + 1] = fmaf(v[i], ww.y, out[4 * f4 + 1]);
                    out[4 * f4 + 2] = fmaf(v[i], ww.z, out[4 * f4 + 2]);
                    out[4 * f4 + 3] = fmaf(v[i], ww.w, out[4 * f4 + 3]);
                }
            }
        }
        if (half == 1) {
#pragma unroll
            for (int f4 = 0; f4 < D2_FP / 4; ++f4)
                reinterpret_cast<float4 *>(comb)[f4] = make_float4(out[4 * f4], out[4 * f4 + 1], out[4 * f4 + 2], out[4 * f4 + 3]);
        }
        ptx::tc_fence_before();
        slot_sync(slot);
        if (half == 0 && row_ok) {
            float s = 0.f;
#pragma unroll
            for (int f = 0; f < D2_FP; ++f) {
                out[f] = (out[f] + comb[f]) + sbo[f];
                if (x && f < F) { const float df = out[f] - __ldg(x + row * F + f); s = fmaf(df, df, s); }
            }
            if (x_recon) store_row16(x_recon + row * F, out, F, vec_f);
            sse_r += (double)s;
        }
    }
    // deterministic CTA reduction (fixed shuffle tree, warps combined in order)
    sse_r = warp_sum_d(sse_r);
    if ((threadIdx.x & 31) == 0) sred[threadIdx.x >> 5] = sse_r;
    ptx::tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0;
        for (int w = 0; w < D2_THR / 32; ++w) a += sred[w];
        part[blockIdx.x] = a;
    }
    if ((threadIdx.x >> 5) == 0) ptx::tmem_dealloc<512>(*tmem_slot);
}

// out[0] = sum_i part[i] in index order
__global__ void sum_partials1_kernel(const double *__restrict__ part, int n, double *__restrict__ out) {
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int i = 0; i < n; ++i) s += part[i];
        out[0] = s;
    }
}

// scratch: [kPartBytes of doubles | K*64 floats of table]
inline size_t dec2_ws_bytes(int K) { return kPartBytes + 256 + (size_t)K * H1 * sizeof(float) + 256; }

// T = E Wd0^T (table mode of dec_kernel), then the per-row decoder; sse_r = device double[1]
inline int decoder2_launch(const lipvq_dims_t *d, const lipvq_params_t *p, const int64_t *idx, const float *x,
                           int64_t n, float *c1, float *d1, float *c2, float *d2, float *x_recon, void *ws,
                           double *sse_r, cudaStream_t st) {
    double *part = (double *)ws;
    float *T = (float *)((char *)ws + align_up(kPartBytes, 256));
    const DecLayout L = dec_layout(d->D);
    LIPVQ_TRY(set_smem(dec_kernel, L.total));
    dec_kernel<<<grid_for(d->K), NTHR, L.total, st>>>(nullptr, p->codebook, nullptr, nullptr, p->dec0_w, p->dec0_b,
                                                      p->dec2_w, p->dec2_b, p->out_w, p->out_b, d->K, d->D, d->K,
                                                      d->F, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
                                                      part, T);
    LIPVQ_LAUNCH_OK();
    LIPVQ_TRY(set_smem(dec2_kernel, Dec2Smem::TOTAL));
    const int64_t pairs = ((n + TM - 1) / TM + 1) / 2;
    const int grid = (int)(pairs < kNumSMs ? pairs : kNumSMs);
    dec2_kernel<<<grid, D2_THR, Dec2Smem::TOTAL, st>>>(idx, T, p->dec0_b, p->dec2_w, p->dec2_b, p->out_w, p->out_b, x,
                                                       n, d->K, d->F, c1, d1, c2, d2, x_recon, part);
    LIPVQ_LAUNCH_OK();
    sum_partials1_kernel<<<1, 32, 0, st>>>(part, grid, sse_r);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

}  // namespace mlp
}  // namespace lipvq
