// misc_kernels.cuh -- HBM-bound passes of the tokenizer: Lipschitz weight normalisation,
// codeword gather + commitment SSE, reconstruction SSE, scatter-add, elementwise backward
// pieces, loss finalisation, AdamW and the EMA extension.
#pragma once
#include "common.cuh"

namespace lipvq {

// ---------------------------------------------------------------- Lipschitz normalisation
// scale_i = min(1, softplus(ci_i) / sum_j |W_ij|);  Wn = W * scale   (v5:6-12)
// one warp per row of W[D,H]
__global__ void lipschitz_prep_kernel(const float *__restrict__ W, const float *__restrict__ ci,
                                      int D, int H, float *__restrict__ scale,
                                      float *__restrict__ Wn) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= D) return;   // (whole warps leave together)
    float s = 0.f;
    for (int j = lane; j < H; j += 32) s += fabsf(W[(size_t)row * H + j]);
    s = warp_sum(s);
    const float sc = fminf(1.0f, softplus_f(ci[row]) / s);
    if (lane == 0) scale[row] = sc;
    for (int j = lane; j < H; j += 32) Wn[(size_t)row * H + j] = W[(size_t)row * H + j] * sc;
}

// Backward through Wn = W * min(1, softplus(ci)/sum|W|)  (autograd of v5:6-12):
//   dot_i = sum_j gWn_ij W_ij;  on the ratio branch (ratio<1; half weight on equality):
//   gW_ij = gWn_ij*scale_i - dot_i*softplus(ci_i)/r_i^2 * sign(W_ij);  gci_i = dot_i/r_i*sigmoid(ci_i)
__global__ void lipschitz_bwd_kernel(const float *__restrict__ gWn, const float *__restrict__ W,
                                     const float *__restrict__ ci, const float *__restrict__ scale,
                                     int D, int H, float *__restrict__ gW, float *__restrict__ gci) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= D) return;
    float r = 0.f, dot = 0.f;
    for (int j = lane; j < H; j += 32) {
        const float w = W[(size_t)row * H + j];
        r += fabsf(w);
        dot = fmaf(gWn[(size_t)row * H + j], w, dot);
    }
    r = warp_sum(r);
    dot = warp_sum(dot);
    const float sp = softplus_f(ci[row]);
    const float ratio = sp / r;
    const float wr = ratio < 1.0f ? 1.0f : (ratio == 1.0f ? 0.5f : 0.0f);
    const float g_ratio = dot * wr;
    const float coef = g_ratio * (-sp / (r * r));
    const float sc = scale[row];
    for (int j = lane; j < H; j += 32) {
        const float w = W[(size_t)row * H + j];
        const float sg = (w > 0.f) ? 1.f : ((w < 0.f) ? -1.f : 0.f);
        gW[(size_t)row * H + j] = gWn[(size_t)row * H + j] * sc + coef * sg;
    }
    if (lane == 0) gci[row] = g_ratio / r * sigmoid_f(ci[row]);
}

// ---------------------------------------------------------------- deterministic SSE reduction
// stage 2: sum per-block partials in index order, in double.
__global__ void partials_to_double_kernel(const float *__restrict__ part, int n, double *__restrict__ out) {
    __shared__ double sm[32];
    double s = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) s += (double)part[i];
    s = warp_sum_d(s);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        double r = threadIdx.x < (blockDim.x >> 5) ? sm[threadIdx.x] : 0.0;
        r = warp_sum_d(r);
        if (threadIdx.x == 0) *out = r;
    }
}

constexpr int kRedBlocks = 4 * kNumSMs;  // grid of the streaming passes (4 CTAs per SM)

// partial[b] = sum (a - b)^2 over a flat range
__global__ void __launch_bounds__(256) sse_kernel(const float *__restrict__ a, const float *__restrict__ b,
                                                  int64_t n, float *__restrict__ partial) {
    __shared__ float red[8];
    float s = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        const float d = a[i] - b[i];
        s += d * d;
    }
    s = block_sum<256>(s, red);
    if (threadIdx.x == 0) partial[blockIdx.x] = s;
}

// ---------------------------------------------------------------- fused quantise + emit pass (k4/k5/k7)
// One streaming pass over the rows (v5:47, 74-81): z_q[n] = E[idx[n]] (this is z_latent), sum (z_q-z_e)^2,
// x_recon[n] = R[idx[n]] with R = decoder(E) the per-code reconstruction table, sum (x_recon-x)^2.
// A CTA walks tiles of QE_ROWS rows: the tile's codes go to shared memory first, so the E / z_e loads of a
// thread are independent (4 float4 pairs in flight); z_e / z_q are streamed past L1 (ld.cs / st.cs) while the
// codebook and R stay cached.  partial: [2][gridDim.x] floats = {sse_q, sse_r} per CTA (summed in order later).
constexpr int QE_ROWS = 64, QE_THR = 256;
constexpr int kQeBlocks = 8 * kNumSMs;   // 8 resident CTAs of 256 threads per SM
// GPRE (training forward): also writes gpre = a_q (z_e - z_q) z_e (1 - z_e), the commitment gradient through the
// sigmoid up to the upstream dL/dloss factor (v5:80; applied by the backward GEMMs' epilogues) -- the row is in
// registers here, so the backward needs no separate pass over z_e and z_q.
template <bool VEC4, bool GPRE = false>
__global__ void __launch_bounds__(QE_THR) quantize_emit_kernel(
    const float *__restrict__ z_e, const float *__restrict__ E, const float *__restrict__ R,
    const int64_t *__restrict__ idx, const float *__restrict__ x, int64_t n_rows, int D, int F, int K,
    float *__restrict__ z_q, float *__restrict__ x_recon, float *__restrict__ partial,
    float *__restrict__ gpre = nullptr, float a_q = 0.f) {
    __shared__ int sidx[QE_ROWS];
    __shared__ float red[8];
    const int tid = threadIdx.x;
    const unsigned dq = VEC4 ? (unsigned)D >> 2 : (unsigned)D;
    // i / dq == umulhi(i, magic) for i < 2^32/dq; a divisor of 1 would need magic = 2^32 (truncates to 0), so it is
    // special-cased at the two call sites (latent_dim 1 or 4, feature_dim 1)
    const unsigned magic = (unsigned)((0x100000000ull + dq - 1) / dq);
    const unsigned fmagic = (unsigned)((0x100000000ull + (unsigned)F - 1) / (unsigned)F);
    auto div_dq = [&](unsigned i) { return dq == 1u ? i : __umulhi(i, magic); };
    auto div_f = [&](unsigned i) { return F == 1 ? i : __umulhi(i, fmagic); };
    float sq = 0.f, sr = 0.f;
    const int64_t n_tiles = (n_rows + QE_ROWS - 1) / QE_ROWS;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t r0 = tile * QE_ROWS;
        const int rows = (int)(n_rows - r0 < QE_ROWS ? n_rows - r0 : QE_ROWS);
        __syncthreads();
        if (tid < rows) {
            const int64_t k = idx[r0 + tid];
            sidx[tid] = (k >= 0 && k < K) ? (int)k : 0;
        }
        __syncthreads();
        const unsigned total = (unsigned)rows * dq;
        if (VEC4) {
            const float4 *zt = reinterpret_cast<const float4 *>(z_e + r0 * D);
            float4 *qt = z_q ? reinterpret_cast<float4 *>(z_q + r0 * D) : nullptr;
            for (unsigned i0 = tid; i0 < total; i0 += 4 * QE_THR) {
                float4 e[4], z[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const unsigned i = i0 + u * QE_THR;
                    if (i < total) {
                        const unsigned r = div_dq(i), c = i - r * dq;
                        e[u] = __ldg(reinterpret_cast<const float4 *>(E + (size_t)sidx[r] * D) + c);
                        z[u] = __ldcs(zt + i);
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const unsigned i = i0 + u * QE_THR;
                    if (i < total) {
                        if (z_q) __stcs(qt + i, e[u]);
                        const float dx = e[u].x - z[u].x, dy = e[u].y - z[u].y, dz = e[u].z - z[u].z,
                                    dw = e[u].w - z[u].w;
                        sq += dx * dx + dy * dy + dz * dz + dw * dw;
                        if (GPRE) {
                            const float4 zz = z[u];
                            __stcs(reinterpret_cast<float4 *>(gpre + r0 * D) + i,
                                   make_float4(-a_q * dx * zz.x * (1.0f - zz.x), -a_q * dy * zz.y * (1.0f - zz.y),
                                               -a_q * dz * zz.z * (1.0f - zz.z), -a_q * dw * zz.w * (1.0f - zz.w)));
                        }
                    }
                }
            }
        } else {
            for (unsigned i = tid; i < total; i += QE_THR) {
                const unsigned r = div_dq(i), c = i - r * dq;
                const float e = __ldg(E + (size_t)sidx[r] * D + c);
                const float zz = __ldcs(z_e + r0 * D + i);
                const float dd = e - zz;
                if (z_q) __stcs(z_q + r0 * D + i, e);
                sq += dd * dd;
                if (GPRE) __stcs(gpre + r0 * D + i, -a_q * dd * zz * (1.0f - zz));
            }
        }
        if (R) {
            const unsigned tf = (unsigned)rows * (unsigned)F;
            for (unsigned i = tid; i < tf; i += QE_THR) {
                const unsigned r = div_f(i), f = i - r * (unsigned)F;
                const float v = __ldg(R + (size_t)sidx[r] * F + f);
                if (x_recon) x_recon[r0 * F + i] = v;
                if (x) { const float dd = v - __ldg(x + r0 * F + i); sr = fmaf(dd, dd, sr); }
            }
        }
    }
    sq = block_sum<QE_THR>(sq, red);
    sr = block_sum<QE_THR>(sr, red);
    if (tid == 0) { partial[blockIdx.x] = sq; partial[gridDim.x + blockIdx.x] = sr; }
}

// sse[0] = sum_i part[n + i] (recon), sse[1] = sum_i part[i] (quantisation); fixed order, double.
// loss_out (nullable): the four loss terms of loss_finalize_kernel, written by the same thread.
__global__ void partials2_to_double_kernel(const float *__restrict__ part, int n, double *__restrict__ sse,
                                           float *__restrict__ loss_out, double inv_nf, double inv_nd) {
    __shared__ double sm[2][8];
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) { a += (double)part[i]; b += (double)part[n + i]; }
    a = warp_sum_d(a); b = warp_sum_d(b);
    if ((threadIdx.x & 31) == 0) { sm[0][threadIdx.x >> 5] = a; sm[1][threadIdx.x >> 5] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double q = 0.0, r = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { q += sm[0][w]; r += sm[1][w]; }
        sse[0] = r; sse[1] = q;
        if (loss_out) {
            const double recon = r * inv_nf, commit = q * inv_nd;
            loss_out[0] = (float)(recon + 0.25 * commit + 0.25 * commit);
            loss_out[1] = (float)recon;
            loss_out[2] = (float)commit;
            loss_out[3] = (float)commit;
        }
    }
}

// the same for a forward that ran its emit pass in up to 8 row chunks (each chunk's partials: [2][grid[c]] floats at
// part + c * stride): chunk by chunk, fixed order inside a chunk, in double.
struct ChunkGrids { int n; int grid[8]; };
__global__ void partials2_chunks_to_double_kernel(const float *__restrict__ part, ChunkGrids cg, int stride,
                                                  double *__restrict__ sse, float *__restrict__ loss_out, double inv_nf,
                                                  double inv_nd) {
    __shared__ double sm[2][8];
    double q = 0.0, r = 0.0;
    for (int c = 0; c < cg.n; ++c) {
        const float *pc = part + (size_t)c * stride;
        const int n = cg.grid[c];
        double a = 0.0, b = 0.0;
        for (int i = threadIdx.x; i < n; i += blockDim.x) { a += (double)pc[i]; b += (double)pc[n + i]; }
        a = warp_sum_d(a); b = warp_sum_d(b);
        __syncthreads();
        if ((threadIdx.x & 31) == 0) { sm[0][threadIdx.x >> 5] = a; sm[1][threadIdx.x >> 5] = b; }
        __syncthreads();
        if (threadIdx.x == 0)
            for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { q += sm[0][w]; r += sm[1][w]; }
    }
    if (threadIdx.x == 0) {
        sse[0] = r; sse[1] = q;
        if (loss_out) {
            const double recon = r * inv_nf, commit = q * inv_nd;
            loss_out[0] = (float)(recon + 0.25 * commit + 0.25 * commit);
            loss_out[1] = (float)recon;
            loss_out[2] = (float)commit;
            loss_out[3] = (float)commit;
        }
    }
}

// search counters between two row chunks of one forward: the list lengths [0] (exact scan) and [4] (rescore) restart,
// their totals accumulate in [5] and [7]; [1] codes rescored, [2] rows resolved directly and [3] the range flag carry on
__global__ void vq_counters_next_chunk_kernel(int *__restrict__ counters) {
    if (threadIdx.x == 0) {
        counters[5] += counters[0]; counters[7] += counters[4];
        counters[0] = 0; counters[4] = 0;
    }
}
__global__ void vq_counters_report_kernel(const int *__restrict__ counters, int *__restrict__ stats) {
    if (threadIdx.x == 0) { stats[0] = counters[5]; stats[1] = counters[1]; stats[2] = counters[2]; stats[3] = counters[3]; }
}

// ---------------------------------------------------------------- per-code decoder table (k7)
// xr_code = to_output(decoder(E)) for all K codes (v5:62-68, 75-76 on the K distinct values of z_q).  K rows are
// too few for the tensor cores to pay off (a 128-row UMMA tile would keep 8 SMs busy for 40 us), so this runs on
// the CUDA cores in exact fp32: a CTA takes CT_CODES codes through the three layers; a warp owns an output
// neuron at a time, its lanes stride the reduction axis (coalesced weight rows), one accumulator per code.
constexpr int CT_CODES = 8;
template <int ACT, int KIN_MAX_UNUSED = 0>
__device__ __forceinline__ void ct_layer(const float *__restrict__ W, const float *__restrict__ bias, int n_out,
                                         int k_in, const float *s_in /* [CT][k_in] */, float *s_out /* [CT][n_out] */,
                                         float *__restrict__ g_pre, float *__restrict__ g_post, int ld, int code0,
                                         int n_codes) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = blockDim.x >> 5;
    constexpr int NJ = 4;   // output neurons in flight per warp: NJ independent weight loads per step hide the L2 latency
    for (int j0 = warp * NJ; j0 < n_out; j0 += nwarp * NJ) {
        float acc[NJ][CT_CODES];
#pragma unroll
        for (int q = 0; q < NJ; ++q)
#pragma unroll
            for (int c = 0; c < CT_CODES; ++c) acc[q][c] = 0.f;
#pragma unroll 2
        for (int k = lane; k < k_in; k += 32) {
            float w[NJ];
#pragma unroll
            for (int q = 0; q < NJ; ++q) w[q] = j0 + q < n_out ? __ldg(W + (size_t)(j0 + q) * k_in + k) : 0.f;
#pragma unroll
            for (int c = 0; c < CT_CODES; ++c) {
                const float e = s_in[c * k_in + k];
#pragma unroll
                for (int q = 0; q < NJ; ++q) acc[q][c] = fmaf(w[q], e, acc[q][c]);
            }
        }
#pragma unroll
        for (int q = 0; q < NJ; ++q) {
            const int j = j0 + q;
            if (j >= n_out) break;          // warp-uniform
#pragma unroll
            for (int c = 0; c < CT_CODES; ++c) acc[q][c] = warp_sum(acc[q][c]);
            if (lane < CT_CODES) {
                float v = 0.f;
#pragma unroll
                for (int c = 0; c < CT_CODES; ++c) v = lane == c ? acc[q][c] : v;
                v += __ldg(bias + j);
                const bool ok = lane < n_codes;
                if (g_pre && ok) g_pre[(size_t)(code0 + lane) * ld + j] = v;
                v = apply_act<ACT>(v);
                if (g_post && ok) g_post[(size_t)(code0 + lane) * ld + j] = v;
                if (s_out) s_out[lane * n_out + j] = v;
            }
        }
    }
}
__global__ void __launch_bounds__(512) decoder_code_table_kernel(
    const float *__restrict__ E, const float *__restrict__ Wd0, const float *__restrict__ bd0,
    const float *__restrict__ Wd1, const float *__restrict__ bd1, const float *__restrict__ Wo,
    const float *__restrict__ bo, int K, int D, int H1, int H, int F, float *__restrict__ c1,
    float *__restrict__ d1, float *__restrict__ c2, float *__restrict__ d2, float *__restrict__ R) {
    extern __shared__ float ct_smem[];
    float *s_e = ct_smem, *s_a = s_e + CT_CODES * D, *s_b = s_a + CT_CODES * H1;
    const int code0 = blockIdx.x * CT_CODES;
    const int n_codes = K - code0 < CT_CODES ? K - code0 : CT_CODES;
    for (int i = threadIdx.x; i < CT_CODES * D; i += blockDim.x) {
        const int c = i / D, k = i - c * D;
        s_e[i] = c < n_codes ? __ldg(E + (size_t)(code0 + c) * D + k) : 0.f;
    }
    __syncthreads();
    ct_layer<ACT_GELU>(Wd0, bd0, H1, D, s_e, s_a, c1, d1, H1, code0, n_codes);
    __syncthreads();
    ct_layer<ACT_GELU>(Wd1, bd1, H, H1, s_a, s_b, c2, d2, H, code0, n_codes);
    __syncthreads();
    ct_layer<ACT_NONE>(Wo, bo, F, H, s_b, nullptr, nullptr, R, F, code0, n_codes);
}

// out = alpha * (a - b)
__global__ void scaled_diff_kernel(const float *__restrict__ a, const float *__restrict__ b,
                                   float alpha, const float *__restrict__ gs, int64_t n,
                                   float *__restrict__ out) {
    if (gs) alpha *= *gs;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x)
        out[i] = alpha * (a[i] - b[i]);
}

// g_pre = alpha * (z_e - z_q) * z_e * (1 - z_e)   (commitment gradient through the sigmoid)
__global__ void commit_sigmoid_bwd_kernel(const float *__restrict__ z_e, const float *__restrict__ z_q,
                                          float alpha, const float *__restrict__ gs, int64_t n,
                                   float *__restrict__ out) {
    if (gs) alpha *= *gs;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        const float ze = z_e[i];
        out[i] = alpha * (ze - z_q[i]) * ze * (1.0f - ze);
    }
}

// classic VQVAE (vq_vae/backbone.py): gradient of mse(x_recon, x) through the ReLU on the decoder output
//   out = gs * alpha * (x_recon - x) * [x_recon > 0]
__global__ void recon_grad_relu_kernel(const float *__restrict__ xr, const float *__restrict__ x, float alpha,
                                       const float *__restrict__ gs, int64_t n, float *__restrict__ out) {
    if (gs) alpha *= *gs;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float v = xr[i];
        out[i] = v > 0.f ? alpha * (v - x[i]) : 0.f;
    }
}
// straight-through estimator + commitment term + the encoder's output ReLU (backbone.py:69, 74):
//   g = (g_zq + gs * alpha * (z_e - z_q)) * [z_e > 0], in place over g_zq
__global__ void ste_commit_relu_kernel(float *__restrict__ g, const float *__restrict__ z_e,
                                       const float *__restrict__ z_q, float alpha, const float *__restrict__ gs,
                                       int64_t n) {
    if (gs) alpha *= *gs;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float ze = z_e[i];
        g[i] = ze > 0.f ? g[i] + alpha * (ze - z_q[i]) : 0.f;
    }
}
// loss = mse(x_recon, x) + (1 + cc) * mse(z_q, z_e)  (backbone.py:49-50, 69-71); out = {loss, recon, cc*mse_q, mse_q}
__global__ void vqvae_loss_kernel(const double *sse_r, const double *sse_q, double inv_nf, double inv_nd, float cc,
                                  float *out) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        const double recon = (*sse_r) * inv_nf, q = (*sse_q) * inv_nd;
        out[0] = (float)(recon + q + (double)cc * q);
        out[1] = (float)recon;
        out[2] = (float)((double)cc * q);
        out[3] = (float)q;
    }
}

__global__ void loss_finalize_kernel(const double *sse_r, const double *sse_q, double inv_nf,
                                     double inv_nd, float *out) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        const double recon = sse_r ? (*sse_r) * inv_nf : 0.0;
        const double commit = (*sse_q) * inv_nd;
        out[0] = (float)(recon + 0.25 * commit + 0.25 * commit);
        out[1] = (float)recon;
        out[2] = (float)commit;
        out[3] = (float)commit;
    }
}

// ---------------------------------------------------------------- scatter-add (k8)
// One CTA takes 256 consecutive rows, sorts their (code,row) keys in shared memory (bitonic),
// and each warp then walks 32 sorted entries: runs of equal code are summed in registers and
// flushed with ONE atomicAdd per (run, column) -- so a collapsed codebook (every row on one code,
// SURVEY 0.9) costs gridDim atomics per address instead of n_rows.
// DIFF = true accumulates alpha * gs * (T[code,:] - src[row,:]) instead of src[row,:] (T = a per-code table):
// the codebook-loss gradient sum 0.5 (z_q - z_e)/(N D) with T = E, and the per-code reconstruction gradient
// sum 2 (x_recon - x)/(N F) with T = decoder(E) -- neither difference tensor is ever materialised.
constexpr int SC_ROWS = 256;
template <bool DIFF>
__global__ void __launch_bounds__(256) scatter_add_kernel(
    const float *__restrict__ src, const int64_t *__restrict__ idx, int64_t n_rows, int D, int K,
    float *__restrict__ sums, float *__restrict__ counts, const float *__restrict__ T, float alpha,
    const float *__restrict__ gs) {
    if (DIFF && gs) alpha *= __ldg(gs);
    __shared__ unsigned long long keys[SC_ROWS];
    const int tid = threadIdx.x;
    const int64_t r0 = (int64_t)blockIdx.x * SC_ROWS;
    {
        const int64_t r = r0 + tid;
        unsigned long long key = ~0ull;   // padding sorts last
        if (r < n_rows) {
            int64_t k = idx[r];
            if (k < 0 || k >= K) k = 0;
            key = ((unsigned long long)(unsigned)k << 32) | (unsigned)tid;
        }
        keys[tid] = key;
    }
    __syncthreads();
    for (int k = 2; k <= SC_ROWS; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            const int p = tid ^ j;
            if (p > tid) {
                const unsigned long long a = keys[tid], b = keys[p];
                const bool up = (tid & k) == 0;
                if ((a > b) == up) { keys[tid] = b; keys[p] = a; }
            }
            __syncthreads();
        }
    const int warp = tid >> 5, lane = tid & 31;
    const int e0 = warp * 32;
    // counts: lane l owns entry e0+l; a run head adds the run length inside this warp's span
    if (counts) {
        const unsigned long long key = keys[e0 + lane];
        const unsigned code = (unsigned)(key >> 32);
        const bool valid = key != ~0ull;
        const unsigned prev = __shfl_up_sync(0xffffffffu, code, 1);
        const bool head = valid && (lane == 0 || prev != code);
        const unsigned m = __match_any_sync(0xffffffffu, valid ? code : 0xffffffffu);
        if (head) atomicAdd(counts + code, (float)__popc(m));
    }
    if ((D & 3) == 0 && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(sums) |
                          (DIFF ? reinterpret_cast<uintptr_t>(T) : 0)) & 15) == 0) {
        // vector path: a lane owns 4 consecutive columns; 4 rows' loads are in flight before the run logic touches
        // them; a run is flushed with one 16-byte vector atomic per lane
        const int D4 = D >> 2;
        for (int c = lane; c < D4; c += 32) {
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f), t = acc;
            unsigned cur = 0xffffffffu;
            auto flush = [&]() {
                const float s = DIFF ? alpha : 1.0f;
                atomicAdd(reinterpret_cast<float4 *>(sums + (size_t)cur * D) + c,
                          make_float4(s * acc.x, s * acc.y, s * acc.z, s * acc.w));
            };
            for (int e = 0; e < 32; e += 4) {
                unsigned long long k4[4];
                float4 v[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    k4[q] = keys[e0 + e + q];
                    v[q] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (k4[q] != ~0ull)
                        v[q] = __ldg(reinterpret_cast<const float4 *>(src + (r0 + (unsigned)(k4[q] & 0xffffffffu)) * D) + c);
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    if (k4[q] == ~0ull) continue;          // (padding sorts last)
                    const unsigned code = (unsigned)(k4[q] >> 32);
                    if (code != cur) {
                        if (cur != 0xffffffffu) flush();
                        cur = code;
                        acc = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (DIFF) t = __ldg(reinterpret_cast<const float4 *>(T + (size_t)code * D) + c);
                    }
                    if (DIFF) { acc.x += t.x - v[q].x; acc.y += t.y - v[q].y; acc.z += t.z - v[q].z; acc.w += t.w - v[q].w; }
                    else { acc.x += v[q].x; acc.y += v[q].y; acc.z += v[q].z; acc.w += v[q].w; }
                }
            }
            if (cur != 0xffffffffu) flush();
        }
        return;
    }
    for (int c = lane; c < D; c += 32) {
        float acc = 0.f, t = 0.f;
        unsigned cur = 0xffffffffu;
        for (int e = 0; e < 32; ++e) {
            const unsigned long long key = keys[e0 + e];
            if (key == ~0ull) break;
            const unsigned code = (unsigned)(key >> 32);
            const int64_t row = r0 + (unsigned)(key & 0xffffffffu);
            if (code != cur) {
                if (cur != 0xffffffffu) atomicAdd(sums + (size_t)cur * D + c, DIFF ? alpha * acc : acc);
                cur = code;
                acc = 0.f;
                if (DIFF) t = __ldg(T + (size_t)code * D + c);
            }
            const float v = __ldg(src + row * D + c);
            acc += DIFF ? t - v : v;
        }
        if (cur != 0xffffffffu) atomicAdd(sums + (size_t)cur * D + c, DIFF ? alpha * acc : acc);
    }
}

// ---------------------------------------------------------------- AdamW (icl.py:886-889)
__global__ void adamw_kernel(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m,
                             float *__restrict__ v, int64_t n, float lr, float b1, float b2, float eps,
                             float wd, float bc1, float bc2_sqrt) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        const float gi = g[i];
        float w = p[i] * (1.0f - lr * wd);
        const float mi = b1 * m[i] + (1.0f - b1) * gi;
        const float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
        m[i] = mi;
        v[i] = vi;
        const float denom = sqrtf(vi) / bc2_sqrt + eps;
        p[i] = w - (lr / bc1) * (mi / denom);
    }
}

// ---------------------------------------------------------------- EMA extension (unpinned)
__global__ void ema_counts_kernel(float *__restrict__ n_ema, const float *__restrict__ counts, int K,
                                  float gamma, float *__restrict__ partial) {
    __shared__ float red[8];
    float s = 0.f;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < K; i += gridDim.x * blockDim.x) {
        const float v = gamma * n_ema[i] + (1.0f - gamma) * counts[i];
        n_ema[i] = v;
        s += v;
    }
    s = block_sum<256>(s, red);
    if (threadIdx.x == 0) partial[blockIdx.x] = s;
}
__global__ void ema_apply_kernel(float *__restrict__ E, const float *__restrict__ n_ema,
                                 float *__restrict__ m_ema, const float *__restrict__ sums,
                                 const double *__restrict__ tot, int K, int D, float gamma, float eps) {
    const float t = (float)(*tot);
    const int64_t total = (int64_t)K * D;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int k = (int)(i / D);
        const float mv = gamma * m_ema[i] + (1.0f - gamma) * sums[i];
        m_ema[i] = mv;
        const float ns = (n_ema[k] + eps) / (t + K * eps) * t;
        E[i] = mv / ns;
    }
}

}  // namespace lipvq
