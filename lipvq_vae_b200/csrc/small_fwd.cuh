// small_fwd.cuh -- the WHOLE no-grad forward (v5:70-84) as ONE kernel for the reference's real call sizes
// (N = B*T = 80 rows in the image config, 500 in the low-dim one: v5:89-93).  At these sizes the
// 12-launch tensor path is pure launch latency; here a (splits x row-groups) grid does everything in exact fp32 on
// the CUDA cores:
//   every CTA of row group g recomputes the encoder for its 4 rows (143 KB of weights from L2, ~36 k MACs per row),
//   scans ITS slice of the codebook in direct form sum (z - e)^2 (the reference's arithmetic, v5:41-45) with
//   lane = code over shared-memory staged tiles, and publishes (distance, code) per row with one 64-bit atomicMax;
//   the LAST CTA of the group to arrive (ticket counter) gathers z_q = E[idx], runs the decoder on its 4 rows and
//   writes the group's two loss partials; the last GROUP sums the partials in group order (deterministic).
// One launch + one memset node; no tensor cores (K-dims of 12/64/128 on 4 rows would idle them), no workspace
// beyond (8 B per row + 12 B per group).
#pragma once
#include "common.cuh"

namespace lipvq {
namespace small {

constexpr int SR = 4;            // rows per row group (CTA)
constexpr int ST = 512;          // threads per CTA (16 warps: the layers are chains of shuffles and MUFU ops, latency bound)
constexpr int TW = 8;            // warps that scan codebook tiles (each owns two 8.7 KB tile buffers)
constexpr int SDC = 64;          // latent columns per staged codebook chunk
constexpr int SPITCH = SDC + 4;  // padded pitch (floats): lane stride 17 float4 -> conflict-free 128-bit reads
constexpr int MAX_ROWS = 768;    // measured crossover with the multi-launch tensor path (~110 us of launches): 47 us at 80
                                 // rows, 66 us at 500, 121 us at 1024 for this kernel

// weight / bias read: a plain generic load -- the pointers are global memory in the direct variant and shared memory in
// the staged one
__device__ __forceinline__ float ldw(const float *p) { return *p; }

// 16-byte asynchronous global -> shared copies (LDGSTS): whole weight matrices and codebook tiles are put in flight at
// once, the kernel being a chain of L2 round trips otherwise.  src_bytes = 0 zero-fills the destination.
__device__ __forceinline__ void cp_async16(void *dst, const void *src, int src_bytes = 16) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src),
                 "r"(src_bytes)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
// n floats, whole CTA; 16-byte path when both sides allow it
__device__ __forceinline__ void stage_floats(float *dst, const float *__restrict__ src, int n) {
    if ((n & 3) == 0 && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0) {
        for (int i = threadIdx.x; i < (n >> 2); i += ST) cp_async16(dst + 4 * i, src + 4 * i);
    } else {
        for (int i = threadIdx.x; i < n; i += ST) dst[i] = __ldg(src + i);
    }
}

// one layer for the CTA's SR rows: warp per output neuron (NJ in flight), lanes stride the reduction axis
template <int ACT>
__device__ __forceinline__ void rows_layer(const float *__restrict__ W, const float *__restrict__ bias, int n_out,
                                           int k_in, const float *s_in /* [SR][k_in] */, float *s_out /* [SR][n_out] */,
                                           float *__restrict__ g_out /* nullable [SR][n_out] global */, int rows_valid) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = ST / 32;
    constexpr int NJ = 4;
    for (int j0 = warp * NJ; j0 < n_out; j0 += nwarp * NJ) {
        float acc[NJ][SR];
#pragma unroll
        for (int q = 0; q < NJ; ++q)
#pragma unroll
            for (int r = 0; r < SR; ++r) acc[q][r] = 0.f;
#pragma unroll 4
        for (int k = lane; k < k_in; k += 32) {      // (the kernel is latency bound: keep 16 weight loads in flight)
            float w[NJ];
#pragma unroll
            for (int q = 0; q < NJ; ++q) w[q] = j0 + q < n_out ? ldw(W + (size_t)(j0 + q) * k_in + k) : 0.f;
#pragma unroll
            for (int r = 0; r < SR; ++r) {
                const float a = s_in[r * k_in + k];
#pragma unroll
                for (int q = 0; q < NJ; ++q) acc[q][r] = fmaf(w[q], a, acc[q][r]);
            }
        }
#pragma unroll
        for (int q = 0; q < NJ; ++q) {
            const int j = j0 + q;
            if (j >= n_out) break;          // warp-uniform
#pragma unroll
            for (int r = 0; r < SR; ++r) acc[q][r] = warp_sum(acc[q][r]);
            if (lane < SR) {
                float v = 0.f;
#pragma unroll
                for (int r = 0; r < SR; ++r) v = lane == r ? acc[q][r] : v;
                v = apply_act<ACT>(v + ldw(bias + j));
                s_out[lane * n_out + j] = v;
                if (g_out && lane < rows_valid) g_out[(size_t)lane * n_out + j] = v;
            }
        }
    }
}

// z = sigmoid(h (W * scale)^T + b) with scale_j = min(1, softplus(ci_j) / sum_k |W_jk|) computed in the same pass
// over the row of W (v5:6-12, 22-24); the product W_jk * scale_j is rounded before the FMA, as in the reference
__device__ __forceinline__ void lipschitz_rows_layer(const float *__restrict__ W, const float *__restrict__ bias,
                                                     const float *__restrict__ ci, int n_out, int k_in,
                                                     const float *s_in, float *s_out, int out_pitch) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = ST / 32;
    constexpr int NJ = 4, KR = 4;       // neurons in flight per warp; weight registers per lane and neuron (k_in <= 128)
    if (k_in <= 32 * KR) {
        // latency-bound at these sizes: all NJ x KR weights of the step are loaded at once (one L2 round trip),
        // the row sums, the scale and the products come from registers
        for (int j0 = warp * NJ; j0 < n_out; j0 += nwarp * NJ) {
            float w[NJ][KR], s[NJ];
#pragma unroll
            for (int q = 0; q < NJ; ++q)
#pragma unroll
                for (int t = 0; t < KR; ++t) {
                    const int k = lane + 32 * t;
                    w[q][t] = (j0 + q < n_out && k < k_in) ? ldw(W + (size_t)(j0 + q) * k_in + k) : 0.f;
                }
#pragma unroll
            for (int q = 0; q < NJ; ++q) {
                float a = 0.f;
#pragma unroll
                for (int t = 0; t < KR; ++t) a += fabsf(w[q][t]);      // lane-strided order == the two-pass loop below
                s[q] = warp_sum(a);
            }
            float acc[NJ][SR];
#pragma unroll
            for (int q = 0; q < NJ; ++q) {
                const float sc = j0 + q < n_out ? fminf(1.0f, softplus_f(ldw(ci + j0 + q)) / s[q]) : 0.f;
#pragma unroll
                for (int r = 0; r < SR; ++r) acc[q][r] = 0.f;
#pragma unroll
                for (int t = 0; t < KR; ++t) {
                    const int k = lane + 32 * t;
                    if (k < k_in) {
                        const float ws = w[q][t] * sc;
#pragma unroll
                        for (int r = 0; r < SR; ++r) acc[q][r] = fmaf(ws, s_in[r * k_in + k], acc[q][r]);
                    }
                }
            }
#pragma unroll
            for (int q = 0; q < NJ; ++q) {
                const int j = j0 + q;
                if (j >= n_out) break;          // warp-uniform
#pragma unroll
                for (int r = 0; r < SR; ++r) acc[q][r] = warp_sum(acc[q][r]);
                if (lane < SR) {
                    float v = 0.f;
#pragma unroll
                    for (int r = 0; r < SR; ++r) v = lane == r ? acc[q][r] : v;
                    s_out[lane * out_pitch + j] = sigmoid_f(v + ldw(bias + j));
                }
            }
        }
        return;
    }
    for (int j = warp; j < n_out; j += nwarp) {
        const float *wr = W + (size_t)j * k_in;
        float s = 0.f;
        for (int k = lane; k < k_in; k += 32) s += fabsf(ldw(wr + k));
        s = warp_sum(s);
        const float sc = fminf(1.0f, softplus_f(ldw(ci + j)) / s);
        float acc[SR];
#pragma unroll
        for (int r = 0; r < SR; ++r) acc[r] = 0.f;
        for (int k = lane; k < k_in; k += 32) {
            const float w = ldw(wr + k) * sc;
#pragma unroll
            for (int r = 0; r < SR; ++r) acc[r] = fmaf(w, s_in[r * k_in + k], acc[r]);
        }
#pragma unroll
        for (int r = 0; r < SR; ++r) acc[r] = warp_sum(acc[r]);
        if (lane < SR) {
            float v = 0.f;
#pragma unroll
            for (int r = 0; r < SR; ++r) v = lane == r ? acc[r] : v;
            s_out[lane * out_pitch + j] = sigmoid_f(v + ldw(bias + j));
        }
    }
}

// ---- thread-per-neuron layers over PADDED shared-memory weight rows (the default when every dimension is a multiple
// of 4).  The warp-per-neuron form above spends 5 shuffle + 5 add instructions per (neuron, row) reduction -- ncu
// (round 2) put 45 % of the kernel's instructions in the 208 x 128 Lipschitz layer alone.  Here a thread owns one
// neuron and a 1 / KS slice of the reduction axis: 128-bit reads of its weight row (row pitch = an ODD number of
// 16-byte chunks: conflict-free across the lanes of a warp), 128-bit broadcast reads of the SR activation rows, 4 * SR
// FMAs per weight chunk and NO shuffles; the KS slices meet in shared memory and are summed in slice order.
__host__ __device__ inline int wpitch(int k_in) {
    int p4 = (k_in + 3) >> 2;
    if (!(p4 & 1)) ++p4;
    return p4 << 2;
}
// [n_rows][k_in] row-major (k_in % 4 == 0) -> padded rows; 16-byte asynchronous copies when the source allows
__device__ __forceinline__ void stage_rows(float *dst, const float *__restrict__ src, int n_rows, int k_in, int pitch) {
    const int c4 = k_in >> 2;
    if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
        for (int i = threadIdx.x; i < n_rows * c4; i += ST) {
            const int j = i / c4, c = i - j * c4;
            cp_async16(dst + j * pitch + 4 * c, src + (size_t)j * k_in + 4 * c);
        }
    } else {
        for (int i = threadIdx.x; i < n_rows * k_in; i += ST) {
            const int j = i / k_in, c = i - j * k_in;
            dst[j * pitch + c] = __ldg(src + i);
        }
    }
}
constexpr int TPN_SCRATCH = SR * ST + 2 * ST;   // floats: slice partials [KS][SR][n_pad] | |W| partials [KS][n_pad] | scale
// LIP: z = act(h (W * scale)^T + b), scale_j = min(1, softplus(ci_j) / sum_k |W_jk|), the product W_jk * scale_j rounded
// before the FMA as in the reference (v5:6-12, 22-24).  Contains block barriers: call from all ST threads.
template <int ACT, bool LIP>
__device__ __forceinline__ void rows_layer_tpn(const float *sW, int P, const float *bias, const float *ci, int n_out,
                                               int k_in, const float *s_in, int in_pitch, float *s_out, int out_pitch,
                                               float *s_scr, float *__restrict__ g_out, int rows_valid) {
    const int tid = threadIdx.x;
    const int n_pad = (n_out + 31) & ~31;
    int KS = ST / n_pad;
    KS = KS < 1 ? 1 : (KS > 4 ? 4 : KS);
    const int k4 = k_in >> 2, kc4 = (k4 + KS - 1) / KS;
    float *s_part = s_scr, *s_abs = s_scr + SR * ST, *s_scale = s_abs + ST;
    const int s = tid / n_pad, j = tid - s * n_pad;
    const bool slot = s < KS, on = slot && j < n_out;
    const int q_lo = s * kc4, q_hi = (q_lo + kc4 < k4) ? q_lo + kc4 : k4;
    const float4 *wrow = reinterpret_cast<const float4 *>(sW + (size_t)(on ? j : 0) * P);
    float sc = 1.f;
    if (LIP) {
        float a = 0.f;
        if (on)
            for (int q = q_lo; q < q_hi; ++q) {
                const float4 w = wrow[q];
                a += (fabsf(w.x) + fabsf(w.y)) + (fabsf(w.z) + fabsf(w.w));
            }
        if (slot) s_abs[s * n_pad + j] = a;
        __syncthreads();
        if (tid < n_out) {
            float t = 0.f;
            for (int s2 = 0; s2 < KS; ++s2) t += s_abs[s2 * n_pad + tid];
            s_scale[tid] = fminf(1.0f, softplus_f(ldw(ci + tid)) / t);
        }
        __syncthreads();
        if (on) sc = s_scale[j];
    }
    float acc[SR];
#pragma unroll
    for (int r = 0; r < SR; ++r) acc[r] = 0.f;
    if (on) {
#pragma unroll 2
        for (int q = q_lo; q < q_hi; ++q) {
            float4 w = wrow[q];
            if (LIP) { w.x *= sc; w.y *= sc; w.z *= sc; w.w *= sc; }
#pragma unroll
            for (int r = 0; r < SR; ++r) {
                const float4 a = *reinterpret_cast<const float4 *>(s_in + r * in_pitch + 4 * q);   // broadcast
                acc[r] = fmaf(w.x, a.x, acc[r]);
                acc[r] = fmaf(w.y, a.y, acc[r]);
                acc[r] = fmaf(w.z, a.z, acc[r]);
                acc[r] = fmaf(w.w, a.w, acc[r]);
            }
        }
    }
    if (slot) {
#pragma unroll
        for (int r = 0; r < SR; ++r) s_part[(s * SR + r) * n_pad + j] = acc[r];
    }
    __syncthreads();
    for (int i = tid; i < SR * n_out; i += ST) {
        const int r = i / n_out, j2 = i - r * n_out;
        float v = 0.f;
        for (int s2 = 0; s2 < KS; ++s2) v += s_part[(s2 * SR + r) * n_pad + j2];
        v = apply_act<ACT>(v + ldw(bias + j2));
        s_out[r * out_pitch + j2] = v;
        if (g_out && r < rows_valid) g_out[(size_t)r * n_out + j2] = v;
    }
}

struct Scratch {
    unsigned long long *best;   // [n_rows] ~((d2 bits << 32) | code), zero-initialised (= no candidate yet)
    int *group_done;            // [G] arrival counters
    int *groups_done;           // [1]
    float *part;                // [G][2] {sse_q, sse_r}
};
inline size_t scratch_bytes(int64_t n) {
    const int64_t G = (n + SR - 1) / SR;
    return align_up((size_t)n * 8, 256) + align_up((size_t)(G + 1) * 4, 256) + align_up((size_t)G * 8, 256) + 256;
}
// everything that must start at zero comes first and is contiguous: best | counters
inline Scratch carve(void *ws, int64_t n, size_t *zero_bytes) {
    const int64_t G = (n + SR - 1) / SR;
    char *b = (char *)align_up((size_t)ws, 256);
    Scratch s;
    s.best = (unsigned long long *)b;
    b += align_up((size_t)n * 8, 256);
    s.group_done = (int *)b;
    s.groups_done = s.group_done + G;
    b += align_up((size_t)(G + 1) * 4, 256);
    *zero_bytes = (size_t)(b - (char *)s.best);
    s.part = (float *)b;
    return s;
}

// STAGED: the encoder (then, in the finishing CTA, the decoder) weights are copied into shared memory with one wave of
// 16-byte asynchronous copies, and the codebook tiles are double buffered, so the kernel pays a handful of L2 round
// trips instead of ~40 dependent ones; needs enc/dec weights <= ~150 KB (latent_dim <= ~256), else the direct variant.
template <int MODE>   // 0: weights read in place; 1: staged, warp-per-neuron layers; 2: staged + padded, thread-per-neuron layers
__global__ void __launch_bounds__(ST, 1) small_forward_kernel(
    lipvq_params_t p, const float *__restrict__ x, int n_rows, int F, int H1, int H, int D, int K, int splits,
    Scratch sc, float *__restrict__ z_latent, int64_t *__restrict__ idx_out, float *__restrict__ x_recon,
    float *__restrict__ loss_out, double inv_nf, double inv_nd, int32_t *__restrict__ stats) {
    constexpr bool STAGED = MODE != 0, TPN = MODE == 2;
    // search statistics of this path are all zero (no tensor-core tiers): written here instead of by a memset node --
    // a call is then two graph nodes, not three, and node boundaries are what a 30 us call is made of
    if (stats && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x < 4) stats[threadIdx.x] = 0;
    extern __shared__ __align__(16) float sm[];
    // [SR][F] x | [SR][H1] | [SR][H] | [SR][ZP] z_e (zero-padded to whole chunks) | [8 warps][32][SPITCH] codebook
    // tiles (re-used as z_q + decoder activations by the finishing CTA)
    const int ZP = (D + SDC - 1) / SDC * SDC;
    float *s_x = sm, *s_h1 = s_x + SR * F, *s_h2 = s_h1 + SR * H1, *s_z = s_h2 + SR * H;
    float *s_scr = s_z + SR * ZP;       // TPN: slice partials of the thread-per-neuron layers
    float *s_tile = s_scr + (TPN ? TPN_SCRATCH : 0);   // big region: STAGED: encoder weights, then 2 x tiles, then decoder weights
    __shared__ float red[ST / 32];
    __shared__ int s_ticket;
    __shared__ int s_idx[SR];
    __shared__ unsigned long long s_key[TW][SR];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = blockIdx.y, split = blockIdx.x;
    const int r0 = g * SR;
    const int rows_valid = min(SR, n_rows - r0);

    for (int i = tid; i < SR * F; i += ST) {
        const int r = i / F;
        s_x[i] = r < rows_valid ? x[(size_t)r0 * F + i] : 0.f;
    }
    for (int i = tid; i < SR * ZP; i += ST) s_z[i] = 0.f;
    const float *w0 = p.enc0_w, *b0 = p.enc0_b, *w1 = p.enc2_w, *b1 = p.enc2_b, *wl = p.lat_W, *bl = p.lat_b, *cl = p.lat_ci;
    if (STAGED) {
        float *q = s_tile;               // (every piece starts 16-byte aligned: sizes are rounded up to 4 floats)
        auto put = [&](const float *src, int n) { float *d = q; stage_floats(d, src, n); q += (n + 3) & ~3; return (const float *)d; };
        auto put_rows = [&](const float *src, int n_out, int k_in) {
            if (!TPN) return put(src, n_out * k_in);
            float *d = q;
            stage_rows(d, src, n_out, k_in, wpitch(k_in));
            q += n_out * wpitch(k_in);
            return (const float *)d;
        };
        w0 = put_rows(p.enc0_w, H1, F); b0 = put(p.enc0_b, H1); w1 = put_rows(p.enc2_w, H, H1); b1 = put(p.enc2_b, H);
        cp_async_commit();               // group 1: the two MLP layers (37 KB) ...
        wl = put_rows(p.lat_W, D, H); bl = put(p.lat_b, D); cl = put(p.lat_ci, D);
        cp_async_commit();               // ... group 2: the Lipschitz layer (108 KB) keeps landing while they run
        cp_async_wait<1>();
    }
    __syncthreads();
    if (TPN) {
        rows_layer_tpn<ACT_GELU, false>(w0, wpitch(F), b0, nullptr, H1, F, s_x, F, s_h1, H1, s_scr, nullptr, 0);
        __syncthreads();
        rows_layer_tpn<ACT_GELU, false>(w1, wpitch(H1), b1, nullptr, H, H1, s_h1, H1, s_h2, H, s_scr, nullptr, 0);
        if (STAGED) cp_async_wait<0>();
        __syncthreads();
        rows_layer_tpn<ACT_SIGMOID, true>(wl, wpitch(H), bl, cl, D, H, s_h2, H, s_z, ZP, s_scr, nullptr, 0);
    } else {
        rows_layer<ACT_GELU>(w0, b0, H1, F, s_x, s_h1, nullptr, 0);
        __syncthreads();
        rows_layer<ACT_GELU>(w1, b1, H, H1, s_h1, s_h2, nullptr, 0);
        if (STAGED) cp_async_wait<0>();
        __syncthreads();
        lipschitz_rows_layer(wl, bl, cl, D, H, s_h2, s_z, ZP);
    }
    __syncthreads();          // (STAGED: the weight region is free from here on: the codebook tiles take it over)

    // ---- this CTA's slice of the codebook: codes [k_lo, k_hi), 32-code tiles dealt to the warps round-robin
    {
        const int per = ((K + splits - 1) / splits + 31) / 32 * 32;
        const int k_lo = split * per, k_hi = min(K, k_lo + per);
        const bool vec = (D & 3) == 0;
        float *tile = s_tile + (warp < TW ? warp : 0) * (STAGED ? 2 : 1) * 32 * SPITCH;
        // STAGED: chunk i+1 of this warp's (tile, column chunk) sequence is in flight while chunk i is used
        const int nchunk = (D + SDC - 1) / SDC;
        auto issue_chunk = [&](int t0, int d0, float *buf) {
            const int dc = min(SDC, D - d0);
            for (int i = lane; i < 32 * (SDC / 4); i += 32) {
                const int c = i / (SDC / 4), q = i - c * (SDC / 4);
                const int code = t0 + c;
                const bool in = code < k_hi && 4 * q < dc;
                cp_async16(buf + c * SPITCH + 4 * q, p.codebook + (in ? (size_t)code * D + d0 + 4 * q : 0), in ? 16 : 0);
            }
            cp_async_commit();
        };
        const bool pipelined = STAGED && vec;
        if (pipelined && warp < TW && k_lo + warp * 32 < k_hi) issue_chunk(k_lo + warp * 32, 0, tile);
        int seq = 0;
        float bd[SR];
        int bk[SR];
#pragma unroll
        for (int r = 0; r < SR; ++r) { bd[r] = INFINITY; bk[r] = 0x7fffffff; }
        for (int t0 = k_lo + warp * 32; warp < TW && t0 < k_hi; t0 += TW * 32) {
            float acc[SR];
#pragma unroll
            for (int r = 0; r < SR; ++r) acc[r] = 0.f;
            for (int d0 = 0; d0 < D; d0 += SDC) {
                const int dc = min(SDC, D - d0);
                // stage 32 codes x SDC columns (zero past the row / the slice): coalesced global reads, padded pitch
                float *cur = tile;
                if (pipelined) {
                    cur = tile + (seq & 1) * 32 * SPITCH;
                    float *nxt = tile + ((seq + 1) & 1) * 32 * SPITCH;
                    __syncwarp();                        // every lane is done reading the buffer about to be refilled
                    const bool last_chunk = d0 + SDC >= D;
                    const int nt0 = last_chunk ? t0 + TW * 32 : t0, nd0 = last_chunk ? 0 : d0 + SDC;
                    if (nt0 < k_hi) { issue_chunk(nt0, nd0, nxt); cp_async_wait<1>(); }
                    else cp_async_wait<0>();
                    ++seq;
                    (void)nchunk;
                } else
                __syncwarp();
                if (pipelined) {
                } else if (vec) {
                    for (int i = lane; i < 32 * (SDC / 4); i += 32) {
                        const int c = i / (SDC / 4), q = i - c * (SDC / 4);
                        const int code = t0 + c;
                        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (code < k_hi && 4 * q < dc)
                            v = __ldg(reinterpret_cast<const float4 *>(p.codebook + (size_t)code * D + d0) + q);
                        *reinterpret_cast<float4 *>(tile + c * SPITCH + 4 * q) = v;
                    }
                } else {
                    for (int i = lane; i < 32 * SDC; i += 32) {
                        const int c = i / SDC, d = i - c * SDC;
                        const int code = t0 + c;
                        tile[c * SPITCH + d] = (code < k_hi && d < dc) ? __ldg(p.codebook + (size_t)code * D + d0 + d) : 0.f;
                    }
                }
                __syncwarp();
                const float4 *tl = reinterpret_cast<const float4 *>(cur + lane * SPITCH);
                const int nq = (dc + 3) >> 2;
#pragma unroll 4
                for (int q = 0; q < nq; ++q) {
                    const float4 e = tl[q];
#pragma unroll
                    for (int r = 0; r < SR; ++r) {
                        const float4 z = *reinterpret_cast<const float4 *>(s_z + r * ZP + d0 + 4 * q);   // broadcast
                        const float a = z.x - e.x, b = z.y - e.y, c2 = z.z - e.z, d3 = z.w - e.w;
                        acc[r] = fmaf(a, a, acc[r]);
                        acc[r] = fmaf(b, b, acc[r]);
                        acc[r] = fmaf(c2, c2, acc[r]);
                        acc[r] = fmaf(d3, d3, acc[r]);
                    }
                }
            }
            const int code = t0 + lane;
            if (code < k_hi) {
#pragma unroll
                for (int r = 0; r < SR; ++r)
                    if (acc[r] < bd[r]) { bd[r] = acc[r]; bk[r] = code; }     // codes ascend per lane: first minimum wins
            }
        }
        // warp argmin per row (lowest code on ties), then one 64-bit atomic per (warp, row)
#pragma unroll
        for (int r = 0; r < SR; ++r) {
            float d = bd[r];
            int k = bk[r];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float od = __shfl_xor_sync(0xffffffffu, d, o);
                const int ok = __shfl_xor_sync(0xffffffffu, k, o);
                if (od < d || (od == d && ok < k)) { d = od; k = ok; }
            }
            // the tile warps meet in shared memory first: ONE global atomic per (CTA, row) -- the same-address atomics of
            // a row group serialise in L2, and their acknowledgement is what the fence below waits for
            if (lane == 0 && warp < TW) {
                const unsigned long long key = ((unsigned long long)__float_as_uint(d) << 32) | (unsigned)k;
                s_key[warp][r] = (k != 0x7fffffff) ? ~key : 0ull;
            }
        }
    }
    __syncthreads();
    if (tid < rows_valid) {
        unsigned long long m = 0ull;
#pragma unroll
        for (int w = 0; w < TW; ++w) m = s_key[w][tid] > m ? s_key[w][tid] : m;
        if (m) atomicMax(sc.best + r0 + tid, m);
    }
    // ---- the last CTA of the row group finishes the rows
    __threadfence();
    __syncthreads();
    if (tid == 0) s_ticket = atomicAdd(sc.group_done + g, 1);
    __syncthreads();
    if (s_ticket != splits - 1) return;
    __threadfence();
    if (tid < SR) {
        int k = 0;
        if (tid < rows_valid) {
            const unsigned long long key = ~(*reinterpret_cast<volatile unsigned long long *>(sc.best + r0 + tid));
            k = (int)(key & 0xffffffffu);
            if (k < 0 || k >= K) k = 0;
            idx_out[r0 + tid] = (int64_t)k;
        }
        s_idx[tid] = k;
    }
    __syncthreads();
    float *s_q = s_tile, *s_c1 = s_q + SR * D, *s_c2 = s_c1 + SR * H1, *s_o = s_c2 + SR * H;
    const float *v0 = p.dec0_w, *c0 = p.dec0_b, *v1 = p.dec2_w, *c1 = p.dec2_b, *vo = p.out_w, *co = p.out_b;
    if (STAGED) {
        float *q = s_o + ((SR * F + 3) & ~3);
        auto put = [&](const float *src, int n) { float *d = q; stage_floats(d, src, n); q += (n + 3) & ~3; return (const float *)d; };
        auto put_rows = [&](const float *src, int n_out, int k_in) {
            if (!TPN) return put(src, n_out * k_in);
            float *d = q;
            stage_rows(d, src, n_out, k_in, wpitch(k_in));
            q += n_out * wpitch(k_in);
            return (const float *)d;
        };
        v0 = put_rows(p.dec0_w, H1, D); c0 = put(p.dec0_b, H1);
        cp_async_commit();               // group 1 lands while the gather below runs ...
        v1 = put_rows(p.dec2_w, H, H1); c1 = put(p.dec2_b, H);
        vo = put_rows(p.out_w, F, H); co = put(p.out_b, F);
        cp_async_commit();               // ... group 2 while the first decoder layer does
    }
    float sq = 0.f;
    for (int i = tid; i < SR * D; i += ST) {
        const int r = i / D, d = i - r * D;
        const float e = __ldg(p.codebook + (size_t)s_idx[r] * D + d);
        s_q[i] = e;
        if (r < rows_valid) {
            if (z_latent) z_latent[(size_t)(r0 + r) * D + d] = e;
            const float df = e - s_z[r * ZP + d];
            sq = fmaf(df, df, sq);
        }
    }
    if (STAGED) cp_async_wait<1>();
    __syncthreads();
    if (TPN) {
        rows_layer_tpn<ACT_GELU, false>(v0, wpitch(D), c0, nullptr, H1, D, s_q, D, s_c1, H1, s_scr, nullptr, 0);
        if (STAGED) cp_async_wait<0>();
        __syncthreads();
        rows_layer_tpn<ACT_GELU, false>(v1, wpitch(H1), c1, nullptr, H, H1, s_c1, H1, s_c2, H, s_scr, nullptr, 0);
        __syncthreads();
        rows_layer_tpn<ACT_NONE, false>(vo, wpitch(H), co, nullptr, F, H, s_c2, H, s_o, F, s_scr,
                                        x_recon ? x_recon + (size_t)r0 * F : nullptr, rows_valid);
    } else {
        rows_layer<ACT_GELU>(v0, c0, H1, D, s_q, s_c1, nullptr, 0);
        if (STAGED) cp_async_wait<0>();
        __syncthreads();
        rows_layer<ACT_GELU>(v1, c1, H, H1, s_c1, s_c2, nullptr, 0);
        __syncthreads();
        rows_layer<ACT_NONE>(vo, co, F, H, s_c2, s_o, x_recon ? x_recon + (size_t)r0 * F : nullptr, rows_valid);
    }
    __syncthreads();
    float sr = 0.f;
    for (int i = tid; i < rows_valid * F; i += ST) {
        const float df = s_o[i] - s_x[i];
        sr = fmaf(df, df, sr);
    }
    sq = block_sum<ST>(sq, red);
    sr = block_sum<ST>(sr, red);
    if (tid == 0) {
        sc.part[2 * g] = sq;
        sc.part[2 * g + 1] = sr;
        __threadfence();
        s_ticket = atomicAdd(sc.groups_done, 1);
    }
    __syncthreads();
    if (s_ticket != (int)gridDim.y - 1 || loss_out == nullptr) return;
    __threadfence();
    if (warp == 0) {       // every group's partials, in group order (lane-strided, fixed shuffle tree): deterministic
        double a = 0.0, b = 0.0;
        const volatile float *pp = sc.part;
        for (int i = lane; i < (int)gridDim.y; i += 32) { a += (double)pp[2 * i]; b += (double)pp[2 * i + 1]; }
        a = warp_sum_d(a);
        b = warp_sum_d(b);
        if (lane == 0) {
            const double recon = b * inv_nf, commit = a * inv_nd;
            loss_out[0] = (float)(recon + 0.25 * commit + 0.25 * commit);
            loss_out[1] = (float)recon;
            loss_out[2] = (float)commit;
            loss_out[3] = (float)commit;
        }
    }
}

inline size_t r4(size_t n) { return (n + 3) & ~(size_t)3; }
inline size_t smem_bytes(const lipvq_dims_t *d, int mode = 0) {
    size_t head = (size_t)SR * (d->F + d->H1 + d->H + (d->D + SDC - 1) / SDC * SDC);
    if (mode == 2) head += TPN_SCRATCH;
    size_t tiles = (size_t)TW * 32 * SPITCH * (mode ? 2 : 1);
    size_t tail = (size_t)SR * (d->D + d->H1 + d->H) + r4((size_t)SR * d->F);
    auto wsz = [&](int n_out, int k_in) { return mode == 2 ? (size_t)n_out * wpitch(k_in) : r4((size_t)n_out * k_in); };
    if (mode) {
        const size_t enc = wsz(d->H1, d->F) + r4(d->H1) + wsz(d->H, d->H1) + r4(d->H) + wsz(d->D, d->H) + 2 * r4(d->D);
        const size_t dec = wsz(d->H1, d->D) + r4(d->H1) + wsz(d->H, d->H1) + r4(d->H) + wsz(d->F, d->H) + r4(d->F);
        tail += dec;
        if (enc > tiles) tiles = enc;
    }
    return (head + (tiles > tail ? tiles : tail)) * sizeof(float) + 64;
}
inline bool staged_ok(const lipvq_dims_t *d) { return smem_bytes(d, 1) <= 200 * 1024; }
// thread-per-neuron layers: 16-byte rows everywhere, every layer's neurons fit one pass of the CTA
inline bool tpn_ok(const lipvq_dims_t *d) {
    const int mx = d->D > d->H ? (d->D > d->H1 ? d->D : d->H1) : (d->H > d->H1 ? d->H : d->H1);
    return ((d->F | d->H1 | d->H | d->D) & 3) == 0 && mx <= ST && d->F <= ST && smem_bytes(d, 2) <= 200 * 1024;
}
inline bool supported(const lipvq_dims_t *d, int64_t n) {
    return n >= 1 && n <= MAX_ROWS && smem_bytes(d) <= 200 * 1024;
}

inline int launch(const lipvq_dims_t *d, const lipvq_params_t *p, const float *x, int64_t n, int64_t n_total,
                  float *z_latent, int64_t *idx, float *x_recon, float *loss_out, int32_t *stats, void *ws,
                  cudaStream_t st) {
    size_t zero_bytes = 0;
    const Scratch sc = carve(ws, n, &zero_bytes);
    LIPVQ_CUDA_OK(cudaMemsetAsync(sc.best, 0, zero_bytes, st));
    const int G = (int)((n + SR - 1) / SR);
    // enough CTAs to cover the chip about twice, at least 32 codes per split
    // ONE wave of CTAs (the kernel is a chain of dependent L2 round trips: a second wave doubles its latency): as many
    // codebook splits per row group as fit on the chip, whole 32-code tiles per split
    int splits = kNumSMs / G;
    if (splits < 1) splits = 1;
    if (splits > 64) splits = 64;
    const int per = ((d->K + splits - 1) / splits + 31) / 32 * 32;     // codes per split
    splits = (d->K + per - 1) / per;                                   // (no empty slices)
    static const bool no_stage = [] { const char *e = getenv("LIPVQ_SMALL_STAGED"); return e && !strcmp(e, "0"); }();
    static const bool no_tpn = [] { const char *e = getenv("LIPVQ_SMALL_LAYERS"); return e && !strcmp(e, "warp"); }();   // A/B
    const int mode = no_stage ? 0 : (!no_tpn && tpn_ok(d)) ? 2 : staged_ok(d) ? 1 : 0;
    const size_t smem = smem_bytes(d, mode);
    auto kern = mode == 2 ? small_forward_kernel<2> : mode == 1 ? small_forward_kernel<1> : small_forward_kernel<0>;
    LIPVQ_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const double nt = (double)(n_total > 0 ? n_total : 1);
    ProfScope prof("small_forward_kernel", st);
    kern<<<dim3(splits, G), ST, smem, st>>>(*p, x, (int)n, d->F, d->H1, d->H, d->D, d->K, splits, sc,
                                                            z_latent, idx, x_recon, loss_out, 1.0 / (nt * d->F),
                                                            1.0 / (nt * d->D), stats);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

}  // namespace small
}  // namespace lipvq
