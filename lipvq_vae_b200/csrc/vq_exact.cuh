// vq_exact.cuh -- exact fp32 nearest-codeword scan in the reference's direct form
// sum_i (z_i - e_ki)^2 (LFQQuantizer.forward, v5:41-46; the sqrt is monotone and dropped).
// Serves as LIPVQ_SEARCH_EXACT and as the tier-2 fallback of the tensor-core search for rows
// whose candidate list overflowed (then `list`/`list_count` select the rows).
#pragma once
#include "common.cuh"

namespace lipvq {

constexpr int XBM = 64, XBN = 64, XBK = 16, XLD = 68, XNT = 256;

// lexicographic (value, index) minimum == torch.argmin's "first minimal index"
__device__ __forceinline__ void argmin_merge(float &v, int &i, float ov, int oi) {
    if (ov < v || (ov == v && oi < i)) { v = ov; i = oi; }
}

template <bool VEC4>
__global__ void __launch_bounds__(XNT) vq_exact_kernel(
    const float *__restrict__ z, const float *__restrict__ E, int64_t n_rows, int D, int K,
    const int *__restrict__ list, const int *__restrict__ list_count,
    int64_t *__restrict__ idx_out, float *__restrict__ d2_out) {
    __shared__ __align__(16) float zs[XBK][XLD];
    __shared__ __align__(16) float es[XBK][XLD];
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    const int64_t total = list ? (int64_t)(*list_count) : n_rows;
    const int64_t n_tiles = (total + XBM - 1) / XBM;

    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int64_t r0 = tile * XBM;
        // the row this thread loads (lr) for the z tile
        const int lr = VEC4 ? (tid >> 2) : (tid >> 4);
        float best[4];
        int bidx[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { best[i] = INFINITY; bidx[i] = 0x7fffffff; }

        for (int c0 = 0; c0 < K; c0 += XBN) {
            float acc[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

            for (int k0 = 0; k0 < D; k0 += XBK) {
                if (VEC4) {
                    // 64 rows x 4 float4 = one tile per pass
                    const int kq = (tid & 3) * 4;
                    float4 vz = make_float4(0.f, 0.f, 0.f, 0.f), ve = vz;
                    const int64_t rr = r0 + lr;
                    if (rr < total && k0 + kq < D) {
                        const int64_t row = list ? (int64_t)list[rr] : rr;
                        vz = __ldg(reinterpret_cast<const float4 *>(z + row * D + k0 + kq));
                    }
                    if (c0 + lr < K && k0 + kq < D)
                        ve = __ldg(reinterpret_cast<const float4 *>(E + (int64_t)(c0 + lr) * D + k0 + kq));
                    zs[kq + 0][lr] = vz.x; zs[kq + 1][lr] = vz.y; zs[kq + 2][lr] = vz.z; zs[kq + 3][lr] = vz.w;
                    es[kq + 0][lr] = ve.x; es[kq + 1][lr] = ve.y; es[kq + 2][lr] = ve.z; es[kq + 3][lr] = ve.w;
                } else {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int e = tid + q * XNT;
                        const int k = e & (XBK - 1), r = e >> 4;
                        const int64_t rr = r0 + r;
                        float vz = 0.f, ve = 0.f;
                        if (rr < total && k0 + k < D) {
                            const int64_t row = list ? (int64_t)list[rr] : rr;
                            vz = __ldg(z + row * D + k0 + k);
                        }
                        if (c0 + r < K && k0 + k < D) ve = __ldg(E + (int64_t)(c0 + r) * D + k0 + k);
                        zs[k][r] = vz;
                        es[k][r] = ve;
                    }
                }
                __syncthreads();
#pragma unroll
                for (int k = 0; k < XBK; ++k) {
                    const float4 a = *reinterpret_cast<const float4 *>(&zs[k][ty * 4]);
                    const float4 b = *reinterpret_cast<const float4 *>(&es[k][tx * 4]);
                    const float av[4] = {a.x, a.y, a.z, a.w};
                    const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float d = av[i] - bv[j];
                            acc[i][j] = fmaf(d, d, acc[i][j]);
                        }
                }
                __syncthreads();
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int code = c0 + tx * 4 + j;
                if (code < K) {
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        if (acc[i][j] < best[i]) { best[i] = acc[i][j]; bidx[i] = code; }
                }
            }
        }
        // merge the 16 threads (tx) that share a row: they are 16 consecutive lanes
#pragma unroll
        for (int i = 0; i < 4; ++i) {
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, best[i], o);
                const int oi = __shfl_xor_sync(0xffffffffu, bidx[i], o);
                argmin_merge(best[i], bidx[i], ov, oi);
            }
            const int64_t rr = r0 + ty * 4 + i;
            if (tx == 0 && rr < total) {
                const int64_t row = list ? (int64_t)list[rr] : rr;
                idx_out[row] = (int64_t)bidx[i];
                if (d2_out) d2_out[row] = best[i];
            }
        }
    }
}

inline int vq_exact_launch(const float *z, const float *E, int64_t n_rows, int D, int K,
                           const int *list, const int *list_count, int64_t *idx, float *d2,
                           cudaStream_t st) {
    if (n_rows <= 0) return LIPVQ_OK;
    int64_t tiles = (n_rows + XBM - 1) / XBM;
    // fallback-list mode: the count lives on the device, so size the grid for a modest share
    // of the rows and let the grid-stride loop cover the rest.
    int64_t grid = list ? (tiles < 4 * kNumSMs ? tiles : 4 * kNumSMs) : (tiles < 65535 * 16 ? tiles : 65535 * 16);
    ProfScope prof("vq_exact_kernel", st);
    if ((D & 3) == 0)
        vq_exact_kernel<true><<<(unsigned)grid, XNT, 0, st>>>(z, E, n_rows, D, K, list, list_count, idx, d2);
    else
        vq_exact_kernel<false><<<(unsigned)grid, XNT, 0, st>>>(z, E, n_rows, D, K, list, list_count, idx, d2);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

}  // namespace lipvq
