// common.cuh -- shared device helpers for the LipVQ-VAE tokenizer kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/lipvq_b200.h"

namespace lipvq {

extern thread_local cudaError_t g_last_cuda_error;
extern long long g_launch_count;   // kernels launched by this library (bench.py's gpu_launches)

#define LIPVQ_CUDA_OK(expr)                                   \
    do {                                                      \
        cudaError_t _e = (expr);                              \
        if (_e != cudaSuccess) {                              \
            ::lipvq::g_last_cuda_error = _e;                  \
            return LIPVQ_ECUDA;                               \
        }                                                     \
    } while (0)

#define LIPVQ_LAUNCH_OK()                                     \
    do {                                                      \
        __atomic_fetch_add(&::lipvq::g_launch_count, 1LL, __ATOMIC_RELAXED); \
        cudaError_t _e = cudaGetLastError();                  \
        if (_e != cudaSuccess) {                              \
            ::lipvq::g_last_cuda_error = _e;                  \
            return LIPVQ_ECUDA;                               \
        }                                                     \
    } while (0)

#define LIPVQ_TRY(expr)                \
    do {                               \
        int _r = (expr);               \
        if (_r != LIPVQ_OK) return _r; \
    } while (0)

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs

enum Act : int { ACT_NONE = 0, ACT_GELU = 1, ACT_SIGMOID = 2 };

// exact-erf GELU, nn.GELU() default (v5:56,58,64,66)
__device__ __forceinline__ float gelu_f(float a) {
    return 0.5f * a * (1.0f + erff(a * 0.70710678118654752f));
}
__device__ __forceinline__ float gelu_grad_f(float a) {
    const float cdf = 0.5f * (1.0f + erff(a * 0.70710678118654752f));
    const float pdf = 0.39894228040143268f * expf(-0.5f * a * a);
    return cdf + a * pdf;
}
__device__ __forceinline__ float sigmoid_f(float a) { return 1.0f / (1.0f + expf(-a)); }
// F.softplus beta=1 threshold=20 (v5:10)
__device__ __forceinline__ float softplus_f(float c) { return c > 20.0f ? c : log1pf(expf(c)); }

template <int ACT>
__device__ __forceinline__ float apply_act(float v) {
    if (ACT == ACT_GELU) return gelu_f(v);
    if (ACT == ACT_SIGMOID) return sigmoid_f(v);
    return v;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Block-wide sum in a fixed order (deterministic); result valid in thread 0.  NT = blockDim.x.
template <int NT>
__device__ __forceinline__ float block_sum(float v, float *smem /* >= NT/32 floats */) {
    v = warp_sum(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) smem[w] = v;
    __syncthreads();
    float r = 0.f;
    if (w == 0) {
        r = (l < NT / 32) ? smem[l] : 0.f;
        r = warp_sum(r);
    }
    __syncthreads();
    return r;
}

static inline bool aligned16(const void *p) { return p != nullptr && (((uintptr_t)p) & 15u) == 0; }
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace lipvq
