// common.cuh -- shared device helpers for the LipVQ-VAE tokenizer kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>
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

// ---- optional per-kernel timing (lipvq_profile_begin / lipvq_profile_end): CUDA events recorded on the
// launching stream right before and after a kernel launch, so bench.py reads every kernel's duration inside
// the real step.  Off by default (one relaxed load per launch site).
extern int g_prof_on;
int prof_open(const char *name, cudaStream_t st);   // -> record index or -1
void prof_close(int rec, cudaStream_t st);
struct ProfScope {
    int rec; cudaStream_t st;
    ProfScope(const char *name, cudaStream_t s) : rec(g_prof_on ? prof_open(name, s) : -1), st(s) {}
    ~ProfScope() { if (rec >= 0) prof_close(rec, st); }
};

// NVTX range around every C-ABI entry point (tracing).  nvtx3 is header-only and resolves the
// tool's injection library lazily: without a profiler attached a push/pop is one predictable branch.
struct NvtxRange {
    explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};
#define LIPVQ_NVTX(name) ::lipvq::NvtxRange _nvtx_range_(name)

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs

// *_FAST: the per-row epilogues of the fused encoder (error at the fp32 rounding floor, see gelu_fast_f)
enum Act : int { ACT_NONE = 0, ACT_GELU = 1, ACT_SIGMOID = 2, ACT_GELU_FAST = 3, ACT_SIGMOID_FAST = 4, ACT_RELU = 5 };

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

__device__ __forceinline__ float ex2_approx(float t) {
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(t));
    return e;
}
__device__ __forceinline__ float rcp_approx(float t) {
    float e;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(t));
    return e;
}
// exact-erf GELU as x * Phi(x) with the upper tail r = 0.5 erfc(|x|/sqrt 2) = 2^q(|x|), q a degree-7 minimax fit
// of log2 r on [0, 6] (weighted for absolute error; fit error 5e-8 in r).  Phi = 1 - r (x >= 0) or r (x < 0), so
// the negative tail keeps its relative accuracy.  Measured against float64 on 2M points of [-12, 12]:
// max |gelu_fast - gelu| = 6.4e-7 (at x ~ 4.2, i.e. the fp32 rounding of x * Phi itself); 11 instructions
// against ~35 for 0.5 x (1 + erff(x / sqrt 2)).
__device__ __forceinline__ float gelu_fast_f(float x) {
    const float u = fminf(fabsf(x), 6.0f);
    float q = 5.184764121e-06f;
    q = fmaf(q, u, -2.271982345e-05f);
    q = fmaf(q, u, -5.336819246e-04f);
    q = fmaf(q, u, 7.583174311e-03f);
    q = fmaf(q, u, -5.285591416e-02f);
    q = fmaf(q, u, -4.590896734e-01f);
    q = fmaf(q, u, -1.151119478e+00f);
    q = fmaf(q, u, -9.999997251e-01f);
    const float r = ex2_approx(q);
    return x * (x >= 0.f ? 1.0f - r : r);
}
// d/dx [x Phi(x)] = Phi(x) + x phi(x) with the same tail fit for Phi and phi(x) = 2^(-x^2 log2(e)/2) / sqrt(2 pi) on
// the ex2 unit: absolute error < 1e-6 on a function of range [-0.13, 1.13]; 16 instructions against ~50 for the
// erff / expf form (the backward's GELU epilogue is instruction-bound)
__device__ __forceinline__ float gelu_grad_fast_f(float x) {
    const float u = fminf(fabsf(x), 6.0f);
    float q = 5.184764121e-06f;
    q = fmaf(q, u, -2.271982345e-05f);
    q = fmaf(q, u, -5.336819246e-04f);
    q = fmaf(q, u, 7.583174311e-03f);
    q = fmaf(q, u, -5.285591416e-02f);
    q = fmaf(q, u, -4.590896734e-01f);
    q = fmaf(q, u, -1.151119478e+00f);
    q = fmaf(q, u, -9.999997251e-01f);
    const float r = ex2_approx(q);
    const float pdf = 0.39894228040143268f * ex2_approx(-0.72134752044448170f * x * x);
    return fmaf(x, pdf, x >= 0.f ? 1.0f - r : r);
}
// 1 / (1 + 2^(-a log2 e)) with the approximate ex2 / rcp units: relative error < 1e-6 (ex2.approx 2^-22, the
// rounding of a*log2e, rcp.approx 1 ulp); saturates to 0 / 1 like the reference for large |a|
__device__ __forceinline__ float sigmoid_fast_f(float a) {
    return rcp_approx(1.0f + ex2_approx(-1.4426950408889634f * a));
}

template <int ACT>
__device__ __forceinline__ float apply_act(float v) {
    if (ACT == ACT_GELU_FAST) return gelu_fast_f(v);
    if (ACT == ACT_SIGMOID_FAST) return sigmoid_fast_f(v);
    if (ACT == ACT_GELU) return gelu_f(v);
    if (ACT == ACT_SIGMOID) return sigmoid_f(v);
    if (ACT == ACT_RELU) return fmaxf(v, 0.f);
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
