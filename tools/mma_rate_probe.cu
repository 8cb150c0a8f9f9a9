// mma_rate_probe.cu -- issue-to-completion cost of back-to-back tcgen05.mma by kind, operand majorness and N, the
// numbers behind the backward-GEMM design notes in DESIGN.md (the backward GEMMs use kind::tf32 with K = 8 per
// instruction, so the PER-INSTRUCTION floor of the tensor core decides them, not the flop rate).
//   mode 1: kind::f16  SS, A and B K-major (the search / encoder form), K = 16
//   mode 2: kind::tf32 SS, A and B K-major, K = 8
//   mode 3: kind::tf32 SS, A and B MN-major (the wgrad form), K = 8
//   mode 4: kind::tf32 TS, A in tensor memory, B MN-major (the dgrad form), K = 8
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/_bin/mma_rate_probe tools/mma_rate_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#include "../lipvq_vae_b200/csrc/ptx_sm100.cuh"

using namespace lipvq;
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint64_t desc_mn(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46) |
           (1ull << 61);
}

__global__ void __launch_bounds__(128, 1) probe(int mode, int n, int count, int n_acc, long long *out) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = ptx::smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t *gen = smem_raw + (base - raw);
    const uint32_t sbar = base + 98304;
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(gen + 98304 + 16);
    for (int i = threadIdx.x; i < 98304 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(gen)[i] = 0u;
    ptx::fence_proxy_async();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { ptx::mbar_init(sbar, 1); ptx::fence_barrier_init(); }
    if (warp == 0) ptx::tmem_alloc<512>(ptx::smem_u32((const void *)tmem_slot));
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem0 = *tmem_slot;
    if (warp == 1 && lane == 0) {
        const int acc_stride = n_acc > 1 ? (n <= 128 ? 128 : 256) : 0;
        uint32_t idesc;
        uint64_t ad, bd;
        if (mode == 1) { idesc = ptx::umma_idesc_f16(128, n); ad = ptx::umma_desc_sw128(base); bd = ptx::umma_desc_sw128(base + 32768); }
        else if (mode == 2) { idesc = ptx::umma_idesc_tf32(128, n); ad = ptx::umma_desc_sw128(base); bd = ptx::umma_desc_sw128(base + 32768); }
        else if (mode == 3) { idesc = ptx::umma_idesc_tf32(128, n) | (1u << 15) | (1u << 16); ad = desc_mn(base, 2048, 512); bd = desc_mn(base + 32768, 2048, 512); }
        else { idesc = ptx::umma_idesc_tf32(128, n) | (1u << 16); ad = 0; bd = desc_mn(base + 32768, 2048, 512); }
        const uint32_t d0 = tmem0, d1 = tmem0 + (uint32_t)(n_acc > 1 ? acc_stride : 0);
        const long long t0 = clock64();
        if (mode == 1) {
#pragma unroll 1
            for (int i = 0; i < count; i += 4) {
                ptx::umma_bf16(d0, ad, bd, idesc, 1u);
                ptx::umma_bf16(d1, ad + 2, bd + 2, idesc, 1u);
                ptx::umma_bf16(d0, ad + 4, bd + 4, idesc, 1u);
                ptx::umma_bf16(d1, ad + 6, bd + 6, idesc, 1u);
            }
        } else if (mode == 2) {
#pragma unroll 1
            for (int i = 0; i < count; i += 4) {
                ptx::umma_tf32_ss(d0, ad, bd, idesc, 1u);
                ptx::umma_tf32_ss(d1, ad + 2, bd + 2, idesc, 1u);
                ptx::umma_tf32_ss(d0, ad + 4, bd + 4, idesc, 1u);
                ptx::umma_tf32_ss(d1, ad + 6, bd + 6, idesc, 1u);
            }
        } else if (mode == 3) {
#pragma unroll 1
            for (int i = 0; i < count; i += 4) {
                ptx::umma_tf32_ss(d0, ad, bd, idesc, 1u);
                ptx::umma_tf32_ss(d1, ad + 64, bd + 64, idesc, 1u);
                ptx::umma_tf32_ss(d0, ad, bd, idesc, 1u);
                ptx::umma_tf32_ss(d1, ad + 64, bd + 64, idesc, 1u);
            }
        } else {
#pragma unroll 1
            for (int i = 0; i < count; i += 4) {
                ptx::umma_tf32_ts(d0, tmem0 + 448u, bd, idesc, 1u);
                ptx::umma_tf32_ts(d1, tmem0 + 456u, bd + 64, idesc, 1u);
                ptx::umma_tf32_ts(d0, tmem0 + 464u, bd, idesc, 1u);
                ptx::umma_tf32_ts(d1, tmem0 + 472u, bd + 64, idesc, 1u);
            }
        }
        ptx::umma_commit(sbar);
        ptx::mbar_wait(sbar, 0);
        out[blockIdx.x] = clock64() - t0;
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc<512>(tmem0);
}

int main() {
    const int G = 148, count = 1024;
    long long *d;
    CK(cudaMalloc(&d, sizeof(long long) * G));
    const int smem = 98304 + 64 + 1024;
    CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const char *names[] = {"", "f16  SS K-major  K=16", "tf32 SS K-major  K=8 ", "tf32 SS MN-major K=8 ", "tf32 TS B MN-major K=8"};
    const int ns[] = {16, 32, 64, 128, 208, 256};
    for (int mode = 1; mode <= 4; ++mode)
        for (int n_acc : {1, 2})
            for (int n : ns) {
                if (n_acc == 2 && n > 256) continue;
                for (int rep = 0; rep < 2; ++rep) {
                    probe<<<G, 128, smem>>>(mode, n, count, n_acc, d);
                    CK(cudaDeviceSynchronize());
                }
                long long h[148];
                CK(cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost));
                double c = 0;
                for (int i = 0; i < G; ++i) c += h[i];
                c /= G;
                const double flop = 2.0 * 128 * n * (mode == 1 ? 16 : 8);
                printf("%s M=128 N=%-3d  %d accumulator(s): %7.1f cycles / MMA  (%6.0f flop/clk/SM)\n", names[mode], n, n_acc,
                       c / count, flop / (c / count));
            }
    return 0;
}
