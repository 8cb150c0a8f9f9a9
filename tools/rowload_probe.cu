// rowload_probe.cu -- how fast can an SM pull fp32 rows when every THREAD walks its own row (the access pattern a
// tcgen05.st-fed A operand forces: TMEM lane = row)?  Compares 128-bit and 256-bit per-thread loads at several depths of
// loads in flight with a warp-coalesced read of the same bytes.  Pitch 832 B (latent_dim 208).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/_bin/rowload_probe tools/rowload_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)
struct __align__(32) F8 { float v[8]; };
__device__ __forceinline__ F8 ldg256(const float *p) {
    F8 r;
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7]) : "l"(p));
    return r;
}
// MODE 0: thread-per-row, 128-bit loads; 1: thread-per-row, 256-bit loads; 2: coalesced 128-bit
template <int MODE, int DEPTH>
__global__ void __launch_bounds__(256) k(const float *__restrict__ a, int64_t rows, int cols, float *out) {
    float s = 0.f;
    if (MODE == 2) {
        const int64_t n4 = rows * cols / 4;
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x * DEPTH) {
            float4 v[DEPTH];
#pragma unroll
            for (int d = 0; d < DEPTH; ++d) {
                const int64_t j = i + (int64_t)d * gridDim.x * blockDim.x;
                v[d] = j < n4 ? __ldg(reinterpret_cast<const float4 *>(a) + j) : make_float4(0, 0, 0, 0);
            }
#pragma unroll
            for (int d = 0; d < DEPTH; ++d) s += v[d].x + v[d].y + v[d].z + v[d].w;
        }
    } else {
        // tile of 128 rows per 128 threads; the two halves of the block take alternate 16-column blocks (as dgrad does)
        const int r = threadIdx.x & 127, h = threadIdx.x >> 7;
        const int nb = cols / 16;
        for (int64_t t = blockIdx.x; t * 128 < rows; t += gridDim.x) {
            const float *row = a + (t * 128 + r) * cols;
            for (int b0 = h; b0 < nb; b0 += 2 * DEPTH) {
                if (MODE == 0) {
                    float4 v[DEPTH][4];
#pragma unroll
                    for (int d = 0; d < DEPTH; ++d)
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            v[d][i] = b0 + 2 * d < nb ? __ldg(reinterpret_cast<const float4 *>(row + (b0 + 2 * d) * 16) + i) : make_float4(0, 0, 0, 0);
#pragma unroll
                    for (int d = 0; d < DEPTH; ++d)
#pragma unroll
                        for (int i = 0; i < 4; ++i) s += v[d][i].x + v[d][i].w;
                } else {
                    F8 v[DEPTH][2];
#pragma unroll
                    for (int d = 0; d < DEPTH; ++d)
#pragma unroll
                        for (int i = 0; i < 2; ++i)
                            if (b0 + 2 * d < nb) v[d][i] = ldg256(row + (b0 + 2 * d) * 16 + 8 * i);
                            else v[d][i].v[0] = v[d][i].v[7] = 0.f;
#pragma unroll
                    for (int d = 0; d < DEPTH; ++d)
#pragma unroll
                        for (int i = 0; i < 2; ++i) s += v[d][i].v[0] + v[d][i].v[7];
                }
            }
        }
    }
    if (s == 1.2345f) out[0] = s;
}
template <int MODE, int DEPTH>
static void run(const float *a, int64_t rows, int cols, float *out, int ctas_per_sm) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const int grid = 148 * ctas_per_sm;
    for (int i = 0; i < 2; ++i) k<MODE, DEPTH><<<grid, 256>>>(a, rows, cols, out);
    CK(cudaEventRecord(e0));
    for (int i = 0; i < 5; ++i) k<MODE, DEPTH><<<grid, 256>>>(a, rows, cols, out);
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= 5;
    const char *names[] = {"thread-per-row 128-bit", "thread-per-row 256-bit", "warp-coalesced 128-bit"};
    printf("%s  loads in flight/thread %d x %d B, %d CTA(s) of 256 threads per SM: %7.1f us  %6.2f TB/s\n", names[MODE], DEPTH,
           MODE == 2 ? 16 : 64, ctas_per_sm, 1e3 * ms, (double)rows * cols * 4 / (ms * 1e-3) / 1e12);
}
int main() {
    const int64_t rows = 655360 * 2;   // 1.09 GB: larger than L2
    const int cols = 208;
    float *a, *out;
    CK(cudaMalloc(&a, rows * cols * 4)); CK(cudaMalloc(&out, 4));
    CK(cudaMemset(a, 0, rows * cols * 4));
    run<2, 4>(a, rows, cols, out, 8);
    run<2, 8>(a, rows, cols, out, 8);
    run<0, 1>(a, rows, cols, out, 1); run<0, 2>(a, rows, cols, out, 1); run<0, 4>(a, rows, cols, out, 1);
    run<0, 2>(a, rows, cols, out, 2); run<0, 4>(a, rows, cols, out, 2); run<0, 4>(a, rows, cols, out, 4);
    run<1, 1>(a, rows, cols, out, 1); run<1, 2>(a, rows, cols, out, 1); run<1, 4>(a, rows, cols, out, 1);
    run<1, 2>(a, rows, cols, out, 2); run<1, 4>(a, rows, cols, out, 2); run<1, 4>(a, rows, cols, out, 4);
    return 0;
}
