// tmem_probe.cu -- microbenchmarks behind the design constants quoted in DESIGN.md / vq_tensor.cuh:
//   (1) the tcgen05.ld (TMEM -> registers) read floor per SM, by load shape and number of reading warps;
//   (2) the issue-to-completion rate of back-to-back tcgen05.mma kind::f16, M=128, by N, with the A operand in
//       shared memory (SS) or in tensor memory (TS);
//   (3) both at once (an MMA stream under a full-rate accumulator read-out), the regime of the search epilogue.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/_bin/tmem_probe tools/tmem_probe.cu
// Run on a B200:  tools/_bin/tmem_probe   (prints one line per case; values in SM clock cycles, clock64)
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#include "../lipvq_vae_b200/csrc/ptx_sm100.cuh"

using namespace lipvq;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ void ld_x64(uint32_t taddr, uint32_t (&r)[64]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,"
        "%32,%33,%34,%35,%36,%37,%38,%39,%40,%41,%42,%43,%44,%45,%46,%47,%48,%49,%50,%51,%52,%53,%54,%55,%56,%57,%58,%59,%60,%61,%62,%63}, [%64];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]), "=r"(r[32]),
          "=r"(r[33]), "=r"(r[34]), "=r"(r[35]), "=r"(r[36]), "=r"(r[37]), "=r"(r[38]), "=r"(r[39]), "=r"(r[40]),
          "=r"(r[41]), "=r"(r[42]), "=r"(r[43]), "=r"(r[44]), "=r"(r[45]), "=r"(r[46]), "=r"(r[47]), "=r"(r[48]),
          "=r"(r[49]), "=r"(r[50]), "=r"(r[51]), "=r"(r[52]), "=r"(r[53]), "=r"(r[54]), "=r"(r[55]), "=r"(r[56]),
          "=r"(r[57]), "=r"(r[58]), "=r"(r[59]), "=r"(r[60]), "=r"(r[61]), "=r"(r[62]), "=r"(r[63])
        : "r"(taddr)
        : "memory");
}
// 16 lanes x 256 bits per repetition: thread t of the warp receives 4 registers per repetition
__device__ __forceinline__ void ld_16x256b_x8(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.16x256b.x8.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n}\n" ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc),
        "r"(acc)
        : "memory");
}

// mode bits: LD_SHAPE (0 none, 1 x16 wait each, 2 x32 wait each, 3 two x32 then wait, 4 x64 wait each, 5 16x256b.x8)
struct Result { long long ld_cycles, mma_cycles; unsigned sink; };

__global__ void __launch_bounds__(576, 1)
probe_kernel(int ld_shape, int ld_warps, int ld_iters, int mma_mode /*0 none 1 SS 2 TS*/, int mma_n, int mma_count,
             Result *out) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = ptx::smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t *gen = smem_raw + (base - raw);
    // [0, 64 KB): operand bytes (zero) | then barrier + tmem slot
    const uint32_t sbar = base + 65536;
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(gen + 65536 + 16);
    for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(gen)[i] = 0x3c003c00u;  // fp16 1.0
    ptx::fence_proxy_async();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { ptx::mbar_init(sbar, 1); ptx::fence_barrier_init(); }
    if (warp == 0) ptx::tmem_alloc<512>(ptx::smem_u32((const void *)tmem_slot));
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem0 = *tmem_slot;
    const int mma_warp = ld_warps;   // the warp after the readers issues the MMAs
    unsigned sink = 0;
    long long t0 = 0, t1 = 0;
    if (warp < ld_warps && ld_shape) {
        const uint32_t lane_addr = tmem0 + ((uint32_t)((warp & 3) * 32) << 16);
        const int col_slices = ld_warps / 4 > 0 ? ld_warps / 4 : 1;
        const int cols_per = 512 / col_slices, c_lo = (warp >> 2) * cols_per;
        __syncwarp();
        t0 = clock64();
        for (int it = 0; it < ld_iters; ++it) {
            if (ld_shape == 1) {
                for (int c = 0; c < cols_per; c += 16) {
                    uint32_t r[16];
                    ptx::tmem_ld_32x32b_x16(lane_addr + c_lo + c, r);
                    ptx::tmem_ld_wait();
                    sink ^= r[0] ^ r[15];
                }
            } else if (ld_shape == 2) {
                for (int c = 0; c < cols_per; c += 32) {
                    uint32_t r[32];
                    ptx::tmem_ld_32x32b_x32(lane_addr + c_lo + c, r);
                    ptx::tmem_ld_wait();
                    sink ^= r[0] ^ r[31];
                }
            } else if (ld_shape == 3) {
                for (int c = 0; c < cols_per; c += 64) {
                    uint32_t r[32], q[32];
                    ptx::tmem_ld_32x32b_x32(lane_addr + c_lo + c, r);
                    ptx::tmem_ld_32x32b_x32(lane_addr + c_lo + c + 32, q);
                    ptx::tmem_ld_wait();
                    sink ^= r[0] ^ q[31];
                }
            } else if (ld_shape == 4) {
                for (int c = 0; c < cols_per; c += 64) {
                    uint32_t r[64];
                    ld_x64(lane_addr + c_lo + c, r);
                    ptx::tmem_ld_wait();
                    sink ^= r[0] ^ r[63];
                }
            } else if (ld_shape == 5) {
                // 16x256b.x8: 16 lanes x 64 columns per instruction; two instructions cover the warp's 32 lanes
                for (int c = 0; c < cols_per; c += 64) {
                    uint32_t r[32], q[32];
                    ld_16x256b_x8(lane_addr + c_lo + c, r);
                    ld_16x256b_x8(lane_addr + (16u << 16) + c_lo + c, q);
                    ptx::tmem_ld_wait();
                    sink ^= r[0] ^ q[31];
                }
            }
        }
        t1 = clock64();
        if (lane == 0 && warp == 0) out[blockIdx.x].ld_cycles = t1 - t0;
    } else if (warp == mma_warp && mma_mode) {
        if (lane == 0) {
            const uint32_t idesc = ptx::umma_idesc_f16(128, mma_n);
            const uint64_t ad = ptx::umma_desc_sw128(base), bd = ptx::umma_desc_sw128(base + 16384);
            t0 = clock64();
            for (int i = 0; i < mma_count; ++i) {
                const uint32_t d = tmem0 + (mma_mode == 1 ? (uint32_t)((i & 1) * 256) : 0u);
                if (mma_mode == 1) ptx::umma_bf16(d, ad + 2 * (i & 3), bd + 2 * (i & 3), idesc, 1u);
                else umma_f16_ts(d, tmem0 + 480u, bd + 2 * (i & 3), idesc, 1u);
            }
            ptx::umma_commit(sbar);
            ptx::mbar_wait(sbar, 0);
            t1 = clock64();
            out[blockIdx.x].mma_cycles = t1 - t0;
        }
    }
    if (sink == 0x12345u) out[blockIdx.x].sink = sink;
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc<512>(tmem0);
}

static void run(const char *name, int ld_shape, int ld_warps, int ld_iters, int mma_mode, int mma_n, int mma_count, int grid) {
    Result *d;
    CK(cudaMalloc(&d, sizeof(Result) * grid));
    CK(cudaMemset(d, 0, sizeof(Result) * grid));
    const int threads = (ld_warps + 1) * 32;
    const int smem = 65536 + 64 + 1024;
    CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    for (int rep = 0; rep < 2; ++rep) {
        probe_kernel<<<grid, threads, smem>>>(ld_shape, ld_warps, ld_iters, mma_mode, mma_n, mma_count, d);
        CK(cudaDeviceSynchronize());
    }
    Result h[148];
    CK(cudaMemcpy(h, d, sizeof(Result) * grid, cudaMemcpyDeviceToHost));
    double ldc = 0, mc = 0;
    for (int i = 0; i < grid; ++i) { ldc += h[i].ld_cycles; mc += h[i].mma_cycles; }
    ldc /= grid; mc /= grid;
    printf("%-44s", name);
    if (ld_shape) {
        // every reading warp covers (512 / col_slices) columns x 32 lanes x 4 B per iteration; all warps together read
        // the whole 256 KB of TMEM (4 quarters x col_slices slices) once per iteration when ld_warps >= 4
        const double bytes = (ld_warps >= 4 ? 262144.0 : 65536.0 * ld_warps) * ld_iters;
        printf(" ld: %9.0f cyc  %7.1f B/clk/SM  (%.0f cyc per 128x256 fp32 accumulator)", ldc, bytes / ldc, 131072.0 / (bytes / ldc));
    }
    if (mma_mode) printf(" mma: %9.0f cyc  %6.1f cyc/MMA (M=128 N=%d K=16; nominal %d)", mc, mc / mma_count, mma_n, mma_n / 2);
    printf("\n");
    CK(cudaFree(d));
}

int main() {
    int dev = 0;
    cudaDeviceProp p;
    CK(cudaGetDeviceProperties(&p, dev));
    printf("device %s, %d SMs, clock %d kHz\n", p.name, p.multiProcessorCount, p.clockRate);
    const int G = 148;
    // (1) read floor
    run("ld 32x32b.x16 wait-each,  4 warps", 1, 4, 64, 0, 0, 0, G);
    run("ld 32x32b.x32 wait-each,  4 warps", 2, 4, 64, 0, 0, 0, G);
    run("ld 32x32b.x32 wait-each,  8 warps", 2, 8, 64, 0, 0, 0, G);
    run("ld 32x32b.x32 wait-each, 16 warps", 2, 16, 64, 0, 0, 0, G);
    run("ld 32x32b.x32 two-in-flight, 16 warps", 3, 16, 64, 0, 0, 0, G);
    run("ld 32x32b.x64 wait-each,  4 warps", 4, 4, 64, 0, 0, 0, G);
    run("ld 32x32b.x64 wait-each,  8 warps", 4, 8, 64, 0, 0, 0, G);
    run("ld 32x32b.x64 wait-each, 16 warps", 4, 16, 64, 0, 0, 0, G);
    run("ld 16x256b.x8 x2 wait-each, 8 warps", 5, 8, 64, 0, 0, 0, G);
    run("ld 16x256b.x8 x2 wait-each, 16 warps", 5, 16, 64, 0, 0, 0, G);
    run("ld 32x32b.x32 wait-each, 16 warps, 1 SM", 2, 16, 64, 0, 0, 0, 1);
    // (2) MMA rate
    const int ns[] = {64, 104, 128, 208, 256};
    for (int n : ns) { char nm[64]; snprintf(nm, 64, "mma SS N=%d", n); run(nm, 0, 4, 0, 1, n, 512, G); }
    for (int n : ns) { char nm[64]; snprintf(nm, 64, "mma TS N=%d", n); run(nm, 0, 4, 0, 2, n, 512, G); }
    // (3) both: 16 reading warps + an MMA stream
    run("ld x32 16 warps + mma SS N=256", 2, 16, 64, 1, 256, 2048, G);
    run("ld x32 16 warps + mma SS N=128", 2, 16, 64, 1, 128, 4096, G);
    run("ld x32  8 warps + mma SS N=256", 2, 8, 64, 1, 256, 2048, G);
    run("ld x32 16 warps + mma TS N=256", 2, 16, 64, 2, 256, 2048, G);
    return 0;
}
