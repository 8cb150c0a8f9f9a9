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

// ---------------------------------------------------------------------------------------------------------------
// (4) cta_group::2: a CTA pair issues M=256 MMAs from the leader CTA; each CTA holds its 128 rows of A and HALF of B
//     (N/2 rows) in its own shared memory, so the per-SM shared-memory read per MMA drops from A 4 KB + B 8 KB to
//     4 KB + 4 KB.  (5) the 1-CTA MMA stream while other warps write shared memory at the rate a TMA-fed B ring does.
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) probe2_kernel(int mma_n, int mma_count, Result *out) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = ptx::smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t *gen = smem_raw + (base - raw);
    const uint32_t sbar = base + 65536;
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(gen + 65536 + 16);
    for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(gen)[i] = 0x3c003c00u;
    ptx::fence_proxy_async();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    if (threadIdx.x == 0) { ptx::mbar_init(sbar, 1); ptx::fence_barrier_init(); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(ptx::smem_u32((const void *)tmem_slot)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    ptx::tc_fence_before();
    cluster_sync_all();
    ptx::tc_fence_after();
    const uint32_t tmem0 = *tmem_slot;
    long long t0 = 0, t1 = 0;
    if (rank == 0 && warp == 1 && lane == 0) {
        const uint32_t idesc = ptx::umma_idesc_f16(256, mma_n);
        const uint64_t ad = ptx::umma_desc_sw128(base), bd = ptx::umma_desc_sw128(base + 16384);
        t0 = clock64();
        for (int i = 0; i < mma_count; ++i) {
            const uint32_t d = tmem0 + (uint32_t)((i & 1) * 256);
            asm volatile(
                "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
                "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(d), "l"(ad + 2 * (i & 3)),
                "l"(bd + 2 * (i & 3)), "r"(idesc), "r"(1u)
                : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(sbar),
                     "h"((uint16_t)3)
                     : "memory");
    }
    if (warp == 1 && lane == 0) {
        ptx::mbar_wait(sbar, 0);
        t1 = clock64();
        if (rank == 0) out[blockIdx.x >> 1].mma_cycles = t1 - t0;
    }
    ptx::tc_fence_before();
    cluster_sync_all();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem0), "r"(512) : "memory");
}

// 1-CTA SS MMA stream + `sts_warps` warps storing 16 B per lane to shared memory as fast as they can issue
__global__ void __launch_bounds__(576, 1) probe3_kernel(int mma_n, int mma_count, int sts_warps, int sts_gap, Result *out) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = ptx::smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t *gen = smem_raw + (base - raw);
    const uint32_t sbar = base + 131072;
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(gen + 131072 + 16);
    volatile int *stop = reinterpret_cast<volatile int *>(gen + 131072 + 32);
    for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(gen)[i] = 0x3c003c00u;
    ptx::fence_proxy_async();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { ptx::mbar_init(sbar, 1); ptx::fence_barrier_init(); *stop = 0; }
    if (warp == 0) ptx::tmem_alloc<512>(ptx::smem_u32((const void *)tmem_slot));
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem0 = *tmem_slot;
    long long bytes = 0;
    if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = ptx::umma_idesc_f16(128, mma_n);
            const uint64_t ad = ptx::umma_desc_sw128(base), bd = ptx::umma_desc_sw128(base + 16384);
            const long long t0 = clock64();
            for (int i = 0; i < mma_count; ++i)
                ptx::umma_bf16(tmem0 + (uint32_t)((i & 1) * 256), ad + 2 * (i & 3), bd + 2 * (i & 3), idesc, 1u);
            ptx::umma_commit(sbar);
            ptx::mbar_wait(sbar, 0);
            out[blockIdx.x].mma_cycles = clock64() - t0;
            *stop = 1;
        }
    } else if (warp >= 2 && warp < 2 + sts_warps) {
        // the second 64 KB of the buffer plays the B ring being refilled
        uint32_t dst = base + 65536 + (uint32_t)((warp - 2) * 4096 + lane * 16);
        int it = 0;
        const long long t0 = clock64();
        while (!*stop) {
            asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(dst + (uint32_t)((it & 7) * 512)), "r"(it) : "memory");
            ++it;
            bytes += 512;
            for (int g = 0; g < sts_gap; ++g) asm volatile("nanosleep.u32 0;" :::);
        }
        if (lane == 0 && warp == 2) out[blockIdx.x].ld_cycles = (long long)((double)bytes * sts_warps / (double)(clock64() - t0) * 1000.0);
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) ptx::tmem_dealloc<512>(tmem0);
}

static void run2(int mma_n, int mma_count) {
    const int pairs = 74;
    Result *d;
    CK(cudaMalloc(&d, sizeof(Result) * pairs));
    CK(cudaMemset(d, 0, sizeof(Result) * pairs));
    const int smem = 65536 + 64 + 1024;
    CK(cudaFuncSetAttribute(probe2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    for (int rep = 0; rep < 2; ++rep) {
        probe2_kernel<<<2 * pairs, 128, smem>>>(mma_n, mma_count, d);
        CK(cudaDeviceSynchronize());
    }
    Result h[74];
    CK(cudaMemcpy(h, d, sizeof(Result) * pairs, cudaMemcpyDeviceToHost));
    double mc = 0;
    for (int i = 0; i < pairs; ++i) mc += h[i].mma_cycles;
    mc /= pairs;
    printf("mma SS cta_group::2 M=256 N=%-3d (74 CTA pairs)   mma: %9.0f cyc  %6.1f cyc/MMA (per 128-row half: nominal %d)\n", mma_n, mc,
           mc / mma_count, mma_n / 2);
    CK(cudaFree(d));
}
static void run3(int mma_n, int mma_count, int sts_warps, int gap) {
    Result *d;
    CK(cudaMalloc(&d, sizeof(Result) * 148));
    CK(cudaMemset(d, 0, sizeof(Result) * 148));
    const int smem = 131072 + 64 + 1024;
    CK(cudaFuncSetAttribute(probe3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    for (int rep = 0; rep < 2; ++rep) {
        probe3_kernel<<<148, (2 + sts_warps) * 32, smem>>>(mma_n, mma_count, sts_warps, gap, d);
        CK(cudaDeviceSynchronize());
    }
    Result h[148];
    CK(cudaMemcpy(h, d, sizeof(Result) * 148, cudaMemcpyDeviceToHost));
    double mc = 0, bw = 0;
    for (int i = 0; i < 148; ++i) { mc += h[i].mma_cycles; bw += h[i].ld_cycles / 1000.0; }
    printf("mma SS N=%d + %2d warps of st.shared.v4 (gap %d)    mma: %6.1f cyc/MMA   shared-memory writes %6.1f B/clk/SM\n", mma_n, sts_warps, gap,
           mc / 148 / mma_count, bw / 148);
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
    // (4) CTA pairs
    run2(256, 512);
    run2(128, 512);
    // (5) shared-memory write traffic beside the 1-CTA MMA stream
    run3(256, 2048, 0, 0);
    run3(256, 2048, 1, 8);
    run3(256, 2048, 2, 4);
    run3(256, 2048, 4, 0);
    run3(256, 2048, 8, 0);
    return 0;
}
