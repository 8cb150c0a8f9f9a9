// lipvq_abi.cu -- extern "C" entry points declared in include/lipvq_b200.h.
// Host-side orchestration only: argument checks, workspace carving, kernel launches.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include <atomic>
#include "common.cuh"
#include "gemm_simt.cuh"
#include "misc_kernels.cuh"
#include "mlp_tc.cuh"
#include "mlp_enc.cuh"
#include "mlp_enc16.cuh"
#include "vq_exact.cuh"
#include "vq_tensor.cuh"
#include "small_fwd.cuh"
#include "seg_reduce.cuh"

namespace lipvq {
thread_local cudaError_t g_last_cuda_error = cudaSuccess;
long long g_launch_count = 0;

// ---- per-kernel event timing
int g_prof_on = 0;
namespace {
struct ProfRec { const char *name; cudaEvent_t beg, end; };
ProfRec *g_prof = nullptr;
int g_prof_cap = 0, g_prof_n = 0;
}  // namespace
int prof_open(const char *name, cudaStream_t st) {
    const int i = __atomic_fetch_add(&g_prof_n, 1, __ATOMIC_RELAXED);
    if (i >= g_prof_cap) return -1;
    g_prof[i].name = name;
    cudaEventRecord(g_prof[i].beg, st);
    return i;
}
void prof_close(int rec, cudaStream_t st) { cudaEventRecord(g_prof[rec].end, st); }

struct Bump {
    char *base; size_t off, cap;
    Bump(void *p, size_t c) : base((char *)p), off(0), cap(c) {}
    template <typename T> T *take(size_t n) {
        off = align_up(off, 256);
        T *r = (T *)(base + off);
        off += n * sizeof(T);
        return r;
    }
    bool ok() const { return off <= cap; }
};

static bool dims_ok(const lipvq_dims_t *d) {
    return d && d->F >= 1 && d->F <= 4096 && d->H1 >= 1 && d->H1 <= 4096 && d->H >= 1 && d->H <= 4096 &&
           d->D >= 1 && d->D <= 8192 && d->K >= 1 && d->K <= (1 << 20);
}

// LIPVQ_MLP=simt forces the fp32 CUDA-core layer kernels (A/B measurements); default = tensor cores
static bool use_mlp_tc(const lipvq_dims_t *d) {
    static const bool simt = [] { const char *e = getenv("LIPVQ_MLP"); return e && !strcmp(e, "simt"); }();
    return !simt && mlp::supported(d);
}

constexpr int kFwdMaxChunks = 8;   // row chunks of the overlapped large-batch forward

static size_t max_splitk_floats(const lipvq_dims_t *d, int64_t n) {
    size_t m = 0;
    auto up = [&](int Nout, int Kin) { size_t v = splitk_ws_floats(n, Nout, Kin); if (v > m) m = v; };
    up(d->F, d->H); up(d->H, d->H1); up(d->H1, d->D); up(d->D, d->H); up(d->H1, d->F);
    return m;
}

// per-code decoder tensors: c1,d1 [K,H1], c2,d2 [K,H], xr_code [K,F]
static size_t code_level_floats(const lipvq_dims_t *d) {
    return (size_t)d->K * (size_t)(2 * d->H1 + 2 * d->H + d->F) + 5 * 64;
}

static size_t workspace_bytes_impl(const lipvq_dims_t *d, int64_t n) {
    const size_t N = (size_t)(n > 0 ? n : 1);
    size_t fl = 0;
    // forward intermediates (fused forward keeps all of them in ws)
    fl += N * (size_t)(2 * d->H1 + 2 * d->H);        // a1,h1,a2,h2 (CUDA-core encoder only)
    fl += N * (size_t)d->D * 2;                      // z_e, z_q
    fl += N * (size_t)d->F;                          // x_recon
    fl += (size_t)d->D * d->H * 2 + d->D;            // Wn, gWn, scale
    fl += code_level_floats(d) * 2;                  // per-code decoder activations + their gradients
    // backward scratch
    fl += N * (size_t)(d->H + d->H1 + d->D);         // g_pre / g_a2 / g_a1
    fl += max_splitk_floats(d, n > d->K ? n : d->K);
    fl += 4 * kRedBlocks + 64;
    size_t bytes = fl * sizeof(float) + 64 * 256;    // alignment slack for the bump allocator
    bytes += vq_tensor_ws_bytes(n, d->D, d->K);
    bytes += N * sizeof(int64_t) + 4096;
    bytes += kFwdMaxChunks * align_up(2 * kQeBlocks * sizeof(float), 256) + 1024;
    bytes += mlp::enc_scratch_bytes(d->D) + 2048;
    bytes += small::scratch_bytes(n > 0 ? n : 1);
    if (seg::supported(d->K, n > 0 ? n : 1)) bytes += seg::plan_bytes(n > 0 ? n : 1, d->K, d->D > d->F ? d->D : d->F);
    return align_up(bytes, 256);
}
}  // namespace lipvq

using namespace lipvq;

extern "C" {

int lipvq_abi_version(void) { return LIPVQ_ABI_VERSION; }

const char *lipvq_strerror(int code) {
    switch (code) {
        case LIPVQ_OK: return "ok";
        case LIPVQ_ESHAPE: return "unsupported shape";
        case LIPVQ_EALIGN: return "null or misaligned pointer";
        case LIPVQ_EARCH: return "device is not sm_100 (B200)";
        case LIPVQ_EWORKSPACE: return "workspace too small";
        case LIPVQ_ECUDA: return "CUDA error";
        case LIPVQ_ELIMIT: return "size exceeds the fixed limit of an allocation-free entry point (static device scratch)";
        default: return "unknown error";
    }
}

int lipvq_last_cuda_error(void) { return (int)g_last_cuda_error; }
const char *lipvq_last_cuda_error_string(void) { return cudaGetErrorString(g_last_cuda_error); }

int64_t lipvq_launch_count(void) { return (int64_t)__atomic_load_n(&g_launch_count, __ATOMIC_RELAXED); }

int lipvq_check_device(void) {
    int dev = 0, major = 0;
    LIPVQ_CUDA_OK(cudaGetDevice(&dev));
    LIPVQ_CUDA_OK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    return major == 10 ? LIPVQ_OK : LIPVQ_EARCH;
}

int lipvq_profile_begin(int32_t max_records) {
    if (max_records < 1 || max_records > (1 << 20)) return LIPVQ_ESHAPE;
    if (g_prof) return LIPVQ_ESHAPE;   // already profiling
    g_prof = (ProfRec *)calloc((size_t)max_records, sizeof(ProfRec));
    if (!g_prof) return LIPVQ_EWORKSPACE;
    for (int i = 0; i < max_records; ++i) {
        LIPVQ_CUDA_OK(cudaEventCreate(&g_prof[i].beg));
        LIPVQ_CUDA_OK(cudaEventCreate(&g_prof[i].end));
    }
    g_prof_cap = max_records;
    g_prof_n = 0;
    __atomic_store_n(&g_prof_on, 1, __ATOMIC_RELEASE);
    return LIPVQ_OK;
}

int lipvq_profile_end(char *report, size_t cap) {
    if (!g_prof) return LIPVQ_ESHAPE;
    __atomic_store_n(&g_prof_on, 0, __ATOMIC_RELEASE);
    const int n = g_prof_n < g_prof_cap ? g_prof_n : g_prof_cap;
    struct Agg { const char *name; double ms; int count; } agg[64];
    int na = 0, rc = LIPVQ_OK;
    for (int i = 0; i < n; ++i) {
        float ms = 0.f;
        if (cudaEventSynchronize(g_prof[i].end) != cudaSuccess ||
            cudaEventElapsedTime(&ms, g_prof[i].beg, g_prof[i].end) != cudaSuccess) { rc = LIPVQ_ECUDA; break; }
        int j = 0;
        while (j < na && strcmp(agg[j].name, g_prof[i].name) != 0) ++j;
        if (j == na) { if (na == 64) continue; agg[na].name = g_prof[i].name; agg[na].ms = 0.0; agg[na].count = 0; ++na; }
        agg[j].ms += ms; agg[j].count += 1;
    }
    size_t off = 0;
    if (report && cap) {
        report[0] = 0;
        for (int j = 0; j < na; ++j) {
            const int w = snprintf(report + off, cap - off, "%s %d %.6f\n", agg[j].name, agg[j].count, agg[j].ms);
            if (w < 0 || (size_t)w >= cap - off) break;
            off += (size_t)w;
        }
    }
    for (int i = 0; i < g_prof_cap; ++i) { cudaEventDestroy(g_prof[i].beg); cudaEventDestroy(g_prof[i].end); }
    free(g_prof);
    g_prof = nullptr; g_prof_cap = 0; g_prof_n = 0;
    return rc;
}

int lipvq_workspace_bytes(const lipvq_dims_t *dims, int64_t n_rows, size_t *out_bytes) {
    if (!dims_ok(dims) || n_rows < 0 || !out_bytes) return LIPVQ_ESHAPE;
    *out_bytes = workspace_bytes_impl(dims, n_rows);
    return LIPVQ_OK;
}

int lipvq_encoder_fwd(const lipvq_dims_t *d, const lipvq_params_t *p, const float *x, int64_t n,
                      const lipvq_acts_t *a, void *ws, size_t ws_bytes, void *stream) {
    LIPVQ_NVTX("lipvq_encoder_fwd");
    if (!dims_ok(d) || n < 0) return LIPVQ_ESHAPE;
    if (!p || !a) return LIPVQ_EALIGN;
    if (n == 0) return LIPVQ_OK;
    const bool tc = use_mlp_tc(d);   // tensor path: a1/h1/a2 are optional outputs (kept only for backward)
    auto opt16 = [](const void *q) { return q == nullptr || aligned16(q); };
    if (!aligned16(x) || !aligned16(a->scale) || !aligned16(a->Wn) || !aligned16(a->z_e)) return LIPVQ_EALIGN;
    if (tc ? !(opt16(a->a1) && opt16(a->h1) && opt16(a->a2) && opt16(a->h2))
           : !(aligned16(a->a1) && aligned16(a->h1) && aligned16(a->a2) && aligned16(a->h2)))
        return LIPVQ_EALIGN;
    if (tc && (ws == nullptr || ws_bytes < mlp::enc_scratch_bytes(d->D) + 1024)) return LIPVQ_EWORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    if (tc) {
        static const bool split = [] { const char *e = getenv("LIPVQ_ENC"); return e && !strcmp(e, "split"); }();
        if (split && a->h2) {
            lipschitz_prep_kernel<<<(d->D + 7) / 8, 256, 0, st>>>(p->lat_W, p->lat_ci, d->D, d->H, a->scale, a->Wn);
            LIPVQ_LAUNCH_OK();
            return mlp::encoder_launch(d, p, x, n, a->a1, a->h1, a->a2, a->h2, a->Wn, a->z_e, st);
        }
        uint8_t *img = (uint8_t *)align_up((size_t)ws, 1024);
        return mlp::encoder_fused16_launch(d, p, x, n, a->a1, a->h1, a->a2, a->h2, a->scale, a->Wn, a->z_e, img, nullptr,
                                           nullptr, nullptr, 0, nullptr, st);
    }
    lipschitz_prep_kernel<<<(d->D + 7) / 8, 256, 0, st>>>(p->lat_W, p->lat_ci, d->D, d->H, a->scale, a->Wn);
    LIPVQ_LAUNCH_OK();
    LIPVQ_TRY(linear_fwd<ACT_GELU>(x, p->enc0_w, p->enc0_b, a->a1, a->h1, n, d->H1, d->F, st));
    LIPVQ_TRY(linear_fwd<ACT_GELU>(a->h1, p->enc2_w, p->enc2_b, a->a2, a->h2, n, d->H, d->H1, st));
    LIPVQ_TRY(linear_fwd<ACT_SIGMOID>(a->h2, a->Wn, p->lat_b, nullptr, a->z_e, n, d->D, d->H, st));
    return LIPVQ_OK;
}

int lipvq_linear_fwd(const float *X, const float *W, const float *b, float *Y, int64_t M, int32_t Nn,
                     int32_t K, int32_t act, void *stream) {
    LIPVQ_NVTX("lipvq_linear_fwd");
    if (M < 0 || Nn < 1 || K < 1) return LIPVQ_ESHAPE;
    if (M == 0) return LIPVQ_OK;
    if (!aligned16(X) || !aligned16(W) || !aligned16(Y)) return LIPVQ_EALIGN;
    cudaStream_t st = (cudaStream_t)stream;
    if (act == ACT_GELU) return linear_fwd<ACT_GELU>(X, W, b, nullptr, Y, M, Nn, K, st);
    if (act == ACT_SIGMOID) return linear_fwd<ACT_SIGMOID>(X, W, b, nullptr, Y, M, Nn, K, st);
    if (act == ACT_NONE) return linear_fwd<ACT_NONE>(X, W, b, nullptr, Y, M, Nn, K, st);
    if (act == ACT_RELU) return linear_fwd<ACT_RELU>(X, W, b, nullptr, Y, M, Nn, K, st);
    return LIPVQ_ESHAPE;
}

size_t lipvq_linear_bwd_ws_bytes(int64_t M, int32_t Nn, int32_t K) {
    if (M < 0 || Nn < 1 || K < 1) return 0;
    return splitk_ws_floats(M > 0 ? M : 1, Nn, K) * sizeof(float) + 1024;
}

int lipvq_linear_bwd(const float *dY, const float *X, const float *W, const float *x_mask, float *dX, float *dW,
                     float *db, int64_t M, int32_t Nn, int32_t K, void *ws, size_t ws_bytes, void *stream) {
    LIPVQ_NVTX("lipvq_linear_bwd");
    if (M < 0 || Nn < 1 || K < 1) return LIPVQ_ESHAPE;
    if (!aligned16(dY) || !aligned16(X) || !aligned16(W) || !dW) return LIPVQ_EALIGN;
    if (ws_bytes < lipvq_linear_bwd_ws_bytes(M, Nn, K) || !aligned16(ws)) return LIPVQ_EWORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    if (M == 0) {
        LIPVQ_CUDA_OK(cudaMemsetAsync(dW, 0, (size_t)Nn * K * sizeof(float), st));
        if (db) LIPVQ_CUDA_OK(cudaMemsetAsync(db, 0, (size_t)Nn * sizeof(float), st));
        return LIPVQ_OK;
    }
    LIPVQ_TRY(linear_bwd_weight(dY, X, dW, db, M, Nn, K, (float *)ws, st));
    if (dX) {
        if (!aligned16(dX)) return LIPVQ_EALIGN;
        if (x_mask) LIPVQ_TRY(linear_bwd_input<ACT_RELU>(dY, W, x_mask, dX, M, Nn, K, nullptr, nullptr, 0.f, nullptr, st));
        else LIPVQ_TRY(linear_bwd_input<ACT_NONE>(dY, W, nullptr, dX, M, Nn, K, nullptr, nullptr, 0.f, nullptr, st));
    }
    return LIPVQ_OK;
}

int lipvq_scatter_add_diff(const float *src, const int64_t *idx, const float *table, float alpha,
                           const float *grad_scale, int64_t n, int32_t D, int32_t K, float *sums, void *stream) {
    LIPVQ_NVTX("lipvq_scatter_add_diff");
    if (n < 0 || D < 1 || K < 1) return LIPVQ_ESHAPE;
    if (n == 0) return LIPVQ_OK;
    if (!aligned16(src) || !idx || !aligned16(table) || !aligned16(sums)) return LIPVQ_EALIGN;
    const int64_t blocks = (n + SC_ROWS - 1) / SC_ROWS;
    scatter_add_kernel<true><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(src, idx, n, D, K, sums, nullptr, table,
                                                                                 alpha, grad_scale);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

int lipvq_sse(const float *a, const float *b, int64_t n, double *out, void *ws, size_t ws_bytes, void *stream) {
    if (n < 0) return LIPVQ_ESHAPE;
    if (!out) return LIPVQ_EALIGN;
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) { LIPVQ_CUDA_OK(cudaMemsetAsync(out, 0, sizeof(double), st)); return LIPVQ_OK; }
    if (!a || !b) return LIPVQ_EALIGN;
    if (ws_bytes < kRedBlocks * sizeof(float) || !aligned16(ws)) return LIPVQ_EWORKSPACE;
    sse_kernel<<<kRedBlocks, 256, 0, st>>>(a, b, n, (float *)ws);
    LIPVQ_LAUNCH_OK();
    partials_to_double_kernel<<<1, 256, 0, st>>>((float *)ws, kRedBlocks, out);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

int lipvq_vqvae_loss(const double *sse_r, const double *sse_q, int64_t n_total, int32_t F, int32_t D,
                     float commitment_cost, float *out, void *stream) {
    if (n_total <= 0 || F < 1 || D < 1) return LIPVQ_ESHAPE;
    if (!sse_r || !sse_q || !out) return LIPVQ_EALIGN;
    vqvae_loss_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(sse_r, sse_q, 1.0 / ((double)n_total * F),
                                                         1.0 / ((double)n_total * D), commitment_cost, out);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

int lipvq_vqvae_recon_grad(const float *x_recon, const float *x, int64_t n_elems, float alpha,
                           const float *grad_scale, float *out, void *stream) {
    if (n_elems < 0) return LIPVQ_ESHAPE;
    if (n_elems == 0) return LIPVQ_OK;
    if (!x_recon || !x || !out) return LIPVQ_EALIGN;
    recon_grad_relu_kernel<<<kRedBlocks, 256, 0, (cudaStream_t)stream>>>(x_recon, x, alpha, grad_scale, n_elems, out);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

int lipvq_vqvae_ste_grad(float *g_zq, const float *z_e, const float *z_q, int64_t n_elems, float alpha,
                         const float *grad_scale, void *stream) {
    if (n_elems < 0) return LIPVQ_ESHAPE;
    if (n_elems == 0) return LIPVQ_OK;
    if (!g_zq || !z_e || !z_q) return LIPVQ_EALIGN;
    ste_commit_relu_kernel<<<kRedBlocks, 256, 0, (cudaStream_t)stream>>>(g_zq, z_e, z_q, alpha, grad_scale, n_elems);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

int lipvq_lipschitz_fwd(const float *h, const float *W, const float *b, const float *ci, float *scale,
                        float *Wn, float *z_e, int64_t n, int32_t D, int32_t H, void *stream) {
    LIPVQ_NVTX("lipvq_lipschitz_fwd");
    if (n < 0 || D < 1 || H < 1) return LIPVQ_ESHAPE;
    if (!aligned16(W) || !ci || !aligned16(scale) || !aligned16(Wn)) return LIPVQ_EALIGN;
    cudaStream_t st = (cudaStream_t)stream;
    lipschitz_prep_kernel<<<(D + 7) / 8, 256, 0, st>>>(W, ci, D, H, scale, Wn);
    LIPVQ_LAUNCH_OK();
    if (n == 0) return LIPVQ_OK;
    if (!aligned16(h) || !aligned16(z_e)) return LIPVQ_EALIGN;
    return linear_fwd<ACT_SIGMOID>(h, Wn, b, nullptr, z_e, n, D, H, st);
}

int lipvq_vq_argmin(const float *z, const float *E, int64_t n, int32_t D, int32_t K, int32_t mode,
                    int64_t *idx, float *best_d2, int32_t *stats, void *ws, size_t ws_bytes,
                    void *stream) {
    LIPVQ_NVTX("lipvq_vq_argmin");
    if (n < 0 || D < 1 || D > 8192 || K < 1 || K > (1 << 20)) return LIPVQ_ESHAPE;
    if (n == 0) return LIPVQ_OK;
    if (!aligned16(z) || !aligned16(E) || !idx) return LIPVQ_EALIGN;
    cudaStream_t st = (cudaStream_t)stream;
    if (mode == LIPVQ_SEARCH_TENSOR && vq_tensor_supported(D, K)) {
        if (ws_bytes < vq_tensor_ws_bytes(n, D, K) || !aligned16(ws)) return LIPVQ_EWORKSPACE;
        return vq_tensor_launch(z, E, n, D, K, idx, best_d2, stats, ws, st);
    }
    if (stats) LIPVQ_CUDA_OK(cudaMemsetAsync(stats, 0, 4 * sizeof(int32_t), st));
    return vq_exact_launch(z, E, n, D, K, nullptr, nullptr, idx, best_d2, st);
}

int lipvq_encode_search_fwd(const lipvq_dims_t *d, const lipvq_params_t *p, const float *x, int64_t n,
                            const lipvq_acts_t *a, int32_t search_mode, int32_t *stats, void *ws, size_t ws_bytes,
                            void *stream) {
    LIPVQ_NVTX("lipvq_encode_search_fwd");
    if (!dims_ok(d) || n < 0) return LIPVQ_ESHAPE;
    if (!p || !a) return LIPVQ_EALIGN;
    if (n == 0) return LIPVQ_OK;
    if (ws_bytes < mlp::enc_scratch_bytes(d->D) + 2048 + vq_tensor_ws_bytes(n, d->D, d->K) || !aligned16(ws))
        return LIPVQ_EWORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const bool fuse = use_mlp_tc(d) && search_mode == LIPVQ_SEARCH_TENSOR && vq_tensor_supported(d->D, d->K) &&
                      n <= (int64_t)0x7fffffff - tc::BM;
    if (!fuse) {
        LIPVQ_TRY(lipvq_encoder_fwd(d, p, x, n, a, ws, ws_bytes, stream));
        return lipvq_vq_argmin(a->z_e, p->codebook, n, d->D, d->K, search_mode, a->idx, nullptr, stats, ws, ws_bytes,
                               stream);
    }
    auto opt16 = [](const void *q) { return q == nullptr || aligned16(q); };
    if (!aligned16(x) || !aligned16(a->scale) || !aligned16(a->Wn) || !aligned16(a->z_e) || !a->idx ||
        !(opt16(a->a1) && opt16(a->h1) && opt16(a->a2) && opt16(a->h2)))
        return LIPVQ_EALIGN;
    // the row operand of the search (Zb, zn) is produced by the encoder's last epilogue: z_e is written once
    // and never re-read for conversion
    uint8_t *img = (uint8_t *)align_up((size_t)ws, 1024);
    void *vq_ws = (void *)align_up((size_t)(img + mlp::enc_scratch_bytes(d->D)), 1024);
    const tc::Geometry g = tc::geometry(n, d->D, d->K);
    const tc::Workspace w = tc::carve(vq_ws, g);
    LIPVQ_TRY(vq_tensor_prep(p->codebook, g, w, st));
    LIPVQ_TRY(mlp::encoder_fused16_launch(d, p, x, n, a->a1, a->h1, a->a2, a->h2, a->scale, a->Wn, a->z_e, img, w.mu,
                                          w.Zb, w.zn, g.DKz, w.counters, st));
    return vq_tensor_search(a->z_e, p->codebook, g, w, a->idx, nullptr, stats, st);
}

int lipvq_gather_loss(const float *z_e, const float *E, const int64_t *idx, int64_t n, int32_t D,
                      int32_t K, float *z_q, double *sse_q, void *ws, size_t ws_bytes, void *stream) {
    LIPVQ_NVTX("lipvq_gather_loss");
    if (n < 0 || D < 1 || K < 1) return LIPVQ_ESHAPE;
    if (!aligned16(z_e) || !aligned16(E) || !idx || !aligned16(z_q) || !sse_q) return LIPVQ_EALIGN;
    if (ws_bytes < 2 * kQeBlocks * sizeof(float) || !aligned16(ws)) return LIPVQ_EWORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) { LIPVQ_CUDA_OK(cudaMemsetAsync(sse_q, 0, sizeof(double), st)); return LIPVQ_OK; }
    float *part = (float *)ws;
    const int64_t tiles = (n + QE_ROWS - 1) / QE_ROWS;
    const int grid = (int)(tiles < kQeBlocks ? tiles : kQeBlocks);
    if ((D & 3) == 0)
        quantize_emit_kernel<true><<<grid, QE_THR, 0, st>>>(z_e, E, nullptr, idx, nullptr, n, D, 1, K, z_q, nullptr, part);
    else
        quantize_emit_kernel<false><<<grid, QE_THR, 0, st>>>(z_e, E, nullptr, idx, nullptr, n, D, 1, K, z_q, nullptr, part);
    LIPVQ_LAUNCH_OK();
    partials_to_double_kernel<<<1, 256, 0, st>>>(part, grid, sse_q);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

int lipvq_decoder_fwd(const lipvq_dims_t *d, const lipvq_params_t *p, const float *z_q, const float *x,
                      int64_t n, const lipvq_acts_t *a, double *sse_r, void *ws, size_t ws_bytes,
                      void *stream) {
    LIPVQ_NVTX("lipvq_decoder_fwd");
    if (!dims_ok(d) || n < 0) return LIPVQ_ESHAPE;
    if (!p || !a) return LIPVQ_EALIGN;
    if (n == 0) return LIPVQ_OK;
    if (!aligned16(z_q) || !aligned16(a->c1) || !aligned16(a->d1) || !aligned16(a->c2) || !aligned16(a->d2) ||
        !aligned16(a->x_recon))
        return LIPVQ_EALIGN;
    cudaStream_t st = (cudaStream_t)stream;
    LIPVQ_TRY(linear_fwd<ACT_GELU>(z_q, p->dec0_w, p->dec0_b, a->c1, a->d1, n, d->H1, d->D, st));
    LIPVQ_TRY(linear_fwd<ACT_GELU>(a->d1, p->dec2_w, p->dec2_b, a->c2, a->d2, n, d->H, d->H1, st));
    LIPVQ_TRY(linear_fwd<ACT_NONE>(a->d2, p->out_w, p->out_b, nullptr, a->x_recon, n, d->F, d->H, st));
    if (x && sse_r) {
        if (ws_bytes < kRedBlocks * sizeof(float) || !aligned16(ws)) return LIPVQ_EWORKSPACE;
        float *part = (float *)ws;
        sse_kernel<<<kRedBlocks, 256, 0, st>>>(a->x_recon, x, n * d->F, part);
        LIPVQ_LAUNCH_OK();
        partials_to_double_kernel<<<1, 256, 0, st>>>(part, kRedBlocks, sse_r);
        LIPVQ_LAUNCH_OK();
    }
    return LIPVQ_OK;
}

// scratch of lipvq_quantize_decode_fwd: [sse partials | xr_code when not supplied]
static size_t qd_ws_bytes(const lipvq_dims_t *d) {
    return kFwdMaxChunks * align_up(2 * kQeBlocks * sizeof(float), 256) +
           code_level_floats(d) * sizeof(float) + 256;
}

// R = decoder(E) for every code (v5:62-68, 75-76 applied to the K distinct values z_q can take)
static int decoder_code_table(const lipvq_dims_t *d, const lipvq_params_t *p, float *c1, float *d1, float *c2,
                              float *d2, float *R, cudaStream_t st) {
    ProfScope prof("decoder_code_table", st);
    const size_t smem = (size_t)CT_CODES * (d->D + d->H1 + d->H) * sizeof(float);
    if (smem > 200 * 1024) return LIPVQ_ESHAPE;
    if (smem > 48 * 1024)
        LIPVQ_CUDA_OK(cudaFuncSetAttribute(decoder_code_table_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)smem));
    decoder_code_table_kernel<<<(d->K + CT_CODES - 1) / CT_CODES, 512, smem, st>>>(
        p->codebook, p->dec0_w, p->dec0_b, p->dec2_w, p->dec2_b, p->out_w, p->out_b, d->K, d->D, d->H1, d->H, d->F, c1,
        d1, c2, d2, R);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

static int quantize_decode_impl(const lipvq_dims_t *d, const lipvq_params_t *p, const float *x, int64_t n,
                                int64_t n_total, const lipvq_acts_t *a, double *sse, float *loss_out, void *ws,
                                size_t ws_bytes, void *stream, bool table_ready);

int lipvq_quantize_decode_fwd(const lipvq_dims_t *d, const lipvq_params_t *p, const float *x, int64_t n,
                              int64_t n_total, const lipvq_acts_t *a, double *sse, float *loss_out, void *ws,
                              size_t ws_bytes, void *stream) {
    LIPVQ_NVTX("lipvq_quantize_decode_fwd");
    return quantize_decode_impl(d, p, x, n, n_total, a, sse, loss_out, ws, ws_bytes, stream, false);
}

// table_ready: acts->xr_code already holds decoder(E) (computed on the helper stream by lipvq_forward_fused)
static int quantize_decode_impl(const lipvq_dims_t *d, const lipvq_params_t *p, const float *x, int64_t n,
                                int64_t n_total, const lipvq_acts_t *a, double *sse, float *loss_out, void *ws,
                                size_t ws_bytes, void *stream, bool table_ready) {
    if (!dims_ok(d) || n < 0 || n_total < n || (loss_out && n_total < 1)) return LIPVQ_ESHAPE;
    if (!p || !a || !sse) return LIPVQ_EALIGN;
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) {
        LIPVQ_CUDA_OK(cudaMemsetAsync(sse, 0, 2 * sizeof(double), st));
        if (loss_out) LIPVQ_CUDA_OK(cudaMemsetAsync(loss_out, 0, 4 * sizeof(float), st));
        return LIPVQ_OK;
    }
    auto opt16 = [](const void *q) { return q == nullptr || aligned16(q); };
    if (!a->idx || !aligned16(a->z_e) || !opt16(a->z_q) || !opt16(a->x_recon) || !opt16(x) || !opt16(a->c1) ||
        !opt16(a->d1) || !opt16(a->c2) || !opt16(a->d2) || !opt16(a->xr_code))
        return LIPVQ_EALIGN;
    if (ws_bytes < qd_ws_bytes(d) || !aligned16(ws)) return LIPVQ_EWORKSPACE;
    Bump b(ws, ws_bytes);
    float *part = b.take<float>(2 * kQeBlocks);
    float *c1 = a->c1, *d1 = a->d1, *c2 = a->c2, *d2 = a->d2, *R = a->xr_code;   // code-level, kept when non-null
    if (!R) R = b.take<float>((size_t)d->K * d->F);
    if (!b.ok()) return LIPVQ_EWORKSPACE;
    // the decoder acts on z_q = E[idx] (v5:47, 75): K distinct inputs, so it runs once per CODE; the per-row
    // work is the streaming pass below (z_q / z_latent write, x_recon = R[idx], both loss sums)
    if (!table_ready) LIPVQ_TRY(decoder_code_table(d, p, c1, d1, c2, d2, R, st));
    const int64_t tiles = (n + QE_ROWS - 1) / QE_ROWS;
    const int grid = (int)(tiles < kQeBlocks ? tiles : kQeBlocks);
    ProfScope prof("quantize_emit_kernel", st);
    const float a_q = (float)(0.5 / ((double)(n_total > 0 ? n_total : 1) * d->D));
    const bool vec = (d->D & 3) == 0 && (a->g_pre == nullptr || aligned16(a->g_pre));
    if (a->g_pre) {
        if (vec)
            quantize_emit_kernel<true, true><<<grid, QE_THR, 0, st>>>(a->z_e, p->codebook, R, a->idx, x, n, d->D, d->F, d->K,
                                                                      a->z_q, a->x_recon, part, a->g_pre, a_q);
        else
            quantize_emit_kernel<false, true><<<grid, QE_THR, 0, st>>>(a->z_e, p->codebook, R, a->idx, x, n, d->D, d->F,
                                                                       d->K, a->z_q, a->x_recon, part, a->g_pre, a_q);
    } else if (vec)
        quantize_emit_kernel<true><<<grid, QE_THR, 0, st>>>(a->z_e, p->codebook, R, a->idx, x, n, d->D, d->F, d->K,
                                                            a->z_q, a->x_recon, part);
    else
        quantize_emit_kernel<false><<<grid, QE_THR, 0, st>>>(a->z_e, p->codebook, R, a->idx, x, n, d->D, d->F, d->K,
                                                             a->z_q, a->x_recon, part);
    LIPVQ_LAUNCH_OK();
    partials2_to_double_kernel<<<1, 256, 0, st>>>(part, grid, sse, x ? loss_out : nullptr,
                                                  1.0 / ((double)(n_total > 0 ? n_total : 1) * d->F),
                                                  1.0 / ((double)(n_total > 0 ? n_total : 1) * d->D));
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

int lipvq_loss_finalize(const double *sse_r, const double *sse_q, int64_t n_total, int32_t F, int32_t D,
                        float *out, void *stream) {
    if (n_total <= 0 || F < 1 || D < 1) return LIPVQ_ESHAPE;
    if (!sse_q || !out) return LIPVQ_EALIGN;
    loss_finalize_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(sse_r, sse_q, 1.0 / ((double)n_total * F),
                                                            1.0 / ((double)n_total * D), out);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

int lipvq_decoder_bwd(const lipvq_dims_t *d, const lipvq_params_t *p, const float *x, int64_t n,
                      int64_t n_total, const float *grad_scale, const lipvq_acts_t *a, const lipvq_grads_t *g,
                      void *ws, size_t ws_bytes, void *stream) {
    LIPVQ_NVTX("lipvq_decoder_bwd");
    if (!dims_ok(d) || n <= 0 || n_total < n) return LIPVQ_ESHAPE;
    if (!p || !a || !g || !aligned16(x) || !a->idx || !aligned16(a->z_e) || !aligned16(a->c1) || !aligned16(a->d1) ||
        !aligned16(a->c2) || !aligned16(a->d2) || !aligned16(a->xr_code) || !aligned16(g->codebook))
        return LIPVQ_EALIGN;
    if (ws_bytes < workspace_bytes_impl(d, n) || !aligned16(ws)) return LIPVQ_EWORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const int K = d->K;
    Bump b(ws, ws_bytes);
    float *g_R = b.take<float>((size_t)K * d->F);
    float *g_c2 = b.take<float>((size_t)K * d->H);
    float *g_c1 = b.take<float>((size_t)K * d->H1);
    float *part = b.take<float>(max_splitk_floats(d, K));
    if (!b.ok()) return LIPVQ_EWORKSPACE;
    const float a_r = (float)(2.0 / ((double)n_total * d->F));
    const float a_q = (float)(0.5 / ((double)n_total * d->D));
    const unsigned blocks = (unsigned)((n + SC_ROWS - 1) / SC_ROWS);
    // deterministic per-code sums (counting sort by code + fixed reduction tree, seg_reduce.cuh); LIPVQ_SCATTER=atomic
    // restores the fp32-atomic kernel (A/B timing), which also serves K > 8192
    static const bool force_atomic = [] { const char *e = getenv("LIPVQ_SCATTER"); return e && !strcmp(e, "atomic"); }();
    const bool det = !force_atomic && seg::supported(K, n);
    seg::Plan plan;
    if (det) {
        const int Dm = d->D > d->F ? d->D : d->F;
        void *pw = b.take<char>(seg::plan_bytes(n, K, Dm));
        if (!b.ok()) return LIPVQ_EWORKSPACE;
        plan = seg::carve(pw, n, K, Dm);
        LIPVQ_TRY(seg::build_plan(a->idx, plan, st));
    }
    // every row with code k sees the same decoder, so the reconstruction gradient is summed per code first:
    // g_R[k] = sum_{idx[n]=k} 2 (x_recon[n] - x[n]) / (n_total F), and the decoder backward runs on K rows
    if (det) {
        LIPVQ_TRY(seg::reduce(x, a->xr_code, d->F, plan, a_r, grad_scale, false, g_R, nullptr, st));
    } else {
        LIPVQ_CUDA_OK(cudaMemsetAsync(g_R, 0, (size_t)K * d->F * sizeof(float), st));
        ProfScope prof("scatter_add_kernel", st);
        scatter_add_kernel<true><<<blocks, 256, 0, st>>>(x, a->idx, n, d->F, K, g_R, nullptr, a->xr_code, a_r, grad_scale);
        LIPVQ_LAUNCH_OK();
    }
    LIPVQ_TRY(linear_bwd_weight(g_R, a->d2, g->out_w, g->out_b, K, d->F, d->H, part, st));
    LIPVQ_TRY(linear_bwd_input<ACT_GELU>(g_R, p->out_w, a->c2, g_c2, K, d->F, d->H, nullptr, nullptr, 0.f, nullptr, st));
    LIPVQ_TRY(linear_bwd_weight(g_c2, a->d1, g->dec2_w, g->dec2_b, K, d->H, d->H1, part, st));
    LIPVQ_TRY(linear_bwd_input<ACT_GELU>(g_c2, p->dec2_w, a->c1, g_c1, K, d->H, d->H1, nullptr, nullptr, 0.f, nullptr, st));
    LIPVQ_TRY(linear_bwd_weight(g_c1, p->codebook, g->dec0_w, g->dec0_b, K, d->H1, d->D, part, st));
    // codebook gradient: decoder path (row k of the code-level batch IS code k) ...
    LIPVQ_TRY(linear_bwd_input<ACT_NONE>(g_c1, p->dec0_w, nullptr, g->codebook, K, d->H1, d->D, nullptr, nullptr, 0.f,
                                         nullptr, st));
    // ... plus the codebook loss: sum_{idx[n]=k} 0.5 (E[k] - z_e[n]) / (n_total D)   (v5:81)
    if (det) {
        LIPVQ_TRY(seg::reduce(a->z_e, p->codebook, d->D, plan, a_q, grad_scale, true, g->codebook, nullptr, st));
    } else {
        ProfScope prof("scatter_add_kernel", st);
        scatter_add_kernel<true><<<blocks, 256, 0, st>>>(a->z_e, a->idx, n, d->D, K, g->codebook, nullptr, p->codebook,
                                                         a_q, grad_scale);
        LIPVQ_LAUNCH_OK();
    }
    return LIPVQ_OK;
}

// Deterministic per-code sums through the C ABI (EMA statistics, tests): see seg_reduce.cuh.
size_t lipvq_segment_sum_ws_bytes(int64_t n, int32_t D, int32_t K) {
    if (n < 1 || D < 1 || !seg::supported(K, n)) return 0;
    return seg::plan_bytes(n, K, D) + 256;
}

int lipvq_segment_sum(const float *src, const int64_t *idx, const float *table, float alpha, const float *grad_scale,
                      int64_t n, int32_t D, int32_t K, int32_t accumulate, float *sums, float *counts, void *ws,
                      size_t ws_bytes, void *stream) {
    LIPVQ_NVTX("lipvq_segment_sum");
    if (n < 0 || D < 1 || K < 1) return LIPVQ_ESHAPE;
    if (n == 0) return LIPVQ_OK;
    if (!seg::supported(K, n)) return LIPVQ_ESHAPE;
    if (!aligned16(src) || !idx || !aligned16(sums) || (table && !aligned16(table))) return LIPVQ_EALIGN;
    if (ws_bytes < lipvq_segment_sum_ws_bytes(n, D, K) || !aligned16(ws)) return LIPVQ_EWORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    const seg::Plan plan = seg::carve(ws, n, K, D);
    LIPVQ_TRY(seg::build_plan(idx, plan, st));
    return seg::reduce(src, table, D, plan, alpha, grad_scale, accumulate != 0, sums, counts, st);
}

int lipvq_scatter_add(const float *src, const int64_t *idx, int64_t n, int32_t D, int32_t K, float *sums,
                      float *counts, void *stream) {
    LIPVQ_NVTX("lipvq_scatter_add");
    if (n < 0 || D < 1 || K < 1) return LIPVQ_ESHAPE;
    if (n == 0) return LIPVQ_OK;
    if (!aligned16(src) || !idx || !aligned16(sums)) return LIPVQ_EALIGN;
    const int64_t blocks = (n + SC_ROWS - 1) / SC_ROWS;
    scatter_add_kernel<false><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(src, idx, n, D, K, sums, counts,
                                                                                  nullptr, 1.f, nullptr);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

int lipvq_encoder_bwd(const lipvq_dims_t *d, const lipvq_params_t *p, const float *x, int64_t n,
                      int64_t n_total, const float *grad_scale, const lipvq_acts_t *a, const lipvq_grads_t *g,
                      void *ws, size_t ws_bytes, void *stream) {
    LIPVQ_NVTX("lipvq_encoder_bwd");
    if (!dims_ok(d) || n <= 0 || n_total < n) return LIPVQ_ESHAPE;
    if (!p || !a || !g || !aligned16(x)) return LIPVQ_EALIGN;
    if (ws_bytes < workspace_bytes_impl(d, n) || !aligned16(ws)) return LIPVQ_EWORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    Bump b(ws, ws_bytes);
    // g_pre comes from the forward's streaming pass when the caller kept it (acts->g_pre, without dL/dloss: the
    // factor is applied in the two GEMM epilogues that consume it); otherwise it is computed here
    const bool saved = a->g_pre != nullptr;
    float *g_pre = saved ? a->g_pre : b.take<float>((size_t)n * d->D);
    float *g_a2 = b.take<float>((size_t)n * d->H);
    float *g_a1 = b.take<float>((size_t)n * d->H1);
    float *gWn = b.take<float>((size_t)d->D * d->H);
    float *part = b.take<float>(max_splitk_floats(d, n));
    if (!b.ok()) return LIPVQ_EWORKSPACE;
    const float a_q = (float)(0.5 / ((double)n_total * d->D));
    const float *late_scale = saved ? grad_scale : nullptr;
    if (!saved) {
        ProfScope prof("commit_sigmoid_bwd_kernel", st);
        commit_sigmoid_bwd_kernel<<<kRedBlocks, 256, 0, st>>>(a->z_e, a->z_q, a_q, grad_scale, n * d->D, g_pre);
        LIPVQ_LAUNCH_OK();
    }
    LIPVQ_TRY(linear_bwd_weight(g_pre, a->h2, gWn, g->lat_b, n, d->D, d->H, part, st, late_scale));
    lipschitz_bwd_kernel<<<(d->D + 7) / 8, 256, 0, st>>>(gWn, p->lat_W, p->lat_ci, a->scale, d->D, d->H,
                                                        g->lat_W, g->lat_ci);
    LIPVQ_LAUNCH_OK();
    LIPVQ_TRY(linear_bwd_input<ACT_GELU>(g_pre, a->Wn, a->a2, g_a2, n, d->D, d->H, nullptr, nullptr, 0.f, nullptr, st,
                                         late_scale));
    LIPVQ_TRY(linear_bwd_weight(g_a2, a->h1, g->enc2_w, g->enc2_b, n, d->H, d->H1, part, st));
    LIPVQ_TRY(linear_bwd_input<ACT_GELU>(g_a2, p->enc2_w, a->a1, g_a1, n, d->H, d->H1, nullptr, nullptr, 0.f, nullptr, st));
    LIPVQ_TRY(linear_bwd_weight(g_a1, x, g->enc0_w, g->enc0_b, n, d->H1, d->F, part, st));
    return LIPVQ_OK;
}


// ---- chunked forward (large no-grad batches): the three big kernels of the forward are bound by three different
// resources -- encoder: instruction issue, search: tensor pipe / TMEM reads, emit: HBM -- but the first two are
// persistent one-CTA-per-SM kernels that cannot share an SM with each other.  The emit pass can: its CTAs are small.
// So the batch CAN be cut into row chunks (whole waves of 128-row tiles), encoder + search of chunk c+1 on the caller's
// stream while the emit pass of chunk c runs on the caller-supplied helper stream (LIPVQ_FWD_CHUNKS=<n>).  MEASURED
// NEGATIVE on B200 (profiles/r02_fwd_chunks.jsonl): the persistent kernels leave the emit CTAs no shared memory to be
// resident beside them, the kernels only interleave at their tails and every chunk pays the encoder's weight-staging
// prologue again -- so the default is ONE chunk and the path stays as an A/B knob.
static int forward_chunks(const lipvq_dims_t *d, int64_t n, int32_t search_mode) {
    static const int env = [] { const char *e = getenv("LIPVQ_FWD_CHUNKS"); return e ? atoi(e) : 0; }();
    const bool fuse = use_mlp_tc(d) && search_mode == LIPVQ_SEARCH_TENSOR && vq_tensor_supported(d->D, d->K) &&
                      n <= (int64_t)0x7fffffff - tc::BM;
    if (!fuse) return 1;
    const int64_t tiles = (n + tc::BM - 1) / tc::BM;
    int c = env > 0 ? env : 1;     // measured (profiles/r02_fwd_chunks.jsonl): 0.80 / 0.85 / 0.94 / 1.05 ms at 1 / 2 / 4 / 8 chunks -> off
    if (c > kFwdMaxChunks) c = kFwdMaxChunks;
    while (c > 1 && tiles / c < 2 * kNumSMs) --c;
    return c;
}

static int forward_chunked(const lipvq_dims_t *d, const lipvq_params_t *p, const float *x, int64_t n, int64_t n_total,
                           const lipvq_acts_t *a, double *sse, float *loss_out, int32_t *stats, float *part, float *R,
                           void *es_ws, size_t es_bytes, int chunks, cudaStream_t st, cudaStream_t side,
                           cudaEvent_t e_fork, cudaEvent_t e_join) {
    if (!aligned16(a->scale) || !aligned16(a->Wn) || !aligned16(a->z_e) || !a->idx) return LIPVQ_EALIGN;
    // row ranges: whole waves of tiles per chunk
    const int64_t tiles = (n + tc::BM - 1) / tc::BM;
    int64_t waves = (tiles + kNumSMs - 1) / kNumSMs;
    const int64_t wpc = (waves + chunks - 1) / chunks;
    const int64_t rows_pc = wpc * kNumSMs * tc::BM;
    chunks = (int)((n + rows_pc - 1) / rows_pc);
    if (es_bytes < mlp::enc_scratch_bytes(d->D) + 2048 + vq_tensor_ws_bytes(rows_pc < n ? rows_pc : n, d->D, d->K))
        return LIPVQ_EWORKSPACE;
    uint8_t *img = (uint8_t *)align_up((size_t)es_ws, 1024);
    void *vq_ws = (void *)align_up((size_t)(img + mlp::enc_scratch_bytes(d->D)), 1024);
    // helper stream: the per-code decoder table first (parameters only)
    LIPVQ_CUDA_OK(cudaEventRecord(e_fork, st));
    LIPVQ_CUDA_OK(cudaStreamWaitEvent(side, e_fork, 0));
    LIPVQ_TRY(decoder_code_table(d, p, nullptr, nullptr, nullptr, nullptr, R, side));
    // codebook operand once (mu, Eb, norms and the counters sit at chunk-independent offsets of the search workspace)
    {
        const tc::Geometry g0 = tc::geometry(rows_pc < n ? rows_pc : n, d->D, d->K);
        const tc::Workspace w0 = tc::carve(vq_ws, g0);
        LIPVQ_TRY(vq_tensor_prep(p->codebook, g0, w0, st));
    }
    const size_t pstride = align_up(2 * kQeBlocks * sizeof(float), 256) / sizeof(float);
    ChunkGrids cg;
    cg.n = chunks;
    const float a_q = (float)(0.5 / ((double)(n_total > 0 ? n_total : 1) * d->D));
    (void)a_q;
    int *counters = nullptr;
    for (int c = 0; c < chunks; ++c) {
        const int64_t r0 = (int64_t)c * rows_pc, nc = (n - r0 < rows_pc ? n - r0 : rows_pc);
        const tc::Geometry g = tc::geometry(nc, d->D, d->K);
        const tc::Workspace w = tc::carve(vq_ws, g);
        counters = w.counters;
        if (c > 0) {
            vq_counters_next_chunk_kernel<<<1, 32, 0, st>>>(w.counters);
            LIPVQ_LAUNCH_OK();
        }
        LIPVQ_TRY(mlp::encoder_fused16_launch(d, p, x + r0 * d->F, nc, nullptr, nullptr, nullptr, nullptr, a->scale, a->Wn,
                                              a->z_e + r0 * d->D, img, w.mu, w.Zb, w.zn, g.DKz, w.counters, st));
        LIPVQ_TRY(vq_tensor_search(a->z_e + r0 * d->D, p->codebook, g, w, a->idx + r0, nullptr, nullptr, st));
        // emit of this chunk on the helper stream, behind its search
        LIPVQ_CUDA_OK(cudaEventRecord(e_fork, st));
        LIPVQ_CUDA_OK(cudaStreamWaitEvent(side, e_fork, 0));
        const int64_t etiles = (nc + QE_ROWS - 1) / QE_ROWS;
        const int grid = (int)(etiles < kQeBlocks ? etiles : kQeBlocks);
        cg.grid[c] = grid;
        float *pc = part + (size_t)c * pstride;
        float *zq = a->z_q ? a->z_q + r0 * d->D : nullptr;
        float *xr = a->x_recon ? a->x_recon + r0 * d->F : nullptr;
        {
            ProfScope prof("quantize_emit_kernel", side);
            if ((d->D & 3) == 0)
                quantize_emit_kernel<true><<<grid, QE_THR, 0, side>>>(a->z_e + r0 * d->D, p->codebook, R, a->idx + r0,
                                                                      x + r0 * d->F, nc, d->D, d->F, d->K, zq, xr, pc);
            else
                quantize_emit_kernel<false><<<grid, QE_THR, 0, side>>>(a->z_e + r0 * d->D, p->codebook, R, a->idx + r0,
                                                                       x + r0 * d->F, nc, d->D, d->F, d->K, zq, xr, pc);
            LIPVQ_LAUNCH_OK();
        }
    }
    LIPVQ_CUDA_OK(cudaEventRecord(e_join, side));
    LIPVQ_CUDA_OK(cudaStreamWaitEvent(st, e_join, 0));
    if (stats) {
        vq_counters_next_chunk_kernel<<<1, 32, 0, st>>>(counters);
        LIPVQ_LAUNCH_OK();
        vq_counters_report_kernel<<<1, 32, 0, st>>>(counters, stats);
        LIPVQ_LAUNCH_OK();
    }
    const double nt = (double)(n_total > 0 ? n_total : 1);
    partials2_chunks_to_double_kernel<<<1, 256, 0, st>>>(part, cg, (int)pstride, sse, loss_out, 1.0 / (nt * d->F),
                                                         1.0 / (nt * d->D));
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

int lipvq_forward_fused(const lipvq_dims_t *d, const lipvq_params_t *p, const float *x, int64_t n,
                        int64_t n_total, int32_t search_mode, float *z_latent, int64_t *idx,
                        float *x_recon, float *loss_out, int32_t *stats, void *ws, size_t ws_bytes,
                        void *stream, void *aux_stream, void *ev_fork, void *ev_join) {
    LIPVQ_NVTX("lipvq_forward_fused");
    if (!dims_ok(d) || n < 0 || n_total < n) return LIPVQ_ESHAPE;
    if (n == 0) return LIPVQ_OK;
    if (!p || !aligned16(x) || (z_latent && !aligned16(z_latent))) return LIPVQ_EALIGN;
    if (ws_bytes < workspace_bytes_impl(d, n) || !aligned16(ws)) return LIPVQ_EWORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    // the reference's real call sizes (80 / 500 rows, v5:89-93): one exact-fp32 kernel instead of 12 launches
    static const bool no_small = [] { const char *e = getenv("LIPVQ_SMALL"); return e && !strcmp(e, "0"); }();
    if (!no_small && idx && small::supported(d, n))
        return small::launch(d, p, x, n, n_total, z_latent, idx, x_recon, loss_out, stats, ws, st);
    Bump b(ws, ws_bytes);
    lipvq_acts_t a;
    const bool tc = use_mlp_tc(d);   // tensor path keeps h1/a*/c*/d* on chip: only h2 and z_e touch HBM
    a.a1 = a.h1 = a.a2 = nullptr;
    if (!tc) {
        a.a1 = b.take<float>((size_t)n * d->H1); a.h1 = b.take<float>((size_t)n * d->H1);
        a.a2 = b.take<float>((size_t)n * d->H);
    }
    a.h2 = tc ? nullptr : b.take<float>((size_t)n * d->H);
    a.scale = b.take<float>(d->D); a.Wn = b.take<float>((size_t)d->D * d->H);
    a.z_e = b.take<float>((size_t)n * d->D);
    a.idx = idx ? idx : b.take<int64_t>(n);
    a.z_q = z_latent;
    a.g_pre = nullptr;
    a.c1 = a.d1 = a.c2 = a.d2 = a.xr_code = nullptr;   // code-level decoder tensors come from the scratch below
    a.x_recon = x_recon;                               // optional output
    double *sse = b.take<double>(2);
    const size_t part_bytes = qd_ws_bytes(d);
    float *part = (float *)b.take<char>(part_bytes);
    if (!b.ok()) return LIPVQ_EWORKSPACE;
    // the rest of the workspace serves the encoder image + the search (checked against their own needs there)
    void *es_ws = (void *)align_up((size_t)(b.base + b.off), 256);
    const size_t es_bytes = ws_bytes - (size_t)((char *)es_ws - (char *)ws);
    // The per-code decoder table depends on the parameters only.  With a caller-supplied helper stream and two
    // events it runs beside the encoder / search kernels (forked here, joined before the emit pass) instead of
    // between them; the library itself owns no stream and no event (re-entrant, graph-capturable).
    cudaStream_t side = (cudaStream_t)aux_stream;
    cudaEvent_t e_fork = (cudaEvent_t)ev_fork, e_join = (cudaEvent_t)ev_join;
    bool forked = false;
    if (side && e_fork && e_join) {
        float *R = (float *)align_up((size_t)part + kFwdMaxChunks * align_up(2 * kQeBlocks * sizeof(float), 256), 256);   // qd scratch tail
        // large batches on the fused tensor path: row chunks, the HBM-bound emit pass of chunk c on the helper stream
        // under the issue-bound encoder / tensor-bound search kernels of chunk c+1
        const int chunks = forward_chunks(d, n, search_mode);
        if (chunks > 1)
            return forward_chunked(d, p, x, n, n_total, &a, sse, loss_out, stats, part, R, es_ws, es_bytes, chunks, st, side,
                                   e_fork, e_join);
        LIPVQ_CUDA_OK(cudaEventRecord(e_fork, st));
        LIPVQ_CUDA_OK(cudaStreamWaitEvent(side, e_fork, 0));
        LIPVQ_TRY(decoder_code_table(d, p, nullptr, nullptr, nullptr, nullptr, R, side));
        LIPVQ_CUDA_OK(cudaEventRecord(e_join, side));
        forked = true;
        a.xr_code = R;
    }
    int rc = lipvq_encode_search_fwd(d, p, x, n, &a, search_mode, stats, es_ws, es_bytes, stream);
    if (forked && cudaStreamWaitEvent(st, e_join, 0) != cudaSuccess && rc == LIPVQ_OK) rc = LIPVQ_ECUDA;   // always re-join
    LIPVQ_TRY(rc);
    return quantize_decode_impl(d, p, x, n, n_total, &a, sse, loss_out, part, part_bytes, stream, forked);
}

int lipvq_forward_small(const lipvq_dims_t *d, const lipvq_params_t *p, const float *x, int64_t n, int64_t n_total,
                        float *z_latent, int64_t *idx, float *x_recon, float *loss_out, int32_t *stats, void *ws,
                        size_t ws_bytes, void *stream) {
    LIPVQ_NVTX("lipvq_forward_small");
    if (!dims_ok(d) || n < 0 || n_total < n || n > (1 << 16) || small::smem_bytes(d) > 200 * 1024) return LIPVQ_ESHAPE;
    if (n == 0) return LIPVQ_OK;
    if (!p || !aligned16(x) || (z_latent && !aligned16(z_latent)) || !idx) return LIPVQ_EALIGN;
    if (ws_bytes < small::scratch_bytes(n) || !aligned16(ws)) return LIPVQ_EWORKSPACE;
    return small::launch(d, p, x, n, n_total, z_latent, idx, x_recon, loss_out, stats, ws, (cudaStream_t)stream);
}

// ---- allocation-free variant of lipvq_forward_small (no counterpart in the reference).
// The scratch lives in a file-scope __device__ array: every GPU that loads the library holds these
// 64 KB until exit, every call on a device shares them (calls that overlap on the GPU race), and the array keeps the
// last call's contents -- small::launch clears the part that must start at zero (best[] + the ticket counters) with
// a memset node on the caller's stream before the kernel.  In host code the array's NAME is only a host-side stand-in:
// the device address comes from cudaGetSymbolAddress, once per device, with that device current.
constexpr int kStaticSmallMaxRows = LIPVQ_FORWARD_SMALL_STATIC_MAX_ROWS;
constexpr size_t kStaticSmallBytes = 64 * 1024;
static __device__ __align__(256) unsigned char g_small_static_scratch[kStaticSmallBytes];
static_assert(LIPVQ_FORWARD_SMALL_STATIC_MAX_ROWS == 4096, "header and scratch size go together");

static int small_static_scratch_ptr(void **out) {
    constexpr int kMaxDev = 64;
    static std::atomic<void *> ptrs[kMaxDev];
    int dev = -1;
    LIPVQ_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDev) return LIPVQ_ESHAPE;
    void *ptr = ptrs[dev].load(std::memory_order_acquire);
    if (!ptr) {
        LIPVQ_CUDA_OK(cudaGetSymbolAddress(&ptr, g_small_static_scratch));   // this device's copy of the array
        if (!ptr) return LIPVQ_ECUDA;
        ptrs[dev].store(ptr, std::memory_order_release);
    }
    *out = ptr;
    return LIPVQ_OK;
}

int lipvq_forward_small_static(const lipvq_dims_t *d, const lipvq_params_t *p, const float *x, int64_t n, int64_t n_total,
                               float *z_latent, int64_t *idx, float *x_recon, float *loss_out, int32_t *stats,
                               void *stream) {
    LIPVQ_NVTX("lipvq_forward_small_static");
    if (!dims_ok(d) || n < 0 || n_total < n || small::smem_bytes(d) > 200 * 1024) return LIPVQ_ESHAPE;
    if (n > kStaticSmallMaxRows || small::scratch_bytes(n) > kStaticSmallBytes) return LIPVQ_ELIMIT;
    if (n == 0) return LIPVQ_OK;
    if (!p || !aligned16(x) || (z_latent && !aligned16(z_latent)) || !idx) return LIPVQ_EALIGN;
    void *scratch = nullptr;
    LIPVQ_TRY(small_static_scratch_ptr(&scratch));
    return small::launch(d, p, x, n, n_total, z_latent, idx, x_recon, loss_out, stats, scratch, (cudaStream_t)stream);
}

int lipvq_adamw_step(float *param, const float *grad, float *exp_avg, float *exp_avg_sq, int64_t n,
                     int64_t step, float lr, float beta1, float beta2, float eps, float weight_decay,
                     void *stream) {
    LIPVQ_NVTX("lipvq_adamw_step");
    if (n < 0 || step < 1) return LIPVQ_ESHAPE;
    if (n == 0) return LIPVQ_OK;
    if (!param || !grad || !exp_avg || !exp_avg_sq) return LIPVQ_EALIGN;
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const double bc2 = 1.0 - pow((double)beta2, (double)step);
    int64_t blocks = (n + 255) / 256;
    if (blocks > 8 * kNumSMs) blocks = 8 * kNumSMs;
    adamw_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(param, grad, exp_avg, exp_avg_sq, n, lr,
                                                                   beta1, beta2, eps, weight_decay,
                                                                   (float)bc1, (float)sqrt(bc2));
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

int lipvq_ema_update(float *E, float *n_ema, float *m_ema, const float *counts, const float *sums,
                     int32_t K, int32_t D, float gamma, float eps, void *ws, size_t ws_bytes, void *stream) {
    LIPVQ_NVTX("lipvq_ema_update");
    if (K < 1 || D < 1) return LIPVQ_ESHAPE;
    if (!E || !n_ema || !m_ema || !counts || !sums) return LIPVQ_EALIGN;
    if (ws_bytes < 64 * sizeof(float) + 256 + sizeof(double) || !aligned16(ws)) return LIPVQ_EWORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    float *part = (float *)ws;
    double *tot = (double *)((char *)ws + 256);
    ema_counts_kernel<<<64, 256, 0, st>>>(n_ema, counts, K, gamma, part);
    LIPVQ_LAUNCH_OK();
    partials_to_double_kernel<<<1, 256, 0, st>>>(part, 64, tot);
    LIPVQ_LAUNCH_OK();
    ema_apply_kernel<<<kRedBlocks, 256, 0, st>>>(E, n_ema, m_ema, sums, tot, K, D, gamma, eps);
    LIPVQ_LAUNCH_OK();
    return LIPVQ_OK;
}

}  // extern "C"
