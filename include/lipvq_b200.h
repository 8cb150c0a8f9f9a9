/*
 * lipvq_b200.h -- C ABI of the B200-native LipVQ-VAE action-tokenizer hot path.
 *
 * The reference has no FFI: its seam is the Python class `LLFQVAE_V4`
 * (robomimic/models/vq_vae/backbone_lfqvae_v5.py:51-84) resolved at
 * robomimic/models/obs_nets.py:1225-1227.  Every entry point below replaces one torch-op
 * sequence of that class (file:line cited per function); the Python mirror of the class
 * (lipvq_vae_b200/tokenizer.py) is the only caller.
 *
 * Conventions
 *   - all data pointers are DEVICE pointers into caller-owned, 16-byte aligned, contiguous
 *     row-major fp32 buffers (indices: int64); nothing is allocated or kept across calls;
 *   - `stream` is a cudaStream_t passed as void*; every call only enqueues work on it;
 *   - scratch memory comes from the caller: `ws` of at least lipvq_workspace_bytes(...) bytes;
 *   - return 0 on success, a negative LIPVQ_E* code otherwise; never throws;
 *   - re-entrant across streams as long as the workspaces differ.
 */
#ifndef LIPVQ_B200_H
#define LIPVQ_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LIPVQ_ABI_VERSION 6

enum {
    LIPVQ_OK = 0,
    LIPVQ_ESHAPE = -1,     /* a dimension is out of the supported range                    */
    LIPVQ_EALIGN = -2,     /* a pointer is null or not 16-byte aligned                      */
    LIPVQ_EARCH = -3,      /* the current device is not sm_100                              */
    LIPVQ_EWORKSPACE = -4, /* ws_bytes is smaller than lipvq_workspace_bytes(...)           */
    LIPVQ_ECUDA = -5,      /* a CUDA runtime/driver call failed (see lipvq_last_cuda_error) */
    LIPVQ_ELIMIT = -6      /* a size exceeds the fixed limit of an allocation-free entry point */
};

/* search modes for lipvq_vq_argmin */
enum {
    LIPVQ_SEARCH_EXACT = 0,  /* fp32 direct-form scan sum_i (z_i-e_ki)^2 on CUDA cores        */
    LIPVQ_SEARCH_TENSOR = 1  /* bf16 tcgen05 GEMM filter + fp32 direct-form rescore (default) */
};

/* dims of the 14 parameter tensors built at v5:52-68 */
typedef struct {
    int32_t F;  /* feature_dim (12 at obs_nets.py:2411)                */
    int32_t H1; /* first hidden width, fixed 64 in the reference       */
    int32_t H;  /* hidden_dim (128)                                    */
    int32_t D;  /* latent_dim                                          */
    int32_t K;  /* num_codes                                           */
} lipvq_dims_t;

/* the 14 tensors, state_dict order (v5:52-68) */
typedef struct {
    const float *enc0_w, *enc0_b; /* encoder.0.{weight[H1,F],bias[H1]}   */
    const float *enc2_w, *enc2_b; /* encoder.2.{weight[H,H1],bias[H]}    */
    const float *lat_W, *lat_b, *lat_ci; /* to_latent.{W[D,H],b[D],ci[D]}   */
    const float *codebook;        /* quantizer.codebook[K,D]             */
    const float *dec0_w, *dec0_b; /* decoder.0.{weight[H1,D],bias[H1]}   */
    const float *dec2_w, *dec2_b; /* decoder.2.{weight[H,H1],bias[H]}    */
    const float *out_w, *out_b;   /* to_output.{weight[F,H],bias[F]}     */
} lipvq_params_t;

/* gradient outputs, same layout; every tensor is OVERWRITTEN (not accumulated) */
typedef struct {
    float *enc0_w, *enc0_b, *enc2_w, *enc2_b;
    float *lat_W, *lat_b, *lat_ci;
    float *codebook;
    float *dec0_w, *dec0_b, *dec2_w, *dec2_b;
    float *out_w, *out_b;
} lipvq_grads_t;

/* activations kept between forward and backward (fp32).  The decoder acts on z_q = E[idx] (v5:47, 75), which
 * takes only K distinct values, so its activations are kept PER CODE ([K, width]), not per row. */
typedef struct {
    float *a1, *h1; /* encoder.0 pre-activation / GELU output   [N,H1]            */
    float *a2, *h2; /* encoder.2 pre-activation / GELU output   [N,H]             */
    float *scale;   /* Lipschitz row scale min(1, softplus(ci)/sum|W|) [D]  v5:6-12 */
    float *Wn;      /* normalised weight W*scale [D,H]                            */
    float *z_e;     /* sigmoid(h2 Wn^T + b)  [N,D]   v5:22-24                      */
    int64_t *idx;   /* nearest code per row [N]      v5:46                        */
    float *z_q;     /* codebook[idx] [N,D]           v5:47                        */
    float *c1, *d1; /* decoder.0 pre-activation / GELU output of E  [K,H1]        */
    float *c2, *d2; /* decoder.2 pre-activation / GELU output of E  [K,H]         */
    float *x_recon; /* to_output(...) [N,F]          v5:76  (= xr_code[idx])       */
    float *xr_code; /* to_output(decoder(E)) [K,F]: reconstruction of every code  */
    float *g_pre;   /* nullable [N,D]: 0.5 (z_e - z_q) z_e (1 - z_e) / (n_total D), the commitment gradient through
                     * the sigmoid up to dL/dloss (v5:80).  Written by lipvq_quantize_decode_fwd while the row is in
                     * registers; lipvq_encoder_bwd then skips its own pass over z_e and z_q.               */
} lipvq_acts_t;

int lipvq_abi_version(void);
const char *lipvq_strerror(int code);
/* last cudaError_t seen by this library on the calling thread (0 = none), and its string */
int lipvq_last_cuda_error(void);
const char *lipvq_last_cuda_error_string(void);
/* 0 if the current device is compute capability 10.x, LIPVQ_EARCH otherwise */
int lipvq_check_device(void);
/* number of kernels this library has launched in this process (bench.py's gpu_launches) */
int64_t lipvq_launch_count(void);

/* Per-kernel timing for measurement (bench.py's roofline): between begin and end every kernel launch of this
 * library is bracketed by CUDA events on its launching stream.  end() synchronises those events, writes one
 * line per kernel name, "name count total_ms\n", into report (NUL-terminated, truncated to cap) and frees
 * the events.  Not re-entrant; off by default. */
int lipvq_profile_begin(int32_t max_records);
int lipvq_profile_end(char *report, size_t cap);

/* scratch bytes needed by any call below for n_rows rows with these dims */
int lipvq_workspace_bytes(const lipvq_dims_t *dims, int64_t n_rows, size_t *out_bytes);

/* Encoder: h=GELU(GELU(x W0^T+b0) W1^T+b1); z_e=sigmoid(h (W*scale)^T + b).
 * Replaces v5:54-60 + normalization v5:6-12 + LipschitzMLP.forward v5:22-24 (called v5:71-72).
 * acts: a1,h1,a2,h2,scale,Wn,z_e are written. */
int lipvq_encoder_fwd(const lipvq_dims_t *dims, const lipvq_params_t *p, const float *x,
                      int64_t n_rows, const lipvq_acts_t *acts, void *ws, size_t ws_bytes,
                      void *stream);

/* One fused linear layer Y[M,N] = act(X[M,K] W[N,K]^T + b[N]) (nn.Linear + nn.GELU of v5:54-68 when a
 * sub-module is called on its own); act: 0 none, 1 exact-erf GELU, 2 sigmoid, 5 ReLU (the classic VQVAE's layers,
 * robomimic/models/vq_vae/backbone.py:17-32).  b nullable. */
int lipvq_linear_fwd(const float *X, const float *W, const float *b, float *Y, int64_t M, int32_t N,
                     int32_t K, int32_t act, void *stream);

/* LipschitzMLP.forward alone (v5:6-12, 22-24): scale[D], Wn[D,H] and z_e[N,D] are written. */
int lipvq_lipschitz_fwd(const float *h, const float *W, const float *b, const float *ci, float *scale,
                        float *Wn, float *z_e, int64_t n_rows, int32_t D, int32_t H, void *stream);

/* Nearest codeword: idx[n] = argmin_k ||z[n]-E[k]||_2, first minimal index on ties
 * (LFQQuantizer.forward v5:37-46; the sign mask v5:39-40 is +-1 and vanishes under the norm).
 * best_d2 (nullable) receives the fp32 direct-form squared distance of the winner.
 * stats (nullable, int32[4]): [0] rows sent to the exact rescan, [1] candidate codes rescored. */
int lipvq_vq_argmin(const float *z, const float *E, int64_t n_rows, int32_t D, int32_t K,
                    int32_t mode, int64_t *idx, float *best_d2, int32_t *stats, void *ws,
                    size_t ws_bytes, void *stream);

/* Encoder + nearest-codeword search in one call (v5:71-73): like lipvq_encoder_fwd followed by
 * lipvq_vq_argmin(acts->z_e -> acts->idx), but on the tensor path the encoder's last epilogue also emits
 * the search's row operand, so z_e is written once and never re-read for conversion.
 * acts: scale, Wn, z_e, idx are written; a1, h1, a2, h2 are written when non-null (kept for backward). */
int lipvq_encode_search_fwd(const lipvq_dims_t *dims, const lipvq_params_t *p, const float *x,
                            int64_t n_rows, const lipvq_acts_t *acts, int32_t search_mode,
                            int32_t *stats, void *ws, size_t ws_bytes, void *stream);

/* z_q[n] = E[idx[n]] (v5:47); sse_q = sum (z_q - z_e)^2 (the value shared by the commitment and
 * codebook losses, v5:80-81).  sse_q: device double[1], overwritten. */
int lipvq_gather_loss(const float *z_e, const float *E, const int64_t *idx, int64_t n_rows,
                      int32_t D, int32_t K, float *z_q, double *sse_q, void *ws, size_t ws_bytes,
                      void *stream);

/* Decoder: x_recon = GELU(GELU(z_q Wd0^T+bd0) Wd1^T+bd1) Wo^T + bo (v5:62-68, 75-76);
 * sse_r = sum (x_recon - x)^2 (v5:79).  x may be null (then sse_r is not written).
 * acts: c1,d1,c2,d2,x_recon are written. */
int lipvq_decoder_fwd(const lipvq_dims_t *dims, const lipvq_params_t *p, const float *z_q,
                      const float *x, int64_t n_rows, const lipvq_acts_t *acts, double *sse_r,
                      void *ws, size_t ws_bytes, void *stream);

/* Gather + decoder + both loss sums (v5:47, 74-81).  The decoder runs once per CODE (xr_code = decoder(E),
 * [K,F]); one streaming pass over the rows then writes z_q[n] = E[idx[n]] to acts->z_q (nullable; this IS z_latent,
 * v5:74), x_recon[n] = xr_code[idx[n]] to acts->x_recon (nullable), and sse[0] = sum (x_recon-x)^2,
 * sse[1] = sum (z_q-z_e)^2 (device double[2], overwritten).  acts: idx, z_e are read; z_q is written;
 * c1,d1,c2,d2,xr_code ([K,.], code level) and g_pre ([N,D]) are written when non-null (kept for backward).
 * x may be null (sse[0] = 0).  loss_out (nullable, needs x): device float[4] = {loss, recon, commit, codebook}
 * as lipvq_loss_finalize would write them for n_total global rows (v5:79-83), from the same reduction. */
int lipvq_quantize_decode_fwd(const lipvq_dims_t *dims, const lipvq_params_t *p, const float *x,
                              int64_t n_rows, int64_t n_total, const lipvq_acts_t *acts, double *sse,
                              float *loss_out, void *ws, size_t ws_bytes, void *stream);

/* loss = sse_r/(n_total*F) + 0.25*sse_q/(n_total*D) + 0.25*sse_q/(n_total*D)  (v5:79-83).
 * out: device float[4] = {loss, recon, commit, codebook}. */
int lipvq_loss_finalize(const double *sse_r, const double *sse_q, int64_t n_total, int32_t F,
                        int32_t D, float *out, void *stream);

/* Backward of the decoder + losses w.r.t. decoder/to_output parameters and the codebook
 * (autograd of v5:75-83 triggered at robomimic/algo/icl.py:969), evaluated per code: the reconstruction
 * gradient is first summed over the rows of each code, g_R[k] = sum_{idx[n]=k} 2 (x_recon[n]-x[n])/(n_total F),
 * the decoder backward then runs on the K code-level rows (acts->c1,d1,c2,d2,xr_code), and
 *   g->codebook[k] = dRecon/dE[k] + sum_{idx[n]=k} 0.5*(E[k]-z_e[n])/(n_total*D)        (complete gradient).
 * Writes g->{dec0_*, dec2_*, out_*, codebook}.  grad_scale: DEVICE scalar dL/dloss (nullable = 1), read by
 * the kernels so the host never synchronises.  n_total = global row count (data parallel). */
int lipvq_decoder_bwd(const lipvq_dims_t *dims, const lipvq_params_t *p, const float *x,
                      int64_t n_rows, int64_t n_total, const float *grad_scale,
                      const lipvq_acts_t *acts, const lipvq_grads_t *g, void *ws, size_t ws_bytes,
                      void *stream);

/* sums[idx[n],:] += src[n,:]; counts[idx[n]] += 1 (counts nullable).  Serves the codebook
 * gradient (index_put_ accumulate of autograd, v5:47) and the EMA cluster statistics.
 * sums/counts are accumulated into: the caller zeroes them. */
int lipvq_scatter_add(const float *src, const int64_t *idx, int64_t n_rows, int32_t D, int32_t K,
                      float *sums, float *counts, void *stream);

/* DETERMINISTIC per-code sums (same results bit for bit on every run and every rank): a counting sort of the rows
 * by code and a fixed reduction tree instead of floating-point atomics.
 *   sums[k,:] (+)= alpha * gs * sum_{idx[n]=k} (table ? table[k,:] - src[n,:] : src[n,:])     gs = *grad_scale or 1
 *   counts[k]   = |{n: idx[n]=k}|   (nullable)
 * accumulate != 0 adds to sums, otherwise sums is overwritten (codes without rows get 0).  K <= 8192.
 * Serves the codebook gradient (autograd's index_put_ accumulate, v5:47) inside lipvq_decoder_bwd and the EMA
 * cluster statistics; ws >= lipvq_segment_sum_ws_bytes(n, D, K). */
size_t lipvq_segment_sum_ws_bytes(int64_t n_rows, int32_t D, int32_t K);
int lipvq_segment_sum(const float *src, const int64_t *idx, const float *table, float alpha,
                      const float *grad_scale, int64_t n_rows, int32_t D, int32_t K, int32_t accumulate,
                      float *sums, float *counts, void *ws, size_t ws_bytes, void *stream);

/* Backward of the encoder from the commitment term only (v5:80; there is no straight-through
 * path): g_ze = 0.5*(z_e - z_q)/(n_total*D) -> sigmoid' -> Lipschitz-norm backward
 * (through scale, sum|W|, softplus) -> MLP backward.  With acts->g_pre (written by the forward for the same
 * n_total) the first two steps are already done; grad_scale is then applied in the GEMM epilogues. */
int lipvq_encoder_bwd(const lipvq_dims_t *dims, const lipvq_params_t *p, const float *x,
                      int64_t n_rows, int64_t n_total, const float *grad_scale,
                      const lipvq_acts_t *acts, const lipvq_grads_t *g, void *ws,
                      size_t ws_bytes, void *stream);

/* Whole forward in one call (inference / no-grad; LLFQVAE_V4.forward v5:70-84):
 * x -> z_latent[N,D] (nullable: tokens-only mode for a consumer that embeds per CODE, the [N,D] write -- 94 % of the
 * forward's mandatory HBM traffic -- is then skipped), idx[N], x_recon[N,F] (nullable), loss_out float[4] (nullable).
 * Intermediates live in ws.
 * aux_stream / ev_fork / ev_join (all three or none; cudaStream_t / cudaEvent_t owned by the CALLER): when given, the
 * parameter-only per-code decoder table is enqueued on aux_stream between ev_fork (recorded on `stream`) and ev_join
 * (waited on by `stream` before the emit pass), so it runs beside the encoder and search kernels.  The library creates
 * no stream and no event of its own: the call is re-entrant across streams and can be captured into a CUDA graph. */
int lipvq_forward_fused(const lipvq_dims_t *dims, const lipvq_params_t *p, const float *x,
                        int64_t n_rows, int64_t n_total, int32_t search_mode, float *z_latent,
                        int64_t *idx, float *x_recon, float *loss_out, int32_t *stats, void *ws,
                        size_t ws_bytes, void *stream, void *aux_stream, void *ev_fork, void *ev_join);

/* The same forward as ONE kernel launch (+ one memset node) for the reference's real call sizes, N = B*T = 80 / 500
 * rows (v5:89-93; obs_nets.py:1298): exact fp32 on the CUDA cores, direct-form distances sum (z-e)^2, first minimal
 * index on ties, deterministic loss reduction.  lipvq_forward_fused routes here by itself for n_rows <= 768 (the measured crossover);
 * this entry point accepts up to 65 536 rows.  ws: 8 B per row + 12 B per 4 rows + 1 KB. */
int lipvq_forward_small(const lipvq_dims_t *dims, const lipvq_params_t *p, const float *x,
                        int64_t n_rows, int64_t n_total, float *z_latent, int64_t *idx, float *x_recon,
                        float *loss_out, int32_t *stats, void *ws, size_t ws_bytes, void *stream);

/* Allocation-free variant of lipvq_forward_small.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  Same kernel, same results; there is no ws argument: the kernel's scratch
 * (8 B per row + 12 B per 4 rows) lives in a 64 KB file-scope `static __device__` array inside the library, which every
 * GPU that loads the library holds until exit.  The call performs no cudaMalloc / cudaFree and no allocating library
 * call; inputs and outputs are the caller's.
 *   limit : n_rows <= LIPVQ_FORWARD_SMALL_STATIC_MAX_ROWS, otherwise LIPVQ_ELIMIT (nothing is launched);
 *   sharing: all calls on one device use the SAME array -- the part that must start at zero is cleared on `stream`
 *            before the kernel, so back-to-back calls on one stream are safe, but calls that overlap on the GPU
 *            (different streams, or concurrent host threads) race and give wrong tokens;
 *   set-up : the first call on each device resolves the array's device address (cudaGetSymbolAddress); make that
 *            first call outside stream capture. */
#define LIPVQ_FORWARD_SMALL_STATIC_MAX_ROWS 4096
int lipvq_forward_small_static(const lipvq_dims_t *dims, const lipvq_params_t *p, const float *x,
                               int64_t n_rows, int64_t n_total, float *z_latent, int64_t *idx, float *x_recon,
                               float *loss_out, int32_t *stats, void *stream);

/* AdamW step over one flat fp32 tensor (torch.optim.AdamW semantics as configured at
 * robomimic/algo/icl.py:886-889): decoupled weight decay, bias-corrected moments.
 * step is the 1-based step count AFTER increment. */
int lipvq_adamw_step(float *param, const float *grad, float *exp_avg, float *exp_avg_sq,
                     int64_t n, int64_t step, float lr, float beta1, float beta2, float eps,
                     float weight_decay, void *stream);

/* EMA codebook update -- NOT in the reference (parity unpinned):
 * n <- g n + (1-g) counts; m <- g m + (1-g) sums; E <- m / laplace(n). */
int lipvq_ema_update(float *E, float *n_ema, float *m_ema, const float *counts,
                     const float *sums, int32_t K, int32_t D, float gamma, float eps, void *ws,
                     size_t ws_bytes, void *stream);

/* ---- secondary variant: the classic VQVAE of robomimic/models/vq_vae/backbone.py:6-76 (ReLU MLPs, nn.Embedding
 * codebook, squared distances, loss = mse(x_recon, x) + mse(z_q, sg z_e) + cc * mse(sg z_q, z_e), straight-through
 * estimator z_e + (z_q - z_e).detach()).  Its call site is commented out at robomimic/models/obs_nets.py:1220-1222;
 * it reuses lipvq_vq_argmin / lipvq_gather_loss / the scatter-add kernel with a different gradient route. */

/* Backward of Y = act(X W^T + b) given dY = dL/d(pre-activation) [M,N]:
 *   dW[N,K] = dY^T X, db[N] = column sums of dY (nullable), dX[M,K] = (dY W) * [x_mask > 0] (dX nullable;
 *   x_mask = the ReLU OUTPUT X itself when X came out of a ReLU, or null for no mask).  autograd of backbone.py:17-32. */
size_t lipvq_linear_bwd_ws_bytes(int64_t M, int32_t N, int32_t K);
int lipvq_linear_bwd(const float *dY, const float *X, const float *W, const float *x_mask, float *dX,
                     float *dW, float *db, int64_t M, int32_t N, int32_t K, void *ws, size_t ws_bytes,
                     void *stream);

/* sums[idx[n],:] += alpha * gs * (table[idx[n],:] - src[n,:]) (gs = *grad_scale, nullable): the embedding-loss
 * gradient 2 (z_q - z_e)/(N D) summed per code (backbone.py:70) without materialising the difference. */
int lipvq_scatter_add_diff(const float *src, const int64_t *idx, const float *table, float alpha,
                           const float *grad_scale, int64_t n_rows, int32_t D, int32_t K, float *sums,
                           void *stream);

/* out[0] = sum (a - b)^2 over n elements (device double), fixed-order reduction; ws >= 4 KB. */
int lipvq_sse(const float *a, const float *b, int64_t n, double *out, void *ws, size_t ws_bytes, void *stream);

/* out float[4] = {loss, recon, commitment_cost * mse_q, mse_q} (backbone.py:49-50, 69-71). */
int lipvq_vqvae_loss(const double *sse_r, const double *sse_q, int64_t n_total, int32_t F, int32_t D,
                     float commitment_cost, float *out, void *stream);

/* out = gs * alpha * (x_recon - x) * [x_recon > 0]: mse gradient through the decoder's output ReLU (backbone.py:31). */
int lipvq_vqvae_recon_grad(const float *x_recon, const float *x, int64_t n_elems, float alpha,
                           const float *grad_scale, float *out, void *stream);

/* g_zq <- (g_zq + gs * alpha * (z_e - z_q)) * [z_e > 0]: straight-through route of the decoder gradient to the
 * encoder + commitment term + the encoder's output ReLU (backbone.py:23, 69, 74). */
int lipvq_vqvae_ste_grad(float *g_zq, const float *z_e, const float *z_q, int64_t n_elems, float alpha,
                         const float *grad_scale, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* LIPVQ_B200_H */
