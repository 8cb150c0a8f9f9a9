"""CPU oracle for the LipVQ-VAE action-tokenizer hot path.  TEST INFRASTRUCTURE ONLY.

This file is a numpy restatement of the reference algorithm.  It is the *checker*: only
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import it.  Nothing under ``lipvq_vae_b200/`` imports it, and the product
path never falls back to it (the product raises when the CUDA library is missing).

Parity status: PINNED.  ``tests/golden/make_golden.py`` imports the real reference module
(``/root/reference/robomimic/models/vq_vae/backbone_lfqvae_v5.py``) in the authoring
container, runs it on numpy-seeded parameters/inputs and stores its outputs under
``tests/golden/*.npz``; ``tests/test_oracle_golden.py`` checks every function below against
those vectors.  The EMA codebook update (``ema_update``) has no reference counterpart and is
"parity unpinned" (SURVEY.md section 8a row 11).

Reference citations use ``v5`` = robomimic/models/vq_vae/backbone_lfqvae_v5.py and
``vq`` = robomimic/models/vq_vae/backbone.py under /root/reference.

Parameter dictionary keys are the reference ``state_dict`` names (v5:52-68):
    encoder.0.weight [64,F]  encoder.0.bias [64]  encoder.2.weight [H,64]  encoder.2.bias [H]
    to_latent.W [D,H]  to_latent.b [D]  to_latent.ci [D]
    quantizer.codebook [K,D]
    decoder.0.weight [64,D]  decoder.0.bias [64]  decoder.2.weight [H,64]  decoder.2.bias [H]
    to_output.weight [F,H]  to_output.bias [F]
"""
from __future__ import annotations

import math
from typing import Dict, Tuple

import numpy as np
from scipy.special import erf as _erf

PARAM_NAMES = (
    "encoder.0.weight", "encoder.0.bias", "encoder.2.weight", "encoder.2.bias",
    "to_latent.W", "to_latent.b", "to_latent.ci",
    "quantizer.codebook",
    "decoder.0.weight", "decoder.0.bias", "decoder.2.weight", "decoder.2.bias",
    "to_output.weight", "to_output.bias",
)

_SQRT1_2 = 0.7071067811865476
_INV_SQRT_2PI = 0.3989422804014327


def param_shapes(feature_dim: int, latent_dim: int, num_codes: int = 1024, hidden_dim: int = 128):
    """Shapes of the 14 tensors built by LLFQVAE_V4.__init__ (v5:52-68)."""
    F, D, K, H = feature_dim, latent_dim, num_codes, hidden_dim
    return {
        "encoder.0.weight": (64, F), "encoder.0.bias": (64,),
        "encoder.2.weight": (H, 64), "encoder.2.bias": (H,),
        "to_latent.W": (D, H), "to_latent.b": (D,), "to_latent.ci": (D,),
        "quantizer.codebook": (K, D),
        "decoder.0.weight": (64, D), "decoder.0.bias": (64,),
        "decoder.2.weight": (H, 64), "decoder.2.bias": (H,),
        "to_output.weight": (F, H), "to_output.bias": (F,),
    }


def init_params(feature_dim, latent_dim, num_codes=1024, hidden_dim=128, seed=0, dtype=np.float32):
    """Random parameters with the reference's init *distributions* (v5:18-20, 32-35, nn.Linear
    default = U(+-1/sqrt(fan_in)) for weight and bias), drawn from numpy's RNG so fixtures can
    be regenerated anywhere."""
    rng = np.random.default_rng(seed)
    p = {}
    for name, shp in param_shapes(feature_dim, latent_dim, num_codes, hidden_dim).items():
        if name == "to_latent.W":
            v = rng.standard_normal(shp)
        elif name == "to_latent.b":
            v = np.zeros(shp)
        elif name == "to_latent.ci":
            v = np.ones(shp)
        elif name == "quantizer.codebook":
            bound = math.sqrt(6.0 / shp[1])          # kaiming_uniform_, a=0, fan_in = D
            v = rng.uniform(-bound, bound, shp)
        else:
            fan_in = shp[1] if len(shp) == 2 else {"encoder.0.bias": feature_dim,
                                                   "encoder.2.bias": 64,
                                                   "decoder.0.bias": latent_dim,
                                                   "decoder.2.bias": 64,
                                                   "to_output.bias": hidden_dim}[name]
            bound = 1.0 / math.sqrt(fan_in)
            v = rng.uniform(-bound, bound, shp)
        p[name] = np.ascontiguousarray(v, dtype=dtype)
    return p


# ----------------------------------------------------------------------------- elementwise
def gelu(a):
    """Exact-erf GELU, nn.GELU() default (v5:56,58,64,66)."""
    a64 = a.astype(np.float64)
    return (0.5 * a64 * (1.0 + _erf(a64 * _SQRT1_2))).astype(a.dtype)


def gelu_grad(a):
    a64 = a.astype(np.float64)
    return (0.5 * (1.0 + _erf(a64 * _SQRT1_2)) + a64 * np.exp(-0.5 * a64 * a64) * _INV_SQRT_2PI).astype(a.dtype)


def softplus(c):
    """F.softplus, beta=1, threshold=20 (v5:10)."""
    c64 = c.astype(np.float64)
    return np.where(c64 > 20.0, c64, np.log1p(np.exp(np.minimum(c64, 20.0)))).astype(c.dtype)


def sigmoid(a):
    a64 = a.astype(np.float64)
    return (1.0 / (1.0 + np.exp(-a64))).astype(a.dtype)


# ----------------------------------------------------------------------------- encoder
def lipschitz_scale(W, ci):
    """scale_i = min(1, softplus(ci_i) / sum_j |W_ij|)  (v5:6-12)."""
    absrowsum = np.sum(np.abs(W), axis=1, dtype=W.dtype)
    return np.minimum(np.asarray(1.0, W.dtype), softplus(ci) / absrowsum)


def encode(p, x, return_cache=False):
    """z_e = sigmoid(GELU(GELU(x W0^T+b0) W1^T+b1) (W*scale)^T + b)  (v5:54-60, 22-24, 71-72)."""
    a1 = x @ p["encoder.0.weight"].T + p["encoder.0.bias"]
    h1 = gelu(a1)
    a2 = h1 @ p["encoder.2.weight"].T + p["encoder.2.bias"]
    h2 = gelu(a2)
    scale = lipschitz_scale(p["to_latent.W"], p["to_latent.ci"])
    Wn = p["to_latent.W"] * scale[:, None]
    z_e = sigmoid(h2 @ Wn.T + p["to_latent.b"])
    if return_cache:
        return z_e, dict(a1=a1, h1=h1, a2=a2, h2=h2, scale=scale, Wn=Wn)
    return z_e


# ----------------------------------------------------------------------------- quantizer
def quantize_reference_form(z_e, E, block_rows=256):
    """Nearest codeword exactly as LFQQuantizer.forward spells it (v5:37-48): sign mask,
    broadcast difference [n,K,D], L2 norm with the sqrt, argmin (first minimal index).
    Evaluated in row blocks because the broadcast temporary is n*K*D*4 bytes."""
    N = z_e.shape[0]
    idx = np.empty(N, dtype=np.int64)
    for s in range(0, N, block_rows):
        z = z_e[s:s + block_rows]
        mask = np.minimum(2 * np.sign(z) + 1, 1).astype(z.dtype)[:, None, :]      # v5:39-40
        diff = mask * (z[:, None, :] - E[None, :, :])                                # v5:41-44
        dist = np.sqrt(np.sum(diff * diff, axis=-1, dtype=z.dtype))                  # v5:43-45
        idx[s:s + block_rows] = np.argmin(dist, axis=-1)                             # v5:46
    return idx


def quantize_gemm_form(z_e, E, block_rows=16384):
    """Same search in GEMM form: argmin_k ||e_k||^2 - 2 z.e_k ("sensible CPU" baseline,
    SURVEY.md section 6).  Equal to the reference form outside the fp32 rounding window."""
    N = z_e.shape[0]
    idx = np.empty(N, dtype=np.int64)
    en = np.sum(E * E, axis=1, dtype=E.dtype)
    Et = np.ascontiguousarray(E.T)
    for s in range(0, N, block_rows):
        sc = en[None, :] - 2.0 * (z_e[s:s + block_rows] @ Et)
        idx[s:s + block_rows] = np.argmin(sc, axis=-1)
    return idx


def exact_top2(z_e, E, block_rows=4096):
    """float64 ground truth: argmin, best and second-best *distances* (with the sqrt) per row.
    Rows whose relative top-2 gap (d2-d1)/d1 is <= 1e-5 are exempt from index parity
    (BASELINE.json north_star)."""
    z = z_e.astype(np.float64)
    Ed = E.astype(np.float64)
    en = np.sum(Ed * Ed, axis=1)
    N = z.shape[0]
    idx = np.empty(N, np.int64)
    d1 = np.empty(N, np.float64)
    d2 = np.empty(N, np.float64)
    for s in range(0, N, block_rows):
        zb = z[s:s + block_rows]
        sq = np.sum(zb * zb, axis=1)[:, None] + en[None, :] - 2.0 * (zb @ Ed.T)
        i = np.argmin(sq, axis=1)
        idx[s:s + block_rows] = i
        # recompute the two smallest in direct form to avoid cancellation
        if sq.shape[1] > 1:
            part = np.argpartition(sq, 1, axis=1)[:, :2]
        else:
            part = np.zeros((sq.shape[0], 2), np.int64)
        da = np.sqrt(np.sum((zb - Ed[part[:, 0]]) ** 2, axis=1))
        db = np.sqrt(np.sum((zb - Ed[part[:, 1]]) ** 2, axis=1)) if sq.shape[1] > 1 else np.full(len(zb), np.inf)
        lo = np.minimum(da, db)
        hi = np.maximum(da, db)
        # make idx consistent with the direct-form winner (lowest index on exact ties)
        win = np.where((da < db) | ((da == db) & (part[:, 0] < part[:, 1])), part[:, 0], part[:, 1])
        idx[s:s + block_rows] = win
        d1[s:s + block_rows] = lo
        d2[s:s + block_rows] = hi
    return idx, d1, d2


def index_parity_mask(d1, d2, rel=1e-5):
    """True where the index must match bit-exactly: top-2 relative distance gap > rel."""
    return (d2 - d1) > rel * np.maximum(d1, np.finfo(np.float64).tiny)


# ----------------------------------------------------------------------------- decoder / loss
def decode(p, z_q, return_cache=False):
    """x_recon = GELU(GELU(z_q Wd0^T+bd0) Wd1^T+bd1) Wo^T + bo  (v5:62-68, 75-76)."""
    c1 = z_q @ p["decoder.0.weight"].T + p["decoder.0.bias"]
    d1 = gelu(c1)
    c2 = d1 @ p["decoder.2.weight"].T + p["decoder.2.bias"]
    d2 = gelu(c2)
    xr = d2 @ p["to_output.weight"].T + p["to_output.bias"]
    if return_cache:
        return xr, dict(c1=c1, d1=d1, c2=c2, d2=d2)
    return xr


def forward(p, x, search="exact"):
    """LLFQVAE_V4.forward (v5:70-84).  Returns dict with z_latent, loss and intermediates.
    loss = mse(x_recon,x) + 0.25*mse(z_q,z_e) + 0.25*mse(z_q,z_e)  (v5:79-83).
    search: "exact" (float64 direct form, the ground truth), "reference" (fp32 broadcast form,
    v5:37-48 literally) or "gemm" (fp32 expanded form; cancellation-prone once ||z-e|| << ||z||,
    e.g. tests/golden case c1_spread -- which is why the CUDA path rescoring is direct-form)."""
    z_e, ce = encode(p, x, return_cache=True)
    E = p["quantizer.codebook"]
    if search == "reference":
        idx = quantize_reference_form(z_e, E)
    elif search == "gemm":
        idx = quantize_gemm_form(z_e, E)
    else:
        idx = exact_top2(z_e, E)[0]
    z_q = E[idx]
    xr, cd = decode(p, z_q, return_cache=True)
    N, F = x.shape
    D = z_e.shape[1]
    sse_r = float(np.sum((xr.astype(np.float64) - x) ** 2))
    sse_q = float(np.sum((z_q.astype(np.float64) - z_e) ** 2))
    recon = sse_r / (N * F)
    commit = sse_q / (N * D)
    loss = recon + 0.25 * commit + 0.25 * commit
    return dict(z_e=z_e, idx=idx, z_q=z_q, z_latent=z_q.copy(), x_recon=xr, loss=loss,
                recon_loss=recon, commit_loss=commit, sse_r=sse_r, sse_q=sse_q, enc=ce, dec=cd)


def backward(p, x, fw, n_total=None) -> Dict[str, np.ndarray]:
    """Gradients of ``loss`` w.r.t. the 14 parameters, following the autograd routing of
    v5:73-83 (verified against the reference in tests/golden): recon -> decoder, to_output and
    the codebook (gather -> scatter-add); commitment -> encoder/to_latent only; codebook loss
    -> codebook only.  No straight-through path.  ``n_total`` is the global row count when
    ``x`` is one data-parallel shard (loss means are over the global batch)."""
    N, F = x.shape
    NT = N if n_total is None else n_total
    z_e, z_q, idx, xr = fw["z_e"], fw["z_q"], fw["idx"], fw["x_recon"]
    D = z_e.shape[1]
    ce, cd = fw["enc"], fw["dec"]
    f = np.float32
    g = {}
    g_xr = ((2.0 / (NT * F)) * (xr - x)).astype(f)
    g["to_output.weight"] = g_xr.T @ cd["d2"]
    g["to_output.bias"] = g_xr.sum(0)
    g_c2 = (g_xr @ p["to_output.weight"]) * gelu_grad(cd["c2"])
    g["decoder.2.weight"] = g_c2.T @ cd["d1"]
    g["decoder.2.bias"] = g_c2.sum(0)
    g_c1 = (g_c2 @ p["decoder.2.weight"]) * gelu_grad(cd["c1"])
    g["decoder.0.weight"] = g_c1.T @ z_q
    g["decoder.0.bias"] = g_c1.sum(0)
    g_zq = g_c1 @ p["decoder.0.weight"] + ((0.5 / (NT * D)) * (z_q - z_e)).astype(f)
    gE = np.zeros_like(p["quantizer.codebook"])
    np.add.at(gE, idx, g_zq)                                   # index_put_(accumulate=True)
    g["quantizer.codebook"] = gE
    g_ze = ((0.5 / (NT * D)) * (z_e - z_q)).astype(f)
    g_pre = g_ze * z_e * (1.0 - z_e)                           # sigmoid'
    g["to_latent.b"] = g_pre.sum(0)
    gWn = g_pre.T @ ce["h2"]
    W, ci = p["to_latent.W"], p["to_latent.ci"]
    sp = softplus(ci)
    r = np.sum(np.abs(W), axis=1, dtype=W.dtype)
    ratio = sp / r
    # torch.minimum routes the gradient to the smaller argument (half each on equality)
    w_ratio = np.where(ratio < 1.0, 1.0, np.where(ratio == 1.0, 0.5, 0.0)).astype(f)
    dot = np.sum(gWn * W, axis=1)                              # dL/dscale_i
    g_ratio = dot * w_ratio
    g["to_latent.W"] = gWn * ce["scale"][:, None] + (g_ratio * (-sp / (r * r)))[:, None] * np.sign(W)
    g["to_latent.ci"] = g_ratio / r * sigmoid(ci)              # softplus' = sigmoid
    g_a2 = (g_pre @ ce["Wn"]) * gelu_grad(ce["a2"])
    g["encoder.2.weight"] = g_a2.T @ ce["h1"]
    g["encoder.2.bias"] = g_a2.sum(0)
    g_a1 = (g_a2 @ p["encoder.2.weight"]) * gelu_grad(ce["a1"])
    g["encoder.0.weight"] = g_a1.T @ x
    g["encoder.0.bias"] = g_a1.sum(0)
    return {k: np.ascontiguousarray(v, dtype=f) for k, v in g.items()}


def adamw_step(p, g, state, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-4):
    """torch.optim.AdamW semantics as configured at robomimic/algo/icl.py:886-889
    (decoupled weight decay, bias-corrected moments).  ``state`` = {"step", "m", "v"}."""
    state["step"] += 1
    t = state["step"]
    b1, b2 = betas
    bc1 = 1.0 - b1 ** t
    bc2 = 1.0 - b2 ** t
    out = {}
    for k in PARAM_NAMES:
        w = p[k].astype(np.float32) * np.float32(1.0 - lr * weight_decay)
        m = state["m"][k] = (b1 * state["m"][k] + (1.0 - b1) * g[k]).astype(np.float32)
        v = state["v"][k] = (b2 * state["v"][k] + (1.0 - b2) * g[k] * g[k]).astype(np.float32)
        denom = np.sqrt(v) / np.float32(math.sqrt(bc2)) + np.float32(eps)
        out[k] = (w - np.float32(lr / bc1) * (m / denom)).astype(np.float32)
    return out


def adamw_init(p):
    return {"step": 0, "m": {k: np.zeros_like(p[k]) for k in PARAM_NAMES},
            "v": {k: np.zeros_like(p[k]) for k in PARAM_NAMES}}


# ----------------------------------------------------------------------------- secondary variant
def vqvae_quantize(z_e, E, commitment_cost=0.25):
    """Classic VQVAE.quantize (vq:55-76): squared distances (no sqrt), argmin, gather,
    quantization loss = mse(z_q, sg z_e) + cc * mse(sg z_q, z_e); STE forward value == z_q."""
    idx = exact_top2(z_e, E)[0]
    z_q = E[idx]
    mse = float(np.mean((z_q.astype(np.float64) - z_e) ** 2))
    return z_q, idx, mse + commitment_cost * mse


VQ_PARAM_NAMES = ("encoder.0.weight", "encoder.0.bias", "encoder.2.weight", "encoder.2.bias",
                  "encoder.4.weight", "encoder.4.bias", "decoder.0.weight", "decoder.0.bias",
                  "decoder.2.weight", "decoder.2.bias", "decoder.4.weight", "decoder.4.bias", "embedding.weight")


def vqvae_param_shapes(feature_dim, latent_dim, num_embeddings=128):
    """state_dict of the classic VQVAE (vq:17-36): ReLU MLPs 12-64-128-D / D-128-64-12, nn.Embedding codebook."""
    F, D, K = feature_dim, latent_dim, num_embeddings
    return {"encoder.0.weight": (64, F), "encoder.0.bias": (64,), "encoder.2.weight": (128, 64),
            "encoder.2.bias": (128,), "encoder.4.weight": (D, 128), "encoder.4.bias": (D,),
            "decoder.0.weight": (128, D), "decoder.0.bias": (128,), "decoder.2.weight": (64, 128),
            "decoder.2.bias": (64,), "decoder.4.weight": (F, 64), "decoder.4.bias": (F,),
            "embedding.weight": (K, D)}


def vqvae_init_params(feature_dim, latent_dim, num_embeddings=128, seed=0):
    """nn.Linear default init; embedding U(+-1/K) (vq:35-36)."""
    rng = np.random.default_rng(seed)
    p = {}
    for name, shp in vqvae_param_shapes(feature_dim, latent_dim, num_embeddings).items():
        if name == "embedding.weight":
            b = 1.0 / num_embeddings
        else:
            fan_in = shp[1] if len(shp) == 2 else vqvae_param_shapes(feature_dim, latent_dim, num_embeddings)[
                name.replace("bias", "weight")][1]
            b = 1.0 / math.sqrt(fan_in)
        p[name] = rng.uniform(-b, b, shp).astype(np.float32)
    return p


def _lin(x, W, b):
    return (x @ W.T + b).astype(np.float32)


def vqvae_forward(p, x, commitment_cost=0.25):
    """VQVAE.forward (vq:38-53) + quantize (vq:55-76), fp32 like the reference."""
    relu = lambda a: np.maximum(a, np.float32(0))
    h1 = relu(_lin(x, p["encoder.0.weight"], p["encoder.0.bias"]))
    h2 = relu(_lin(h1, p["encoder.2.weight"], p["encoder.2.bias"]))
    z_e = relu(_lin(h2, p["encoder.4.weight"], p["encoder.4.bias"]))
    E = p["embedding.weight"]
    idx = exact_top2(z_e, E)[0]
    z_q = E[idx]
    d1 = relu(_lin(z_q, p["decoder.0.weight"], p["decoder.0.bias"]))
    d2 = relu(_lin(d1, p["decoder.2.weight"], p["decoder.2.bias"]))
    xr = relu(_lin(d2, p["decoder.4.weight"], p["decoder.4.bias"]))
    recon = float(np.mean((xr.astype(np.float64) - x) ** 2))
    mse_q = float(np.mean((z_q.astype(np.float64) - z_e) ** 2))
    return dict(h1=h1, h2=h2, z_e=z_e, idx=idx, z_q=z_q, z_latent=z_q.copy(), d1=d1, d2=d2, x_recon=xr,
                recon=recon, mse_q=mse_q, loss=recon + mse_q + commitment_cost * mse_q)


def vqvae_backward(p, x, fw, commitment_cost=0.25):
    """autograd of vqvae_forward: recon -> decoder -> (straight-through, vq:74) -> encoder; commitment term
    -> encoder; embedding loss -> codebook only (the decoder does NOT reach the codebook through the STE)."""
    f64 = np.float64
    N, F = x.shape
    D = fw["z_e"].shape[1]
    g = (2.0 / (N * F)) * (fw["x_recon"].astype(f64) - x) * (fw["x_recon"] > 0)
    gr = {}

    def back(g_out, inp, W, mask_src):
        dW = g_out.T @ inp.astype(f64)
        db = g_out.sum(0)
        g_in = g_out @ W.astype(f64)
        if mask_src is not None:
            g_in = g_in * (mask_src > 0)
        return dW, db, g_in
    gr["decoder.4.weight"], gr["decoder.4.bias"], g = back(g, fw["d2"], p["decoder.4.weight"], fw["d2"])
    gr["decoder.2.weight"], gr["decoder.2.bias"], g = back(g, fw["d1"], p["decoder.2.weight"], fw["d1"])
    gr["decoder.0.weight"], gr["decoder.0.bias"], g = back(g, fw["z_q"], p["decoder.0.weight"], None)
    diff = fw["z_e"].astype(f64) - fw["z_q"]
    g = (g + commitment_cost * (2.0 / (N * D)) * diff) * (fw["z_e"] > 0)
    gE = np.zeros(p["embedding.weight"].shape, f64)
    np.add.at(gE, fw["idx"], -(2.0 / (N * D)) * diff)
    gr["embedding.weight"] = gE
    gr["encoder.4.weight"], gr["encoder.4.bias"], g = back(g, fw["h2"], p["encoder.4.weight"], fw["h2"])
    gr["encoder.2.weight"], gr["encoder.2.bias"], g = back(g, fw["h1"], p["encoder.2.weight"], fw["h1"])
    gr["encoder.0.weight"], gr["encoder.0.bias"], _ = back(g, x, p["encoder.0.weight"], None)
    return {k: v.astype(np.float32) for k, v in gr.items()}


# ----------------------------------------------------------------------------- EMA extension
def ema_update(E, n_ema, m_ema, z_e, idx, gamma=0.99, eps=1e-5):
    """EMA codebook update -- NOT IN THE REFERENCE (parity unpinned).  Standard VQ-VAE EMA:
    n <- g n + (1-g) count; m <- g m + (1-g) sum_{idx=k} z_e; E <- m / laplace(n)."""
    K, D = E.shape
    counts = np.bincount(idx, minlength=K).astype(np.float32)
    sums = np.zeros((K, D), np.float32)
    np.add.at(sums, idx, z_e)
    n_new = (gamma * n_ema + (1 - gamma) * counts).astype(np.float32)
    m_new = (gamma * m_ema + (1 - gamma) * sums).astype(np.float32)
    tot = n_new.sum()
    n_s = (n_new + eps) / (tot + K * eps) * tot
    return (m_new / n_s[:, None]).astype(np.float32), n_new, m_new, counts, sums


# ----------------------------------------------------------------------------- comparisons
def rel_err(a, b):
    """max|a-b| / max(|b|) -- the relative error used for latents / reconstructions / grads."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    den = max(float(np.max(np.abs(b))) if b.size else 0.0, 1e-30)
    return float(np.max(np.abs(a - b))) / den if b.size else 0.0
