"""B200-native drop-in for the reference's classic ``VQVAE`` (robomimic/models/vq_vae/backbone.py:6-76), the
secondary tokenizer variant whose call site is commented out at robomimic/models/obs_nets.py:1220-1222
(SURVEY 8a row 12).  Same constructor, same ``state_dict`` keys (``encoder.{0,2,4}``, ``decoder.{0,2,4}``,
``embedding.weight``), same ``forward(x) -> (z_latent detached, loss)``; differences from ``LLFQVAE_V4``:
ReLU MLPs (also on both outputs), squared distances, ``loss = mse(x_recon, x) + mse(z_q, sg z_e) +
commitment_cost * mse(sg z_q, z_e)`` and a straight-through estimator, so the reconstruction gradient reaches the
ENCODER (not the codebook) and the codebook only sees the embedding loss.

It reuses the path's kernels through the C ABI -- ``lipvq_vq_argmin`` (tcgen05 search + fp32 rescore),
``lipvq_gather_loss``, the scatter-add kernel (``lipvq_scatter_add_diff``), the linear layers -- with the other
gradient route; there is no torch or CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn as nn

from . import _native as N
from ._native import LipVQNativeError
from .tokenizer import NativeLinear, _check_cuda, _p, _stream_ptr, _workspace

ACT_NONE, ACT_RELU = 0, 5


class _NativeReLULinear(NativeLinear):
    _act = ACT_RELU


def _lin_fwd(lib, x, w, b, act, st):
    y = torch.empty(x.shape[0], w.shape[0], dtype=torch.float32, device=x.device)
    N.check(lib.lipvq_linear_fwd(_p(x), _p(w), _p(b), _p(y), x.shape[0], w.shape[0], w.shape[1], act, st),
            "lipvq_linear_fwd")
    return y


class _VQVAEFn(torch.autograd.Function):
    """inputs: x, n_total, search_mode, commitment_cost, the 13 parameters (reference state_dict order)."""

    @staticmethod
    def forward(ctx, x, n_total, search_mode, cc, *params):
        lib = N.lib()
        dev = x.device
        st = _stream_ptr(dev)
        params = [p if p.is_contiguous() else p.contiguous() for p in params]
        (w0, b0, w1, b1, w2, b2, v0, c0, v1, c1, v2, c2, E) = params
        n, F_ = x.shape
        K, D = E.shape
        f32 = dict(dtype=torch.float32, device=dev)
        stats = torch.zeros(4, **f32)
        idx = torch.empty(n, dtype=torch.int64, device=dev)
        z_q = torch.empty(n, D, **f32)
        if n == 0:
            ctx.empty = True
            ctx.mark_non_differentiable(z_q, idx)
            return z_q, stats, idx
        ctx.empty = False
        ws = _workspace(dev, N.Dims(F_, 64, 128, D, K), n)
        h1 = _lin_fwd(lib, x, w0, b0, ACT_RELU, st)                    # vq:17-24
        h2 = _lin_fwd(lib, h1, w1, b1, ACT_RELU, st)
        z_e = _lin_fwd(lib, h2, w2, b2, ACT_RELU, st)
        vq_stats = torch.zeros(4, dtype=torch.int32, device=dev)
        sse = torch.zeros(2, dtype=torch.float64, device=dev)         # [recon, quantisation]
        N.check(lib.lipvq_vq_argmin(_p(z_e), _p(E), n, D, K, search_mode, _p(idx), None, _p(vq_stats), _p(ws),
                                    ws.numel(), st), "lipvq_vq_argmin")               # vq:57-63 (same argmin)
        N.check(lib.lipvq_gather_loss(_p(z_e), _p(E), _p(idx), n, D, K, _p(z_q), C.c_void_p(sse.data_ptr() + 8),
                                      _p(ws), ws.numel(), st), "lipvq_gather_loss")   # vq:66, 69-70
        d1 = _lin_fwd(lib, z_q, v0, c0, ACT_RELU, st)                  # vq:25-32 (STE forward value == z_q, vq:74)
        d2 = _lin_fwd(lib, d1, v1, c1, ACT_RELU, st)
        xr = _lin_fwd(lib, d2, v2, c2, ACT_RELU, st)
        N.check(lib.lipvq_sse(_p(xr), _p(x), n * F_, _p(sse), _p(ws), ws.numel(), st), "lipvq_sse")
        N.check(lib.lipvq_vqvae_loss(_p(sse), C.c_void_p(sse.data_ptr() + 8), n_total, F_, D, float(cc), _p(stats),
                                     st), "lipvq_vqvae_loss")
        ctx.save_for_backward(x, *params, h1, h2, z_e, idx, z_q, d1, d2, xr)
        ctx.n_total, ctx.cc = n_total, float(cc)
        ctx.mark_non_differentiable(z_q, idx)
        return z_q, stats, idx

    @staticmethod
    def backward(ctx, g_zq_unused, g_stats, g_idx):
        none4 = (None, None, None, None)
        if ctx.empty or g_stats is None:
            return none4 + (None,) * 13
        lib = N.lib()
        saved = ctx.saved_tensors
        x = saved[0]
        (w0, b0, w1, b1, w2, b2, v0, c0, v1, c1, v2, c2, E) = saved[1:14]
        h1, h2, z_e, idx, z_q, d1, d2, xr = saved[14:]
        dev = x.device
        st = _stream_ptr(dev)
        n, F_ = x.shape
        K, D = E.shape
        gs = g_stats[0:1].contiguous().float()                        # dL/dloss stays on the device
        f32 = dict(dtype=torch.float32, device=dev)
        need = max(lib.lipvq_linear_bwd_ws_bytes(n, a, b) for a, b in ((F_, 64), (64, 128), (128, D), (D, 128),
                                                                        (128, 64), (64, F_)))
        ws = torch.empty(int(need) + 4096, dtype=torch.uint8, device=dev)
        grads = {k: torch.empty_like(t) for k, t in zip("w0 b0 w1 b1 w2 b2 v0 c0 v1 c1 v2 c2".split(),
                                                        (w0, b0, w1, b1, w2, b2, v0, c0, v1, c1, v2, c2))}
        gE = torch.zeros_like(E)

        def lin_bwd(dY, X, W, mask, want_dx, gw, gb):
            dX = torch.empty_like(X) if want_dx else None
            N.check(lib.lipvq_linear_bwd(_p(dY), _p(X), _p(W), _p(mask), _p(dX), _p(gw), _p(gb), dY.shape[0],
                                         W.shape[0], W.shape[1], _p(ws), ws.numel(), st), "lipvq_linear_bwd")
            return dX
        a_r = 2.0 / (float(ctx.n_total) * F_)
        a_q = 2.0 / (float(ctx.n_total) * D)
        g = torch.empty_like(xr)
        N.check(lib.lipvq_vqvae_recon_grad(_p(xr), _p(x), n * F_, a_r, _p(gs), _p(g), st), "lipvq_vqvae_recon_grad")
        g = lin_bwd(g, d2, v2, d2, True, grads["v2"], grads["c2"])     # decoder.4 (input d2 = ReLU output)
        g = lin_bwd(g, d1, v1, d1, True, grads["v1"], grads["c1"])     # decoder.2
        g = lin_bwd(g, z_q, v0, None, True, grads["v0"], grads["c0"])  # decoder.0: dL/dz_q
        # straight-through: dL/dz_q goes to z_e (vq:74), plus the commitment term, through the encoder's ReLU
        N.check(lib.lipvq_vqvae_ste_grad(_p(g), _p(z_e), _p(z_q), n * D, ctx.cc * a_q, _p(gs), st),
                "lipvq_vqvae_ste_grad")
        # embedding loss -> codebook: sum_{idx=k} 2 (E[k] - z_e[n]) / (N D)   (vq:70)
        N.check(lib.lipvq_scatter_add_diff(_p(z_e), _p(idx), _p(E), a_q, _p(gs), n, D, K, _p(gE), st),
                "lipvq_scatter_add_diff")
        g = lin_bwd(g, h2, w2, h2, True, grads["w2"], grads["b2"])     # encoder.4
        g = lin_bwd(g, h1, w1, h1, True, grads["w1"], grads["b1"])     # encoder.2
        lin_bwd(g, x, w0, None, False, grads["w0"], grads["b0"])       # encoder.0
        order = "w0 b0 w1 b1 w2 b2 v0 c0 v1 c1 v2 c2".split()
        out = [grads[k] if ctx.needs_input_grad[4 + i] else None for i, k in enumerate(order)]
        out.append(gE if ctx.needs_input_grad[16] else None)
        return none4 + tuple(out)


class VQVAE(nn.Module):
    """Drop-in for ``robomimic.models.vq_vae.backbone.VQVAE`` (vq:6-76)."""

    def __init__(self, feature_dim, latent_dim, num_embeddings=128, commitment_cost=0.25):
        super().__init__()
        self.feature_dim, self.latent_dim = feature_dim, latent_dim
        self.num_embeddings, self.commitment_cost = num_embeddings, commitment_cost
        self.encoder = nn.Sequential(_NativeReLULinear(feature_dim, 64), nn.ReLU(), _NativeReLULinear(64, 128), nn.ReLU(),
                                     _NativeReLULinear(128, latent_dim), nn.ReLU())
        self.decoder = nn.Sequential(_NativeReLULinear(latent_dim, 128), nn.ReLU(), _NativeReLULinear(128, 64), nn.ReLU(),
                                     _NativeReLULinear(64, feature_dim), nn.ReLU())
        self.embedding = nn.Embedding(num_embeddings, latent_dim)
        self.embedding.weight.data.uniform_(-1 / num_embeddings, 1 / num_embeddings)      # vq:36
        self.search_mode = N.SEARCH_TENSOR
        self.n_total_rows = None
        self.last_indices = None
        self.last_loss_terms = None

    def _param_list(self):
        e, d = self.encoder, self.decoder
        return [e[0].weight, e[0].bias, e[2].weight, e[2].bias, e[4].weight, e[4].bias,
                d[0].weight, d[0].bias, d[2].weight, d[2].bias, d[4].weight, d[4].bias, self.embedding.weight]

    def forward(self, x):
        _check_cuda(x, "x")
        if x.dim() != 2 or x.shape[1] != self.feature_dim:
            raise RuntimeError("expected x of shape [N, %d], got %s" % (self.feature_dim, tuple(x.shape)))
        params = self._param_list()
        if not torch.is_grad_enabled():
            params = [p.detach() for p in params]
        n_total = int(self.n_total_rows) if self.n_total_rows is not None else x.shape[0]
        z_latent, stats, idx = _VQVAEFn.apply(x.detach().contiguous(), max(n_total, 1), int(self.search_mode),
                                              float(self.commitment_cost), *params)
        self.last_indices = idx
        self.last_loss_terms = stats.detach()
        return z_latent, stats[0]

    @torch.no_grad()
    def quantize(self, z_e):
        """(z_q, quantization_loss) = VQVAE.quantize (vq:55-76); forward value of the STE output is z_q."""
        _check_cuda(z_e, "z_e")
        z_e = z_e.contiguous()
        E = self.embedding.weight.detach().contiguous()
        n, D = z_e.shape
        K = E.shape[0]
        dev = z_e.device
        st = _stream_ptr(dev)
        ws = _workspace(dev, N.Dims(1, 1, 1, D, K), n)
        idx = torch.empty(n, dtype=torch.int64, device=dev)
        z_q = torch.empty(n, D, dtype=torch.float32, device=dev)
        sse = torch.zeros(1, dtype=torch.float64, device=dev)
        vq_stats = torch.zeros(4, dtype=torch.int32, device=dev)
        lib = N.lib()
        N.check(lib.lipvq_vq_argmin(_p(z_e), _p(E), n, D, K, int(self.search_mode), _p(idx), None, _p(vq_stats), _p(ws),
                                    ws.numel(), st), "lipvq_vq_argmin")
        N.check(lib.lipvq_gather_loss(_p(z_e), _p(E), _p(idx), n, D, K, _p(z_q), _p(sse), _p(ws), ws.numel(), st),
                "lipvq_gather_loss")
        self.last_indices = idx
        mse = (sse / float(max(n, 1) * D)).float()[0]
        return z_q, mse * (1.0 + self.commitment_cost)
