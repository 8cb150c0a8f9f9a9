"""B200-native drop-in for the reference tokenizer ``LLFQVAE_V4``
(robomimic/models/vq_vae/backbone_lfqvae_v5.py:51-84; instantiated at
robomimic/models/obs_nets.py:1225-1227 behind ``algo.transformer.vq_vae_enabled``).

Same constructor, same 14 parameter names/shapes/initialisers (so ``state_dict``s interchange
and ``optim.AdamW(model.parameters())`` at robomimic/algo/icl.py:886-889 sees the same
tensors), same ``forward(x) -> (z_latent, loss)`` contract.  All arithmetic runs in the
hand-written sm_100a kernels behind the C ABI (include/lipvq_b200.h); there is no torch or CPU
fallback -- CPU tensors raise ``LipVQNativeError``.
"""
from __future__ import annotations

import ctypes as C
import math

import torch
import torch.nn as nn

from . import _native as N
from ._native import LipVQNativeError

import weakref

_WS = {}  # (device index, stream) -> uint8 workspace tensor
# host-side caches of the no-grad fast path, kept OUTSIDE the module instances (ctypes structs do not deepcopy / pickle)
_PSLOTS = weakref.WeakKeyDictionary()   # model -> [(module._parameters, name)] * 14
_FAST = weakref.WeakKeyDictionary()     # model -> (data_ptrs, Params struct, Dims struct, {rows: workspace bytes})
_AUX = {}  # device index -> (helper stream, fork event, join event) lent to lipvq_forward_fused
AUX_MIN_ROWS = 16384   # below this the fork/join costs more than the 18 us decoder table it hides


def _aux(dev):
    """Helper stream + two events owned by THIS module (the C library owns none): the parameter-only decoder table of
    a large no-grad forward runs on the helper stream beside the encoder / search kernels."""
    a = _AUX.get(dev.index)
    if a is None:
        lo, _hi = 0, 0
        s = torch.cuda.Stream(device=dev, priority=0)
        ef, ej = torch.cuda.Event(), torch.cuda.Event()
        with torch.cuda.device(dev):
            ef.record(s); ej.record(s)             # materialise the cudaEvent_t handles
        a = _AUX[dev.index] = (s, ef, ej)
    return a


def _stream_ptr(dev):
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def _workspace(dev, dims, n_rows):
    need = C.c_size_t(0)
    N.check(N.lib().lipvq_workspace_bytes(C.byref(dims), n_rows, C.byref(need)), "lipvq_workspace_bytes")
    key = (dev.index, torch.cuda.current_stream(dev).cuda_stream)
    ws = _WS.get(key)
    if ws is None or ws.numel() < need.value:
        ws = torch.empty(int(need.value * 1.25) + 4096, dtype=torch.uint8, device=dev)
        _WS[key] = ws
    return ws


def _check_cuda(t, name):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise LipVQNativeError("%s must be a CUDA tensor: lipvq_vae_b200 has no CPU path" % name)
    if t.dtype != torch.float32:
        raise TypeError("%s must be float32 (the reference runs .float(), icl.py:883), got %s" % (name, t.dtype))


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _param_struct(cls, tensors):
    s = cls()
    for f, t in zip(N._PARAM_FIELDS, tensors):
        setattr(s, f, t.data_ptr() if t is not None else 0)
    return s


def _acts_struct(**kw):
    a = N.Acts()
    for f in N._ACT_FIELDS:
        t = kw.get(f)
        setattr(a, f, t.data_ptr() if t is not None else 0)
    return a


def _contig(ts):
    return [t if t.is_contiguous() else t.contiguous() for t in ts]


def _empty_acts(dev, n, dims, with_gpre=True):
    """Activation tensors kept between a training forward and its backward (lipvq_acts_t)."""
    f32 = dict(dtype=torch.float32, device=dev)
    F_, H1, H, D, K = dims.F, dims.H1, dims.H, dims.D, dims.K
    t = dict(a1=torch.empty(n, H1, **f32), h1=torch.empty(n, H1, **f32),
             a2=torch.empty(n, H, **f32), h2=torch.empty(n, H, **f32),
             scale=torch.empty(D, **f32), Wn=torch.empty(D, H, **f32), z_e=torch.empty(n, D, **f32),
             # decoder activations are kept per CODE (z_q takes K distinct values), not per row
             c1=torch.empty(K, H1, **f32), d1=torch.empty(K, H1, **f32),
             c2=torch.empty(K, H, **f32), d2=torch.empty(K, H, **f32), xr_code=torch.empty(K, F_, **f32))
    if with_gpre:
        t["g_pre"] = torch.empty(n, D, **f32)   # commitment gradient through the sigmoid, written by the emit pass
    return t


def train_forward(x, n_total, search_mode, dims, params):
    """Training-mode forward of LLFQVAE_V4.forward (v5:70-84) through the C ABI, every activation the backward
    needs kept.  Returns ``(z_q, stats float[4], idx, vq_stats, x_recon, saved)``; ``saved`` feeds
    ``train_backward_decoder`` / ``train_backward_encoder``.  No autograd involved (``_TokenizerFn`` and
    ``dist.TokenizerTrainer`` are the two callers)."""
    lib = N.lib()
    dev = x.device
    n = x.shape[0]
    f32 = dict(dtype=torch.float32, device=dev)
    st = _stream_ptr(dev)
    ws = _workspace(dev, dims, n)
    P = _param_struct(N.Params, params)
    stats = torch.empty(4, **f32)
    idx = torch.empty(n, dtype=torch.int64, device=dev)
    z_q = torch.empty(n, dims.D, **f32)
    vq_stats = torch.empty(4, dtype=torch.int32, device=dev)
    x_recon = torch.empty(n, dims.F, **f32)
    acts = _empty_acts(dev, n, dims)
    sse = torch.empty(2, dtype=torch.float64, device=dev)
    A = _acts_struct(idx=idx, z_q=z_q, x_recon=x_recon, **acts)
    wsz = ws.numel()
    N.check(lib.lipvq_encode_search_fwd(C.byref(dims), C.byref(P), _p(x), n, C.byref(A), search_mode,
                                        _p(vq_stats), _p(ws), wsz, st), "lipvq_encode_search_fwd")
    N.check(lib.lipvq_quantize_decode_fwd(C.byref(dims), C.byref(P), _p(x), n, n_total, C.byref(A), _p(sse),
                                          _p(stats), _p(ws), wsz, st), "lipvq_quantize_decode_fwd")
    saved = dict(x=x, params=list(params), idx=idx, z_q=z_q, dims=dims, n_total=n_total, **acts)
    return z_q, stats, idx, vq_stats, x_recon, saved


def _bwd_args(saved, grads):
    dims, x = saved["dims"], saved["x"]
    dev = x.device
    n = x.shape[0]
    ws = _workspace(dev, dims, n)
    P = _param_struct(N.Params, saved["params"])
    G = _param_struct(N.Grads, grads)
    A = _acts_struct(**{k: saved[k] for k in N._ACT_FIELDS if k in saved})
    return dims, x, n, ws, P, G, A, _stream_ptr(dev)


def train_backward_decoder(saved, gs, grads):
    """Gradients of the decoder / to_output parameters and the COMPLETE codebook gradient (struct fields 7..13 of
    ``grads``, a list of 14 tensors in state_dict order, overwritten).  ``gs``: device float[1] = dL/dloss."""
    dims, x, n, ws, P, G, A, st = _bwd_args(saved, grads)
    N.check(N.lib().lipvq_decoder_bwd(C.byref(dims), C.byref(P), _p(x), n, saved["n_total"], _p(gs), C.byref(A),
                                      C.byref(G), _p(ws), ws.numel(), st), "lipvq_decoder_bwd")


def train_backward_encoder(saved, gs, grads):
    """Gradients of encoder.* and to_latent.* (struct fields 0..6 of ``grads``)."""
    dims, x, n, ws, P, G, A, st = _bwd_args(saved, grads)
    N.check(N.lib().lipvq_encoder_bwd(C.byref(dims), C.byref(P), _p(x), n, saved["n_total"], _p(gs), C.byref(A),
                                      C.byref(G), _p(ws), ws.numel(), st), "lipvq_encoder_bwd")


class _TokenizerFn(torch.autograd.Function):
    """forward + backward of LLFQVAE_V4.forward (v5:70-84) through the C ABI.

    inputs : x, n_total, search_mode, the 14 parameters (reference state_dict order)
    outputs: z_latent [N,D] (non-differentiable, v5:74), stats float[4] = loss/recon/commit/codebook,
             idx int64 [N] (non-differentiable)
    """

    @staticmethod
    def forward(ctx, x, n_total, search_mode, dims, *params):
        lib = N.lib()
        dev = x.device
        n = x.shape[0]
        params = _contig(params)
        f32 = dict(dtype=torch.float32, device=dev)
        need_grad = any(ctx.needs_input_grad[4:])
        if n == 0:
            stats = torch.zeros(4, **f32)
            vq_stats = torch.zeros(4, dtype=torch.int32, device=dev)
            idx = torch.empty(0, dtype=torch.int64, device=dev)
            z_q = torch.empty(0, dims.D, **f32)
            x_recon = torch.empty(0, dims.F, **f32)
            ctx.mark_non_differentiable(z_q, idx, vq_stats, x_recon)
            ctx.saved = None
            return z_q, stats, idx, vq_stats, x_recon, None
        if not need_grad:
            st = _stream_ptr(dev)
            ws = _workspace(dev, dims, n)
            P = _param_struct(N.Params, params)
            stats = torch.empty(4, **f32)
            idx = torch.empty(n, dtype=torch.int64, device=dev)
            z_q = torch.empty(n, dims.D, **f32)
            vq_stats = torch.empty(4, dtype=torch.int32, device=dev)
            x_recon = torch.empty(n, dims.F, **f32)
            aux = (None, None, None)
            if n >= AUX_MIN_ROWS and not torch.cuda.is_current_stream_capturing():
                s_aux, ef, ej = _aux(dev)
                aux = (C.c_void_p(s_aux.cuda_stream), C.c_void_p(ef.cuda_event), C.c_void_p(ej.cuda_event))
            N.check(lib.lipvq_forward_fused(C.byref(dims), C.byref(P), _p(x), n, n_total, search_mode, _p(z_q),
                                            _p(idx), _p(x_recon), _p(stats), _p(vq_stats), _p(ws), ws.numel(), st,
                                            *aux), "lipvq_forward_fused")
            ctx.mark_non_differentiable(z_q, idx, vq_stats, x_recon)
            ctx.saved = None
            return z_q, stats, idx, vq_stats, x_recon, None
        z_q, stats, idx, vq_stats, x_recon, saved = train_forward(x, n_total, search_mode, dims, params)
        ctx.saved = saved           # plain tensors created here: nothing autograd needs to version-check
        ctx.mark_non_differentiable(z_q, idx, vq_stats, x_recon, saved["z_e"])
        return z_q, stats, idx, vq_stats, x_recon, saved["z_e"]

    @staticmethod
    def backward(ctx, g_zlat, g_stats, g_idx, g_vq, g_xr, g_ze=None):
        none4 = (None, None, None, None)
        if ctx.saved is None or g_stats is None:
            return none4 + (None,) * 14
        saved = ctx.saved
        # dL/dloss lives on the device; the kernels read it there (no host sync)
        gs = g_stats[0:1].contiguous().float()
        grads = [torch.empty_like(p) for p in saved["params"]]
        train_backward_decoder(saved, gs, grads)
        train_backward_encoder(saved, gs, grads)
        out = [g if ctx.needs_input_grad[4 + i] else None for i, g in enumerate(grads)]
        return none4 + tuple(out)


def _linear_native(x, weight, bias, act):
    """act(x W^T + b) through lipvq_linear_fwd; no autograd."""
    _check_cuda(x, "x")
    x = x.contiguous()
    w = weight.detach().contiguous()
    b = bias.detach().contiguous() if bias is not None else None
    lead = x.shape[:-1]
    x2 = x.reshape(-1, x.shape[-1])
    y = torch.empty(x2.shape[0], w.shape[0], dtype=torch.float32, device=x.device)
    N.check(N.lib().lipvq_linear_fwd(_p(x2), _p(w), _p(b), _p(y), x2.shape[0], w.shape[0], w.shape[1], act,
                                     _stream_ptr(x.device)), "lipvq_linear_fwd")
    return y.reshape(*lead, w.shape[0])


class NativeLinear(nn.Linear):
    """``nn.Linear`` parameters/initialiser; calling it alone runs the native kernel without
    autograd (training goes through ``LLFQVAE_V4.forward``)."""
    _act = 0

    @torch.no_grad()
    def forward(self, x):
        return _linear_native(x, self.weight, self.bias, self._act)


class _NativeMLP(nn.Sequential):
    """``nn.Sequential(Linear, GELU, Linear, GELU)`` key layout (v5:54-59, 62-67).  Calling it on
    its own fuses each Linear with the GELU that follows and runs the native kernels, no autograd."""

    @torch.no_grad()
    def forward(self, x):
        mods = list(self)
        i = 0
        while i < len(mods):
            m = mods[i]
            if isinstance(m, nn.Linear):
                gelu = i + 1 < len(mods) and isinstance(mods[i + 1], nn.GELU)
                x = _linear_native(x, m.weight, m.bias, 1 if gelu else 0)
                i += 2 if gelu else 1
            else:
                raise LipVQNativeError("unsupported layer %r in tokenizer MLP" % (m,))
        return x


class LipschitzMLP(nn.Module):
    """Lipschitz-normalised sigmoid layer (v5:15-24): parameters W[out,in]~N(0,1), b=0, ci=1."""

    def __init__(self, in_dim, out_dim):
        super().__init__()
        self.W = nn.Parameter(torch.randn(out_dim, in_dim))
        self.b = nn.Parameter(torch.zeros(out_dim))
        self.ci = nn.Parameter(torch.ones(out_dim))

    @torch.no_grad()
    def forward(self, h):
        _check_cuda(h, "h")
        h = h.contiguous()
        D, H = self.W.shape
        n = h.shape[0]
        dev = h.device
        scale = torch.empty(D, dtype=torch.float32, device=dev)
        Wn = torch.empty(D, H, dtype=torch.float32, device=dev)
        z_e = torch.empty(n, D, dtype=torch.float32, device=dev)
        N.check(N.lib().lipvq_lipschitz_fwd(_p(h), _p(self.W.detach().contiguous()), _p(self.b.detach().contiguous()),
                                            _p(self.ci.detach().contiguous()), _p(scale), _p(Wn), _p(z_e), n, D, H,
                                            _stream_ptr(dev)), "lipvq_lipschitz_fwd")
        return z_e


class LFQQuantizer(nn.Module):
    """Codebook holder + nearest-codeword search (v5:27-48): returns ``(z_q, indices)``."""

    def __init__(self, num_codes, code_dim):
        super().__init__()
        self.num_codes = num_codes
        self.code_dim = code_dim
        self.codebook = nn.Parameter(torch.randn(num_codes, code_dim))
        nn.init.kaiming_uniform_(self.codebook)          # v5:35
        self.search_mode = N.SEARCH_TENSOR

    @torch.no_grad()
    def forward(self, z_e):
        _check_cuda(z_e, "z_e")
        z_e = z_e.contiguous()
        E = self.codebook.detach().contiguous()
        n, D = z_e.shape
        K = E.shape[0]
        dev = z_e.device
        dims = N.Dims(1, 1, 1, D, K)
        ws = _workspace(dev, dims, n)
        idx = torch.empty(n, dtype=torch.int64, device=dev)
        z_q = torch.empty(n, D, dtype=torch.float32, device=dev)
        sse = torch.zeros(1, dtype=torch.float64, device=dev)
        self.last_search_stats = torch.zeros(4, dtype=torch.int32, device=dev)
        st = _stream_ptr(dev)
        lib = N.lib()
        N.check(lib.lipvq_vq_argmin(_p(z_e), _p(E), n, D, K, self.search_mode, _p(idx), None,
                                    _p(self.last_search_stats), _p(ws), ws.numel(), st), "lipvq_vq_argmin")
        N.check(lib.lipvq_gather_loss(_p(z_e), _p(E), _p(idx), n, D, K, _p(z_q), _p(sse), _p(ws), ws.numel(), st),
                "lipvq_gather_loss")
        return z_q, idx


class LLFQVAE_V4(nn.Module):
    """Drop-in for the reference ``LLFQVAE_V4`` (v5:51-84).

    ``forward(x[N, feature_dim]) -> (z_latent[N, latent_dim] detached, loss 0-dim)`` exactly as
    consumed at robomimic/models/obs_nets.py:1335-1337.  Extra, reference-compatible surface:
    ``encode / quantize / decode`` (the submodule calls v5:71-76 spelled as methods),
    ``last_indices``, ``last_reconstruction``, ``last_loss_terms`` and ``last_search_stats`` (device tensors,
    no sync).

    Knobs are attributes, not config keys (robomimic's config is locked, train.py:494-497):
      ``search_mode``  N.SEARCH_TENSOR (default) or N.SEARCH_EXACT
      ``n_total_rows`` global row count when ``x`` is one data-parallel shard (set by dist.py)
    """

    def __init__(self, feature_dim, latent_dim, num_codes=1024, hidden_dim=128):
        super().__init__()
        self.encoder = _NativeMLP(
            NativeLinear(feature_dim, 64), nn.GELU(), NativeLinear(64, hidden_dim), nn.GELU())
        self.to_latent = LipschitzMLP(hidden_dim, latent_dim)
        self.quantizer = LFQQuantizer(num_codes, latent_dim)
        self.decoder = _NativeMLP(
            NativeLinear(latent_dim, 64), nn.GELU(), NativeLinear(64, hidden_dim), nn.GELU())
        self.to_output = NativeLinear(hidden_dim, feature_dim)
        self.feature_dim, self.latent_dim = feature_dim, latent_dim
        self.num_codes, self.hidden_dim = num_codes, hidden_dim
        self.search_mode = N.SEARCH_TENSOR
        self.n_total_rows = None
        self.last_indices = None
        self.last_reconstruction = None
        self.last_loss_terms = None
        self.last_search_stats = None
        self.last_z_e = None
        # EMA codebook state (extension, NOT in the reference): plain attributes created on first use -- the module
        # keeps the reference's "14 parameters, no buffers" surface (algo.py:323-337 serialises state_dict only)
        self.ema_cluster_size = None
        self.ema_cluster_sum = None
        # bumped by every parameter update that bypasses autograd's version counters (FlatAdamW.step, ema_update,
        # TokenizerTrainer.step, the data-parallel broadcast): PromptTokenCache keys on it
        self._param_epoch = 0

    # ------------------------------------------------------------------ EMA codebook update (extension)
    @torch.no_grad()
    def codebook_statistics(self, z_e=None, indices=None):
        """Per-code ``counts[K]`` and ``sums[K, D] = sum_{idx[n]=k} z_e[n]`` of a batch (the EMA statistics; the
        same scatter-add access pattern as the codebook gradient).  Defaults to the last training forward."""
        z_e = self.last_z_e if z_e is None else z_e
        indices = self.last_indices if indices is None else indices
        if z_e is None or indices is None:
            raise LipVQNativeError("codebook_statistics needs a training-mode forward first (or explicit z_e, indices)")
        _check_cuda(z_e, "z_e")
        z_e = z_e.contiguous()
        indices = indices.contiguous()
        K, D = self.num_codes, self.latent_dim
        n = z_e.shape[0]
        lib = N.lib()
        need = lib.lipvq_segment_sum_ws_bytes(n, D, K) if n else 0
        if need:
            # deterministic: counting sort by code + fixed reduction tree (bit-identical on every run and every rank)
            counts = torch.empty(K, dtype=torch.float32, device=z_e.device)
            sums = torch.empty(K, D, dtype=torch.float32, device=z_e.device)
            ws = torch.empty(need, dtype=torch.uint8, device=z_e.device)
            N.check(lib.lipvq_segment_sum(_p(z_e), _p(indices), None, 1.0, None, n, D, K, 0, _p(sums), _p(counts), _p(ws),
                                          ws.numel(), _stream_ptr(z_e.device)), "lipvq_segment_sum")
            return counts, sums
        counts = torch.zeros(K, dtype=torch.float32, device=z_e.device)
        sums = torch.zeros(K, D, dtype=torch.float32, device=z_e.device)
        N.check(lib.lipvq_scatter_add(_p(z_e), _p(indices), n, D, K, _p(sums), _p(counts),
                                      _stream_ptr(z_e.device)), "lipvq_scatter_add")
        return counts, sums

    @torch.no_grad()
    def ema_update(self, counts, sums, gamma=0.99, eps=1e-5):
        """``n <- g n + (1-g) counts; m <- g m + (1-g) sums; codebook <- m / laplace(n)`` in place.  ``counts`` /
        ``sums`` must already be summed over the data-parallel ranks (``DataParallelTokenizer.allreduce_codebook_stats``)
        so every replica applies the identical update.  Parity unpinned: the reference trains the codebook with AdamW."""
        E = self.quantizer.codebook.data
        if self.ema_cluster_size is None or self.ema_cluster_size.device != E.device:
            self.ema_cluster_sum = E.detach().clone()     # m_0 = E, n_0 = 1: the first update starts from the codebook
            self.ema_cluster_size = torch.ones(self.num_codes, dtype=torch.float32, device=E.device)
        ws = _workspace(E.device, self._dims(), 1)
        N.check(N.lib().lipvq_ema_update(_p(E), _p(self.ema_cluster_size), _p(self.ema_cluster_sum), _p(counts.contiguous()),
                                         _p(sums.contiguous()), self.num_codes, self.latent_dim, float(gamma), float(eps),
                                         _p(ws), ws.numel(), _stream_ptr(E.device)), "lipvq_ema_update")
        self._param_epoch += 1

    # ------------------------------------------------------------------ helpers
    def _dims(self):
        return N.Dims(self.feature_dim, 64, self.hidden_dim, self.latent_dim, self.num_codes)

    def _param_list(self):
        e, d = self.encoder, self.decoder
        return [e[0].weight, e[0].bias, e[2].weight, e[2].bias,
                self.to_latent.W, self.to_latent.b, self.to_latent.ci,
                self.quantizer.codebook,
                d[0].weight, d[0].bias, d[2].weight, d[2].bias,
                self.to_output.weight, self.to_output.bias]

    # ------------------------------------------------------------------ no-grad fast path
    # The reference's real call sizes are 80 .. 500 rows (v5:89-93), where the kernel is 45 .. 65 us and the HOST side
    # of the call decides the latency: under no_grad the forward therefore skips torch.autograd.Function, looks the 14
    # parameters up through cached ``_parameters`` slots, reuses the ctypes parameter struct while the storages are
    # unchanged, and caches the workspace size per row count.  Same C-ABI call, same outputs and ``last_*`` fields.
    def _param_slots(self):
        slots = _PSLOTS.get(self)
        if slots is None:
            e, d = self.encoder, self.decoder
            mods = [(e[0], "weight"), (e[0], "bias"), (e[2], "weight"), (e[2], "bias"),
                    (self.to_latent, "W"), (self.to_latent, "b"), (self.to_latent, "ci"), (self.quantizer, "codebook"),
                    (d[0], "weight"), (d[0], "bias"), (d[2], "weight"), (d[2], "bias"),
                    (self.to_output, "weight"), (self.to_output, "bias")]
            slots = _PSLOTS[self] = [(m._parameters, k) for m, k in mods]
        return slots

    def _forward_nograd(self, x):
        dev, n = x.device, x.shape[0]
        ps = [d[k] for d, k in self._param_slots()]
        ptrs = tuple([q.data_ptr() for q in ps])
        fast = _FAST.get(self)
        if fast is None or fast[0] != ptrs:
            if not all(q.is_contiguous() and q.is_cuda and q.dtype == torch.float32 for q in ps):
                return None                          # the general path copes (and raises where it has to)
            fast = _FAST[self] = (ptrs, _param_struct(N.Params, ps), self._dims(), {})
        _, P, dims, need_by_n = fast
        if dims.F != self.feature_dim or dims.D != self.latent_dim or dims.K != self.num_codes:
            _FAST.pop(self, None)
            return None
        stream = torch.cuda.current_stream(dev).cuda_stream
        need = need_by_n.get(n)
        lib = N.lib()
        if need is None:
            c_need = C.c_size_t(0)
            N.check(lib.lipvq_workspace_bytes(C.byref(dims), n, C.byref(c_need)), "lipvq_workspace_bytes")
            if len(need_by_n) > 64:
                need_by_n.clear()
            need = need_by_n[n] = int(c_need.value)
        key = (dev.index, stream)
        ws = _WS.get(key)
        if ws is None or ws.numel() < need:
            ws = _WS[key] = torch.empty(int(need * 1.25) + 4096, dtype=torch.uint8, device=dev)
        stats = torch.empty(4, dtype=torch.float32, device=dev)
        idx = torch.empty(n, dtype=torch.int64, device=dev)
        z_q = torch.empty((n, dims.D), dtype=torch.float32, device=dev)
        vq_stats = torch.empty(4, dtype=torch.int32, device=dev)
        x_recon = torch.empty((n, dims.F), dtype=torch.float32, device=dev)
        aux = (None, None, None)
        if n >= AUX_MIN_ROWS and not torch.cuda.is_current_stream_capturing():
            s_aux, ef, ej = _aux(dev)
            aux = (s_aux.cuda_stream, ef.cuda_event, ej.cuda_event)
        n_total = int(self.n_total_rows) if self.n_total_rows is not None else n
        rc = lib.lipvq_forward_fused(C.byref(dims), C.byref(P), x.data_ptr(), n, max(n_total, 1), int(self.search_mode),
                                     z_q.data_ptr(), idx.data_ptr(), x_recon.data_ptr(), stats.data_ptr(),
                                     vq_stats.data_ptr(), ws.data_ptr(), ws.numel(), stream, *aux)
        if rc:
            N.check(rc, "lipvq_forward_fused")
        sd = self.__dict__
        sd["last_z_e"], sd["last_indices"], sd["last_reconstruction"] = None, idx, x_recon
        sd["last_loss_terms"], sd["last_search_stats"] = stats, vq_stats
        return z_q, stats[0]

    # ------------------------------------------------------------------ reference interface
    def forward(self, x):
        _check_cuda(x, "x")
        if x.dim() != 2 or x.shape[1] != self.feature_dim:
            raise RuntimeError("expected x of shape [N, %d], got %s" % (self.feature_dim, tuple(x.shape)))
        x = x.contiguous()
        if not torch.is_grad_enabled() and x.shape[0] > 0:
            out = self._forward_nograd(x)
            if out is not None:
                return out
        n_total = int(self.n_total_rows) if self.n_total_rows is not None else x.shape[0]
        params = self._param_list()
        if not torch.is_grad_enabled():
            params = [p.detach() for p in params]
        z_latent, stats, idx, vq_stats, x_recon, z_e = _TokenizerFn.apply(
            x.detach(), max(n_total, 1), int(self.search_mode), self._dims(), *params)
        self.last_z_e = z_e                         # encoder output of this batch (training forward only), no graph
        self.last_indices = idx
        self.last_reconstruction = x_recon          # decode(z_q) of this batch (v5:75-76), no graph
        self.last_loss_terms = stats.detach()
        self.last_search_stats = vq_stats
        return z_latent, stats[0]

    @torch.no_grad()
    def forward_tokens(self, x):
        """Tokens-only no-grad forward: ``(indices int64[N], loss 0-dim)`` with the [N, latent_dim] ``z_latent`` write
        skipped (94 % of the forward's mandatory HBM traffic).  For a consumer that embeds per CODE --
        ``embed_latent`` -- instead of per row; ``last_reconstruction`` / ``last_loss_terms`` are set as by forward."""
        _check_cuda(x, "x")
        if x.dim() != 2 or x.shape[1] != self.feature_dim:
            raise RuntimeError("expected x of shape [N, %d], got %s" % (self.feature_dim, tuple(x.shape)))
        x = x.contiguous()
        dev, n = x.device, x.shape[0]
        dims = self._dims()
        n_total = max(int(self.n_total_rows) if self.n_total_rows is not None else n, 1)
        stats = torch.zeros(4, dtype=torch.float32, device=dev)
        idx = torch.empty(n, dtype=torch.int64, device=dev)
        vq_stats = torch.zeros(4, dtype=torch.int32, device=dev)
        x_recon = torch.empty(n, self.feature_dim, dtype=torch.float32, device=dev)
        if n:
            params = _contig([p.detach() for p in self._param_list()])
            P = _param_struct(N.Params, params)
            ws = _workspace(dev, dims, n)
            aux = (None, None, None)
            if n >= AUX_MIN_ROWS and not torch.cuda.is_current_stream_capturing():
                s_aux, ef, ej = _aux(dev)
                aux = (C.c_void_p(s_aux.cuda_stream), C.c_void_p(ef.cuda_event), C.c_void_p(ej.cuda_event))
            N.check(N.lib().lipvq_forward_fused(C.byref(dims), C.byref(P), _p(x), n, n_total, int(self.search_mode), None,
                                                _p(idx), _p(x_recon), _p(stats), _p(vq_stats), _p(ws), ws.numel(),
                                                _stream_ptr(dev), *aux), "lipvq_forward_fused")
        self.last_z_e, self.last_indices, self.last_reconstruction = None, idx, x_recon
        self.last_loss_terms, self.last_search_stats = stats, vq_stats
        return idx, stats[0]

    def embed_latent(self, embed: nn.Linear, indices=None):
        """``embed(z_latent)`` for the policy's input embedding (``nets["embed_encoder"]``, a Linear(latent_dim -> 512):
        robomimic/models/obs_nets.py:2537-2543, called :2582) WITHOUT the [N, latent_dim] tensor: ``z_latent`` takes
        only K distinct values (v5:47, 74), so ``embed(z_latent)[n] = (E W^T + b)[idx[n]]`` -- a [K, 512] table (K rows
        of GEMM instead of N) and a gather.  Differentiable w.r.t. ``embed``'s parameters (the codebook is detached,
        exactly like ``z_latent``); stock torch ops on the caller's side of the boundary."""
        indices = self.last_indices if indices is None else indices
        if indices is None:
            raise LipVQNativeError("embed_latent needs indices (call forward / forward_tokens first)")
        table = torch.nn.functional.linear(self.quantizer.codebook.detach(), embed.weight, embed.bias)    # [K, 512]
        return table.index_select(0, indices)

    @torch.no_grad()
    def encode(self, x):
        """z_e = to_latent(encoder(x))  (v5:71-72)"""
        return self.to_latent(self.encoder(x))

    @torch.no_grad()
    def quantize(self, z_e):
        """(z_q, indices) = quantizer(z_e)  (v5:73, 37-48)"""
        self.quantizer.search_mode = self.search_mode
        return self.quantizer(z_e)

    @torch.no_grad()
    def decode(self, z_q):
        """x_recon = to_output(decoder(z_q))  (v5:75-76)"""
        return self.to_output(self.decoder(z_q))

    @torch.no_grad()
    def tokenize(self, x):
        """indices only: the tokens of each action row."""
        return self.quantize(self.encode(x))[1]

    @torch.no_grad()
    def codes_to_latent(self, indices):
        E = self.quantizer.codebook.detach()
        z = torch.zeros(indices.shape[0], self.latent_dim, dtype=torch.float32, device=E.device)
        dummy = torch.zeros(1, dtype=torch.float64, device=E.device)
        ws = _workspace(E.device, self._dims(), indices.shape[0])
        N.check(N.lib().lipvq_gather_loss(_p(z), _p(E.contiguous()), _p(indices.contiguous()), indices.shape[0],
                                          self.latent_dim, self.num_codes, _p(z), _p(dummy), _p(ws), ws.numel(),
                                          _stream_ptr(E.device)), "lipvq_gather_loss")
        return z


def install_into_robomimic():
    """Rebind ``LLFQVAE_V4`` in robomimic's ``obs_nets`` namespace (imported there at
    obs_nets.py:47, looked up at call time :1225) so ``vq_vae_enabled`` builds this module."""
    import robomimic.models.obs_nets as obs_nets  # noqa: WPS433 (optional dependency)
    from .vqvae_classic import VQVAE
    obs_nets.LLFQVAE_V4 = LLFQVAE_V4
    obs_nets.VQVAE = VQVAE                        # the commented-out alternative at obs_nets.py:1220-1222
    return obs_nets
