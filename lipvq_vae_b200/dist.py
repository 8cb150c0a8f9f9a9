"""Data-parallel plumbing for the tokenizer: rows shard across ranks, the 14
parameters are replicated, and the only collective is ONE all-reduce per training step over a
flat fp32 bucket (gradients, or EMA counts/sums).  The reference has no distributed code
(single process on cuda:0, robomimic/utils/torch_utils.py:48-50); this wrapper adds the one
strategy the path supports.  Works on any ``torch.distributed`` backend (NCCL on the GPUs; the
host-side logic is covered with gloo on CPU in tests/test_dist_cpu.py).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

CHUNK_LEN = 10  # rows per action chunk (obs_nets.py:1298)


def _bump_param_epoch(model):
    """Parameters were rewritten through raw pointers / ``.data`` (no autograd ``_version`` bump): tell the caches
    that key on parameter identity (``cache.PromptTokenCache``) that their entries are stale."""
    model._param_epoch = getattr(model, "_param_epoch", 0) + 1


def shard_rows(n_total: int, rank: int, world: int, align: int = CHUNK_LEN):
    """Contiguous, chunk-aligned [start, stop) row range of ``rank``; the ranges tile [0, n_total)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world %r/%r" % (rank, world))
    units = (n_total + align - 1) // align
    base, rem = divmod(units, world)
    u0 = rank * base + min(rank, rem)
    u1 = u0 + base + (1 if rank < rem else 0)
    return min(u0 * align, n_total), min(u1 * align, n_total)


class FlatBucket:
    """One contiguous fp32 buffer viewed as a list of tensors (a single collective per step)."""

    def __init__(self, shapes, device, extra=0):
        self.shapes = [tuple(s) for s in shapes]
        self.sizes = [int(torch.Size(s).numel()) for s in self.shapes]
        self.flat = torch.zeros(sum(self.sizes) + extra, dtype=torch.float32, device=device)
        self.views, off = [], 0
        for s, n in zip(self.shapes, self.sizes):
            self.views.append(self.flat[off:off + n].view(s))
            off += n
        self.extra = self.flat[off:]

    def pack(self, tensors):
        for v, t in zip(self.views, tensors):
            if t is None:
                v.zero_()
            else:
                v.copy_(t)

    def unpack_into(self, tensors):
        for v, t in zip(self.views, tensors):
            if t is not None:
                t.copy_(v)


class DataParallelTokenizer:
    """Wraps an ``LLFQVAE_V4``: call it on the rank's row shard, then ``allreduce_grads()``.

    The kernels take the GLOBAL row count (``n_total_rows``), so each rank's loss and gradients
    already carry the 1/N_global factor and the collective is a plain SUM: the result equals the
    single-process gradient on the concatenated batch up to fp32 summation order.
    """

    def __init__(self, model, process_group=None, broadcast_params=True):
        self.model = model
        self.pg = process_group
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(process_group) if dist.is_initialized() else 0
        self._bucket = None
        self._stat_bucket = None
        if broadcast_params and self.world > 1:
            for p in model.parameters():
                dist.broadcast(p.data, src=0, group=process_group)
            _bump_param_epoch(model)

    # ------------------------------------------------------------------ forward
    def global_rows(self, n_local: int, device) -> int:
        if self.world == 1:
            return n_local
        t = torch.tensor([n_local], dtype=torch.int64, device=device)
        dist.all_reduce(t, group=self.pg)
        return int(t.item())

    def __call__(self, x_local, n_total=None):
        """forward on this rank's rows; ``n_total`` (global rows) skips the count all-reduce."""
        if n_total is None:
            n_total = self.global_rows(x_local.shape[0], x_local.device)
        prev = self.model.n_total_rows
        self.model.n_total_rows = n_total            # only for the duration of this call (later direct calls of the
        try:                                         # model must see their own row count again)
            return self.model(x_local)
        finally:
            self.model.n_total_rows = prev

    # ------------------------------------------------------------------ collectives
    def _params(self):
        return [p for p in self.model.parameters()]

    def allreduce_grads(self, loss=None):
        """SUM all gradients (and optionally the scalar loss, same bucket) across ranks."""
        params = self._params()
        if self._bucket is None:
            self._bucket = FlatBucket([p.shape for p in params], params[0].device, extra=1)
        b = self._bucket
        b.pack([p.grad for p in params])
        if loss is not None:
            b.extra.copy_(loss.detach().reshape(1))
        else:
            b.extra.zero_()
        if self.world > 1:
            dist.all_reduce(b.flat, group=self.pg)
        for p, v in zip(params, b.views):
            if p.grad is None:
                p.grad = v.clone()
            else:
                p.grad.copy_(v)
        return b.extra[0] if loss is not None else None

    def allreduce_codebook_stats(self, counts, sums):
        """EMA mode: SUM counts[K] and sums[K,D] in one collective (in place)."""
        if self._stat_bucket is None or self._stat_bucket.shapes != [tuple(counts.shape), tuple(sums.shape)]:
            self._stat_bucket = FlatBucket([counts.shape, sums.shape], counts.device)
        b = self._stat_bucket
        b.pack([counts, sums])
        if self.world > 1:
            dist.all_reduce(b.flat, group=self.pg)
        b.unpack_into([counts, sums])
        return counts, sums


class FlatAdamW:
    """``torch.optim.AdamW(model.parameters(), lr, weight_decay)`` (robomimic/algo/icl.py:886-889) as ONE kernel
    launch per step: the 14 parameters are re-pointed at views of one flat fp32 buffer, the moments are flat too,
    and ``lipvq_adamw_step`` updates everything at once from a flat gradient (the all-reduce bucket)."""

    def __init__(self, model, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-4):
        self.model = model
        self.params = [p for p in model.parameters()]
        dev = self.params[0].device
        self.numel = sum(p.numel() for p in self.params)
        self.flat = torch.empty(self.numel, dtype=torch.float32, device=dev)
        off = 0
        for p in self.params:
            n = p.numel()
            self.flat[off:off + n].copy_(p.data.reshape(-1))
            p.data = self.flat[off:off + n].view(p.shape)      # same values, storage now inside the flat buffer
            off += n
        self.exp_avg = torch.zeros_like(self.flat)
        self.exp_avg_sq = torch.zeros_like(self.flat)
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.steps = 0

    def step(self, flat_grad):
        """flat_grad: fp32 [numel] in ``model.parameters()`` order (FlatBucket.flat[:numel])."""
        import ctypes as C
        from . import _native as N
        if flat_grad.numel() < self.numel or not flat_grad.is_cuda:
            raise ValueError("flat_grad must be a CUDA tensor with at least %d elements" % self.numel)
        self.steps += 1
        st = C.c_void_p(torch.cuda.current_stream(self.flat.device).cuda_stream)
        N.check(N.lib().lipvq_adamw_step(self.flat.data_ptr(), flat_grad.data_ptr(), self.exp_avg.data_ptr(),
                                         self.exp_avg_sq.data_ptr(), self.numel, self.steps, self.lr, self.betas[0],
                                         self.betas[1], self.eps, self.weight_decay, st), "lipvq_adamw_step")
        _bump_param_epoch(self.model)            # written through raw pointers: no autograd version bump


class TokenizerTrainer:
    """One data-parallel training step of the tokenizer (BASELINE config 3): forward + backward on this rank's
    rows, the gradient all-reduce, ONE fused AdamW launch.  Mirrors what ``ICLTransformer_GMM`` does with
    ``self.vq_vae_model`` each iteration (icl.py:913-914, 968-970).

    The step calls the C ABI directly (``tokenizer.train_forward`` / ``train_backward_*``), no autograd graph: the
    backward kernels write the 14 gradients straight into views of the flat all-reduce bucket (``p.grad`` ARE those
    views), so nothing is packed or unpacked around the collective: ONE all-reduce of the 1.09 MB bucket (+ the loss
    scalar) after the backward, then the fused AdamW.

    ``overlap=True`` splits the collective in two and issues the first half under the backward:
      1. decoder backward -> struct fields 7..13 (codebook, decoder.*, to_output.*: 80 % of the bucket, one contiguous
         slice) + the loss scalar are all-reduced on the communication stream while
      2. the encoder backward (the N-row GEMMs, ~45 % of the step) runs; its slice (fields 0..6) follows.
    Measured no better and therefore not the default (65 536 chunks per GPU, 5 warm-up steps: 2.21 ms per step against
    2.17 ms for the single all-reduce on 8 x B200, 2.17 against 2.16 on 2): the collective costs ~40 us when it runs alone,
    and the encoder-side backward kernels it would hide under are persistent, one CTA per SM with a static row partition,
    so an NCCL kernel that holds an SM while they run delays their last CTA.  (A first measurement with 2 warm-up steps
    showed 3.6 ms for the split schedule: NCCL's first calls on the communication stream were inside the timed steps.)"""

    def __init__(self, model, process_group=None, lr=1e-3, weight_decay=1e-4, ema_codebook=False, ema_gamma=0.99,
                 overlap=False, concurrent_backward=True):
        self.dp = DataParallelTokenizer(model, process_group)
        self.opt = FlatAdamW(model, lr=lr, weight_decay=weight_decay)
        self.model = model
        self.overlap = overlap
        # The decoder-side backward (per-code GEMMs + the per-code segment sums) and the encoder-side backward (the
        # N-row GEMMs) depend only on the forward and write disjoint gradient slices: with concurrent_backward they run
        # on two streams (each with its own workspace), the small decoder-side kernels filling the SM and HBM slack the
        # persistent encoder-side kernels leave (measured at 65 536 chunks, 1 GPU: 2.94 -> 2.52 ms per step).
        self.concurrent_backward = concurrent_backward
        self._side = None
        # EMA mode (extension): the codebook follows the EMA of its clusters instead of AdamW --
        # its gradient is zeroed in the bucket and counts[K] + sums[K,D] are all-reduced in one more collective
        self.ema_codebook, self.ema_gamma = ema_codebook, ema_gamma
        params = self.opt.params
        self._cb_index = [i for i, p in enumerate(params) if p is model.quantizer.codebook][0]
        if [id(p) for p in params] != [id(p) for p in model._param_list()]:
            raise RuntimeError("parameter order differs from the C ABI's struct order")
        # gradient bucket: 14 views in struct order + 1 float for the loss; the decoder-side slice starts at the codebook
        self.bucket = FlatBucket([p.shape for p in params], params[0].device, extra=1)
        self.dp._bucket = self.bucket
        self._split = sum(self.bucket.sizes[:self._cb_index])
        self._gs = torch.ones(1, dtype=torch.float32, device=params[0].device)      # dL/dloss = 1

    def _all_reduce(self, t, async_op):
        if self.dp.world > 1:
            return dist.all_reduce(t, group=self.dp.pg, async_op=async_op)
        return None

    def step(self, x_local, n_total=None):
        from . import tokenizer as T
        m, b = self.model, self.bucket
        if not x_local.is_cuda:
            raise T.LipVQNativeError("TokenizerTrainer.step needs a CUDA batch: there is no CPU path")
        x = x_local.detach().contiguous()
        n = x.shape[0]
        if n_total is None:
            n_total = self.dp.global_rows(n, x.device)
        if n == 0:
            raise ValueError("every rank needs at least one row (shard_rows keeps shards non-empty for N >= world)")
        params = [p.detach() for p in self.opt.params]
        z_q, stats, idx, vq_stats, x_recon, saved = T.train_forward(x, max(int(n_total), 1), int(m.search_mode),
                                                                   m._dims(), params)
        m.last_z_e, m.last_indices, m.last_reconstruction = saved["z_e"], idx, x_recon
        m.last_loss_terms, m.last_search_stats = stats, vq_stats
        if self.ema_codebook:
            counts, sums = m.codebook_statistics()
            self.dp.allreduce_codebook_stats(counts, sums)
        b.extra.copy_(stats[0:1])
        if self.concurrent_backward:
            main = torch.cuda.current_stream(x.device)
            if self._side is None:
                self._side = torch.cuda.Stream(device=x.device)
            self._side.wait_stream(main)
            with torch.cuda.stream(self._side):
                T.train_backward_decoder(saved, self._gs, b.views)
                w1 = self._all_reduce(b.flat[self._split:], True) if self.overlap else None
            T.train_backward_encoder(saved, self._gs, b.views)
            main.wait_stream(self._side)
        else:
            T.train_backward_decoder(saved, self._gs, b.views)
            w1 = self._all_reduce(b.flat[self._split:], True) if self.overlap else None     # codebook .. to_output + loss
            T.train_backward_encoder(saved, self._gs, b.views)
        if self.overlap:
            w0 = self._all_reduce(b.flat[:self._split], True)
            for w in (w1, w0):
                if w is not None:
                    w.wait()                      # stream-level wait: the host does not block
        else:
            self._all_reduce(b.flat, False)
        for p, v in zip(self.opt.params, b.views):
            p.grad = v                            # the reduced (global-mean) gradients, no copy
        if self.ema_codebook:
            cb = m.quantizer.codebook
            keep = cb.data.clone()                              # AdamW must not move (or decay) the codebook
            b.views[self._cb_index].zero_()
            self.opt.step(b.flat)
            cb.data.copy_(keep)
            m.ema_update(counts, sums, gamma=self.ema_gamma)
        else:
            self.opt.step(b.flat)
        _bump_param_epoch(m)
        return z_q, b.extra[0].clone()          # (a copy: the bucket is overwritten by the next step)
