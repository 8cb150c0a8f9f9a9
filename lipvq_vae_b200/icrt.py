"""The caller side of the tokenizer, for BASELINE config 5: an in-context transformer policy (ICRT)
with the LipVQ tokenizer inline, mirroring -- module for module and state_dict key for key -- what the reference
builds as ``ICLTransformerGMMActorNetwork`` (robomimic/models/policy_nets.py:2375-2590) on top of
``ICL_MIMO_Transformer`` (robomimic/models/obs_nets.py:2336-2611), ``ICLObservationGroupEncoder``
(obs_nets.py:1120-1372, low-dim observations: flatten + concatenate, no learned obs net) and ``GPT_Backbone``
(robomimic/models/transformers.py:321-439), plus the training step of ``ICLTransformer_GMM``
(robomimic/algo/icl.py:891-974).

Why it exists: the reference tree cannot travel to the GPU box, so the drop-in can only be shown end to end there
against a policy that consumes ``z_latent`` exactly like the reference's.  The policy part is STOCK torch (Linear,
LayerNorm, scaled_dot_product_attention, torch.distributions) -- SURVEY 8f-4: "stock torch SDPA is adequate" -- and
its results are pinned on golden vectors produced by the real reference network (tests/golden/make_icrt_golden.py:
same state_dict loaded strictly, same inputs -> same GMM parameters, same ``_vq_vae_loss``, same tokenizer gradients).
Only the tokenizer (``action_network``) is this repo's product; ``tokenizer_cls`` selects it.
"""
from __future__ import annotations

import math
from collections import OrderedDict

import torch
import torch.distributions as D
import torch.nn as nn
import torch.nn.functional as F

from .tokenizer import LLFQVAE_V4

__all__ = ["ICRTPolicy", "ICRTTrainer", "synthetic_icrt_batch"]


class _LowDimObsEncoder(nn.Module):
    """``ObservationEncoder`` with ``core_class=None`` for every key (obs_nets.py:240-420): each key is flattened and
    the keys are concatenated in ``obs_shapes`` order; no parameters, ``feature_activation=None``
    (obs_nets.py:2413-2419)."""

    def __init__(self, obs_shapes):
        super().__init__()
        self.obs_shapes = OrderedDict((k, tuple(v)) for k, v in obs_shapes.items())

    def output_dim(self):
        return int(sum(math.prod(s) for s in self.obs_shapes.values()))

    def forward(self, obs):
        return torch.cat([obs[k].reshape(obs[k].shape[0], -1) for k in self.obs_shapes], dim=-1)


class _GroupEncoder(nn.Module):
    """``ICLObservationGroupEncoder`` (obs_nets.py:1120-1372), vq_vae branch (:1219-1227, :1335-1337)."""

    def __init__(self, obs_shapes, action_input_shape, tokenizer_cls):
        super().__init__()
        self.nets = nn.ModuleDict({"obs": _LowDimObsEncoder(obs_shapes)})
        d = self.nets["obs"].output_dim()
        self.action_network = tokenizer_cls(feature_dim=action_input_shape, latent_dim=d)     # obs_nets.py:1225-1227
        self._vq_vae_loss = None

    def output_dim(self):
        return self.nets["obs"].output_dim()

    def forward(self, obs, prompt_obs, prompt_actions):
        o = self.nets["obs"](obs)
        c = self.nets["obs"](prompt_obs)
        context_actions, loss = self.action_network(prompt_actions)                          # obs_nets.py:1336
        self._vq_vae_loss = loss                                                             # obs_nets.py:1337
        return o, c, context_actions


class _SelfAttention(nn.Module):
    """transformers.py:100-222: bias-free qkv projection, softmax(QK^T/sqrt(d)) V, output projection."""

    def __init__(self, embed_dim, num_heads, context_length, causal, attn_dropout, output_dropout):
        super().__init__()
        self.num_heads, self.attn_p, self.causal = num_heads, attn_dropout, causal
        self.nets = nn.ModuleDict({
            "qkv": nn.Linear(embed_dim, 3 * embed_dim, bias=False),
            "attn_dropout": nn.Dropout(attn_dropout),
            "output": nn.Linear(embed_dim, embed_dim),
            "output_dropout": nn.Dropout(output_dropout),
        })
        mask = torch.tril(torch.ones(context_length, context_length)).view(1, 1, context_length, context_length)
        if not causal:
            mask[:] = 1
        self.register_buffer("mask", mask)

    def forward(self, x):
        B, T, Dm = x.shape
        q, k, v = torch.chunk(self.nets["qkv"](x), 3, dim=-1)
        sh = lambda t: t.view(B, T, self.num_heads, Dm // self.num_heads).transpose(1, 2)
        y = F.scaled_dot_product_attention(sh(q), sh(k), sh(v), dropout_p=self.attn_p if self.training else 0.0,
                                           is_causal=self.causal)
        y = y.transpose(1, 2).contiguous().view(B, T, Dm)
        return self.nets["output_dropout"](self.nets["output"](y))


class _Block(nn.Module):
    """transformers.py:225-318."""

    def __init__(self, embed_dim, num_heads, context_length, causal, attn_dropout, output_dropout):
        super().__init__()
        self.nets = nn.ModuleDict({
            "attention": _SelfAttention(embed_dim, num_heads, context_length, causal, attn_dropout, output_dropout),
            "mlp": nn.Sequential(nn.Linear(embed_dim, 4 * embed_dim), nn.GELU(), nn.Linear(4 * embed_dim, embed_dim),
                                 nn.Dropout(output_dropout)),
            "ln1": nn.LayerNorm(embed_dim),
            "ln2": nn.LayerNorm(embed_dim),
        })

    def forward(self, x):
        x = x + self.nets["attention"](self.nets["ln1"](x))
        return x + self.nets["mlp"](self.nets["ln2"](x))


class _GPTBackbone(nn.Module):
    """transformers.py:321-439 (normal(0, 0.02) weights, zero biases, unit LayerNorms)."""

    def __init__(self, embed_dim, context_length, causal, attn_dropout, block_output_dropout, num_layers, num_heads):
        super().__init__()
        self.nets = nn.ModuleDict({
            "transformer": nn.Sequential(*[_Block(embed_dim, num_heads, context_length, causal, attn_dropout,
                                                  block_output_dropout) for _ in range(num_layers)]),
            "output_ln": nn.LayerNorm(embed_dim),
        })
        self.apply(self._init_weights)

    @staticmethod
    def _init_weights(m):
        if isinstance(m, (nn.Linear, nn.Embedding)):
            m.weight.data.normal_(mean=0.0, std=0.02)
            if isinstance(m, nn.Linear) and m.bias is not None:
                m.bias.data.zero_()
        elif isinstance(m, nn.LayerNorm):
            m.bias.data.zero_()
            m.weight.data.fill_(1.0)

    def forward(self, x):
        return self.nets["output_ln"](self.nets["transformer"](x))


class _Decoder(nn.Module):
    """``ObservationDecoder`` (obs_nets.py:700-771): one Linear per output key, reshaped to the key's shape."""

    def __init__(self, decode_shapes, input_feat_dim):
        super().__init__()
        self.obs_shapes = OrderedDict(decode_shapes)
        self.nets = nn.ModuleDict({k: nn.Linear(input_feat_dim, int(math.prod(s))) for k, s in self.obs_shapes.items()})

    def forward(self, feats):
        return {k: self.nets[k](feats).reshape(-1, *self.obs_shapes[k]) for k in self.obs_shapes}


class ICRTPolicy(nn.Module):
    """``ICLTransformerGMMActorNetwork`` for low-dim observations with ``transformer_vq_vae_enabled=True``.

    ``forward_train(obs_dict, context_obs, actions)`` takes ``[B, T, ...]`` tensors and returns the GMM over
    ``[B, T]`` exactly as policy_nets.py:2518-2574; ``_vq_vae_loss`` is the tokenizer loss of the call
    (obs_nets.py:2576-2577).  ``state_dict()`` carries the reference's keys, including the ``vq_vae_model.*`` alias
    of ``nets.encoder.action_network.*`` (obs_nets.py:2421-2422)."""

    def __init__(self, obs_shapes, ac_dim=12, embed_dim=512, num_layers=6, num_heads=8, context_length=10,
                 causal=False, emb_dropout=0.1, attn_dropout=0.1, block_output_dropout=0.1,
                 nn_parameter_for_timesteps=True, num_modes=5, min_std=0.01, tokenizer_cls=LLFQVAE_V4,
                 fuse_action_embedding=False):
        super().__init__()
        # opt-in: embed the prompt tokens per CODE ([K, 512] table + gather, ``embed_latent``) instead of
        # running embed_encoder over the [B*T, latent_dim] ``z_latent``; under no_grad the tokenizer then skips the
        # ``z_latent`` write altogether (``forward_tokens``).  Same numbers up to fp32 summation order.
        self.fuse_action_embedding = fuse_action_embedding
        self.ac_dim, self.num_modes, self.min_std = ac_dim, num_modes, min_std
        self.context_length, self.embed_dim = context_length, embed_dim
        self.nn_parameter_for_timesteps = nn_parameter_for_timesteps
        self.nets = nn.ModuleDict()
        self.params = nn.ParameterDict()
        self.nets["encoder"] = _GroupEncoder(obs_shapes, 12, tokenizer_cls)        # action_input_shape=12, obs_nets.py:2411
        self.vq_vae_model = self.nets["encoder"].action_network                    # obs_nets.py:2421-2422
        d = self.nets["encoder"].output_dim()
        self.nets["embed_encoder"] = nn.Linear(d, embed_dim)
        if nn_parameter_for_timesteps:
            self.params["embed_timestep"] = nn.Parameter(torch.zeros(1, context_length, embed_dim))
        else:
            self.nets["embed_timestep"] = nn.Embedding(context_length, embed_dim)
        self.nets["embed_ln"] = nn.LayerNorm(embed_dim)
        self.nets["embed_drop"] = nn.Dropout(emb_dropout)
        self.nets["transformer"] = _GPTBackbone(embed_dim, 3 * context_length, causal, attn_dropout,
                                                block_output_dropout, num_layers, num_heads)
        self.nets["decoder"] = _Decoder(OrderedDict(mean=(num_modes, ac_dim), scale=(num_modes, ac_dim),
                                                    logits=(num_modes,)), embed_dim)
        self._vq_vae_loss = None

    # obs_nets.py:2524-2543
    def input_embedding(self, x, pre_embedded=False):
        e = x if pre_embedded else self.nets["embed_encoder"](x)
        if self.nn_parameter_for_timesteps:
            t = self.params["embed_timestep"]
        else:
            t = self.nets["embed_timestep"](torch.arange(e.shape[1], device=e.device)).unsqueeze(0)
        return self.nets["embed_drop"](self.nets["embed_ln"](e + t))

    def forward_train(self, obs_dict, context_obs, actions, goal_dict=None, low_noise_eval=False):
        B, T = actions.shape[:2]
        flat = lambda dct: {k: v.reshape(B * T, *v.shape[2:]) for k, v in dct.items()}     # icl_time_distributed
        enc = self.nets["encoder"]
        tok = enc.action_network
        if self.fuse_action_embedding and hasattr(tok, "embed_latent"):
            o, c = enc.nets["obs"](flat(obs_dict)), enc.nets["obs"](flat(context_obs))
            rows = actions.reshape(B * T, -1)
            if torch.is_grad_enabled():
                _, loss = tok(rows)
                idx = tok.last_indices
            else:
                idx, loss = tok.forward_tokens(rows)               # no [B*T, latent_dim] tensor at all
            enc._vq_vae_loss = self._vq_vae_loss = loss
            ae = self.input_embedding(tok.embed_latent(self.nets["embed_encoder"], idx).reshape(B, T, -1), pre_embedded=True)
            o, c = o.reshape(B, T, -1), c.reshape(B, T, -1)
        else:
            o, c, a = enc(flat(obs_dict), flat(context_obs), actions.reshape(B * T, -1))
            self._vq_vae_loss = enc._vq_vae_loss
            o, c, a = (t.reshape(B, T, -1) for t in (o, c, a))
            ae = self.input_embedding(a)
        oe, ce = self.input_embedding(o), self.input_embedding(c)
        ctx = torch.stack([ce, ae], dim=2).view(B, 2 * T, -1)                                # obs_nets.py:2586-2591
        h = self.nets["transformer"](torch.cat([ctx, oe], dim=1))[:, -T:]                     # :2594-2603
        out = self.nets["decoder"](h.reshape(B * T, -1))
        means = torch.tanh(out["mean"].reshape(B, T, self.num_modes, self.ac_dim))           # policy_nets.py:2545-2547
        if low_noise_eval and not self.training:
            scales = torch.ones_like(means) * 1e-4
        else:
            scales = F.softplus(out["scale"].reshape(B, T, self.num_modes, self.ac_dim)) + self.min_std
        comp = D.Independent(D.Normal(loc=means, scale=scales), 1)
        mix = D.Categorical(logits=out["logits"].reshape(B, T, self.num_modes))
        return D.MixtureSameFamily(mixture_distribution=mix, component_distribution=comp)

    def policy_parameters(self):
        """every parameter that is NOT the tokenizer's (the tokenizer has its own AdamW, icl.py:885-889)."""
        tok = {id(p) for p in self.vq_vae_model.parameters()}
        return [p for p in self.parameters() if id(p) not in tok]


class ICRTTrainer:
    """One training iteration of ``ICLTransformer_GMM`` (icl.py:891-974 + torch_utils.py:217-218): the batch is split
    into a prompt half and a query half, the tokenizer loss runs its own backward + AdamW(1e-3, wd 1e-4) step BEFORE
    the policy's negative-log-likelihood backward + AdamW step.  Data parallel: every rank works on its own
    sequences; the tokenizer's and the policy's gradients are averaged with one flat all-reduce each."""

    def __init__(self, policy: ICRTPolicy, lr=1e-4, vq_lr=1e-3, vq_weight_decay=1e-4, process_group=None):
        import torch.distributed as dist
        self.policy = policy
        self.vq_optimizer = torch.optim.AdamW(policy.vq_vae_model.parameters(), lr=vq_lr, weight_decay=vq_weight_decay)
        self.pol_params = policy.policy_parameters()
        self.optimizer = torch.optim.AdamW(self.pol_params, lr=lr, weight_decay=0.01)
        self.pg = process_group
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1

    def _allreduce_mean(self, params):
        if self.world == 1:
            return
        import torch.distributed as dist
        grads = [p.grad for p in params if p.grad is not None]
        flat = torch._utils._flatten_dense_tensors(grads)
        dist.all_reduce(flat, group=self.pg)
        flat.div_(self.world)
        for g, f in zip(grads, torch._utils._unflatten_dense_tensors(flat, grads)):
            g.copy_(f)

    def step(self, batch):
        """batch: {"obs": {key: [2B, T, ...]}, "actions": [2B, T, 12]}; first half = prompt, second half = query."""
        mid = batch["actions"].shape[0] // 2                                       # icl.py:905
        ctx_obs = {k: v[:mid] for k, v in batch["obs"].items()}
        trn_obs = {k: v[mid:] for k, v in batch["obs"].items()}
        ctx_act, trn_act = batch["actions"][:mid], batch["actions"][mid:]
        self.vq_optimizer.zero_grad()                                              # icl.py:913-914
        dists = self.policy.forward_train(obs_dict=trn_obs, context_obs=ctx_obs, actions=ctx_act)
        vq_loss = self.policy._vq_vae_loss
        action_loss = -dists.log_prob(trn_act).mean()                              # icl.py:948, 963
        vq_loss.backward()                                                         # icl.py:968-970
        self._allreduce_mean(list(self.policy.vq_vae_model.parameters()))
        self.vq_optimizer.step()
        self.optimizer.zero_grad()                                                 # torch_utils.py:211-218
        action_loss.backward()
        self._allreduce_mean(self.pol_params)
        self.optimizer.step()
        m = self.policy.vq_vae_model
        if hasattr(m, "_param_epoch"):
            m._param_epoch += 1
        return action_loss.detach(), vq_loss.detach()


def synthetic_icrt_batch(obs_shapes, n_seq, T=10, ac_dim=12, device="cpu", seed=0):
    """Synthetic prompt/query sequences of BASELINE config 5: ``n_seq`` sequences (first half prompts, second half
    queries, icl.py:905-911), observations ~ N(0,1), actions uniform in [-1, 1]."""
    g = torch.Generator().manual_seed(seed)
    obs = {k: torch.randn(n_seq, T, *s, generator=g).to(device) for k, s in obs_shapes.items()}
    act = (2.0 * torch.rand(n_seq, T, ac_dim, generator=g) - 1.0).to(device)
    return {"obs": obs, "actions": act}
