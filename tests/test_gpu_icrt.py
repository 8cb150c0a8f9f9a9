"""ICRT end to end on the GPU (BASELINE config 5) against golden vectors produced by the REAL
reference network (tests/golden/make_icrt_golden.py: ``ICLTransformerGMMActorNetwork`` +
reference ``LLFQVAE_V4``, run in the authoring container).  The policy mirror (lipvq_vae_b200/icrt.py, pinned to
the reference network in tests/test_icrt_integration_cpu.py) is built around THIS repo's tokenizer, the reference's
state_dict is loaded strictly, and one ``ICLTransformer_GMM`` iteration (icl.py:891-974) is replayed:
forward_train -> ``_vq_vae_loss.backward()`` -> the reference's own ``optim.AdamW(1e-3, wd 1e-4).step()``."""
import os
from collections import OrderedDict

import numpy as np
import pytest
import torch

from oracle import lipvq_oracle as orc

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden", "icrt_small.npz")
OBS_SHAPES = OrderedDict(robot0_eef_pos=[3], robot0_eef_quat=[4], robot0_gripper_qpos=[2], object=[27])
TOK = "nets.encoder.action_network."


def _golden():
    g = dict(np.load(G))
    sd = OrderedDict((k[4:], torch.from_numpy(v.copy())) for k, v in g.items() if k.startswith("sd::"))
    for k in list(sd):
        if k.startswith(TOK):
            sd["vq_vae_model." + k[len(TOK):]] = sd[k]
    return g, sd


def _policy(sd, dev="cuda:0"):
    from lipvq_vae_b200 import LLFQVAE_V4
    from lipvq_vae_b200.icrt import ICRTPolicy
    pol = ICRTPolicy(OBS_SHAPES, ac_dim=12, embed_dim=64, num_layers=2, num_heads=4, context_length=10, causal=False,
                     emb_dropout=0.0, attn_dropout=0.0, block_output_dropout=0.0, nn_parameter_for_timesteps=True,
                     num_modes=5, min_std=0.0001, tokenizer_cls=LLFQVAE_V4).to(dev)
    pol.load_state_dict(sd, strict=True)              # the reference checkpoint, unmodified (algo.py:337)
    return pol.train()


@pytest.mark.parametrize("mode", [1, 0])
def test_icrt_iteration_matches_reference_network(mode):
    g, sd = _golden()
    pol = _policy(sd)
    tok = pol.vq_vae_model
    tok.search_mode = mode
    dev = "cuda:0"
    obs = {k: torch.from_numpy(g["obs::" + k]).to(dev) for k in OBS_SHAPES}
    ctx = {k: torch.from_numpy(g["ctx::" + k]).to(dev) for k in OBS_SHAPES}
    pa, ta = torch.from_numpy(g["prompt_actions"]).to(dev), torch.from_numpy(g["train_actions"]).to(dev)
    vq_opt = torch.optim.AdamW(pol.vq_vae_model.parameters(), lr=1e-3, weight_decay=1e-4)      # icl.py:886-889, as is
    vq_opt.zero_grad()
    d = pol.forward_train(obs_dict=obs, context_obs=ctx, actions=pa)
    # --- the tokenizer inside the policy forward (obs_nets.py:1335-1337)
    pre = "sd::" + TOK
    p_tok = {k[len(pre):]: g[k] for k in g if k.startswith(pre)}
    z_e = orc.encode(p_tok, g["prompt_actions"].reshape(-1, 12))
    _, d1, d2 = orc.exact_top2(z_e, p_tok["quantizer.codebook"])
    must = orc.index_parity_mask(d1, d2)
    idx = tok.last_indices.cpu().numpy()
    assert np.array_equal(idx[must], g["idx"][must])
    same = np.array_equal(idx, g["idx"])
    assert same or must.mean() > 0.99
    if same:
        assert abs(pol._vq_vae_loss.item() - float(g["vq_loss"])) < 1e-5 * float(g["vq_loss"])
        # --- the GMM head (policy_nets.py:2540-2574) and the log-likelihood (icl.py:948)
        for name, got in (("means", d.component_distribution.base_dist.loc), ("scales", d.component_distribution.base_dist.scale),
                          ("logits", d.mixture_distribution.logits), ("log_probs", d.log_prob(ta))):
            ref = g[name]
            assert np.abs(got.detach().cpu().numpy() - ref).max() <= 1e-4 * max(1.0, np.abs(ref).max()), name
    assert not tok.last_indices.requires_grad
    # --- tokenizer loss backward + the reference's AdamW step, then the policy loss backward (icl.py:968-970)
    action_loss = -d.log_prob(ta).mean()
    pol._vq_vae_loss.backward()
    if same:
        for k, p in tok.named_parameters():
            assert orc.rel_err(p.grad.cpu().numpy(), g["vqgrad::" + k]) < 5e-5, k
    vq_opt.step()
    if same:
        for k, p in tok.named_parameters():
            ref = g["vqstep::" + k]
            dd = np.abs(p.detach().cpu().numpy() - ref)
            # AdamW's first step moves every element by +-lr (sign of the gradient): elements whose gradient is at
            # rounding-noise level may flip, all others agree to fp32 rounding
            assert dd.max() <= 2.02e-3 and np.quantile(dd, 0.99) < 2e-6, (k, dd.max())
    pol.zero_grad()
    action_loss.backward()
    assert all(p.grad is None for p in tok.parameters()), "z_latent must carry no graph (v5:74)"
    if same:
        named = dict(pol.named_parameters())
        for k in [k for k in g if k.startswith("polgrad::")]:
            ref = g[k]
            got = named[k[9:]].grad.cpu().numpy()
            assert np.abs(got - ref).max() <= 2e-4 * np.abs(ref).max() + 1e-7, k


def test_icrt_trainer_steps_and_cache_wiring():
    """ICRTTrainer (the icl.py:891-974 iteration) runs, losses are finite and decrease on a fixed batch; wrapping the
    CALLED module (``nets.encoder.action_network``) in PromptTokenCache leaves training untouched and replays rollout
    prompts."""
    from lipvq_vae_b200 import PromptTokenCache
    from lipvq_vae_b200.icrt import ICRTTrainer, synthetic_icrt_batch
    g, sd = _golden()
    pol = _policy(sd)
    keys = list(pol.state_dict().keys())
    pol.nets["encoder"].action_network = PromptTokenCache(pol.vq_vae_model)
    assert list(pol.state_dict().keys()) == keys
    tr = ICRTTrainer(pol, lr=1e-3)
    batch = synthetic_icrt_batch(OBS_SHAPES, 32, device="cuda:0", seed=3)
    losses = [tuple(t.item() for t in tr.step(batch)) for _ in range(8)]
    assert all(np.isfinite(l).all() for l in losses)
    assert losses[-1][0] < losses[0][0] and losses[-1][1] < losses[0][1], losses
    pol.eval()
    cache = pol.nets["encoder"].action_network
    h0 = cache.hits
    with torch.no_grad():
        obs = {k: v[:4] for k, v in batch["obs"].items()}
        for _ in range(3):                      # a rollout re-tokenises the same prompt every environment step
            pol.forward_train(obs_dict=obs, context_obs=obs, actions=batch["actions"][:4].contiguous())
    assert cache.hits - h0 >= 1


def test_fused_action_embedding_equals_plain_path():
    """embed_encoder(z_latent) as a [K, 512] per-code table + gather (``embed_latent``), with the tokenizer
    in tokens-only mode under no_grad (``forward_tokens``: no z_latent write) -- same GMM, same tokenizer loss, same
    gradients on embed_encoder as the reference-shaped path."""
    g, sd = _golden()
    dev = "cuda:0"
    obs = {k: torch.from_numpy(g["obs::" + k]).to(dev) for k in OBS_SHAPES}
    ctx = {k: torch.from_numpy(g["ctx::" + k]).to(dev) for k in OBS_SHAPES}
    pa, ta = torch.from_numpy(g["prompt_actions"]).to(dev), torch.from_numpy(g["train_actions"]).to(dev)
    outs = {}
    for fuse in (False, True):
        pol = _policy(sd)
        pol.fuse_action_embedding = fuse
        d = pol.forward_train(obs_dict=obs, context_obs=ctx, actions=pa)
        nll = -d.log_prob(ta).mean()
        nll.backward()
        with torch.no_grad():
            d0 = pol.forward_train(obs_dict=obs, context_obs=ctx, actions=pa)
            lp0 = d0.log_prob(ta)
            vq0 = pol._vq_vae_loss.clone()
        outs[fuse] = (d.component_distribution.base_dist.loc.detach(), d.mixture_distribution.logits.detach(),
                      pol.nets["embed_encoder"].weight.grad.clone(), lp0, vq0, pol.vq_vae_model.last_indices.clone())
    a, b = outs[False], outs[True]
    assert torch.equal(a[5], b[5])
    for i, name in enumerate(("means", "logits", "embed_encoder.weight.grad", "log_probs (no_grad)")):
        scale = max(1.0, a[i].abs().max().item())
        assert (a[i] - b[i]).abs().max().item() <= 2e-5 * scale, name
    assert abs(a[4].item() - b[4].item()) <= 1e-6 * abs(a[4].item())
    # tokens-only mode returns the same tokens and loss as the full forward
    tok = _policy(sd).vq_vae_model
    rows = pa.reshape(-1, 12).contiguous()
    with torch.no_grad():
        z, loss = tok(rows)
        idx_full = tok.last_indices.clone()
        idx, loss_t = tok.forward_tokens(rows)
    assert torch.equal(idx, idx_full) and abs(loss.item() - loss_t.item()) <= 1e-6 * abs(loss.item())
    emb = torch.nn.Linear(36, 64).to(dev)
    assert (tok.embed_latent(emb, idx) - emb(z)).abs().max().item() < 1e-5
