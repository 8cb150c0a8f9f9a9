"""Golden vectors for the ICRT integration (BASELINE config 5), produced by the REAL reference
network: ``ICLTransformerGMMActorNetwork`` (robomimic/models/policy_nets.py:2375) with
``transformer_vq_vae_enabled=True`` around the reference ``LLFQVAE_V4`` -- built here, in the authoring container,
where /root/reference exists (the GPU box has no reference tree).  Run:  python -m tests.golden.make_icrt_golden

Stored in tests/golden/icrt_small.npz:
  sd::<key>        the network's state_dict before the step (reference key names; the vq_vae_model.* alias keys
                   duplicate nets.encoder.action_network.* and are re-created by the loader)
  obs::<k>, ctx::<k>, prompt_actions, train_actions     the synthetic batch ([B, T, ...])
  vq_loss, context_actions, idx                          the tokenizer's outputs inside the policy forward
  means, scales, logits, log_probs, action_loss          the GMM head (policy_nets.py:2540-2574) and icl.py:948-963
  vqgrad::<key>    gradients of the 14 tokenizer tensors after ``_vq_vae_loss.backward()`` (icl.py:968-969)
  vqstep::<key>    the 14 tensors after the reference's own ``optim.AdamW(lr=1e-3, weight_decay=1e-4).step()``
  polgrad::<key>   gradients of a few policy tensors after ``action_loss.backward()``
"""
import os
import sys
from collections import OrderedDict

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

OBS_SHAPES = OrderedDict(robot0_eef_pos=[3], robot0_eef_quat=[4], robot0_gripper_qpos=[2], object=[27])   # D = 36
NET_KW = dict(ac_dim=12, transformer_embed_dim=64, transformer_num_layers=2, transformer_num_heads=4,
              transformer_context_length=10, transformer_causal=False, transformer_emb_dropout=0.0,
              transformer_attn_dropout=0.0, transformer_block_output_dropout=0.0,
              transformer_nn_parameter_for_timesteps=True, transformer_vq_vae_enabled=True, num_modes=5, min_std=0.0001)
B, T = 16, 10


def build_reference_net():
    from tests.stubs import robomimic_stubs
    robomimic_stubs.install()
    import robomimic.utils.obs_utils as ObsUtils
    import robomimic.models.policy_nets as PN
    ObsUtils.initialize_obs_modality_mapping_from_dict({"low_dim": list(OBS_SHAPES.keys())})
    enc_kwargs = {"low_dim": dict(core_class=None, core_kwargs={}, obs_randomizer_class=None, obs_randomizer_kwargs={})}
    torch.manual_seed(0)
    return PN.ICLTransformerGMMActorNetwork(obs_shapes=OBS_SHAPES, encoder_kwargs=enc_kwargs, **NET_KW)


def spread_tokenizer_(tok, seed=3):
    """random init collapses every row onto one code: soften the Lipschitz layer, use both ``min``
    branches and take the codebook from the encoder's own outputs so the indices are diverse."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        tok.to_latent.W.mul_(0.08)
        tok.to_latent.ci[::3] = 5.0
        K = tok.quantizer.codebook.shape[0]
        xs = 2.0 * torch.rand(K, 12, generator=g) - 1.0
        z = tok.to_latent(tok.encoder(xs))
        tok.quantizer.codebook.copy_(z + 1e-3 * torch.randn(z.shape, generator=g))
        # a learned timestep parameter of zeros would hide ordering mistakes
    return tok


def batch(seed=1):
    g = torch.Generator().manual_seed(seed)
    obs = {k: torch.randn(B, T, *s, generator=g) for k, s in OBS_SHAPES.items()}
    ctx = {k: torch.randn(B, T, *s, generator=g) for k, s in OBS_SHAPES.items()}
    pa = 2.0 * torch.rand(B, T, 12, generator=g) - 1.0
    ta = 2.0 * torch.rand(B, T, 12, generator=g) - 1.0
    return obs, ctx, pa, ta


def main():
    net = build_reference_net()
    tok = net.nets["encoder"].action_network
    assert type(tok).__module__.endswith("backbone_lfqvae_v5"), type(tok)
    spread_tokenizer_(tok)
    with torch.no_grad():
        net.params["embed_timestep"].normal_(0.0, 0.02, generator=torch.Generator().manual_seed(5))
    net.train()
    out = {}
    for k, v in net.state_dict().items():
        if not k.startswith("vq_vae_model."):        # alias of nets.encoder.action_network.* (obs_nets.py:2421-2422)
            out["sd::" + k] = v.detach().numpy().copy()
    obs, ctx, pa, ta = batch()
    for k in OBS_SHAPES:
        out["obs::" + k] = obs[k].numpy()
        out["ctx::" + k] = ctx[k].numpy()
    out["prompt_actions"], out["train_actions"] = pa.numpy(), ta.numpy()
    cap = {}
    h = tok.quantizer.register_forward_hook(lambda m, i, o: cap.__setitem__("idx", o[1].detach().numpy().copy()))
    h2 = tok.register_forward_hook(lambda m, i, o: cap.__setitem__("z", o[0].detach().numpy().copy()))
    vq_opt = torch.optim.AdamW(net.vq_vae_model.parameters(), lr=1e-3, weight_decay=1e-4)       # icl.py:886-889
    vq_opt.zero_grad()
    dists = net.forward_train(obs_dict=obs, context_obs=ctx, actions=pa, goal_dict=None, low_noise_eval=False)
    h.remove(); h2.remove()
    vq_loss = net._vq_vae_loss
    log_probs = dists.log_prob(ta)
    action_loss = -log_probs.mean()
    out["vq_loss"] = np.float32(vq_loss.item())
    out["context_actions"], out["idx"] = cap["z"], cap["idx"]
    out["means"] = dists.component_distribution.base_dist.loc.detach().numpy()
    out["scales"] = dists.component_distribution.base_dist.scale.detach().numpy()
    out["logits"] = dists.mixture_distribution.logits.detach().numpy()
    out["log_probs"] = log_probs.detach().numpy()
    out["action_loss"] = np.float32(action_loss.item())
    vq_loss.backward()
    for k, p in tok.named_parameters():
        out["vqgrad::" + k] = p.grad.detach().numpy().copy()
    vq_opt.step()
    for k, p in tok.named_parameters():
        out["vqstep::" + k] = p.detach().numpy().copy()
    net.zero_grad()
    action_loss.backward()
    for k in ("nets.embed_encoder.weight", "nets.decoder.nets.mean.weight", "params.embed_timestep",
              "nets.transformer.nets.transformer.0.nets.attention.nets.qkv.weight"):
        out["polgrad::" + k] = dict(net.named_parameters())[k].grad.detach().numpy().copy()
    assert all(p.grad is None for p in tok.parameters()), "z_latent must carry no graph (v5:74)"
    path = os.path.join(HERE, "icrt_small.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes; unique codes", len(np.unique(cap["idx"])), "vq_loss", vq_loss.item(),
          "action_loss", action_loss.item())


if __name__ == "__main__":
    main()
