"""The drop-in seam, executed (VERDICT r1 missing 6).  These tests need the reference tree and run in
the authoring container only (skipped where /root/reference is absent, e.g. on the GPU box):

* ``install_into_robomimic()`` really makes robomimic's own ``ICLObservationGroupEncoder`` (obs_nets.py:1219-1227)
  build THIS repo's ``LLFQVAE_V4`` under ``vq_vae_enabled``, and the resulting network's ``state_dict`` interchanges
  strictly, both ways, with one built around the reference class (algo.py:323-337);
* the policy mirror used on the GPU box (lipvq_vae_b200/icrt.py) reproduces the real reference network: same
  state_dict keys, and -- with the REFERENCE tokenizer class plugged in -- the same GMM parameters, log-probs,
  tokenizer loss and policy gradients as the committed golden vectors.
The golden file itself is checked for freshness against a live run of the reference network."""
import importlib.util
import os
from collections import OrderedDict

import numpy as np
import pytest
import torch

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "robomimic")), reason="reference tree not present")
G = os.path.join(os.path.dirname(__file__), "golden", "icrt_small.npz")


def _golden():
    g = dict(np.load(G))
    sd = OrderedDict((k[4:], torch.from_numpy(v.copy())) for k, v in g.items() if k.startswith("sd::"))
    for k in list(sd):
        if k.startswith("nets.encoder.action_network."):
            sd["vq_vae_model." + k[len("nets.encoder.action_network."):]] = sd[k]
    return g, sd


def _ref_tokenizer_cls():
    spec = importlib.util.spec_from_file_location(
        "ref_v5", os.path.join(REF, "robomimic/models/vq_vae/backbone_lfqvae_v5.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.LLFQVAE_V4


def test_install_into_robomimic_builds_the_replacement_and_state_dicts_interchange():
    from tests.golden import make_icrt_golden as mk
    ref_net = mk.build_reference_net()
    import robomimic.models.obs_nets as obs_nets
    ref_cls = obs_nets.LLFQVAE_V4
    import lipvq_vae_b200
    try:
        lipvq_vae_b200.install_into_robomimic()
        assert obs_nets.LLFQVAE_V4 is lipvq_vae_b200.LLFQVAE_V4
        new_net = mk.build_reference_net()                      # robomimic's own constructor chain, vq_vae_enabled=True
    finally:
        obs_nets.LLFQVAE_V4 = ref_cls
    tok = new_net.nets["encoder"].action_network
    assert type(tok) is lipvq_vae_b200.LLFQVAE_V4 and new_net.vq_vae_model is tok
    assert (tok.feature_dim, tok.latent_dim, tok.num_codes) == (12, 36, 1024)
    assert list(new_net.state_dict().keys()) == list(ref_net.state_dict().keys())
    assert [tuple(v.shape) for v in new_net.state_dict().values()] == [tuple(v.shape) for v in ref_net.state_dict().values()]
    new_net.load_state_dict(ref_net.state_dict(), strict=True)          # reference checkpoint -> replacement
    ref_net2 = mk.build_reference_net()
    ref_net2.load_state_dict(new_net.state_dict(), strict=True)         # replacement checkpoint -> reference
    for (k, a), b in zip(ref_net.state_dict().items(), ref_net2.state_dict().values()):
        assert torch.equal(a, b), k
    # the two optimisers of the algo see the same tensors (icl.py:886-889, algo.py:206-209)
    assert len(list(new_net.vq_vae_model.parameters())) == 14
    assert sum(p.numel() for p in new_net.parameters()) == sum(p.numel() for p in ref_net.parameters())
    # PromptTokenCache around the module that is CALLED (obs_nets.py:1336) keeps the checkpoint keys
    new_net.nets["encoder"].action_network = lipvq_vae_b200.PromptTokenCache(tok)
    assert list(new_net.state_dict().keys()) == list(ref_net.state_dict().keys())
    new_net.load_state_dict(ref_net.state_dict(), strict=True)


def test_policy_mirror_equals_reference_network_on_golden():
    from lipvq_vae_b200.icrt import ICRTPolicy
    from tests.golden import make_icrt_golden as mk
    g, sd = _golden()
    kw = mk.NET_KW
    pol = ICRTPolicy(mk.OBS_SHAPES, ac_dim=12, embed_dim=kw["transformer_embed_dim"], num_layers=kw["transformer_num_layers"],
                     num_heads=kw["transformer_num_heads"], context_length=10, causal=False, emb_dropout=0.0,
                     attn_dropout=0.0, block_output_dropout=0.0, nn_parameter_for_timesteps=True, num_modes=5,
                     min_std=kw["min_std"], tokenizer_cls=_ref_tokenizer_cls())
    ref_net = mk.build_reference_net()
    assert list(pol.state_dict().keys()) == list(ref_net.state_dict().keys())
    pol.load_state_dict(sd, strict=True)
    pol.train()
    obs = {k: torch.from_numpy(g["obs::" + k]) for k in mk.OBS_SHAPES}
    ctx = {k: torch.from_numpy(g["ctx::" + k]) for k in mk.OBS_SHAPES}
    pa, ta = torch.from_numpy(g["prompt_actions"]), torch.from_numpy(g["train_actions"])
    d = pol.forward_train(obs_dict=obs, context_obs=ctx, actions=pa)
    assert abs(pol._vq_vae_loss.item() - float(g["vq_loss"])) < 1e-6
    for name, got in (("means", d.component_distribution.base_dist.loc), ("scales", d.component_distribution.base_dist.scale),
                      ("logits", d.mixture_distribution.logits), ("log_probs", d.log_prob(ta))):
        ref = g[name]
        assert np.abs(got.detach().numpy() - ref).max() <= 2e-5 * max(1.0, np.abs(ref).max()), name
    loss = -d.log_prob(ta).mean()
    loss.backward()
    named = dict(pol.named_parameters())
    for k in [k for k in g if k.startswith("polgrad::")]:
        ref = g[k]
        got = named[k[9:]].grad.numpy()
        assert np.abs(got - ref).max() <= 1e-4 * np.abs(ref).max() + 1e-7, k


def test_golden_is_fresh():
    """the committed vectors equal a live run of the reference network (guards against a stale file)."""
    from tests.golden import make_icrt_golden as mk
    g, _ = _golden()
    net = mk.build_reference_net()
    mk.spread_tokenizer_(net.nets["encoder"].action_network)
    assert np.array_equal(net.state_dict()["nets.encoder.action_network.quantizer.codebook"].numpy(),
                          g["sd::nets.encoder.action_network.quantizer.codebook"])
