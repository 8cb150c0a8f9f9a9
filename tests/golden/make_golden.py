#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ by running the REAL reference module.

Run in the authoring container only (needs /root/reference, CPU torch):
    python tests/golden/make_golden.py

The reference (`robomimic/models/vq_vae/backbone_lfqvae_v5.py`, `.../backbone.py`) is imported
by file path -- it depends only on torch -- and is fed parameters and inputs drawn from numpy's
`default_rng` (stable across versions), so the tests can regenerate the *inputs* anywhere and only
the reference's *outputs* are stored.  Nothing here is needed at test time on the GPU box.
"""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import lipvq_oracle as orc  # noqa: E402  (input generators only)

REF = "/root/reference/robomimic/models/vq_vae"


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


_REF_MODS = {}


def _ref(name):
    """Import the reference module lazily: only the generator (run in the authoring container)
    touches /root/reference; the tests import this file for the input generators alone."""
    if name not in _REF_MODS:
        fn = {"v5": "backbone_lfqvae_v5.py", "vq": "backbone.py"}[name]
        _REF_MODS[name] = _load("ref_" + name, os.path.join(REF, fn))
    return _REF_MODS[name]


# ------------------------------------------------------------------ shared input generators
def case_inputs(name):
    """Deterministic (params, x) for each model-level case; imported by the tests too."""
    if name == "c1_init":          # BASELINE config 1: random init, 256 chunks x 10 rows
        p = orc.init_params(12, 208, 1024, 128, seed=0)
        x = np.random.default_rng(100).standard_normal((2560, 12)).astype(np.float32)
    elif name == "c1_spread":      # spread codebook + both branches of the Lipschitz min
        p = orc.init_params(12, 208, 1024, 128, seed=1)
        rng = np.random.default_rng(101)
        x = (2.0 * rng.random((2560, 12)) - 1.0).astype(np.float32)
        p["to_latent.W"] = (0.05 * p["to_latent.W"]).astype(np.float32)
        ci = p["to_latent.ci"].copy()
        ci[::2] = 6.0              # softplus(6) > sum|W| on these rows -> scale = 1 branch
        p["to_latent.ci"] = ci
        z = orc.encode(p, (2.0 * rng.random((1024, 12)) - 1.0).astype(np.float32))
        p["quantizer.codebook"] = (z + 1e-3 * rng.standard_normal(z.shape)).astype(np.float32)
    elif name == "small":
        p = orc.init_params(12, 24, 64, 128, seed=2)
        rng = np.random.default_rng(102)
        x = rng.standard_normal((80, 12)).astype(np.float32)
        p["to_latent.W"] = (0.1 * p["to_latent.W"]).astype(np.float32)
        p["to_latent.ci"][1::3] = 4.0
        z = orc.encode(p, rng.standard_normal((64, 12)).astype(np.float32))
        p["quantizer.codebook"] = (z + 1e-2 * rng.standard_normal(z.shape)).astype(np.float32)
    elif name == "d787":           # low-dim + lang_emb width measured in SURVEY section 8c
        p = orc.init_params(12, 787, 128, 128, seed=3)
        rng = np.random.default_rng(103)
        x = rng.standard_normal((80, 12)).astype(np.float32)
        z = orc.encode(p, rng.standard_normal((128, 12)).astype(np.float32))
        p["quantizer.codebook"] = (z + 1e-4 * rng.standard_normal(z.shape)).astype(np.float32)
    else:
        raise KeyError(name)
    return p, x


def quant_inputs(name):
    """Deterministic (z, E) for quantizer-only cases."""
    if name == "q_uniform64":      # BASELINE config 2 shape, small N
        rng = np.random.default_rng(200)
        return rng.random((4096, 64), dtype=np.float32), rng.random((1024, 64), dtype=np.float32)
    if name == "q_neartie":        # adversarial near-ties at relative gaps 1e-3 .. 1e-8 + exact duplicates
        rng = np.random.default_rng(201)
        D, K = 64, 256
        E = rng.random((K, D), dtype=np.float32)
        z = rng.random((512, D), dtype=np.float32)
        for r in range(512):
            a = int(rng.integers(0, K - 1))
            b = int(rng.integers(a + 1, K))
            mid = 0.5 * (E[a] + E[b])
            gap = 10.0 ** (-3 - (r % 6))
            z[r] = mid + (gap * (r % 2 * 2 - 1)) * (E[b] - E[a])   # just off the bisector
        E[7] = E[3]                # exact duplicates: lowest index must win
        E[200] = E[100]
        z[0] = E[3] + 1e-3
        z[1] = E[100] - 1e-3
        return z.astype(np.float32), E
    if name == "q_k4096_d128":
        rng = np.random.default_rng(202)
        return rng.random((1000, 128), dtype=np.float32), rng.random((4096, 128), dtype=np.float32)
    if name == "q_d256_k128":
        rng = np.random.default_rng(203)
        return (rng.standard_normal((777, 256)).astype(np.float32),
                rng.standard_normal((128, 256)).astype(np.float32))
    if name == "q_collapsed":      # reference-init regime: every row maps to one code
        rng = np.random.default_rng(204)
        z = (0.5 + 0.02 * rng.standard_normal((2000, 208))).astype(np.float32)
        b = np.sqrt(6.0 / 208)
        return z, rng.uniform(-b, b, (1024, 208)).astype(np.float32)
    raise KeyError(name)


VQVAE_CASES = ("vq_small", "vq_ref")


def vqvae_case_inputs(name):
    """Deterministic (params, x) for the classic-VQVAE cases (vq:6-76)."""
    if name == "vq_small":
        F, D, K, n, seed = 12, 24, 64, 80, 300
    elif name == "vq_ref":         # the widths of the commented-out call site (obs_nets.py:1220-1222): K=128 default
        F, D, K, n, seed = 12, 208, 128, 2560, 301
    else:
        raise KeyError(name)
    p = orc.vqvae_init_params(F, D, K, seed=seed)
    rng = np.random.default_rng(seed + 1000)
    x = rng.random((n, F)).astype(np.float32)                # the output ReLU can only reconstruct x >= 0
    for k in ("encoder.0.bias", "encoder.2.bias", "encoder.4.bias"):
        p[k] = (p[k] + 0.05).astype(np.float32)              # keep a healthy share of ReLU units alive
    z = orc.vqvae_forward(p, rng.random((K, F)).astype(np.float32))["z_e"]
    p["embedding.weight"] = (z + 1e-3 * rng.standard_normal(z.shape)).astype(np.float32)   # spread codebook
    return p, x


def run_vqvae_case(name):
    p, x = vqvae_case_inputs(name)
    K, D = p["embedding.weight"].shape
    m = _ref("vq").VQVAE(x.shape[1], D, num_embeddings=K)
    m.load_state_dict({k: torch.from_numpy(v.copy()) for k, v in p.items()}, strict=True)
    assert tuple(m.state_dict().keys()) == orc.VQ_PARAM_NAMES
    xt = torch.from_numpy(x)
    z_latent, loss = m(xt)                                                    # vq:38-53
    loss.backward()
    with torch.no_grad():
        z_e = m.encoder(xt)
        d = (z_e.unsqueeze(1) - m.embedding.weight).pow(2).sum(-1)
        idx = torch.argmin(d, dim=1)
        x_rec = m.decoder(m.embedding(idx))
    out = dict(loss=np.float64(loss.item()), idx=idx.numpy().astype(np.int32), z_e_head=z_e[:32].numpy(),
               x_recon_head=x_rec[:32].numpy(), z_latent_sum=np.float64(z_latent.double().sum().item()),
               z_latent_abs=np.float64(z_latent.double().abs().sum().item()))
    for k, v in m.named_parameters():
        out["grad/" + k] = v.grad.detach().numpy().copy()
    np.savez_compressed(os.path.join(HERE, f"vqvae_{name}.npz"), **out)
    print(name, "loss", loss.item(), "unique idx", len(np.unique(out["idx"])))


MODEL_CASES = ("c1_init", "c1_spread", "small", "d787")
QUANT_CASES = ("q_uniform64", "q_neartie", "q_k4096_d128", "q_d256_k128", "q_collapsed")


def _ref_model(p):
    F = p["encoder.0.weight"].shape[1]
    K, D = p["quantizer.codebook"].shape
    H = p["encoder.2.weight"].shape[0]
    m = _ref("v5").LLFQVAE_V4(F, D, num_codes=K, hidden_dim=H)
    m.load_state_dict({k: torch.from_numpy(v.copy()) for k, v in p.items()}, strict=True)
    return m.float()


def run_model_case(name):
    p, x = case_inputs(name)
    m = _ref_model(p)
    xt = torch.from_numpy(x)
    opt = torch.optim.AdamW(m.parameters(), lr=1e-3, weight_decay=1e-4)      # icl.py:886-889
    opt.zero_grad()
    z_latent, loss = m(xt)                                                     # v5:70-84
    with torch.no_grad():
        z_e = m.to_latent(m.encoder(xt))
        z_q, idx = m.quantizer(z_e)
        x_rec = m.to_output(m.decoder(z_q))
    assert torch.equal(z_q, z_latent) and not z_latent.requires_grad
    loss.backward()                                                            # icl.py:969
    grads = {k: v.grad.detach().numpy().copy() for k, v in m.named_parameters()}
    opt.step()                                                                 # icl.py:970
    after = {k: v.detach().numpy().copy() for k, v in m.named_parameters()}
    out = dict(loss=np.float64(loss.item()), idx=idx.numpy().astype(np.int32),
               z_e_head=z_e[:32].numpy(), x_recon_head=x_rec[:32].numpy(),
               z_e_sum=np.float64(z_e.double().sum().item()),
               z_e_abs=np.float64(z_e.double().abs().sum().item()),
               z_latent_sum=np.float64(z_latent.double().sum().item()),
               x_recon_sum=np.float64(x_rec.double().sum().item()),
               x_recon_abs=np.float64(x_rec.double().abs().sum().item()))
    full = name in ("small",)
    for k, v in grads.items():
        if full or v.size <= 30000:
            out["grad/" + k] = v
        else:
            out["gradsum/" + k] = np.array([v.astype(np.float64).sum(), np.abs(v.astype(np.float64)).sum(),
                                            (v.astype(np.float64) ** 2).sum()])
            out["gradhead/" + k] = v.reshape(-1)[:4096].copy()
            if k == "quantizer.codebook":
                nz = np.flatnonzero(np.abs(v).sum(1) > 0)[:64]
                out["gradrows_idx/" + k] = nz.astype(np.int32)
                out["gradrows/" + k] = v[nz]
    for k, v in after.items():
        if full:
            out["after/" + k] = v
        else:
            out["aftersum/" + k] = np.array([v.astype(np.float64).sum(), np.abs(v.astype(np.float64)).sum()])
            out["afterhead/" + k] = v.reshape(-1)[:2048].copy()
    np.savez_compressed(os.path.join(HERE, f"model_{name}.npz"), **out)
    print(name, "loss", loss.item(), "unique idx", len(np.unique(out["idx"])))


def run_quant_case(name):
    z, E = quant_inputs(name)
    q = _ref("v5").LFQQuantizer(E.shape[0], E.shape[1])
    with torch.no_grad():
        q.codebook.copy_(torch.from_numpy(E))
        idx = np.empty(len(z), np.int32)
        for s in range(0, len(z), 256):                                      # rows are independent
            _, i = q(torch.from_numpy(z[s:s + 256]))
            idx[s:s + 256] = i.numpy()
        # classic VQVAE.quantize (vq:55-76) on the same data
        m = _ref("vq").VQVAE(4, E.shape[1], num_embeddings=E.shape[0])
        m.embedding.weight.copy_(torch.from_numpy(E))
        zq2, qloss = m.quantize(torch.from_numpy(z[:256]))
        d = (torch.from_numpy(z[:256]).unsqueeze(1) - m.embedding.weight).pow(2).sum(-1)
        idx2 = torch.argmin(d, dim=1).numpy().astype(np.int32)
    np.savez_compressed(os.path.join(HERE, f"quant_{name}.npz"), idx=idx, vq_idx=idx2,
                        vq_loss=np.float64(qloss.item()), vq_zq_head=zq2[:8].numpy())
    print(name, "unique idx", len(np.unique(idx)))


if __name__ == "__main__":
    torch.manual_seed(0)
    torch.set_num_threads(8)
    if len(sys.argv) > 1 and sys.argv[1] == "vqvae":      # only the classic-VQVAE vectors
        for c in VQVAE_CASES:
            run_vqvae_case(c)
        sys.exit(0)
    for c in VQVAE_CASES:
        run_vqvae_case(c)
    for c in MODEL_CASES:
        run_model_case(c)
    for c in QUANT_CASES:
        run_quant_case(c)
