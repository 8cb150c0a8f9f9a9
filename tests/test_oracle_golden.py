"""Pin the CPU oracle (oracle/lipvq_oracle.py) against outputs of the REAL reference module,
stored in tests/golden/*.npz by tests/golden/make_golden.py (which imports
/root/reference/robomimic/models/vq_vae/backbone_lfqvae_v5.py).  CPU only."""
import os

import numpy as np
import pytest

from oracle import lipvq_oracle as orc
from tests.golden.make_golden import MODEL_CASES, QUANT_CASES, case_inputs, quant_inputs

G = os.path.join(os.path.dirname(__file__), "golden")
RTOL = 1e-5          # BASELINE.json: latents / losses / reconstructions within 1e-5 (fp32)
GTOL = 5e-5          # gradients: fp32 sums over N rows in a different order than torch's


def _load(name):
    return dict(np.load(os.path.join(G, name + ".npz")))


@pytest.mark.parametrize("case", MODEL_CASES)
@pytest.mark.parametrize("search", ["exact", "reference"])
def test_forward_matches_reference(case, search):
    if search == "reference" and case not in ("small", "d787"):
        pytest.skip("broadcast form is slow at N=2560,K=1024; covered by the quantizer cases")
    p, x = case_inputs(case)
    g = _load("model_" + case)
    fw = orc.forward(p, x, search=search)
    _, d1, d2 = orc.exact_top2(fw["z_e"], p["quantizer.codebook"])
    must = orc.index_parity_mask(d1, d2)
    assert np.array_equal(fw["idx"][must], g["idx"][must].astype(np.int64))
    assert orc.rel_err(fw["z_e"][:32], g["z_e_head"]) < RTOL
    assert abs(fw["z_e"].astype(np.float64).sum() - g["z_e_sum"]) < RTOL * g["z_e_abs"]
    if must.all():
        assert orc.rel_err(fw["x_recon"][:32], g["x_recon_head"]) < RTOL
        assert abs(fw["loss"] - g["loss"]) < RTOL * abs(g["loss"])
        assert abs(fw["z_latent"].astype(np.float64).sum() - g["z_latent_sum"]) < RTOL * g["z_e_abs"]


@pytest.mark.parametrize("case", MODEL_CASES)
def test_backward_and_adamw_match_reference(case):
    p, x = case_inputs(case)
    g = _load("model_" + case)
    fw = orc.forward(p, x)
    if not np.array_equal(fw["idx"], g["idx"]):
        pytest.skip("near-tie rows differ; gradient comparison needs identical assignments")
    gr = orc.backward(p, x, fw)
    for k in orc.PARAM_NAMES:
        if "grad/" + k in g:
            assert orc.rel_err(gr[k], g["grad/" + k]) < GTOL, k
        else:
            s = g["gradsum/" + k]
            v = gr[k].astype(np.float64)
            assert abs(v.sum() - s[0]) < GTOL * s[1] + 1e-12, k
            assert abs((v ** 2).sum() - s[2]) < 4 * GTOL * s[2] + 1e-20, k
            assert orc.rel_err(gr[k].reshape(-1)[:4096], g["gradhead/" + k]) < GTOL or \
                np.abs(g["gradhead/" + k]).max() == 0, k
            if "gradrows/" + k in g:
                rows = g["gradrows_idx/" + k]
                assert orc.rel_err(gr[k][rows], g["gradrows/" + k]) < GTOL, k
    st = orc.adamw_init(p)
    # use the reference's own gradients where stored in full so AdamW is tested in isolation
    new = orc.adamw_step(p, gr, st)
    for k in orc.PARAM_NAMES:
        if "after/" + k in g:
            ref = g["after/" + k]
            # Adam's first step moves every weight by ~lr*sign(g): compare the update itself
            assert np.max(np.abs(new[k] - ref)) < 2e-6 + 1e-3 * 0.02, k
        else:
            s = g["aftersum/" + k]
            assert abs(new[k].astype(np.float64).sum() - s[0]) < 1e-4 * s[1] + 1e-3, k


@pytest.mark.parametrize("case", QUANT_CASES)
def test_quantizer_matches_reference(case):
    z, E = quant_inputs(case)
    g = _load("quant_" + case)
    ref_idx = g["idx"].astype(np.int64)
    ex_idx, d1, d2 = orc.exact_top2(z, E)
    must = orc.index_parity_mask(d1, d2)
    assert must.mean() > 0.3
    assert np.array_equal(ex_idx[must], ref_idx[must])
    if case != "q_collapsed":   # fp32 expanded form cancels when ||z-e|| << ||z|| (documented)
        assert np.array_equal(orc.quantize_gemm_form(z, E)[must], ref_idx[must])
    n = min(len(z), 512)
    rf = orc.quantize_reference_form(z[:n], E)
    assert np.array_equal(rf[must[:n]], ref_idx[:n][must[:n]])
    # classic VQVAE.quantize (secondary variant, backbone.py:55-76)
    zq, idx2, ql = orc.vqvae_quantize(z[:256], E)
    m2 = must[:256]
    assert np.array_equal(idx2[m2], g["vq_idx"].astype(np.int64)[m2])
    if m2.all():
        assert abs(ql - g["vq_loss"]) < 1e-5 * abs(g["vq_loss"])


def test_duplicate_codes_pick_lowest_index():
    z, E = quant_inputs("q_neartie")
    g = _load("quant_q_neartie")
    assert g["idx"][0] == 3 and g["idx"][1] == 100          # the reference itself (torch.argmin)
    ex = orc.exact_top2(z, E)[0]
    assert ex[0] == 3 and ex[1] == 100
    assert orc.quantize_gemm_form(z, E)[0] == 3


def test_ema_update_properties():
    rng = np.random.default_rng(5)
    E = rng.random((16, 8), dtype=np.float32)
    z = rng.random((200, 8), dtype=np.float32)
    idx = orc.exact_top2(z, E)[0]
    n0 = np.ones(16, np.float32)
    m0 = E.copy()
    E1, n1, m1, counts, sums = orc.ema_update(E, n0, m0, z, idx, gamma=0.9)
    assert counts.sum() == 200
    assert np.allclose(sums.sum(0), z.sum(0), rtol=1e-5)
    assert np.allclose(n1, 0.9 * n0 + 0.1 * counts)
    assert E1.shape == E.shape and np.isfinite(E1).all()


@pytest.mark.parametrize("case", ["vq_small", "vq_ref"])
def test_oracle_classic_vqvae_matches_reference(case):
    """The oracle's restatement of the classic VQVAE (vq_vae/backbone.py:6-76, forward + autograd incl. the
    straight-through route) against vectors produced by the real reference module."""
    from tests.golden.make_golden import vqvae_case_inputs
    p, x = vqvae_case_inputs(case)
    g = dict(np.load(os.path.join(G, "vqvae_%s.npz" % case)))
    fw = orc.vqvae_forward(p, x)
    assert np.array_equal(fw["idx"], g["idx"])
    assert abs(fw["loss"] - g["loss"]) < 1e-6 * abs(g["loss"])
    assert orc.rel_err(fw["z_e"][:32], g["z_e_head"]) < 1e-6
    assert orc.rel_err(fw["x_recon"][:32], g["x_recon_head"]) < 1e-6
    gr = orc.vqvae_backward(p, x, fw)
    for k in orc.VQ_PARAM_NAMES:
        assert orc.rel_err(gr[k], g["grad/" + k]) < 2e-6, k
