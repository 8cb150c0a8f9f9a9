"""CPU-only checks of the drop-in boundary: the C-ABI library builds for sm_100a, loads without a
GPU, exports every symbol include/lipvq_b200.h declares, and the Python mirror keeps the
reference's constructor / parameter names / shapes (v5:52-68).  No compute is called."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def native():
    from lipvq_vae_b200 import _native
    _native.build_native()
    return _native


def test_header_symbols_exported(native):
    hdr = open(os.path.join(ROOT, "include", "lipvq_b200.h")).read()
    declared = set(re.findall(r"\b(lipvq_[a-z0-9_]+)\s*\(", hdr))
    declared = {d for d in declared if not d.endswith("_t")}
    assert declared == set(native.SIGNATURES), declared ^ set(native.SIGNATURES)
    h = ctypes.CDLL(native.LIB_PATH)
    for name in declared:
        assert hasattr(h, name), name
    h.lipvq_abi_version.restype = ctypes.c_int
    assert h.lipvq_abi_version() == 6
    h.lipvq_strerror.restype = ctypes.c_char_p
    assert h.lipvq_strerror(-4) == b"workspace too small"
    assert b"limit" in h.lipvq_strerror(-6)          # LIPVQ_ELIMIT: the allocation-free entry point's size cap


def test_workspace_query_and_arg_checks(native):
    lib = native.lib()
    d = native.Dims(12, 64, 128, 208, 1024)
    out = ctypes.c_size_t(0)
    assert lib.lipvq_workspace_bytes(ctypes.byref(d), 2560, ctypes.byref(out)) == 0
    assert out.value > 2560 * 208 * 4
    bad = native.Dims(12, 64, 128, 0, 1024)
    assert lib.lipvq_workspace_bytes(ctypes.byref(bad), 2560, ctypes.byref(out)) == -1   # LIPVQ_ESHAPE
    # null / misaligned pointers are rejected before anything touches the device
    assert lib.lipvq_vq_argmin(None, None, 16, 64, 128, 0, None, None, None, None, 0, None) == -2
    assert lib.lipvq_vq_argmin(None, None, 16, 0, 128, 0, None, None, None, None, 0, None) == -1
    assert lib.lipvq_scatter_add(8, 16, 4, 8, 8, 32, None, None) == -2                    # misaligned src


def test_module_mirrors_reference_interface():
    from lipvq_vae_b200 import LLFQVAE_V4
    from oracle import lipvq_oracle as orc
    torch.manual_seed(0)
    m = LLFQVAE_V4(12, 208)
    sd = m.state_dict()
    assert tuple(sd.keys()) == orc.PARAM_NAMES                       # names AND order (v5:52-68)
    for k, shp in orc.param_shapes(12, 208).items():
        assert tuple(sd[k].shape) == shp, k
    assert sum(p.numel() for p in m.parameters()) == 272428
    assert len(list(m.buffers())) == 0
    # initialisers (v5:18-20, 32-35)
    assert torch.all(sd["to_latent.b"] == 0) and torch.all(sd["to_latent.ci"] == 1)
    assert abs(sd["to_latent.W"].std().item() - 1.0) < 0.05
    assert sd["quantizer.codebook"].abs().max().item() <= (6.0 / 208) ** 0.5 + 1e-6
    m2 = LLFQVAE_V4(12, 24, num_codes=64, hidden_dim=32)
    assert m2.state_dict()["encoder.2.weight"].shape == (32, 64)
    # optimiser construction used by the reference algo (icl.py:886-889)
    torch.optim.AdamW(m.parameters(), lr=1e-3, weight_decay=1e-4)
    import copy
    import io
    copy.deepcopy(m)
    torch.save(m.state_dict(), io.BytesIO())


def test_no_cpu_fallback():
    from lipvq_vae_b200 import LLFQVAE_V4, LipVQNativeError
    m = LLFQVAE_V4(12, 16, num_codes=8)
    with pytest.raises(LipVQNativeError):
        m(torch.randn(3, 12))
    with pytest.raises(LipVQNativeError):
        m.encode(torch.randn(3, 12))


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "lipvq_vae_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                assert "oracle" not in open(os.path.join(dp, f)).read().replace("oracle.", "oracle."), f
