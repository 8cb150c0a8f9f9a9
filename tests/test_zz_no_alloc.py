"""lipvq_forward_small_static -- the allocation-free variant of lipvq_forward_small (static __device__ scratch inside the
library; no counterpart in the reference) -- against the existing entry
point, through the C ABI, on two calls in a row (the second with other inputs and another row count, so stale scratch
contents of the first call would show).  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
RTOL = 1e-5          # the tolerance of the other parity tests (BASELINE.json north_star)


def _setup(seed, D=208, K=1024):
    from lipvq_vae_b200 import LLFQVAE_V4
    from lipvq_vae_b200 import _native as N
    from lipvq_vae_b200.tokenizer import _param_struct
    torch.manual_seed(seed)
    dev = torch.device("cuda:0")
    m = LLFQVAE_V4(12, D, num_codes=K).to(dev)
    with torch.no_grad():   # index diversity (random init collapses every row onto one code): a milder Lipschitz layer
        m.to_latent.W.mul_(0.08)                        # and a codebook sampled from the encoder's own outputs
        m.to_latent.ci[::3] = 5.0
        z = m.encode(2.0 * torch.rand(K, 12, device=dev) - 1.0)
        m.quantizer.codebook.copy_(z + 1e-3 * torch.randn_like(z))
    params = [p.detach().contiguous() for p in m._param_list()]
    return N, m._dims(), _param_struct(N.Params, params), params, dev


def _outputs(n, D, dev):
    return dict(z=torch.full((n, D), float("nan"), device=dev), idx=torch.full((n,), -7, dtype=torch.int64, device=dev),
                xr=torch.full((n, 12), float("nan"), device=dev), loss=torch.full((4,), float("nan"), device=dev),
                stats=torch.full((4,), -1, dtype=torch.int32, device=dev))


def _call_ws(N, dims, P, x, o, dev):
    lib = N.lib()
    n = x.shape[0]
    ws = torch.empty(8 * n + 12 * ((n + 3) // 4) + 4096, dtype=torch.uint8, device=dev)
    ws.fill_(0xFF)
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    N.check(lib.lipvq_forward_small(C.byref(dims), C.byref(P), x.data_ptr(), n, n, o["z"].data_ptr(), o["idx"].data_ptr(),
                                    o["xr"].data_ptr(), o["loss"].data_ptr(), o["stats"].data_ptr(), ws.data_ptr(),
                                    ws.numel(), st), "lipvq_forward_small")


def _call_static(N, dims, P, x, o, dev):
    lib = N.lib()
    n = x.shape[0]
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    return lib.lipvq_forward_small_static(C.byref(dims), C.byref(P), x.data_ptr(), n, n, o["z"].data_ptr(),
                                          o["idx"].data_ptr(), o["xr"].data_ptr(), o["loss"].data_ptr(),
                                          o["stats"].data_ptr(), st)


def test_static_scratch_forward_equals_workspace_forward_twice_in_a_row():
    N, dims, P, params, dev = _setup(5)
    g = torch.Generator(device="cpu").manual_seed(17)
    # two calls in a row on one stream, the second smaller and with other inputs: the shared array keeps the first
    # call's contents and the entry point has to clear what must start at zero
    for n in (500, 80, 4096):
        x = torch.randn(n, 12, generator=g).to(dev)
        ref, got = _outputs(n, dims.D, dev), _outputs(n, dims.D, dev)
        _call_ws(N, dims, P, x, ref, dev)
        before = torch.cuda.memory_allocated(dev)
        rc = _call_static(N, dims, P, x, got, dev)
        assert rc == 0, N.lib().lipvq_strerror(rc)
        assert torch.cuda.memory_allocated(dev) == before          # (torch's allocator saw nothing either)
        torch.cuda.synchronize()
        assert torch.equal(got["idx"], ref["idx"])                  # integer work: bit-exact
        assert int(got["idx"].min()) >= 0 and len(torch.unique(got["idx"])) > 8
        assert torch.equal(got["z"], ref["z"])                      # z_latent = E[idx]
        for k in ("xr", "loss"):
            a, b = got[k].double(), ref[k].double()
            assert torch.isfinite(a).all()
            assert ((a - b).abs().max() / b.abs().max()).item() <= RTOL, k


def test_static_scratch_forward_rejects_sizes_beyond_its_limit():
    N, dims, P, params, dev = _setup(6, D=24, K=64)
    lib = N.lib()
    n = 4097                                                        # LIPVQ_FORWARD_SMALL_STATIC_MAX_ROWS + 1
    x = torch.randn(n, 12, device=dev)
    o = _outputs(n, dims.D, dev)
    rc = _call_static(N, dims, P, x, o, dev)
    torch.cuda.synchronize()
    assert rc == -6 and b"limit" in lib.lipvq_strerror(rc)          # LIPVQ_ELIMIT, nothing launched
    assert int(o["idx"][0]) == -7 and torch.isnan(o["loss"]).all()
    with pytest.raises(N.LipVQNativeError):
        N.check(rc, "lipvq_forward_small_static")
