#!/usr/bin/env python
"""Time the encoder / decoder entry points alone (CUDA events), N rows, D, K from argv."""
import ctypes as C
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from lipvq_vae_b200 import LLFQVAE_V4, _native as N
from lipvq_vae_b200.tokenizer import _workspace, _param_struct, _acts_struct, _p

n, D, K = (int(a) for a in (sys.argv[1:4] + ["655360", "208", "1024"][len(sys.argv) - 1:]))
dev = torch.device("cuda:0")
m = LLFQVAE_V4(12, D, num_codes=K).to(dev)
lib = N.lib()
dims = m._dims()
params = [p.detach() for p in m._param_list()]
P = _param_struct(N.Params, params)
x = torch.randn(n, 12, device=dev)
ws = _workspace(dev, dims, n)
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
f32 = dict(dtype=torch.float32, device=dev)
scale, Wn, z_e = torch.empty(D, **f32), torch.empty(D, 128, **f32), torch.empty(n, D, **f32)
z_q, xr = torch.empty(n, D, **f32), torch.empty(n, 12, **f32)
idx = torch.randint(0, K, (n,), device=dev)
sse = torch.zeros(2, dtype=torch.float64, device=dev)
A = _acts_struct(scale=scale, Wn=Wn, z_e=z_e, idx=idx, z_q=z_q, x_recon=xr)


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); s.record()
    for _ in range(iters):
        fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


enc = lambda: N.check(lib.lipvq_encoder_fwd(C.byref(dims), C.byref(P), _p(x), n, C.byref(A), _p(ws), ws.numel(), st))
dec = lambda: N.check(lib.lipvq_quantize_decode_fwd(C.byref(dims), C.byref(P), _p(x), n, n, C.byref(A), _p(sse), None, _p(ws), ws.numel(), st))
vqs = torch.zeros(4, dtype=torch.int32, device=dev)
A2 = _acts_struct(scale=scale, Wn=Wn, z_e=z_e, idx=idx, z_q=z_q, x_recon=xr)
es = lambda: N.check(lib.lipvq_encode_search_fwd(C.byref(dims), C.byref(P), _p(x), n, C.byref(A2), 1, _p(vqs), _p(ws), ws.numel(), st))
if os.environ.get("ENC_PROBE_ONCE"):
    es(); torch.cuda.synchronize(); sys.exit(0)
print("encode_search_fwd %.1f us" % (1e3 * timeit(es)))
print("encoder_fwd %.1f us   quantize_decode_fwd %.1f us   (n=%d D=%d K=%d, extra=%r)"
      % (1e3 * timeit(enc), 1e3 * timeit(dec), n, D, K, os.environ.get("LIPVQ_NVCC_EXTRA", "")))
