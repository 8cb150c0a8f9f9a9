#!/usr/bin/env python
"""A few calls of the standalone search (lipvq_vq_argmin) for profiling: python tools/vq_one.py [n D K calls]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
import torch
from lipvq_vae_b200 import _native as N
a = [int(v) for v in sys.argv[1:]] + [None] * 4
n, D, K, calls = a[0] or (1 << 20), a[1] or 64, a[2] or 1024, a[3] or 3
lib = N.lib(); dev = torch.device("cuda", 0)
torch.manual_seed(0)
z = torch.rand(n, D, device=dev); E = torch.rand(K, D, device=dev)
idx = torch.empty(n, dtype=torch.int64, device=dev)
need = C.c_size_t(0)
N.check(lib.lipvq_workspace_bytes(C.byref(N.Dims(1, 1, 1, D, K)), n, C.byref(need)))
ws = torch.empty(need.value + 4096, dtype=torch.uint8, device=dev)
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for _ in range(calls):
    N.check(lib.lipvq_vq_argmin(z.data_ptr(), E.data_ptr(), n, D, K, 1, idx.data_ptr(), None, None, ws.data_ptr(), ws.numel(), st))
torch.cuda.synchronize()
print("ok", idx[:4].tolist())
