#!/usr/bin/env python
"""vq_tc_kernel time by latent width (K blocks of the GEMM): python tools/vq_shape_probe.py [n K]
Prints us per call and SM cycles per 128x256 accumulator (1965 MHz, 148 SMs)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
import torch
from lipvq_vae_b200 import _native as N
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
K = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
lib = N.lib(); dev = torch.device("cuda", 0)
for D in [int(v) for v in os.environ.get('PROBE_D', '61,125,189,208,253,300').split(',')]:
    torch.manual_seed(0)
    z = torch.rand(n, D, device=dev); E = torch.rand(K, D, device=dev)
    idx = torch.empty(n, dtype=torch.int64, device=dev)
    need = C.c_size_t(0)
    N.check(lib.lipvq_workspace_bytes(C.byref(N.Dims(1, 1, 1, D, K)), n, C.byref(need)))
    ws = torch.empty(need.value + 4096, dtype=torch.uint8, device=dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    def run(i):
        N.check(lib.lipvq_vq_argmin(z.data_ptr(), E.data_ptr(), n, D, K, 1, idx.data_ptr(), None, None, ws.data_ptr(), ws.numel(), st))
    for i in range(3): run(i)
    rep = N.profile_kernels(run, 5)
    us = 1e3 * rep["vq_tc_kernel"][1] / 5
    acc_per_sm = ((n + 127) // 128) * ((K + 255) // 256) / 148.0
    nk16 = (D + 3 + 15) // 16
    print("D=%3d  K blocks %d  MMAs/acc %2d (nominal %4d cyc)  vq_tc %7.1f us  = %5.0f cycles per accumulator" %
          (D, (D + 3 + 63) // 64, nk16, nk16 * 128, us, us * 1965.0 / acc_per_sm))
