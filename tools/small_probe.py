#!/usr/bin/env python
"""Latency of the single-kernel small-N forward: python tools/small_probe.py [n]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench
from lipvq_vae_b200 import LLFQVAE_V4, _native as N
n = int(sys.argv[1]) if len(sys.argv) > 1 else 80
dev = torch.device("cuda", 0)
p = bench.spread_params()
m = LLFQVAE_V4(12, 208).to(dev); m.load_state_dict({k: torch.from_numpy(v) for k, v in p.items()})
x = torch.randn(n, 12, device=dev)
def fwd(i):
    with torch.no_grad(): m(x)
for i in range(20): fwd(i)
rep = N.profile_kernels(fwd, 50)
print(n, {k: round(1e3 * v[1] / 50, 1) for k, v in rep.items()})
