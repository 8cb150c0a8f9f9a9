#!/usr/bin/env python
"""One forward+backward of the tokenizer at bench size (for ncu launch lists of the training step)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from lipvq_vae_b200 import LLFQVAE_V4
n = int(sys.argv[1]) if len(sys.argv) > 1 else 655360
m = LLFQVAE_V4(12, 208).cuda()
x = torch.randn(n, 12, device="cuda")
for it in range(3):
    for p in m.parameters():
        p.grad = None
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    z, loss = m(x)
    loss.backward()
    e.record(); torch.cuda.synchronize()
    print("fwd+bwd %.3f ms" % s.elapsed_time(e))

from lipvq_vae_b200 import _native as N
def step(i):
    for p in m.parameters():
        p.grad = None
    z, loss = m(x)
    loss.backward()
prof = N.profile_kernels(step, 3)
tot = 0.0
for k, (c, ms) in sorted(prof.items(), key=lambda kv: -kv[1][1]):
    print("  %-28s %3d launches/step %9.1f us/step" % (k, c // 3, 1e3 * ms / 3)); tot += 1e3 * ms / 3
print("  profiled kernels total %.1f us/step" % tot)
