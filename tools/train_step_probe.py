#!/usr/bin/env python
"""TokenizerTrainer.step at bench size: ms per step (CUDA events) and the per-kernel split of one serial step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from lipvq_vae_b200 import LLFQVAE_V4, TokenizerTrainer
from lipvq_vae_b200 import _native as N
n = int(sys.argv[1]) if len(sys.argv) > 1 else 655360
m = LLFQVAE_V4(12, 208).cuda()
tr = TokenizerTrainer(m, lr=1e-3, weight_decay=1e-4)
xs = [torch.randn(n, 12, device="cuda") for _ in range(4)]
for i in range(5):
    tr.step(xs[i % 4])
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for i in range(20):
    tr.step(xs[i % 4])
e.record(); torch.cuda.synchronize()
print("train step %.3f ms" % (s.elapsed_time(e) / 20))
tr.concurrent_backward = False
prof = N.profile_kernels(lambda i: tr.step(xs[i % 4]), 5)
tot = 0.0
for k, (c, ms) in sorted(prof.items(), key=lambda kv: -kv[1][1]):
    print("  %-44s %3d launches/step %9.1f us/step" % (k, c // 5, 1e3 * ms / 5)); tot += 1e3 * ms / 5
print("  profiled kernels total %.1f us/step" % tot)
