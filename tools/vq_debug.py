#!/usr/bin/env python
"""Debug aid: rows where the tensor-path search disagrees with the exact scan beyond the parity tolerance."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from lipvq_vae_b200 import LFQQuantizer
torch.manual_seed(0)
n, D, K = 1 << 20, 64, 1024
z = torch.rand(n, D, device="cuda")
q = LFQQuantizer(K, D).cuda()
with torch.no_grad():
    q.codebook.copy_(torch.rand(K, D, device="cuda"))
E = q.codebook.detach()
out = {}
for mode in (0, 1):
    q.search_mode = mode
    out[mode] = q(z)[1]
    print("mode", mode, "stats", q.last_search_stats.tolist())
i0, i1 = out[0], out[1]
d0 = (z - E[i0]).pow(2).sum(1); d1 = (z - E[i1]).pow(2).sum(1)
bad = ((i0 != i1) & ((d1 - d0) > 2.5e-5 * d0)).nonzero().flatten()
print("differ", int((i0 != i1).sum()), "bad", bad.numel())
for r in bad[:12].tolist():
    a, b = int(i0[r]), int(i1[r])
    print("row %d tile %d r_in %d | exact code %d (part %d chunk %d el %d) d2 %.6f | tensor code %d (part %d chunk %d el %d) d2 %.6f" % (
        r, r // 128, r % 128, a, (a % 256) // 64, a // 8, a % 8, d0[r].item(), b, (b % 256) // 64, b // 8, b % 8, d1[r].item()))
