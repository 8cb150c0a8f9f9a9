#!/usr/bin/env python
"""Every GPU kernel (name, count, total us) launched by ONE timed bench step of each kind, via torch.profiler (CUPTI):
shows what is inside the timed region besides this library's kernels (stray torch fill kernels)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

import bench  # noqa: E402
from lipvq_vae_b200 import LLFQVAE_V4, HostTokenizer, TokenizerTrainer  # noqa: E402

rows = bench.CHUNKS * bench.CHUNK_LEN
dev = torch.device("cuda", 0)
p = bench._make_params()
m = LLFQVAE_V4(bench.F_DIM, bench.D_LAT, num_codes=bench.K_CODES, hidden_dim=bench.H_DIM).to(dev)
m.load_state_dict({k: torch.from_numpy(v) for k, v in p.items()})
x = torch.randn(rows, bench.F_DIM, device=dev)
xh = torch.randn(rows, bench.F_DIM).pin_memory()


def fwd():
    with torch.no_grad():
        m(x)


tr = TokenizerTrainer(LLFQVAE_V4(bench.F_DIM, bench.D_LAT).to(dev))
ht = HostTokenizer(m, rows, depth=2)


def e2e():
    ht.submit(xh)
    ht.collect()


for name, fn in (("forward", fwd), ("train_step", lambda: tr.step(x)), ("e2e", e2e)):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        fn()
        torch.cuda.synchronize()
    print("== %s" % name)
    for ev in sorted(prof.key_averages(), key=lambda e: -e.device_time_total):
        if ev.device_time_total > 0:
            print("  %-90s x%-3d %9.1f us" % (ev.key[:90], ev.count, ev.device_time_total))
