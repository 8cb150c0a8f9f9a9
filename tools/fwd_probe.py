#!/usr/bin/env python
"""Forward-only timing of the benchmarked workload (655 360 rows, D=208, K=1024) with the library's per-kernel split;
run once per setting of an A/B environment knob, e.g.
    for c in 1 2 4 8; do LIPVQ_FWD_CHUNKS=$c python tools/fwd_probe.py; done
Prints one JSON line: ms per forward (CUDA events, 20 iterations after 5 warm-ups), tokens-only variant, kernel split."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from lipvq_vae_b200 import LLFQVAE_V4, _native as N  # noqa: E402

rows = int(os.environ.get("ROWS", bench.CHUNKS * bench.CHUNK_LEN))
dev = torch.device("cuda", 0)
p = bench._make_params() if os.environ.get("WEIGHTS", "init") == "init" else bench.spread_params()
m = LLFQVAE_V4(bench.F_DIM, bench.D_LAT, num_codes=bench.K_CODES, hidden_dim=bench.H_DIM).to(dev)
m.load_state_dict({k: torch.from_numpy(v) for k, v in p.items()})
g = torch.Generator().manual_seed(1)
xs = [torch.randn(rows, bench.F_DIM, generator=g).to(dev) for _ in range(6)]


def step(i):
    with torch.no_grad():
        m(xs[i % 6])


def step_tokens(i):
    m.forward_tokens(xs[i % 6])


out = {"rows": rows, "env": {k: v for k, v in os.environ.items() if k.startswith("LIPVQ_")}}
for name, fn in (("forward", step), ("forward_tokens", step_tokens)):
    for i in range(5):
        fn(i)
    out[name + "_ms"] = bench._time_events(fn, 20, torch)
rep = N.profile_kernels(step, 5)
out["kernels_us"] = {k: round(1e3 * v[1] / 5, 1) for k, v in rep.items()}
print(json.dumps(out))
