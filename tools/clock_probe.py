#!/usr/bin/env python
"""SM clock / power while ONE kernel family runs back to back for ~3 s (nvidia-smi at 100 ms): tells a tensor-bound
kernel's real clock (the 1 kW cap pulls a B200 to ~1.3-1.4 GHz under sustained MMA load) from its nominal one."""
import os, subprocess, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
import torch
import bench
from lipvq_vae_b200 import _native as N

lib = N.lib()
dev = torch.device("cuda", 0)
which = sys.argv[1] if len(sys.argv) > 1 else "search208"
n, D, K = (655360, 208, 1024) if which == "search208" else (1 << 20, 64, 1024)
z = torch.rand(n, D, device=dev); E = torch.rand(K, D, device=dev)
idx = torch.empty(n, dtype=torch.int64, device=dev)
need = C.c_size_t(0)
N.check(lib.lipvq_workspace_bytes(C.byref(N.Dims(1, 1, 1, D, K)), n, C.byref(need)))
ws = torch.empty(need.value + 4096, dtype=torch.uint8, device=dev)
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
def call():
    N.check(lib.lipvq_vq_argmin(z.data_ptr(), E.data_ptr(), n, D, K, 1, idx.data_ptr(), None, None, ws.data_ptr(), ws.numel(), st))
for _ in range(3): call()
torch.cuda.synchronize()
p = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap", "--format=csv,noheader,nounits", "-lms", "100"],
                     stdout=subprocess.PIPE, text=True)
t0 = time.time(); k = 0
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
while time.time() - t0 < 3.0:
    for _ in range(50): call()
    k += 50
    torch.cuda.synchronize()
e.record(); torch.cuda.synchronize()
p.terminate()
rows = [l.strip() for l in p.stdout.read().splitlines() if l.strip()]
clk = sorted(float(r.split(",")[0]) for r in rows[3:])
print(which, "calls", k, "ms/call", s.elapsed_time(e) / k, "sm MHz median", clk[len(clk) // 2] if clk else None, "min", clk[0] if clk else None,
      "max", clk[-1] if clk else None, "samples", len(clk))
print(rows[5:12])
