#!/usr/bin/env python
"""BASELINE config 4 on ONE B200: large-codebook sweep of the nearest-codeword search (tensor mode: 16-bit
tcgen05 distances + fp32 near-tie rescore), N = 2^20 rows, D in {128, 256}, K in {4096, 16384, 65536}.
Prints one JSON line per case with the whole call, the tcgen05 kernel alone (library CUDA events) and a
parity check against the exact fp32 scan on the first 8192 rows."""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from lipvq_vae_b200 import _native as N

PEAK = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
dev = torch.device("cuda:0")
lib = N.lib()
n = 1 << 20
cases = [(128, 4096), (128, 16384), (128, 65536), (256, 4096), (256, 16384), (256, 65536), (64, 1024), (208, 1024)]
if len(sys.argv) > 1:
    cases = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]]
for D, K in cases:
    g = torch.Generator(device=dev).manual_seed(0)
    z = torch.rand(n, D, device=dev, generator=g)
    E = torch.rand(K, D, device=dev, generator=g)
    need = C.c_size_t(0)
    N.check(lib.lipvq_workspace_bytes(C.byref(N.Dims(1, 1, 1, D, K)), n, C.byref(need)))
    ws = torch.empty(need.value + 4096, dtype=torch.uint8, device=dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    idx = torch.empty(n, dtype=torch.int64, device=dev)
    stats = torch.zeros(4, dtype=torch.int32, device=dev)

    def call(i=0, mode=1, rows=n, out=idx):
        N.check(lib.lipvq_vq_argmin(z.data_ptr(), E.data_ptr(), rows, D, K, mode, out.data_ptr(), None,
                                    stats.data_ptr(), ws.data_ptr(), ws.numel(), st), "vq_argmin")
    for _ in range(2):
        call()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); s.record()
    iters = 3
    for _ in range(iters):
        call()
    e.record(); torch.cuda.synchronize()
    ms = s.elapsed_time(e) / iters
    prof = N.profile_kernels(call, iters)
    st_now = stats.cpu().tolist()
    sub = 8192
    ex = torch.empty(sub, dtype=torch.int64, device=dev)
    call(mode=0, rows=sub, out=ex)
    torch.cuda.synchronize()
    zs = z[:sub]
    d_t = (zs - E[idx[:sub]]).pow(2).sum(1)
    d_e = (zs - E[ex]).pow(2).sum(1)
    mism = int((idx[:sub] != ex).sum())
    bad = int(((idx[:sub] != ex) & ((d_t - d_e).abs() > 2.5e-5 * d_e)).sum())
    flop = 2.0 * n * D * K
    tc_ms = prof["vq_tc_kernel"][1] / iters if "vq_tc_kernel" in prof else None
    print(json.dumps({"N": n, "D": D, "K": K, "call_ms": ms, "call_tflops": flop / ms / 1e9,
                      "vq_tc_kernel_ms": tc_ms, "vq_tc_kernel_tflops": flop / tc_ms / 1e9 if tc_ms else None,
                      "frac_of_sustained_bf16": (flop / tc_ms / 1e9) / PEAK["bf16_tflops_sustained"] if tc_ms else None,
                      "kernels_ms": {k: v[1] / iters for k, v in prof.items()},
                      "fallback_rows": st_now[0], "rescored_codes": st_now[1],
                      "index_mismatch_vs_exact_first_8192": mism, "outside_tie_window": bad}), flush=True)
    del z, E, ws, idx
    torch.cuda.empty_cache()
