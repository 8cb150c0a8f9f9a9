#!/usr/bin/env python
"""Probe the nearest-codeword search on a B200: tensor vs exact mode, timing + diagnostics.
usage: python tools/vq_probe.py [N D K [dist]] ...   (dist: uniform | collapsed | spread)"""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from lipvq_vae_b200 import _native as N


def make(n, D, K, dist, dev):
    g = torch.Generator(device=dev).manual_seed(0)
    if dist == "collapsed":
        z = 0.5 + 0.02 * torch.randn(n, D, device=dev, generator=g)
        b = (6.0 / D) ** 0.5
        E = (torch.rand(K, D, device=dev, generator=g) * 2 - 1) * b
    elif dist == "spread":
        base = torch.rand(1, D, device=dev, generator=g)
        lat = torch.randn(K, 12, device=dev, generator=g)
        P = torch.randn(12, D, device=dev, generator=g) * 0.05
        E = torch.sigmoid(base + lat @ P)
        z = torch.sigmoid(base + torch.randn(n, 12, device=dev, generator=g) @ P)
    else:
        z = torch.rand(n, D, device=dev, generator=g)
        E = torch.rand(K, D, device=dev, generator=g)
    return z.contiguous(), E.contiguous()


def run(n, D, K, dist="uniform", iters=5):
    dev = torch.device("cuda:0")
    lib = N.lib()
    z, E = make(n, D, K, dist, dev)
    dims = N.Dims(1, 1, 1, D, K)
    need = C.c_size_t(0)
    N.check(lib.lipvq_workspace_bytes(C.byref(dims), n, C.byref(need)))
    ws = torch.empty(need.value + 4096, dtype=torch.uint8, device=dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    out = {}
    for mode in (0, 1):
        idx = torch.full((n,), -1, dtype=torch.int64, device=dev)
        d2 = torch.zeros(n, device=dev)
        stats = torch.zeros(4, dtype=torch.int32, device=dev)
        def call(with_d2=False):
            N.check(lib.lipvq_vq_argmin(z.data_ptr(), E.data_ptr(), n, D, K, mode, idx.data_ptr(),
                                        d2.data_ptr() if with_d2 else None,
                                        stats.data_ptr(), ws.data_ptr(), ws.numel(), st), "vq_argmin")
        call(True)
        torch.cuda.synchronize()
        idx_d2 = idx.clone()
        call()
        torch.cuda.synchronize()
        assert torch.equal(idx, idx_d2), "index differs between the d2 and the no-d2 path"
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(iters):
            call()
        e.record()
        torch.cuda.synchronize()
        out[mode] = (idx.clone(), d2.clone(), stats.cpu().tolist(), s.elapsed_time(e) / iters)
    i0, d0, _, t0 = out[0]
    i1, d1, s1, t1 = out[1]
    diff = (i0 != i1)
    nd = int(diff.sum())
    rel = ((d1 - d0).abs() / d0.clamp_min(1e-30))
    bad = int((diff & (rel > 2.5e-5)).sum())
    tf = 2.0 * n * D * K / 1e12
    print("N=%d D=%d K=%d %-9s exact %.3f ms (%.1f TF/s) | tensor %.3f ms (%.1f TF/s) | idx diff %d (outside tie "
          "window: %d) maxrel %.2e | fallback rows %d rescored codes %d (%.2f/row) unambiguous rows %.1f%%"
          % (n, D, K, dist, t0, tf / t0 * 1e3, t1, tf / t1 * 1e3, nd, bad, float(rel.max()), s1[0], s1[1],
             s1[1] / max(n, 1), 100.0 * s1[2] / max(n, 1)), flush=True)
    if bad:
        w = torch.nonzero(diff & (rel > 2.5e-5))[:8, 0]
        print("  first bad rows", w.tolist(), "exact", i0[w].tolist(), "tensor", i1[w].tolist(),
              "d0", d0[w].tolist(), "d1", d1[w].tolist())
    return bad


if __name__ == "__main__":
    a = sys.argv[1:]
    cases = []
    while a:
        n, D, K = int(a[0]), int(a[1]), int(a[2])
        dist = a[3] if len(a) > 3 and not a[3].isdigit() else "uniform"
        a = a[4:] if len(a) > 3 and not a[3].isdigit() else a[3:]
        cases.append((n, D, K, dist))
    if not cases:
        cases = [(1000, 64, 256, "uniform"), (4096, 64, 1024, "uniform"), (5000, 208, 1024, "uniform"),
                 (5000, 208, 1024, "collapsed"), (5000, 208, 1024, "spread"), (3000, 128, 4096, "uniform"),
                 (3000, 256, 1024, "uniform"), (3000, 300, 512, "uniform"), (2000, 24, 64, "uniform"),
                 (1 << 20, 64, 1024, "uniform"), (655360, 208, 1024, "uniform"), (655360, 208, 1024, "collapsed"),
                 (655360, 208, 1024, "spread"), (1 << 18, 128, 4096, "uniform"), (1 << 18, 256, 16384, "uniform")]
    bad = 0
    for c in cases:
        bad += run(*c)
    sys.exit(1 if bad else 0)
