"""GPU probe: lipvq_linear_bwd (dW = dY^T X, db, dX = (dY W) [* mask]) vs float64 torch, with per-kernel timing.
Usage: python tools/bwd_probe.py [rows]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from lipvq_vae_b200 import _native as N

lib = N.lib()
dev = torch.device("cuda:0")
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 655360
st = torch.cuda.current_stream().cuda_stream
ok = True
for (Nout, Kin, scale) in [(208, 128, 1.0), (64, 12, 1.0)] if os.environ.get('PROBE_V') else ((208, 128, 1e-6), (128, 64, 1e-6), (64, 12, 1e-6), (208, 128, 1.0), (64, 128, 1.0), (256, 128, 1.0), (12, 128, 1.0)):
    for M in (rows, 4096 + 40, 100003):
        g = torch.Generator(device="cpu").manual_seed(Nout * 7 + Kin)
        dY = (torch.randn(M, Nout, generator=g) * scale).to(dev)
        X = torch.randn(M, Kin, generator=g).to(dev)
        W = (torch.randn(Nout, Kin, generator=g) / Kin ** 0.5).to(dev)
        dX = torch.empty(M, Kin, device=dev); dW = torch.empty(Nout, Kin, device=dev); db = torch.empty(Nout, device=dev)
        ws = torch.empty(lib.lipvq_linear_bwd_ws_bytes(M, Nout, Kin), dtype=torch.uint8, device=dev)
        mask = torch.relu(torch.randn(M, Kin, generator=g)).to(dev) if (Nout + Kin + M) % 2 else None
        def run(_i=0):
            N.check(lib.lipvq_linear_bwd(dY.data_ptr(), X.data_ptr(), W.data_ptr(), mask.data_ptr() if mask is not None else None, dX.data_ptr(), dW.data_ptr(),
                                         db.data_ptr(), M, Nout, Kin, ws.data_ptr(), ws.numel(), st), "linear_bwd")
        run(); torch.cuda.synchronize()
        rW = dY.double().t() @ X.double(); rb = dY.double().sum(0); rX = dY.double() @ W.double()
        if mask is not None: rX = rX * (mask > 0)
        def err(a, r): return ((a.double() - r).abs().max() / r.abs().max()).item()
        eW, eb, eX = err(dW, rW), err(db, rb), err(dX, rX)
        good = eW < 4e-6 and eb < 2e-6 and eX < 5e-6
        ok &= good
        line = "Nout=%d Kin=%d M=%d  err dW %.2e db %.2e dX %.2e %s" % (Nout, Kin, M, eW, eb, eX, "ok" if good else "BAD")
        if M == rows:
            rep = N.profile_kernels(run, 5)
            line += "  | " + ", ".join("%s %.0fus" % (k, 1e3 * v[1] / 5) for k, v in rep.items())
        print(line, flush=True)
        if not good and M != rows and os.environ.get("PROBE_V"):
            print("   dW[0,:6]", dW[0, :6].tolist(), "\n   ref     ", rW[0, :6].tolist(), "\n   db[:4]", db[:4].tolist(), "ref", rb[:4].tolist())
            print("   dW[130 % Nout,:4]", dW[130 % Nout, :4].tolist(), "ref", rW[130 % Nout, :4].tolist(), "nnz", int((dW != 0).sum()), "nan", int(dW.isnan().sum()))
print("ALL OK" if ok else "FAILED")
