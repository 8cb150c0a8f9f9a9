"""Parity at the BENCHMARKED sizes (VERDICT r1 "What's weak" 1): the full forward at BASELINE config 3's batch
(655 360 rows, D=208, K=1024) and the search kernel at config 2 (N=2^20, D=64, K=1024) against the oracle's
float64 ground truth (``oracle.exact_top2``, row-blocked), under BASELINE.json's rule: indices bit-exact on every
row whose top-2 distance gap exceeds 1e-5 relative, everything else within 1e-5.  Plus the tiny-dimension shapes
ADVICE r1 flagged (latent_dim 4 / 1, feature_dim 1), the prompt-token cache on the device and the two-GPU trainer."""
import os
import socket

import numpy as np
import pytest
import torch

from oracle import lipvq_oracle as orc

pytestmark = pytest.mark.gpu
RTOL = 1e-5
GTOL = 5e-5


def _model(p, mode=1, dev="cuda:0"):
    from lipvq_vae_b200 import LLFQVAE_V4
    F = p["encoder.0.weight"].shape[1]
    K, D = p["quantizer.codebook"].shape
    H = p["encoder.2.weight"].shape[0]
    m = LLFQVAE_V4(F, D, num_codes=K, hidden_dim=H).to(dev)
    m.load_state_dict({k: torch.from_numpy(v.copy()) for k, v in p.items()}, strict=True)
    m.search_mode = mode
    return m


def _poison_workspaces():
    """Fill every cached scratch workspace with 0xFF bytes (NaN floats, -1 ints): a kernel that reads scratch before it
    wrote it -- or a value a previous call left there -- then fails loudly instead of passing by luck."""
    from lipvq_vae_b200 import tokenizer
    for ws in tokenizer._WS.values():
        ws.fill_(0xFF)
    torch.cuda.synchronize()


def spread_params(F=12, D=208, K=1024, seed=11):
    """Weights with index diversity (random init collapses onto one code): a milder Lipschitz layer with
    both ``min`` branches in use and a codebook sampled from the encoder's own output distribution."""
    rng = np.random.default_rng(seed)
    p = orc.init_params(F, D, K, 128, seed=seed)
    p["to_latent.W"] = (0.08 * p["to_latent.W"]).astype(np.float32)
    p["to_latent.ci"][::3] = 5.0
    z = orc.encode(p, (2.0 * rng.random((K, F)) - 1.0).astype(np.float32))
    p["quantizer.codebook"] = (z + 1e-3 * rng.standard_normal(z.shape)).astype(np.float32)
    return p


@pytest.mark.timeout(900)
def test_forward_at_bench_size_vs_oracle():
    """655 360 rows x 12 -> D=208, K=1024 (the workload bench.py times), spread weights, no_grad forward."""
    n, F, D, K = 655360, 12, 208, 1024
    p = spread_params(F, D, K)
    x = (2.0 * np.random.default_rng(5).random((n, F)) - 1.0).astype(np.float32)
    m = _model(p)
    _poison_workspaces()
    with torch.no_grad():
        z_lat, loss = m(torch.from_numpy(x).cuda())
    torch.cuda.synchronize()
    idx = m.last_indices.cpu().numpy()
    terms = m.last_loss_terms.cpu().numpy().astype(np.float64)
    xr = m.last_reconstruction.cpu().numpy()
    z_lat = z_lat.cpu().numpy()
    # oracle, row-blocked
    E = p["quantizer.codebook"]
    z_e = np.concatenate([orc.encode(p, x[s:s + 65536]) for s in range(0, n, 65536)])
    ref_idx, d1, d2 = orc.exact_top2(z_e, E)
    must = orc.index_parity_mask(d1, d2)
    assert must.mean() > 0.99 and len(np.unique(ref_idx)) > K // 2          # the case is not vacuous
    assert np.array_equal(idx[must], ref_idx[must]), "index mismatch outside the 1e-5 gap window"
    assert np.array_equal(z_lat, E[idx]), "z_latent is not the gathered codeword"
    # rows inside the window may legitimately pick the other code: evaluate the oracle's loss terms on the kernel's idx
    xr_code = orc.decode(p, E)
    assert orc.rel_err(xr, xr_code[idx]) < RTOL
    sse_r = float(np.sum((xr_code[idx].astype(np.float64) - x) ** 2))
    sse_q = float(np.sum((E[idx].astype(np.float64) - z_e) ** 2))
    recon, commit = sse_r / (n * F), sse_q / (n * D)
    ref_loss = recon + 0.5 * commit
    assert abs(loss.item() - ref_loss) < RTOL * abs(ref_loss), (loss.item(), ref_loss)
    assert abs(terms[1] - recon) < RTOL * recon and abs(terms[2] - commit) < RTOL * commit


@pytest.mark.timeout(900)
def test_config2_search_vs_oracle():
    """BASELINE config 2: the search API alone, N = 2^20 latents, D=64, K=1024, fp32, both codebook distributions."""
    from lipvq_vae_b200 import LFQQuantizer
    n, D, K = 1 << 20, 64, 1024
    rng = np.random.default_rng(0)
    z = rng.random((n, D), dtype=np.float32)
    zt = torch.from_numpy(z).cuda()
    for name, E in (("spread", rng.random((K, D), dtype=np.float32)),
                    ("reference-init", ((2.0 * rng.random((K, D)) - 1.0) * np.sqrt(6.0 / D)).astype(np.float32))):
        q = LFQQuantizer(K, D).cuda()
        with torch.no_grad():
            q.codebook.copy_(torch.from_numpy(E))
        ref_idx, d1, d2 = orc.exact_top2(z, E, block_rows=16384)
        must = orc.index_parity_mask(d1, d2)
        for mode in (1, 0):
            q.search_mode = mode
            _poison_workspaces()
            z_q, idx = q(zt)
            idx = idx.cpu().numpy()
            assert np.array_equal(idx[must], ref_idx[must]), (name, mode, int((idx != ref_idx)[must].sum()))
            assert np.array_equal(z_q.cpu().numpy(), E[idx])


@pytest.mark.parametrize("F,D,K,n", [(12, 4, 16, 300), (12, 1, 8, 130), (1, 4, 16, 257), (1, 1, 4, 64), (3, 8, 32, 1000)])
def test_tiny_dims_forward_backward(F, D, K, n):
    """latent_dim 4 / 1 and feature_dim 1: the emit pass's reciprocal-multiply division has a divisor of 1 there
    (2^32/1 truncated to a zero multiplier -> every element read row 0)."""
    rng = np.random.default_rng(100 * F + 10 * D + K)
    p = orc.init_params(F, D, K, 128, seed=F + D + K)
    p["to_latent.W"] = (0.05 * p["to_latent.W"]).astype(np.float32)
    z = orc.encode(p, (2.0 * rng.random((K, F)) - 1.0).astype(np.float32))
    p["quantizer.codebook"] = (z + 1e-2 * rng.standard_normal(z.shape)).astype(np.float32)
    x = (2.0 * rng.random((n, F)) - 1.0).astype(np.float32)
    fw = orc.forward(p, x)
    _, d1, d2 = orc.exact_top2(fw["z_e"], p["quantizer.codebook"])
    must = orc.index_parity_mask(d1, d2)
    for mode in (0, 1):
        m = _model(p, mode)
        xt = torch.from_numpy(x).cuda()
        with torch.no_grad():
            z0, loss0 = m(xt)
        z_lat, loss = m(xt)
        loss.backward()
        torch.cuda.synchronize()
        idx = m.last_indices.cpu().numpy()
        assert np.array_equal(idx[must], fw["idx"][must])
        mt = torch.from_numpy(must).cuda()       # (the no-grad call runs the single-kernel exact path)
        assert np.array_equal(z_lat.cpu().numpy(), p["quantizer.codebook"][idx]) and torch.equal(z0[mt], z_lat[mt])
        if np.array_equal(idx, fw["idx"]):
            assert abs(loss.item() - fw["loss"]) < RTOL * abs(fw["loss"]) and abs(loss0.item() - fw["loss"]) < RTOL * abs(fw["loss"])
            assert orc.rel_err(m.last_reconstruction.cpu().numpy(), fw["x_recon"]) < RTOL
            gr = orc.backward(p, x, fw)
            # latent_dim 1: the whole encoder gradient hangs on one column of ~1e-8 magnitudes through the 3-term split
            # GEMMs of the fallback path (shapes outside the tcgen05 kernels); 2e-4 of the tensor's largest element
            tol = 2e-4 if D == 1 else GTOL
            for k, t in m.named_parameters():
                assert orc.rel_err(t.grad.cpu().numpy(), gr[k]) < tol, (k, mode)


def test_prompt_cache_on_device_hit_miss_invalidate():
    """PromptTokenCache around the real module on the GPU: replay on an unchanged prompt, miss after a
    TokenizerTrainer step (raw-pointer parameter update -> parameter epoch), reference state_dict keys intact."""
    from lipvq_vae_b200 import PromptTokenCache, TokenizerTrainer
    from tests.golden.make_golden import case_inputs
    p, x = case_inputs("small")
    m = _model(p)
    c = PromptTokenCache(m, capacity=4)
    assert list(c.state_dict().keys()) == list(p.keys())                      # no "model." prefix
    c.load_state_dict({k: torch.from_numpy(v.copy()) for k, v in p.items()}, strict=True)
    xt = torch.from_numpy(x).cuda()
    xh = torch.from_numpy(x).pin_memory()
    with torch.no_grad():
        z0, l0 = c(xt)
        z1, l1 = c(xt)                       # device key: same storage, same version
        z2, l2 = c(xh)                       # host key: content hash (uploads on a miss)
        z3, l3 = c(xh.clone())
    assert (c.hits, c.misses) == (2, 2)
    assert z1 is z0 and z3 is z2 and torch.equal(z0, z2) and torch.equal(l0, l2)
    assert torch.equal(c.last_indices, m.last_indices)
    fw = orc.forward(p, x)
    assert np.array_equal(c.last_indices.cpu().numpy(), fw["idx"])
    tr = TokenizerTrainer(m)
    tr.step(xt)                              # parameters move through the flat AdamW kernel: no autograd version bump
    with torch.no_grad():
        z4, l4 = c(xt)
    assert c.misses == 3 and not torch.equal(l4, l0)
    xt.add_(0.25)                            # in-place prompt edit bumps the tensor version: miss
    with torch.no_grad():
        c(xt)
    assert c.misses == 4
    _, loss = c(xt)                          # autograd on: straight through, graph intact
    assert loss.requires_grad and c.misses == 4


# ----------------------------------------------------------------------------- two GPUs
def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _dp_worker(rank, world, port, n_total, steps, out_dir, overlap):
    import torch.distributed as dist
    from lipvq_vae_b200 import TokenizerTrainer, shard_rows
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        p = spread_params(12, 64, 256, seed=21)
        if rank != 0:
            p = {k: v + 0.5 for k, v in p.items()}            # the constructor's broadcast must overwrite these
        m = _model(p, 1, dev="cuda:%d" % rank)
        tr = TokenizerTrainer(m, overlap=overlap)
        rng = np.random.default_rng(77)
        losses = []
        for s in range(steps):
            x = (2.0 * rng.random((n_total, 12)) - 1.0).astype(np.float32)     # same stream on every rank
            s0, s1 = shard_rows(n_total, rank, world)
            _, loss = tr.step(torch.from_numpy(x[s0:s1]).cuda(), n_total)
            losses.append(loss)
        flat = tr.opt.flat.clone()
        gathered = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat)
        torch.cuda.synchronize()
        if rank == 0:
            np.savez(os.path.join(out_dir, "dp_%d.npz" % int(overlap)), losses=np.array([l.item() for l in losses]),
                     **{"flat%d" % r: g.cpu().numpy() for r, g in enumerate(gathered)})
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("overlap", [True, False])
def test_two_gpu_trainer_replicas_identical_and_match_one_gpu(tmp_path, overlap):
    """after 3 TokenizerTrainer steps on sharded rows the replicas' flat parameter buffers are
    bit-identical, and they match a 1-GPU run on the concatenated batch."""
    import torch.multiprocessing as mp
    from lipvq_vae_b200 import TokenizerTrainer
    n_total, steps = 20480, 3
    mp.spawn(_dp_worker, args=(2, _free_port(), n_total, steps, str(tmp_path), overlap), nprocs=2, join=True)
    got = np.load(os.path.join(str(tmp_path), "dp_%d.npz" % int(overlap)))
    assert np.array_equal(got["flat0"], got["flat1"]), "replicas diverged"
    m = _model(spread_params(12, 64, 256, seed=21), 1)
    tr = TokenizerTrainer(m)
    rng = np.random.default_rng(77)
    for s in range(steps):
        x = (2.0 * rng.random((n_total, 12)) - 1.0).astype(np.float32)
        _, loss = tr.step(torch.from_numpy(x).cuda(), n_total)
        assert abs(loss.item() - got["losses"][s]) < RTOL * abs(loss.item()), (s, loss.item(), got["losses"][s])
    d = np.abs(tr.opt.flat.cpu().numpy().astype(np.float64) - got["flat0"])
    # Adam's step is +-lr wherever |g| sits at its own rounding noise: bound by the total possible movement, and
    # require fp32-rounding agreement on 99 % of the elements
    assert d.max() <= steps * 1e-3 * 1.01 and np.quantile(d, 0.99) < 2e-6, (d.max(), np.quantile(d, 0.99))


# ----------------------------------------------------------------------------- the single-kernel small-N forward
def _small_forward(p, x, n_total=None):
    """lipvq_forward_small through the raw C ABI."""
    import ctypes as C
    from lipvq_vae_b200 import _native as N
    F = p["encoder.0.weight"].shape[1]
    K, D = p["quantizer.codebook"].shape
    H = p["encoder.2.weight"].shape[0]
    dims = N.Dims(F, 64, H, D, K)
    dev = "cuda:0"
    ts = [torch.from_numpy(p[k].copy()).to(dev) for k in N.PARAM_KEYS]
    P = N.Params(*[t.data_ptr() for t in ts])
    n = x.shape[0]
    xt = torch.from_numpy(x).to(dev)
    z = torch.empty(n, D, device=dev)
    idx = torch.empty(n, dtype=torch.int64, device=dev)
    xr = torch.empty(n, F, device=dev)
    loss = torch.empty(4, device=dev)
    stats = torch.empty(4, dtype=torch.int32, device=dev)
    ws = torch.empty(16 * n + 65536, dtype=torch.uint8, device=dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    N.check(N.lib().lipvq_forward_small(C.byref(dims), C.byref(P), xt.data_ptr(), n, n_total or n, z.data_ptr(),
                                        idx.data_ptr(), xr.data_ptr(), loss.data_ptr(), stats.data_ptr(), ws.data_ptr(),
                                        ws.numel(), st), "lipvq_forward_small")
    torch.cuda.synchronize()
    return z.cpu().numpy(), idx.cpu().numpy(), xr.cpu().numpy(), loss.cpu().numpy().astype(np.float64)


@pytest.mark.parametrize("case", ["small", "d787", "c1_spread", "c1_init"])
def test_small_forward_kernel_vs_reference_golden(case):
    """csrc/small_fwd.cuh against the vectors of the REAL reference module (tests/golden/model_*.npz)."""
    from tests.golden.make_golden import case_inputs
    p, x = case_inputs(case)
    g = dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "model_%s.npz" % case)))
    z, idx, xr, loss = _small_forward(p, x)
    fw = orc.forward(p, x)
    _, d1, d2 = orc.exact_top2(fw["z_e"], p["quantizer.codebook"])
    must = orc.index_parity_mask(d1, d2)
    ref_idx = g["idx"].astype(np.int64)
    assert np.array_equal(idx[must], ref_idx[must])
    assert np.array_equal(z, p["quantizer.codebook"][idx])
    if np.array_equal(idx, ref_idx):
        assert abs(loss[0] - g["loss"]) < RTOL * abs(g["loss"]), (loss, g["loss"])
        assert orc.rel_err(xr[:32], g["x_recon_head"]) < RTOL
        assert abs(xr.astype(np.float64).sum() - g["x_recon_sum"]) < RTOL * g["x_recon_abs"]
        assert abs(loss[1] - fw["recon_loss"]) < RTOL * fw["recon_loss"] and abs(loss[2] - fw["commit_loss"]) < RTOL * fw["commit_loss"]
    # run-to-run determinism of the whole output, loss reduction included
    z2, idx2, xr2, loss2 = _small_forward(p, x)
    assert np.array_equal(idx, idx2) and np.array_equal(xr, xr2) and np.array_equal(loss, loss2)


@pytest.mark.parametrize("n,F,D,K", [(1, 12, 208, 1024), (80, 12, 208, 1024), (500, 12, 208, 1024), (3, 12, 50, 33),
                                     (257, 5, 7, 100), (2048, 12, 36, 64), (130, 1, 4, 16), (64, 12, 1, 8)])
def test_small_forward_kernel_shapes_vs_oracle(n, F, D, K):
    rng = np.random.default_rng(n + F + D + K)
    p = orc.init_params(F, D, K, 128, seed=n + D)
    p["to_latent.W"] = (0.08 * p["to_latent.W"]).astype(np.float32)
    p["to_latent.ci"][::3] = 5.0
    zc = orc.encode(p, (2.0 * rng.random((K, F)) - 1.0).astype(np.float32))
    p["quantizer.codebook"] = (zc + 1e-3 * rng.standard_normal(zc.shape)).astype(np.float32)
    x = (2.0 * rng.random((n, F)) - 1.0).astype(np.float32)
    fw = orc.forward(p, x)
    _, d1, d2 = orc.exact_top2(fw["z_e"], p["quantizer.codebook"])
    must = orc.index_parity_mask(d1, d2) if K > 1 else np.ones(n, bool)
    n_total = 3 * n                                     # a data-parallel shard: the loss carries 1/n_total
    z, idx, xr, loss = _small_forward(p, x, n_total)
    assert np.array_equal(idx[must], fw["idx"][must])
    assert np.array_equal(z, p["quantizer.codebook"][idx])
    if np.array_equal(idx, fw["idx"]):
        assert orc.rel_err(xr, fw["x_recon"]) < RTOL
        ref = fw["loss"] * n / n_total
        assert abs(loss[0] - ref) < RTOL * abs(ref), (loss[0], ref)
    # and the module routes no-grad calls of this size here: same tokens as the raw call
    m = _model(p)
    with torch.no_grad():
        z_m, loss_m = m(torch.from_numpy(x).cuda())
    idx_m = m.last_indices.cpu().numpy()
    if n <= 768:
        assert np.array_equal(idx_m, idx) and np.array_equal(z_m.cpu().numpy(), z)
    else:                                   # larger calls take the tensor path: equal outside the near-tie window
        assert np.array_equal(idx_m[must], idx[must])


# ----------------------------------------------------------------------------- deterministic per-code sums
@pytest.mark.parametrize("n,D,K,collapsed", [(100003, 208, 1024, True), (100003, 208, 1024, False), (5000, 50, 100, False),
                                             (70000, 12, 1024, False), (1, 208, 1024, False), (300000, 64, 8192, False),
                                             (655360, 208, 1024, True)])
def test_segment_sum_deterministic_and_exact(n, D, K, collapsed):
    """lipvq_segment_sum (csrc/seg_reduce.cuh): sums, differences against a table, counts and accumulation against
    float64; bit-identical results over repeated runs (the property the atomic scatter did not have)."""
    import ctypes as C
    from lipvq_vae_b200 import _native as N
    lib = N.lib()
    rng = np.random.default_rng(n + D + K)
    src = rng.random((n, D), dtype=np.float32)
    idx = np.full(n, 7 % K, np.int64) if collapsed else rng.integers(0, K, n)
    if not collapsed and n > 10:
        idx[: n // 3] = idx[0]                      # one heavy code next to many light ones
    T = rng.random((K, D), dtype=np.float32)
    gs = torch.tensor([0.5], device="cuda")
    ts, ti, tT = torch.from_numpy(src).cuda(), torch.from_numpy(idx).cuda(), torch.from_numpy(T).cuda()
    ws = torch.empty(lib.lipvq_segment_sum_ws_bytes(n, D, K), dtype=torch.uint8, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def run(table, alpha, scale, accumulate, init):
        sums = torch.full((K, D), init, device="cuda")
        counts = torch.empty(K, device="cuda")
        N.check(lib.lipvq_segment_sum(ts.data_ptr(), ti.data_ptr(), table.data_ptr() if table is not None else None, alpha,
                                      scale.data_ptr() if scale is not None else None, n, D, K, int(accumulate),
                                      sums.data_ptr(), counts.data_ptr(), ws.data_ptr(), ws.numel(), st))
        torch.cuda.synchronize()
        return sums.cpu().numpy(), counts.cpu().numpy()

    ref = np.zeros((K, D), np.float64)
    np.add.at(ref, idx, src.astype(np.float64))
    cnt = np.bincount(idx, minlength=K).astype(np.float64)
    s1, c1 = run(None, 1.0, None, False, 123.0)
    assert np.array_equal(c1, cnt)
    assert orc.rel_err(s1, ref) < 1e-6
    for _ in range(3):
        s2, _ = run(None, 1.0, None, False, -5.0)
        assert np.array_equal(s1, s2), "not bit-identical run to run"
    # difference mode + device-side scale + accumulation: sums = 2 + 0.25 * 0.5 * sum (T[k] - src[n])
    s3, _ = run(tT, 0.25, gs, True, 2.0)
    ref3 = 2.0 + 0.125 * (cnt[:, None] * T.astype(np.float64) - ref)
    assert np.abs(s3 - ref3).max() <= 2e-6 * max(1.0, np.abs(ref3).max())


def test_backward_is_bit_identical_run_to_run():
    """all 14 gradients of two identical training forwards/backwards agree bit for bit (deterministic codebook
    gradient + the fixed-order wgrad reductions)."""
    p = spread_params(12, 208, 1024, seed=4)
    x = torch.from_numpy((2.0 * np.random.default_rng(9).random((30000, 12)) - 1.0).astype(np.float32)).cuda()
    grads = []
    for _ in range(2):
        m = _model(p)
        _, loss = m(x)
        loss.backward()
        torch.cuda.synchronize()
        grads.append({k: t.grad.clone() for k, t in m.named_parameters()})
    for k in grads[0]:
        assert torch.equal(grads[0][k], grads[1][k]), k


@pytest.mark.parametrize("n", [80, 500, 5000])
def test_nograd_fast_path_equals_general_path(n):
    """The no-grad host fast path (no autograd.Function, cached parameter struct / workspace size) returns exactly what
    the general path returns, follows in-place parameter updates, and notices re-allocated parameter storage."""
    p = spread_params()
    m = _model(p)
    x = torch.from_numpy(np.random.default_rng(3).standard_normal((n, 12)).astype(np.float32)).cuda()

    def general(xx):
        z, loss = m(xx)                      # grad enabled: the autograd.Function path
        return z.clone(), loss.detach().clone(), m.last_indices.clone(), m.last_reconstruction.clone()

    def fast(xx):
        with torch.no_grad():
            z, loss = m(xx)
        from lipvq_vae_b200 import tokenizer
        assert tokenizer._FAST.get(m) is not None, "fast path not taken"
        import copy
        copy.deepcopy(m)                      # the caches live outside the instance: modules stay copyable
        return z.clone(), loss.clone(), m.last_indices.clone(), m.last_reconstruction.clone()

    for rnd in range(3):
        zg, lg, ig, rg = general(x)
        zf, lf, i_f, rf = fast(x)
        assert torch.equal(ig, i_f) and torch.equal(zg, zf)
        assert torch.allclose(rg, rf, rtol=1e-6, atol=1e-7)
        assert abs(lg.item() - lf.item()) <= 1e-6 * abs(lg.item())
        with torch.no_grad():
            if rnd == 0:
                m.quantizer.codebook.mul_(1.5)            # in place: same storage, new values
            else:
                w = m.encoder[0].weight
                w.data = (w.data * 0.9).clone()           # new storage: the cached struct must be rebuilt
    # and against the oracle once, after both kinds of update
    sd = {k: v.detach().cpu().numpy() for k, v in m.state_dict().items()}
    fw = orc.forward(sd, x.cpu().numpy())
    zf, lf, i_f, rf = fast(x)
    assert abs(lf.item() - fw["loss"]) <= RTOL * abs(fw["loss"])
    assert (i_f.cpu().numpy() == fw["idx"]).mean() > 0.999


def test_pair_search_matches_oracle():
    """The opt-in CTA-pair search kernel (cta_group::2, LIPVQ_VQ_PAIR=1; read once per process, hence the subprocess):
    bit-exact indices outside the 1e-5 gap window against the float64 oracle, odd and even tile counts."""
    import subprocess, sys, textwrap
    code = textwrap.dedent("""
        import ctypes as C, numpy as np, torch
        from lipvq_vae_b200 import _native as N
        from oracle import lipvq_oracle as orc
        lib = N.lib(); dev = torch.device("cuda", 0)
        for n, D, K in ((128 * 5 + 17, 208, 1024), (4096, 64, 300), (128 * 2, 253, 512)):
            g = np.random.default_rng(n + D)
            z = g.random((n, D), dtype=np.float32); E = g.random((K, D), dtype=np.float32)
            zt, Et = torch.from_numpy(z).to(dev), torch.from_numpy(E).to(dev)
            idx = torch.full((n,), -1, dtype=torch.int64, device=dev)
            need = C.c_size_t(0)
            N.check(lib.lipvq_workspace_bytes(C.byref(N.Dims(1, 1, 1, D, K)), n, C.byref(need)))
            ws = torch.full((need.value + 4096,), 0xFF, dtype=torch.uint8, device=dev)
            st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
            N.check(lib.lipvq_vq_argmin(zt.data_ptr(), Et.data_ptr(), n, D, K, 1, idx.data_ptr(), None, None, ws.data_ptr(), ws.numel(), st))
            torch.cuda.synchronize()
            i1, d1, d2 = orc.exact_top2(z, E)
            must = (d2 - d1) > 1e-5 * np.maximum(d1, 1e-30)
            got = idx.cpu().numpy()
            assert must.mean() > 0.9 and np.array_equal(got[must], i1[must]), (n, D, K)
        print("PAIR_OK")
    """)
    env = dict(os.environ, LIPVQ_VQ_PAIR="1")
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300,
                       cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert r.returncode == 0 and "PAIR_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
