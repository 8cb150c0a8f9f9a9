"""World-size-2 gloo tests of the data-parallel host logic (lipvq_vae_b200/dist.py) on CPU.
The CUDA kernels are not called here: the oracle stands in for each rank's local forward/backward
so the test checks exactly what the wrapper adds -- row sharding, n_total scaling and the single
flat all-reduce reproduce the single-process gradient."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

from lipvq_vae_b200.dist import DataParallelTokenizer, FlatBucket, shard_rows
from oracle import lipvq_oracle as orc


def test_shard_rows_tiles_and_aligns():
    for n in (0, 7, 10, 2560, 100003, 655360):
        for w in (1, 2, 3, 8):
            spans = [shard_rows(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
                assert a1 == b0 and a0 <= a1
            for s0, s1 in spans[:-1]:
                assert (s0 % 10 == 0 or s0 == n) and (s1 % 10 == 0 or s1 == n)
    with pytest.raises(ValueError):
        shard_rows(10, 2, 2)


def test_flat_bucket_roundtrip():
    shapes = [(3, 4), (5,), (2, 2, 2)]
    b = FlatBucket(shapes, "cpu", extra=1)
    ts = [torch.randn(*s) for s in shapes]
    b.pack(ts)
    assert b.flat.numel() == 12 + 5 + 8 + 1
    out = [torch.zeros(*s) for s in shapes]
    b.unpack_into(out)
    for a, c in zip(ts, out):
        assert torch.equal(a, c)
    b.pack([None, ts[1], None])
    assert b.views[0].abs().sum() == 0 and torch.equal(b.views[1], ts[1])


class _Holder(nn.Module):
    """parameter container with the reference names; forward is never called on CPU"""

    def __init__(self, p):
        super().__init__()
        self.ps = nn.ParameterList([nn.Parameter(torch.from_numpy(p[k].copy())) for k in orc.PARAM_NAMES])
        self.n_total_rows = None


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n_total, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        p = orc.init_params(12, 24, 64, 32, seed=3)
        if rank != 0:  # broadcast_params must overwrite whatever a non-zero rank holds
            p_bad = {k: v + 1.0 for k, v in p.items()}
        m = _Holder(p if rank == 0 else p_bad)
        dp = DataParallelTokenizer(m)
        for t, k in zip(m.ps, orc.PARAM_NAMES):
            assert np.array_equal(t.detach().numpy(), p[k]), "broadcast failed for " + k
        x = np.random.default_rng(5).standard_normal((n_total, 12)).astype(np.float32)
        s0, s1 = shard_rows(n_total, rank, world)
        xl = x[s0:s1]
        assert dp.global_rows(xl.shape[0], "cpu") == n_total
        fw = orc.forward(p, xl)
        gr = orc.backward(p, xl, fw, n_total=n_total)           # local grads carry 1/N_global
        for t, k in zip(m.ps, orc.PARAM_NAMES):
            t.grad = torch.from_numpy(gr[k].astype(np.float32))
        # local loss contribution = local SSEs over the global denominators
        loss_local = torch.tensor(fw["sse_r"] / (n_total * 12) + 0.5 * fw["sse_q"] / (n_total * 24),
                                  dtype=torch.float32)
        loss = dp.allreduce_grads(loss_local)
        counts = torch.from_numpy(np.bincount(fw["idx"], minlength=64).astype(np.float32))
        sums = torch.zeros(64, 24)
        sums.index_add_(0, torch.from_numpy(fw["idx"]), torch.from_numpy(fw["z_e"]))
        dp.allreduce_codebook_stats(counts, sums)
        if rank == 0:
            np.savez(os.path.join(out_dir, "dp.npz"), loss=loss.numpy(), counts=counts.numpy(), sums=sums.numpy(),
                     **{k: t.grad.numpy() for t, k in zip(m.ps, orc.PARAM_NAMES)})
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_two_rank_gloo_matches_single_process(tmp_path):
    n_total = 250                                               # ragged: 130 + 120 rows
    port = _free_port()
    mp.spawn(_worker, args=(2, port, n_total, str(tmp_path)), nprocs=2, join=True)
    got = np.load(os.path.join(str(tmp_path), "dp.npz"))
    p = orc.init_params(12, 24, 64, 32, seed=3)
    x = np.random.default_rng(5).standard_normal((n_total, 12)).astype(np.float32)
    fw = orc.forward(p, x)
    gr = orc.backward(p, x, fw)
    assert abs(float(got["loss"]) - fw["loss"]) < 1e-5 * abs(fw["loss"])
    for k in orc.PARAM_NAMES:
        assert orc.rel_err(got[k], gr[k]) < 1e-5, k
    assert np.array_equal(got["counts"], np.bincount(fw["idx"], minlength=64).astype(np.float32))
    sums = np.zeros((64, 24), np.float64)
    np.add.at(sums, fw["idx"], fw["z_e"].astype(np.float64))
    assert orc.rel_err(got["sums"], sums) < 1e-5
