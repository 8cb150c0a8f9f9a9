import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from tests.test_gpu_scale_parity import spread_params, _model
from lipvq_vae_b200 import TokenizerTrainer, shard_rows
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
n_total = 20480
p = spread_params(12, 64, 256, seed=21)
x = (2.0 * np.random.default_rng(77).random((n_total, 12)) - 1.0).astype(np.float32)
m = _model(p, 1, dev="cuda:%d" % rank)
s0, s1 = shard_rows(n_total, rank, world)
xl = torch.from_numpy(x[s0:s1]).cuda()
m.n_total_rows = n_total
with torch.no_grad():
    m(xl); t_ng = m.last_loss_terms.clone()
_, l = m(xl); t_g = m.last_loss_terms.clone()
m.n_total_rows = None
full = _model(p, 1, dev="cuda:%d" % rank)
xf = torch.from_numpy(x).cuda()
with torch.no_grad():
    full(xf); f_ng = full.last_loss_terms.clone()
_, lf = full(xf); f_g = full.last_loss_terms.clone()
tot = t_g.clone(); dist.all_reduce(tot)
print(rank, "local nograd", t_ng.tolist(), "local grad", t_g.tolist(), "sum", tot.tolist(), "full nograd", f_ng.tolist(), "full grad", f_g.tolist(), flush=True)
pp = p if rank == 0 else {k: v + 0.5 for k, v in p.items()}
mt = _model(pp, 1, dev="cuda:%d" % rank)
tr = TokenizerTrainer(mt)
for k, t in mt.named_parameters():
    d = (t.detach().cpu() - torch.from_numpy(p[k])).abs().max().item()
    if d > 0:
        print(rank, "param differs after broadcast", k, d, flush=True)
_, loss = tr.step(xl, n_total)
print(rank, "trainer loss", loss.item(), flush=True)
dist.destroy_process_group()
