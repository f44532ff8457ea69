"""world_size-2 gloo test of the data-parallel training step (CPU): sharding the batch over two ranks with the Trainer's
bucketed flat-buffer gradient all-reduce (hooks fire during backward, one collective per bucket) gives the same (averaged)
gradients as a single process on the whole batch
(the loss is a ratio of batch means, so each rank's loss is computed on its shard and gradients are averaged:
equality holds when both shards share the denominator, which the test arranges)."""
import os, sys, subprocess, tempfile
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = r'''
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, os.environ['CNO_ROOT'])
import pregenerating_pde_b200 as pkg
from pregenerating_pde_b200 import train, cno_model
rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
dist.init_process_group('gloo', rank=rank, world_size=world)
cfg = dict(in_dim=7, in_size=16, N_layers=1, N_res=1, N_res_neck=1, channel_multiplier=4, out_dim=3, is_time=1, nl_dim=[2, 3], latent_lift_proj_dim=8, lr=1e-2)
torch.manual_seed(0)
tr = train.Trainer(cfg, device='cpu', ddp=world > 1, buckets=3)
assert (len(tr.buckets) == 3 and tr.flat_grad is not None) if world > 1 else (tr.flat_grad is None and not tr.buckets)   # a single rank keeps no flat buffer
g = torch.Generator().manual_seed(3)
with torch.no_grad():
    for p in tr.model.parameters():
        p.add_(0.05 * torch.randn(p.shape, generator=g))
t, x, y = train.synthetic_fpo_batch(4, size=16, device='cpu', seed=5)
y[2:] = y[:2].flip(0).abs() * y[2:].sign()            # both shards have the same mean |y| -> identical loss denominators
n = 4 // world
sl = slice(rank * n, (rank + 1) * n)
for p in tr.params:
    p.grad = None
loss = tr.model.training_loss(tr.model(x[sl], t[sl]), y[sl])
loss.backward()                                        # the hooks launch each bucket's all-reduce as soon as it is complete
if world > 1:
    assert any(tr._launched), 'no bucket was reduced from inside backward'
    tr._finish_allreduce()
grads = {k: (p.grad.detach().clone() if p.grad is not None else torch.zeros_like(p)) for k, p in tr.model.named_parameters() if p.requires_grad}   # no gradient: a convolution bias in front of a CIN (cin.conv_before_film)
if world > 1:                                          # after the reduce every gradient is a view of the flat buffer (AdamW reads it in place)
    assert all(p.grad.untyped_storage().data_ptr() == tr.flat_grad.untyped_storage().data_ptr() for p in tr.params)
l2 = tr.step(t[sl], x[sl], y[sl])                       # a whole step (grads to None, forward, backward, gather + reduce, AdamW)
if rank == 0:
    torch.save(grads, os.environ['CNO_OUT'])
dist.destroy_process_group()
'''


def _run(world, out):
    env = dict(os.environ, CNO_ROOT=ROOT, CNO_OUT=out, MASTER_ADDR='127.0.0.1', MASTER_PORT=str(29600 + world), WORLD_SIZE=str(world),
               OMP_NUM_THREADS='2')
    procs = [subprocess.Popen([sys.executable, '-c', WORKER], env=dict(env, RANK=str(r))) for r in range(world)]
    for p in procs:
        assert p.wait(timeout=600) == 0


def test_ddp_two_ranks_equals_single_process():
    with tempfile.TemporaryDirectory() as d:
        a, b = os.path.join(d, 'one.pt'), os.path.join(d, 'two.pt')
        _run(1, a); _run(2, b)
        one, two = torch.load(a), torch.load(b)
        assert one.keys() == two.keys()
        # gradients that are pure round-off (e.g. conv biases in front of an instance norm) are compared on the global scale
        scale = max(float(v.abs().max()) for v in one.values())
        worst = max(float((one[k] - two[k]).abs().max()) for k in one) / scale
        assert worst < 1e-5, worst
