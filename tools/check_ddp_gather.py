"""2+ GPU check (torchrun): the multi-rank Trainer with the bucket gather on the parameter-gradient side stream produces the same reduced
gradients as with the gather on the compute stream (and as the mean of single-rank gradients), eager and CUDA-graph, over several steps.
usage: python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/check_ddp_gather.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import pregenerating_pde_b200 as cno
from pregenerating_pde_b200 import train

rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
torch.cuda.set_device(int(os.environ.get('LOCAL_RANK', rank)))
dist.init_process_group('nccl')
torch.backends.cudnn.allow_tf32 = False
SMALL = dict(in_dim=7, in_size=32, N_layers=2, N_res=2, N_res_neck=2, channel_multiplier=8, out_dim=3, is_time=1, nl_dim=[2, 3],
             latent_lift_proj_dim=16, lr=0.0, weight_decay=0.0)
worst = 0.0
for graph in (False, True):
    torch.manual_seed(0)
    train.GATHER_ON_SIDE_STREAM = False
    a = train.Trainer(SMALL, device='cuda', ddp=True, graph=graph)
    train.GATHER_ON_SIDE_STREAM = True
    b = train.Trainer(SMALL, device='cuda', ddp=True, graph=graph)
    b.model.load_state_dict(a.model.state_dict())
    for step in range(6):
        bt = train.synthetic_fpo_batch(4, size=32, device='cuda', seed=100 * rank + step)
        train.GATHER_ON_SIDE_STREAM = False
        la, ga = float(a.step(*bt)), a.grad_vector()
        train.GATHER_ON_SIDE_STREAM = True
        lb, gb = float(b.step(*bt)), b.grad_vector()
        err = float((ga - gb).abs().max() / ga.abs().max())
        # every rank holds the same reduced gradient
        ref = gb.clone(); dist.broadcast(ref, 0)
        same = float((gb - ref).abs().max())
        worst = max(worst, err)
        if rank == 0:
            print(f'graph={int(graph)} step {step}: loss {la:.6f} / {lb:.6f}, gradient difference {err:.2e}, across ranks {same:.1e}', flush=True)
        assert err < 1e-4 and same == 0.0, (err, same)
    a.close(); b.close()
if rank == 0:
    print('OK, worst relative gradient difference', worst)
dist.destroy_process_group()
