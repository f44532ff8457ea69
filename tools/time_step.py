"""eager vs CUDA-graph timing of the public-API step (fwd + autograd bwd) at the bench shape"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pregenerating_pde_b200 as cno
act = cno.cno_model.lrelu_for(64, 128, 128).cuda()
F = cno.filtered_lrelu.filtered_lrelu
x = torch.randn(32, 64, 128, 128, device='cuda'); b = torch.randn(64, device='cuda'); dy = torch.randn_like(x)
kw = dict(up=2, down=2, padding=act.padding)
def step():
    xd, bd = x.detach().requires_grad_(True), b.detach().requires_grad_(True)
    y = F(xd, act.up_filter, act.down_filter, bd, **kw)
    return torch.autograd.grad(y, [xd, bd], dy)
def fwd():
    return F(x.detach().requires_grad_(True), act.up_filter, act.down_filter, b, **kw)
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
print('eager step us', t(step), 'eager fwd us', t(fwd))
s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    g1 = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g1): o1 = step()
    g2 = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g2): o2 = fwd()
torch.cuda.current_stream().wait_stream(s)
print('graph step us', t(g1.replay), 'graph fwd us', t(g2.replay))
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3): g1.replay()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by='cuda_time_total', row_limit=8, max_name_column_width=60))
