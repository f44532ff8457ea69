"""Kernel-only timing of the fused FILM/CIN forward + backward.  usage: python tools/time_cin.py [N C S dtype iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pregenerating_pde_b200 as cno
from pregenerating_pde_b200 import cno_model
a = sys.argv[1:]
N, C, S = (int(a[0]), int(a[1]), int(a[2])) if len(a) >= 3 else (32, 64, 128)
dt = {'f32': torch.float32, 'bf16': torch.bfloat16}[a[3] if len(a) > 3 else 'f32']
iters = int(a[4]) if len(a) > 4 else 10
film = cno_model.FILM(C).cuda()
xs = [torch.randn(N, C, S, S, device='cuda', dtype=dt).requires_grad_(True) for _ in range(4)]
t = torch.rand(N, device='cuda')
dy = torch.randn(N, C, S, S, device='cuda', dtype=dt)
for i in range(iters):
    y = film(xs[i % 4], t)
    torch.autograd.grad(y, [xs[i % 4]] + list(film.parameters()), dy)
torch.cuda.synchronize()
print('cin ok', N, C, S, dt)
