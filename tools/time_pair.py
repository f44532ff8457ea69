"""LReLu(FILM(x, t)) forward + backward through cin.film_lrelu (the prologue-fused pair).  usage: python tools/time_pair.py [N C S dtype iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pregenerating_pde_b200 as cno
from pregenerating_pde_b200 import cno_model
a = sys.argv[1:]
N, C, S = (int(a[0]), int(a[1]), int(a[2])) if len(a) >= 3 else (32, 64, 128)
dt = {'f32': torch.float32, 'bf16': torch.bfloat16}[a[3] if len(a) > 3 else 'f32']
iters = int(a[4]) if len(a) > 4 else 6
film = cno_model.FILM(C).cuda()
act = cno_model.lrelu_for(C, S, S).cuda()
xs = [torch.randn(N, C, S, S, device='cuda', dtype=dt).requires_grad_(True) for _ in range(4)]
t = torch.rand(N, device='cuda')
dy = torch.randn(N, C, S, S, device='cuda', dtype=dt)
for i in range(iters):
    y = cno.cin.film_lrelu(film, act, xs[i % 4], t)
    assert 'FilmLRelu' in type(y.grad_fn).__name__
    torch.autograd.grad(y, [xs[i % 4]] + list(film.parameters()) + [act.bias], dy)
torch.cuda.synchronize()
print('pair ok', N, C, S, dt)
