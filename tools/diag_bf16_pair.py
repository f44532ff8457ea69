"""diagnostic: where does the 16-bit film_lrelu backward deviate from the fp32 composition?"""
import sys, math, os, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pregenerating_pde_b200 as cno
from pregenerating_pde_b200 import cno_model
from oracle import cno_oracle as O

rel = lambda a, b: float((a.detach().double().cpu() - b.detach().double().cpu()).abs().max() / b.detach().double().cpu().abs().max())
for dtype in (torch.bfloat16, torch.float16):
    torch.manual_seed(9)
    N, C, S = 3, 6, 32
    film = cno_model.FILM(C, dim=[2, 3], s=S)
    act = cno_model.LReLu(C, C, S, S, **cno_model._filter_args(S, S, 2.0001, 0.8))
    with torch.no_grad():
        for p_ in list(film.parameters()) + [act.bias]:
            p_.add_(0.3 * torch.randn_like(p_))
    x = (torch.randn(N, C, S, S) * 1.5 + 0.3).to(dtype)
    t = torch.rand(N)
    xr = x.float().requires_grad_(True)
    z = film(xr, t)
    y_ref = O.filtered_lrelu(z, act.up_filter, act.down_filter, act.bias, up=act.up_factor, down=act.down_factor, padding=act.padding,
                             gain=math.sqrt(2), slope=0.2)
    dy = torch.randn_like(y_ref).to(dtype)
    dz_ref, = torch.autograd.grad(y_ref, z, dy.float(), retain_graph=True)
    dx_ref, = torch.autograd.grad(z, xr, dz_ref)
    fg, ag = film.cuda(), act.cuda()
    for fuse in (True, False):
        cno.cin.FUSE_LRELU = fuse
        xg = x.cuda().requires_grad_(True)
        if fuse:
            y = cno.cin.film_lrelu(fg, ag, xg, t.cuda())
            zg = None
        else:
            zg = fg(xg, t.cuda())
            y = ag(zg)
        dx, = torch.autograd.grad(y, xg, dy.cuda(), retain_graph=not fuse)
        line = f'{dtype} fuse={fuse}: y {rel(y, y_ref):.2e} dx {rel(dx, dx_ref):.2e}'
        if zg is not None:
            dz, = torch.autograd.grad(y, zg, dy.cuda())
            line += f' z {rel(zg, z):.2e} dz {rel(dz, dz_ref):.2e}'
        e = (dx.float().cpu() - dx_ref).abs()
        idx = e.flatten().argmax()
        line += f' worst at {tuple(int(v) for v in torch.unravel_index(idx, e.shape))} got {float(dx.flatten()[idx]):.4f} want {float(dx_ref.flatten()[idx]):.4f}'
        print(line)
        print('   per-plane max err / plane max:', (e.amax(dim=(2, 3)) / dx_ref.abs().amax(dim=(2, 3))).flatten().tolist()[:8])
