"""Kernel-only timing of the fused filtered_lrelu (forward WRITE / backward READ) through the plugin layer.
usage: python tools/time_flr.py [N C S_in S_out dtype iters]   (env CNO_FLR_CTAS forces the persistent grid)"""
import os, sys, math
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pregenerating_pde_b200 as cno
from oracle import cno_oracle as O

def main():
    a = sys.argv[1:]
    N, C, si, so = (int(a[0]), int(a[1]), int(a[2]), int(a[3])) if len(a) >= 4 else (32, 64, 128, 128)
    dt = {'f32': torch.float32, 'bf16': torch.bfloat16, 'f16': torch.float16}[a[4] if len(a) > 4 else 'f32']
    iters = int(a[5]) if len(a) > 5 else 20
    dev = 'cuda'
    cfg = O.lrelu_config(si, so)
    fu, fd = cfg['fu'].to(dev), cfg['fd'].to(dev)
    plug = cno.custom_ops.get_plugin('filtered_lrelu_plugin')
    px0, px1, py0, py1 = cfg['padding']; up, down = cfg['up'], cfg['down']
    e = torch.empty(0)
    nbuf = max(2, int(600e6 // (N * C * si * si * dt.itemsize)) + 1)       # rotate over > 4x L2
    xs = [torch.randn(N, C, si, si, device=dev, dtype=dt) for _ in range(min(nbuf, 8))]
    b = torch.randn(C, device=dev, dtype=dt)
    dys = [torch.randn(N, C, so, so, device=dev, dtype=dt) for _ in range(len(xs))]
    gain, slope = math.sqrt(2), 0.2
    def fwd(i):
        return plug.filtered_lrelu(xs[i % len(xs)], fu, fd, b, e, up, down, px0, px1, py0, py1, 0, 0, gain, slope, float('inf'), False, True)
    y, so_, rc = fwd(0)
    assert rc == 0, 'fused kernel refused'
    tu, td = len(cfg['fu']), len(cfg['fd'])
    pp = ((tu - 1) + (td - 1) - px0, si * up - so * down + px0 - (up - 1), (tu - 1) + (td - 1) - py0, si * up - so * down + py0 - (up - 1))
    gg = gain * up * up / (down * down)
    zb = torch.zeros(C, device=dev, dtype=dt)
    dbt = torch.zeros(C, device=dev) if os.environ.get('TIME_DB') else None
    def bwd(i):
        return plug.filtered_lrelu(dys[i % len(dys)], fd, fu, zb, so_, down, up, pp[0], pp[1], pp[2], pp[3], px0 - (tu - 1), py0 - (tu - 1), gg, slope, float('inf'), True, False, db=dbt, signs_only=not os.environ.get('TIME_CLAMPBITS'))
    assert bwd(0)[2] == 0
    res = {}
    for name, fn in (('fwd', fwd), ('bwd', bwd)):
        for i in range(5): fn(i)
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
        ev[0].record()
        for i in range(iters):
            fn(i); ev[i + 1].record()
        torch.cuda.synchronize()
        ts = sorted(ev[i].elapsed_time(ev[i + 1]) * 1e3 for i in range(iters))
        res[name] = (ts[len(ts) // 2], ts[0])
    es = dt.itemsize
    sh = so * down - (down - 1) + (td - 1); swb = ((sh + 15) // 16 * 16) // 4
    byt = N * C * (si * si * es + so * so * es + sh * swb)
    print(f"[{N},{C},{si}->{so}] {a[4] if len(a) > 4 else 'f32'} ctas={os.environ.get('CNO_FLR_CTAS', 'auto')}: "
          + ', '.join(f"{k} {v[0]:.1f} us (min {v[1]:.1f}) {byt / v[0] / 1e3:.0f} GB/s" for k, v in res.items()))

main()
