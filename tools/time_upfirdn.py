"""Kernel-only timing of upfirdn2d (fast path).  usage: python tools/time_upfirdn.py [up|down] [N C S dtype iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pregenerating_pde_b200 as cno
from oracle import cno_oracle as O
a = sys.argv[1:]
mode = a[0] if a else 'up'
N, C, S = (int(a[1]), int(a[2]), int(a[3])) if len(a) >= 4 else (32, 64, 128)
dt = {'f32': torch.float32, 'bf16': torch.bfloat16}[a[4] if len(a) > 4 else 'f32']
iters = int(a[5]) if len(a) > 5 else 20
f = O.design_lowpass_filter(12, 64 / 2.0001, 2 * (0.8 * 64 - 64 / 2.0001), 256).cuda()
kw = dict(up=2, padding=[6, 5, 6, 5], gain=4) if mode == 'up' else dict(down=2, padding=[5, 5, 5, 5])
xs = [torch.randn(N, C, S, S, device='cuda', dtype=dt) for _ in range(6)]
fn = lambda i: cno.upfirdn2d.upfirdn2d(xs[i % 6], f, **kw)
for i in range(5): y = fn(i)
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
ev[0].record()
for i in range(iters):
    fn(i); ev[i + 1].record()
torch.cuda.synchronize()
ts = sorted(ev[i].elapsed_time(ev[i + 1]) * 1e3 for i in range(iters))
byt = (xs[0].numel() + y.numel()) * dt.itemsize
print(f'upfirdn2d {mode} [{N},{C},{S}] {dt}: median {ts[len(ts)//2]:.1f} us (min {ts[0]:.1f}) {byt / ts[len(ts)//2] / 1e3:.0f} GB/s')
