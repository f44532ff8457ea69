"""CIN at 256^2 planes: cluster kernel vs streaming kernel (env CNO_CIN_NO_CLUSTER), CUDA-graph timed (no host gaps)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pregenerating_pde_b200 as cno

def tgraph(fn, reps=8):
    for i in range(2): fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(reps): fn(i)
    best = 1e9
    for _ in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3 / reps)
    return best

S = int(sys.argv[1]) if len(sys.argv) > 1 else 256
for dt in (torch.float32, torch.bfloat16):
    N, C = 8, 64
    xs = [torch.randn(N, C, S, S, device='cuda', dtype=dt) for _ in range(4)]
    gamma, beta = torch.randn(C, device='cuda'), torch.randn(C, device='cuda')
    sc, sh = torch.randn(N, C, device='cuda'), torch.randn(N, C, device='cuda')
    dy = torch.randn(N, C, S, S, device='cuda', dtype=dt)
    for mode in ('cluster', 'stream'):
        if mode == 'stream': os.environ['CNO_CIN_NO_CLUSTER'] = '1'
        else: os.environ.pop('CNO_CIN_NO_CLUSTER', None)
        def fwd(i):
            with torch.no_grad():
                return cno.cin.cin(xs[i % 4], gamma, beta, sc, sh)
        def both(i):
            x = xs[i % 4].detach().requires_grad_(True)
            y = cno.cin.cin(x, gamma, beta, sc, sh)
            return torch.autograd.grad(y, [x], dy)
        f = tgraph(fwd); fb = tgraph(both)
        nb = xs[0].numel() * xs[0].element_size()
        print(f'[{N},{C},{S},{S}] {dt} {mode}: fwd {f:.1f} us ({2 * nb / f / 1e3:.0f} GB/s)  bwd {fb - f:.1f} us ({3 * nb / (fb - f) / 1e3:.0f} GB/s)')
