"""Per-warp cycle accounting of the fused filtered_lrelu streaming loop (diagnostic variant built with -DFLR_TIMING):
   CNO_B200_LIB=tools/variants/<name>.so python tools/flr_timing.py [fwd|bwd]"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
sys.argv = [sys.argv[0]] + sys.argv[1:]
which = sys.argv[1] if len(sys.argv) > 1 else 'fwd'
import pregenerating_pde_b200 as cno
import bench_sweep as B
lib = cno._lib.load()
rows = []
import math
act = cno.cno_model.lrelu_for(64, 128, 128).to('cuda')
plug = cno.custom_ops.get_plugin('filtered_lrelu_plugin')
px0, px1, py0, py1 = act.padding; fu, fd = act.up_filter, act.down_filter
x = torch.randn(32, 64, 128, 128, device='cuda'); b = torch.randn(64, device='cuda'); e = torch.empty(0)
dy = torch.randn_like(x)
gain, slope = math.sqrt(2), 0.2
y, so, rc = plug.filtered_lrelu(x, fu, fd, b, e, 2, 2, px0, px1, py0, py1, 0, 0, gain, slope, float('inf'), False, True)
tu = td = 12
pp = (22 - px0, px0 - 1, 22 - py0, py0 - 1)
if which == 'bwd':
    for _ in range(3):
        plug.filtered_lrelu(dy, fd, fu, torch.zeros_like(b), so, 2, 2, pp[0], pp[1], pp[2], pp[3], px0 - 11, py0 - 11, gain, slope, float('inf'), True, False, signs_only=True)
else:
    for _ in range(3):
        plug.filtered_lrelu(x, fu, fd, b, e, 2, 2, px0, px1, py0, py1, 0, 0, gain, slope, float('inf'), False, True)
torch.cuda.synchronize()
buf = np.zeros(148 * 16 * 8, np.int64)
assert lib.cno_flr_prof_read(buf.ctypes.data_as(ctypes.c_void_p), buf.size) == 0
t = buf.reshape(148, 16, 8)[:, :15, :]
names = ['seg set-up', 'wait other warps', 'tma issue/wait', 'body', 'h-down + mask out', 'drain', 'total']
tot = t[..., 6].mean()
print(which, 'mean cycles per warp:', ', '.join(f'{n} {t[..., i].mean():.0f} ({100 * t[..., i].mean() / tot:.1f}%)' for i, n in enumerate(names)))
for w in range(15):
    print(f' warp {w:2d} (sched {w % 4}):', ' '.join(f'{t[:, w, i].mean():9.0f}' for i in range(7)))
print('CTA total: min %.0f max %.0f' % (t[..., 6].max(axis=1).min(), t[..., 6].max(axis=1).max()))
