"""times cno_nchw_to_nhwc / cno_nhwc_to_nchw against torch's memory-format copy on the CNO activation shapes (us, GB/s of read + write)"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pregenerating_pde_b200 as cno
from pregenerating_pde_b200 import layout

cno._lib.load()
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')


def t(fn, reps=20):
    fn(); torch.cuda.synchronize()
    tot = 0.0
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / reps * 1e3


for shape in [(32, 64, 128, 128), (32, 128, 64, 64), (32, 256, 32, 32), (32, 512, 16, 16), (32, 1024, 8, 8), (32, 7, 128, 128)]:
    x = torch.randn(shape, device='cuda')
    xc = x.contiguous(memory_format=torch.channels_last)
    b = 2 * x.numel() * 4
    a, c = t(lambda: layout.to_nhwc(x)), t(lambda: x.contiguous(memory_format=torch.channels_last))
    d, e = t(lambda: layout.to_nchw(xc)), t(lambda: xc.contiguous())
    print(f'{str(shape):24s} to_nhwc {a:7.1f} us {b / a / 1e3:6.0f} GB/s (torch {c:7.1f} us) | to_nchw {d:7.1f} us {b / d / 1e3:6.0f} GB/s (torch {e:7.1f} us)')
