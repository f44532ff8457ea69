"""torch.profiler (CUPTI) timeline of the CNO training step: real (warm, overlapped) kernel durations by name + GPU busy time.
usage: python tools/prof_train_torch.py [graph 0|1] [batch] [train|finetune] [fp32|bf16]"""
import os, sys, collections, re
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pregenerating_pde_b200 as cno
from torch.profiler import profile, ProfilerActivity
graph = int(sys.argv[1]) if len(sys.argv) > 1 else 1
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
torch.backends.cudnn.benchmark = True
mode = sys.argv[3] if len(sys.argv) > 3 else 'train'
ac = torch.bfloat16 if (len(sys.argv) > 4 and sys.argv[4] == 'bf16') else None
if mode == 'finetune':
    ft = cno.finetune.wide_finetune_config()
    tr = cno.train.Trainer(device='cuda', graph=bool(graph), mode='finetune', ft=ft, autocast=ac)
    b = cno.train.synthetic_finetune_batch(B, ft, device='cuda')
else:
    tr = cno.train.Trainer(device='cuda', graph=bool(graph), autocast=ac)
    b = cno.train.synthetic_fpo_batch(B, device='cuda')
for _ in range(5): tr.step(*b)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3): tr.step(*b)
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
agg = collections.defaultdict(lambda: [0, 0.0])
t0 = min(e.time_range.start for e in ev); t1 = max(e.time_range.end for e in ev)
for e in ev:
    k = re.sub(r'^void ', '', e.name).replace('(anonymous namespace)::', ''); k = re.sub(r'\(.*', '', k)[:90]
    agg[k][0] += 1; agg[k][1] += e.time_range.end - e.time_range.start
tot = sum(v[1] for v in agg.values())
print(f'graph={graph} batch={B}: 3 steps span {(t1 - t0) / 3e3:.2f} ms/step, kernel time sum {tot / 3e3:.2f} ms/step, {len(ev) // 3} kernels/step')
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:32]:
    print(f'{t / 3e3:8.3f} ms {100 * t / tot:5.1f}%  x{n // 3:4d}  {k}')
