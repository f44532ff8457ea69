"""Per-kernel and per-category breakdown of ONE training step from an ncu launch list
(`ncu --metrics gpu__time_duration.sum --csv`) of `bench.py --workload cno_train --graph 0 --steps 1 --warmup 3`.
The step is found as the period of the kernel-name sequence at the end of the list.  usage: launch_breakdown.py launches.csv"""
import csv, sys, collections, re

rows = []
with open(sys.argv[1], newline='') as f:
    lines = [l for l in f if not l.startswith('==')]
rd = csv.reader(lines)
hdr = next(rd)
ki, vi, ui = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
for r in rd:
    if len(r) <= vi:
        continue
    v = float(r[vi].replace(',', ''))
    u = r[ui]
    us = v / 1e3 if u in ('ns', 'nsecond') else v * (1e3 if u in ('ms', 'msecond') else 1.0)
    rows.append((r[ki], us))
names = [n for n, _ in rows]
period, whole = None, False
for P in range(200, len(names) // 2):
    if names[-P:] == names[-2 * P:-P]:
        period = P
        break
if period is None:                     # e.g. a list cut by ncu's -c limit: summarise everything that was captured
    print('no repeating step found: summarising the whole list')
    period = len(rows)
    whole = True
step = rows[-period:]
tot = sum(us for _, us in step)
print(('whole launch list' if whole else 'one CNO_time training step (eager)') +
      f' - ncu per-launch times (cold cache, serialised): {period} launches, sum {tot / 1e3:.2f} ms')
agg = collections.defaultdict(lambda: [0.0, 0])
for n, us in step:
    agg[n][0] += us; agg[n][1] += 1
for n, (us, c) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:28]:
    print(f'  {us / 1e3:7.3f} ms {100 * us / tot:5.1f}%  x{c:4d}  {n[:110]}')
CATS = [('flr fused (ours)', r'flr_fused'), ('cin / film (ours)', r'cin_|film_'), ('upfirdn2d / bias_act (ours)', r'upfirdn2d|bias_act'),
        ('cudnn layout transposes', r'nchwToNhwc|nhwcToNchw'), ('conv kernels', r'wgrad|dgrad|fprop|conv|implicit_gemm|xmma|sgemm_largek'),
        ('optimizer', r'multi_tensor_apply'), ('torch reductions', r'reduce_kernel'), ('gemm / gemv', r'gemm|gemv|cublas'),
        ('torch elementwise / copy / cat', r'elementwise|CatArray|copy|fill')]
cat = collections.defaultdict(lambda: [0.0, 0])
for n, us in step:
    for c, pat in CATS:
        if re.search(pat, n):
            cat[c][0] += us; cat[c][1] += 1
            break
    else:
        cat['other'][0] += us; cat['other'][1] += 1
print('by category:')
for c, (us, k) in sorted(cat.items(), key=lambda kv: -kv[1][0]):
    print(f'  {us / 1e3:7.3f} ms {100 * us / tot:5.1f}%  x{k:4d}  {c}')
