"""Summarise an ncu source-page CSV: executed-instruction mix by opcode and stall samples by reason, plus the top
stalled instructions.  usage: python tools/ncu_src_summary.py src.csv [top_n]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
# the page may hold several kernels: "Kernel Name" row, header row, instruction rows; pick section argv[3] (default 0)
starts = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
sec = int(sys.argv[3]) if len(sys.argv) > 3 else 0
hi = starts[sec]; end = starts[sec + 1] - 1 if sec + 1 < len(starts) else len(rows)
print('kernel:', rows[hi - 1][1][:110])
h = rows[hi]; col = {c: i for i, c in enumerate(h)}
rows = rows[:end]
ex, smp = collections.Counter(), collections.Counter()
stall = collections.Counter()
recs = []
for r in rows[hi + 1:]:
    if len(r) < len(h): continue
    src = r[col['Source']].strip()
    toks = src.split()
    op = toks[1] if toks[0].startswith('@') else toks[0]
    op = op.split('.')[0]
    n = int(r[col['Instructions Executed']]); s = int(r[col['# Samples']])
    ex[op] += n; smp[op] += s
    for c, i in col.items():
        if c.startswith('stall_') and 'Not Issued' not in c:
            stall[c] += int(r[i] or 0)
    recs.append((s, n, r[col['Address']][-5:], src))
tot = sum(ex.values()); ts = sum(smp.values())
print(f'total warp instructions {tot}, samples {ts}')
for op, n in ex.most_common(22):
    print(f'  {op:10s} {n:12d} {100*n/tot:5.1f}% exec   {100*smp[op]/max(ts,1):5.1f}% samples')
print('stall reasons:', ', '.join(f'{k[6:]}={100*v/max(ts,1):.1f}%' for k, v in stall.most_common(8)))
print('top stalled instructions:')
for s, n, a, src in sorted(recs, reverse=True)[:topn]:
    print(f'  {100*s/max(ts,1):5.2f}%  x{n:9d}  {a}  {src[:90]}')
