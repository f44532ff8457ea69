"""Execution-count histogram of an ncu source-page CSV (groups SASS instructions by how often they ran: body, h-down loop,
per-body overhead, spin loops ...), with the opcode mix of each group.  usage: ncu_exec_hist.py src.csv [section] [groups]"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
sec = int(sys.argv[2]) if len(sys.argv) > 2 else 0
ng = int(sys.argv[3]) if len(sys.argv) > 3 else 12
starts = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
hi = starts[sec]; end = starts[sec + 1] - 1 if sec + 1 < len(starts) else len(rows)
h = rows[hi]; col = {c: i for i, c in enumerate(h)}
print('kernel:', rows[hi - 1][1][:120])
groups = collections.defaultdict(list)
for r in rows[hi + 1:end]:
    if len(r) < len(h): continue
    src = r[col['Source']].strip(); toks = src.split()
    op = (toks[1] if toks[0].startswith('@') else toks[0]).split('.')[0]
    groups[int(r[col['Instructions Executed']])].append((op, r[col['Address']][-5:]))
tot = sum(n * len(v) for n, v in groups.items())
print('total warp instructions', tot)
for n, v in sorted(groups.items(), key=lambda kv: -kv[0] * len(kv[1]))[:ng]:
    c = collections.Counter(op for op, _ in v)
    print(f'x{n:8d}  {len(v):4d} instrs {100 * n * len(v) / tot:5.1f}%  [{v[0][1]}..{v[-1][1]}]  ' + ', '.join(f'{k} {m}' for k, m in c.most_common(9)))
