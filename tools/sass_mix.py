"""Static instruction mix of one kernel from `cuobjdump -sass`: splits the function at branch targets and prints an opcode
histogram per straight-line region (largest FFMA2 holders first).  Used to count the non-FMA instructions of the streaming
body before spending GPU time.   usage: sass_mix.py file.sass [min_ffma2]"""
import re, sys, collections
lines = open(sys.argv[1]).read().splitlines()
minf = int(sys.argv[2]) if len(sys.argv) > 2 else 20
ins = []
for l in lines:
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m:
        addr = int(m.group(1), 16); txt = m.group(2).strip()
        ins.append((addr, txt))
targets = set()
for a, t in ins:
    m = re.search(r'\b(BRA|BRA\.U|BRA\.U\.ANY|BSSY\S*|CALL\S*)\b.*?(0x[0-9a-f]+)\s*$', t)
    if m and 'BRA' in t:
        targets.add(int(m.group(2), 16))
cuts = sorted(targets)
def opcode(t):
    t = re.sub(r'^@!?U?P\d+\s+', '', t)
    return t.split()[0].split('.')[0]
regions = []
cur = []
for a, t in ins:
    if a in targets and cur:
        regions.append(cur); cur = []
    cur.append((a, t))
    if re.search(r'\bBRA\b', t) or t.startswith('EXIT'):
        regions.append(cur); cur = []
if cur: regions.append(cur)
tot = collections.Counter(opcode(t) for _, t in ins)
print('whole function:', len(ins), 'instructions;', dict(tot.most_common(12)))
for r in sorted(regions, key=lambda r: -sum('FFMA2' in t for _, t in r)):
    nf = sum('FFMA2' in t for _, t in r)
    if nf < minf: break
    c = collections.Counter(opcode(t) for _, t in r)
    print(f'region {r[0][0]:#06x}..{r[-1][0]:#06x}: {len(r)} instr, FFMA2 {nf}, other {len(r) - nf}')
    print('   ', ', '.join(f'{k} {v}' for k, v in c.most_common()))
