"""Per-source-line instruction / sample attribution from `ncu --page source --csv --print-source sass,cuda`.
usage: python tools/ncu_line_summary.py file.csv <substring of the kernel's function name> [top_n]"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2]; topn = int(sys.argv[3]) if len(sys.argv) > 3 else 40
cur_file = None; line = None; fn = ''
per = collections.Counter(); perff = collections.Counter(); samples = collections.Counter(); txt = {}
for r in rows:
    if not r: continue
    if r[0] == 'Function Name': fn = r[1]; continue
    if r[0] == 'File Path': cur_file = r[1].split('/')[-1]; continue
    if r[0] == 'Line No': continue
    if r[0] != '' and r[0].isdigit():
        line = (cur_file, int(r[0])); txt[line] = r[1].strip()[:100]; continue
    if r[0] == '' and len(r) > 8 and r[2].startswith('0x') and want in fn:
        n = int(r[7]); per[line] += n; samples[line] += int(r[6])
        if 'FFMA2' in r[3]: perff[line] += n
tot = sum(per.values()); ts = sum(samples.values())
print('total warp instructions', tot, 'samples', ts)
for l, n in per.most_common(topn):
    print(f'{l[0]}:{l[1]:4d} {n:10d} {100*n/tot:5.1f}% (ffma2 {perff[l]:9d}) smp {100*samples[l]/max(ts,1):5.1f}% | {txt.get(l, "")}')
