#!/usr/bin/env python
"""Hot spots of one kernel from `ncu -i X.ncu-rep --page source --csv`: stall totals, top SASS instructions by stall
samples, opcode mix.   python profiles/ncu_hot.py src.csv [top] > summary.txt"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1], errors='replace')))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
h = next(i for i, r in enumerate(rows) if 'Source' in r and '# Samples' in r)
hdr = rows[h]; ix = {n: i for i, n in enumerate(hdr)}
body = [r for r in rows[h + 1:] if len(r) == len(hdr)]
S, E = ix['# Samples'], ix['Instructions Executed']
tot_s = sum(int(r[S]) for r in body); tot_e = sum(int(r[E]) for r in body)
print(f'# {rows[0][1] if len(rows[0]) > 1 else ""}')
print(f'SASS instructions {len(body)}   stall samples {tot_s}   warp-level instructions executed {tot_e}')
stalls = [n for n in hdr if n.startswith('stall_') and 'Not Issued' not in n]
agg = {n[6:]: sum(int(r[ix[n]]) for r in body) for n in stalls}
print('stall samples by reason:', {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
print(f'--- top {top} instructions by stall samples')
for n, r in sorted(enumerate(body), key=lambda nr: -int(nr[1][S]))[:top]:
    st = {m[6:]: int(r[ix[m]]) for m in stalls if int(r[ix[m]])}
    print(f'{n:5d} {int(r[S]):7d} {100 * int(r[S]) / max(1, tot_s):5.1f}%  exec {int(r[E]):9d}  {r[ix["Source"]].strip()[:64]:64s} {st}')
mix = collections.Counter()
for r in body:
    t = r[ix['Source']].split()
    if t:
        mix[(t[1] if t[0].startswith('@') and len(t) > 1 else t[0]).split('.')[0]] += int(r[E])
print('--- opcode mix (warp-level instructions executed):', mix.most_common(24))
