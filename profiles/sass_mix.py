import csv,io,collections,re,sys
def load(f):
    lines=open(f,errors='replace').read().split('\n')
    for i,l in enumerate(lines):
        if l.startswith('"Address"'):
            start=i;break
    return list(csv.reader(io.StringIO('\n'.join(lines[start:]))))
rows=load(sys.argv[1]); hdr=rows[0]; col={h:i for i,h in enumerate(hdr)}
ops=collections.defaultdict(lambda:[0,0])
stalls=collections.Counter()
tot_i=tot_s=0
for r in rows[1:]:
    if len(r)<len(hdr) or not r[col["# Samples"]].isdigit(): continue
    src=r[1].strip()
    m=re.match(r'(@!?U?P\d+\s+)?([A-Z0-9_.]+)',src)
    op=m.group(2) if m else src
    op='.'.join(op.split('.')[:2])
    n=int(r[col['Instructions Executed']] or 0); s=int(r[col['# Samples']] or 0)
    ops[op][0]+=n; ops[op][1]+=s; tot_i+=n; tot_s+=s
    for h in hdr:
        if h.startswith('stall_') and 'Not Issued' not in h:
            stalls[h]+=int(r[col[h]] or 0)
print('total warp-instr',tot_i,'samples',tot_s)
for k,v in sorted(ops.items(),key=lambda kv:-kv[1][0])[:28]:
    print(f'{k:22s} {v[0]:12d} {v[0]/tot_i:6.3f}   samples {v[1]/tot_s:6.3f}')
print({k:round(v/tot_s,3) for k,v in stalls.most_common(8)})
