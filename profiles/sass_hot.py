"""Top SASS instructions by stall samples from an `ncu --page source --csv` export (one kernel)."""
import csv, io, sys
lines = open(sys.argv[1], errors='replace').read().split('\n')
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rows = list(csv.reader(io.StringIO('\n'.join(lines[start:]))))
hdr = rows[0]; col = {h: i for i, h in enumerate(hdr)}
st = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
body = [r for r in rows[1:] if len(r) >= len(hdr) and r[col['# Samples']].isdigit()]
tot = sum(int(r[col['# Samples']]) for r in body) or 1
top = sorted(range(len(body)), key=lambda i: -int(body[i][col['# Samples']]))[:int(sys.argv[2]) if len(sys.argv) > 2 else 25]
for i in sorted(top):
    r = body[i]
    s = int(r[col['# Samples']])
    why = sorted(((int(r[col[h]] or 0), h[6:]) for h in st), reverse=True)[:2]
    print(f'{i:5d} {s / tot:6.3f} {r[col["Instructions Executed"]]:>9s}  {r[1].strip()[:70]:70s} {why[0][1]}:{why[0][0]} {why[1][1]}:{why[1][0]}')
