import json,sys
for f in sys.argv[1:]:
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); k=d['kernels']
        print(f, round(d['ms_per_step'],4), {n: round(k[n]['ms_per_step'],4) for n in ('enc.dx','enc.dx_rows','enc.dwqkv') if n in k})
    except Exception as e: print(f,'ERR',e)
