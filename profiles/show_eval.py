"""print the eval leg(s) of a bench.py JSON line:  python profiles/show_eval.py [file]"""
import json, sys
d = json.loads(open(sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/b_eval.json').read().strip().splitlines()[-1])
for name, leg in (('main', d), ('naml', d.get('naml', {}))):
    e = leg.get('eval')
    if e:
        print(name, {k: (round(v, 3) if isinstance(v, float) else v) for k, v in e.items() if k in ('news_per_s', 'impressions_per_s', 'ms_cache', 'ms_score')},
              [round(x, 4) for x in e['metrics_mean_scored']])
        print('   ', {k: round(v['ms'], 3) for k, v in e['kernels'].items()})
