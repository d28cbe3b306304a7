"""Pretty-print the JSON line(s) of bench.py logs: python profiles/show_bench.py <log>..."""
import json
import sys

for f in sys.argv[1:]:
    for l in open(f):
        if l.startswith('{'):
            d = json.loads(l)
            print(f, 'value %.0f  ms/step %.3f  e2e %.0f  launches %s  clocks %s' % (
                d['value'], d['ms_per_step'], d['e2e']['value'], d.get('gpu_launches'), d.get('clocks')))
            if d.get('eval'):
                print('  eval', d['eval'])
            if d.get('cpu_baseline'):
                print('  cpu', d['cpu_baseline'])
            print('  roofline', d.get('roofline'))
            for k, v in (d.get('kernels') or {}).items():
                print('  %-16s %8.3f ms  x%-3g share %.3f  %s' % (k, v['ms_per_step'], v['launches_per_step'], v['share'],
                                                                  ('%.1f TF' % v['tflops']) if 'tflops' in v else ''))
        elif 'Error' in l or 'error' in l:
            print(l[:300])
