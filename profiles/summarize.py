"""Summarise ncu outputs brought back in gpurun_out/ into small committed text files.
  python profiles/summarize.py launches gpurun_out/X_launches.csv profiles/X_launches.txt
  python profiles/summarize.py full gpurun_out/X.ncu-rep profiles/X_full.txt
  python profiles/summarize.py traffic gpurun_out/X.ncu-rep NRMS profiles/r1_traffic.json
      (X = `ncu --set full --profile-from-start off ... bench.py --profile-step`: exactly one train step;
       bench.py reads the json for roofline.traffic)
"""
import json
import os
import csv
import io
import re
import subprocess
import sys
from collections import defaultdict


def launches(src, dst):
    rows = [l for l in open(src, errors='replace') if l.startswith('"')]
    rd = csv.DictReader(io.StringIO(''.join(rows)))
    agg = defaultdict(lambda: [0, 0.0])
    total = 0.0
    for r in rd:
        if r.get('Metric Name') != 'gpu__time_duration.sum':
            continue
        v = float(r['Metric Value'].replace(',', ''))
        unit = r.get('Metric Unit', 'ns')
        ns = v * {'ns': 1, 'us': 1e3, 'usecond': 1e3, 'ms': 1e6, 'msecond': 1e6, 'nsecond': 1}.get(unit, 1)
        name = re.sub(r'\(.*', '', r['Kernel Name'])[:110]
        agg[name][0] += 1
        agg[name][1] += ns
        total += ns
    with open(dst, 'w') as f:
        f.write(f'# ncu --metrics gpu__time_duration.sum --clock-control none ; source {src}\n')
        f.write(f'# per-launch times are cold-cache and serialised: compare SHARES\n')
        f.write(f'{"kernel":110s} {"launches":>8s} {"total_us":>12s} {"avg_us":>10s} {"share":>7s}\n')
        for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f'{k:110s} {n:8d} {ns / 1e3:12.1f} {ns / 1e3 / n:10.2f} {ns / total:7.3f}\n')
        f.write(f'{"TOTAL":110s} {sum(v[0] for v in agg.values()):8d} {total / 1e3:12.1f}\n')


_OLD_KEYS = ('dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__time_duration.sum', 'gpu__dram_throughput',
        'dram__cycles_active', 'sm__pipe_tensor', 'sm__warps_active.avg.pct', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'sm__throughput.avg.pct', 'smsp__inst_executed.sum',
        'launch__occupancy_limit', 'sm__inst_executed_pipe_tensor', 'lts__t_bytes.sum', 'l1tex__t_bytes.sum',
        'smsp__cycles_active.avg', 'sm__cycles_elapsed.max', 'launch__shared_mem_per_block')


def raw_page(src):
    """`ncu -i X.ncu-rep --page raw --csv`, or that output already saved as X.csv on the GPU box."""
    if src.endswith('.csv'):
        return ''.join(l for l in open(src, errors='replace') if l.startswith('"'))
    return subprocess.run(['ncu', '-i', src, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout


COLS = (('time_us', 'gpu__time_duration.sum'), ('dramR_MB', 'dram__bytes_read.sum'), ('dramW_MB', 'dram__bytes_write.sum'),
        ('dram%', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'),
        ('l2%', 'lts__throughput.avg.pct_of_peak_sustained_elapsed'),
        ('sm%', 'sm__throughput.avg.pct_of_peak_sustained_elapsed'),
        ('tensor%', 'sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active'),
        ('issue%', 'smsp__issue_active.avg.pct_of_peak_sustained_active'),
        ('warps%', 'sm__warps_active.avg.pct_of_peak_sustained_active'),
        ('regs', 'launch__registers_per_thread'), ('grid', 'launch__grid_size'), ('block', 'launch__block_size'))
UNIT = {'ns': 1e-3, 'nsecond': 1e-3, 'us': 1, 'usecond': 1, 'ms': 1e3, 'msecond': 1e3,
        'byte': 1e-6, 'Kbyte': 1e-3, 'Mbyte': 1, 'Gbyte': 1e3}


def full(src, dst):
    """One line per profiled launch: the `ncu --set full` numbers the roofline discussion uses."""
    rd = list(csv.reader(io.StringIO(raw_page(src))))
    hdr, units = rd[0], rd[1]
    col = {h: i for i, h in enumerate(hdr)}
    with open(dst, 'w') as f:
        f.write(f'# ncu --set full --clock-control none ; source {src}\n')
        f.write(f'{"kernel":64s}' + ''.join(f'{n:>10s}' for n, _ in COLS) + '\n')
        for row in rd[2:]:
            kn = re.sub(r'^void ', '', row[col['Kernel Name']])
            kn = re.sub(r'\(.*', '', kn).replace('nr::', '').replace('<unnamed>::', '')[:63]
            out = f'{kn:64s}'
            for n, m in COLS:
                if m not in col:
                    out += f'{"-":>10s}'
                    continue
                v = float(row[col[m]].replace(',', '') or 0) * UNIT.get(units[col[m]], 1)
                out += f'{v:10.1f}' if n not in ('regs', 'grid', 'block') else f'{int(v):10d}'
            f.write(out + '\n')


# kernel function (demangled, as ncu prints it) -> the name the library's event profiler reports
NAMES = (('fused_enc_fwd_kernel', 'enc.fused_fwd'), ('adam_dev_kernel', 'adam'), ('pack_weights_multi_kernel', 'enc.pack_w'),
         ('stage_rows_kernel', 'enc.stage'), ('pack_weights_kernel', 'enc.pack_w'), ('EpiTanhLogit', 'enc.fc1'),
         ('EpiPoolDx', 'enc.pool_dx'), ('EpiDx', 'enc.dx'), ('TnEpiPlain', 'enc.dw1'), ('TnEpiQkv', 'enc.dwqkv'),
         ('TnEpiConv', 'enc.dwconv'), ('attn_fwd', 'enc.attn_fwd'), ('attn_bwd', 'enc.attn_bwd'),
         ('pool_tail_fwd', 'enc.pool_fwd'), ('pool_tail_bwd', 'enc.pool_bwd'), ('dx_rowlist_kernel', 'enc.dx_rows'),
         ('zero_tail_rows_kernel', 'enc.dx_rows'))


def traffic(src, model, dst):
    out = raw_page(src)
    rd = list(csv.reader(io.StringIO(out)))
    hdr, units = rd[0], rd[1]
    col = {h: i for i, h in enumerate(hdr)}
    scale = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
    agg = defaultdict(lambda: {'launches': 0, 'dram_bytes_per_step': 0.0, 'ncu_time_us': 0.0})
    for row in rd[2:]:
        kn = row[col['Kernel Name']]
        name = next((n for pat, n in NAMES if pat in kn), None)
        if name is None and 'TcEpiStoreBf16' in kn:
            name = 'enc.qkv' if model == 'NRMS' else 'enc.conv'
        if name is None:
            name = re.sub(r'[<(].*', '', kn)
        b = 0.0
        for m in ('dram__bytes_read.sum', 'dram__bytes_write.sum'):
            b += float(row[col[m]].replace(',', '')) * scale.get(units[col[m]], 1)
        t = float(row[col['gpu__time_duration.sum']].replace(',', ''))
        t *= {'ns': 1e-3, 'us': 1, 'ms': 1e3, 'usecond': 1, 'nsecond': 1e-3, 'msecond': 1e3}.get(units[col['gpu__time_duration.sum']], 1)
        a = agg[name]
        a['launches'] += 1
        a['dram_bytes_per_step'] += b
        a['ncu_time_us'] += t
    old = json.load(open(dst)) if os.path.exists(dst) else {}
    old[model] = dict(sorted(agg.items(), key=lambda kv: -kv[1]['ncu_time_us']))
    old.setdefault('_how', 'ncu --set full --clock-control none --profile-from-start off python bench.py --model M '
                           '--profile-step : one eager train step (batch 128, bf16); dram__bytes_read.sum + '
                           'dram__bytes_write.sum summed over the launches of each kernel in that step')
    json.dump(old, open(dst, 'w'), indent=1)


if __name__ == '__main__':
    {'launches': launches, 'full': full, 'traffic': traffic}[sys.argv[1]](*sys.argv[2:])
