"""bench.py's one-JSON-line contract, checked on the host: the committed B200 lines under profiles/ carry every key
the driver reads, and the CPU reference arm (`--impl reference`, the oracle port of src/main.py:105-110) prints a
line of the same shape."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {'metric', 'value', 'unit', 'n_gpus', 'steps', 'warmup', 'ms_per_step', 'higher_is_better', 'scaling',
             'vs_baseline', 'dtype', 'data', 'config', 'e2e'}


def _line(path):
    lines = [l for l in open(path) if l.startswith('{')]
    assert len(lines) == 1, f'{path}: expected ONE JSON line'
    return json.loads(lines[0])


@pytest.mark.parametrize('name', ['r1_bench_bf16_nrms.json', 'r1_bench_bf16_naml.json'])
def test_committed_single_gpu_lines(name):
    d = _line(os.path.join(ROOT, 'profiles', name))
    assert BASE_KEYS <= d.keys()
    assert d['n_gpus'] == 1 and d['unit'] == 'samples/s' and d['higher_is_better'] is True and d['scaling'] == 'weak'
    assert d['warmup'] >= 3 and d['data'] == 'synthetic' and d['dtype'] == 'bf16' and d['vs_baseline'] is None
    assert 'workload' in d['config'] and 'l2_policy' in d['config']
    assert abs(d['value'] - d['config']['per_gpu_batch'] * 1e3 / d['ms_per_step']) < 1e-6 * d['value']
    e = d['e2e']
    assert e['h2d_bytes_per_step'] > 0 and e['d2h_bytes_per_step'] > 0 and 0 < e['value'] < d['value']
    assert d['gpu_launches'] > 0
    c = d['clocks']
    assert c['sm_mhz'] > 0.8 * c['sm_max_mhz']
    assert not set(c['reasons']) & {'hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown'}
    r = d['roofline']
    assert r['bound'] in ('hbm', 'tensor') and r['unit'] in ('GB/s', 'TFLOP/s')
    assert abs(r['frac'] - r['achieved'] / r['peak']) < 1e-9 and 0 < r['frac'] < 1
    assert r['traffic'] is None or r['traffic'] > 0
    b = d['cpu_baseline']
    assert b['kind'] in ('reference', 'port') and b['cores'] >= 1 and b['value'] > 0 and b['sample']


def _leg_ok(d):
    r = d['roofline']
    assert r['bound'] in ('hbm', 'tensor') and abs(r['frac'] - r['achieved'] / r['peak']) < 1e-9 and 0 < r['frac'] < 1.2
    e = d['e2e']
    assert e['h2d_bytes_per_step'] > 0 and e['d2h_bytes_per_step'] > 0 and 0 < e['value'] < d['value']
    assert d['gpu_launches'] > 0
    b = d['cpu_baseline']
    assert b['kind'] == 'reference' and b['cores'] >= 1 and b['value'] > 0 and 'oracle/_ref' in b['sample']
    g = d['gpu_eager']
    assert g['value'] > 0 and g['value'] < d['value'] and 'TF32 off' in g['what']
    ev = d['eval']
    assert ev['reps'] >= 10 and ev['impressions_per_s'] > 0 and ev['news_per_s'] > 0
    assert ev['roofline']['bound'] == 'hbm' and 0 < ev['roofline']['frac'] < 1
    cb = ev['cpu_baseline']
    assert cb['kind'] == 'reference' and cb['unit'] == 'impressions/s' and cb['value'] > 0 and 'sklearn' in cb['sample']
    # the device metrics and the reference loop score the same synthetic population: means agree to ~1e-2
    assert all(abs(a - b) < 3e-2 for a, b in zip(ev['metrics_mean_scored'], cb['metrics_mean_scored']))


def test_round2_default_line_carries_both_families_eval_and_baselines():
    """the driver-parsed line holds the NRMS and NAML legs, each with e2e, roofline, the reference's own
    modules as CPU baseline and eager-GPU comparator, and an eval leg timed over >= 10 repetitions with its own
    roofline and CPU baseline; roofline.traffic comes from the ncu capture of the same round."""
    d = _line(os.path.join(ROOT, 'profiles', 'r2', 'r2_bench_1gpu.json'))
    assert BASE_KEYS <= d.keys() and d['n_gpus'] == 1 and d['dtype'] == 'bf16' and d['config']['legs'] == ['NRMS', 'NAML']
    assert abs(d['value'] - d['config']['per_gpu_batch'] * 1e3 / d['ms_per_step']) < 1e-6 * d['value']
    _leg_ok(d)
    _leg_ok(d['naml'])
    assert d['roofline']['traffic'] and 0.8 < d['roofline']['traffic'] / d['roofline']['algorithmic_per_launch'] * (
        d['roofline']['intensity_flop_per_byte'] if d['roofline']['bound'] == 'tensor' else 1.0) < 1.3
    c = d['clocks']
    assert c['sm_mhz'] > 0.8 * c['sm_max_mhz'] and not set(c['reasons']) & {'hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown'}
    r = _line(os.path.join(ROOT, 'profiles', 'r2', 'r2_bench_reference_arm.json'))
    assert r['impl'] == 'reference' and r['cpu_baseline']['kind'] == 'reference' and r['native_so_mapped'] == []


def test_round2_scaling_lines():
    one = _line(os.path.join(ROOT, 'profiles', 'r2', 'r2f_1gpu_nrms.json'))['value']
    for name, world in (('r2f_2gpu_nrms.json', 2), ('r2f_4gpu_nrms.json', 4), ('r2f_8gpu_default.json', 8)):
        d = _line(os.path.join(ROOT, 'profiles', 'r2', name))
        assert d['n_gpus'] == world and d['params_identical_across_ranks'] is True
        assert abs(d['value'] - d['config']['global_batch'] * 1e3 / d['ms_per_step']) < 1e-6 * d['value']
        assert d['value'] / (world * one) > 0.97
        assert 'cpu_baseline' not in d


@pytest.mark.parametrize('name,world', [('r1_bench_bf16_nrms_2gpu.json', 2), ('r1_bench_bf16_nrms_8gpu.json', 8),
                                        ('r1_bench_bf16_naml_8gpu.json', 8)])
def test_committed_multi_gpu_lines(name, world):
    d = _line(os.path.join(ROOT, 'profiles', name))
    assert BASE_KEYS <= d.keys() and d['n_gpus'] == world
    assert d['config']['global_batch'] == world * d['config']['per_gpu_batch']
    assert abs(d['value'] - d['config']['global_batch'] * 1e3 / d['ms_per_step']) < 1e-6 * d['value']
    assert 'cpu_baseline' not in d                 # rank 0 at N = 1 only


def test_reference_arm_prints_the_contract_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--impl', 'reference', '--steps', '1',
                          '--warmup', '1', '--cpu-batch', '4'], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith('{')]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d['impl'] == 'reference' and d['unit'] == 'samples/s' and d['value'] > 0
    assert d['e2e'] == {'value': d['value'], 'unit': d['unit'], 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}
    assert d['cpu_baseline']['value'] == d['value'] and d['cpu_baseline']['kind'] in ('reference', 'port')
    assert d['native_so_mapped'] == []          # the CPU arm maps no libnr_sm100.so
