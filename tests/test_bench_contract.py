"""The bench.py JSON line contract, checked on the lines recorded from the GPU box (profiles/): the keys the driver reads
must be there with sane values, at N = 1 and for the sharded runs.  bench.py itself needs a GPU; this keeps the recorded
evidence and the contract from drifting apart."""
import glob
import json
import os

import pytest

from conftest import ROOT

REQUIRED = ('metric', 'value', 'unit', 'n_gpus', 'steps', 'warmup', 'ms_per_step', 'higher_is_better', 'scaling',
            'vs_baseline', 'dtype', 'data', 'config', 'e2e', 'gpu_launches', 'clocks', 'roofline', 'cpu_baseline')


def _line(path):
    return json.loads(open(path).read().strip().splitlines()[-1])


@pytest.mark.parametrize('name', ['r02j_bench.json', 'r02g_bench_1gpu.json', 'r02g_bench_2gpu.json', 'r02g_bench_4gpu.json',
                                  'r02g_bench_8gpu.json'])
def test_recorded_bench_lines_keep_the_contract(name):
    d = _line(os.path.join(ROOT, 'profiles', name))
    for k in REQUIRED:
        assert k in d, k
    assert d['unit'] == 'sub-beam gates/s' and d['higher_is_better'] is True and d['data'] == 'synthetic'
    assert d['scaling'] == 'strong' and d['vs_baseline'] is None and 'workload' in d['config']
    n = d['n_gpus']
    assert d['config']['rays_total'] == 7200 and d['config']['rays_per_gpu'] == 7200 // n
    # value = all sub-beam gates of the ONE volume / the step time of the slowest rank
    assert abs(d['value'] - 7200 * 500 * 9 / (d['ms_per_step'] * 1e-3)) < 1e-6 * d['value']
    e = d['e2e']
    for k in ('value', 'unit', 'h2d_bytes_per_step', 'd2h_bytes_per_step'):
        assert k in e, k
    assert e['h2d_bytes_per_step'] > 0 and e['d2h_bytes_per_step'] >= 7200 * 500 * 11 * 4 and e['value'] < d['value']
    assert d['gpu_launches'] >= d['steps']
    c = d['clocks']
    assert c['sm_mhz'] > 0.9 * c['sm_max_mhz'] and not set(c['reasons']) & {'hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown'}
    r = d['roofline']
    for k in ('bound', 'achieved', 'peak', 'unit', 'frac', 'traffic', 'kernel'):
        assert k in r, k
    assert r['kernel'] == 'k_gate_fused' and abs(r['frac'] - r['achieved'] / r['peak']) < 1e-9
    cb = d['cpu_baseline']
    for k in ('value', 'unit', 'cores', 'kind', 'sample'):
        assert k in cb, k
    if n == 1:
        assert cb['kind'] == 'port' and cb['cores'] >= 1 and cb['value'] > 0


def test_recorded_scaling_is_what_the_docs_say():
    v = {n: _line(os.path.join(ROOT, 'profiles', 'r02g_bench_{}gpu.json'.format(n)))['value'] for n in (1, 2, 4, 8)}
    eff = {n: v[n] / (n * v[1]) for n in (2, 4, 8)}
    assert eff[2] > 0.95 and eff[4] > 0.9 and eff[8] > 0.85, eff


def test_reference_arm_line():
    d = _line(os.path.join(ROOT, 'profiles', 'r02h_bench_reference_arm.json'))
    assert d['impl'] == 'reference' and d['unit'] == 'sub-beam gates/s' and d['value'] > 0
    assert d['cpu_baseline']['kind'] == 'port' and d['e2e']['h2d_bytes_per_step'] == 0 and d['gpu_launches'] == 0
