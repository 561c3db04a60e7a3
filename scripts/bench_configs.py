#!/usr/bin/env python
# -*- coding: utf-8 -*-
"""Timings of the other BASELINE.json configurations (1, 2, 4, 5) on one GPU.  They are parity-test
cases, not bench lines (bench.py measures configuration 3); this script records what they cost,
for profiles/.  One JSON object per configuration on stdout.

    python scripts/bench_configs.py [cfg1 cfg2 cfg4 cfg5] [--steps K]
"""
from __future__ import print_function

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..'))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def base_conf(freq=5.6, nh=3, nv=3, doppler=1, fft=128):
    c = bench.volume_config()
    c['radar']['frequency'] = freq
    c['radar']['FFT_length'] = fft
    c['integration'].update({'nh_GH': nh, 'nv_GH': nv})
    c['doppler'] = {'scheme': doppler}
    return c


# ncu --set full of k_spectrum at cfg5s (profiles/r02l_k_spectrum_full.txt): 1.707e10 warp instructions for 2.661e7 pairs
SPECTRUM_NCU = {'warp_instructions_per_pair': 1.707e10 / 2.661e7, 'issue_active_pct': 73.2, 'sm_mhz': 1962.0, 'dram_bytes': 2.09e9,
                'source': 'profiles/r02l_k_spectrum_full.txt'}

CONFIGS = {
    'cfg1': dict(conf=base_conf(), scheme='1mom', rays='ppi', what='PPI 1.0 deg, C-band, 1-mom, 360 x 500 x 3x3'),
    'cfg2': dict(conf=base_conf(freq=9.41), scheme='2mom', rays='ppi', what='PPI 1.0 deg, X-band, 2-mom R/S/G/H/I, 360 x 500 x 3x3'),
    'cfg4': dict(conf=base_conf(nh=29, nv=29), scheme='1mom', rays='rhi', what='RHI 0..90 deg, 29x29 Gauss-Hermite, attenuation on, 91 x 500 x 841'),
    'cfg5': dict(conf=base_conf(doppler=3), scheme='1mom', rays='volume', model=dict(nz=80, ny=774, nx=1158, res=0.01, coarsen=3),
                 what='20-elevation volume, Doppler spectrum (scheme 3, FFT 128), COSMO-1 sized grid 80x774x1158'),
    'cfg5s': dict(conf=base_conf(doppler=3), scheme='1mom', rays='volume',
                  what='20-elevation volume, Doppler spectrum (scheme 3, FFT 128), standard 60x400x400 grid'),
}


def rays_of(kind):
    if kind == 'ppi':
        return np.arange(360.), np.full(360, 1.0)
    if kind == 'rhi':
        return np.full(91, 0.0), np.arange(91.)
    return bench.volume_rays()


def run(name, steps):
    import torch
    from cosmo_pol_b200 import RadarOperator, synthetic
    spec = CONFIGS[name]
    quiet, stdout = open(os.devnull, 'w'), sys.stdout
    sys.stdout = quiet
    t0 = time.time()
    model = synthetic.make_model(scheme=spec['scheme'], **spec.get('model', {}))
    luts = synthetic.make_all_luts(spec['conf']['radar']['frequency'], spec['scheme'], with_ice=True)
    t_gen = time.time() - t0
    op = RadarOperator(spec['conf'], output_variables='only_radar', device=0)
    op.set_lut(luts)
    op.load_model(model)
    t0 = time.time()
    eng = op._prepare_engine()
    t_up = time.time() - t0
    az, el = rays_of(spec['rays'])
    scan = op.build_scan(az, el)
    eng.stage(scan)
    for _ in range(2):
        eng.execute()
    eng.synchronize()
    eng.enable_timing(True)
    stream = torch.cuda.ExternalStream(eng.stream_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        eng.execute()
    e1.record(stream)
    eng.synchronize()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    st = eng.stats()
    eng.enable_timing(False)
    arrays = eng.simulate(scan)
    t0 = time.perf_counter()
    for _ in range(steps):
        arrays = eng.simulate(scan)
    te = (time.perf_counter() - t0) / steps
    sys.stdout = stdout
    sbg = scan.n_subbeam_gates
    line = {'config': name, 'what': spec['what'], 'sub_beam_gates': sbg, 'n_sub': scan.n_sub, 'rays': scan.n_rays,
            'ms_per_scan_device': ms, 'sub_beam_gates_per_s': sbg / (ms * 1e-3),
            'ms_per_scan_e2e': te * 1e3, 'e2e_sub_beam_gates_per_s': sbg / te,
            'd2h_bytes': int(sum(a.nbytes for a in arrays.values())),
            'stages_ms': {k: st[k] for k in ('ms_interp', 'ms_sort', 'ms_integrate', 'ms_reduce', 'ms_spectrum', 'ms_total')},
            'active_pairs': int(st['n_active_pairs']), 'launches': int(st['launches']),
            'algorithmic_tflops_integrate': 28672.0 * st['n_active_pairs'] / (st['ms_integrate'] * 1e-3) / 1e12 if st['ms_integrate'] else None,
            'host_s': {'synthetic_inputs': t_gen, 'one_time_upload': t_up}, 'steps': steps,
            'checksum_ZH': float(np.nansum(arrays['ZH'][::7]))}
    if st['ms_spectrum'] > 0.01 and spec['conf']['doppler']['scheme'] == 3:
        # k_spectrum (K4b) dominates scheme 3.  Its bound is instruction issue + the shared-memory pipe, not HBM or FP32:
        # figures of the ncu --set full capture named below (same command, cfg5s) beside the live stage time.
        pairs = float(st['n_active_pairs'])
        ncu = SPECTRUM_NCU
        issue_floor_ms = ncu['warp_instructions_per_pair'] * pairs / (148 * 4 * ncu['sm_mhz'] * 1e3)
        line['roofline_spectrum'] = {
            'kernel': 'k_spectrum (+ k_beam_attenuation, k_spectrum_finish in the stage time)', 'bound': 'issue',
            'stage_ms': st['ms_spectrum'], 'pairs': pairs, 'ns_per_pair': st['ms_spectrum'] * 1e6 / pairs,
            'warp_instructions_per_pair': ncu['warp_instructions_per_pair'], 'issue_active_pct': ncu['issue_active_pct'],
            'issue_floor_ms': issue_floor_ms, 'frac_of_issue_floor': issue_floor_ms / st['ms_spectrum'],
            'shared_bytes_per_pair': 8192 + 4096 * 0.35 * 2, 'dram_bytes_per_launch': ncu['dram_bytes'],
            'algorithmic': 'per pair: 1024 samples N(D) sigma(D) dD prefix-summed (1 FFMA + 1 store each), <= 129 velocity-bin '
                           'edges mapped to diameters (float64 power law), <= 128 float64 atomics into [ray][gate][n_vel]',
            'source': ncu['source']}
    print(json.dumps(line))
    sys.stdout.flush()
    op.close()


if __name__ == '__main__':
    ap = argparse.ArgumentParser()
    ap.add_argument('names', nargs='*', default=['cfg1', 'cfg2', 'cfg4', 'cfg5'])
    ap.add_argument('--steps', type=int, default=5)
    a = ap.parse_args()
    for n in a.names:
        run(n, a.steps)
