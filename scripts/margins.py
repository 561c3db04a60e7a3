"""Largest deviations of the device path from the float64 oracle and from every reference golden, per observable,
beside the tolerance of the parity tests (how much of each tolerance is used).  Runs on a GPU box:

    python scripts/margins.py > gpurun_out/margins.json        (copied to profiles/r02_margins.json)
"""
import copy, json, os, sys
import numpy as np
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..'))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'oracle')); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import radop_oracle as O
from conftest import plain, db, load_golden_case, synthetic_inputs
from test_gpu_parity import CONF_1MOM, make_operator, oracle_cfg

TOL = {'ZH': ('dB', 0.01), 'ZV': ('dB', 0.01), 'ZDR': ('dB', 0.01), 'KDP': ('rel', 1e-3), 'RHOHV': ('abs', 1e-4),
       'ATT_H': ('rel', 1e-3), 'ATT_V': ('rel', 1e-3), 'PHIDP': ('rel', 1e-3), 'RVEL': ('abs', 2e-3)}


def margins(out, ref, nyq=0.0):
    rep = {}
    for k, (kind, tol) in TOL.items():
        if k not in ref or k not in out:
            continue
        a, b = np.asarray(out[k], dtype=np.float64), np.asarray(ref[k], dtype=np.float64)
        ok = np.isfinite(b) & np.isfinite(a)
        if not ok.any():
            continue
        with np.errstate(all='ignore'):
            if kind == 'dB':
                ok &= (b > 0) & (a > 0)
                e = np.abs(db(a[ok]) - db(b[ok]))
            elif kind == 'rel':
                ok &= np.abs(b) > 1e-30
                e = np.abs(a[ok] - b[ok]) / np.abs(b[ok])
            else:
                d = a[ok] - b[ok]
                if k == 'RVEL' and nyq > 0:
                    d = (d + nyq) % (2 * nyq) - nyq
                e = np.abs(d)
        if e.size == 0:
            continue
        rep[k] = {'kind': kind, 'tol': tol, 'max': float('%.3g' % e.max()), 'q99': float('%.3g' % np.quantile(e, 0.99)),
                  'n': int(e.size), 'nan_pattern_equal': bool(np.array_equal(np.isnan(a), np.isnan(b)))}
    return rep


_stdout, sys.stdout = sys.stdout, sys.stderr          # the host layer prints like the reference does: keep stdout for the JSON
report = {'oracle_f64': {}, 'goldens': {}}
for scheme, freq in (('1mom', 5.6), ('2mom', 9.41)):
    for dop in (1, 2):
        conf = copy.deepcopy(CONF_1MOM)
        conf['doppler'] = {'scheme': dop}
        op, model, luts = make_operator(conf, scheme, freq)
        az = np.arange(0., 360., 10.)
        els = [1.0, 8.5, 35.0]
        rays_az, rays_el = np.tile(az, 3), np.repeat(els, len(az))
        out = op.simulate_rays(rays_az, rays_el)
        pm, pl = plain(model, luts)
        ref = O.simulate_scan(pm, pl, oracle_cfg(op), rays_az, rays_el, accum='f64', sensitivity_cut=True)
        report['oracle_f64']['{}_doppler{}'.format(scheme, dop)] = {'fused': bool(op._engine.used_fused()) if hasattr(op._engine, 'used_fused') else None,
                                                                    'margins': margins(out, ref)}
        op.close()

from cosmo_pol_b200 import RadarOperator
for name, scheme, freq in (('c1mom', '1mom', 5.6), ('c2mom', '2mom', 9.41), ('c1mom_thr', '1mom', 5.6),
                           ('c1mom_nyq', '1mom', 5.6), ('c1mom_dop2', '1mom', 5.6), ('c1mom_quad2', '1mom', 5.6),
                           ('c1mom_quad3', '1mom', 5.6), ('c1mom_quad5', '1mom', 5.6),
                           ('c1mom_melt', '1mom', 5.6), ('c1mom_melt_dop2', '1mom', 5.6), ('c1mom_ml', '1mom', 5.6),
                           ('c1mom_dop3', '1mom', 5.6), ('c1mom_dop3_turb', '1mom', 5.6), ('c1mom_dop3_mot', '1mom', 5.6)):
    try:
        g, cfg, kw = load_golden_case(name)
        model, luts = synthetic_inputs(scheme, freq, with_melting=bool(cfg['microphysics'].get('with_melting', 0)), **kw)
        conf = copy.deepcopy(cfg)
        conf['radar']['sensitivity'] = [-200, 10000]
        op = RadarOperator(conf, output_variables='only_radar', device=0)
        op.set_lut(luts)
        op.load_model(model)
        rays = g['rays']
        out = op.simulate_rays(rays[:, 0], rays[:, 1])
        ref = {k[4:]: g[k] for k in g.files if k.startswith('obs_')}
        nyq = float(g['nyquist']) if 'nyquist' in g.files else 0.0
        report['goldens'][name] = margins(out, ref, nyq)
        op.close()
    except Exception as e:                     # a golden whose configuration the plain call does not cover
        report['goldens'][name] = {'skipped': '{}: {}'.format(type(e).__name__, e)[:200]}
sys.stdout = _stdout
print(json.dumps(report, indent=1))
