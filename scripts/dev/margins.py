"""Largest deviations from the float64 oracle per observable (how much of each tolerance is used)."""
import copy, sys, os
import numpy as np
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..', '..'))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'oracle')); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import radop_oracle as O
from conftest import plain, db
from test_gpu_parity import CONF_1MOM, make_operator, oracle_cfg
for scheme, freq in (('1mom', 5.6), ('2mom', 9.41)):
    op, model, luts = make_operator(CONF_1MOM, scheme, freq)
    az = np.arange(0., 360., 10.)
    els = [1.0, 8.5, 35.0]
    rays_az, rays_el = np.tile(az, 3), np.repeat(els, len(az))
    out = op.simulate_rays(rays_az, rays_el)
    pm, pl = plain(model, luts)
    ref = O.simulate_scan(pm, pl, oracle_cfg(op), rays_az, rays_el, accum='f64', sensitivity_cut=True)
    rep = {}
    for k in ('ZH', 'ZV', 'ZDR'):
        ok = np.isfinite(ref[k]); rep[k + ' dB (tol 0.01)'] = float(np.max(np.abs(db(out[k][ok].astype(float)) - db(ref[k][ok]))))
    ok = np.isfinite(ref['KDP']); e = np.abs(out['KDP'][ok] - ref['KDP'][ok]) / np.abs(ref['KDP'][ok])
    rep['KDP rel (tol 1e-3)'] = float(e.max()); rep['KDP rel q99'] = float(np.quantile(e, 0.99))
    ok = np.isfinite(ref['RHOHV']); rep['RHOHV abs (tol 1e-4)'] = float(np.max(np.abs(out['RHOHV'][ok] - ref['RHOHV'][ok])))
    for k in ('ATT_H', 'ATT_V'):
        ok = np.isfinite(ref[k]); rep[k + ' rel (tol 1e-3)'] = float(np.max(np.abs(out[k][ok] - ref[k][ok]) / np.abs(ref[k][ok])))
    ok = np.isfinite(ref['RVEL']); rep['RVEL abs (tol 2e-3)'] = float(np.max(np.abs(out['RVEL'][ok] - ref['RVEL'][ok])))
    print(scheme, {k: float('%.3g' % v) for k, v in rep.items()}, file=sys.stderr)
