import copy, sys, os
import numpy as np
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..', '..'))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'oracle')); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import radop_oracle as O
from conftest import synthetic_inputs, plain
from test_gpu_parity import CONF_1MOM, make_operator, oracle_cfg
for dop in (1, 2):
    conf = copy.deepcopy(CONF_1MOM); conf['doppler'] = {'scheme': dop}
    op, model, luts = make_operator(conf, '2mom', 9.41)
    az = np.arange(5., 360., 45.); els = [0.4, 6.5, 30.0, 89.0]
    rays_az, rays_el = np.tile(az, len(els)), np.repeat(els, len(az))
    out = op.simulate_rays(rays_az, rays_el, keep_debug=True)
    pm, pl = plain(model, luts)
    ref = O.simulate_scan(pm, pl, oracle_cfg(op), rays_az, rays_el, accum='f64', sensitivity_cut=True, debug=True)
    a, b = out['ZH'], ref['ZH']
    bad = np.argwhere(np.isnan(a) != np.isnan(b))
    print('doppler', dop, 'nan mismatches', len(bad), file=sys.stderr)
    for r, g in bad[:6]:
        print('  ray', r, 'az', rays_az[r], 'el', rays_el[r], 'gate', g, 'gpu', a[r, g], 'ref', b[r, g],
              'KDP', out['KDP'][r, g], ref['KDP'][r, g], 'ATT', out['ATT_H'][r, g], ref['ATT_H'][r, g], file=sys.stderr)
        print('   ref sz_total', ref['sz_total'][r, g], file=sys.stderr)
