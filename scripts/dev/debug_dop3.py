import copy, sys, os
import numpy as np
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..', '..'))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'oracle')); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import radop_oracle as O
from conftest import synthetic_inputs, plain
from test_gpu_parity import CONF_1MOM, make_operator, oracle_cfg
e = sys.stderr
for scheme, freq in (('1mom', 5.6), ('2mom', 9.41)):
    conf = copy.deepcopy(CONF_1MOM); conf['doppler'] = {'scheme': 3}
    conf['radar'].update({'FFT_length': 64, 'sensitivity': [-300, 10000]}); conf['microphysics'] = {'with_attenuation': int(os.environ.get('ATT', '1'))}
    op, model, luts = make_operator(conf, scheme, freq)
    az = np.array([5., 95., 185., 275.]); els = [1.0, 8.5, 35.0]
    rays_az, rays_el = np.tile(az, len(els)), np.repeat(els, len(az))
    out = op.simulate_rays(rays_az, rays_el)
    pm, pl = plain(model, luts)
    ref = O.simulate_scan(pm, pl, oracle_cfg(op), rays_az, rays_el, accum='f64', sensitivity_cut=True)
    a, b = np.nan_to_num(out['DSPECTRUM']).astype(np.float64), ref['DSPECTRUM']
    ta, tb = a.sum(-1), b.sum(-1)
    bad = np.argwhere(((tb > 1e-25) & ~(ta > 0)) | ((tb == 0) & (ta != 0)))
    print(scheme, 'bad gates', len(bad), file=e)
    for r, g in bad[:8]:
        print('  ray', r, rays_az[r], rays_el[r], 'gate', g, 'gpu tot', ta[r, g], 'ref tot', tb[r, g], 'ZH', out['ZH'][r, g], ref['ZH'][r, g], file=e)
        print('    ref nz', np.nonzero(b[r, g])[0][:10], b[r, g][np.nonzero(b[r, g])[0][:5]], file=e)
        print('    gpu nz', np.nonzero(a[r, g])[0][:10], a[r, g][np.nonzero(a[r, g])[0][:5]], file=e)
    ok = tb > 1e-25
    ca, cb = np.cumsum(a, -1)[ok], np.cumsum(b, -1)[ok]
    err = np.max(np.abs(ca - cb), -1) / tb[ok]
    print('  cum err q50 q99 max', np.quantile(err, .5), np.quantile(err, .99), err.max(), 'total err', np.max(np.abs(ca[:, -1] - cb[:, -1]) / tb[ok]), file=e)
    ex = np.abs(a[ok] - b[ok]) <= 1e-4 * np.max(b[ok], axis=-1, keepdims=True)
    print('  exact frac', ex.mean(), file=e)
    d = np.abs(out['RVEL'] - ref['RVEL']); print('  rvel', np.nanquantile(d, .99), np.nanmax(d), file=e)
    if scheme == '1mom':
        tot_err = np.abs(ta - tb) / np.where(tb > 0, tb, 1)
        tot_err[~(tb > 1e-25)] = 0
        for r, g in np.argwhere(tot_err > 1e-3)[:6]:
            print('  worst ray', r, rays_az[r], rays_el[r], 'gate', g, 'gpu tot', ta[r, g], 'ref tot', tb[r, g], 'ZH', ref['ZH'][r, g], 'ATT_H', ref['ATT_H'][r, g], file=e)
            nz = np.nonzero(b[r, g])[0]
            print('    bins', nz, file=e); print('    ref', b[r, g][nz], file=e); print('    gpu', a[r, g][nz], file=e)
