import copy, sys, os
import numpy as np
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..', '..'))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'oracle')); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import radop_oracle as O
from conftest import synthetic_inputs, plain
from test_gpu_parity import CONF_1MOM, make_operator, oracle_cfg
e = sys.stderr
conf = copy.deepcopy(CONF_1MOM); conf['doppler'] = {'scheme': 3}
conf['integration'] = {'scheme': 1, 'nh_GH': 1, 'nv_GH': 1}
conf['radar'].update({'FFT_length': 64, 'sensitivity': [-300, 10000]}); conf['microphysics'] = {'with_attenuation': 0}
op, model, luts = make_operator(conf, '1mom', 5.6)
rays_az, rays_el = np.array([5., 95.]), np.array([8.5, 8.5])
out = op.simulate_rays(rays_az, rays_el)
pm, pl = plain(model, luts)
cfg = oracle_cfg(op)
for r in range(2):
    ref = O.simulate_ray(pm, pl, cfg, rays_az[r], rays_el[r], accum='f64', debug=True)
    sub = ref['subs'][0]
    a, b = np.nan_to_num(out['DSPECTRUM'][r]).astype(np.float64), ref['values']['DSPECTRUM']
    for g in range(a.shape[0]):
        nz = np.nonzero(b[g])[0]
        if len(nz) == 0: continue
        err = np.abs(a[g][nz] - b[g][nz]) / b[g][nz]
        pres = [h for h in 'RSGI' if sub['values']['Q' + h + '_v'][g] > 0]
        if err.max() > 1e-4:
            print('ray', r, 'gate', g, pres, 'bins', nz, 'relerr', np.round(err, 5), 'ref', b[g][nz], file=e)
