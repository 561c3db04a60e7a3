import copy, sys, os
import numpy as np
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..', '..'))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'oracle')); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import radop_oracle as O
from conftest import synthetic_inputs, plain
from test_gpu_parity import CONF_1MOM, make_operator, oracle_cfg
e = sys.stderr
conf = copy.deepcopy(CONF_1MOM)
conf['doppler'] = {'scheme': 3, 'motion_correction': 0, 'turbulence_correction': 1}
conf['integration'] = {'scheme': 1, 'nh_GH': 1, 'nv_GH': 1}
conf['radar'].update({'FFT_length': 64, 'sensitivity': [-300, 10000], 'antenna_speed': 0.35})
op, model, luts = make_operator(conf, '1mom', 5.6, with_edr=True, top=15000.0)
rays_az, rays_el = np.array([5.]), np.array([1.0])
out = op.simulate_rays(rays_az, rays_el)
pm, pl = plain(model, luts)
cfg = oracle_cfg(op)
consts = O.derived_constants(cfg)
ref = O.simulate_ray(pm, pl, cfg, rays_az[0], rays_el[0], accum='f64', debug=True)
s = ref['subs'][0]
std = O.spectral_width_turb(consts['RANGE_RADAR'], s['values']['EDR'], cfg)
print('EDR', s['values']['EDR'][:6], 'std', std[:6], 'vres', consts['VRES'], file=e)
a, b = np.nan_to_num(out['DSPECTRUM'][0]).astype(np.float64), np.nan_to_num(ref['values']['DSPECTRUM'])
for g in range(0, 6):
    na, nb = np.nonzero(a[g] > 1e-6 * a[g].max())[0], np.nonzero(b[g] > 1e-6 * b[g].max())[0]
    print('gate', g, 'gpu bins', na.min() if len(na) else None, na.max() if len(na) else None, 'ref bins', nb.min() if len(nb) else None, nb.max() if len(nb) else None,
          'gpu 2nd moment', np.sqrt(np.sum(a[g] * (np.arange(65) - np.sum(a[g] * np.arange(65)) / a[g].sum()) ** 2) / a[g].sum()) if a[g].sum() else None,
          'ref', np.sqrt(np.sum(b[g] * (np.arange(65) - np.sum(b[g] * np.arange(65)) / b[g].sum()) ** 2) / b[g].sum()) if b[g].sum() else None, file=e)
op2, _, _ = make_operator(conf, '1mom', 5.6, output_variables='all', with_edr=True, top=15000.0)
o2 = op2.simulate_rays(rays_az, rays_el)
print(sorted(o2.keys()), file=e)
print('gpu EDR', o2.get('EDR', np.zeros((1, 6)))[0][:6], 'T', o2['T'][0][:3], file=e)
print('params', op2._engine.params.spectrum_turbulence, op2._engine.params.spectrum_motion, op2._engine.params.beamwidth_deg, op2._engine.uploaded_vars, file=e)
