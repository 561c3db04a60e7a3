import copy, sys, os
import numpy as np
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..', '..'))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'oracle')); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import radop_oracle as O
from conftest import synthetic_inputs, plain
from test_gpu_parity import CONF_1MOM, make_operator, oracle_cfg
e = sys.stderr
turb = int(sys.argv[1]) if len(sys.argv) > 1 else 0
conf = copy.deepcopy(CONF_1MOM)
conf['doppler'] = {'scheme': 3, 'motion_correction': 1, 'turbulence_correction': turb}
conf['radar'].update({'FFT_length': 64, 'sensitivity': [-300, 10000], 'antenna_speed': 0.35})
op, model, luts = make_operator(conf, '1mom', 5.6, with_edr=bool(turb), top=15000.0)
az = np.array([5., 95., 185., 275.]); els = [1.0, 8.5, 40.0]
rays_az, rays_el = np.tile(az, len(els)), np.repeat(els, len(az))
out = op.simulate_rays(rays_az, rays_el)
pm, pl = plain(model, luts)
cfg = oracle_cfg(op)
consts = O.derived_constants(cfg)
for r in range(len(rays_az)):
    ref = O.simulate_ray(pm, pl, cfg, rays_az[r], rays_el[r], accum='f64', debug=True)
    araw = out['DSPECTRUM'][r].astype(np.float64)
    a, b = np.nan_to_num(araw), ref['values']['DSPECTRUM']
    bn = np.isnan(b).all(axis=1)
    b = np.nan_to_num(b)
    ta, tb = a.sum(1), b.sum(1)
    bad = np.nonzero(np.abs(ta - tb) > 1e-3 * np.maximum(tb, 1e-300))[0]
    bad = [g for g in bad if not bn[g]]
    if not bad: continue
    print('ray', r, rays_az[r], rays_el[r], 'bad gates', bad, file=e)
    for g in bad[:4]:
        print('  gate', g, 'gpu tot', ta[g], 'gpu allnan', np.isnan(araw[g]).all(), 'ref tot', tb[g], file=e)
        for q, s in enumerate(ref['subs']):
            hyd = None
            print('    sub', q, 'mask', s['mask'][g], 'Q', [float(s['values']['Q%s_v' % h][g]) for h in 'RSGI'], 'w', s['quad_weight'], file=e)
        hydrom = O.species_list(cfg)
        hyd = {}
        for h in hydrom:
            hyd[h] = O.Hydro(h, cfg['microphysics']['scheme'])
            d_axis = pl[h]['axes'][2]
            hyd[h].nbins_D = pl[h]['value_table'].shape[-2]; hyd[h].d_min = d_axis[0]; hyd[h].d_max = d_axis[-1]
        for q, s in enumerate(ref['subs']):
            spec = O.get_doppler_spectrum(s, hyd, pl, list(hydrom), cfg, consts)
            for g in bad[:4]:
                row = spec[g]
                print('    sub', q, 'gate', g, 'row sum', row.sum(), 'nonzero bins', np.nonzero(row)[0][:6], 'std',
                      O.spectral_width_motion(s['elev'], cfg, consts)[g], 'elev', s['elev'][g], file=e)
