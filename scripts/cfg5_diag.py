"""Diagnostics of cfg 5 (COSMO-1 sized domain, Doppler spectra): sample rays, device spectrum against the oracle's."""
import os, sys
import numpy as np
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..'))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'oracle')); sys.path.insert(0, os.path.join(ROOT, 'tests'))
import bench
import radop_oracle as O
from conftest import plain
from test_gpu_parity import oracle_cfg
from cosmo_pol_b200 import RadarOperator, synthetic

conf = bench.volume_config()
conf['doppler'] = {'scheme': 3}
conf['radar']['FFT_length'] = 128
conf['radar']['sensitivity'] = [-300, 10000]
model = synthetic.make_model(nz=80, ny=774, nx=1158, res=0.01, coarsen=3)
luts = synthetic.make_all_luts(5.6, '1mom', with_ice=True)
op = RadarOperator(conf, output_variables='only_radar', device=0)
op.set_lut(luts)
op.load_model(model)
raz, rel = np.array([25., 200., 311., 0., 143.]), np.array([1.0, 4.5, 20.0, -0.2, 7.5])
out = op.simulate_rays(raz, rel)
pm, pl = plain(model, luts)
ref = O.simulate_scan(pm, pl, oracle_cfg(op), raz, rel, accum='f64', sensitivity_cut=True)
gs, rs = out['DSPECTRUM'].astype(np.float64), ref['DSPECTRUM'].astype(np.float64)
gt, rt = np.nansum(gs, axis=2), np.nansum(rs, axis=2)
bad = (rt > 1e-25) & (gt == 0)
bad2 = (rt == 0) & (gt > 0)
print('gates with oracle power and no device power:', int(bad.sum()), ' reverse:', int(bad2.sum()))
for r, g in zip(*np.where(bad | bad2)):
    jr, jg = np.where(rs[r, g] > 0)[0], np.where(gs[r, g] > 0)[0]
    print('ray %d gate %d: oracle tot %g (bins %s, nan %d) device tot %g (bins %s, nan %d) ZH %g / %g RVEL %g / %g' % (
        r, g, rt[r, g], jr[:6], np.isnan(rs[r, g]).sum(), gt[r, g], jg[:6], np.isnan(gs[r, g]).sum(), ref['ZH'][r, g], out['ZH'][r, g],
        ref['RVEL'][r, g], out['RVEL'][r, g]))
    if len(jr):
        print('    oracle bins', rs[r, g][jr[:6]], ' device same bins', gs[r, g][jr[:6]])
ok = (rt > 1e-25) & (gt > 0)
rel_err = np.abs(gt[ok] - rt[ok]) / rt[ok]
print('total power relative error: max %g q99 %g over %d gates' % (rel_err.max(), np.quantile(rel_err, 0.99), ok.sum()))
