"""Diagnostics of the cfg-5 full-domain run: where do ZH (NaN pattern) and the returned Doppler spectrum disagree?"""
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
az, el = bench.volume_rays()
out = op.simulate_rays(az, el, copy=False)
sp, zh, rv, mask = out['DSPECTRUM'], out['ZH'], out['RVEL'], out['mask']
tot = np.nansum(sp, axis=2, dtype=np.float64)
nanbins = np.isnan(sp).sum(axis=2)
has = tot > 0
A = np.isnan(zh) & has
B = ~np.isnan(zh) & ~has
print('gates', zh.size, 'zh nan', int(np.isnan(zh).sum()), 'has', int(has.sum()), 'A (zh nan, power)', int(A.sum()), 'B (zh finite, no power)', int(B.sum()))
els = np.unique(el)
for e in els[:6].tolist() + els[-2:].tolist():
    r = el == e
    print('el %5.1f  A %7d  B %7d  zh-nan %7d has %7d' % (e, A[r].sum(), B[r].sum(), np.isnan(zh[r]).sum(), has[r].sum()))
for name, M in (('A', A), ('B', B)):
    if not M.any():
        continue
    ri, gi = np.where(M)
    print(name, 'gate index quantiles', np.quantile(gi, [0, .1, .5, .9, 1]), 'mask values', np.unique(mask[M])[:8])
    for k in range(0, len(ri), max(1, len(ri) // 6)):
        r, g = ri[k], gi[k]
        print('  ray %d (az %.0f el %.1f) gate %d: ZH %g tot %g nan-bins %d rvel %g mask %g  spectrum max %g' % (
            r, az[r], el[r], g, zh[r, g], tot[r, g], nanbins[r, g], rv[r, g], mask[r, g], np.nanmax(np.nan_to_num(sp[r, g], nan=0))))
# oracle on the first ray with each kind of mismatch
pm, pl = plain(model, luts)
for name, M in (('A', A), ('B', B)):
    if not M.any():
        continue
    r = int(np.where(M.any(axis=1))[0][0])
    ref = O.simulate_scan(pm, pl, oracle_cfg(op), az[r:r + 1], el[r:r + 1], accum='f64', sensitivity_cut=True)
    g = np.where(M[r])[0]
    rt = np.nansum(ref['DSPECTRUM'][0], axis=1)
    print(name, 'ray', r, 'gates', g[:5], '... oracle ZH', ref['ZH'][0][g[:5]], 'oracle tot', rt[g[:5]], 'device ZH', zh[r][g[:5]], 'device tot', tot[r][g[:5]])
    print('   oracle nan(ZH) == device nan(ZH):', np.array_equal(np.isnan(ref['ZH'][0]), np.isnan(zh[r])), ' oracle has == device has:', np.array_equal(rt > 0, has[r]))
