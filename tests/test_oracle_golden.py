"""The CPU oracle against golden vectors produced by RUNNING THE REFERENCE ITSELF in the build
container (tests/golden/make_golden_from_reference.py): the whole per-ray path, its units and
its constants."""
import json
import os

import numpy as np
import pytest

import radop_oracle as O
from conftest import GOLDEN, load_golden_case, plain, synthetic_inputs

CASES = {'c1mom': ('1mom', 5.6), 'c2mom': ('2mom', 9.41), 'c1mom_thr': ('1mom', 5.6), 'c1mom_nyq': ('1mom', 5.6),
         'c1mom_dop2': ('1mom', 5.6), 'c1mom_dop3': ('1mom', 5.6), 'c1mom_dop3_mot': ('1mom', 5.6),
         'c1mom_dop3_turb': ('1mom', 5.6), 'c1mom_melt': ('1mom', 5.6), 'c1mom_melt_dop2': ('1mom', 5.6),
         'c1mom_quad2': ('1mom', 5.6), 'c1mom_ml': ('1mom', 5.6), 'c1mom_quad3': ('1mom', 5.6),
         'c1mom_quad5': ('1mom', 5.6)}


def test_constants_match_reference():
    ref = json.load(open(os.path.join(GOLDEN, 'ref_constants.json')))
    names = {'a': 'AM', 'b': 'BM', 'alpha': 'AV', 'beta': 'BV', 'mu': 'MU', 'lambda_factor': 'LAMBDA_FACTOR',
             'vel_factor': 'VEL_FACTOR', 'ntot_factor': 'NTOT_FACTOR'}
    for k in 'RSGI':
        for n, a in names.items():
            assert O.C1[k][n] == ref['1mom'][a + '_' + k], (k, n)
    assert O.C1['R']['N0'] == ref['1mom']['N0_R'] and O.C1['G']['N0'] == ref['1mom']['N0_G']
    names2 = dict(names, nu='NU', x_min='X_MIN', x_max='X_MAX')
    for k in 'RSGHI':
        for n, a in names2.items():
            assert O.C2[k][n] == ref['2mom'][a + '_' + k], (k, n)
    assert O.KE == ref['global']['KE'] and O.T0 == ref['global']['T0'] and O.C_LIGHT == ref['global']['C']


def test_units_match_reference():
    g = np.load(os.path.join(GOLDEN, 'ref_units.npz'))
    cfg = O.default_config()
    cfg['radar'].update({'frequency': 5.6, 'range': 150000, 'radial_resolution': 300, 'FFT_length': 128})
    d = O.derived_constants(cfg)
    assert d['WAVELENGTH'] == float(g['derived_wavelength'])
    assert np.array_equal(d['VARRAY'], g['derived_varray'])
    assert np.array_equal(d['RANGE_RADAR'], g['derived_range'])
    for i, el in enumerate(g['traj_elev']):
        s, h, e = O.ref_4_3(d['RANGE_RADAR'], el, cfg['radar']['coords'])
        assert np.array_equal(s, g['traj_s'][i]) and np.array_equal(h, g['traj_h'][i]) and np.array_equal(e, g['traj_e'][i])
    assert np.array_equal(np.array([O.earth_radius(x) for x in g['earth_radius_in']]), g['earth_radius'])
    assert np.array_equal(O.aliasing(g['alias_v'], 8.3), g['alias_out'])
    pol = O.get_pol_from_sz(g['pol_sz'], d['WAVELENGTH'], 0.93)
    assert np.allclose(np.stack([np.asarray(p, dtype=np.float64) for p in pol]), g['pol_out'], rtol=1e-6, equal_nan=True)


@pytest.mark.parametrize('scheme,kinds', [('1mom', 'RSGI'), ('2mom', 'RSGHI')])
def test_psd_matches_reference(scheme, kinds):
    g = np.load(os.path.join(GOLDEN, 'ref_units.npz'))
    for k in kinds:
        tag = 'psd_{}_{}_'.format(scheme, k)
        hy = O.Hydro(k, scheme)
        D = g[tag + 'D']
        hy.d_min, hy.d_max = D[0], D[-1]
        if scheme == '1mom':
            if k in 'SI':
                hy.set_psd(g[tag + 'T'], g[tag + 'QM'])
            else:
                hy.set_psd(g[tag + 'QM'])
        else:
            hy.set_psd(g[tag + 'QN'], g[tag + 'QM'])
        # the reference mixes float32 inputs into its float64 arithmetic differently under
        # numpy >= 2; agreement is at float32 resolution of lambda (amplified through exp)
        assert np.allclose(hy.lambda_, g[tag + 'lambda'], rtol=2e-6)
        assert np.allclose(np.broadcast_to(hy.N0, g[tag + 'N0'].shape), g[tag + 'N0'], rtol=2e-5)
        assert np.allclose(hy.get_N(D), g[tag + 'N'], rtol=5e-4, atol=1e-300)
        v, n = hy.integrate_V()
        assert np.allclose(v, g[tag + 'v'], rtol=5e-5) and np.allclose(n, g[tag + 'n'], rtol=5e-5)


def test_lookup_line_matches_reference():
    from cosmo_pol_b200 import synthetic
    g = np.load(os.path.join(GOLDEN, 'ref_units.npz'))
    lut = synthetic.lut_as_plain(synthetic.make_lut('S', 5.6, '1mom', n_e=10, n_t=12, n_d=64))
    line, ie, it = O.lookup_line(lut, g['lut_e'], g['lut_t'])
    assert np.array_equal(line, g['lut_line'])
    assert ie.min() == 0 and ie.max() == 9 and it.min() == 0 and it.max() == 11      # clamping exercised


@pytest.mark.parametrize('name', list(CASES.keys()))
def test_path_matches_reference(name):
    scheme, freq = CASES[name]
    g, cfg, kw = load_golden_case(name)
    model, luts = synthetic_inputs(scheme, freq, with_melting=bool(cfg['microphysics'].get('with_melting', 0)), **kw)
    # the seeded generator must still produce the inputs the goldens were made from
    assert float(np.sum(model['T'].data.astype(np.float64))) == float(g['fp_T'])
    assert float(np.sum(model['QR_v'].data.astype(np.float64))) == float(g['fp_QR'])
    assert float(np.sum(luts['R'].value_table[::5, ::5, ::64])) == float(g['fp_lut_R'])
    pm, pl = plain(model, luts)
    for i, (az, el) in enumerate(g['rays']):
        r = O.simulate_ray(pm, pl, cfg, az, el, accum='ref', debug=True, output_variables='all')
        # interpolation half: bit-exact
        for q, s in enumerate(r['subs']):
            assert np.array_equal(s['mask'], g['sub_mask'][i][q])
            assert np.array_equal(np.asarray(s['quad_weight']), g['sub_w'][i][q])
            assert np.array_equal(np.asarray(s['quad_pt']), g['sub_pt'][i][q])
            assert np.array_equal(s['lats'], g['sub_lats'][i][q]) and np.array_equal(s['lons'], g['sub_lons'][i][q])
            for k, v in s['values'].items():
                assert np.array_equal(v, g['sub_' + k][i][q], equal_nan=True), (name, i, q, k)
        assert np.array_equal(r['mask'], g['mask'][i])
        # scattering half: float32-rounding level (the reference's own dtype promotion differs
        # between numpy generations); KDP carries the reference's float32 cancellation noise
        for k in [f for f in g.files if f.startswith('obs_')]:
            a = np.asarray(r['values'][k[4:]], dtype=np.float64)
            b = np.asarray(g[k][i], dtype=np.float64)
            assert np.array_equal(np.isnan(a), np.isnan(b)), (name, i, k)
            ok = np.isfinite(b)
            if ok.any():
                rtol = 2e-3 if k == 'obs_KDP' else 5e-5
                assert np.allclose(a[ok], b[ok], rtol=rtol, atol=1e-30), (name, i, k, np.max(np.abs(a[ok] - b[ok]) / (np.abs(b[ok]) + 1e-300)))
        for k in [f for f in g.files if f.startswith('int_')]:
            assert np.allclose(r['values'][k[4:]], g[k][i], rtol=1e-6, atol=1e-12, equal_nan=True), (name, i, k)
        if 'cut_ZH' in g.files:
            vals = {k[4:]: np.array(g[k][i], copy=True) for k in g.files if k.startswith('obs_')}
            O.cut_at_sensitivity(vals, cfg)
            for k in vals:
                assert np.array_equal(vals[k], g['cut_' + k][i], equal_nan=True), (name, k)


def test_gaussian_filter_restates_scipy():
    """The oracle's numpy restatement of scipy.ndimage.gaussian_filter (the one third-party call
    of the spectrum broadening, doppler_scatter.py:796) against scipy itself, including radii
    beyond the array (multiple reflections), sigma = 0 and NaN (axis skipped)."""
    ndi = pytest.importorskip('scipy.ndimage')
    rng = np.random.default_rng(3)
    for n in (65, 129):
        for sigma in (0.0, float('nan'), 1e-16, 0.12, 0.9, 3.3, 17.0, 40.0, 150.0):
            x = (rng.random(n) ** 8).astype(np.float32)
            a = O.gaussian_filter1d(x, sigma)
            b = ndi.gaussian_filter(x, sigma)
            assert a.dtype == b.dtype
            assert np.allclose(a, b, rtol=2e-6, atol=1e-12), (n, sigma, np.max(np.abs(a - b)))
