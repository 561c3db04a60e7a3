"""Host-side logic of the package (no GPU): configuration, lookup-table container and file
format, quadrature, refraction, hydrometeor classes, constants, the C-ABI exports."""
import ctypes
import json
import os
import sys
import re

import numpy as np
import pytest

from conftest import GOLDEN, ROOT


def test_constants_match_reference_values():
    from cosmo_pol_b200.constants import constants_1mom as c1, constants_2mom as c2
    ref = json.load(open(os.path.join(GOLDEN, 'ref_constants.json')))
    for k, v in ref['1mom'].items():
        if hasattr(c1, k):
            assert float(getattr(c1, k)) == v, k
    for k, v in ref['2mom'].items():
        if hasattr(c2, k):
            assert float(getattr(c2, k)) == v, k
    must = ['N0_G', 'AM_G', 'BM_G', 'AV_G', 'BV_G', 'LAMBDA_FACTOR_R', 'VEL_FACTOR_S', 'NTOT_FACTOR_I', 'D_MAX_S']
    assert all(hasattr(c1, k) for k in must)
    assert all(hasattr(c2, k + '_' + s) for k in ('AM', 'BM', 'AV', 'BV', 'NU', 'MU', 'X_MIN', 'X_MAX', 'LAMBDA_FACTOR')
               for s in 'RSGHI')


def test_config_defaults_and_validation(tmp_path):
    from cosmo_pol_b200.config import cfg
    ref = json.load(open(os.path.join(GOLDEN, 'ref_constants.json')))['defaults']
    for sec, d in ref.items():
        for k, v in d.items():
            assert cfg.DEFAULTS[sec][k] == v, (sec, k)
    yml = tmp_path / 'opt.yml'
    yml.write_text('radar:\n    coords: [47.28436, 8.51201, 891]\n    frequency: 5.6\n    range: 150000\n'
                   '    radial_resolution: 150\n    3dB_beamwidth: 1.\nintegration:\n    nh_GH: 4\n    nv_GH: 3\n'
                   'doppler:\n    scheme: 2\n')
    c = cfg.sanity_check(cfg.init(str(yml)))
    assert c['radar']['frequency'] == 5.6 and c['doppler']['scheme'] == 2
    assert c['integration']['nh_GH'] == 3          # 4 is invalid (odd values only) -> default
    assert c['microphysics']['with_ice_crystals'] == 1
    with pytest.raises(ValueError):
        cfg.sanity_check({'radar': {'frequency': 5.6}})      # coords are mandatory
    assert cfg.init(str(tmp_path / 'missing.yml')) is None and cfg.CONFIG['radar']['range'] == 150000


def test_lut_file_roundtrip_and_floor_lookup(tmp_path):
    from cosmo_pol_b200 import synthetic
    from cosmo_pol_b200.lookup import load_lut, save_lut, load_all_lut
    lut = synthetic.make_lut('S', 5.6, '1mom', n_e=10, n_t=12, n_d=64)
    f = tmp_path / 'lut_tmatrix_masc' / 'lut_SZ_S_5_6_1mom.lut'
    os.makedirs(f.parent)
    save_lut(lut, str(f))
    import tarfile
    assert sorted(m.name for m in tarfile.open(str(f)).getmembers()) == sorted(
        ['axes.npy', 'axes_limits.npy', 'axes_names.npy', 'axes_step.npy', 'value_table.npy'])
    back = load_all_lut('1mom', ['S'], 5.6, 'tmatrix_masc', folder=str(tmp_path))['S']
    assert np.array_equal(back.value_table, lut.value_table) and back.axes_names == lut.axes_names
    g = np.load(os.path.join(GOLDEN, 'ref_units.npz'))
    assert np.array_equal(back.lookup_line(e=g['lut_e'], t=g['lut_t']), g['lut_line'])     # reference output
    with pytest.raises(IOError):
        load_all_lut('1mom', ['G'], 5.6, 'tmatrix_masc', folder=str(tmp_path))


def test_reads_a_table_file_written_by_the_reference():
    """tests/golden/ref_saved_by_reference.lut was written by the reference's own save_lut (cosmo_pol/lookup/lut.py:78-124)
    from a table assembled with its own Lookup_table (make_golden_from_reference.py golden_lut_file); the loader of this
    package must read it into the same arrays and answer lookup_line like the reference's load_lut + lookup_line."""
    from cosmo_pol_b200.lookup import load_lut
    g = np.load(os.path.join(GOLDEN, 'ref_lut_file.npz'))
    lut = load_lut(os.path.join(GOLDEN, 'ref_saved_by_reference.lut'))
    assert dict(lut.axes_names) == json.loads(str(g['names']))
    assert np.array_equal(np.asarray(lut.value_table), g['value_table'])
    assert np.array_equal(np.asarray(lut.axes_limits, dtype=np.float64), g['axes_limits'])
    assert np.array_equal(np.asarray(lut.axes_step, dtype=np.float64), g['axes_step'])
    assert np.array_equal(np.asarray(lut.axes[lut.axes_names['d']], dtype=np.float64), g['axis_d'])
    assert np.array_equal(lut.lookup_line(e=g['e'], t=g['t']), g['line'])


def test_quadrature_fan():
    from cosmo_pol_b200.interpolation.quadrature import subbeam_fan
    import radop_oracle as O
    c = O.default_config()
    c['integration'].update({'nh_GH': 3, 'nv_GH': 5, 'weight_threshold': 0.9})
    fan = subbeam_fan(c)
    ph, pv, w, keep = O.gauss_hermite_subbeams(c)
    assert np.array_equal(fan.pts_hor, ph) and np.array_equal(fan.pts_ver, pv) and np.array_equal(fan.weights, w)
    assert fan.n_sub == keep.sum() < 15 and abs(w.sum() - 1) < 1e-15
    c['integration'].update({'nh_GH': 1, 'nv_GH': 1, 'weight_threshold': 1.})
    fan = subbeam_fan(c)
    assert fan.n_sub == 1 and fan.weight[0] == 1.0 and fan.daz[0] == 0.0 and fan.dele[0] == 0.0
    # the reference's own sub-beam list (golden): order and weights
    g = np.load(os.path.join(GOLDEN, 'ref_path_c1mom_thr.npz'))
    cfg = json.loads(str(g['config_json']))
    fan = subbeam_fan(cfg)
    az, el = g['rays'][0]
    assert np.array_equal(fan.weight, g['sub_w'][0])
    assert np.array_equal(np.stack([az + fan.daz, el + fan.dele], axis=1), g['sub_pt'][0])


def test_quadrature_fans_on_an_antenna_diagram():
    """Integration schemes 3 and 5 (interpolation.py:232-258, 276-303) against the sub-beam lists the reference itself
    produced (tests/golden/ref_path_c1mom_quad3/5.npz): weights and pointing angles bit for bit; and the configuration-time
    sum-of-Gaussians fit of the diagram (config/antenna_fit.py)."""
    from conftest import load_golden_case
    from cosmo_pol_b200.interpolation.quadrature import subbeam_fan
    from cosmo_pol_b200.config import antenna_fit, cfg as cfgmod
    for name, n_min in (('c1mom_quad3', 10), ('c1mom_quad5', 10)):
        g, cfg, kw = load_golden_case(name)
        fan = subbeam_fan(cfg)
        az, el = g['rays'][0]
        assert fan.n_sub == g['sub_w'].shape[1] >= n_min
        assert np.array_equal(fan.weight, g['sub_w'][0]) and abs(fan.weight.sum() - 1) < 1e-3
        assert np.array_equal(np.stack([az + fan.daz, el + fan.dele], axis=1), g['sub_pt'][0])
    cfg['integration']['antenna_diagram'] = None
    with pytest.raises(ValueError):
        subbeam_fan(cfg)
    # the fit: a diagram made of two known Gaussians comes back
    x = np.round(np.arange(-3., 3.001, 0.05), 2)
    truth = np.array([[0., 0., 0.3], [-30., 1.5, 0.2], [-30., -1.5, 0.2]])
    y = antenna_fit.gaussian_sum_db(x, truth)
    got = antenna_fit.optimize_gaussians(x, y, 3)
    start = np.array([[0., 0., 0.5], [-30., 1.5, 0.5], [-30., -1.5, 0.5]])
    err = lambda p: np.sqrt(np.sum((antenna_fit.gaussian_sum_db(x, p) - y) ** 2))
    assert got.shape == (3, 3) and got[0, 0] == 0 and got[0, 1] == 0 and err(got) < err(start)    # main lobe pinned
    one = antenna_fit.optimize_gaussians(x, antenna_fit.gaussian_sum_db(x, truth[:1]), 1)
    assert one.shape == (1, 3) and abs(one[0, 2] - 0.3) < 1e-3
    lin = antenna_fit.antenna_diagram(x, [1.0], [0.0], [0.0])
    assert abs(lin[60] - 1) < 1e-12 and abs(lin[np.argmin(np.abs(x - 0.5))] - 0.5) < 1e-12        # 3 dB width
    # and through the configuration check: a named diagram yields antenna_params for scheme 2
    c = cfgmod.sanity_check({'radar': {'coords': [46., 8., 1000.], 'range': 30000, 'radial_resolution': 1000, 'frequency': 5.6},
                             'integration': {'scheme': 2, 'n_gaussians': 3,
                                             'antenna_diagram': os.path.join(GOLDEN, 'antenna_diagram.csv')}})
    assert np.asarray(c['integration']['antenna_params']).shape == (3, 3)
    assert subbeam_fan(c).n_sub > 9


def test_refraction_matches_reference_bitwise():
    from cosmo_pol_b200.interpolation.atm_refraction import compute_trajectory_radial
    g = np.load(os.path.join(GOLDEN, 'ref_units.npz'))
    coords = [47.28436, 8.51201, 891.]
    for i, el in enumerate(g['traj_elev']):
        s, h, e = compute_trajectory_radial(g['derived_range'], el, coords, 1)
        assert s.dtype == np.float32 and np.array_equal(s, g['traj_s'][i])
        assert np.array_equal(h, g['traj_h'][i]) and np.array_equal(e, g['traj_e'][i])
    with pytest.raises(AttributeError):      # scheme 2 without a refractivity field (N = 0), as in the reference
        compute_trajectory_radial(g['derived_range'], 1.0, coords, 2)


def test_ode_refraction_matches_reference_bitwise():
    """Refraction scheme 2 (Zeng and Blahak 2014, atm_refraction.py:56-148) through the synthetic
    refractivity field: host implementation and oracle against vectors produced by the reference."""
    import radop_oracle as O
    from cosmo_pol_b200 import synthetic
    from cosmo_pol_b200.interpolation.atm_refraction import compute_trajectory_radial
    g = np.load(os.path.join(GOLDEN, 'ref_units.npz'), allow_pickle=True)
    model = synthetic.make_model(scheme='1mom', with_refractivity=True, nz=24, ny=64, nx=64, res=0.04)
    pm = synthetic.model_as_plain(model)
    N = {'data': pm['refractivity'], 'z_levels': pm['z_levels'], 'proj': pm['proj'], 'resolution': pm['resolution']}
    coords = list(synthetic.ALBIS)
    for i, el in enumerate(g['ode_elev']):
        for s, h, e in (compute_trajectory_radial(g['ode_range'], el, coords, 2, model['N']),
                        O.ref_ode(g['ode_range'], el, coords, N)):
            assert s.dtype == np.float32 and s[0] == 0
            assert np.array_equal(s, g['ode_s'][i]) and np.array_equal(h, g['ode_h'][i]) and np.array_equal(e, g['ode_e'][i])


def test_melting_hydrometeor_classes_match_oracle():
    """Host mirror of MeltingSnow / MeltingGraupel (hydrometeors.py:303-500, 1393-1481) against the
    oracle's restatement, which the reference goldens pin (test_oracle_golden.py, c1mom_melt*)."""
    import radop_oracle as O
    from cosmo_pol_b200.hydrometeors import create_hydrometeor
    rng = np.random.default_rng(11)
    # float32 values in float64 arrays, as melting() leaves them in the sub-radial (melting.py:62-80)
    q = rng.uniform(1e-5, 2e-3, 7).astype(np.float32).astype(np.float64)
    fw = rng.uniform(0.02, 0.97, 7).astype(np.float32).astype(np.float64)
    T = rng.uniform(268, 276, 7).astype(np.float32)
    for kind in ('mS', 'mG'):
        a, b = create_hydrometeor(kind, '1mom'), O.MeltHydro(kind, '1mom')
        args = (T, q, fw) if kind == 'mS' else (q, fw)
        a.set_psd(*args)
        b.set_psd(*args)
        assert np.allclose(a.d_min, b.d_min, rtol=1e-7) and np.allclose(a.d_max, b.d_max, rtol=1e-7)
        D = O.vlinspace(b.d_min, b.d_max, 1024)
        assert np.allclose(a.get_N(D), b.get_N(D), rtol=1e-9)
        assert np.allclose(a.get_V(D), b.get_V(D), rtol=1e-12)
        for x, y in zip(a.integrate_V(), b.integrate_V()):
            assert np.allclose(x, y, rtol=1e-9)
        # the PSD carries the prescribed mass
        assert np.allclose(a.integrate_M(), q, rtol=1e-6)
        f = a.abi_fields()
        assert f['psd'] == 4 and len(f['melt']) == 8


@pytest.mark.parametrize('scheme,kinds', [('1mom', 'RSGI'), ('2mom', 'RSGHI')])
def test_hydrometeor_classes_match_reference(scheme, kinds):
    from cosmo_pol_b200.hydrometeors import create_hydrometeor
    g = np.load(os.path.join(GOLDEN, 'ref_units.npz'))
    for k in kinds:
        tag = 'psd_{}_{}_'.format(scheme, k)
        hy = create_hydrometeor(k, scheme)
        D = g[tag + 'D']
        hy.d_min, hy.d_max = D[0], D[-1]
        if scheme == '1mom':
            hy.set_psd(*([g[tag + 'T'], g[tag + 'QM']] if k in 'SI' else [g[tag + 'QM']]))
        else:
            hy.set_psd(g[tag + 'QN'], g[tag + 'QM'])
        assert np.allclose(hy.lambda_, g[tag + 'lambda'], rtol=1e-5)
        assert np.allclose(hy.get_N(D), g[tag + 'N'], rtol=2e-3, atol=1e-300)
        f = hy.abi_fields()
        assert f['nu'] == hy.nu and f['species'] in range(5) and f['psd'] in range(4)
    with pytest.raises(ValueError):
        create_hydrometeor('X', scheme)


def test_abi_library_exports_every_declared_symbol():
    """libradop.so loads without a GPU and exports exactly what include/radop.h declares."""
    from cosmo_pol_b200 import _abi
    header = open(os.path.join(ROOT, 'include', 'radop.h')).read()
    declared = sorted(set(re.findall(r'^(?:int|const char \*)\s*\*?(radop_\w+)\s*\(', header, flags=re.M)))
    assert len(declared) >= 20
    assert sorted(_abi.EXPORTS) == declared
    lib = _abi.load_library()
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.radop_abi_version() == _abi.RADOP_ABI_VERSION
    # struct layouts agree with the header (sizes computed by the C compiler)
    src = '#include "radop.h"\n#include <stdio.h>\nint main(){printf("%zu %zu %zu %zu %zu", sizeof(RadopSpecies), sizeof(RadopParams), sizeof(RadopScan), sizeof(RadopOut), sizeof(RadopStats));return 0;}'
    import subprocess, tempfile
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, 't.c'), 'w').write(src)
        subprocess.check_call(['gcc', '-I', os.path.join(ROOT, 'include'), '-o', os.path.join(d, 't'), os.path.join(d, 't.c')])
        sizes = [int(x) for x in subprocess.check_output([os.path.join(d, 't')]).split()]
    assert sizes == [ctypes.sizeof(_abi.RadopSpecies), ctypes.sizeof(_abi.RadopParams), ctypes.sizeof(_abi.RadopScan),
                     ctypes.sizeof(_abi.RadopOut), ctypes.sizeof(_abi.RadopStats)]


def test_integration_stub_matches_library():
    """The reference-side ctypes stub documented in INTEGRATION.md is generated from include/radop.h; execute it
    against the built library: its own assertions compare every struct size with radop_sizeof()."""
    import subprocess
    from cosmo_pol_b200 import _abi
    doc = open(os.path.join(ROOT, 'INTEGRATION.md')).read()
    gen = subprocess.check_output([sys.executable, os.path.join(ROOT, 'scripts', 'gen_integration_stub.py'), '--print'],
                                  universal_newlines=True)
    assert gen in doc, 'INTEGRATION.md stub is stale: run python scripts/gen_integration_stub.py'
    ns = {}
    exec(gen.replace("C.CDLL('libradop.so')", 'C.CDLL({!r})'.format(_abi.LIB_PATH)), ns)
    for name in ('RadopParams', 'RadopSpecies', 'RadopScan', 'RadopOut', 'RadopStats'):
        assert ctypes.sizeof(ns[name]) == ctypes.sizeof(getattr(_abi, name)), name
    assert ns['lib'].radop_sizeof(99) == -1
    with pytest.raises(IndexError):
        ns['check'](None, 3)


def test_no_cpu_fallback_without_gpu():
    from conftest import has_gpu
    if has_gpu():
        pytest.skip('GPU present')
    from cosmo_pol_b200 import _abi
    with pytest.raises(_abi.RadopError):
        _abi.Engine(0)


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, 'cosmo_pol_b200')):
        for f in files:
            if f.endswith('.py'):
                src = open(os.path.join(dirpath, f)).read()
                assert 'radop_oracle' not in src and 'oracle/' not in src.replace('oracle/_ref', ''), f


def test_output_adapter_matches_per_radial_assembly():
    """RadopVolume against the reference's own assembly order, restated here from
    pyart_wrapper.py:205-327 over the list (sweeps) of lists (rays) of Radial objects."""
    from cosmo_pol_b200.radar_operator import SimulatedScan
    rng = np.random.default_rng(5)
    ns, nr, ng, nv = 3, 4, 6, 5
    fields = {k: rng.uniform(0.1, 50, (ns, nr, ng)).astype(np.float32) for k in ('ZH', 'ZV', 'ZDR', 'KDP', 'RVEL')}
    fields['ZH'][0, 1, 2] = 0.0                      # exact zero -> NaN before the dB conversion
    fields['KDP'][1, 0, 3] = np.nan
    fields['DSPECTRUM'] = rng.uniform(0, 1, (ns, nr, ng, nv)).astype(np.float32)
    lats, lons = rng.uniform(46, 48, (ns, nr, ng)), rng.uniform(7, 9, (ns, nr, ng))
    mask = np.zeros((ns, nr, ng), np.float32)
    els, azs, ranges = np.array([1.0, 5.0, 9.0]), np.array([0., 90., 180., 270.]), np.arange(ng) * 300. + 150.
    pos = {'latitude': 47.28, 'longitude': 8.51, 'altitude': 891., 'time': '2026-01-01 00:00:00'}
    scan = SimulatedScan('ppi', els, azs, ranges, pos, fields, mask, lats, lons, lats * 0, lats * 0)
    vol = scan.to_volume()
    # the reference's loop
    data = scan.data
    exp, start, stop, el_all, az_all, i0 = {}, [], [], [], [], 0
    for i in range(ns):
        sweep = {}
        for k in fields:
            a = np.array([it.values[k] for it in data[i]])
            if k in ('ZDR', 'ZV', 'ZH', 'DSpectrum'):
                a[a == 0] = float('nan')
                with np.errstate(invalid='ignore'):
                    a = 10 * np.log10(a)
            sweep[k] = a
        sweep['Latitude'] = np.array([it.lats_profile for it in data[i]])
        sweep['Longitude'] = np.array([it.lons_profile for it in data[i]])
        start.append(i0); stop.append(i0 + nr - 1); i0 += nr
        el_all.extend([els[i]] * nr); az_all.extend(list(azs))
        for k, a in sweep.items():
            exp[k] = a if k not in exp else np.vstack([exp[k], a])
    for k, a in exp.items():
        got = vol.fields[k]['data']
        assert got.shape == a.shape, k
        assert np.array_equal(np.ma.getmaskarray(got), np.isnan(a)), k
        assert np.allclose(np.ma.filled(got, 0), np.nan_to_num(a), rtol=1e-6), k
    assert np.array_equal(vol.sweep_start_ray_index['data'], start) and np.array_equal(vol.sweep_end_ray_index['data'], stop)
    assert np.array_equal(vol.elevation['data'], el_all) and np.array_equal(vol.azimuth['data'], az_all)
    assert np.array_equal(vol.fields['rangearray']['data'], np.tile(ranges, (ns * nr, 1)))
    assert vol.get_field(1, 'ZH').shape == (nr, ng) and vol.nrays == ns * nr and vol.ngates == ng
    assert np.array_equal(vol.fixed_angle['data'], els)
    rhi = SimulatedScan('rhi', np.array([0., 1., 2., 3.]), np.array([10., 20., 30.]), ranges, pos, fields, mask, lats, lons,
                        lats * 0, lats * 0).to_volume()
    assert np.array_equal(rhi.azimuth['data'], np.repeat([10., 20., 30.], nr))
    assert np.array_equal(rhi.elevation['data'], np.tile([0., 1., 2., 3.], ns))
    with pytest.raises(ImportError):
        vol.to_pyart()
