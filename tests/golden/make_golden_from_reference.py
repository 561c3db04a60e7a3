#!/usr/bin/env python
# -*- coding: utf-8 -*-
"""
make_golden_from_reference.py -- generate golden vectors by RUNNING THE REFERENCE ITSELF.

Run in the build container only (needs /root/reference; the GPU box never runs this):

    python tests/golden/make_golden_from_reference.py

The reference package does not import as-is under python 3.12 / numpy 2 / scipy 1.18 and
three of its dependencies are absent (pycosmo, pyproj; pytmatrix for the tables).  This
script imports the reference's OWN modules from /root/reference, file by file, after
  (a) restoring removed aliases:  np.float, np.warnings, scipy.integrate.trapz,
      scipy.interpolate.interp2d (never called);
  (b) applying four py2->py3 source patches IN MEMORY (nothing is written, nothing is copied
      into this repo):   dict.keys()/values() -> list(...) in interpolation.py (the code
      indexes them), see PATCHES below;
  (c) providing stand-ins for the absent third-party calls:
        pyproj.Geod.fwd        -> Vincenty direct geodesic on WGS84 (oracle.geod_fwd)
        pycosmo.WGS_to_COSMO   -> rotated-pole rotation            (oracle.wgs_to_rotated)
      and binding the two SWIG shims (interpolation_c / doppler_c, built for python 2.7) to
      the reference's own C sources compiled unmodified into oracle/_ref (ctypes).
Everything else -- quadrature, refraction, masks, PSDs, lookup, einsum, Doppler, observables,
scans -- is the reference's code executing.

Outputs (committed): tests/golden/ref_constants.json, ref_units.npz, ref_path_<case>.npz
"""
import importlib.util
import json
import os
import sys
import types
import warnings

import numpy as np
import scipy.integrate
import scipy.interpolate

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.abspath(os.path.join(HERE, '..', '..'))
REF = '/root/reference'
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'oracle'))

import radop_oracle as O                      # noqa: E402  (stand-ins + reference C via ctypes)
from cosmo_pol_b200 import synthetic          # noqa: E402  (seeded inputs only)

# ---------------------------------------------------------------------------------------- (a)
np.float = float
np.warnings = warnings
if not hasattr(scipy.integrate, 'trapz'):
    scipy.integrate.trapz = np.trapezoid
if not hasattr(np, 'trapz'):
    np.trapz = np.trapezoid
if not hasattr(scipy.interpolate, 'interp2d'):
    scipy.interpolate.interp2d = object

# ---------------------------------------------------------------------------------------- (b)
PATCHES = {
    'cosmo_pol/interpolation/interpolation.py': [
        ('    list_variables = dic_variables.values()\n    keys = dic_variables.keys()\n',
         '    list_variables = list(dic_variables.values())\n    keys = list(dic_variables.keys())\n'),
    ],
    # only exercised by golden_lut_file: dict.has_key is python 2; np.save of the ragged list of axes made an
    # object array by itself up to numpy 1.23 and raises since; see also the third patch
    'cosmo_pol/lookup/lut.py': [
        ('if self.axes_names.has_key(name):', 'if name in self.axes_names:'),
        ("np.save(tmp_dir + 'axes',lut.axes)", "np.save(tmp_dir + 'axes', np.array(lut.axes, dtype=object))"),
        # ndarray.all() of a 0-d object array returned the object (the dict) up to numpy 1.x
        ('            data = data.all()\n', '            data = data.item()\n'),
    ],
}


def _load(modname, relpath, package=None):
    path = os.path.join(REF, relpath)
    src = open(path).read()
    for old, new in PATCHES.get(relpath, []):
        assert old in src, (relpath, old)
        src = src.replace(old, new)
    mod = types.ModuleType(modname)
    mod.__file__ = path
    mod.__package__ = package if package is not None else modname.rpartition('.')[0]
    sys.modules[modname] = mod
    exec(compile(src, path, 'exec'), mod.__dict__)
    return mod


def _pkg(name, relpath):
    m = types.ModuleType(name)
    m.__path__ = [os.path.join(REF, relpath)]
    m.__package__ = name
    sys.modules[name] = m
    return m


# ---------------------------------------------------------------------------------------- (c)
pyproj = types.ModuleType('pyproj')


class _Geod(object):
    def __init__(self, ellps='WGS84'):
        assert ellps == 'WGS84'

    def fwd(self, lon, lat, az, dist):
        lo, la = O.geod_fwd(lon, lat, az, np.array([dist], dtype=np.float64))
        return float(lo[0]), float(la[0]), 0.0


pyproj.Geod = _Geod
sys.modules['pyproj'] = pyproj

pycosmo = types.ModuleType('pycosmo')


def _wgs_to_cosmo(coords, sp):
    lats, lons = coords[0], coords[1]      # _ref_ODE passes [lat, lon, alt] (atm_refraction.py:100-103)
    r = O.wgs_to_rotated(lats, lons, sp[0], sp[1])
    return r[0] if np.ndim(lats) == 0 else r          # scalar position -> (rlat, rlon)


pycosmo.WGS_to_COSMO = _wgs_to_cosmo
sys.modules['pycosmo'] = pycosmo


def _get_all_radar_pts(n, coords, heights, data, model_heights, llc, res):
    return [None, O.interp_ref(coords, heights, data, model_heights, llc, res)]


def _get_refl(n, Da, Db, rcs, D, N, step_D, D_min):
    return [None, O.spectrum_bins_ref(Da, Db, rcs, D, N, step_D, D_min)]


def import_reference():
    cp = _pkg('cosmo_pol', 'cosmo_pol')
    # utilities
    _pkg('cosmo_pol.utilities', 'cosmo_pol/utilities')
    u = _load('cosmo_pol.utilities.utilities', 'cosmo_pol/utilities/utilities.py')
    sys.modules['cosmo_pol.utilities'].__dict__.update({k: getattr(u, k) for k in dir(u) if not k.startswith('__')})
    # config
    _pkg('cosmo_pol.config', 'cosmo_pol/config')
    _load('cosmo_pol.config.antenna_fit', 'cosmo_pol/config/antenna_fit.py')
    nf = types.ModuleType('cosmo_pol.config.nyquist_fit')
    nf.nyquist_interpolator = lambda f: None
    sys.modules['cosmo_pol.config.nyquist_fit'] = nf
    cfg = _load('cosmo_pol.config.cfg', 'cosmo_pol/config/cfg.py', package='cosmo_pol.config')
    sys.modules['cosmo_pol.config'].cfg = cfg
    sys.modules['cosmo_pol.config'].CONFIG = None
    # constants
    _pkg('cosmo_pol.constants', 'cosmo_pol/constants')
    gc = _load('cosmo_pol.constants.global_constants', 'cosmo_pol/constants/global_constants.py',
               package='cosmo_pol.constants')
    # the package __init__ rebinds the name to the instance
    sys.modules['cosmo_pol.constants'].global_constants = gc.global_constants
    c1 = _load('cosmo_pol.constants.constants_1mom', 'cosmo_pol/constants/constants_1mom.py', package='cosmo_pol.constants')
    c2 = _load('cosmo_pol.constants.constants_2mom', 'cosmo_pol/constants/constants_2mom.py', package='cosmo_pol.constants')
    sys.modules['cosmo_pol.constants'].constants_1mom = c1
    sys.modules['cosmo_pol.constants'].constants_2mom = c2
    # hydrometeors
    _pkg('cosmo_pol.hydrometeors', 'cosmo_pol/hydrometeors')
    _load('cosmo_pol.hydrometeors.dielectric', 'cosmo_pol/hydrometeors/dielectric.py')
    hy = _load('cosmo_pol.hydrometeors.hydrometeors', 'cosmo_pol/hydrometeors/hydrometeors.py')
    sys.modules['cosmo_pol.hydrometeors'].create_hydrometeor = hy.create_hydrometeor
    # lookup
    _pkg('cosmo_pol.lookup', 'cosmo_pol/lookup')
    lut = _load('cosmo_pol.lookup.lut', 'cosmo_pol/lookup/lut.py')
    # interpolation
    ip = _pkg('cosmo_pol.interpolation', 'cosmo_pol/interpolation')
    rad = _load('cosmo_pol.interpolation.radial', 'cosmo_pol/interpolation/radial.py')
    ip.Radial = rad.Radial
    atm = _load('cosmo_pol.interpolation.atm_refraction', 'cosmo_pol/interpolation/atm_refraction.py')
    ip.compute_trajectory_GPM = atm.compute_trajectory_GPM
    ip.compute_trajectory_radial = atm.compute_trajectory_radial
    mel = _load('cosmo_pol.interpolation.melting', 'cosmo_pol/interpolation/melting.py')
    ip.melting = mel.melting
    qd = _load('cosmo_pol.interpolation.quadrature', 'cosmo_pol/interpolation/quadrature.py')
    ip.gautschi_points_and_weights = qd.gautschi_points_and_weights
    ip.get_all_radar_pts = _get_all_radar_pts
    interp = _load('cosmo_pol.interpolation.interpolation', 'cosmo_pol/interpolation/interpolation.py')
    ip.get_interpolated_radial = interp.get_interpolated_radial
    ip.integrate_radials = interp.integrate_radials
    # scatter
    sc = _pkg('cosmo_pol.scatter', 'cosmo_pol/scatter')
    sc.get_refl = _get_refl
    ds = _load('cosmo_pol.scatter.doppler_scatter', 'cosmo_pol/scatter/doppler_scatter.py')
    return dict(cfg=cfg, constants=gc.global_constants, gc_module=gc, c1=c1, c2=c2, hydro=hy, lut=lut, interp=interp,
                scatter=ds, atm=atm, utilities=u, radial=rad)


def set_config(R, config):
    R['cfg'].CONFIG = config
    sys.modules['cosmo_pol.config'].CONFIG = config
    # doppler_scatter.py:26 binds CONFIG at import time (None), and spectral_width_turb / _motion
    # (:741-742, 771-772) read that stale module global: as shipped, the spectrum-broadening options
    # raise TypeError.  The goldens of those options rebind the global to the live configuration.
    R['scatter'].CONFIG = config
    R['constants'].update()


def base_config(**over):
    c = O.default_config()      # = the reference DEFAULTS + radar position (verified below)
    for sec, d in over.items():
        c[sec].update(d)
    return c


def ref_lut(R, lut):
    """Lookup_table of the reference filled the way load_lut does (setattr of the arrays)."""
    L = R['lut'].Lookup_table()
    L.value_table = lut.value_table
    L.axes = lut.axes
    L.axes_names = dict(lut.axes_names)
    L.axes_limits = np.asarray(lut.axes_limits)      # what np.save/np.load of the list gives
    try:
        L.axes_step = np.asarray(lut.axes_step)
    except ValueError:                                # melting tables: the 'd' step is a vector (object array)
        L.axes_step = np.empty(len(lut.axes_step), dtype=object)
        for i, v in enumerate(lut.axes_step):
            L.axes_step[i] = v
    return L


# --------------------------------------------------------------------------------------------
MODEL_KW = dict(nz=24, ny=64, nx=64, res=0.04)


def golden_constants(R):
    out = {}
    for tag, mod in (('1mom', R['c1']), ('2mom', R['c2'])):
        out[tag] = {k: float(getattr(mod, k)) for k in dir(mod)
                    if k.isupper() and isinstance(getattr(mod, k), (int, float, np.floating))}
    gcst = R['constants']
    out['global'] = {k: float(getattr(gcst, k)) for k in ('EPS', 'C', 'RHO_W', 'RHO_I', 'A', 'RHO_0', 'KE',
                                                          'MAX_MODEL_HEIGHT', 'T0', 'T_K_SQUARED')}
    out['defaults'] = json.loads(json.dumps(R['cfg'].DEFAULTS))
    with open(os.path.join(HERE, 'ref_constants.json'), 'w') as f:
        json.dump(out, f, indent=1, sort_keys=True)
    # the oracle's default_config must equal the reference DEFAULTS on shared keys
    d = O.default_config()
    for sec, dd in R['cfg'].DEFAULTS.items():
        for k, v in dd.items():
            if k in d[sec]:
                assert d[sec][k] == v, (sec, k)


def golden_units(R):
    rng = np.random.default_rng(777)
    g = {}
    cfgd = base_config(radar={'frequency': 5.6, 'range': 150000, 'radial_resolution': 300, 'FFT_length': 128})
    set_config(R, cfgd)
    cst = R['constants']
    g['derived_wavelength'] = np.array(cst.WAVELENGTH)
    g['derived_varray'] = np.asarray(cst.VARRAY)
    g['derived_range'] = np.asarray(cst.RANGE_RADAR)
    # refraction
    els = np.array([-0.2, 0.4, 1.0, 5.5, 20.0, 40.0, 89.0, 90.0])
    g['traj_elev'] = els
    tr = [R['atm'].compute_trajectory_radial(cst.RANGE_RADAR, e, cfgd['radar']['coords'], 1) for e in els]
    g['traj_s'] = np.stack([t[0] for t in tr])
    g['traj_h'] = np.stack([t[1] for t in tr])
    g['traj_e'] = np.stack([t[2] for t in tr])
    # ODE refraction (Zeng and Blahak 2014, atm_refraction.py:56-148) through a synthetic refractivity field
    mN = synthetic.make_model(scheme='1mom', with_refractivity=True, **MODEL_KW)['N']
    els2 = np.array([-0.2, 0.4, 1.0, 5.5, 20.0, 40.0])
    rr2 = np.arange(150., 60000., 300.)
    g['ode_elev'] = els2
    g['ode_range'] = rr2
    # numpy >= 1.24 refuses the ragged [scalar, array(1)] list _deriv_z returns (atm_refraction.py:75-77);
    # same arithmetic, elements squeezed to floats
    _orig_deriv = R['atm']._deriv_z
    R['atm']._deriv_z = lambda z, r, n_h, dn_dh, RE: [float(np.squeeze(v)) for v in _orig_deriv(z, r, n_h, dn_dh, RE)]
    tr = [R['atm'].compute_trajectory_radial(rr2, e, cfgd['radar']['coords'], 2, mN) for e in els2]
    g['ode_s'] = np.stack([t[0] for t in tr])
    g['ode_h'] = np.stack([t[1] for t in tr])
    g['ode_e'] = np.stack([t[2] for t in tr])
    g['earth_radius_in'] = np.array([0., 8.51201, 47.3, 90.])
    g['earth_radius'] = np.array([R['utilities'].get_earth_radius(x) for x in g['earth_radius_in']])
    # aliasing, nansum
    v = rng.uniform(-40, 40, 64)
    g['alias_v'] = v
    g['alias_out'] = R['utilities'].aliasing(v, 8.3)
    # get_pol_from_sz
    sz = np.abs(rng.standard_normal((32, 12))).astype('float32') * np.array(
        [5, .2, .2, 5, -4, .05, -.05, -4, 1, .01, 1.03, .011], dtype='float32')
    g['pol_sz'] = sz
    pol = R['scatter'].get_pol_from_sz(sz, 0.93)
    g['pol_out'] = np.stack([np.asarray(p, dtype=np.float64) for p in pol])
    # quadrature (through the private code path is not callable separately: replicate by
    # running get_interpolated_radial in the path goldens; here store hermgauss inputs only)
    # hydrometeors
    n = 24
    T = rng.uniform(205., 300., n).astype('float32')
    for scheme, kinds in (('1mom', 'RSGI'), ('2mom', 'RSGHI')):
        for k in kinds:
            hy = R['hydro'].create_hydrometeor(k, scheme)
            lim = synthetic.diameter_limits(k, scheme)
            D = np.linspace(lim[0], lim[1], 1024).astype('float32')
            hy.d_min, hy.d_max, hy.nbins_D = D[0], D[-1], 1024
            QM = (10 ** rng.uniform(-7, -2.5, n)).astype('float32')
            tag = 'psd_{}_{}_'.format(scheme, k)
            g[tag + 'QM'] = QM
            g[tag + 'T'] = T
            if scheme == '1mom':
                if k in 'SI':
                    hy.set_psd(T, QM)
                else:
                    hy.set_psd(QM)
            else:
                lo, hi = (getattr(R['c2'], 'X_MIN_' + k), getattr(R['c2'], 'X_MAX_' + k))
                xb = np.exp(rng.uniform(np.log(lo) - 1, np.log(hi) + 1, n))
                QN = (QM / xb).astype('float32')
                g[tag + 'QN'] = QN
                hy.set_psd(QN, QM)
            g[tag + 'lambda'] = np.asarray(hy.lambda_, dtype=np.float64)
            g[tag + 'N0'] = np.broadcast_to(np.asarray(hy.N0, dtype=np.float64), (n,)).copy()
            g[tag + 'N'] = np.asarray(hy.get_N(D), dtype=np.float64)
            vv, nn = hy.integrate_V()
            g[tag + 'v'] = np.asarray(vv, dtype=np.float64)
            g[tag + 'n'] = np.asarray(nn, dtype=np.float64)
            g[tag + 'D'] = D
    # lookup_line
    lut = synthetic.make_lut('S', 5.6, '1mom', n_e=10, n_t=12, n_d=64)
    L = ref_lut(R, lut)
    e = rng.uniform(-1, 25, 40).astype('float32')
    t = rng.uniform(195, 230, 40).astype('float32')
    g['lut_e'] = e
    g['lut_t'] = t
    g['lut_line'] = L.lookup_line(e=e, t=t)
    np.savez_compressed(os.path.join(HERE, 'ref_units.npz'), **g)


def golden_lut_file(R):
    """A lookup-table file written by the reference's OWN save_lut (lut.py:78-124) from a table assembled with its own
    Lookup_table.add_axis / set_value_table, read back with its own load_lut: tests/golden/ref_saved_by_reference.lut
    (committed, 10 KB) is what cosmo_pol_b200.lookup.load_lut must be able to read."""
    lut = synthetic.make_lut('S', 5.6, '1mom', n_e=3, n_t=4, n_d=8)
    L = R['lut'].Lookup_table()
    for name in ('e', 't', 'd', 'sz'):
        L.add_axis(name, lut.axes[lut.axes_names[name]])
    L.set_value_table(lut.value_table)
    path = os.path.join(HERE, 'ref_saved_by_reference.lut')
    R['lut'].save_lut(L, path)
    back = R['lut'].load_lut(path)
    e = np.array([-0.5, 1.0, 3.9, 9.0], dtype='float32')
    t = np.array([199.0, 202.5, 205.9, 300.0], dtype='float32')
    np.savez_compressed(os.path.join(HERE, 'ref_lut_file.npz'), value_table=np.asarray(back.value_table),
                        axes_limits=np.asarray(back.axes_limits, dtype=np.float64),
                        axes_step=np.asarray(back.axes_step, dtype=np.float64),
                        axis_d=np.asarray(back.axes[back.axes_names['d']], dtype=np.float64),
                        names=np.array(json.dumps(back.axes_names)), e=e, t=t, line=back.lookup_line(e=e, t=t))
    print('lutfile done:', os.path.getsize(path), 'bytes')


CASES = {
    # name: (scheme, freq, config overrides, rays [(az, el)], model kwargs)
    'c1mom': ('1mom', 5.6, dict(radar={'range': 45000, 'radial_resolution': 900}, integration={'nv_GH': 3, 'nh_GH': 3}),
              [(10.0, 1.0), (123.0, -0.2), (250.5, 8.5), (300.0, 40.0), (77.0, 89.5)], {}),
    'c2mom': ('2mom', 9.41, dict(radar={'range': 45000, 'radial_resolution': 900}, integration={'nv_GH': 3, 'nh_GH': 3}),
              [(10.0, 1.0), (200.0, 16.0)], {}),
    'c1mom_thr': ('1mom', 5.6, dict(radar={'range': 30000, 'radial_resolution': 1000, 'sensitivity': [5, 10000]},
                                    integration={'nv_GH': 5, 'nh_GH': 3, 'weight_threshold': 0.9},
                                    microphysics={'with_attenuation': 0, 'with_ice_crystals': 0}),
                  [(45.0, 3.5)], {}),
    'c1mom_nyq': ('1mom', 5.6, dict(radar={'range': 30000, 'radial_resolution': 1000, 'nyquist_velocity': 8.3},
                                    integration={'nv_GH': 1, 'nh_GH': 1}),
                  [(45.0, 3.5)], {}),
    'c1mom_dop2': ('1mom', 5.6, dict(radar={'range': 30000, 'radial_resolution': 1000}, doppler={'scheme': 2},
                                     integration={'nv_GH': 3, 'nh_GH': 1}),
                   [(45.0, 3.5)], {}),
    'c1mom_dop3': ('1mom', 5.6, dict(radar={'range': 20000, 'radial_resolution': 1000, 'FFT_length': 64},
                                     doppler={'scheme': 3}, integration={'nv_GH': 1, 'nh_GH': 3}),
                   [(45.0, 3.5)], {}),
    # spectrum broadening (doppler_scatter.py:357-367, 727-801): antenna motion alone, and turbulence +
    # motion with an EDR field.  Ray 2 of the second case leaves the model top, so its upper sub-beams hold
    # NaN EDR gates and are not broadened (np.sum(NaN) > 0 is False)
    'c1mom_dop3_mot': ('1mom', 5.6, dict(radar={'range': 20000, 'radial_resolution': 1000, 'FFT_length': 64},
                                         doppler={'scheme': 3, 'motion_correction': 1},
                                         integration={'nv_GH': 3, 'nh_GH': 1}),
                       [(45.0, 3.5)], {}),
    'c1mom_dop3_turb': ('1mom', 5.6, dict(radar={'range': 20000, 'radial_resolution': 1000, 'FFT_length': 64},
                                          doppler={'scheme': 3, 'motion_correction': 1, 'turbulence_correction': 1},
                                          integration={'nv_GH': 3, 'nh_GH': 1}),
                        [(45.0, 3.5), (200.0, 60.0)], {'with_edr': True, 'top': 15000.0}),
    # melting scheme (melting.py:19-90, hydrometeors.py:303-500, 1393-1481): wet snow / wet graupel
    # with per-gate diameter grids and the wet-fraction table axis; Doppler schemes 1 and 2
    'c1mom_melt': ('1mom', 5.6, dict(radar={'range': 30000, 'radial_resolution': 500},
                                     microphysics={'with_melting': 1}, integration={'nv_GH': 3, 'nh_GH': 1}),
                   [(220.0, 3.5), (170.0, 3.5)], {}),
    'c1mom_melt_dop2': ('1mom', 5.6, dict(radar={'range': 30000, 'radial_resolution': 500}, doppler={'scheme': 2},
                                          microphysics={'with_melting': 1}, integration={'nv_GH': 1, 'nh_GH': 3}),
                        [(220.0, 3.5)], {}),
    # integration scheme 'ml' (interpolation.py:168-193, 423-436; doppler_scatter.py:127-129, 186-189, 259-262): ten
    # times the vertical points, the extra ones weighted only around the edges of the melting layer of their sub-beam
    'c1mom_ml': ('1mom', 5.6, dict(radar={'range': 30000, 'radial_resolution': 500},
                                   microphysics={'with_melting': 1}, integration={'scheme': 'ml', 'nv_GH': 1, 'nh_GH': 3}),
                 [(220.0, 3.5), (170.0, 3.5)], {}),
    # integration scheme 2 (interpolation.py:194-228, 440-494): irregular (azimuth, elevation) points of a
    # two-Gaussian antenna pattern (main lobe + first side-lobe ring), one trajectory per point
    'c1mom_quad2': ('1mom', 5.6, dict(radar={'range': 30000, 'radial_resolution': 1000},
                                      integration={'scheme': 2, 'nr_GH': 3, 'na_GL': 5, 'weight_threshold': 0.999,
                                                   'antenna_params': [[0.0, 0.0, 0.4247], [-24.0, 1.6, 0.3]]}),
                    [(45.0, 3.5), (250.5, 8.5)], {}),
    # integration schemes 3 and 5 (interpolation.py:232-258, 276-303) on the antenna diagram of tests/golden/antenna_diagram.csv
    # (1 deg main lobe, side-lobe ring at 1.8 deg, -25 dB).  The reference's configuration check does not list scheme 5
    # (cfg.py:101); the generator sets CONFIG directly, as for every other case.
    'c1mom_quad3': ('1mom', 5.6, dict(radar={'range': 30000, 'radial_resolution': 1000},
                                      integration={'scheme': 3, 'nh_GH': 9, 'nv_GH': 11, 'weight_threshold': 0.99999,
                                                   'antenna_diagram': os.path.join(HERE, 'antenna_diagram.csv')}),
                    [(45.0, 3.5), (250.5, 8.5)], {}),
    'c1mom_quad5': ('1mom', 5.6, dict(radar={'range': 30000, 'radial_resolution': 1000},
                                      integration={'scheme': 5, 'nh_GH': 3, 'nv_GH': 5, 'weight_threshold': 0.9999,
                                                   'antenna_diagram': os.path.join(HERE, 'antenna_diagram.csv')}),
                    [(45.0, 3.5), (250.5, 8.5)], {}),
}
MELTING_N_E = 8          # elevation bins of the synthetic melting tables (0..14 deg): 78 MB per species


def golden_path(R, name):
    scheme, freq, over, rays, mkw = CASES[name]
    kw = dict(MODEL_KW)
    kw.update(mkw)
    model = synthetic.make_model(scheme=scheme, **kw)
    melt = bool(over.get('microphysics', {}).get('with_melting', 0))
    luts = synthetic.make_all_luts(freq, scheme, with_ice=True, with_melting=melt, n_e_melting=MELTING_N_E)
    over = {k: dict(v) for k, v in over.items()}
    over.setdefault('radar', {})['frequency'] = freq
    over.setdefault('microphysics', {})['scheme'] = scheme
    cfgd = base_config(**over)
    nyq = cfgd['radar'].get('nyquist_velocity')
    if nyq is not None:
        cfgd['radar']['nyquist_velocity'] = (lambda e, a, _v=nyq: _v)
    if 'antenna_params' in cfgd['integration']:      # interpolation.py:200-204 slices it as an array
        cfgd['integration']['antenna_params'] = np.asarray(cfgd['integration']['antenna_params'], dtype=float)
    set_config(R, cfgd)
    ref_luts = {h: ref_lut(R, l) for h, l in luts.items()}
    out = {'rays': np.array(rays)}
    per_ray = []
    for az, el in rays:
        # the worker passes N = self.N, 0 unless a refractivity field was loaded (radar_operator.py:76, 210, 417)
        subs = R['interp'].get_interpolated_radial(model, az, el, 0)
        sub_vals = {k: np.stack([np.array(s.values[k], copy=True) for s in subs]) for k in subs[0].values.keys()}
        sub_mask = np.stack([s.mask for s in subs])
        sub_w = np.array([s.quad_weight for s in subs])
        sub_pt = np.array([s.quad_pt for s in subs])
        lats = np.stack([s.lats_profile for s in subs])
        lons = np.stack([s.lons_profile for s in subs])
        integ = R['interp'].integrate_radials(subs)
        rad = R['scatter'].get_radar_observables(subs, ref_luts)
        rec = {'sub_mask': sub_mask, 'sub_w': sub_w, 'sub_pt': sub_pt, 'sub_lats': lats, 'sub_lons': lons,
               'mask': np.asarray(rad.mask), 'imask': np.asarray(integ.mask)}
        for k, v in sub_vals.items():
            rec['sub_' + k] = v
        for k, v in rad.values.items():
            rec['obs_' + k] = np.asarray(v)
        for k, v in integ.values.items():
            rec['int_' + k] = np.asarray(v)
        # censored copy (cut_at_sensitivity works in place on the list)
        if name in ('c1mom', 'c1mom_thr'):
            import copy as _copy
            cut = R['scatter'].cut_at_sensitivity([_copy.deepcopy(rad)])[0]
            for k, v in cut.values.items():
                rec['cut_' + k] = np.asarray(v)
        per_ray.append(rec)
    for k in per_ray[0].keys():
        out[k] = np.stack([r[k] for r in per_ray])
    out['config_json'] = np.array(json.dumps({s: {k: (v.tolist() if isinstance(v, np.ndarray) else v if not callable(v) else 'callable') for k, v in d.items()}
                                              for s, d in cfgd.items()}, default=float))
    out['nyquist'] = np.array(-1.0 if nyq is None else float(nyq))
    out['model_kw_json'] = np.array(json.dumps(kw))
    # fingerprint of the seeded inputs, so that a drifting generator is detected
    out['fp_T'] = np.array(float(np.sum(model['T'].data.astype(np.float64))))
    out['fp_QR'] = np.array(float(np.sum(model['QR_v'].data.astype(np.float64))))
    out['fp_lut_R'] = np.array(float(np.sum(luts['R'].value_table[::5, ::5, ::64])))
    np.savez_compressed(os.path.join(HERE, 'ref_path_{}.npz'.format(name)), **out)
    print(name, 'done:', {k: v.shape for k, v in out.items() if hasattr(v, 'shape') and k.startswith('obs_')})


if __name__ == '__main__':
    R = import_reference()
    golden_constants(R)
    golden_units(R)
    golden_lut_file(R)
    which = sys.argv[1:] or list(CASES.keys())
    for name in which:
        if name != 'lutfile':
            golden_path(R, name)
