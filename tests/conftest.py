import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..'))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA GPU (B200); run with -m gpu')


@pytest.fixture(scope='session', autouse=True)
def _native_built():
    """The oracle's C libraries (and libradop.so) must exist; build them once if missing."""
    need = [os.path.join(ROOT, 'oracle', 'liboracle_port.so'), os.path.join(ROOT, 'cosmo_pol_b200', 'libradop.so')]
    if not all(os.path.exists(p) for p in need):
        import __graft_entry__ as g
        g.build()


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


_CACHE = {}


MELTING_N_E = 8      # elevation bins of the synthetic melting tables, as in make_golden_from_reference.py


def synthetic_inputs(scheme='1mom', freq=5.6, with_melting=False, **model_kw):
    """Seeded synthetic model + lookup tables (cached per session)."""
    from cosmo_pol_b200 import synthetic
    key = (scheme, freq, with_melting, tuple(sorted(model_kw.items())))
    if key not in _CACHE:
        kw = dict(nz=24, ny=64, nx=64, res=0.04)
        kw.update(model_kw)
        model = synthetic.make_model(scheme=scheme, **kw)
        lkey = ('lut', scheme, freq)
        if lkey not in _CACHE:
            _CACHE[lkey] = synthetic.make_all_luts(freq, scheme, with_ice=True)
        luts = _CACHE[lkey]
        if with_melting:
            mkey = ('lut_melting', scheme, freq)
            if mkey not in _CACHE:
                _CACHE[mkey] = {h: synthetic.make_melting_lut(h, freq, scheme, n_e=MELTING_N_E) for h in ('mS', 'mG')}
            luts = dict(luts)
            luts.update(_CACHE[mkey])
        _CACHE[key] = (model, luts)
    return _CACHE[key]


def load_golden_case(name):
    g = np.load(os.path.join(GOLDEN, 'ref_path_{}.npz'.format(name)))
    cfg = json.loads(str(g['config_json']))
    nyq = float(g['nyquist'])
    cfg['radar']['nyquist_velocity'] = nyq if nyq > 0 else None
    if cfg.get('integration', {}).get('antenna_diagram'):     # the generator recorded its own absolute path
        cfg['integration']['antenna_diagram'] = os.path.join(GOLDEN, os.path.basename(cfg['integration']['antenna_diagram']))
    kw = json.loads(str(g['model_kw_json']))
    return g, cfg, kw


def plain(model, luts):
    from cosmo_pol_b200 import synthetic
    return synthetic.model_as_plain(model), {h: synthetic.lut_as_plain(l) for h, l in luts.items()}


def db(x):
    with np.errstate(divide='ignore', invalid='ignore'):
        return 10 * np.log10(x)


def assert_observables_close(got, ref, sz_scale=None, what=''):
    """Tolerances of BASELINE.json north_star: ZH/ZV/ZDR 0.01 dB, KDP 1e-3 relative,
    RHOHV 1e-4 absolute, attenuation 1e-3 relative.  NaN patterns must be identical."""
    for k in ('ZH', 'ZV', 'ZDR', 'KDP', 'RHOHV', 'ATT_H', 'ATT_V', 'PHIDP', 'DELTA_HV'):
        a = np.asarray(got[k], dtype=np.float64)
        b = np.asarray(ref[k], dtype=np.float64)
        assert a.shape == b.shape, (what, k, a.shape, b.shape)
        assert np.array_equal(np.isnan(a), np.isnan(b)), '{} {}: NaN pattern differs'.format(what, k)
        ok = np.isfinite(b)
        if not ok.any():
            continue
        if k in ('ZH', 'ZV', 'ZDR'):
            err = np.max(np.abs(db(a[ok]) - db(b[ok])))
            assert err < 0.01, '{} {}: {} dB'.format(what, k, err)
        elif k == 'KDP':
            tol = 1e-3 * np.abs(b[ok])
            if sz_scale is not None:
                tol = tol + sz_scale[ok]      # float32 cancellation floor of the reference itself
            assert np.all(np.abs(a[ok] - b[ok]) <= tol), '{} KDP: {}'.format(what, np.max(np.abs(a[ok] - b[ok]) / np.abs(b[ok])))
        elif k == 'RHOHV':
            assert np.max(np.abs(a[ok] - b[ok])) < 1e-4, (what, k)
        elif k in ('ATT_H', 'ATT_V'):
            assert np.max(np.abs(a[ok] - b[ok]) / np.abs(b[ok])) < 1e-3, (what, k)
        elif k == 'PHIDP':
            assert np.max(np.abs(a[ok] - b[ok]) / (np.abs(b[ok]) + 1e-3)) < 2e-3, (what, k)
        else:
            assert np.max(np.abs(a[ok] - b[ok])) < 1e-4, (what, k)
