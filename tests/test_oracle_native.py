"""The oracle's C restatement (oracle/oracle_port.c) against the reference's own C sources
compiled unmodified (oracle/_ref): bit-exact, including the NaN / -9999 sentinels."""
import os

import numpy as np
import pytest

import radop_oracle as O
from conftest import synthetic_inputs, plain

needs_ref = pytest.mark.skipif(O.native().ref_interp is None, reason='oracle/_ref not built (no /root/reference)')


def _random_gates(model, n, seed):
    rng = np.random.default_rng(seed)
    proj, res = model['proj'], model['resolution']
    nz, ny, nx = model['z_levels'].shape
    rlat = proj['La1'] + rng.uniform(0.2, ny - 1.2, n) * res[1]
    rlon = proj['Lo1'] + rng.uniform(0.2, nx - 1.2, n) * res[0]
    h = rng.uniform(100., 26000., n)
    h[: n // 8] = rng.uniform(0., 900., n // 8)          # plenty below topography
    coords = np.stack([rlat, rlon], axis=1).astype(np.float32)
    return coords, h.astype(np.float32)


@needs_ref
def test_port_matches_reference_c_bit_exact():
    model, _ = plain(*synthetic_inputs())
    llc = np.asarray([model['proj']['Lo1'], model['proj']['La1']], dtype=np.float32)
    coords, h = _random_gates(model, 4000, 1)
    for name in ('T', 'QR_v', 'U'):
        a = O.interp_ref(coords, h, model['vars'][name], model['z_levels'], llc, model['resolution'])
        b, cell, level, status = O.interp_port(coords, h, model['vars'][name], model['z_levels'], llc,
                                               model['resolution'], debug=True)
        assert np.array_equal(a.view(np.uint32) & 0x7fffffff, b.view(np.uint32) & 0x7fffffff) or \
            np.array_equal(a, b, equal_nan=True)
        assert np.array_equal(np.isnan(a), status == -1)
        assert np.array_equal(a == -9999, status == 1)
        assert (status == 0).sum() > 1000 and (status == -1).sum() > 10 and (status == 1).sum() > 10
        ok = status == 0
        assert np.all(level[ok] >= 0) and np.all(level[ok] <= model['z_levels'].shape[0] - 3)


@needs_ref
def test_lowest_layer_never_used_and_linear_field_reproduced():
    """interpolation_c.c:74-82 (quirk Q2) and exactness on a field linear in height."""
    model, _ = plain(*synthetic_inputs())
    z = model['z_levels']
    lin = (0.001 * z + 3.0).astype(np.float32)
    llc = np.asarray([model['proj']['Lo1'], model['proj']['La1']], dtype=np.float32)
    coords, h = _random_gates(model, 2000, 2)
    out, cell, level, status = O.interp_port(coords, h, lin, z, llc, model['resolution'], debug=True)
    ok = status == 0
    assert level[ok].max() <= z.shape[0] - 3
    # bilinear average of columns that are each linear in h: 0.001*h + 3 up to rounding
    assert np.allclose(out[ok], 0.001 * h[ok] + 3.0, rtol=2e-5, atol=2e-4)


@needs_ref
def test_spectrum_bins_port_matches_reference_c():
    rng = np.random.default_rng(3)
    nh, nd, nb = 3, 1024, 40
    D = np.stack([np.linspace(0.1 * (j + 1), 8. + j, nd) for j in range(nh)], axis=1).astype(np.float32)
    step = (D[1] - D[0]).astype(np.float32)
    dmin = D[0].copy()
    N = rng.uniform(0, 1e3, (nd, nh)).astype(np.float32)
    rcs = rng.uniform(0, 1e-3, (nd, nh)).astype(np.float32)
    a = np.sort(rng.uniform(0, 1, (nb + 1, nh)), axis=0)
    edges = (dmin + a * (D[-1] - dmin)).astype(np.float32)
    Da, Db = edges[:-1], edges[1:]
    x = O.spectrum_bins_ref(Da, Db, rcs, D, N, step, dmin)
    y = O.spectrum_bins_port(Da, Db, rcs, N, step, dmin)
    assert np.array_equal(x, y)
