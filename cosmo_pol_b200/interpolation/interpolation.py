# -*- coding: utf-8 -*-
"""Single-ray entry points with the names of cosmo_pol/interpolation/interpolation.py.
They run a one-ray batch through the same GPU path the scans use (no CPU fallback)."""
import numpy as np

from .radial import Radial

_ACTIVE_OPERATOR = None


def _set_active_operator(op):
    global _ACTIVE_OPERATOR
    _ACTIVE_OPERATOR = op


def get_interpolated_radial(dic_variables, azimuth, elevation, N=None, list_refraction=None):
    """Sub-radials (one per kept quadrature point) of one ray, interpolation.py:91-496.
    ``dic_variables`` must be the model dictionary loaded in the active RadarOperator."""
    op = _ACTIVE_OPERATOR
    if op is None:
        raise RuntimeError('create a cosmo_pol_b200.RadarOperator and load a model first')
    if dic_variables is not None and dic_variables is not op.dic_vars:
        op.load_model(dic_variables)
    return op._single_ray_subradials(azimuth, elevation)


def subradial_digest(sub):
    """Digest of everything get_radar_observables / integrate_radials of the reference would read from a
    sub-radial (doppler_scatter.py:49-489, interpolation.py:36-89)."""
    import hashlib
    hsh = hashlib.blake2b(digest_size=16)
    for k in sorted(sub.values):
        hsh.update(k.encode())
        hsh.update(np.ascontiguousarray(sub.values[k]).tobytes())
    for a in (sub.mask, sub.elev_profile, sub.dist_profile, sub.heights_profile, sub.quad_pt, sub.quad_weight):
        hsh.update(np.ascontiguousarray(a).tobytes())
    return hsh.hexdigest()


def seal_subradials(subs, result, model, lut_sz):
    """Mark the sub-radials of one ray as the ones a fused GPU pass produced `result` / `model` from."""
    token = object()
    for i, sub in enumerate(subs):
        sub._ray = (token, i, len(subs), subradial_digest(sub), result, model, lut_sz)
    return subs


def check_sealed(list_subradials, what, lut_sz=None):
    """The interpolation and the scattering half of a ray are ONE batched GPU launch sequence here, so the
    reference's second call returns what the first one computed.  That is only right when the caller passes the
    unmodified, complete sub-radial list (and the tables the operator holds); anything else raises instead of
    returning a stale answer."""
    if not len(list_subradials) or getattr(list_subradials[0], '_ray', None) is None:
        raise ValueError('{} needs the sub-radials returned by cosmo_pol_b200.interpolation.get_interpolated_radial '
                         '(GPU path, no CPU fallback)'.format(what))
    token, _, count, _, result, model, luts = list_subradials[0]._ray
    if len(list_subradials) != count:
        raise ValueError('{}: pass the complete list of sub-radials of the ray ({} expected, {} given); the GPU path '
                         'computes a ray in one pass'.format(what, count, len(list_subradials)))
    for i, sub in enumerate(list_subradials):
        r = getattr(sub, '_ray', None)
        if r is None or r[0] is not token or r[1] != i:
            raise ValueError('{}: the sub-radials must be those of ONE get_interpolated_radial call, in its order'.format(what))
        if subradial_digest(sub) != r[3]:
            raise ValueError('{}: sub-radial {} was modified after get_interpolated_radial returned it; the GPU path '
                             'cannot recompute from edited profiles -- edit the model variables and call '
                             'get_interpolated_radial again'.format(what, i))
    if lut_sz is not None and luts is not None and lut_sz is not luts:
        raise ValueError('{}: lut_sz is not the table dictionary installed in the active RadarOperator (set_lut)'.format(what))
    return result, model


def integrate_radials(list_subradials):
    """Quadrature-weighted model variables of one ray (interpolation.py:36-89)."""
    _, model = check_sealed(list_subradials, 'integrate_radials')
    if model is None:
        raise ValueError("integrate_radials needs a RadarOperator with output_variables 'all' or 'only_model'")
    return model
