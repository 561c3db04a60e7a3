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


def integrate_radials(list_subradials):
    """Quadrature-weighted model variables of one ray (interpolation.py:36-89)."""
    res = getattr(list_subradials[0], '_ray_model', None) if len(list_subradials) else None
    if res is None:
        raise ValueError('integrate_radials needs the sub-radials returned by get_interpolated_radial of this package')
    return res
