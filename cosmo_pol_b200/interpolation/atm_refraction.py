# -*- coding: utf-8 -*-
"""Beam trajectories (host side, float64 then float32 exactly like the reference so that the
heights and elevation angles handed to the kernels are bit-identical):
4/3 effective earth radius model, cosmo_pol/interpolation/atm_refraction.py:181-220."""
import numpy as np

from ..constants import global_constants as constants
from ..utilities.utilities import get_earth_radius


def _ref_4_3(range_vec, elevation_angle, coords_radar):
    theta = np.deg2rad(elevation_angle)
    KE = constants.KE
    RE = get_earth_radius(coords_radar[1])    # sic: index 1, in degrees (reference quirk Q1)
    alt = coords_radar[2]
    # evaluation order kept term by term: the float64 results are rounded to float32 and
    # feed bit-exact index decisions downstream
    h = np.sqrt(range_vec ** 2 + (KE * RE) ** 2 + 2 * range_vec * KE * RE * np.sin(theta)) - KE * RE + alt
    s = KE * RE * np.arcsin((range_vec * np.cos(theta)) / (KE * RE + h))
    e = theta + np.arctan(range_vec * np.cos(theta) / (range_vec * np.sin(theta) + KE * RE + alt))
    return s.astype('float32'), h.astype('float32'), np.rad2deg(e.astype('float32'))


def compute_trajectory_radial(range_vec, elevation_angle, coords_radar, refraction_method=1, N=0):
    """(s, h, e): ground distance [m], height [m], incidence elevation [deg] per gate."""
    if refraction_method == 2:
        raise NotImplementedError('ODE refraction (Zeng and Blahak 2014) is not part of this build (SURVEY.md 8f N3)')
    return _ref_4_3(range_vec, elevation_angle, coords_radar)
