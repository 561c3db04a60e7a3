# -*- coding: utf-8 -*-
"""Beam trajectories (host side, float64 then float32 exactly like the reference so that the
heights and elevation angles handed to the kernels are bit-identical):
4/3 effective earth radius model (cosmo_pol/interpolation/atm_refraction.py:181-220) and the ODE scheme of
Zeng and Blahak (2014) through the model's refractivity (:56-148)."""
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


def _profile_interpolator(x, y):
    """Linear interpolation inside the profile, linear extrapolation from the end segments outside
    (atm_refraction.py:151-179)."""
    from scipy.interpolate import interp1d
    inner = interp1d(x, y)
    xs, ys = inner.x, inner.y
    lo = (xs[0], ys[0], (ys[1] - ys[0]) / (xs[1] - xs[0]))
    hi = (xs[-1], ys[-1], (ys[-1] - ys[-2]) / (xs[-1] - xs[-2]))

    def value(v):
        if v < lo[0]:
            return lo[1] + (v - lo[0]) * lo[2]
        if v > hi[0]:
            return hi[1] + (v - hi[0]) * hi[2]
        return inner(v)
    return value


def _ref_ODE(range_vec, elevation_angle, coords_radar, N, radar_type='ground'):
    """Zeng and Blahak (2014): (h, sin e) integrated along the range through the refractive index
    profile n(h) = 1 + N 1e-6 of the model column at the radar (atm_refraction.py:56-148).  Host-side
    set-up code, once per elevation and vertical quadrature point; scipy's odeint / interp1d are the
    reference's own dependencies here.  The reference indexes the column with the latitude index on both
    horizontal axes (:113-117), keeps the profile in float32 and starts the arc distance at 0."""
    from scipy.integrate import odeint
    from ..utilities.utilities import WGS_to_COSMO
    proj = N.attributes['proj_info']
    rlat, rlon = WGS_to_COSMO((coords_radar[0], coords_radar[1]),
                              [proj['Latitude_of_southern_pole'], proj['Longitude_of_southern_pole']])
    res = N.attributes['resolution']
    row = int(np.round((rlat - float(proj['La1'])) / res[1]))
    n_profile = 1 + N.data[:, row, row] * 1E-6
    levels = N.attributes['z-levels'][:, row, row]
    if radar_type == 'ground':                      # model level 0 is the top
        levels, n_profile = levels[::-1], n_profile[::-1]
    RE = get_earth_radius(coords_radar[0])
    n_of_h = _profile_interpolator(levels, n_profile)
    dn_of_h = _profile_interpolator(levels[0:-1], np.diff(n_profile) / np.diff(levels))

    def rhs(state, _r):
        h, u = state
        curv = (1. / n_of_h(h)) * dn_of_h(h) + 1. / (RE + h)
        return [u, float(np.squeeze(-u ** 2 * curv + curv))]

    sol = odeint(rhs, [coords_radar[2], np.sin(np.deg2rad(elevation_angle))], range_vec)
    h, e = sol[:, 0], np.arcsin(sol[:, 1])
    dR = range_vec[1] - range_vec[0]
    s = np.zeros(h.shape)
    for i in range(1, len(s)):
        s[i] = s[i - 1] + RE * np.arcsin((np.cos(e[i - 1]) * dR) / (RE + h[i]))
    return s.astype('float32'), h.astype('float32'), np.rad2deg(e.astype('float32'))


def compute_trajectory_radial(range_vec, elevation_angle, coords_radar, refraction_method=1, N=0, radar_type='ground'):
    """(s, h, e): ground distance [m], height [m], incidence elevation [deg] per gate."""
    if refraction_method == 2:
        return _ref_ODE(range_vec, elevation_angle, coords_radar, N, radar_type)
    return _ref_4_3(range_vec, elevation_angle, coords_radar)
