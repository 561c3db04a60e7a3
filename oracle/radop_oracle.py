# -*- coding: utf-8 -*-
"""
radop_oracle.py -- CPU ORACLE.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import this module.  The product
(``cosmo_pol_b200``) never does; it fails loudly without its CUDA library.

This is a numpy restatement of the per-ray hot path of wolfidan/cosmo_pol
(``worker`` in cosmo_pol/radar_operator.py:407-427), function by function, with
the reference's own two C kernels used through ctypes (``oracle/_ref``: the
reference sources compiled unmodified; ``oracle/liboracle_port.so``: our C
restatement of them, used when ``_ref`` is absent and cross-checked against it).

Parity status: **pinned against the reference itself, run in the build container**
(tests/golden/make_golden_from_reference.py imports the reference's own modules from
/root/reference and runs get_interpolated_radial / get_radar_observables /
integrate_radials / cut_at_sensitivity, its hydrometeor classes, lookup_line,
_ref_4_3 ... on seeded synthetic inputs; the outputs are committed under
tests/golden/ and tests/test_oracle_golden.py holds this oracle to them: the
interpolation half bit-exact, the scattering half to float32 rounding).  The
reference ships no tests or golden vectors of its own.
What stays **unpinned** is third-party arithmetic absent from this container,
restated from its published algorithm and used as a stand-in in BOTH the golden
generation and this oracle:
  - pyproj.Geod(ellps='WGS84').fwd  (cosmo_pol/interpolation/interpolation.py:526-536,
    version unpinned in setup.py:53)  -> Vincenty's direct solution on WGS84
    (agrees with Karney's GeographicLib solution to < 0.1 mm below 500 km;
    checked in tests against a numerically integrated geodesic);
  - pycosmo.WGS_to_COSMO (interpolation.py:566-568, undeclared dependency)
    -> the standard rotated-pole rotation (COSMO documentation part I, 3.3);
  - pytmatrix (table CONTENTS): synthetic tables in the reference's own format.
The two native kernels (rows a4 / a16) are bit-exact against the reference C
compiled unmodified (oracle/_ref, tests/test_oracle_native.py).

No function in this file imports anything from ``cosmo_pol_b200``.
"""
import ctypes
import math
import os

import numpy as np
from numpy.polynomial.hermite import hermgauss
from scipy.special import gamma as _gamma

_HERE = os.path.dirname(os.path.abspath(__file__))

# --------------------------------------------------------------------------- #
# native kernels
# --------------------------------------------------------------------------- #
_F = ctypes.POINTER(ctypes.c_float)
_I = ctypes.POINTER(ctypes.c_int)


def _fp(a):
    return a.ctypes.data_as(_F)


class _Native(object):
    """ctypes access to oracle/_ref (reference C, unmodified) and the C port."""

    def __init__(self):
        self.port = None
        self.ref_interp = None
        self.ref_doppler = None
        p = os.path.join(_HERE, 'liboracle_port.so')
        if os.path.exists(p):
            self.port = ctypes.CDLL(p)
            self.port.oracle_interp_gates.restype = None
            self.port.oracle_spectrum_bins.restype = None
        p = os.path.join(_HERE, '_ref', 'libinterp_ref.so')
        if os.path.exists(p):
            self.ref_interp = ctypes.CDLL(p)
            self.ref_interp.get_all_radar_pts.restype = ctypes.c_void_p
        p = os.path.join(_HERE, '_ref', 'libdoppler_ref.so')
        if os.path.exists(p):
            self.ref_doppler = ctypes.CDLL(p)
            self.ref_doppler.get_refl.restype = ctypes.c_void_p


_native = None


def native():
    global _native
    if _native is None:
        _native = _Native()
    return _native


def interp_ref(coords, heights_gates, data, model_heights, llc, res):
    """The reference's get_all_radar_pts (interpolation_c.c:12-104), compiled
    unmodified into oracle/_ref, called with the argument order of the SWIG
    wrapper (interpolation.py:586-595)."""
    nat = native()
    if nat.ref_interp is None:
        raise RuntimeError('oracle/_ref/libinterp_ref.so missing (run make -C oracle)')
    coords = np.ascontiguousarray(coords, dtype=np.float32)
    hg = np.ascontiguousarray(heights_gates, dtype=np.float32)
    data = np.ascontiguousarray(data, dtype=np.float32)
    mh = np.ascontiguousarray(model_heights, dtype=np.float32)
    llc = np.ascontiguousarray(llc, dtype=np.float32)
    res = np.ascontiguousarray(res, dtype=np.float32)
    n = coords.shape[0]
    out = np.zeros(n, dtype=np.float32)
    nat.ref_interp.get_all_radar_pts(
        _fp(out), ctypes.c_int(n), _fp(coords), ctypes.c_int(n), ctypes.c_int(2),
        _fp(hg), ctypes.c_int(n),
        _fp(data), ctypes.c_int(data.shape[0]), ctypes.c_int(data.shape[1]), ctypes.c_int(data.shape[2]),
        _fp(mh), ctypes.c_int(mh.shape[0]), ctypes.c_int(mh.shape[1]), ctypes.c_int(mh.shape[2]),
        _fp(llc), ctypes.c_int(2), _fp(res), ctypes.c_int(2))
    return out


def interp_port(coords, heights_gates, data, model_heights, llc, res, debug=False):
    """Our C restatement (oracle/oracle_port.c) of the same kernel; also returns
    the integer decisions when debug=True."""
    nat = native()
    if nat.port is None:
        raise RuntimeError('oracle/liboracle_port.so missing (run make -C oracle)')
    coords = np.ascontiguousarray(coords, dtype=np.float32)
    hg = np.ascontiguousarray(heights_gates, dtype=np.float32)
    data = np.ascontiguousarray(data, dtype=np.float32)
    mh = np.ascontiguousarray(model_heights, dtype=np.float32)
    llc = np.ascontiguousarray(llc, dtype=np.float32)
    res = np.ascontiguousarray(res, dtype=np.float32)
    n = coords.shape[0]
    nz, ny, nx = data.shape
    out = np.zeros(n, dtype=np.float32)
    cell = np.zeros((n, 2), dtype=np.int32)
    level = np.zeros((n, 4), dtype=np.int32)
    status = np.zeros(n, dtype=np.int32)
    nat.port.oracle_interp_gates(
        ctypes.c_int(n), _fp(coords), _fp(hg), _fp(data), _fp(mh),
        ctypes.c_int(nz), ctypes.c_int(ny), ctypes.c_int(nx), _fp(llc), _fp(res),
        _fp(out), cell.ctypes.data_as(_I), level.ctypes.data_as(_I),
        status.ctypes.data_as(_I))
    if debug:
        return out, cell, level, status
    return out


def interp_gates(coords, heights_gates, data, model_heights, llc, res):
    """Reference C when available, else the port (they are bit-identical)."""
    if native().ref_interp is not None:
        return interp_ref(coords, heights_gates, data, model_heights, llc, res)
    return interp_port(coords, heights_gates, data, model_heights, llc, res)


def spectrum_bins_ref(Da, Db, rcs, D, N, step_D, D_min):
    """The reference's get_refl (doppler_c.c:11-32) compiled unmodified."""
    nat = native()
    Da = np.ascontiguousarray(Da, dtype=np.float32)
    Db = np.ascontiguousarray(Db, dtype=np.float32)
    rcs = np.ascontiguousarray(rcs, dtype=np.float32)
    D = np.ascontiguousarray(D, dtype=np.float32)
    N = np.ascontiguousarray(N, dtype=np.float32)
    step_D = np.ascontiguousarray(step_D, dtype=np.float32)
    D_min = np.ascontiguousarray(D_min, dtype=np.float32)
    nb, nh = Da.shape
    refl = np.zeros(nb, dtype=np.float32)
    c = ctypes.c_int
    nat.ref_doppler.get_refl(_fp(refl), c(nb), _fp(Da), c(nb), c(nh), _fp(Db), c(nb), c(nh),
                             _fp(rcs), c(rcs.shape[0]), c(nh), _fp(D), c(D.shape[0]), c(nh),
                             _fp(N), c(N.shape[0]), c(nh), _fp(step_D), c(nh), _fp(D_min), c(nh))
    return refl


def spectrum_bins_port(Da, Db, rcs, N, step_D, D_min):
    nat = native()
    Da = np.ascontiguousarray(Da, dtype=np.float32)
    Db = np.ascontiguousarray(Db, dtype=np.float32)
    rcs = np.ascontiguousarray(rcs, dtype=np.float32)
    N = np.ascontiguousarray(N, dtype=np.float32)
    step_D = np.ascontiguousarray(step_D, dtype=np.float32)
    D_min = np.ascontiguousarray(D_min, dtype=np.float32)
    nb, nh = Da.shape
    refl = np.zeros(nb, dtype=np.float32)
    nat.port.oracle_spectrum_bins(ctypes.c_int(nb), ctypes.c_int(nh), _fp(Da), _fp(Db), _fp(rcs),
                                  _fp(N), _fp(step_D), _fp(D_min), _fp(refl))
    return refl


def spectrum_bins(Da, Db, rcs, D, N, step_D, D_min):
    if native().ref_doppler is not None:
        return spectrum_bins_ref(Da, Db, rcs, D, N, step_D, D_min)
    return spectrum_bins_port(Da, Db, rcs, N, step_D, D_min)


# --------------------------------------------------------------------------- #
# configuration (the subset of cfg.CONFIG the hot path reads)
# --------------------------------------------------------------------------- #
C_LIGHT = 299792458.0   # global_constants.py:121
KE = 4. / 3.            # global_constants.py:126
T0 = 273.15             # global_constants.py:128
EPS = np.finfo(np.float64).eps  # global_constants.py:116
SIMULATED_VARIABLES = ['ZH', 'DSPECTRUM', 'RVEL', 'ZV', 'PHIDP', 'ZDR', 'RHOHV', 'KDP']


def default_config():
    """cfg.DEFAULTS (config/cfg.py:43-76) plus the mandatory radar position."""
    return {
        'radar': {'coords': [47.28436, 8.51201, 891.], 'type': 'ground', 'range': 150000,
                  'radial_resolution': 500, 'PRI': 700, 'FFT_length': 256,
                  'sensitivity': [-5, 10000], '3dB_beamwidth': 1., 'K_squared': 0.93,
                  'antenna_speed': 0.2, 'nyquist_velocity': None, 'frequency': 9.41},
        'refraction': {'scheme': 1},
        'integration': {'scheme': 1, 'nv_GH': 9, 'nh_GH': 3, 'weight_threshold': 1.},
        'doppler': {'scheme': 1, 'turbulence_correction': 0, 'motion_correction': 0},
        'microphysics': {'scheme': '1mom', 'with_melting': 0, 'with_ice_crystals': 1,
                         'with_attenuation': 1, 'scattering': 'tmatrix_masc'},
    }


def derived_constants(cfg):
    """global_constants.py:163-190."""
    r = cfg['radar']
    out = {}
    out['NVEL'] = C_LIGHT / (4 * 1E-6 * r['PRI'] * r['frequency'] * 1E09)
    out['VRES'] = 2 * out['NVEL'] / r['FFT_length']
    out['VARRAY'] = np.arange(-out['NVEL'], out['NVEL'] + out['VRES'], out['VRES'])
    out['WAVELENGTH'] = C_LIGHT / (r['frequency'] * 1E09) * 1000
    out['RANGE_RADAR'] = np.arange(r['radial_resolution'] / 2., r['range'], r['radial_resolution'])
    return out


# --------------------------------------------------------------------------- #
# a1: quadrature  (interpolation.py:148-166, 343-354)
# --------------------------------------------------------------------------- #
def gauss_hermite_subbeams(cfg):
    bw = cfg['radar']['3dB_beamwidth']
    nh = int(cfg['integration']['nh_GH'])
    nv = int(cfg['integration']['nv_GH'])
    sigma = bw / (2 * np.sqrt(2 * np.log(2)))
    pts_hor, w_hor = hermgauss(nh)
    pts_hor = pts_hor * sigma
    pts_ver, w_ver = hermgauss(nv)
    pts_ver = pts_ver * sigma
    weights = np.outer(w_hor * sigma, w_ver * sigma)
    weights *= np.abs(np.cos(np.deg2rad(pts_ver)))
    weights /= np.sum(weights.ravel())
    beam_broadening = nh > 1 or nv > 1
    # threshold, interpolation.py:343-352
    w_sort = np.sort(np.array(weights).ravel())[::-1]
    w_cum = np.cumsum(w_sort / np.sum(w_sort))
    w_cum[-1] = 1.
    idx_above = np.where(w_cum >= cfg['integration']['weight_threshold'])[0][0]
    threshold = w_sort[idx_above]
    keep = np.logical_or(weights >= threshold, not beam_broadening)
    return pts_hor, pts_ver, weights, keep


def multi_gaussian_points(cfg, azimuth, elevation):
    """Integration scheme 2 (interpolation.py:194-228): Gauss-Hermite in the radial and Gauss-Legendre in
    the angular direction around the beam axis for an antenna pattern given as a sum of Gaussians
    (antenna_params rows: amplitude [dB], mean, sigma [deg]).  An irregular list of (azimuth, elevation)
    points; every point has its own refracted trajectory (:440-494)."""
    from numpy.polynomial.legendre import leggauss
    nr_GH = int(cfg['integration']['nr_GH'])
    na_GL = int(cfg['integration']['na_GL'])
    ap = np.asarray(cfg['integration']['antenna_params'], dtype=np.float64)
    pts_ang, w_ang = leggauss(na_GL)
    pts_rad, w_rad = hermgauss(nr_GH)
    a_dB, mu, sigma = ap[:, 0], ap[:, 1], ap[:, 2]
    list_pts, weights, sum_weights = [], [], 0
    for i in range(nr_GH):
        for j in range(len(sigma)):
            for k in range(na_GL):
                r = mu[j] + np.sqrt(2) * sigma[j] * pts_rad[i]
                theta = np.pi * pts_ang[k] + np.pi
                weight = (np.pi * w_ang[k] * w_rad[i] * 10 ** (0.1 * a_dB[j]) * np.sqrt(2) * sigma[j] * abs(r))
                weight *= np.cos(r * np.sin(theta))
                weights.append(weight)
                sum_weights += weight
                list_pts.append([r * np.cos(theta) + azimuth, r * np.sin(theta) + elevation])
    weights = np.array(weights)
    weights /= sum_weights
    w_sort = np.sort(weights.ravel())[::-1]
    w_cum = np.cumsum(w_sort / np.sum(w_sort))
    w_cum[-1] = 1.
    threshold = w_sort[np.where(w_cum >= cfg['integration']['weight_threshold'])[0][0]]
    return list_pts, weights, weights >= threshold


# --------------------------------------------------------------------------- #
# a2: trajectory (atm_refraction.py:181-220, utilities.py:126-139)
# --------------------------------------------------------------------------- #
def earth_radius(latitude):
    a = 6378.1370 * 1000
    b = 6356.7523 * 1000
    num = ((a ** 2 * np.cos(latitude)) ** 2 + (b ** 2 * np.sin(latitude)) ** 2)
    den = ((a * np.cos(latitude)) ** 2 + (b * np.sin(latitude)) ** 2)
    return np.sqrt(num / den)


def ref_4_3(range_vec, elevation_angle, coords_radar):
    elevation_angle = np.deg2rad(elevation_angle)
    altitude_radar = coords_radar[2]
    latitude_radar = coords_radar[1]          # sic (quirk Q1): the longitude, in degrees
    RE = earth_radius(latitude_radar)
    temp = np.sqrt(range_vec ** 2 + (KE * RE) ** 2 + 2 * range_vec * KE * RE * np.sin(elevation_angle))
    h = temp - KE * RE + altitude_radar
    s = KE * RE * np.arcsin((range_vec * np.cos(elevation_angle)) / (KE * RE + h))
    e = elevation_angle + np.arctan(range_vec * np.cos(elevation_angle) /
                                    (range_vec * np.sin(elevation_angle) + KE * RE + altitude_radar))
    s = s.astype('float32')
    h = h.astype('float32')
    e = np.rad2deg(e.astype('float32'))
    return s, h, e


def ref_ode(range_vec, elevation_angle, coords_radar, N, radar_type='ground'):
    """atm_refraction.py:56-148 (Zeng and Blahak 2014): beam height and elevation from the second-order
    ODE through the refractive-index profile of the model column at the radar.  N: dict(data [nz,ny,nx]
    refractivity, z_levels, proj, resolution).  Third-party arithmetic of the reference on this path:
    scipy.integrate.odeint (LSODA) and scipy.interpolate.interp1d (scipy 1.18.1 in this image), called
    here exactly as the reference calls them.  Quirks kept: the latitude index of the radar is used for
    both horizontal axes (:113-117); the profile stays float32; the arc distance of the first gate is 0."""
    from scipy.integrate import odeint
    from scipy.interpolate import interp1d
    proj = N['proj']
    rot = wgs_to_rotated(coords_radar[0], coords_radar[1], proj['Latitude_of_southern_pole'],
                         proj['Longitude_of_southern_pole'])[0]
    llc = (float(proj['Lo1']), float(proj['La1']))
    res = N['resolution']
    pos = [(rot[0] - llc[1]) / res[1], (rot[1] - llc[0]) / res[0]]
    i = int(np.round(pos[0]))
    n_prof = 1 + (N['data'][:, i, i]) * 1E-6
    h = N['z_levels'][:, i, i]
    RE = earth_radius(coords_radar[0])
    if radar_type == 'ground':
        h = h[::-1]
        n_prof = n_prof[::-1]

    def piecewise(x, y):
        f = interp1d(x, y)
        xs, ys = f.x, f.y

        def at(v):
            if v < xs[0]:
                return ys[0] + (v - xs[0]) * (ys[1] - ys[0]) / (xs[1] - xs[0])
            if v > xs[-1]:
                return ys[-1] + (v - xs[-1]) * (ys[-1] - ys[-2]) / (xs[-1] - xs[-2])
            return f(v)
        return at

    n_of_h = piecewise(h, n_prof)
    dn_of_h = piecewise(h[0:-1], np.diff(n_prof) / np.diff(h))

    def rhs(z, r):
        hh, u = z
        k = (1. / n_of_h(hh)) * dn_of_h(hh) + 1. / (RE + hh)
        return [u, float(np.squeeze(-u ** 2 * k + k))]

    Z = odeint(rhs, [coords_radar[2], np.sin(np.deg2rad(elevation_angle))], range_vec)
    hh = Z[:, 0]
    e = np.arcsin(Z[:, 1])
    s = np.zeros(hh.shape)
    dR = range_vec[1] - range_vec[0]
    for k in range(1, len(s)):
        s[k] = s[k - 1] + RE * np.arcsin((np.cos(e[k - 1]) * dR) / (RE + hh[k]))
    return s.astype('float32'), hh.astype('float32'), np.rad2deg(e.astype('float32'))


# --------------------------------------------------------------------------- #
# a3: geolocation (third-party arithmetic restated)
# --------------------------------------------------------------------------- #
WGS84_A = 6378137.0
WGS84_F = 1 / 298.257223563
VINCENTY_ITERS = 5


def geod_fwd(lon0, lat0, azimuth, dist):
    """Direct geodesic on WGS84 (stand-in for pyproj.Geod.fwd,
    interpolation.py:526-536).  Vincenty (1975) direct formula with a fixed
    number of fixed-point iterations (converges by ~1e-3 per iteration for the
    < 500 km lines a radar produces).  All float64.  Returns lon, lat [deg]."""
    a = WGS84_A
    f = WGS84_F
    b = (1 - f) * a
    dist = np.asarray(dist, dtype=np.float64)
    phi1 = np.deg2rad(np.float64(lat0))
    alpha1 = np.deg2rad(np.float64(azimuth))
    sinA1 = np.sin(alpha1)
    cosA1 = np.cos(alpha1)
    tanU1 = (1 - f) * np.tan(phi1)
    cosU1 = 1.0 / np.sqrt(1 + tanU1 * tanU1)
    sinU1 = tanU1 * cosU1
    sigma1 = np.arctan2(tanU1, cosA1)
    sinAlpha = cosU1 * sinA1
    cosSqAlpha = 1 - sinAlpha * sinAlpha
    uSq = cosSqAlpha * (a * a - b * b) / (b * b)
    A = 1 + uSq / 16384 * (4096 + uSq * (-768 + uSq * (320 - 175 * uSq)))
    B = uSq / 1024 * (256 + uSq * (-128 + uSq * (74 - 47 * uSq)))
    sig0 = dist / (b * A)
    sigma = sig0
    for _ in range(VINCENTY_ITERS):
        cos2SigmaM = np.cos(2 * sigma1 + sigma)
        sinSigma = np.sin(sigma)
        cosSigma = np.cos(sigma)
        deltaSigma = B * sinSigma * (cos2SigmaM + B / 4 * (
            cosSigma * (-1 + 2 * cos2SigmaM * cos2SigmaM) -
            B / 6 * cos2SigmaM * (-3 + 4 * sinSigma * sinSigma) * (-3 + 4 * cos2SigmaM * cos2SigmaM)))
        sigma = sig0 + deltaSigma
    cos2SigmaM = np.cos(2 * sigma1 + sigma)
    sinSigma = np.sin(sigma)
    cosSigma = np.cos(sigma)
    tmp = sinU1 * sinSigma - cosU1 * cosSigma * cosA1
    phi2 = np.arctan2(sinU1 * cosSigma + cosU1 * sinSigma * cosA1,
                      (1 - f) * np.sqrt(sinAlpha * sinAlpha + tmp * tmp))
    lam = np.arctan2(sinSigma * sinA1, cosU1 * cosSigma - sinU1 * sinSigma * cosA1)
    C = f / 16 * cosSqAlpha * (4 + f * (4 - 3 * cosSqAlpha))
    L = lam - (1 - C) * f * sinAlpha * (sigma + C * sinSigma * (
        cos2SigmaM + C * cosSigma * (-1 + 2 * cos2SigmaM * cos2SigmaM)))
    lon2 = np.float64(lon0) + np.rad2deg(L)
    lat2 = np.rad2deg(phi2)
    return lon2, lat2


def geod_fwd_karney(lon0, lat0, azimuth, dist):
    """Direct geodesic on WGS84 by the algorithm pyproj.Geod.fwd runs (PROJ's geodesic.c = GeographicLib):
    C. F. F. Karney, "Algorithms for geodesics", J. Geodesy 87 (2013), sections 2-4, series truncated at order 6
    in the expansion parameter epsilon (eqs. 17, 18, 21, 24, 25).  The reference calls it at
    interpolation.py:526-536; pyproj is absent here and its version is unpinned (setup.py:53), so this is a
    restatement of the PUBLISHED algorithm, not of PROJ's source, used to measure how far the Vincenty stand-in
    (geod_fwd) is from it.  All float64, vectorised with numpy broadcasting.  Returns lon, lat [deg]."""
    a, f = WGS84_A, WGS84_F
    b = (1 - f) * a
    n = f / (2 - f)
    ep2 = f * (2 - f) / ((1 - f) ** 2)                  # second eccentricity squared
    dist = np.asarray(dist, dtype=np.float64)
    phi1 = np.deg2rad(np.float64(lat0))
    alp1 = np.deg2rad(np.asarray(azimuth, dtype=np.float64))
    salp1, calp1 = np.sin(alp1), np.cos(alp1)
    # reduced latitude
    sbet1, cbet1 = (1 - f) * np.sin(phi1), np.cos(phi1)
    nrm = np.hypot(sbet1, cbet1)
    sbet1, cbet1 = sbet1 / nrm, cbet1 / nrm
    # azimuth at the equator crossing (eq. 10), arc length and longitude on the auxiliary sphere at point 1 (11, 12)
    salp0 = salp1 * cbet1
    calp0 = np.hypot(calp1, salp1 * sbet1)
    ssig1, csig1 = sbet1, calp1 * cbet1
    somg1, comg1 = salp0 * sbet1, calp1 * cbet1
    nrm = np.hypot(ssig1, csig1)
    ssig1, csig1 = ssig1 / nrm, csig1 / nrm
    sig1 = np.arctan2(ssig1, csig1)
    k2 = ep2 * calp0 ** 2
    eps = k2 / (2 * (1 + np.sqrt(1 + k2)) + k2)         # eq. 16
    e2_ = eps * eps
    A1 = (1 + e2_ * (1. / 4 + e2_ * (1. / 64 + e2_ / 256))) / (1 - eps)          # eq. 17
    C1 = [eps * (-1. / 2 + e2_ * (3. / 16 - e2_ / 32)),                           # eq. 18
          e2_ * (-1. / 16 + e2_ * (1. / 32 - e2_ * 9. / 2048)),
          eps * e2_ * (-1. / 48 + e2_ * 3. / 256),
          e2_ * e2_ * (-5. / 512 + e2_ * 3. / 512),
          eps * e2_ * e2_ * (-7. / 1280),
          e2_ * e2_ * e2_ * (-7. / 2048)]
    C1p = [eps * (1. / 2 + e2_ * (-9. / 32 + e2_ * 205. / 1536)),                 # eq. 21 (reverted series)
           e2_ * (5. / 16 + e2_ * (-37. / 96 + e2_ * 1335. / 4096)),
           eps * e2_ * (29. / 96 - e2_ * 75. / 128),
           e2_ * e2_ * (539. / 1536 - e2_ * 2391. / 2560),
           eps * e2_ * e2_ * (3467. / 7680),
           e2_ * e2_ * e2_ * (38081. / 61440)]
    A3 = 1 - eps * ((1 - n) / 2 + eps * ((1. / 4 + n / 8 - 3 * n * n / 8) + eps * ((1. / 16 + 3 * n / 16 + n * n / 16) +
               eps * ((3. / 64 + n / 32) + eps * 3. / 128))))                     # eq. 24
    C3 = [eps * ((1 - n) / 4 + eps * ((1 - n * n) / 8 + eps * ((3. / 64 + 3 * n / 64 - n * n / 64) +
                 eps * ((5. / 128 + n / 64) + eps * 3. / 128)))),                 # eq. 25
          e2_ * ((1. / 16 - 3 * n / 32 + n * n / 32) + eps * ((3. / 64 - n / 32 - 3 * n * n / 64) +
                 eps * ((3. / 128 + n / 128) + eps * 5. / 256))),
          eps * e2_ * ((5. / 192 - 3 * n / 64 + 5 * n * n / 192) + eps * ((3. / 128 - 5 * n / 192) + eps * 7. / 512)),
          e2_ * e2_ * ((7. / 512 - 7 * n / 256) + eps * 7. / 512),
          eps * e2_ * e2_ * (21. / 2560)]

    def fourier(coef, sig):
        return sum(c * np.sin(2 * (l + 1) * sig) for l, c in enumerate(coef))

    B11 = fourier(C1, sig1)
    tau1 = sig1 + B11                                   # I1(sigma1) / A1
    tau2 = tau1 + dist / (b * A1)                       # eq. 7: s / b = I1(sigma)
    sig2 = tau2 + fourier(C1p, tau2)                    # eq. 20
    ssig2, csig2 = np.sin(sig2), np.cos(sig2)
    sbet2 = calp0 * ssig2
    cbet2 = np.hypot(salp0, calp0 * csig2)
    somg2, comg2 = salp0 * ssig2, csig2
    # longitude difference on the auxiliary sphere, then eq. 8 with I3 (eq. 23)
    omg12 = np.arctan2(somg2 * comg1 - comg2 * somg1, comg2 * comg1 + somg2 * somg1)
    lam12 = omg12 - f * salp0 * A3 * ((sig2 - sig1) + (fourier(C3, sig2) - fourier(C3, sig1)))
    lat2 = np.rad2deg(np.arctan2(sbet2, (1 - f) * cbet2))
    lon2 = np.float64(lon0) + np.rad2deg(lam12)
    return lon2, lat2


def wgs_to_rotated(lat, lon, sp_lat, sp_lon):
    """Geographic -> rotated-pole coordinates given the rotated SOUTH pole
    (stand-in for pycosmo.WGS_to_COSMO, interpolation.py:566-568).  Rotation about
    z by the pole longitude then about y by 90deg + pole latitude.
    Returns array [n,2] = (rlat, rlon) in degrees, float64."""
    lon = np.deg2rad(np.asarray(lon, dtype=np.float64))
    lat = np.deg2rad(np.asarray(lat, dtype=np.float64))
    theta = np.deg2rad(90. + sp_lat)
    phi = np.deg2rad(sp_lon)
    x = np.cos(lon) * np.cos(lat)
    y = np.sin(lon) * np.cos(lat)
    z = np.sin(lat)
    x_new = np.cos(theta) * np.cos(phi) * x + np.cos(theta) * np.sin(phi) * y + np.sin(theta) * z
    y_new = -np.sin(phi) * x + np.cos(phi) * y
    z_new = -np.sin(theta) * np.cos(phi) * x - np.sin(theta) * np.sin(phi) * y + np.cos(theta) * z
    rlon = np.rad2deg(np.arctan2(y_new, x_new))
    rlat = np.rad2deg(np.arcsin(z_new))
    return np.vstack((rlat, rlon)).T


# --------------------------------------------------------------------------- #
# a3-a5: sub-beam interpolation (interpolation.py:381-438, 498-597)
# --------------------------------------------------------------------------- #
class OutOfDomain(IndexError):
    pass


def trilin_interp_radial(model, names, azimuth, s, h, cfg, debug=False):
    radar_pos = cfg['radar']['coords']
    lons, lats = geod_fwd(radar_pos[1], radar_pos[0], azimuth, s)
    proj = model['proj']
    llc = np.asarray((float(proj['Lo1']), float(proj['La1']))).astype('float32')
    urc = np.asarray((float(proj['Lo2']), float(proj['La2']))).astype('float32')
    res = model['resolution']
    coords = wgs_to_rotated(lats, lons, proj['Latitude_of_southern_pole'],
                            proj['Longitude_of_southern_pole'])
    if (np.any(coords[:, 1] < llc[0]) or np.any(coords[:, 0] < llc[1]) or
            np.any(coords[:, 1] > urc[0]) or np.any(coords[:, 0] > urc[1])):
        raise OutOfDomain('radar domain is not entirely contained in the model domain')
    out = []
    for n in names:
        out.append(interp_gates(coords, h, model['vars'][n], model['z_levels'], llc, res))
    if debug:
        dbg = interp_port(coords, h, model['vars'][names[0]], model['z_levels'], llc, res, debug=True)
        return lats, lons, out, coords, dbg
    return lats, lons, out


def melting(sub):
    """melting.py:19-90."""
    v = sub['values']
    T = v['T']
    QR, QS, QG = v['QR_v'], v['QS_v'], v['QG_v']
    QS_QG = QS + QG
    mask_ml = np.logical_and(QR > 0, QS_QG > 0)
    n = len(T)
    if not np.any(mask_ml):
        has = False
        for k in ('QmS_v', 'QmG_v', 'fwet_mS', 'fwet_mG'):
            v[k] = np.zeros(n)
    else:
        has = True
        QRm, QSm, QGm = QR[mask_ml], QS[mask_ml], QG[mask_ml]
        QSG = QSm + QGm
        v['QmS_v'] = np.zeros(n)
        v['QmS_v'][mask_ml] = QSm + QRm * (QSm / QSG)
        v['QmG_v'] = np.zeros(n)
        v['QmG_v'][mask_ml] = QGm + QRm * (QGm / QSG)
        mwm = np.logical_or(v['QmS_v'] > 0, v['QmG_v'] > 0)
        v['QS_v'][mwm] = 0
        v['QG_v'][mwm] = 0
        v['QR_v'][mwm] = 0
        with np.errstate(invalid='ignore', divide='ignore'):      # 0 / 0 where only one of QS, QG is present
            v['fwet_mS'] = np.zeros(n)
            v['fwet_mS'][mask_ml] = (QRm * QSm / QSG) / v['QmS_v'][mask_ml]
            v['fwet_mG'] = np.zeros(n)
            v['fwet_mG'][mask_ml] = (QRm * QGm / QSG) / v['QmG_v'][mask_ml]
    sub['has_melting'] = has
    sub['mask_ml'] = mask_ml
    return sub


def _trajectory(cfg, model, rranges, elevation, radar_pos):
    if cfg['refraction']['scheme'] == 2:
        N = {'data': model['refractivity'], 'z_levels': model['z_levels'], 'proj': model['proj'],
             'resolution': model['resolution']}
        return ref_ode(rranges, elevation, radar_pos, N, cfg['radar'].get('type', 'ground'))
    return ref_4_3(rranges, elevation, radar_pos)


def _irregular_radial(model, names, azimuth, elevation, cfg, rranges, radar_pos, debug):
    """interpolation.py:440-494: one trajectory per quadrature point."""
    list_pts, weights, keep = multi_gaussian_points(cfg, azimuth, elevation)
    subs = []
    for i, pt in enumerate(list_pts):
        if not keep[i]:
            continue
        s, h, e = _trajectory(cfg, model, rranges, pt[1], radar_pos)
        r = trilin_interp_radial(model, names, pt[0], s, h, cfg, debug=debug)
        values = {}
        for k, bi in enumerate(r[2]):
            if k == 0:
                mask = np.zeros(len(bi))
                mask[bi == -9999] = 1
                mask[np.isnan(bi)] = -1
            bi[mask != 0] = np.nan
            values[names[k]] = bi
        sub = {'values': values, 'mask': mask, 'lats': r[0], 'lons': r[1], 'dist': s, 'heights': h, 'elev': e,
               'quad_pt': list(pt), 'quad_weight': weights[i], 'has_melting': False, 'ij': (i, i)}
        if debug:
            sub['coords'] = r[3]
            sub['dbg'] = r[4]
        if cfg['microphysics']['with_melting']:
            sub = melting(sub)
        subs.append(sub)
    return subs


def _threshold_keep(weights, cfg, beam_broadening):
    # interpolation.py:343-352
    w_sort = np.sort(np.array(weights).ravel())[::-1]
    w_cum = np.cumsum(w_sort / np.sum(w_sort))
    w_cum[-1] = 1.
    threshold = w_sort[np.where(w_cum >= cfg['integration']['weight_threshold'])[0][0]]
    return np.logical_or(weights >= threshold, not beam_broadening)


def _polar2cartesian(r, t, grid, x, y, order=3):
    """utilities.py:285-316"""
    from scipy.interpolate import interp1d
    from scipy.ndimage import map_coordinates
    X, Y = np.meshgrid(x, y)
    new_r = np.sqrt(X * X + Y * Y)
    new_t = np.arctan2(X, Y) + np.pi
    new_ir = interp1d(r, np.arange(len(r)), bounds_error=False)(new_r.ravel())
    new_it = interp1d(t, np.arange(len(t)))(new_t.ravel())
    new_ir[new_r.ravel() > r.max()] = len(r) - 1
    new_ir[new_r.ravel() < r.min()] = 0
    return map_coordinates(grid, np.array([new_it, new_ir]), order=order).reshape(new_r.shape)


def _vector_1d_to_polar(angles, values, x, y):
    """utilities.py:318-341"""
    midpt = int(np.floor(len(angles) / 2.))
    r = angles[midpt:]
    thet = [0, np.pi, 2 * np.pi]
    pol = np.zeros((len(thet), len(r)))
    pol[0, :] = values[midpt:]
    pol[1, :] = values[0:midpt + 1][::-1]
    pol[2, :] = pol[0, :]
    return _polar2cartesian(r, thet, pol, x, y)


def gauss_legendre_antenna_subbeams(cfg):
    """Integration scheme 3, interpolation.py:232-258."""
    from numpy.polynomial.legendre import leggauss
    nh, nv = int(cfg['integration']['nh_GH']), int(cfg['integration']['nv_GH'])
    antenna = np.genfromtxt(cfg['integration']['antenna_diagram'], delimiter=',')
    angles = antenna[:, 0]
    power_sq = (10 ** (0.1 * antenna[:, 1])) ** 2
    bounds = np.max(angles)
    pts_hor, w_hor = leggauss(nh)
    pts_hor = pts_hor * bounds
    pts_ver, w_ver = leggauss(nv)
    pts_ver = pts_ver * bounds
    weights = _vector_1d_to_polar(angles, power_sq, pts_hor, pts_ver).T * np.outer(w_hor, w_ver)
    weights *= np.abs(np.cos(np.deg2rad(pts_ver)))
    weights *= 2 * bounds
    weights /= np.sum(weights.ravel())
    return pts_hor, pts_ver, weights, _threshold_keep(weights, cfg, nh > 1 or nv > 1)


def _discrete_gautschi(z, w, n_iter):
    """interpolation/quadrature.py:33-54"""
    p = np.ones(z.shape)
    p /= np.sqrt(np.dot(p, p))
    p_prev = np.zeros(z.shape)
    wz = z * w
    a, b = np.empty(n_iter), np.empty(n_iter)
    for j in range(n_iter):
        p_norm = np.dot(w * p, p)
        a[j] = np.dot(wz * p, p) / p_norm
        b[j] = 0.0 if j == 0 else p_norm / np.dot(w * p_prev, p_prev)
        p, p_prev = (z - a[j]) * p - b[j] * p_prev, p
    return a, b[1:]


def _gautschi_points_and_weights(w_func, left, right, num_points, n=8192):
    """interpolation/quadrature.py:57-101"""
    import scipy.linalg
    dx = (float(right) - left) / n
    z = np.hstack(np.linspace(left + 0.5 * dx, right - 0.5 * dx, n))
    w = dx * w_func(z)
    a, b = _discrete_gautschi(z, w, num_points)
    beta = np.sqrt(b)
    J = np.diag(a)
    J += np.diag(beta, k=-1)
    J += np.diag(beta, k=1)
    J[np.isnan(J)] = 0.
    J[np.isinf(J)] = 0.
    points, v = scipy.linalg.eigh(J)
    ind = points.argsort()
    return points[ind], (v[0, :] ** 2 * w.sum())[ind]


def gautschi_antenna_subbeams(cfg):
    """Integration scheme 5, interpolation.py:276-303."""
    from scipy.interpolate import interp1d
    nh, nv = int(cfg['integration']['nh_GH']), int(cfg['integration']['nv_GH'])
    antenna = np.genfromtxt(cfg['integration']['antenna_diagram'], delimiter=',')
    angles = antenna[:, 0]
    power_sq = (10 ** (0.1 * antenna[:, 1])) ** 2
    bounds = np.max(angles)
    fit = interp1d(angles, power_sq, fill_value=0)
    fit_w = interp1d(angles, power_sq * np.cos(np.abs(np.deg2rad(angles))), fill_value=0)
    pts_hor, w_hor = _gautschi_points_and_weights(fit, -bounds, bounds, nh)
    pts_ver, w_ver = _gautschi_points_and_weights(fit_w, -bounds, bounds, nv)
    weights = np.outer(w_hor, w_ver)
    weights /= np.sum(weights.ravel())
    return pts_hor, pts_ver, weights, _threshold_keep(weights, cfg, nh > 1 or nv > 1)


def get_interpolated_radial(model, azimuth, elevation, cfg, debug=False):
    """interpolation.py:91-438 for integration scheme 1, refraction schemes 1 and 2 (the refractivity field is
    model['refractivity'], kept apart from the interpolated variables like the reference's self.N).
    A sub-radial is a dict with the attributes of the reference's Radial
    (radial.py:16-54)."""
    names = list(model['vars'].keys())
    rranges = derived_constants(cfg)['RANGE_RADAR']
    radar_pos = cfg['radar']['coords']
    if cfg['integration']['scheme'] == 2:
        return _irregular_radial(model, names, azimuth, elevation, cfg, rranges, radar_pos, debug)
    ml = cfg['integration']['scheme'] == 'ml'
    if cfg['integration']['scheme'] == 3:
        pts_hor, pts_ver, weights, keep = gauss_legendre_antenna_subbeams(cfg)
    elif cfg['integration']['scheme'] == 5:
        pts_hor, pts_ver, weights, keep = gautschi_antenna_subbeams(cfg)
    else:
        pts_hor, pts_ver, weights, keep = melting_layer_subbeams(cfg) if ml else gauss_hermite_subbeams(cfg)
    if cfg['refraction']['scheme'] == 2:
        N = {'data': model['refractivity'], 'z_levels': model['z_levels'], 'proj': model['proj'],
             'resolution': model['resolution']}
        list_refraction = [ref_ode(rranges, pt + elevation, radar_pos, N, cfg['radar'].get('type', 'ground'))
                           for pt in pts_ver]
    else:
        list_refraction = [ref_4_3(rranges, pt + elevation, radar_pos) for pt in pts_ver]
    subs = []
    for i in range(len(pts_hor)):
        for j in range(len(pts_ver)):
            if not keep[i, j]:
                continue
            s, h, e = list_refraction[j]
            r = trilin_interp_radial(model, names, pts_hor[i] + azimuth, s, h, cfg, debug=debug)
            lats, lons, list_vars = r[0], r[1], r[2]
            values = {}
            for k, bi in enumerate(list_vars):
                if k == 0:
                    mask = np.zeros(len(bi))
                    mask[bi == -9999] = 1
                    mask[np.isnan(bi)] = -1
                bi[mask != 0] = np.nan
                values[names[k]] = bi
            sub = {'values': values, 'mask': mask, 'lats': lats, 'lons': lons,
                   'dist': s, 'heights': h, 'elev': e,     # e is shared between i's (quirk Q3)
                   'quad_pt': [pts_hor[i] + azimuth, pts_ver[j] + elevation],
                   'quad_weight': weights[i, j], 'has_melting': False, 'ij': (i, j)}
            if debug:
                sub['coords'] = r[3]
                sub['dbg'] = r[4]
            if cfg['microphysics']['with_melting']:
                sub = melting(sub)
            if ml:
                # interpolation.py:423-436 (sub['mask_ml'] only exists with the melting scheme: KeyError here,
                # AttributeError in the reference)
                if j > int(cfg['integration']['nv_GH']):
                    smoothing = np.zeros(len(lats))
                    idx_ml = np.where(sub['mask_ml'])[0]
                    if len(idx_ml):
                        smoothing[idx_ml[0]] = 1
                        smoothing[idx_ml[-1]] = 1
                        from scipy.ndimage import gaussian_filter
                        smoothing = gaussian_filter(smoothing, 2)
                    sub['quad_weight'] = sub['quad_weight'] * smoothing
                else:
                    sub['quad_weight'] = sub['quad_weight'] * np.ones(len(lats))
            subs.append(sub)
    return subs


def sweep_indices(model, luts, cfg, azimuths, elevation, geod=None):
    """The integer decisions of every sub-beam gate of one sweep (all `azimuths` at one `elevation`), vectorised
    over the rays: what get_interpolated_radial (interpolation.py:381-438, 498-597; interpolation_c.c:42-103) and
    lookup_line (lut.py:336-342) decide for integration scheme 1, refraction scheme 1, without the melting scheme.
    Same arithmetic as get_interpolated_radial / get_radar_observables above, element for element (numpy's
    elementwise kernels do not depend on the array shape), so that a whole 7200-ray volume can be checked in
    minutes.  Returns dict of arrays [n_rays, n_sub, n_gates(, ...)]: cell [.., 2], level [.., 4], status,
    lut_e / lut_t [.., n_species] (-1 where the species is inactive), rot [.., 2] float64 rotated coordinates.
    geod: alternative direct-geodesic function (lon0, lat0, azimuth, dist) -> (lon, lat), default geod_fwd."""
    geod = geod_fwd if geod is None else geod
    names = list(model['vars'].keys())
    rranges = derived_constants(cfg)['RANGE_RADAR']
    radar_pos = cfg['radar']['coords']
    pts_hor, pts_ver, weights, keep = gauss_hermite_subbeams(cfg)
    refr = [ref_4_3(rranges, pt + elevation, radar_pos) for pt in pts_ver]
    species = species_list(cfg)
    proj = model['proj']
    llc = np.asarray((float(proj['Lo1']), float(proj['La1']))).astype('float32')
    urc = np.asarray((float(proj['Lo2']), float(proj['La2']))).astype('float32')
    res = model['resolution']
    azimuths = np.asarray(azimuths, dtype=np.float64)
    nr, ng = len(azimuths), len(rranges)
    subs = [(i, j) for i in range(len(pts_hor)) for j in range(len(pts_ver)) if keep[i, j]]
    out = {'cell': np.zeros((nr, len(subs), ng, 2), np.int32), 'level': np.zeros((nr, len(subs), ng, 4), np.int32),
           'status': np.zeros((nr, len(subs), ng), np.int32), 'rot': np.zeros((nr, len(subs), ng, 2), np.float64),
           'lut_e': np.full((nr, len(subs), ng, len(species)), -1, np.int32),
           'lut_t': np.full((nr, len(subs), ng, len(species)), -1, np.int32)}
    for q, (i, j) in enumerate(subs):
        s, h, e = refr[j]
        lons, lats = geod(radar_pos[1], radar_pos[0], (pts_hor[i] + azimuths)[:, None], np.asarray(s, np.float64)[None, :])
        coords = wgs_to_rotated(lats.ravel(), lons.ravel(), proj['Latitude_of_southern_pole'],
                                proj['Longitude_of_southern_pole'])
        if (np.any(coords[:, 1] < llc[0]) or np.any(coords[:, 0] < llc[1]) or
                np.any(coords[:, 1] > urc[0]) or np.any(coords[:, 0] > urc[1])):
            raise OutOfDomain('radar domain is not entirely contained in the model domain')
        hh = np.tile(h, nr)
        first, cell, level, status = interp_port(coords, hh, model['vars'][names[0]], model['z_levels'], llc, res, debug=True)
        mask = np.zeros(len(first))
        mask[first == -9999] = 1
        mask[np.isnan(first)] = -1
        out['cell'][:, q] = cell.reshape(nr, ng, 2)
        out['level'][:, q] = level.reshape(nr, ng, 4)
        out['status'][:, q] = status.reshape(nr, ng)
        out['rot'][:, q] = coords.reshape(nr, ng, 2)
        ef = np.array(e, copy=True)
        ef[ef > 90] = 180 - ef[ef > 90]
        ef[ef < 0] = -ef[ef < 0]
        ef = np.tile(ef, nr)
        T = interp_gates(coords, hh, model['vars']['T'], model['z_levels'], llc, res)
        T[mask != 0] = np.nan
        for k, hname in enumerate(species):
            QM = interp_gates(coords, hh, model['vars']['Q' + hname + '_v'], model['z_levels'], llc, res)
            QM[mask != 0] = np.nan
            with np.errstate(invalid='ignore'):
                valid = QM > 0
            lut = luts[hname]
            v = lut['value_table']
            ie = lut_index(ef[valid], lut['axes_limits'][0][0], lut['axes_step'][0], v.shape[0])
            it = lut_index(T[valid], lut['axes_limits'][1][0], lut['axes_step'][1], v.shape[1])
            le = np.full(nr * ng, -1, np.int32)
            lt = np.full(nr * ng, -1, np.int32)
            le[valid] = ie
            lt[valid] = it
            out['lut_e'][:, q, :, k] = le.reshape(nr, ng)
            out['lut_t'][:, q, :, k] = lt.reshape(nr, ng)
    return out


def melting_layer_subbeams(cfg):
    """Integration scheme 'ml' (interpolation.py:168-193): ten times the vertical Gauss-Hermite points (made odd),
    sorted by decreasing weight; threshold as for scheme 1 (:343-352)."""
    bw = cfg['radar']['3dB_beamwidth']
    nh = int(cfg['integration']['nh_GH'])
    nv = int(cfg['integration']['nv_GH'])
    nv_ml = 10 * nv
    if not nv_ml % 2:
        nv_ml += 1
    sigma = bw / (2 * np.sqrt(2 * np.log(2)))
    pts_hor, w_hor = hermgauss(nh)
    pts_hor = pts_hor * sigma
    pts_ver, w_ver = hermgauss(nv_ml)
    pts_ver = pts_ver * sigma
    idx_sort = np.argsort(w_ver)[::-1]
    w_ver = w_ver[idx_sort]
    pts_ver = pts_ver[idx_sort]
    weights = np.outer(w_hor * sigma, w_ver * sigma)
    weights *= np.abs(np.cos(np.deg2rad(pts_ver)))
    weights /= np.sum(weights.ravel())
    beam_broadening = nh > 1 or nv > 1
    w_sort = np.sort(np.array(weights).ravel())[::-1]
    w_cum = np.cumsum(w_sort / np.sum(w_sort))
    w_cum[-1] = 1.
    threshold = w_sort[np.where(w_cum >= cfg['integration']['weight_threshold'])[0][0]]
    keep = np.logical_or(weights >= threshold, not beam_broadening)
    return pts_hor, pts_ver, weights, keep


def _nansum2(x, y):
    """utilities.nansum_arr (utilities.py:231-261) for equal shapes or y shorter
    (zero padded at the end)."""
    x = np.array(x)
    y = np.array(y)
    if x.shape != y.shape:
        pad_x = []
        pad_y = []
        for d in (np.array(x.shape) - np.array(y.shape)):
            if d < 0:
                pad_x.append((0, -d)); pad_y.append((0, 0))
            else:
                pad_y.append((0, d)); pad_x.append((0, 0))
        x = np.pad(x, pad_x, 'constant', constant_values=0)
        y = np.pad(y, pad_y, 'constant', constant_values=0)
    return np.nansum([x, y], axis=0)


def integrate_radials(subs):
    """interpolation.py:36-89."""
    n = len(subs)
    out = {}
    sum_w = 0
    for s in subs:
        sum_w += s['quad_weight']
    for k in subs[0]['values'].keys():
        acc = np.array([np.nan])
        for s in subs:
            acc = _nansum2(acc, s['values'][k] * s['quad_weight'] / sum_w)
        out[k] = acc
    mask = np.zeros(len(subs[0]['mask']))
    for s in subs:
        mask = mask + s['mask']
    mask /= float(n)
    mask[np.logical_and(mask > -1, mask <= 0)] = 0
    return out, mask


# --------------------------------------------------------------------------- #
# a7-a9, a12: hydrometeors (hydrometeors.py, constants_1mom.py, constants_2mom.py)
# --------------------------------------------------------------------------- #
RHO_W = 1000. / (1000 ** 3)


def _c1():
    """constants_1mom.py:45-104."""
    c = {}
    g = {'N0': 4 * 1E3, 'b': 3.1, 'beta': 0.89, 'mu': 0.0}
    g['a'] = 169.6 * (1000 ** -g['b'])
    g['alpha'] = 442.0 * (1000 ** -g['beta'])
    g['lambda_factor'] = g['a'] * g['N0'] * _gamma(g['b'] + 1)
    g['vel_factor'] = _gamma(g['mu'] + g['beta'] + 1)
    g['ntot_factor'] = _gamma(g['mu'] + 1)
    c['G'] = g
    s = {'N0': None, 'b': 2., 'beta': 0.25, 'mu': 0.0}
    s['a'] = 0.038 * (1000 ** -s['b'])
    s['alpha'] = 4.9 * (1000 ** -s['beta'])
    s['lambda_factor'] = _gamma(s['b'] + 1)
    s['vel_factor'] = _gamma(s['mu'] + s['beta'] + 1)
    s['ntot_factor'] = _gamma(s['mu'] + 1)
    c['S'] = s
    r = {'mu': 0.5, 'b': 3., 'beta': 0.5}
    n00 = 8E6 / (1000 ** (1 + r['mu'])) * (0.01) ** (-r['mu'])
    r['N0'] = 0.1 * n00 * np.exp(3.2 * r['mu'])
    r['a'] = np.pi / 6. * RHO_W
    r['alpha'] = 130 * (1000 ** -r['beta'])
    r['lambda_factor'] = r['a'] * r['N0'] * _gamma(1. + r['b'] + r['mu'])
    r['vel_factor'] = _gamma(r['mu'] + r['beta'] + 1)
    r['ntot_factor'] = _gamma(r['mu'] + 1)
    c['R'] = r
    i = {'N0': None, 'b': 3, 'mu': 0.0, 'alpha': 0.9655930341942476, 'beta': 1.2019867549668874}
    i['a'] = 130 * (1000 ** -i['b'])
    i['lambda_factor'] = _gamma(i['b'] + 1)
    i['ntot_factor'] = _gamma(i['mu'] + 1)
    i['vel_factor'] = _gamma(i['mu'] + i['beta'] + 1)
    c['I'] = i
    for k in c:
        c[k]['nu'] = 1.0
    return c


def phi_23_I(x):
    """constants_1mom.py:103-104."""
    return 490.6 * np.exp(-20.78 * x) + 17.46 * x ** (0.6357) * np.exp(-3.290 * x)


def _c2():
    """constants_2mom.py:16-177."""
    raw = {
        'G': dict(am=0.142, bm=0.314, av=86.89371, bv=0.268325, nu=1. / 3., mu=1.0, x_min=1E-09, x_max=5E-04),
        'S': dict(am=2.4, bm=0.455, av=4.2, bv=0.092, nu=0.5, mu=0.0, x_min=1E-10, x_max=2E-05),
        'R': dict(am=0.124, bm=1. / 3., av=114.0137, bv=0.23437, nu=1. / 3., mu=0.0, x_min=2.6E-10, x_max=3E-06),
        'H': dict(am=0.1366, bm=1. / 3., av=39.3, bv=1. / 6., nu=1. / 3., mu=1.0, x_min=2.6E-09, x_max=5E-04),
        'I': dict(am=0.124, bm=0.302, av=317, bv=0.363, nu=1. / 3., mu=0.0, x_min=1E-12, x_max=1E-6),
    }
    c = {}
    for k, r in raw.items():
        d = {}
        b = 1. / r['bm']
        a = r['am'] ** (-1 / r['bm'])
        beta = r['bv'] / r['bm']
        alpha = r['av'] * r['am'] ** (-r['bv'] / r['bm'])
        nu = r['nu'] / r['bm']
        mu = (r['mu'] + 1) / r['bm'] - 1
        d['lambda_factor'] = 1. / a * _gamma((mu + 1) / nu) / _gamma((mu + b + 1) / nu)
        d['a'] = a * 1000 ** (-b)
        d['alpha'] = alpha * 1000 ** (-beta)
        d['b'], d['beta'], d['nu'], d['mu'] = b, beta, nu, mu
        d['vel_factor'] = _gamma((mu + beta + 1) / nu)
        d['ntot_factor'] = _gamma((mu + 1) / nu)
        d['x_min'], d['x_max'] = r['x_min'], r['x_max']
        d['N0'] = None
        c[k] = d
    return c


C1 = _c1()
C2 = _c2()


class Hydro(object):
    """PSD of one species: hydrometeors.py `_Hydrometeor` (:81-255), Rain (:747-779),
    Snow (:879-914), Graupel (:1025-1056), Hail (:1124-1160), IceParticle (:1181-1373)."""

    def __init__(self, kind, scheme):
        self.kind = kind
        self.scheme = scheme
        c = (C1 if scheme == '1mom' else C2)[kind]
        self.c = c
        self.a, self.b = c['a'], c['b']
        self.alpha, self.beta = c['alpha'], c['beta']
        self.mu, self.nu = c['mu'], c['nu']
        self.N0 = c['N0']
        self.lambda_ = None
        self.ntot = None
        self.nbins_D = 1024
        self.d_min = None
        self.d_max = None
        self.dtype = np.float64

    # --- set_psd -------------------------------------------------------------
    def set_psd(self, *args):
        f = self.dtype
        c = self.c
        if self.scheme == '2mom':
            qn = np.asarray(args[0]).astype(f)
            q = np.asarray(args[1]).astype(f)
            with np.errstate(divide='ignore', invalid='ignore'):
                x_mean = np.minimum(np.maximum(q * 1.0 / (qn + EPS), c['x_min']), c['x_max'])
                lam = np.asarray((c['lambda_factor'] * x_mean) ** (-self.nu / self.b))
                lam[q == 0] = np.nan
                N0 = np.asarray((self.nu / c['ntot_factor']) * qn * lam ** ((self.mu + 1) / self.nu))
                N0 = N0 * 1000 ** (-(1 + self.mu))
                lam = lam * 1000 ** (-self.nu)
            self.N0 = N0
            self.lambda_ = lam
            if self.kind == 'I':
                self.ntot = c['ntot_factor'] * self.N0 * self.lambda_ ** (-self.mu - 1)
            else:
                self.ntot = qn
            return
        # one moment
        if self.kind in ('R', 'G'):
            q = np.asarray(args[0]).astype(f)
            with np.errstate(divide='ignore'):
                lam = np.array((c['lambda_factor'] / q) ** (1. / (4. + self.mu)))
                lam[q == 0] = np.nan
            self.lambda_ = lam
            self.N0 = c['N0']
        elif self.kind == 'S':
            T = np.asarray(args[0])          # stays float32: N0 is evaluated in the input dtype
            q = np.asarray(args[1]).astype(f)
            self.N0 = 13.5 * (5.65 * 10 ** 5 * np.exp(-0.107 * (T - 273.15))) / 1000
            with np.errstate(divide='ignore'):
                lam = np.array((self.a * self.N0 * c['lambda_factor'] / q) ** (1. / (self.b + 1)))
                lam[q == 0] = np.nan
            self.lambda_ = lam
        elif self.kind == 'I':
            T = np.asarray(args[0])          # stays float32 inside get_mom_2 (hydrometeors.py:1321)
            QM = np.asarray(args[1]).astype(np.float64)
            Q2 = self.get_mom_2(T, QM / C1['I']['b'])
            N0 = Q2 ** ((self.b + 1) / (self.b - 2)) * QM ** ((2 + 1) / (2 - self.b))
            N0 /= 10 ** 5
            lam = (Q2 / QM) ** (1 / (self.b - 2))
            D = np.linspace(self.d_min, self.d_max, self.nbins_D)
            x = lam[:, None] * D.T / 1000
            N = N0[:, None] * phi_23_I(x)
            QM_est = np.nansum(self.a * D ** self.b * N, axis=1) * (D[1] - D[0])
            N0 = N0 / QM_est * QM
            self.N0 = N0
            self.lambda_ = lam
            self.ntot = np.nansum(N, axis=1) * (D[1] - D[0])
        else:
            raise ValueError(self.kind)

    @staticmethod
    def get_mom_2(T, QM):
        """hydrometeors.py:1277-1299 (Field et al. 2005)."""
        n = 3
        T = T - T0
        a = (5.065339 - 0.062659 * T - 3.032362 * n + 0.029469 * T * n
             - 0.000285 * T ** 2 + 0.312550 * n ** 2 + 0.000204 * T ** 2 * n
             + 0.003199 * T * n ** 2 - 0.015952 * n ** 3)
        a = 10 ** (a)
        b = (0.476221 - 0.015896 * T + 0.165977 * n + 0.007468 * T * n
             - 0.000141 * T ** 2 + 0.060366 * n ** 2 + 0.000079 * T ** 2 * n
             + 0.000594 * T * n ** 2 - 0.003577 * n ** 3)
        return (QM / a) ** (1 / b)

    # --- N(D), V(D) ------------------------------------------------------------
    def get_N(self, D):
        D = np.asarray(D, dtype=np.float64)
        if self.kind == 'I' and self.scheme == '1mom':
            x = self.lambda_[:, None] * D / 1000.
            return self.N0[:, None] * phi_23_I(x)
        lam = np.atleast_1d(self.lambda_)
        if np.isscalar(self.N0):
            return self.N0 * D ** self.mu * np.exp(-np.outer(lam, D ** self.nu))
        return np.outer(self.N0, D ** self.mu) * np.exp(-np.outer(lam, D ** self.nu))

    def get_V(self, D):
        return self.alpha * np.asarray(D, dtype=np.float64) ** self.beta

    def get_D_from_V(self, V):
        return (V / self.alpha) ** (1. / self.beta)

    def integrate_V(self):
        """hydrometeors.py:178-199; IceParticle :1256-1275 (no axis in np.sum:
        the total over ALL gates is returned as a length-1 array -- mirrored)."""
        if self.kind == 'I':
            D = np.linspace(self.d_min, self.d_max, self.nbins_D)
            dD = D[1] - D[0]
            N = self.get_N(D)
            v = np.sum(N * self.get_V(D)) * dD
            n = np.sum(N) * dD
            return np.array([v]), np.array([n])
        c = self.c
        v = (c['vel_factor'] * self.N0 * self.alpha / self.nu *
             self.lambda_ ** (-(self.beta + self.mu + 1) / self.nu))
        if self.scheme == '2mom':
            n = self.ntot
        else:
            n = c['ntot_factor'] * self.N0 / self.nu * self.lambda_ ** (-(self.mu + 1) / self.nu)
        return np.atleast_1d(v), np.atleast_1d(n)


# --------------------------------------------------------------------------- #
# a6 / N1: melting snow and graupel (hydrometeors.py:303-500, 1393-1481)
# --------------------------------------------------------------------------- #
D_LIMITS_1MOM = {'R': (0.1, 8.), 'S': (0.2, 20.), 'G': (0.2, 15.), 'I': (0.05, 2.)}   # constants_1mom.py:52-91


def vlinspace(a, b, N):
    """utilities.py:158-174."""
    a, b = np.asanyarray(a), np.asanyarray(b)
    return a[..., None] + (b - a)[..., None] / (N - 1) * np.arange(N)


class MeltHydro(object):
    """_MeltingHydrometeor with its two subclasses, one-moment scheme (the two-moment call
    set_psd(QN, QM) of doppler_scatter.py:209-211 raises ValueError in the reference).
    The PSD follows from the flux of the equivalent rain (get_N, :372-390); the solid
    counterpart only enters through its mass and fall-speed power laws."""

    def __init__(self, kind, scheme):
        if scheme != '1mom':
            raise ValueError('Invalid parameters, cannot set psd')
        self.kind = kind
        self.scheme = scheme
        self.rain = Hydro('R', scheme)
        self.solid = Hydro('S' if kind == 'mS' else 'G', scheme)
        self.rain.d_min, self.rain.d_max = D_LIMITS_1MOM['R']
        self.solid.d_min, self.solid.d_max = D_LIMITS_1MOM[self.solid.kind]
        self.nbins_D = 1024
        self.prop_factor = None
        self.d_min = self.d_max = None

    def set_fwet(self, fw):
        self.f_wet = fw
        self.d_max = fw * self.rain.d_max + (1 - fw) * self.solid.d_max
        self.d_min = fw * self.rain.d_min + (1 - fw) * self.solid.d_min

    def set_psd(self, *args):
        self.prop_factor = None
        if self.kind == 'mS':
            T, q, fw = np.array(args[0]), np.array(args[1]), np.array(args[2])
            self.solid.set_psd(T, q)
        else:
            q, fw = np.array(args[0]), np.array(args[1])
            self.solid.set_psd(q)
        self.rain.set_psd(q)
        self.set_fwet(fw)
        with np.errstate(divide='ignore', invalid='ignore'):
            self.prop_factor = q / self.integrate_M()

    def get_M(self, D):
        fw2 = self.f_wet ** 2
        return fw2[:, None] * (self.rain.a * D ** self.rain.b) + (1 - fw2)[:, None] * (self.solid.a * D ** self.solid.b)

    def _D_r(self, d):
        rho_m = self.get_M(d) / (np.pi / 6 * d ** 3)
        return (rho_m / RHO_W) ** (1 / 3.) * d

    def get_V(self, D):
        V_rain = self.rain.alpha * self._D_r(D) ** self.rain.beta
        V_dry = self.solid.alpha * D ** self.solid.beta
        phi = 0.246 * self.f_wet + (1 - 0.246) * self.f_wet ** 7
        return phi[:, None] * V_rain + (1 - phi)[:, None] * V_dry

    def get_N(self, D):
        Dr = self._D_r(D)
        dDr = (self._D_r(D + 0.01) - Dr) / 0.01
        lam = self.rain.lambda_
        Nr = self.rain.N0 * Dr ** self.rain.mu * np.exp(-lam[:, None] * Dr ** self.rain.nu)
        if self.prop_factor is not None:
            Nr = self.prop_factor[:, None] * Nr
        return Nr * (self.rain.alpha * Dr ** self.rain.beta) / self.get_V(D) * dDr

    def integrate_M(self):
        D = vlinspace(self.d_min, self.d_max, self.nbins_D)
        dD = D[:, 1] - D[:, 0]
        return np.sum(self.get_N(D) * self.get_M(D), axis=1) * dD

    def integrate_V(self):
        D = vlinspace(self.d_min, self.d_max, self.nbins_D)
        dD = D[:, 1] - D[:, 0]
        N = self.get_N(D)
        return np.sum(N * self.get_V(D), axis=1) * dD, np.sum(N, axis=1) * dD


# --------------------------------------------------------------------------- #
# a10: lookup (lut.py:309-344)
# --------------------------------------------------------------------------- #
def lut_index(x, lo, step, dim):
    """floor((x - lo)/step) in float32, clamped to [0, dim-1]."""
    x = np.asarray(x, dtype=np.float32)
    closest = np.floor((x - np.float32(lo)) / np.float32(step))
    closest = np.array(closest, dtype=int)
    closest[closest < 0] = 0
    closest[closest >= dim] = dim - 1
    return closest


def lookup_line(lut, e, t):
    """lut: dict(value_table [ne,nt,nd,12] f64, axes_limits, axes_step, axes).
    Axis 0 is 'e', axis 1 is 't' (or 'wc' for the melting species)."""
    v = lut['value_table']
    ie = lut_index(e, lut['axes_limits'][0][0], lut['axes_step'][0], v.shape[0])
    it = lut_index(t, lut['axes_limits'][1][0], lut['axes_step'][1], v.shape[1])
    return v[ie, it], ie, it


# --------------------------------------------------------------------------- #
# a13: observables from the 12 integrated scattering numbers (doppler_scatter.py:491-544)
# --------------------------------------------------------------------------- #
def get_pol_from_sz(sz, wavelength, K_squared):
    rxh = 2 * np.pi * (sz[:, 0] - sz[:, 1] - sz[:, 2] + sz[:, 3])
    z_h = wavelength ** 4 / (np.pi ** 5 * K_squared) * rxh
    rxv = 2 * np.pi * (sz[:, 0] + sz[:, 1] + sz[:, 2] + sz[:, 3])
    z_v = wavelength ** 4 / (np.pi ** 5 * K_squared) * rxv
    zdr = rxh / rxv
    kdp = 1e-3 * (180.0 / np.pi) * wavelength * (sz[:, 10] - sz[:, 8])
    ah = 4.343e-3 * (2 * wavelength * sz[:, 11])
    av = 4.343e-3 * (2 * wavelength * sz[:, 9])
    a = (sz[:, 4] + sz[:, 7]) ** 2 + (sz[:, 6] - sz[:, 5]) ** 2
    b = (sz[:, 0] - sz[:, 1] - sz[:, 2] + sz[:, 3])
    c = (sz[:, 0] + sz[:, 1] + sz[:, 2] + sz[:, 3])
    with np.errstate(invalid='ignore', divide='ignore'):
        rhohv = np.sqrt(a / (b * c))
    delta_hv = np.arctan2(sz[:, 5] - sz[:, 6], -sz[:, 4] - sz[:, 7])
    return z_h, z_v, zdr, rhohv, kdp, ah, av, delta_hv


def aliasing(v, nyquist):
    """utilities.py:142-155."""
    theta = (v + nyquist) / (2 * nyquist) * np.pi - np.pi / 2.
    theta_fold = np.arctan(np.tan(theta))
    return (theta_fold + np.pi / 2) * (2 * nyquist) / np.pi - nyquist


def nan_cumsum(x):
    x = np.array(x)
    x[np.isnan(x)] = 0
    return np.cumsum(x)


def nan_cumprod(x):
    x = np.array(x)
    x[np.isnan(x)] = 1
    return np.cumprod(x)


def proj_vel(U, V, W, vf, theta, phi):
    """doppler_scatter.py:30-47."""
    return ((U * np.sin(phi) + V * np.cos(phi)) * np.cos(theta) + (W - vf) * np.sin(theta))


def species_list(cfg):
    """doppler_scatter.py:99-106."""
    h = ['R', 'S', 'G']
    if cfg['microphysics']['with_melting']:
        h.extend(['mS', 'mG'])
    if cfg['microphysics']['scheme'] == '2mom':
        h.extend(['H'])
    if cfg['microphysics']['with_ice_crystals']:
        h.extend('I')
    return h


# --------------------------------------------------------------------------- #
# a16: Doppler spectrum (doppler_scatter.py:546-725)
# --------------------------------------------------------------------------- #
def get_diameter_from_rad_vel(hyd, order, phi, theta, U, V, W, rho_corr, varray):
    theta = np.deg2rad(theta)
    phi = np.deg2rad(phi)
    with np.errstate(divide='ignore', invalid='ignore'):
        wh = (1. / rho_corr * (W + (U * np.sin(phi) + V * np.cos(phi)) / np.tan(theta)
                               - varray / np.sin(theta)))
    idx = np.where(wh >= 0)[0]
    wh = wh[idx]
    D = np.zeros((len(idx), len(order)), dtype='float32')
    for i, h in enumerate(order):
        D[:, i] = hyd[h].get_D_from_V(wh)
        D[D >= hyd[h].d_max] = hyd[h].d_max     # sic: thresholds the whole matrix
        D[D <= hyd[h].d_min] = hyd[h].d_min
    Da = np.minimum(D[0:-1, :], D[1:, :])
    Db = np.maximum(D[0:-1, :], D[1:, :])
    mask = np.where(np.sum((Db - Da) == 0.0, axis=1) < len(order))[0]
    return Da[mask, :], Db[mask, :], idx[mask]


def get_doppler_spectrum(sub, hyd, luts, order, cfg, consts):
    scheme = cfg['microphysics']['scheme']
    KW = cfg['radar']['K_squared']
    varray = consts['VARRAY']
    n_gates = len(sub['dist'])
    refl = np.zeros((n_gates, len(varray)), dtype='float32')
    elev = sub['elev']
    elev_lut = np.array(elev, copy=True)
    elev_lut[elev_lut > 90] = 180 - elev_lut[elev_lut > 90]
    elev_lut[elev_lut < 0] = -elev_lut[elev_lut < 0]
    phi = sub['quad_pt'][0]
    v = sub['values']
    with np.errstate(invalid='ignore'):
        rho_corr = (v['RHO'] / v['RHO'][0]) ** 0.5
    for i in range(n_gates):
        if sub['mask'][i] != 0:
            continue
        present = []
        for h in order:
            Q = v['Q' + h + '_v'][i]
            T = v['T'][i]
            if Q > 0:
                present.append(h)
                if scheme == '1mom':
                    if h in ('S', 'I'):
                        hyd[h].set_psd(np.array([T]), np.array([Q]))
                    else:
                        hyd[h].set_psd(np.array([Q]))
                else:
                    hyd[h].set_psd(np.array([v['QN' + h + '_v'][i]]), np.array([Q]))
        nh = len(present)
        nd = 1024
        rcs = np.zeros((nd, nh), dtype='float32') + np.nan
        N = np.zeros((nd, nh), dtype='float32') + np.nan
        D = np.zeros((nd, nh), dtype='float32') + np.nan
        D_min = np.zeros((nh), dtype='float32') + np.nan
        step_D = np.zeros((nh), dtype='float32') + np.nan
        for j, h in enumerate(present):
            D[:, j] = np.linspace(hyd[h].d_min, hyd[h].d_max, hyd[h].nbins_D)
            D_min[j] = D[0, j]
            step_D[j] = D[1, j] - D[0, j]
            N[:, j] = hyd[h].get_N(D[:, j])
        for j, h in enumerate(present):
            sz, _, _ = lookup_line(luts[h], elev_lut[i:i + 1], v['T'][i:i + 1])
            sz = sz[0]
            rcs[:, j] = (2 * np.pi * (sz[:, 0] - sz[:, 1] - sz[:, 2] + sz[:, 3])).T
        if nh == 0:
            continue
        Da, Db, idx = get_diameter_from_rad_vel(hyd, present, phi, elev[i], v['U'][i], v['V'][i],
                                                v['W'][i], rho_corr[i], varray)
        if len(idx) == 0:
            continue
        refl[i, idx] = spectrum_bins(Da, Db, rcs, D, N, step_D, D_min)
        refl[i, idx] *= consts['WAVELENGTH'] ** 4 / (np.pi ** 5 * KW ** 2)
    return refl


# --------------------------------------------------------------------------- #
# a16: spectrum broadening (doppler_scatter.py:357-367, 727-801)
# --------------------------------------------------------------------------- #
TURB_A = 1.6          # global_constants.py: turbulence constant A (Doviak & Zrnic p. 409)


def spectral_width_turb(ranges, EDR, cfg):
    """doppler_scatter.py:727-758 (python-3 division: 3/2 = 1.5)."""
    sigma_r = 0.35 * cfg['radar']['radial_resolution']
    sigma_theta = np.deg2rad(cfg['radar']['3dB_beamwidth']) / (4. * np.sqrt(np.log(2)))
    std_turb = np.zeros((len(EDR),))
    with np.errstate(invalid='ignore'):
        idx_r = sigma_r < 0.1 * ranges * sigma_theta
        std_turb[idx_r] = ((ranges[idx_r] * EDR[idx_r] * sigma_theta * TURB_A ** (3 / 2.)) / 0.72) ** (1 / 3.)
        idx_r = sigma_r >= 0.1 * ranges * sigma_theta
        std_turb[idx_r] = (((EDR[idx_r] * sigma_r * (1.35 * TURB_A) ** (3 / 2)) /
                            (11. / 15. + 4. / 15. * (ranges[idx_r] * sigma_theta / sigma_r) ** 2)
                            ** (-3 / 2.)) ** (1 / 3.))
    return std_turb


def spectral_width_motion(elevations, cfg, consts):
    """doppler_scatter.py:760-777 (the wavelength in mm divided by 100, sic)."""
    wavelength = consts['WAVELENGTH'] / 100.
    bandwidth_3dB = cfg['radar']['3dB_beamwidth']
    ang_vel = cfg['radar']['antenna_speed']
    return ((wavelength * ang_vel * np.cos(np.deg2rad(elevations))) /
            (2 * np.pi * np.deg2rad(bandwidth_3dB)))


def gaussian_filter1d(x, sigma, truncate=4.0):
    """scipy.ndimage.gaussian_filter of a 1-D array with the defaults the reference uses
    (scipy 1.18.1 here; _filters.py: gaussian_filter -> gaussian_filter1d -> correlate1d):
    axes with sigma <= 1e-15 (or NaN) are skipped, radius int(4 sigma + 0.5), weights
    exp(-x^2 / (2 sigma^2)) normalised to 1, mode 'reflect' (d c b a | a b c d | d c b a),
    float64 accumulation, result in the dtype of the input."""
    sd = float(sigma)
    if not sd > 1e-15:
        return np.array(x, copy=True)
    lw = int(truncate * sd + 0.5)
    k = np.arange(-lw, lw + 1)
    w = np.exp(-0.5 / (sd * sd) * k ** 2)
    w = w / w.sum()
    n = len(x)
    idx = np.mod(np.arange(n)[:, None] + k[None, :], 2 * n)
    idx = np.where(idx >= n, 2 * n - 1 - idx, idx)
    return (np.asarray(x, dtype=np.float64)[idx] @ w).astype(np.asarray(x).dtype)


def broaden_spectrum(spectrum, std, consts):
    """doppler_scatter.py:779-801: rows with no power become NaN (0 / 0)."""
    v = consts['VARRAY']
    v_res = v[2] - v[1]
    original_power = np.sum(spectrum, 1)
    for i, t in enumerate(std):
        spectrum[i, :] = gaussian_filter1d(spectrum[i, :], t / v_res)
    convolved_power = np.sum(spectrum, 1)
    with np.errstate(invalid='ignore', divide='ignore'):
        spectrum = spectrum / convolved_power[:, None] * original_power[:, None]
    return spectrum


# --------------------------------------------------------------------------- #
# a7-a15: get_radar_observables (doppler_scatter.py:49-489)
# --------------------------------------------------------------------------- #
def get_radar_observables(subs, luts, cfg, accum='ref', debug=False):
    """accum='ref' stores the per-species accumulator in float32 exactly like the
    reference (doppler_scatter.py:133-134); accum='f64' keeps everything in
    float64 (attribution variant)."""
    consts = derived_constants(cfg)
    scheme = cfg['microphysics']['scheme']
    doppler_scheme = cfg['doppler']['scheme']
    att_corr = cfg['microphysics']['with_attenuation']
    radial_res = cfg['radar']['radial_resolution']
    KW = cfg['radar']['K_squared']
    nyq = cfg['radar'].get('nyquist_velocity', None)
    wavelength = consts['WAVELENGTH']
    num_beams = len(subs)
    idx_0 = int(num_beams / 2)
    n_gates = max([len(s['dist']) for s in subs])
    elev_0 = subs[idx_0]['quad_pt'][1]
    az_0 = subs[idx_0]['quad_pt'][0]
    hydrom_types = species_list(cfg)
    hyd = {}
    for h in hydrom_types:
        if h in ('mS', 'mG'):
            # the limits taken from the table (doppler_scatter.py:117-123) are overwritten by the
            # f_wet setter at every set_psd (hydrometeors.py:333-340)
            hyd[h] = MeltHydro(h, scheme)
            hyd[h].nbins_D = luts[h]['value_table'].shape[-2]
            continue
        hyd[h] = Hydro(h, scheme)
        d_axis = luts[h]['axes'][2]
        hyd[h].nbins_D = luts[h]['value_table'].shape[-2]
        hyd[h].d_min = d_axis[0]
        hyd[h].d_max = d_axis[-1]
    acc_t = np.float32 if accum == 'ref' else np.float64
    sz_integ = np.zeros((n_gates, len(hydrom_types), 12), dtype=acc_t) + np.nan
    rvel_avg = np.zeros(n_gates,) + np.nan
    total_weight_rvel = np.zeros(n_gates,)
    if doppler_scheme == 3:
        doppler_spectrum = np.zeros((n_gates, len(consts['VARRAY'])))
    dbg = {'e_idx': {}, 't_idx': {}, 'valid': {}}
    if cfg['integration']['scheme'] == 'ml':            # doppler_scatter.py:127-129
        total_weight_at_gates = np.sum(np.array([b['quad_weight'] for b in subs]), axis=0)

    for i, sub in enumerate(subs):
        v_integ = np.zeros(n_gates,)
        n_integ = np.zeros(n_gates,)
        if doppler_scheme == 3:
            ah_per_beam = np.zeros((n_gates,), dtype='float32') + np.nan
        for j, h in enumerate(hydrom_types):
            wet = h in ('mS', 'mG')
            if wet and not sub.get('has_melting', False):
                continue                                    # doppler_scatter.py:162-165
            if wet and doppler_scheme == 3:
                raise NotImplementedError('Doppler spectrum of melting species')
            elev_lut = sub['elev']                          # alias, folded in place (Q3)
            elev_lut[elev_lut > 90] = 180 - elev_lut[elev_lut > 90]
            elev_lut[elev_lut < 0] = -elev_lut[elev_lut < 0]
            T = sub['values']['T']
            QM = sub['values']['Q' + h + '_v']
            with np.errstate(invalid='ignore'):
                valid = QM > 0
            if not np.isscalar(sub['quad_weight']):          # doppler_scatter.py:186-189
                valid = np.logical_and(valid, sub['quad_weight'] > 0)
            if not np.any(valid):
                continue
            if scheme == '1mom':
                if h == 'mG':
                    fwet = sub['values']['fwet_' + h]
                    hyd[h].set_psd(QM[valid], fwet[valid])
                elif h == 'mS':
                    fwet = sub['values']['fwet_' + h]
                    hyd[h].set_psd(T[valid], QM[valid], fwet[valid])
                elif h in ('S', 'I'):
                    hyd[h].set_psd(T[valid], QM[valid])
                else:
                    hyd[h].set_psd(QM[valid])
            else:
                QN = sub['values']['QN' + h + '_v']
                hyd[h].set_psd(QN[valid], QM[valid])
            if wet:
                # per-gate diameter grids between the wet-fraction dependent limits (:216-221)
                list_D = vlinspace(hyd[h].d_min, hyd[h].d_max, hyd[h].nbins_D)
                dD = (list_D[:, 1] - list_D[:, 0])[:, None]
            else:
                list_D = luts[h]['axes'][2]
                dD = list_D[1] - list_D[0]
            N = hyd[h].get_N(list_D)
            if len(N.shape) == 1:
                N = np.reshape(N, [len(N), 1])
            sz, ie, it = lookup_line(luts[h], elev_lut[valid], fwet[valid] if wet else T[valid])
            if debug:
                dbg['e_idx'][(i, h)] = ie
                dbg['t_idx'][(i, h)] = it
                dbg['valid'][(i, h)] = valid.copy()
            sz_psd_integ = np.einsum('ijk,ij->ik', sz, N) * dD
            if not np.isscalar(sub['quad_weight']):          # doppler_scatter.py:259-265
                wts = sub['quad_weight'][valid] / total_weight_at_gates[valid]
                sz_integ[valid, j, :] = _nansum2(sz_integ[valid, j, :], wts[:, None] * sz_psd_integ)
            else:
                sz_integ[valid, j, :] = _nansum2(sz_integ[valid, j, :], sz_psd_integ * sub['quad_weight'])
            if doppler_scheme == 1:
                vh, n = hyd[h].integrate_V()
                v_integ[valid] = _nansum2(v_integ[valid], vh)
                n_integ[valid] = _nansum2(n_integ[valid], n)
            elif doppler_scheme == 2:
                rcs = 2 * np.pi * (sz[:, :, 0] - sz[:, :, 1] - sz[:, :, 2] + sz[:, :, 3])
                v_f = hyd[h].get_V(list_D)
                trap = getattr(np, 'trapezoid', None) or np.trapz
                vh_w = trap(np.multiply(v_f, N * rcs), axis=1)
                n_w = trap(N * rcs, axis=1)
                v_integ[valid] = _nansum2(v_integ[valid], vh_w)
                n_integ[valid] = _nansum2(n_integ[valid], n_w)
            elif doppler_scheme == 3:
                ah = 4.343e-3 * 2 * wavelength * sz_psd_integ[:, 11]
                ah *= radial_res / 1000.
                # sic: result has the length of the valid subset, aligned at gate 0
                ah_per_beam = _nansum2(ah_per_beam, ah)
        if doppler_scheme in (1, 2):
            with np.errstate(invalid='ignore', divide='ignore'):
                v_hydro = v_integ / n_integ
            theta = np.deg2rad(sub['elev'])
            phi = np.deg2rad(sub['quad_pt'][0])
            pw = proj_vel(sub['values']['U'], sub['values']['V'], sub['values']['W'], v_hydro, theta, phi)
            total_weight_rvel = total_weight_rvel + (~np.isnan(pw)) * sub['quad_weight']
            rvel_avg = _nansum2(rvel_avg, pw * sub['quad_weight'])
        elif doppler_scheme == 3:
            spec = get_doppler_spectrum(sub, hyd, luts, list(hydrom_types), cfg, consts)
            add_specwidth = np.zeros(len(spec))
            if cfg['doppler'].get('turbulence_correction', 0):
                add_specwidth += spectral_width_turb(consts['RANGE_RADAR'], sub['values']['EDR'], cfg)
            if cfg['doppler'].get('motion_correction', 0):
                add_specwidth += spectral_width_motion(sub['elev'], cfg, consts)
            if np.sum(add_specwidth) > 0:          # False when a gate of the sub-beam is NaN (sic)
                spec = broaden_spectrum(spec, add_specwidth, consts)
            if att_corr:
                ahc = nan_cumsum(ah_per_beam)
                sum_power = np.nansum(spec, axis=1)
                idv = sum_power > 0
                with np.errstate(divide='ignore'):
                    spdb = 10 * np.log10(sum_power)
                frac = sum_power[idv] / (10 ** (0.1 * (spdb[idv] - ahc[idv])))
                spec[idv, :] /= frac[:, None]
            spec = spec * sub['quad_weight']
            doppler_spectrum += spec

    if debug:
        dbg['sz_species'] = sz_integ.copy()
    sz_tot = np.nansum(sz_integ, axis=1)
    sz_tot[sz_tot == 0] = np.nan
    ZH, ZV, ZDR, RHOHV, KDP, AH, AV, DELTA_HV = get_pol_from_sz(sz_tot, wavelength, KW)
    PHIDP = nan_cumsum(2 * KDP) * radial_res / 1000. + DELTA_HV
    ZV_ATT = ZV.copy()
    ZH_ATT = ZH.copy()
    if att_corr:
        ZV_ATT *= nan_cumprod(10 ** (-0.1 * AV * (radial_res / 1000.)))
        ZH_ATT *= nan_cumprod(10 ** (-0.1 * AH * (radial_res / 1000.)))
        ZDR = ZH_ATT / ZV_ATT
    if doppler_scheme in (1, 2):
        with np.errstate(invalid='ignore', divide='ignore'):
            rvel_avg = rvel_avg / total_weight_rvel
    elif doppler_scheme == 3:
        with np.errstate(invalid='ignore', divide='ignore'):
            rvel_avg = np.nansum(np.tile(consts['VARRAY'], (n_gates, 1)) * doppler_spectrum, axis=1)
            rvel_avg /= np.nansum(doppler_spectrum, axis=1)
    if nyq is not None:
        nv = nyq(elev_0, az_0) if callable(nyq) else float(nyq)
        with np.errstate(invalid='ignore'):
            rvel_avg = aliasing(rvel_avg, nv)
    out = {'ZH': ZH, 'ZDR': ZDR, 'ZV': ZV, 'KDP': KDP, 'DELTA_HV': DELTA_HV, 'PHIDP': PHIDP,
           'RHOHV': RHOHV, 'ATT_H': AH, 'ATT_V': AV, 'RVEL': rvel_avg}
    if doppler_scheme == 3:
        out['DSPECTRUM'] = doppler_spectrum
    mask = np.zeros(n_gates,)
    for s in subs:
        mask = mask + s['mask'][0:n_gates]
    mask /= float(num_beams)
    mask[np.logical_and(mask > -1, mask <= 0)] = 0
    res = {'values': out, 'mask': mask, 'lats': subs[idx_0]['lats'], 'lons': subs[idx_0]['lons'],
           'dist': subs[idx_0]['dist'], 'heights': subs[idx_0]['heights']}
    if debug:
        dbg['sz_total'] = sz_tot
        res['dbg'] = dbg
    return res


def cut_at_sensitivity(values, cfg):
    """doppler_scatter.py:804-862 for one ray (dict of 1-D arrays), in place."""
    sens = cfg['radar']['sensitivity']
    if not isinstance(sens, list):
        sens = [sens]
    n = len(values['ZH'])
    rr = cfg['radar']['radial_resolution'] * np.arange(n)
    with np.errstate(divide='ignore', invalid='ignore'):
        if len(sens) == 2:
            thr = (sens[0] - 20 * np.log10(sens[1] / 1000.)) + 20 * np.log10(rr / 1000.)
        elif len(sens) == 1:
            thr = sens[0] + 0 * rr
        else:
            raise ValueError('3-element sensitivity not restated')
        mask = 10 * np.log10(values['ZH']) < thr
    for k in list(values.keys()):
        if k in SIMULATED_VARIABLES and k != 'DSPECTRUM':
            values[k] = np.array(values[k])
            values[k][mask] = np.nan
    return values


# --------------------------------------------------------------------------- #
# the worker (radar_operator.py:407-427) and a scan
# --------------------------------------------------------------------------- #
def simulate_ray(model, luts, cfg, azimuth, elevation, accum='ref', debug=False,
                 output_variables='only_radar', sensitivity_cut=False):
    subs = get_interpolated_radial(model, azimuth, elevation, cfg, debug=debug)
    res = None
    if output_variables in ('all', 'only_radar'):
        res = get_radar_observables(subs, luts, cfg, accum=accum, debug=debug)
    if output_variables in ('all', 'only_model'):
        mv, mask = integrate_radials(subs)
        if res is None:
            i0 = int(len(subs) / 2)
            res = {'values': {}, 'mask': mask, 'lats': subs[i0]['lats'], 'lons': subs[i0]['lons'],
                   'dist': subs[i0]['dist'], 'heights': subs[i0]['heights']}
        res['values'].update(mv)
    if sensitivity_cut and 'ZH' in res['values']:
        cut_at_sensitivity(res['values'], cfg)
    if debug:
        res['subs'] = subs
    return res


def simulate_scan(model, luts, cfg, azimuths, elevations, accum='ref', output_variables='only_radar',
                  sensitivity_cut=False, debug=False):
    """List of rays given as parallel arrays (az[r], el[r]); returns dict of
    [n_rays, n_gates] arrays."""
    rays = [simulate_ray(model, luts, cfg, a, e, accum=accum, output_variables=output_variables,
                         sensitivity_cut=sensitivity_cut, debug=debug)
            for a, e in zip(azimuths, elevations)]
    out = {}
    for k in rays[0]['values'].keys():
        out[k] = np.stack([np.asarray(r['values'][k]) for r in rays])
    out['mask'] = np.stack([r['mask'] for r in rays])
    out['lats'] = np.stack([r['lats'] for r in rays])
    out['lons'] = np.stack([r['lons'] for r in rays])
    out['dist'] = np.stack([r['dist'] for r in rays])
    out['heights'] = np.stack([r['heights'] for r in rays])
    if debug and 'dbg' in rays[0]:
        out['sz_total'] = np.stack([r['dbg']['sz_total'] for r in rays])
    return out
