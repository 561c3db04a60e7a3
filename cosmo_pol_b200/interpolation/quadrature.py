# -*- coding: utf-8 -*-
"""Antenna quadrature: the fan of sub-beams of one ray.

Scheme 1 of cosmo_pol/interpolation/interpolation.py:148-166 (Gauss-Hermite nodes scaled by
sigma = bw/(2 sqrt(2 ln 2)), weights multiplied by |cos(pts_ver)| and normalised) followed by
the cumulative weight threshold of interpolation.py:343-354; scheme 2 (multi-Gaussian antenna pattern,
irregular points, interpolation.py:194-228); schemes 3 and 5 on a measured antenna diagram (:232-258, 276-303);
scheme 'ml' (:168-193)."""
import numpy as np
from numpy.polynomial.hermite import hermgauss


class SubbeamFan(object):
    """pts_hor [nh], pts_ver [nv], weights [nh, nv] and the kept sub-beams in the reference's
    order (horizontal index major): i_hor, j_ver, weight."""

    def __init__(self, pts_hor, pts_ver, weights, keep):
        self.pts_hor, self.pts_ver, self.weights = pts_hor, pts_ver, weights
        ii, jj = np.nonzero(keep)           # row-major == horizontal major
        self.i_hor = ii.astype(np.int32)
        self.j_ver = jj.astype(np.int32)
        self.daz = pts_hor[ii].astype(np.float64)
        self.dele = pts_ver[jj].astype(np.float64)
        self.weight = weights[ii, jj].astype(np.float64)
        self.n_sub = len(ii)
        self.center = int(self.n_sub / 2)
        # integration scheme 'ml': sub-beams whose weight only counts near the edges of the melting layer
        self.ml_only = np.zeros(self.n_sub, dtype=np.int32)


    @classmethod
    def irregular(cls, daz, dele, weight):
        """Fan of unrelated points (integration scheme 2): every sub-beam has its own elevation offset and
        therefore its own trajectory row (interpolation.py:440-494)."""
        fan = cls.__new__(cls)
        n = len(daz)
        fan.pts_hor = np.asarray(daz, dtype=np.float64)
        fan.pts_ver = np.asarray(dele, dtype=np.float64)
        fan.weights = np.asarray(weight, dtype=np.float64)
        fan.i_hor = np.arange(n, dtype=np.int32)
        fan.j_ver = np.arange(n, dtype=np.int32)
        fan.daz, fan.dele, fan.weight = fan.pts_hor, fan.pts_ver, fan.weights
        fan.n_sub = n
        fan.center = int(n / 2)
        fan.ml_only = np.zeros(n, dtype=np.int32)
        return fan


def _threshold(weights, fraction):
    """Smallest weight kept so that the kept ones hold `fraction` of the total (interpolation.py:343-352)."""
    srt = np.sort(np.asarray(weights).ravel())[::-1]
    cum = np.cumsum(srt / np.sum(srt))
    cum[-1] = 1.
    return srt[np.where(cum >= fraction)[0][0]]


def multi_gaussian_fan(config):
    """Integration scheme 2 (interpolation.py:194-228): an antenna pattern given as a sum of Gaussians
    (``antenna_params`` rows: amplitude [dB], mean [deg], sigma [deg]) sampled with Gauss-Hermite nodes
    along the radius and Gauss-Legendre nodes around the beam axis; weight = polar area element times the
    cosine of the elevation offset.  Offsets are relative to the ray direction."""
    from numpy.polynomial.legendre import leggauss
    integ = config['integration']
    ap = np.atleast_2d(np.asarray(integ['antenna_params'], dtype=np.float64))
    xa, wa = leggauss(int(integ['na_GL']))
    xr, wr = hermgauss(int(integ['nr_GH']))
    amp, mean, sig = 10 ** (0.1 * ap[:, 0]), ap[:, 1], ap[:, 2]
    # index order of the reference's triple loop: radial node, Gaussian, angular node
    r = (mean[None, :, None] + np.sqrt(2) * sig[None, :, None] * xr[:, None, None]) * np.ones((1, 1, len(xa)))
    theta = (np.pi * xa + np.pi)[None, None, :] * np.ones_like(r)
    w = (np.pi * wa[None, None, :] * wr[:, None, None] * amp[None, :, None] * np.sqrt(2) * sig[None, :, None] * np.abs(r))
    w = w * np.cos(r * np.sin(theta))
    r, theta, w = r.ravel(), theta.ravel(), w.ravel()
    total = 0.0
    for x in w:                      # the reference accumulates the normalisation sequentially (:216)
        total += x
    w = w / total
    keep = w >= _threshold(w, integ['weight_threshold'])
    return SubbeamFan.irregular((r * np.cos(theta))[keep], (r * np.sin(theta))[keep], w[keep])


def melting_layer_fan(config):
    """Integration scheme 'ml' (interpolation.py:168-193, 423-436): Gauss-Hermite with ten times the vertical points
    (made odd), sorted by decreasing weight.  The first nv_GH + 1 vertical points of that order keep their weight
    at every gate; the others only count within a few gates of the first and the last melting-layer gate of their
    own sub-beam (their weight profile is a Gaussian-smoothed pair of spikes, made on the device: k_ml_weights)."""
    bw = config['radar']['3dB_beamwidth']
    nh, nv = int(config['integration']['nh_GH']), int(config['integration']['nv_GH'])
    nv_ml = 10 * nv
    if not nv_ml % 2:
        nv_ml += 1
    sigma = bw / (2 * np.sqrt(2 * np.log(2)))
    xh, wh = hermgauss(nh)
    xv, wv = hermgauss(nv_ml)
    pts_hor, pts_ver = xh * sigma, xv * sigma
    order = np.argsort(wv)[::-1]
    wv, pts_ver = wv[order], pts_ver[order]
    weights = np.outer(wh * sigma, wv * sigma) * np.abs(np.cos(np.deg2rad(pts_ver)))
    weights /= np.sum(weights.ravel())
    keep = weights >= _threshold(weights, config['integration']['weight_threshold']) if (nh > 1 or nv > 1) \
        else np.ones_like(weights, dtype=bool)
    fan = SubbeamFan(pts_hor, pts_ver, weights, keep)
    fan.ml_only = (fan.j_ver > nv).astype(np.int32)
    return fan


def _regular_fan(pts_hor, pts_ver, weights, fraction, broadening):
    """Normalised weights on a regular (horizontal x vertical) grid and the weight threshold of interpolation.py:343-354."""
    weights = weights / np.sum(weights.ravel())
    keep = weights >= _threshold(weights, fraction) if broadening else np.ones_like(weights, dtype=bool)
    return SubbeamFan(np.asarray(pts_hor, dtype=np.float64), np.asarray(pts_ver, dtype=np.float64), weights, keep)


def read_antenna_diagram(path):
    """Two comma-separated columns: off-axis angle [deg] (symmetric about 0) and one-way power [dB]; returns the
    angles and the two-way power in linear units (interpolation.py:236-240, 280-284)."""
    table = np.genfromtxt(path, delimiter=',')
    return table[:, 0], (10 ** (0.1 * table[:, 1])) ** 2


def pattern_on_grid(angles, power, x, y):
    """The measured cut through the antenna pattern, rotated about the beam axis and sampled at the offsets (x, y):
    what utilities.vector_1d_to_polar / polar2cartesian (utilities.py:285-341) do -- both halves of the cut become the
    rows theta = 0 and theta = pi of a three-row polar grid and scipy's cubic `map_coordinates` reads it at the
    fractional (theta, radius) index of every point.  Returns [len(y), len(x)]."""
    from scipy.interpolate import interp1d
    from scipy.ndimage import map_coordinates
    mid = int(np.floor(len(angles) / 2.))
    radius = angles[mid:]
    theta = np.array([0., np.pi, 2 * np.pi])
    polar = np.zeros((3, len(radius)))
    polar[0] = power[mid:]
    polar[1] = power[:mid + 1][::-1]
    polar[2] = polar[0]
    gx, gy = np.meshgrid(x, y)
    r_pt = np.sqrt(gx * gx + gy * gy).ravel()
    t_pt = (np.arctan2(gx, gy) + np.pi).ravel()
    i_r = interp1d(radius, np.arange(len(radius)), bounds_error=False)(r_pt)
    i_t = interp1d(theta, np.arange(3))(t_pt)
    i_r[r_pt > radius.max()] = len(radius) - 1
    i_r[r_pt < radius.min()] = 0
    return map_coordinates(polar, np.array([i_t, i_r]), order=3).reshape(gx.shape)


def legendre_antenna_fan(config):
    """Integration scheme 3 (interpolation.py:232-258): Gauss-Legendre nodes over the angular extent of a measured
    antenna diagram, weighted with the two-way pattern at the nodes and the cosine of the elevation offset."""
    from numpy.polynomial.legendre import leggauss
    integ = config['integration']
    nh, nv = int(integ['nh_GH']), int(integ['nv_GH'])
    angles, power = read_antenna_diagram(integ['antenna_diagram'])
    bounds = np.max(angles)
    xh, wh = leggauss(nh)
    xv, wv = leggauss(nv)
    pts_hor, pts_ver = xh * bounds, xv * bounds
    weights = pattern_on_grid(angles, power, pts_hor, pts_ver).T * np.outer(wh, wv)
    weights = weights * np.abs(np.cos(np.deg2rad(pts_ver)))
    weights = weights * (2 * bounds)
    return _regular_fan(pts_hor, pts_ver, weights, integ['weight_threshold'], nh > 1 or nv > 1)


def gautschi_nodes(w_func, left, right, n_points, n_samples=8192):
    """Gauss nodes and weights for an arbitrary weight function on (left, right) -- cosmo_pol/interpolation/
    quadrature.py:33-101: the weight is sampled at `n_samples` midpoints, the recurrence coefficients of its orthogonal
    polynomials come from the discretised Stieltjes procedure, nodes and weights from the eigen-decomposition of the
    Jacobi matrix (Golub-Welsch); NaN / infinite matrix entries are zeroed as the reference does."""
    import scipy.linalg
    step = (float(right) - left) / n_samples
    z = np.linspace(left + 0.5 * step, right - 0.5 * step, n_samples)
    w = step * w_func(z)
    wz = z * w
    cur = np.ones(z.shape)
    cur /= np.sqrt(np.dot(cur, cur))
    old = np.zeros(z.shape)
    diag, off = np.empty(n_points), np.empty(n_points)
    norm_old = 1.0
    for j in range(n_points):
        norm = np.dot(w * cur, cur)
        diag[j] = np.dot(wz * cur, cur) / norm
        off[j] = 0.0 if j == 0 else norm / norm_old
        cur, old, norm_old = (z - diag[j]) * cur - off[j] * old, cur, norm
    sub = np.sqrt(off[1:])
    jacobi = np.diag(diag) + np.diag(sub, k=-1) + np.diag(sub, k=1)
    jacobi[~np.isfinite(jacobi)] = 0.
    nodes, vec = scipy.linalg.eigh(jacobi)
    order = nodes.argsort()
    return nodes[order], (vec[0, :] ** 2 * w.sum())[order]


def gautschi_fan(config):
    """Integration scheme 5 (interpolation.py:276-303; present in the reference's code, but its configuration check,
    cfg.py:101, does not list it): discrete Gautschi quadrature whose weight function IS the measured two-way antenna
    pattern (times the cosine of the offset in the vertical), so few nodes integrate the pattern exactly."""
    from scipy.interpolate import interp1d
    integ = config['integration']
    nh, nv = int(integ['nh_GH']), int(integ['nv_GH'])
    angles, power = read_antenna_diagram(integ['antenna_diagram'])
    bounds = np.max(angles)
    pattern = interp1d(angles, power, fill_value=0)
    pattern_cos = interp1d(angles, power * np.cos(np.abs(np.deg2rad(angles))), fill_value=0)
    pts_hor, wh = gautschi_nodes(pattern, -bounds, bounds, nh)
    pts_ver, wv = gautschi_nodes(pattern_cos, -bounds, bounds, nv)
    return _regular_fan(pts_hor, pts_ver, np.outer(wh, wv), integ['weight_threshold'], nh > 1 or nv > 1)


def subbeam_fan(config):
    bw = config['radar']['3dB_beamwidth']
    scheme = config['integration']['scheme']
    if scheme == 2:
        return multi_gaussian_fan(config)
    nh, nv = int(config['integration']['nh_GH']), int(config['integration']['nv_GH'])
    if scheme == 'ml':
        return melting_layer_fan(config)
    if scheme in (3, 5):
        if not config['integration'].get('antenna_diagram'):
            raise ValueError('integration scheme {} needs integration.antenna_diagram: a comma-separated file of off-axis '
                             'angles [deg] and one-way power [dB] (interpolation.py:236, 280)'.format(scheme))
        return legendre_antenna_fan(config) if scheme == 3 else gautschi_fan(config)
    if scheme != 1:
        raise NotImplementedError("integration schemes 1 (Gauss-Hermite), 2 (multi-Gaussian antenna), 3 (Gauss-Legendre on an "
                                  "antenna diagram), 5 (Gautschi on an antenna diagram) and 'ml' (melting-layer oversampling) "
                                  "are part of this build; 4 reads a pickle from the authors' disk (interpolation.py:262) and 6 "
                                  "needs the SpectralToolbox package (:307)")
    sigma = bw / (2 * np.sqrt(2 * np.log(2)))
    xh, wh = hermgauss(nh)
    xv, wv = hermgauss(nv)
    pts_hor, pts_ver = xh * sigma, xv * sigma
    weights = np.outer(wh * sigma, wv * sigma) * np.abs(np.cos(np.deg2rad(pts_ver)))
    weights /= np.sum(weights.ravel())
    srt = np.sort(weights.ravel())[::-1]
    cum = np.cumsum(srt / np.sum(srt))
    cum[-1] = 1.
    threshold = srt[np.where(cum >= config['integration']['weight_threshold'])[0][0]]
    keep = weights >= threshold if (nh > 1 or nv > 1) else np.ones_like(weights, dtype=bool)
    return SubbeamFan(pts_hor, pts_ver, weights, keep)
