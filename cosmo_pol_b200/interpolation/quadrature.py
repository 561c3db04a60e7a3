# -*- coding: utf-8 -*-
"""Antenna quadrature: the fan of sub-beams of one ray.

Scheme 1 of cosmo_pol/interpolation/interpolation.py:148-166 (Gauss-Hermite nodes scaled by
sigma = bw/(2 sqrt(2 ln 2)), weights multiplied by |cos(pts_ver)| and normalised) followed by
the cumulative weight threshold of interpolation.py:343-354; scheme 2 (multi-Gaussian antenna pattern,
irregular points, interpolation.py:194-228)."""
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


def subbeam_fan(config):
    bw = config['radar']['3dB_beamwidth']
    if config['integration']['scheme'] == 2:
        return multi_gaussian_fan(config)
    nh, nv = int(config['integration']['nh_GH']), int(config['integration']['nv_GH'])
    if config['integration']['scheme'] == 'ml':
        return melting_layer_fan(config)
    if config['integration']['scheme'] != 1:
        raise NotImplementedError("integration schemes 1 (Gauss-Hermite), 2 (multi-Gaussian antenna) and 'ml' (melting-layer "
                                  "oversampling) are part of this build; 3 and 4 need an antenna diagram file / a pickle on the "
                                  "authors' disk, 5 and 6 are rejected by the reference's own configuration check (cfg.py:101)")
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
