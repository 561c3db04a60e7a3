# -*- coding: utf-8 -*-
"""Antenna quadrature: the fan of sub-beams of one ray.

Scheme 1 of cosmo_pol/interpolation/interpolation.py:148-166 (Gauss-Hermite nodes scaled by
sigma = bw/(2 sqrt(2 ln 2)), weights multiplied by |cos(pts_ver)| and normalised) followed by
the cumulative weight threshold of interpolation.py:343-354."""
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


def subbeam_fan(config):
    bw = config['radar']['3dB_beamwidth']
    nh, nv = int(config['integration']['nh_GH']), int(config['integration']['nv_GH'])
    if config['integration']['scheme'] != 1:
        raise NotImplementedError('only integration scheme 1 (Gauss-Hermite) is part of this build (SURVEY.md 8f N3)')
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
