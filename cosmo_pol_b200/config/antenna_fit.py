# -*- coding: utf-8 -*-
"""Sum-of-Gaussians fit of a measured antenna diagram (cosmo_pol/config/antenna_fit.py:21-120): what the
reference's configuration loader runs when ``integration.antenna_diagram`` names a file, to obtain the
``antenna_params`` rows (amplitude [dB], mean [deg], sigma [deg]) that integration scheme 2 samples
(interpolation.py:194-228).  Host side, configuration time; nothing here is on the per-gate path.

The reference's objective function reshapes with ``len(params) / 3`` (a float under Python 3), so its fit only runs
under Python 2; this restatement does what it was written to do."""
import numpy as np


def gaussian_sum_db(x, params):
    """10 log10 of the sum of the Gaussians; params rows: amplitude [dB], mean, sigma."""
    p = np.atleast_2d(np.asarray(params, dtype=np.float64))
    lin = np.sum(10 ** (0.1 * p[:, 0:1]) * np.exp(-(np.asarray(x)[None, :] - p[:, 1:2]) ** 2 / (2 * p[:, 2:3] ** 2)), axis=0)
    with np.errstate(divide='ignore'):
        return 10 * np.log10(lin)


def antenna_diagram(x, bw, ampl, mu):
    """Linear power of a sum of Gaussians given by their 3 dB widths, amplitudes [dB] and offsets."""
    sigma = np.asarray(bw, dtype=np.float64) / (2 * np.sqrt(2 * np.log(2)))
    y = np.zeros(np.shape(x))
    for s, a, m in zip(sigma, ampl, mu):
        y = y + 10 ** (0.1 * a) * np.exp(-(x - m) ** 2 / (2 * s ** 2))
    return y


def optimize_gaussians(x, y, n_gaussians):
    """Fit `n_gaussians` Gaussians to the diagram y [dB] at the angles x [deg]: the lobes (local maxima of y, plus
    the axis) ranked by level seed amplitudes and means, sigma starts at 0.5 deg and stays in [0.1, 2]; the main lobe
    is pinned at 0 dB, 0 deg; SLSQP on the root of the summed squared dB error."""
    from scipy.optimize import minimize
    from scipy.signal import argrelextrema
    x, y = np.asarray(x, dtype=np.float64), np.asarray(y, dtype=np.float64)
    peaks = argrelextrema(y, np.greater)
    level, where = y[peaks], x[peaks]
    if 0 not in where:
        where, level = np.append(where, 0), np.append(level, 0)
    order = np.argsort(level)[::-1][:n_gaussians]
    n = len(order)
    start = np.column_stack((level[order], where[order], np.full(n, 0.5)))
    bounds = [[None, None], [None, None], [0.1, 2]] * n
    bounds[0], bounds[1] = [0, 0], [0, 0]

    def cost(flat):
        return np.sqrt(np.sum((gaussian_sum_db(x, flat.reshape(n, 3)) - y) ** 2))

    return minimize(cost, start.ravel(), bounds=bounds, method='SLSQP')['x'].reshape(n, 3)
