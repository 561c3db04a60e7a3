# -*- coding: utf-8 -*-
"""Nyquist-velocity table lookup (cosmo_pol/config/nyquist_fit.py:14-30): nearest elevation
row of a csv with columns elevation, azimuth, nyquist."""
import numpy as np


def nyquist_interpolator(nyquist_file):
    data = np.genfromtxt(nyquist_file, delimiter=',', names=True)
    elev = np.atleast_1d(data['elevation']).astype(float)
    azim = np.atleast_1d(data['azimuth']).astype(float)
    nyq = np.atleast_1d(data['nyquist']).astype(float)

    def interpolator(elevation, azimuth):
        closest_el = elev[np.argmin(np.abs(elev - elevation))]
        idx_el = np.where(elev == closest_el)[0]
        az_at_el = azim[idx_el]
        # The reference selects the azimuth row with pandas label lookups that take the azimuth VALUES of the
        # elevation's rows as row labels of the whole table, and an argmin over (az - azimuth) without abs()
        # (nyquist_fit.py:22-28).  Restated as is: with one azimuth (0) per elevation -- every shipped table --
        # it is that elevation's row; a table whose azimuth values are not valid row labels raises KeyError
        # there and here.
        rows = az_at_el.astype(int)
        if np.any(rows != az_at_el) or np.any(rows < 0) or np.any(rows >= len(azim)):
            raise KeyError('azimuth values {} are not row labels of the Nyquist table'.format(list(az_at_el)))
        sub = azim[rows]
        closest_az = azim[int(np.argmin(sub - azimuth))]
        j = np.where(closest_az == sub)[0][0]
        return float(nyq[idx_el[j]])

    interpolator.table = (elev, azim, nyq)
    return interpolator
