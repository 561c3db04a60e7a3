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
        # the reference picks the azimuth row through an argmin over (az - azimuth) without
        # abs(); every shipped table has a single azimuth (0) per elevation, so it is row 0
        return float(nyq[idx_el[0]])

    interpolator.table = (elev, azim, nyq)
    return interpolator
