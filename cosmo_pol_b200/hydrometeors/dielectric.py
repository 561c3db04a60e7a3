# -*- coding: utf-8 -*-
"""Dielectric factor of liquid water (Liebe et al. 1991 double-Debye model), used for the
config-time scalar K_squared when the user leaves it empty (cosmo_pol/hydrometeors/
dielectric.py:21-126, cfg.py:276-279).  Not on the per-gate path."""
import numpy as np

T_K_SQUARED = 283.15


def dielectric_water(T, f):
    """Complex refractive index of water at temperature T [K] and frequency f [GHz]."""
    theta = 1 - 300. / T
    e0 = 77.66 - 103.3 * theta
    e1 = 0.0671 * e0
    e2 = 3.52 + 7.52 * theta
    gamma1 = 20.20 + 146.5 * theta + 316 * theta * theta
    gamma2 = 39.8 * gamma1
    term1 = e0 - e1
    term2 = 1 + (f / gamma1) ** 2
    term3 = 1 + (f / gamma2) ** 2
    term4 = e1 - e2
    real = term1 / term2 + term4 / term3 + e2
    imag = (term1 / term2) * (f / gamma1) + (term4 / term3) * (f / gamma2)
    return np.sqrt(complex(real, imag))


def K_squared(f, T=T_K_SQUARED):
    m = dielectric_water(T, f)
    k = (m ** 2 - 1) / (m ** 2 + 2)
    return float(np.abs(k) ** 2)
