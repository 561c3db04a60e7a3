# -*- coding: utf-8 -*-
"""One-moment (operational COSMO) microphysics constants, names as in
cosmo_pol/constants/constants_1mom.py:45-104.

N0 [mm-1 m-3], MU [-], m(D) = AM * D**BM [kg, D in mm], v(D) = AV * D**BV [m/s],
D_MIN / D_MAX [mm]; *_FACTOR = gamma-function factors of the analytic moments."""
import numpy as np
from scipy.special import gamma as _G

RHO_W = 1000. / (1000 ** 3)


def _mm(coef, expo):
    """SI (m based) power-law intercept -> mm based."""
    return coef * (1000 ** -expo)


# graupel
N0_G, BM_G, BV_G, MU_G = 4 * 1E3, 3.1, 0.89, 0.0
AM_G, AV_G = _mm(169.6, BM_G), _mm(442.0, BV_G)
D_MIN_G, D_MAX_G = 0.2, 15
LAMBDA_FACTOR_G = AM_G * N0_G * _G(BM_G + 1)
VEL_FACTOR_G = _G(MU_G + BV_G + 1)
NTOT_FACTOR_G = _G(MU_G + 1)

# snow (N0 is temperature dependent, see hydrometeors.Snow.set_psd)
BM_S, BV_S, MU_S = 2., 0.25, 0.0
AM_S, AV_S = _mm(0.038, BM_S), _mm(4.9, BV_S)
D_MIN_S, D_MAX_S = 0.2, 20
LAMBDA_FACTOR_S = _G(BM_S + 1)
VEL_FACTOR_S = _G(MU_S + BV_S + 1)
NTOT_FACTOR_S = _G(MU_S + 1)

# rain (gamma DSD of COSMO 4.21)
RAIN_FACTOR, MU_R = 0.1, 0.5
N00_r = 8E6 / (1000 ** (1 + MU_R)) * (0.01) ** (-MU_R)
N0_R = RAIN_FACTOR * N00_r * np.exp(3.2 * MU_R)
BM_R, BV_R = 3., 0.5
AM_R, AV_R = np.pi / 6. * RHO_W, _mm(130, BV_R)
D_MIN_R, D_MAX_R = 0.1, 8
LAMBDA_FACTOR_R = AM_R * N0_R * _G(1. + BM_R + MU_R)
VEL_FACTOR_R = _G(MU_R + BV_R + 1)
NTOT_FACTOR_R = _G(MU_R + 1)

# ice crystals (fall-speed law borrowed from the two-moment scheme)
BM_I, MU_I = 3, 0.0
AM_I = _mm(130, BM_I)
D_MIN_I, D_MAX_I = 0.05, 2
AV_I, BV_I = 0.9655930341942476, 1.2019867549668874
LAMBDA_FACTOR_I = _G(BM_I + 1)
NTOT_FACTOR_I = _G(MU_I + 1)
VEL_FACTOR_I = _G(MU_I + BV_I + 1)

# double-normalised PSD shape for moments (2,3), Field et al. (2005)
PHI_23_COEFFS = (490.6, 20.78, 17.46, 0.6357, 3.290)


def PHI_23_I(x):
    c = PHI_23_COEFFS
    return c[0] * np.exp(-c[1] * x) + c[2] * x ** c[3] * np.exp(-c[4] * x)
