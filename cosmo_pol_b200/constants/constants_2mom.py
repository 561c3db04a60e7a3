# -*- coding: utf-8 -*-
"""Two-moment (Seifert-Beheng) microphysics constants, names as in
cosmo_pol/constants/constants_2mom.py:16-177.  The scheme is formulated in
particle mass x (D = a x**b, v = av x**bv, f(x) ~ x**mu exp(-lam x**nu)); the
diameter based equivalents are derived here and converted to mm."""
from scipy.special import gamma as _G

_MASS_BASED = {
    #        D = am x^bm       v = av x^bv        nu     mu    x_min    x_max    D range [mm]
    'G': (0.142, 0.314, 86.89371, 0.268325, 1. / 3., 1.0, 1E-09, 5E-04, 0.2, 15),
    'S': (2.4, 0.455, 4.2, 0.092, 0.5, 0.0, 1E-10, 2E-05, 0.2, 20),
    'R': (0.124, 1. / 3., 114.0137, 0.23437, 1. / 3., 0.0, 2.6E-10, 3E-06, 0.2, 8),
    'H': (0.1366, 1. / 3., 39.3, 1. / 6., 1. / 3., 1.0, 2.6E-09, 5E-04, 0.2, 15),
    'I': (0.124, 0.302, 317, 0.363, 1. / 3., 0.0, 1E-12, 1E-6, 0.05, 2),
}

# raindrop sedimentation v-D relation (unused by default: sedi=False)
C_1, C_2, C_3 = 9.65, 10.3, 600. * 1000. ** (-1)
TAU_1, TAU_2, D_EQ = 4, 1, 1.1


def _derive(tag, am_, bm_, av_, bv_, nu_, mu_, x_min, x_max, d_min, d_max):
    g = globals()
    bm = 1. / bm_
    am = am_ ** (-1 / bm_)
    bv = bv_ / bm_
    av = av_ * am_ ** (-bv_ / bm_)
    nu = nu_ / bm_
    mu = (mu_ + 1) / bm_ - 1
    g['X_MIN_' + tag], g['X_MAX_' + tag] = x_min, x_max
    g['D_MIN_' + tag], g['D_MAX_' + tag] = d_min, d_max
    g['BM_' + tag], g['BV_' + tag], g['NU_' + tag], g['MU_' + tag] = bm, bv, nu, mu
    g['LAMBDA_FACTOR_' + tag] = 1. / am * _G((mu + 1) / nu) / _G((mu + bm + 1) / nu)
    g['AM_' + tag] = am * 1000 ** (-bm)
    g['AV_' + tag] = av * 1000 ** (-bv)
    g['VEL_FACTOR_' + tag] = _G((mu + bv + 1) / nu)
    g['NTOT_FACTOR_' + tag] = _G((mu + 1) / nu)


for _tag, _row in _MASS_BASED.items():
    _derive(_tag, *_row)
