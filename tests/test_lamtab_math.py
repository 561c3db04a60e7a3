# -*- coding: utf-8 -*-
"""CPU pin of the lambda-transformed tables (csrc/radop_kernels.cu: k_build_lamtab / lt_horner).

The CUDA path evaluates the diameter integral of a pair, sum_k tab[k][c] * shape(lambda, D_k)
(hydrometeors.py:128-152 get_N, :1231-1254 ice; doppler_scatter.py:246-268 einsum), from degree-5
polynomials in log2(lambda) fitted per eighth-octave piece to the envelope exponent L_c and the signed
ratio R_c.  This file restates that construction in numpy (same nodes, same scaling, same verification
rule) and holds it to the oracle's own N(D) on every species family, without a GPU."""
import numpy as np
import pytest
from numpy.polynomial import chebyshev as C

import radop_oracle as O
from cosmo_pol_b200 import synthetic

U0, OCT, PPO, TOL = -10, 22, 8, 2e-7          # LT_U0, LT_OCT, LT_PPO, LT_TOL


def combined_columns(lut, ie, it, mu, ice1):
    """k_pack_lut: the 7 linear combinations get_pol_from_sz needs, premultiplied, rounded to float32."""
    t = lut.value_table[ie, it].astype(np.float64)
    D = lut.axes[lut.axes_names['d']].astype(np.float64)
    dD = float(np.float32(D[1]) - np.float32(D[0]))
    pre = np.full_like(D, dD) if ice1 else D ** mu * dD
    c = np.stack([t[:, 0] - t[:, 1] - t[:, 2] + t[:, 3], t[:, 0] + t[:, 1] + t[:, 2] + t[:, 3], t[:, 4] + t[:, 7],
                  t[:, 6] - t[:, 5], t[:, 10] - t[:, 8], t[:, 9], t[:, 11]], 1)
    return (c * pre[:, None]).astype(np.float32).astype(np.float64), D, dD


def shape_weights(ice1, lam, D, nu):
    """lt_weight / lt_shift: shape(lambda, D_k) with the common factor 2^shift taken out."""
    lam = np.atleast_1d(lam)[:, None]
    if ice1:
        x = lam * D[None, :]
        w = 490.6 * np.exp(-lam * (20.78 * D[None, :] - 3.29 * D[0])) + \
            17.46 * x ** 0.6357 * np.exp(-3.29 * lam * (D[None, :] - D[0]))
        return w, -3.29 * lam[:, 0] * D[0] * np.log2(np.e)
    Y = D ** nu
    return np.exp(-lam * (Y[None, :] - Y[0])), -lam[:, 0] * Y[0] * np.log2(np.e)


def build_piece(cols, D, ice1, nu, piece):
    """One (bin, piece) of k_build_lamtab: coefficients (power basis in x), exponents, accepted?"""
    xj = np.concatenate([np.cos(np.pi * (np.arange(6) + 0.5) / 6), [0.0, 1.0]])
    lam = 2.0 ** (U0 + (piece + 0.5 * (xj + 1)) / PPO)
    w, sh = shape_weights(ice1, lam, D, nu)
    A = w @ np.abs(cols)
    L = np.log2(A) + sh[:, None]
    R = (w @ cols) / A
    nc = cols.shape[1]
    coef = np.zeros((2 * nc, 6))
    err = np.zeros(2 * nc)
    for f, v in enumerate(np.concatenate([L, R], 1).T):
        p = C.cheb2poly(C.chebfit(xj[:6], v[:6], 5))
        p = np.pad(p, (0, 6 - len(p)))
        e = max(abs(p[0] - v[6]), abs(p.sum() - v[7]))
        err[f] = e * (np.log(2) if f < nc else 1.0) / max(1.0, np.abs(p[1:]).sum())
        coef[f] = p
    ok = np.all(np.isfinite(coef)) and np.all(err[:nc] + err[nc:] <= TOL)
    L0 = np.rint(coef[:nc, 0]).astype(int)
    coef[:nc, 0] -= L0
    return coef.astype(np.float32), L0, ok


def evaluate(coef, L0, x):
    """lt_horner + lt_scale in float32 arithmetic."""
    nc = len(L0)
    x = np.float32(x)
    out = np.zeros(nc)
    for k in range(nc):
        Lv, Rv = coef[k, 5], coef[nc + k, 5]
        for j in range(4, -1, -1):
            Lv = np.float32(Lv * x + coef[k, j])
            Rv = np.float32(Rv * x + coef[nc + k, j])
        out[k] = float(np.float32(np.exp2(Lv)) * Rv) * 2.0 ** int(L0[k])
    return out


CASES = [('1mom', 'R', 5.6), ('1mom', 'S', 5.6), ('1mom', 'G', 5.6), ('1mom', 'I', 5.6),
         ('2mom', 'R', 9.41), ('2mom', 'S', 9.41), ('2mom', 'G', 9.41), ('2mom', 'H', 9.41), ('2mom', 'I', 9.41)]


@pytest.mark.parametrize('scheme,kind,freq', CASES)
def test_lambda_table_reproduces_the_oracle_integral(scheme, kind, freq):
    consts = (O.C1 if scheme == '1mom' else O.C2)[kind]
    ice1 = scheme == '1mom' and kind == 'I'
    lut = synthetic.make_lut(kind, freq, scheme, n_e=3, n_t=2)
    # elevation index 2 is the steepest of the truncated axis; (0, 0) the flattest
    rng = np.random.default_rng(5)
    for ie, it in ((0, 0), (2, 1)):
        cols, D, dD = combined_columns(lut, ie, it, consts['mu'], ice1)
        # the oracle's own PSD evaluated on the table's diameter axis
        hyd = O.Hydro(kind, scheme)
        hyd.d_min, hyd.d_max = float(D[0]), float(D[-1])
        for u in rng.uniform(-2.5, 7.5, 12):
            lam = 2.0 ** u
            piece = int(np.floor((u - U0) * PPO))
            coef, L0, ok = build_piece(cols, D, ice1, consts['nu'], piece)
            assert ok, (scheme, kind, u)
            x = 2 * ((u - U0) * PPO - piece) - 1
            got = evaluate(coef, L0, x)
            # direct sum with the oracle's N(D) (N0 = 1): get_N of the generalised gamma / Field shape
            if ice1:
                N = O.phi_23_I(lam * D)
                ref = (N[:, None] * cols).sum(0)
                env = (np.abs(N)[:, None] * np.abs(cols)).sum(0)
            else:
                hyd.N0, hyd.lambda_ = np.array([1.0]), np.array([lam])
                N = np.asarray(hyd.get_N(D)).reshape(-1)
                # the table columns carry D^mu dD already: divide the oracle's N(D) by D^mu
                N = N / D ** consts['mu']
                ref = (N[:, None] * cols).sum(0)
                env = (np.abs(N)[:, None] * np.abs(cols)).sum(0)
            assert np.all(np.abs(got - ref) <= 1.5e-6 * env), (scheme, kind, u, np.max(np.abs(got - ref) / env))


def test_rejection_rule_catches_a_bad_fit():
    """A column that flips sign sharply inside a piece cannot be fitted to 2e-7 at degree 5 on a wide
    piece: the verification nodes must say so (the kernel then integrates those pairs directly)."""
    global PPO
    lut = synthetic.make_lut('R', 5.6, '1mom', n_e=2, n_t=2)
    cols, D, dD = combined_columns(lut, 0, 0, 0.5, False)
    cols = cols.copy()
    cols[:, 4] *= np.where(D < 2.0, 1.0, -1.0)           # s10 - s8 changing sign at D = 2 mm
    old = PPO
    try:
        PPO = 1                                           # octave-wide pieces
        accepted = [build_piece(cols, D, False, 1.0, p)[2] for p in range(10, 16)]
    finally:
        PPO = old
    assert not all(accepted)
    assert all(build_piece(cols, D, False, 1.0, p)[2] for p in range(8 * 10, 8 * 16))   # fine at 8 per octave
