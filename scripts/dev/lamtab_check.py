"""CPU emulation of k_build_lamtab's verification for one bin per species (which pieces get rejected and why)."""
import sys, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'oracle')
import radop_oracle as O
from cosmo_pol_b200 import synthetic
from numpy.polynomial import chebyshev as C
U0, OCT, PPO = -10, 22, 4
def cols_of(lut, ie, it, mu, mode):
    t = lut.value_table[ie, it].astype(np.float64)
    D = lut.axes[lut.axes_names['d']].astype(np.float64)
    dD = float(np.float32(D[1]) - np.float32(D[0]))
    pre = np.full_like(D, dD) if mode == "ice1" else D ** mu * dD
    c = np.stack([t[:, 0]-t[:, 1]-t[:, 2]+t[:, 3], t[:, 0]+t[:, 1]+t[:, 2]+t[:, 3], t[:, 4]+t[:, 7], t[:, 6]-t[:, 5], t[:, 10]-t[:, 8], t[:, 9], t[:, 11]], 1)
    return (c * pre[:, None]).astype(np.float32).astype(np.float64), D
def weights(mode, lam, D, nu):
    if mode == 'ice1':
        x = lam[:, None] * D[None, :]
        y0 = D[0]
        return 490.6*np.exp(-lam[:, None]*(20.78*D[None, :]-3.29*y0)) + 17.46*x**0.6357*np.exp(-3.29*lam[:, None]*(D[None, :]-y0)), -3.29*lam*y0*np.log2(np.e)
    Y = D ** nu
    return np.exp(-lam[:, None]*(Y[None, :]-Y[0])), -lam*Y[0]*np.log2(np.e)
def check(cols, D, mode, nu):
    xj = np.concatenate([np.cos(np.pi*(np.arange(6)+0.5)/6), [0.0, 1.0]])
    res = []
    for piece in range(OCT*PPO):
        u = U0 + (piece + 0.5*(xj+1))/PPO
        lam = 2.0**u
        w, sh = weights(mode, lam, D, nu)
        F = w@cols; A = w@np.abs(cols)
        with np.errstate(divide='ignore', invalid='ignore'):
            L = np.log2(A) + sh[:, None]; R = F/A
        worst = 0
        for f, v in enumerate(np.concatenate([L, R], 1).T):
            if not np.all(np.isfinite(v)): worst = np.inf; break
            cf = C.cheb2poly(C.chebfit(xj[:6], v[:6], 5))
            e = max(abs(np.polyval(cf[::-1], 0)-v[6]), abs(np.polyval(cf[::-1], 1)-v[7]))
            rng = max(1.0, np.abs(cf[1:]).sum())
            e = e*(np.log(2) if f < cols.shape[1] else 1.0)/rng
            worst = max(worst, e) if f >= cols.shape[1] else max(worst, e)
        res.append(worst)
    return np.array(res)
if __name__ == '__main__':
    for scheme in ('1mom', '2mom'):
        C_ = O.C1 if scheme == '1mom' else O.C2
        for kind in (['R', 'S', 'G', 'I'] if scheme == '1mom' else ['R', 'S', 'G', 'H', 'I']):
            lut = synthetic.make_lut(kind, 5.6 if scheme == '1mom' else 9.41, scheme, n_e=2, n_t=2)
            mode = 'ice1' if (scheme == '1mom' and kind == 'I') else 'gen'
            mu, nu = C_[kind]['mu'], C_[kind]['nu']
            cols, D = cols_of(lut, 1, 1, mu, mode)
            e = check(cols, D, mode, nu)
            bad = np.flatnonzero(~(e <= 2e-7))
            print(scheme, kind, 'mu %.2f nu %.3f' % (mu, nu), 'rejected pieces:', len(bad), 'u of rejected:', np.round(U0 + bad/PPO, 2)[:40], 'max err in live range', e[(U0+np.arange(len(e))/PPO > -3) & (U0+np.arange(len(e))/PPO < 8)].max())
