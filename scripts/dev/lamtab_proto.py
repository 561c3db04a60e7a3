"""Prototype: piecewise-Chebyshev tabulation of F_c(u) = sum_k tab[k][c] w(lambda, D_k), u = log2(lambda)."""
import sys, numpy as np
sys.path.insert(0, '.')
from cosmo_pol_b200 import synthetic
from numpy.polynomial import chebyshev as C

def combined(lut, ie, it, mu):
    t = lut.value_table[ie, it].astype(np.float64)   # [nd,12]
    D = lut.axes[lut.axes_names['d']].astype(np.float64)
    dD = D[1]-D[0]
    pre = D**mu*dD
    cols = np.stack([t[:,0]-t[:,1]-t[:,2]+t[:,3], t[:,0]+t[:,1]+t[:,2]+t[:,3], t[:,4]+t[:,7], t[:,6]-t[:,5], t[:,10]-t[:,8], t[:,9], t[:,11]],1)
    return cols*pre[:,None], D

def exact(cols, D, lam, nu=1.0):
    w = np.exp(-lam[:,None]*(D[None,:]**nu))     # [n, nd]
    return w@cols, w@np.abs(cols)

def build(cols, D, u0, u1, width, deg, nu=1.0):
    npc = int(np.ceil((u1-u0)/width))
    xj = np.cos(np.pi*(np.arange(deg+1)+0.5)/(deg+1))
    co = []
    for p in range(npc):
        a = u0+p*width
        u = a+0.5*width*(xj+1)
        F, A = exact(cols, D, 2.0**u, nu)
        L = np.log2(A); R = F/A
        cl = C.chebfit(xj, L, deg); cr = C.chebfit(xj, R, deg)
        co.append((cl, cr))
    return co

def evaluate(co, u, u0, width):
    p = np.clip(((u-u0)/width).astype(int), 0, len(co)-1)
    x = 2*((u-u0)/width-p)-1
    out = np.zeros((len(u), co[0][0].shape[1]))
    for i in range(len(u)):
        cl, cr = co[p[i]]
        out[i] = 2.0**C.chebval(x[i], cl)*C.chebval(x[i], cr)
    return out

if __name__ == '__main__':
    rng = np.random.default_rng(0)
    for hyd, mu in (('R', 0.5), ('S', 0.0), ('G', 0.0)):
        lut = synthetic.make_lut(hyd, 5.6, '1mom', n_e=3, n_t=3)
        cols, D = combined(lut, 1, 1, mu)
        for rough in (False, True):
            c = cols*(1+0.5*rng.standard_normal(cols.shape)) if rough else cols
            u0, u1 = -6.0, 8.0
            ut = rng.uniform(u0, u1, 400)
            F, A = exact(c, D, 2.0**ut)
            for width, deg in ((1.0, 12), (0.5, 8), (0.5, 10), (0.25, 6), (0.25, 8), (0.125, 5), (0.125, 6)):
                co = build(c, D, u0, u1, width, deg)
                G = evaluate(co, ut, u0, width)
                err = np.max(np.abs(G-F)/A, axis=0)
                print(hyd, 'rough' if rough else 'smooth', 'D=[%.2f,%.2f]' % (D[0], D[-1]), width, deg, 'max err/A %.1e' % err.max())
