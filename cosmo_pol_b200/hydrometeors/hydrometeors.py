# -*- coding: utf-8 -*-
"""hydrometeors.py -- hydrometeor species of the one- and two-moment schemes.

Host-side, vectorised counterpart of cosmo_pol/hydrometeors/hydrometeors.py: the same
class names and methods (``create_hydrometeor``, ``set_psd``, ``get_N``, ``get_V``,
``get_D_from_V``, ``integrate_V``) so user code keeps working, built on one table-driven
class instead of seven hand-written ones.  On the GPU path these classes only *describe*
the species (``abi_struct()`` -> ``RadopSpecies``); the per-gate PSD evaluation runs in
csrc/radop_kernels.cu (k_interp_psd: parameters, k_integrate: N(D) over the 1024 bins).

Shape / dielectric helpers of the reference (aspect ratios, canting angles, m(D) functions)
serve the offline T-matrix table generation only and are out of scope (SURVEY.md section 2,
row 13)."""
import numpy as np

from ..constants import constants_1mom, constants_2mom
from ..constants import global_constants as constants

_PSD_1MOM_FIXED, _PSD_1MOM_SNOW, _PSD_1MOM_ICE, _PSD_2MOM, _PSD_MELTING = 0, 1, 2, 3, 4
_SPECIES_ID = {'R': 0, 'S': 1, 'G': 2, 'H': 3, 'I': 4, 'mS': 5, 'mG': 6}
_CANTING = {'R': 10., 'S': 20., 'G': 40., 'H': 40., 'I': 5.}


def create_hydrometeor(hydrom_type, scheme='1mom', sedi=False):
    """Factory with the reference's signature (hydrometeors.py:39-74)."""
    cls = {'R': Rain, 'S': Snow, 'G': Graupel, 'H': Hail, 'I': IceParticle,
           'mS': MeltingSnow, 'mG': MeltingGraupel}.get(hydrom_type)
    if cls is None:
        raise ValueError('Invalid hydrometeor type, must be R, S, G, H, I, mS or mG')
    return cls(scheme)


class _Hydrometeor(object):
    kind = None

    def __init__(self, scheme='1mom'):
        if scheme not in ('1mom', '2mom'):
            scheme = '1mom'
        if self.kind == 'H':
            scheme = '2mom'
        self.scheme = scheme
        c = constants_1mom if scheme == '1mom' else constants_2mom
        k = self.kind
        self.nbins_D = 1024
        self.d_min = getattr(c, 'D_MIN_' + k)
        self.d_max = getattr(c, 'D_MAX_' + k)
        self.a = getattr(c, 'AM_' + k)
        self.b = getattr(c, 'BM_' + k)
        self.alpha = getattr(c, 'AV_' + k)
        self.beta = getattr(c, 'BV_' + k)
        self.mu = getattr(c, 'MU_' + k)
        self.nu = 1.0 if scheme == '1mom' else getattr(c, 'NU_' + k)
        self.N0 = getattr(c, 'N0_' + k, None) if scheme == '1mom' else None
        self.lambda_factor = getattr(c, 'LAMBDA_FACTOR_' + k)
        self.vel_factor = getattr(c, 'VEL_FACTOR_' + k)
        self.ntot_factor = getattr(c, 'NTOT_FACTOR_' + k)
        self.canting_angle_std = _CANTING[k]
        self.lambda_ = None
        self.ntot = None
        if scheme == '2mom':
            self.x_min = getattr(c, 'X_MIN_' + k)
            self.x_max = getattr(c, 'X_MAX_' + k)
        else:
            self.x_min = self.x_max = 0.0

    # -- description for the C-ABI ---------------------------------------------------------
    def psd_family(self):
        if self.scheme == '2mom':
            return _PSD_2MOM
        return {'R': _PSD_1MOM_FIXED, 'G': _PSD_1MOM_FIXED, 'S': _PSD_1MOM_SNOW, 'I': _PSD_1MOM_ICE}[self.kind]

    def abi_fields(self):
        return dict(species=_SPECIES_ID[self.kind], psd=self.psd_family(),
                    numeric_vn=1 if self.kind == 'I' else 0,
                    a=float(self.a), b=float(self.b), alpha=float(self.alpha), beta=float(self.beta),
                    mu=float(self.mu), nu=float(self.nu), N0=float(self.N0 if np.isscalar(self.N0) and self.N0 is not None else 0.0),
                    lambda_factor=float(self.lambda_factor), vel_factor=float(self.vel_factor),
                    ntot_factor=float(self.ntot_factor), x_min=float(self.x_min), x_max=float(self.x_max))

    # -- PSD ---------------------------------------------------------------------------------
    def set_psd(self, *args):
        """1-mom: (QM) for rain/graupel, (T, QM) for snow; 2-mom: (QN, QM)."""
        with np.errstate(divide='ignore', invalid='ignore'):
            if self.scheme == '2mom':
                qn = np.atleast_1d(np.asarray(args[0], dtype=np.float64))
                q = np.atleast_1d(np.asarray(args[1], dtype=np.float64))
                x_mean = np.clip(q / (qn + constants.EPS), self.x_min, self.x_max)
                lam = (self.lambda_factor * x_mean) ** (-self.nu / self.b)
                lam[q == 0] = np.nan
                n0 = (self.nu / self.ntot_factor) * qn * lam ** ((self.mu + 1) / self.nu)
                self.N0 = n0 * 1000 ** (-(1 + self.mu))
                self.lambda_ = lam * 1000 ** (-self.nu)
                self.ntot = qn
            elif self.kind in ('R', 'G'):
                q = np.atleast_1d(np.asarray(args[0], dtype=np.float64))
                lam = (self.lambda_factor / q) ** (1. / (4. + self.mu))
                lam[q == 0] = np.nan
                self.lambda_ = lam
                self.ntot = self.ntot_factor * self.N0 * lam ** (self.mu - 1)
            elif self.kind == 'S':
                T = np.atleast_1d(np.asarray(args[0], dtype=np.float64))
                q = np.atleast_1d(np.asarray(args[1], dtype=np.float64))
                self.N0 = 13.5 * (5.65 * 10 ** 5 * np.exp(-0.107 * (T - 273.15))) / 1000
                lam = (self.a * self.N0 * self.lambda_factor / q) ** (1. / (self.b + 1))
                lam[q == 0] = np.nan
                self.lambda_ = lam
                self.ntot = self.ntot_factor * self.N0 * lam ** (self.mu - 1)
            else:
                raise ValueError('invalid call to set_psd')

    def get_N(self, D):
        """N(D) = N0 D^mu exp(-lambda D^nu) [mm-1 m-3]; rows = gates, columns = diameters."""
        D = np.asarray(D, dtype=np.float64)
        lam = np.atleast_1d(self.lambda_)
        n0 = np.atleast_1d(self.N0) if not np.isscalar(self.N0) else np.full(lam.shape, self.N0)
        return n0[:, None] * D[None, :] ** self.mu * np.exp(-lam[:, None] * D[None, :] ** self.nu)

    def get_V(self, D):
        return self.alpha * np.asarray(D, dtype=np.float64) ** self.beta

    def get_D_from_V(self, V):
        return (V / self.alpha) ** (1. / self.beta)

    def get_M(self, D):
        return self.a * np.asarray(D, dtype=np.float64) ** self.b

    def integrate_V(self):
        """Analytic moments: (sum of v(D) N(D) dD, sum of N(D) dD) (hydrometeors.py:178-199)."""
        v = (self.vel_factor * self.N0 * self.alpha / self.nu *
             self.lambda_ ** (-(self.beta + self.mu + 1) / self.nu))
        if self.scheme == '2mom':
            n = self.ntot
        else:
            n = self.ntot_factor * self.N0 / self.nu * self.lambda_ ** (-(self.mu + 1) / self.nu)
        return np.atleast_1d(v), np.atleast_1d(n)


class Rain(_Hydrometeor):
    kind = 'R'


class Snow(_Hydrometeor):
    kind = 'S'


class Graupel(_Hydrometeor):
    kind = 'G'


class Hail(_Hydrometeor):
    kind = 'H'


class IceParticle(_Hydrometeor):
    """Ice crystals.  1-mom: double-normalised Field et al. (2005) PSD with moments (2,3)
    (hydrometeors.py:1231-1373)."""
    kind = 'I'

    @staticmethod
    def get_mom_2(T, QM):
        n = 3
        T = np.asarray(T, dtype=np.float64) - constants.T0
        a = (5.065339 - 0.062659 * T - 3.032362 * n + 0.029469 * T * n - 0.000285 * T ** 2 + 0.312550 * n ** 2
             + 0.000204 * T ** 2 * n + 0.003199 * T * n ** 2 - 0.015952 * n ** 3)
        b = (0.476221 - 0.015896 * T + 0.165977 * n + 0.007468 * T * n - 0.000141 * T ** 2 + 0.060366 * n ** 2
             + 0.000079 * T ** 2 * n + 0.000594 * T * n ** 2 - 0.003577 * n ** 3)
        return (QM / 10 ** a) ** (1 / b)

    def _grid(self):
        return np.linspace(self.d_min, self.d_max, self.nbins_D)

    def set_psd(self, arg1, arg2):
        if self.scheme == '2mom':
            return _Hydrometeor.set_psd(self, arg1, arg2)
        T = np.atleast_1d(np.asarray(arg1, dtype=np.float64))
        QM = np.atleast_1d(np.asarray(arg2, dtype=np.float64))
        Q2 = self.get_mom_2(T, QM / constants_1mom.BM_I)
        lam = (Q2 / QM) ** (1 / (self.b - 2))
        D = self._grid()
        shape = constants_1mom.PHI_23_I(lam[:, None] * D / 1000.)
        # intercept fixed by the mass integral on the discrete grid
        self.N0 = QM / (np.nansum(self.a * D ** self.b * shape, axis=1) * (D[1] - D[0]))
        self.lambda_ = lam
        self.ntot = np.nansum(self.N0[:, None] * shape, axis=1) * (D[1] - D[0])

    def get_N(self, D):
        if self.scheme == '2mom':
            return _Hydrometeor.get_N(self, D)
        D = np.asarray(D, dtype=np.float64)
        return self.N0[:, None] * constants_1mom.PHI_23_I(self.lambda_[:, None] * D / 1000.)

    def integrate_V(self):
        """Numeric, per gate.  (The reference sums over all gates at once,
        hydrometeors.py:1256-1275; the CUDA path mirrors that, this host helper returns
        the per-gate values.)"""
        D = self._grid()
        N = self.get_N(D)
        dD = D[1] - D[0]
        return np.sum(N * self.get_V(D), axis=1) * dD, np.sum(N, axis=1) * dD


class _MeltingHydrometeor(object):
    """Wet snow / wet graupel between their dry counterpart and the equivalent rain
    (hydrometeors.py:303-500): mass quadratic in the wet fraction, fall speed after
    Frick et al. (2013), PSD from the flux of the equivalent rain, scaled so that the mass is
    conserved.  One-moment scheme (the reference's two-moment call raises ValueError).
    On the GPU path the class describes the species (``abi_fields``); the per-gate
    evaluation is k_integrate_melting in csrc/radop_kernels.cu."""
    kind = None
    solid_cls = None

    def __init__(self, scheme='1mom'):
        if scheme != '1mom':
            raise ValueError('Invalid parameters, cannot set psd')
        self.scheme = scheme
        self.nbins_D = 1024
        self.equivalent_rain = Rain(scheme)
        self.equivalent_solid = self.solid_cls(scheme)
        self.prop_factor = None
        self.canting_angle_std = self.equivalent_solid.canting_angle_std
        self._f_wet = None

    @property
    def f_wet(self):
        return self._f_wet

    @f_wet.setter
    def f_wet(self, fw):
        r, d = self.equivalent_rain, self.equivalent_solid
        self._f_wet = fw
        self.d_max = fw * r.d_max + (1 - fw) * d.d_max
        self.d_min = fw * r.d_min + (1 - fw) * d.d_min

    def abi_fields(self):
        r, d = self.equivalent_rain, self.equivalent_solid
        return dict(species=_SPECIES_ID[self.kind], psd=_PSD_MELTING, numeric_vn=0,
                    a=float(d.a), b=float(d.b), alpha=float(d.alpha), beta=float(d.beta),
                    mu=float(r.mu), nu=float(r.nu), N0=float(r.N0), lambda_factor=float(r.lambda_factor),
                    vel_factor=0.0, ntot_factor=0.0, x_min=0.0, x_max=0.0,
                    melt=[float(r.a), float(r.b), float(r.alpha), float(r.beta),
                          float(r.d_min), float(r.d_max), float(d.d_min), float(d.d_max)])

    def set_psd(self, *args):
        """(T, QM, f_wet) for melting snow, (QM, f_wet) for melting graupel."""
        q = np.atleast_1d(np.asarray(args[-2], dtype=np.float64))
        fw = np.atleast_1d(np.asarray(args[-1], dtype=np.float64))
        self.prop_factor = None
        self.equivalent_rain.set_psd(q)
        self.f_wet = fw
        with np.errstate(divide='ignore', invalid='ignore'):
            self.prop_factor = q / self.integrate_M()

    def _grid(self):
        from ..utilities import vlinspace
        D = vlinspace(self.d_min, self.d_max, self.nbins_D)
        return D, D[:, 1] - D[:, 0]

    def get_M(self, D):
        fw2 = self.f_wet[:, None] ** 2
        return fw2 * self.equivalent_rain.get_M(D) + (1 - fw2) * self.equivalent_solid.get_M(D)

    def _equivalent_diameter(self, D):
        return (self.get_M(D) / (np.pi / 6 * constants.RHO_W)) ** (1 / 3.)

    def get_V(self, D):
        fw = self.f_wet[:, None]
        phi = 0.246 * fw + (1 - 0.246) * fw ** 7
        return (phi * self.equivalent_rain.get_V(self._equivalent_diameter(D)) +
                (1 - phi) * self.equivalent_solid.get_V(D))

    def get_N(self, D):
        D = np.asarray(D, dtype=np.float64)
        r = self.equivalent_rain
        Dr = self._equivalent_diameter(D)
        slope = (self._equivalent_diameter(D + 0.01) - Dr) / 0.01
        n_rain = r.N0 * Dr ** r.mu * np.exp(-np.atleast_1d(r.lambda_)[:, None] * Dr ** r.nu)
        if self.prop_factor is not None:
            n_rain = self.prop_factor[:, None] * n_rain
        return n_rain * r.get_V(Dr) / self.get_V(D) * slope

    def integrate_M(self):
        D, dD = self._grid()
        return np.sum(self.get_N(D) * self.get_M(D), axis=1) * dD

    def integrate_V(self):
        D, dD = self._grid()
        N = self.get_N(D)
        return np.sum(N * self.get_V(D), axis=1) * dD, np.sum(N, axis=1) * dD


class MeltingSnow(_MeltingHydrometeor):
    kind = 'mS'
    solid_cls = Snow


class MeltingGraupel(_MeltingHydrometeor):
    kind = 'mG'
    solid_cls = Graupel
