# -*- coding: utf-8 -*-
"""
synthetic.py -- seeded synthetic inputs shaped like the reference's inputs.

The reference needs a COSMO GRIB file (read through pycosmo,
cosmo_pol/radar_operator.py:217-309) and pre-computed T-matrix lookup tables
(cosmo_pol/lookup/compute_lut_sz.py:371-437, pytmatrix); neither is shipped nor
can be produced offline.  This module generates stand-ins with the same shapes,
dtypes, axes and units:

  * ``make_model``  : rotated-pole grid [nz,ny,nx] float32, level 0 = model top,
    terrain-following descending ``z-levels`` (last level = topography), smooth
    T / RHO / U / V / W and blob-shaped hydrometeor mass (and number) densities
    with exact zeros outside the blobs.
  * ``make_lut``    : one scattering table per species with the axes of
    compute_lut_sz.py:60-69 -- e = 0..90 step 2, t = 262..314 (rain, ice) or
    200..276 (snow, graupel, hail) step 2, 1024 diameters, 12 columns in the
    order of ``_flatten_matrices`` (compute_lut_sz.py:265-297) -- filled by an
    analytic Rayleigh-spheroid model (realistic magnitudes and signs), float64.
"""
import numpy as np
from scipy.ndimage import gaussian_filter

from .constants import constants_1mom, constants_2mom
from .lookup.lut import Lookup_table

ALBIS = (47.28436, 8.51201, 891.)          # option_files/CH_PPI_albis_alias.yml:2
SOUTH_POLE = (-43.0, 10.0)                 # COSMO-CH rotated south pole (lat, lon)
BASE_VARIABLES = ['U', 'V', 'W', 'QR_v', 'QS_v', 'QG_v', 'QI_v', 'RHO', 'T']   # radar_operator.py:35
BASE_VARIABLES_2MOM = ['QH_v', 'QNH_v', 'QNR_v', 'QNS_v', 'QNG_v', 'QNI_v']    # radar_operator.py:36


class ModelVariable(object):
    """Duck-type of the pycosmo variable the reference consumes: ``.data`` and
    ``.attributes`` with 'z-levels', 'proj_info', 'resolution', 'time'
    (interpolation.py:547-568, radar_operator.py:191)."""

    def __init__(self, name, data, attributes):
        self.name = name
        self.data = data
        self.attributes = attributes


def _wgs_to_rot(lat, lon, sp_lat, sp_lon):
    lon = np.deg2rad(lon); lat = np.deg2rad(lat)
    th = np.deg2rad(90. + sp_lat); ph = np.deg2rad(sp_lon)
    x = np.cos(lon) * np.cos(lat); y = np.sin(lon) * np.cos(lat); z = np.sin(lat)
    xn = np.cos(th) * np.cos(ph) * x + np.cos(th) * np.sin(ph) * y + np.sin(th) * z
    yn = -np.sin(ph) * x + np.cos(ph) * y
    zn = -np.sin(th) * np.cos(ph) * x - np.sin(th) * np.sin(ph) * y + np.cos(th) * z
    return np.rad2deg(np.arcsin(zn)), np.rad2deg(np.arctan2(yn, xn))


def _smooth_noise(rng, shape, sigma, coarsen=1):
    """Gaussian-filtered white noise with unit variance.  coarsen > 1 draws it on a horizontally coarser grid and
    interpolates linearly (for COSMO-1 sized domains, where filtering 7e7 points per variable takes minutes)."""
    if coarsen > 1:
        small = (shape[0], -(-shape[1] // coarsen) + 1, -(-shape[2] // coarsen) + 1)
        g = gaussian_filter(rng.standard_normal(small).astype(np.float32), (sigma[0], sigma[1] / coarsen, sigma[2] / coarsen),
                            mode='wrap')
        yy = np.arange(shape[1], dtype=np.float32) / coarsen
        xx = np.arange(shape[2], dtype=np.float32) / coarsen
        y0, x0 = yy.astype(int), xx.astype(int)
        fy, fx = (yy - y0)[None, :, None], (xx - x0)[None, None, :]
        f = ((1 - fy) * (1 - fx) * g[:, y0][:, :, x0] + (1 - fy) * fx * g[:, y0][:, :, x0 + 1] +
             fy * (1 - fx) * g[:, y0 + 1][:, :, x0] + fy * fx * g[:, y0 + 1][:, :, x0 + 1]).astype(np.float32)
    else:
        f = gaussian_filter(rng.standard_normal(shape).astype(np.float32), sigma, mode='wrap')
    f /= (np.std(f) + 1e-12)
    return f


def make_model(nz=60, ny=400, nx=400, res=0.02, radar=ALBIS, scheme='1mom', seed=1234,
               south_pole=SOUTH_POLE, top=22000.0, with_edr=False, with_refractivity=False, coarsen=1):
    """Returns dict name -> ModelVariable (all float32, C-contiguous).  coarsen: see _smooth_noise (1 = the generator
    every seeded fixture was made with)."""
    rng = np.random.default_rng(seed)
    _plain_noise = _smooth_noise

    def _smooth_noise_c(r, shp, sg):
        return _plain_noise(r, shp, sg, coarsen)
    rlat0, rlon0 = _wgs_to_rot(radar[0], radar[1], south_pole[0], south_pole[1])
    Lo1 = float(np.round(rlon0 - res * (nx - 1) / 2., 4))
    La1 = float(np.round(rlat0 - res * (ny - 1) / 2., 4))
    proj = {'Lo1': Lo1, 'La1': La1, 'Lo2': Lo1 + res * (nx - 1), 'La2': La1 + res * (ny - 1),
            'Latitude_of_southern_pole': south_pole[0], 'Longitude_of_southern_pole': south_pole[1]}
    jj, ii = np.meshgrid(np.arange(ny), np.arange(nx), indexing='ij')
    topo = np.full((ny, nx), 400.0)
    for _ in range(6):
        amp = rng.uniform(200, 2500)
        wid = rng.uniform(5, 40) * (min(ny, nx) / 400.0 + 0.25)
        cy, cx = rng.uniform(0, ny), rng.uniform(0, nx)
        topo += amp * np.exp(-((jj - cy) ** 2 + (ii - cx) ** 2) / (2 * wid ** 2))
    # keep the radar site above ground
    ry, rx = int(round((ny - 1) / 2.)), int(round((nx - 1) / 2.))
    scale = min(1.0, (radar[2] - 150.0) / max(topo[max(ry - 2, 0):ry + 3, max(rx - 2, 0):rx + 3].max(), 1.0))
    topo = 200.0 + (topo - 200.0) * scale if scale < 1.0 else topo
    k = np.arange(nz, dtype=np.float64)
    frac = ((nz - 1 - k) / (nz - 1)) ** 1.8
    z = (topo[None] + (top - topo[None]) * frac[:, None, None]).astype(np.float32)
    shape = (nz, ny, nx)
    sig = (1.5, 8 * ny / 400. + 1, 8 * nx / 400. + 1)
    zz = z.astype(np.float64)
    T = 288.15 - 0.0065 * zz + 0.5 * rng.standard_normal(shape)
    T = np.clip(T, 200., 314.).astype(np.float32)
    RHO = (1.225 * np.exp(-zz / 8500.)).astype(np.float32)
    U = (15. * _smooth_noise_c(rng, shape, sig)).astype(np.float32)
    V = (15. * _smooth_noise_c(rng, shape, sig)).astype(np.float32)
    W = (2. * _smooth_noise_c(rng, shape, sig)).astype(np.float32)

    def blobs(peak, cond, thr=0.3):
        f = _smooth_noise_c(rng, shape, sig)
        f = np.maximum(f - thr, 0.0)
        f = f / (f.max() + 1e-12) * peak
        f[~cond] = 0.0
        f[f < peak * 1e-4] = 0.0
        return f.astype(np.float32)

    data = {'U': U, 'V': V, 'W': W, 'RHO': RHO, 'T': T}
    data['QR_v'] = blobs(2e-3, T > 271.)
    data['QS_v'] = blobs(1e-3, T < 275.)
    data['QG_v'] = blobs(2e-3, np.logical_and(T > 268., T < 283.))
    data['QI_v'] = blobs(1e-4, T < 258., thr=0.0)
    names = list(BASE_VARIABLES)
    if scheme == '2mom':
        data['QH_v'] = blobs(5e-3, T > 255., thr=1.2)
        lim = {'R': (constants_2mom.X_MIN_R, constants_2mom.X_MAX_R),
               'S': (constants_2mom.X_MIN_S, constants_2mom.X_MAX_S),
               'G': (constants_2mom.X_MIN_G, constants_2mom.X_MAX_G),
               'H': (constants_2mom.X_MIN_H, constants_2mom.X_MAX_H),
               'I': (constants_2mom.X_MIN_I, constants_2mom.X_MAX_I)}
        for h in ('H', 'R', 'S', 'G', 'I'):
            lo, hi = lim[h]
            u = 0.5 + 0.25 * np.clip(_smooth_noise_c(rng, shape, sig), -1.8, 1.8)
            xbar = np.exp(np.log(lo) + (np.log(hi) - np.log(lo)) * u)
            data['QN' + h + '_v'] = (data['Q' + h + '_v'].astype(np.float64) / xbar).astype(np.float32)
        names += BASE_VARIABLES_2MOM
    if with_edr:
        # eddy dissipation rate [m2 s-3] (radar_operator.py:253-254), drawn last so that the other
        # variables do not depend on the flag
        edr = 10.0 ** (-3.0 + 0.8 * np.clip(_smooth_noise_c(rng, shape, sig), -2.0, 2.0))
        data['EDR'] = edr.astype(np.float32)
        names.append('EDR')
    if with_refractivity:
        # atmospheric refractivity N = (n - 1) 1e6 (radar_operator.py:239, 294-296): exponential reference
        # atmosphere with a seeded ducting-like perturbation near 1.5 km; drawn after everything else
        nn = 315.0 * np.exp(-zz / 7350.0) * (1.0 + 0.02 * np.exp(-((zz - 1500.0) / 400.0) ** 2))
        data['N'] = (nn + 0.3 * _smooth_noise_c(rng, shape, sig)).astype(np.float32)
        names.append('N')
    attrs = {'z-levels': z, 'proj_info': proj, 'resolution': [res, res], 'time': '2026-01-01 00:00:00'}
    return {n: ModelVariable(n, np.ascontiguousarray(data[n]), attrs) for n in names}


# --------------------------------------------------------------------------- #
# lookup tables
# --------------------------------------------------------------------------- #
ELEVATIONS = np.arange(0, 91, 2)           # compute_lut_sz.py:60
TEMPERATURES_LIQ = np.arange(262, 316, 2)  # compute_lut_sz.py:61
TEMPERATURES_SOL = np.arange(200, 278, 2)  # compute_lut_sz.py:62
NUM_DIAMETERS = 1024                       # compute_lut_sz.py:69

_SPECIES = {
    # |K|^2-like factor, oblateness slope, canting std [deg], absorption, t axis
    'R': dict(k2=0.93, obl=0.062, cant=10., absorb=0.020, liq=True),
    'S': dict(k2=0.035, obl=0.030, cant=20., absorb=0.0006, liq=False),
    'G': dict(k2=0.12, obl=0.020, cant=40., absorb=0.0015, liq=False),
    'H': dict(k2=0.19, obl=0.015, cant=40., absorb=0.0030, liq=False),
    'I': dict(k2=0.17, obl=0.400, cant=5., absorb=0.0010, liq=True),   # generator gives ice the liquid axis
}


def diameter_limits(hydrom, scheme):
    c = constants_1mom if scheme == '1mom' else constants_2mom
    return getattr(c, 'D_MIN_' + hydrom), getattr(c, 'D_MAX_' + hydrom)


W_CONTENTS = np.linspace(1E-3, 0.999, 100)     # compute_lut_sz.py:63


def make_melting_lut(hydrom, frequency=5.6, scheme='1mom', seed=4321, n_e=None, n_d=NUM_DIAMETERS):
    """Synthetic table of a melting species ('mS', 'mG') in the layout of sz_lut_melting
    (compute_lut_sz.py:299-369): axes e, wc (100 wet fractions), d = [n_wc, n_d] diameters between
    wet-fraction dependent limits (hydrometeors.py:333-340), 12 columns.  The scattering blends the
    solid species into rain with the wet fraction (wet particles scatter like water well before
    they are liquid)."""
    solid = 'S' if hydrom == 'mS' else 'G'
    sw, sd = _SPECIES['R'], _SPECIES[solid]
    rng = np.random.default_rng(seed + sum(ord(ch) for ch in hydrom))
    elev = ELEVATIONS if n_e is None else ELEVATIONS[:n_e]
    wc = W_CONTENTS
    rmin, rmax = diameter_limits('R', scheme)
    smin, smax = diameter_limits(solid, scheme)
    D = np.stack([np.linspace(w * rmin + (1 - w) * smin, w * rmax + (1 - w) * smax, n_d).astype('float32') for w in wc])
    lam = 299792458. / (frequency * 1e9) * 1000.
    Dd = D.astype(np.float64)[None, :, :]
    e = np.deg2rad(elev.astype(np.float64))[:, None, None]
    w = wc[None, :, None]
    blend = w ** 0.4
    mix = lambda a, b: (1 - blend) * a + blend * b
    kfac = mix(np.sqrt(sd['k2']), np.sqrt(sw['k2']))
    xsz = np.pi * Dd / lam
    amp = np.pi ** 2 * kfac * Dd ** 3 / (2 * lam ** 2) * (1.0 - 0.18 * xsz ** 2 / (1.0 + xsz ** 2))
    oblc = mix(sd['obl'], 1.6 * sw['obl'])
    cant = mix(sd['cant'], sw['cant'])
    obl = oblc * Dd / (1.0 + 0.08 * Dd) * np.exp(-2 * np.deg2rad(cant) ** 2) * np.cos(e) ** 2
    wig = 1.0 + 0.01 * np.sin(0.7 * Dd + 3.0 * w)
    shh = amp * (1.0 + obl) * wig
    svv = amp * (1.0 - obl) / wig
    dphi = 0.04 * xsz ** 2 * np.cos(e) ** 2 * (1 + 0.2 * w)
    rho0 = 0.9985 - 0.004 * obl / (oblc + 1e-12) - 0.002 * xsz ** 2 / (1 + xsz ** 2) - 0.01 * w * (1 - w)
    shape = (len(elev), len(wc), n_d)
    tab = np.zeros(shape + (12,), dtype=np.float64)
    hh2 = np.broadcast_to(shh ** 2, shape)
    vv2 = np.broadcast_to(svv ** 2, shape)
    co = np.broadcast_to(shh * svv * rho0, shape)
    tab[..., 0] = 0.5 * (hh2 + vv2)
    tab[..., 1] = 0.5 * (vv2 - hh2)
    tab[..., 2] = tab[..., 1]
    tab[..., 3] = tab[..., 0] * (1.0 + 1e-3)
    tab[..., 4] = -co * np.cos(dphi)
    tab[..., 5] = 0.5 * co * np.sin(dphi)
    tab[..., 6] = -0.5 * co * np.sin(dphi)
    tab[..., 7] = tab[..., 4]
    fwd = np.pi ** 2 * kfac * Dd ** 3 / (2 * lam ** 2)
    loss = mix(sd['absorb'], 2.0 * sw['absorb']) * fwd + (2. / 3.) * (2 * np.pi / lam) * fwd ** 2
    tab[..., 8] = np.broadcast_to(fwd * (1.0 - 0.8 * obl), shape)
    tab[..., 9] = np.broadcast_to(loss * (1.0 - 0.8 * obl), shape)
    tab[..., 10] = np.broadcast_to(fwd * (1.0 + 0.8 * obl), shape)
    tab[..., 11] = np.broadcast_to(loss * (1.0 + 0.8 * obl), shape)
    tab *= 1.0 + 0.01 * rng.standard_normal((len(elev), len(wc), 1, 12))
    lut = Lookup_table()
    lut.add_axis('e', elev)
    lut.add_axis('wc', wc)
    lut.add_axis('d', D)
    lut.add_axis('sz', np.arange(12))
    lut.set_value_table(tab)
    return lut


def make_lut(hydrom, frequency=5.6, scheme='1mom', seed=4321, n_e=None, n_t=None, n_d=NUM_DIAMETERS, family='smooth'):
    """Synthetic scattering table for one species; same axes/layout the reference
    generator writes (compute_lut_sz.py:428-433).  ``n_e``/``n_t`` truncate the
    e / t axes (tests use small tables).

    family='smooth': Rayleigh-Gans-like columns with a 1 % ripple (every column but s10 - s8 of one sign).
    family='mie': the same magnitudes with what real T-matrix columns of large particles show at C / X band
    (compute_lut_sz.py:265-297 tabulates them up to size parameters of several units for graupel and hail):
    resonance ripple along the diameter axis with a period of about one unit of the size parameter x = pi D / lambda
    (i.e. lambda / pi in D) and an amplitude that grows with x, and SIGN CHANGES of the differential quantities
    -- Z12 = Z21, Z34 = -Z43, Re and Im of S_hh - S_vv -- once x exceeds ~1.  It exists to stress the
    lambda-transformed tables (sign-changing columns are the hard case of their envelope / ratio representation)."""
    sp = _SPECIES[hydrom]
    rng = np.random.default_rng(seed + sum(ord(ch) for ch in hydrom))
    elev = ELEVATIONS if n_e is None else ELEVATIONS[:n_e]
    temp = TEMPERATURES_LIQ if sp['liq'] else TEMPERATURES_SOL
    temp = temp if n_t is None else temp[:n_t]
    dmin, dmax = diameter_limits(hydrom, scheme)
    D = np.linspace(dmin, dmax, n_d).astype('float32')
    lam = 299792458. / (frequency * 1e9) * 1000.           # wavelength [mm]
    Dd = D.astype(np.float64)[None, None, :]
    e = np.deg2rad(elev.astype(np.float64))[:, None, None]
    t = temp.astype(np.float64)[None, :, None]
    # dielectric factor drifts slowly with temperature; Rayleigh amplitude with a
    # mild size-parameter (resonance) modulation
    kfac = np.sqrt(sp['k2']) * (1.0 + 0.0015 * (t - 273.15))
    xsz = np.pi * Dd / lam
    amp = np.pi ** 2 * kfac * Dd ** 3 / (2 * lam ** 2) * (1.0 - 0.18 * xsz ** 2 / (1.0 + xsz ** 2))
    # oblateness seen by the beam: grows with D, vanishes at vertical incidence
    obl = sp['obl'] * Dd / (1.0 + 0.08 * Dd) * np.exp(-2 * np.deg2rad(sp['cant']) ** 2) * np.cos(e) ** 2
    wig = 1.0 + 0.01 * np.sin(0.7 * Dd + 0.05 * (t - 250.))
    dphi = 0.04 * xsz ** 2 * np.cos(e) ** 2 * (1 + 0.002 * (t - 273.15))     # backscatter phase difference
    if family == 'mie':
        # resonances: one ripple per unit of the size parameter (period lambda / pi in D; 1.5 ripples per unit for the
        # liquid species), amplitude saturating at 45 %; the differential terms oscillate through zero with a
        # slightly different frequency and a phase that drifts with elevation and temperature
        kres = 2.0 * np.pi * (1.5 if sp['liq'] else 1.0)
        grow = xsz ** 2 / (1.0 + xsz ** 2)
        amp = amp * (1.0 + 0.45 * grow * np.sin(kres * xsz + 0.3 + 0.002 * (t - 250.)))
        flip = np.cos(0.8 * kres * np.maximum(xsz - 0.6, 0.0) + 0.4 * np.cos(e) + 0.003 * (t - 250.))
        obl = obl * flip                                                     # Z12 / Z21, S_hh - S_vv change sign
        dphi = dphi * np.cos(0.6 * kres * xsz + 0.2)                         # Z34 / Z43 change sign
    elif family != 'smooth':
        raise ValueError('unknown table family {!r}'.format(family))
    shh = amp * (1.0 + obl) * wig
    svv = amp * (1.0 - obl) / wig
    rho0 = 0.9985 - 0.004 * obl / (sp['obl'] + 1e-12) - 0.002 * xsz ** 2 / (1 + xsz ** 2)
    shape = (len(elev), len(temp), n_d)
    tab = np.zeros(shape + (12,), dtype=np.float64)
    hh2 = np.broadcast_to(shh ** 2, shape)
    vv2 = np.broadcast_to(svv ** 2, shape)
    co = np.broadcast_to(shh * svv * rho0, shape)
    tab[..., 0] = 0.5 * (hh2 + vv2)                         # Z11
    tab[..., 1] = 0.5 * (vv2 - hh2)                         # Z12
    tab[..., 2] = tab[..., 1]                               # Z21
    tab[..., 3] = tab[..., 0] * (1.0 + 1e-3)                # Z22
    tab[..., 4] = -co * np.cos(dphi)                        # Z33
    tab[..., 5] = 0.5 * co * np.sin(dphi)                   # Z34
    tab[..., 6] = -0.5 * co * np.sin(dphi)                  # Z43
    tab[..., 7] = tab[..., 4]                               # Z44
    fwd = np.pi ** 2 * kfac * Dd ** 3 / (2 * lam ** 2)
    loss = sp['absorb'] * (1 + 0.01 * (273.15 - t)) * fwd + (2. / 3.) * (2 * np.pi / lam) * fwd ** 2
    tab[..., 8] = np.broadcast_to(fwd * (1.0 - 0.8 * obl), shape)     # Re S_vv (forward)
    tab[..., 9] = np.broadcast_to(loss * (1.0 - 0.8 * obl), shape)    # Im S_vv
    tab[..., 10] = np.broadcast_to(fwd * (1.0 + 0.8 * obl), shape)    # Re S_hh
    tab[..., 11] = np.broadcast_to(loss * (1.0 + 0.8 * obl), shape)   # Im S_hh
    # seeded smooth perturbation (1 %) so that neighbouring (e,t) bins never coincide
    pert = 1.0 + 0.01 * rng.standard_normal((len(elev), len(temp), 1, 12))
    tab *= pert
    lut = Lookup_table()
    lut.add_axis('e', elev)
    lut.add_axis('t', temp)
    lut.add_axis('d', D)
    lut.add_axis('sz', np.arange(12))
    lut.set_value_table(tab)
    return lut


def species_for(scheme, with_ice=True):
    h = ['R', 'S', 'G']
    if scheme == '2mom':
        h.append('H')
    if with_ice:
        h.append('I')
    return h


def make_all_luts(frequency=5.6, scheme='1mom', with_ice=True, seed=4321, n_e=None, n_t=None, with_melting=False,
                  n_e_melting=None, family='smooth'):
    luts = {h: make_lut(h, frequency, scheme, seed, n_e, n_t, family=family) for h in species_for(scheme, with_ice)}
    if with_melting:
        for h in ('mS', 'mG'):
            luts[h] = make_melting_lut(h, frequency, scheme, seed, n_e_melting if n_e_melting else n_e)
    return luts


def lut_as_plain(lut):
    """Plain-dict view of a Lookup_table (what the CPU oracle consumes)."""
    return {'value_table': lut.value_table, 'axes': lut.axes, 'axes_limits': lut.axes_limits,
            'axes_step': lut.axes_step, 'axes_names': dict(lut.axes_names)}


def model_as_plain(dic_vars):
    """Plain-dict view of a model dictionary (what the CPU oracle consumes)."""
    first = next(iter(dic_vars.values()))
    extra = {'refractivity': dic_vars['N'].data} if 'N' in dic_vars else {}
    return {'vars': {k: v.data for k, v in dic_vars.items() if k != 'N'}, **extra,
            'z_levels': first.attributes['z-levels'],
            'proj': first.attributes['proj_info'],
            'resolution': first.attributes['resolution']}
