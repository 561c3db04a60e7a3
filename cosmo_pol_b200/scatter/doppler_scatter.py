# -*- coding: utf-8 -*-
"""Host-side entry points of the scattering half of the hot path, with the names of
cosmo_pol/scatter/doppler_scatter.py.  The arithmetic itself (PSD evaluation, lookup,
diameter integral, quadrature sum, polarimetric variables, along-ray scans) runs on the GPU:
csrc/radop_kernels.cu (k_interp_psd, k_integrate, k_reduce_subbeams, k_ray_epilogue)."""
import numpy as np

from ..constants import global_constants as constants


def get_radar_observables(list_subradials, lut_sz=None):
    """Radar observables of one ray (doppler_scatter.py:49-489).

    The sub-radials must come from ``get_interpolated_radial`` of this package: that call
    already ran the fused GPU path for the ray (interpolation and scattering are one batched
    launch sequence here), so this returns the ray's integrated Radial -- after checking that the
    list is complete, in order and unmodified, and that ``lut_sz`` is the operator's own dictionary;
    a caller that edited a sub-radial (the reference would recompute from the edit) gets a ValueError,
    never the unedited answer."""
    from ..interpolation.interpolation import check_sealed
    result, _ = check_sealed(list_subradials, 'get_radar_observables', lut_sz)
    return result


def get_pol_from_sz(sz, KW=None):
    """Polarimetric variables from the 12 integrated scattering numbers
    (doppler_scatter.py:491-544); host helper, linear units."""
    from ..config import cfg
    wavelength = constants.WAVELENGTH
    K_squared = cfg.CONFIG['radar']['K_squared'] if KW is None else KW
    sz = np.asarray(sz)
    b = sz[:, 0] - sz[:, 1] - sz[:, 2] + sz[:, 3]
    c = sz[:, 0] + sz[:, 1] + sz[:, 2] + sz[:, 3]
    zfac = wavelength ** 4 / (np.pi ** 5 * K_squared)
    z_h, z_v = zfac * 2 * np.pi * b, zfac * 2 * np.pi * c
    kdp = 1e-3 * (180.0 / np.pi) * wavelength * (sz[:, 10] - sz[:, 8])
    ah, av = 4.343e-3 * 2 * wavelength * sz[:, 11], 4.343e-3 * 2 * wavelength * sz[:, 9]
    with np.errstate(invalid='ignore', divide='ignore'):
        rhohv = np.sqrt(((sz[:, 4] + sz[:, 7]) ** 2 + (sz[:, 6] - sz[:, 5]) ** 2) / (b * c))
        zdr = b / c
    delta_hv = np.arctan2(sz[:, 5] - sz[:, 6], -sz[:, 4] - sz[:, 7])
    return z_h, z_v, zdr, rhohv, kdp, ah, av, delta_hv


def sensitivity_threshold(config, n_gates):
    """Censoring threshold in dBZ per gate (doppler_scatter.py:816-834)."""
    sens = config['radar']['sensitivity']
    sens = list(sens) if isinstance(sens, (list, tuple)) else [sens]
    r = config['radar']['radial_resolution'] * np.arange(n_gates)
    with np.errstate(divide='ignore'):
        if len(sens) == 3:
            return sens[0] + constants.RADAR_CONSTANT_DB + sens[2] + 20 * np.log10(r / 1000.)
        if len(sens) == 2:
            return (sens[0] - 20 * np.log10(sens[1] / 1000.)) + 20 * np.log10(r / 1000.)
        if len(sens) == 1:
            return sens[0] + 0. * r
    return None


def cut_at_sensitivity(list_subradials, config=None):
    """Host version for lists (or lists of lists) of Radial; the batched GPU path applies the
    same censoring inside k_ray_epilogue."""
    from ..config import cfg
    config = cfg.CONFIG if config is None else config
    if not len(list_subradials):
        return list_subradials
    nested = isinstance(list_subradials[0], list)
    for sweep in (list_subradials if nested else [list_subradials]):
        for b in sweep:
            thr = sensitivity_threshold(config, len(b.dist_profile))
            if thr is None:
                print('Sensitivity parameters are invalid, cannot cut at specified sensitivity')
                return list_subradials
            with np.errstate(divide='ignore', invalid='ignore'):
                mask = 10 * np.log10(b.values['ZH']) < thr
            for k in b.values.keys():
                if k in constants.SIMULATED_VARIABLES and k != 'DSPECTRUM':
                    b.values[k][mask] = np.nan
    return list_subradials
