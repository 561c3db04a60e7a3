# -*- coding: utf-8 -*-
"""Host-side helpers with the semantics of cosmo_pol/utilities/utilities.py (names kept
for user code).  The per-gate versions of these operations (NaN-as-zero sums, inclusive
scans, aliasing) live in the CUDA kernels; these numpy versions serve the thin Python
layers (result assembly, config validation) only."""
import numpy as np


class Range(object):
    """Closed interval usable with ``in`` (utilities.py:26-48)."""

    def __init__(self, x0, x1):
        self.x0, self.x1 = (x0, x1) if x0 <= x1 else (x1, x0)

    def __contains__(self, z):
        return isinstance(z, (int, float)) and not isinstance(z, bool) and self.x0 <= z <= self.x1

    def __str__(self):
        return 'Continuous range from {} to {}'.format(self.x0, self.x1)


class TypeList(object):
    """List of values of given types with an optional fixed shape (utilities.py:69-124)."""

    def __init__(self, types, dim=()):
        self.types = list(types) if isinstance(types, (list, tuple)) else [types]
        self.dim = tuple(dim)

    def __eq__(self, array):
        try:
            arr = np.array(array)
            if self.dim and tuple(arr.shape) != self.dim:
                return False
            ok = []
            for v in arr.ravel():
                kind = int if np.issubdtype(type(v), np.integer) else (float if np.issubdtype(type(v), np.floating) else type(v))
                ok.append(kind in self.types)
            return all(ok)
        except Exception:
            return False

    def __str__(self):
        return 'Array of {}, dimensions {}'.format(self.types, self.dim or 'arbitrary')


def get_earth_radius(latitude):
    """Earth radius at a latitude (utilities.py:126-139; the trigonometric functions are
    applied to the argument as given, which the refraction code passes in degrees)."""
    a = 6378.1370 * 1000
    b = 6356.7523 * 1000
    c, s = np.cos(latitude), np.sin(latitude)
    return np.sqrt(((a * a * c) ** 2 + (b * b * s) ** 2) / ((a * c) ** 2 + (b * s) ** 2))


def aliasing(v, nyquist):
    """Fold a velocity into the Nyquist interval (utilities.py:142-155)."""
    theta = (v + nyquist) / (2 * nyquist) * np.pi - np.pi / 2.
    return (np.arctan(np.tan(theta)) + np.pi / 2) * (2 * nyquist) / np.pi - nyquist


def vlinspace(a, b, N, endpoint=True):
    a, b = np.asanyarray(a), np.asanyarray(b)
    return a[..., None] + (b - a)[..., None] / (N - endpoint) * np.arange(N)


def nan_cumprod(x):
    x = np.array(x, copy=True)
    x[np.isnan(x)] = 1
    return np.cumprod(x)


def nan_cumsum(x):
    x = np.array(x, copy=True)
    x[np.isnan(x)] = 0
    return np.cumsum(x)


def _pad_pair(x, y, cst):
    x, y = np.asarray(x), np.asarray(y)
    if x.shape == y.shape:
        return x, y
    px, py = [], []
    for d in (np.array(x.shape) - np.array(y.shape)):
        px.append((0, max(-d, 0)))
        py.append((0, max(d, 0)))
    return (np.pad(x, px, 'constant', constant_values=cst), np.pad(y, py, 'constant', constant_values=cst))


def sum_arr(x, y, cst=0):
    x, y = _pad_pair(x, y, cst)
    return np.sum([x, y], axis=0)


def nansum_arr(x, y, cst=0):
    x, y = _pad_pair(x, y, cst)
    return np.nansum([x, y], axis=0)


def combine_subradials(list_of_subradials):
    """Merge the variables of radials defined on the same gates (utilities.py:263-285)."""
    x = list_of_subradials[0]
    for l in list_of_subradials:
        if np.all(np.asarray(l.dist_profile) == np.asarray(x.dist_profile)):
            x.values.update(l.values)
        else:
            print('Beams are not defined on the same set of coordinates, aborting')
            return
    return x


def WGS_to_COSMO(coords, south_pole):
    """Geographic (lat, lon) [deg] -> rotated-pole (rlat, rlon) [deg] for the rotated SOUTH pole
    (lat, lon): rotation about z by the pole longitude, then about y by 90 deg + pole latitude
    (the transform pycosmo.WGS_to_COSMO provides to the reference, interpolation.py:566-568;
    on the device the same rotation is in k_interp_psd)."""
    lat, lon = np.deg2rad(np.asarray(coords[0], dtype=np.float64)), np.deg2rad(np.asarray(coords[1], dtype=np.float64))
    th, ph = np.deg2rad(90. + south_pole[0]), np.deg2rad(south_pole[1])
    x, y, z = np.cos(lon) * np.cos(lat), np.sin(lon) * np.cos(lat), np.sin(lat)
    xn = np.cos(th) * np.cos(ph) * x + np.cos(th) * np.sin(ph) * y + np.sin(th) * z
    yn = -np.sin(ph) * x + np.cos(ph) * y
    zn = -np.sin(th) * np.cos(ph) * x - np.sin(th) * np.sin(ph) * y + np.cos(th) * z
    return np.rad2deg(np.arcsin(zn)), np.rad2deg(np.arctan2(yn, xn))
