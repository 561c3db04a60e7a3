# -*- coding: utf-8 -*-
"""lut.py -- scattering lookup tables: the container class, the on-disk format and the
query rule of the reference (cosmo_pol/lookup/lut.py:27-344).

On-disk format (lut.py:78-154): an uncompressed tar holding ``value_table.npy``
[ne, nt|nwc, nd, 12] float64, ``axes.npy``, ``axes_names.npy`` (pickled dict),
``axes_step.npy`` and ``axes_limits.npy``.

Query rule (``lookup_line``, lut.py:309-344): FLOOR-index slice, no interpolation:
``idx = floor((x - axis_min) / axis_step)`` clamped to the axis.  The CUDA path applies
exactly this rule in float32 (csrc/radop_kernels.cu: ``lut_bin``); the numpy method here
serves host-side users of the class."""
import glob
import os
import pickle
import shutil
import tarfile
import tempfile
from io import BytesIO

import numpy as np
from scipy import ndimage

LUT_FOLDERS = {'tmatrix': 'lut_tmatrix', 'tmatrix_masc': 'lut_tmatrix_masc', 'dda': 'lut_dda'}


class Error(Exception):
    """Lookup Table Error"""


class Lookup_table(object):
    def __init__(self):
        self.axes = []          # independent variable values along each axis
        self.axes_names = {}    # name -> axis index
        self.axes_limits = []   # [min, max] per axis
        self.axes_step = []     # x1 - x0 per axis
        self.axes_len = []
        self.value_table = []
        self.int_transform = False

    def add_axis(self, name, axis_values=None):
        if name in self.axes_names:
            raise Error("Axis already exists with name: '%s'" % name)
        vals = np.asarray(axis_values).astype('float32')
        self.axes_names[name] = len(self.axes)
        self.axes_limits.append([np.min(vals), np.max(vals)])
        self.axes_step.append(vals[1] - vals[0])
        self.axes_len.append(len(vals[0]) if isinstance(vals[0], np.ndarray) else len(vals))
        self.axes.append(vals)

    def set_axis_values(self, axis_name, axis_values):
        i = self.axes_names[axis_name]
        vals = np.asarray(axis_values).astype('float32')
        self.axes_limits[i] = [np.min(vals), np.max(vals)]
        self.axes[i] = vals

    def set_value_table(self, value_table):
        value_table = np.asarray(value_table)
        if value_table.shape != tuple(self.axes_len):
            return ValueError('The shape of the data does not match the length of the axes')
        self.value_table = value_table

    def get_axis_name(self, axis_i):
        for name, i in self.axes_names.items():
            if i == axis_i:
                return name
        return None

    def lookup_pts(self, coords, order=1):
        """Interpolated query (lut.py:276-307; not used on the hot path)."""
        if not len(self.value_table):
            raise Error('No values set for lookup table')
        if coords.shape[0] != len(self.axes_names):
            raise Error('Invalid coordinates dimension')
        coords = [(c - lo) * (n - 1) / (hi - lo)
                  for (lo, hi), c, n in zip(self.axes_limits, coords, self.value_table.shape)]
        return ndimage.map_coordinates(self.value_table, coords, order=order, mode='constant', cval=np.nan)

    def bin_index(self, axis_name, values):
        """The floor-index rule for one axis, evaluated in float32."""
        i = self.axes_names[axis_name]
        x = np.asarray(values, dtype=np.float32)
        idx = np.floor((x - np.float32(self.axes_limits[i][0])) / np.float32(self.axes_step[i]))
        idx = np.array(idx, dtype=int)
        return np.clip(idx, 0, self.value_table.shape[i] - 1)

    def lookup_line(self, **kwargs):
        sel = [slice(None)] * self.value_table.ndim
        for k, v in kwargs.items():
            if k in self.axes_names:
                sel[self.axes_names[k]] = self.bin_index(k, v)
        return self.value_table[tuple(sel)]


def save_lut(lut, filename):
    tmp_dir = os.path.join(tempfile.mkdtemp(), os.path.basename(filename).split('.')[0])
    os.makedirs(tmp_dir)
    np.save(os.path.join(tmp_dir, 'value_table'), lut.value_table)
    for name in ('axes', 'axes_step', 'axes_limits'):
        obj = getattr(lut, name)
        try:
            np.save(os.path.join(tmp_dir, name), np.asarray(obj))
        except Exception:
            np.save(os.path.join(tmp_dir, name), np.asarray(obj, dtype=object), allow_pickle=True)
    np.save(os.path.join(tmp_dir, 'axes_names'), np.array(dict(lut.axes_names), dtype=object), allow_pickle=True)
    with tarfile.open(filename, 'w') as tar:
        for name in sorted(glob.glob(os.path.join(tmp_dir, '*.npy'))):
            tar.add(name, arcname=os.path.basename(name))
    shutil.rmtree(os.path.dirname(tmp_dir))


def load_lut(filename):
    lut = Lookup_table()
    with tarfile.open(filename, 'r') as tar:
        for member in tar.getmembers():
            buf = BytesIO(tar.extractfile(member).read())
            name = member.name.replace('.npy', '')
            try:
                data = np.load(buf, allow_pickle=True, encoding='latin1')
            except Exception:
                buf.seek(0)
                data = pickle.load(buf)
            if name == 'axes_names':
                data = data.item() if isinstance(data, np.ndarray) else data   # 0-d object array -> dict
            setattr(lut, name, data)
    return lut


def load_all_lut(scheme, list_hydrom, frequency, scattering_method, folder=None):
    """Load ``lut_SZ_<hydrom>_<freq>_<scheme>.lut`` for every species (lut.py:27-76).
    ``folder`` overrides the package's lookup directory."""
    base = folder if folder is not None else os.path.dirname(os.path.realpath(__file__))
    sub = LUT_FOLDERS.get(scattering_method, LUT_FOLDERS['tmatrix_masc'])
    default = os.path.join(base, LUT_FOLDERS['tmatrix_masc'])
    freq_str = str(frequency).replace('.', '_')
    out = {}
    for h in list_hydrom:
        name = 'lut_SZ_' + h + '_' + freq_str + '_' + scheme + '.lut'
        where = default if (scattering_method == 'dda' and h in ('R', 'H')) else os.path.join(base, sub)
        path = os.path.join(where, name)
        if not os.path.exists(path) and folder is not None and os.path.exists(os.path.join(folder, name)):
            path = os.path.join(folder, name)
        try:
            out[h] = load_lut(path)
        except Exception:
            raise IOError('Could not find lookup table for scheme = {:s}, hydrometeor = {:s}, frequency = {} '
                          'and scattering method = {:s} ({:s})'.format(scheme, h, frequency, scattering_method, path))
    return out
