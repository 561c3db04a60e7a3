# -*- coding: utf-8 -*-
"""_abi.py -- ctypes binding of libradop.so (include/radop.h) and the ``Engine`` that the
Python host layer drives.  There is NO CPU fallback: a missing library or a missing GPU
raises, loudly."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# RADOP_LIB: another build of the same sources (launch-shape experiments, scripts/build_variants.sh)
LIB_PATH = os.environ.get('RADOP_LIB') or os.path.join(_HERE, 'libradop.so')

RADOP_ABI_VERSION = 6
NSPECIES_MAX = 7
NVAR_MAX = 16
N_OUT = 10
OUT_NAMES = ['ZH', 'ZDR', 'ZV', 'KDP', 'DELTA_HV', 'PHIDP', 'RHOHV', 'ATT_H', 'ATT_V', 'RVEL']
VAR_IDS = {'U': 0, 'V': 1, 'W': 2, 'T': 3, 'QR_v': 4, 'QS_v': 5, 'QG_v': 6, 'QI_v': 7, 'RHO': 8, 'QH_v': 9,
           'QNR_v': 10, 'QNS_v': 11, 'QNG_v': 12, 'QNI_v': 13, 'QNH_v': 14, 'EDR': 15}
VAR_NAMES = {v: k for k, v in VAR_IDS.items()}
# one-moment slots that carry the melting re-partition when with_melting is set (melting.py:74-90)
MELT_VAR_SLOTS = {'QmS_v': 9, 'QmG_v': 10, 'fwet_mS': 11, 'fwet_mG': 12}
SPECIES_IDS = {'R': 0, 'S': 1, 'G': 2, 'H': 3, 'I': 4, 'mS': 5, 'mG': 6}

EXPORTS = ['radop_abi_version', 'radop_create', 'radop_destroy', 'radop_last_error', 'radop_set_params',
           'radop_upload_grid', 'radop_upload_field', 'radop_upload_lut', 'radop_simulate', 'radop_stage_scan',
           'radop_execute', 'radop_fetch', 'radop_synchronize', 'radop_device_outputs', 'radop_get_stream',
           'radop_get_stats', 'radop_enable_timing', 'radop_debug_indices', 'radop_measure_fp32_peak',
           'radop_get_all_radar_pts', 'radop_selftest_division', 'radop_set_chunk', 'radop_set_overlap_threshold',
           'radop_set_geodesy', 'radop_get_geodesy', 'radop_set_integration', 'radop_get_integration', 'radop_debug_integration',
           'radop_set_search', 'radop_sizeof', 'radop_device_buffer', 'radop_get_refl', 'radop_get_refl_noalloc', 'radop_set_output_blocks', 'radop_host_register', 'radop_host_unregister', 'radop_set_pipeline', 'radop_get_pipeline', 'radop_set_tuning']

BUF_OBS, BUF_LAT, BUF_LON, BUF_MODEL, BUF_DSPECTRUM = range(5)
NOALLOC_SCRATCH_BYTES = 4 << 20      # RADOP_NOALLOC_SCRATCH_BYTES

(RADOP_OK, RADOP_ERR_CUDA, RADOP_ERR_ARG, RADOP_ERR_OUT_OF_DOMAIN, RADOP_ERR_LUT, RADOP_ERR_STATE,
 RADOP_ERR_UNSUPPORTED) = range(7)


class RadopSpecies(C.Structure):
    _fields_ = [('species', C.c_int32), ('psd', C.c_int32), ('numeric_vn', C.c_int32), ('reserved', C.c_int32),
                ('a', C.c_double), ('b', C.c_double), ('alpha', C.c_double), ('beta', C.c_double),
                ('mu', C.c_double), ('nu', C.c_double), ('N0', C.c_double), ('lambda_factor', C.c_double),
                ('vel_factor', C.c_double), ('ntot_factor', C.c_double), ('x_min', C.c_double), ('x_max', C.c_double),
                ('melt', C.c_double * 8)]


class RadopParams(C.Structure):
    _fields_ = [('abi_version', C.c_int32), ('n_gates', C.c_int32),
                ('radar_lat', C.c_double), ('radar_lon', C.c_double), ('radar_alt', C.c_double),
                ('wavelength_mm', C.c_double), ('k_squared', C.c_double), ('radial_resolution_m', C.c_double),
                ('two_moment', C.c_int32), ('with_attenuation', C.c_int32), ('doppler_scheme', C.c_int32),
                ('sensitivity_mode', C.c_int32), ('sensitivity', C.c_double * 3),
                ('output_model_vars', C.c_int32), ('keep_debug', C.c_int32), ('n_species', C.c_int32),
                ('reserved', C.c_int32), ('species', RadopSpecies * NSPECIES_MAX),
                ('n_vel', C.c_int32), ('reserved2', C.c_int32), ('vel_min', C.c_double), ('vel_res', C.c_double),
                ('spectrum_turbulence', C.c_int32), ('spectrum_motion', C.c_int32),
                ('antenna_speed', C.c_double), ('beamwidth_deg', C.c_double),
                ('with_melting', C.c_int32), ('reserved3', C.c_int32)]


class RadopScan(C.Structure):
    _fields_ = [('n_rays', C.c_int32), ('n_sub', C.c_int32), ('n_traj', C.c_int32), ('n_vert', C.c_int32),
                ('ray_az', C.c_void_p), ('ray_traj', C.c_void_p), ('ray_nyquist', C.c_void_p),
                ('sub_daz', C.c_void_p), ('sub_jv', C.c_void_p), ('sub_weight', C.c_void_p),
                ('traj_s', C.c_void_p), ('traj_h', C.c_void_p), ('traj_e', C.c_void_p),
                ('traj_sin', C.c_void_p), ('traj_tan', C.c_void_p), ('sub_ml_only', C.c_void_p)]


class RadopOut(C.Structure):
    _fields_ = [('obs', C.c_void_p * N_OUT), ('mask', C.c_void_p), ('lat', C.c_void_p), ('lon', C.c_void_p),
                ('model', C.c_void_p * NVAR_MAX), ('dspectrum', C.c_void_p)]


class RadopStats(C.Structure):
    _fields_ = [('n_subbeam_gates', C.c_int64), ('n_active_pairs', C.c_int64), ('n_lut_bins', C.c_int64),
                ('n_tiles', C.c_int64), ('launches', C.c_int64), ('ms_interp', C.c_float), ('ms_sort', C.c_float),
                ('ms_integrate', C.c_float), ('ms_reduce', C.c_float), ('ms_total', C.c_float),
                ('n_timed', C.c_int32), ('ms_spectrum', C.c_float),
                ('pairs_by_species', C.c_int64 * NSPECIES_MAX), ('exec_flop_per_pair', C.c_double * NSPECIES_MAX)]


# order of RADOP_SIZEOF_* in include/radop.h
ABI_STRUCTS = (RadopParams, RadopSpecies, RadopScan, RadopOut, RadopStats)


class RadopError(RuntimeError):
    pass


_lib = None


def load_library():
    """dlopen libradop.so; raises if it has not been built (python __graft_entry__.py / build())."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RadopError('libradop.so is missing: build it with `python -c "import __graft_entry__ as g; g.build()"` '
                         '(nvcc, sm_100a). There is no CPU fallback.')
    lib = C.CDLL(LIB_PATH)
    for name in EXPORTS:
        getattr(lib, name)           # AttributeError if an export is missing
        getattr(lib, name).restype = C.c_int
    lib.radop_last_error.restype = C.c_char_p
    lib.radop_last_error.argtypes = [C.c_void_p]
    lib.radop_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    lib.radop_destroy.argtypes = [C.c_void_p]
    lib.radop_set_params.argtypes = [C.c_void_p, C.POINTER(RadopParams)]
    lib.radop_upload_grid.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p]
    lib.radop_upload_field.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    lib.radop_upload_lut.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float,
                                     C.c_float, C.c_float, C.c_void_p]
    lib.radop_simulate.argtypes = [C.c_void_p, C.POINTER(RadopScan), C.POINTER(RadopOut)]
    lib.radop_stage_scan.argtypes = [C.c_void_p, C.POINTER(RadopScan)]
    lib.radop_execute.argtypes = [C.c_void_p]
    lib.radop_fetch.argtypes = [C.c_void_p, C.POINTER(RadopOut)]
    lib.radop_synchronize.argtypes = [C.c_void_p]
    lib.radop_device_outputs.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    lib.radop_get_stream.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
    lib.radop_get_stats.argtypes = [C.c_void_p, C.POINTER(RadopStats)]
    lib.radop_enable_timing.argtypes = [C.c_void_p, C.c_int]
    lib.radop_debug_indices.argtypes = [C.c_void_p] + [C.c_void_p] * 7
    lib.radop_measure_fp32_peak.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
    lib.radop_set_chunk.argtypes = [C.c_void_p, C.c_int64]
    lib.radop_set_overlap_threshold.argtypes = [C.c_void_p, C.c_int64]
    lib.radop_set_geodesy.argtypes = [C.c_void_p, C.c_int]
    lib.radop_get_geodesy.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
    lib.radop_set_integration.argtypes = [C.c_void_p, C.c_int]
    lib.radop_set_search.argtypes = [C.c_void_p, C.c_int]
    lib.radop_debug_integration.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int64), C.c_void_p, C.c_int64]
    lib.radop_get_integration.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                          C.POINTER(C.c_int64), C.POINTER(C.c_float)]
    lib.radop_selftest_division.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64)]
    lib.radop_get_all_radar_pts.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.radop_set_pipeline.argtypes = [C.c_void_p, C.c_int]
    lib.radop_get_pipeline.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
    lib.radop_set_tuning.argtypes = [C.c_void_p, C.c_int, C.c_int64]
    lib.radop_sizeof.argtypes = [C.c_int]
    lib.radop_device_buffer.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
    lib.radop_get_refl.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 8
    lib.radop_set_output_blocks.argtypes = [C.c_void_p, C.c_int, C.c_int]
    lib.radop_host_register.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
    lib.radop_host_unregister.argtypes = [C.c_void_p, C.c_void_p]
    lib.radop_get_refl_noalloc.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 8
    if lib.radop_abi_version() != RADOP_ABI_VERSION:
        raise RadopError('libradop.so ABI version mismatch')
    # the struct layouts of this binding against the ones the library was compiled with (radop_sizeof):
    # a short RadopParams would make radop_set_params read past the caller's struct
    for which, st in enumerate(ABI_STRUCTS):
        if lib.radop_sizeof(which) != C.sizeof(st):
            raise RadopError('libradop.so: sizeof({}) is {} in the library and {} in this binding'.format(
                st.__name__, lib.radop_sizeof(which), C.sizeof(st)))
    _lib = lib
    return lib


def _ptr(a):
    return C.c_void_p(a.ctypes.data) if a is not None else C.c_void_p(None)


class Scan(object):
    """Host arrays of one scan (RadopScan) plus bookkeeping for the result assembly."""

    def __init__(self, ray_az, ray_el, ray_traj, ray_nyquist, fan, traj_s, traj_h, traj_e):
        self.ray_az = np.ascontiguousarray(ray_az, dtype=np.float64)
        self.ray_el = np.ascontiguousarray(ray_el, dtype=np.float64)
        self.ray_traj = np.ascontiguousarray(ray_traj, dtype=np.int32)
        self.ray_nyquist = None if ray_nyquist is None else np.ascontiguousarray(ray_nyquist, dtype=np.float32)
        self.fan = fan
        self.sub_daz = np.ascontiguousarray(fan.daz, dtype=np.float64)
        self.sub_jv = np.ascontiguousarray(fan.j_ver, dtype=np.int32)
        self.sub_weight = np.ascontiguousarray(fan.weight, dtype=np.float64)
        ml = np.ascontiguousarray(getattr(fan, 'ml_only', np.zeros(fan.n_sub)), dtype=np.int32)
        self.sub_ml_only = ml if ml.any() else None
        self.traj_s = np.ascontiguousarray(traj_s, dtype=np.float32)
        self.traj_h = np.ascontiguousarray(traj_h, dtype=np.float32)
        self.traj_e = np.ascontiguousarray(traj_e, dtype=np.float32)
        # float32 sin / tan of the folded elevation exactly as numpy forms them in
        # get_diameter_from_rad_vel (doppler_scatter.py:571-575); read by Doppler scheme 3 only
        ef = self.traj_e.copy()
        ef[ef > 90] = 180 - ef[ef > 90]
        ef[ef < 0] = -ef[ef < 0]
        th = np.deg2rad(ef)
        self.traj_sin = np.ascontiguousarray(np.sin(th), dtype=np.float32)
        self.traj_tan = np.ascontiguousarray(np.tan(th), dtype=np.float32)
        self.n_rays = len(self.ray_az)
        self.n_sub = fan.n_sub
        self.n_traj, self.n_vert, self.n_gates = self.traj_s.shape
        self._geom_key = None
        self._center = None

    def geometry_key(self):
        """Digest of everything the gate positions depend on (ray list, fan, trajectories)."""
        if self._geom_key is None:
            import hashlib
            hsh = hashlib.blake2b(digest_size=16)
            for a in (self.ray_az, self.ray_traj, self.sub_daz, self.sub_jv, self.traj_s):
                hsh.update(a.tobytes())
            self._geom_key = hsh.hexdigest()
        return self._geom_key

    def center_profiles(self):
        """(ground distance, height) of the central sub-beam at every gate of every ray, [n_rays, n_gates]: geometry
        only, made once per scan (Radial.dist_profile / heights_profile of the central sub-radial, radial.py:40-44)."""
        if self._center is None:
            j = self.fan.j_ver[self.fan.center]
            self._center = (self.traj_s[self.ray_traj, j], self.traj_h[self.ray_traj, j])
        return self._center

    def as_struct(self):
        s = RadopScan()
        s.n_rays, s.n_sub, s.n_traj, s.n_vert = self.n_rays, self.n_sub, self.n_traj, self.n_vert
        s.ray_az, s.ray_traj, s.ray_nyquist = _ptr(self.ray_az), _ptr(self.ray_traj), _ptr(self.ray_nyquist)
        s.sub_daz, s.sub_jv, s.sub_weight = _ptr(self.sub_daz), _ptr(self.sub_jv), _ptr(self.sub_weight)
        s.traj_s, s.traj_h, s.traj_e = _ptr(self.traj_s), _ptr(self.traj_h), _ptr(self.traj_e)
        s.traj_sin, s.traj_tan = _ptr(self.traj_sin), _ptr(self.traj_tan)
        s.sub_ml_only = _ptr(self.sub_ml_only)
        return s

    @property
    def n_subbeam_gates(self):
        return self.n_rays * self.n_sub * self.n_gates

    def bytes_in(self):
        return sum(a.nbytes for a in (self.ray_az, self.ray_traj, self.sub_daz, self.sub_jv, self.sub_weight,
                                      self.traj_s, self.traj_h, self.traj_e, self.traj_sin, self.traj_tan)) + (
            self.ray_nyquist.nbytes if self.ray_nyquist is not None else 0)


class Engine(object):
    """One handle = one GPU.  Owns the device-resident model fields and lookup tables."""

    def __init__(self, device=0):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.radop_create(int(device), C.byref(h))
        if rc != RADOP_OK or not h:
            raise RadopError('radop_create failed on device {} (code {}): no usable CUDA GPU. '
                             'The hot path has no CPU fallback.'.format(device, rc))
        self.h = h
        self.device = device
        self.params = None
        self.uploaded_vars = []
        self._pinned = {}
        self._keep = []
        self._latlon_key = None      # geometry the pinned lat / lon buffers currently hold
        self._staged = None          # (Scan object, parameter epoch) the library currently has staged
        self.reuse_staged = True     # False: upload the scan with every call (bench.py: inputs travel every step)
        self._epoch = 0
        self.cache_latlon = True

    def close(self):
        if getattr(self, 'h', None):
            self.lib.radop_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------------------------
    def _check(self, rc):
        if rc == RADOP_OK:
            return
        msg = self.lib.radop_last_error(self.h)
        msg = msg.decode() if msg else ''
        if rc == RADOP_ERR_OUT_OF_DOMAIN:
            # same exception type as interpolation.py:572-580
            raise IndexError('ERROR: RADAR DOMAIN IS NOT ENTIRELY CONTAINED IN COSMO SIMULATION DOMAIN: ABORTING (' + msg + ')')
        if rc in (RADOP_ERR_ARG, RADOP_ERR_UNSUPPORTED):
            raise ValueError(msg)
        raise RadopError('libradop error {}: {}'.format(rc, msg))

    # ------------------------------------------------------------------------------------
    def set_params(self, config, constants, hydrometeors, n_gates, output_model_vars=False, keep_debug=False,
                   sensitivity_cut=True):
        """config: cfg.CONFIG-like dict; constants: global_constants; hydrometeors: ordered
        list of species objects (summation order of doppler_scatter.py:99-106)."""
        p = RadopParams()
        p.abi_version = RADOP_ABI_VERSION
        p.n_gates = int(n_gates)
        r = config['radar']
        p.radar_lat, p.radar_lon, p.radar_alt = float(r['coords'][0]), float(r['coords'][1]), float(r['coords'][2])
        p.wavelength_mm = float(constants.WAVELENGTH)
        p.k_squared = float(r['K_squared'])
        p.radial_resolution_m = float(r['radial_resolution'])
        p.two_moment = 1 if config['microphysics']['scheme'] == '2mom' else 0
        p.with_attenuation = int(config['microphysics']['with_attenuation'])
        p.doppler_scheme = int(config['doppler']['scheme'])
        p.with_melting = 1 if config['microphysics'].get('with_melting', 0) else 0
        if p.doppler_scheme == 3:
            d = config['doppler']
            # broadening of the sub-beam spectra (doppler_scatter.py:357-367, 727-801)
            p.spectrum_turbulence = 1 if d.get('turbulence_correction', 0) else 0
            p.spectrum_motion = 1 if d.get('motion_correction', 0) else 0
            p.antenna_speed = float(r.get('antenna_speed', 0.2))
            p.beamwidth_deg = float(r['3dB_beamwidth'])
            # constants.VARRAY = np.arange(-NVEL, NVEL + VRES, VRES), global_constants.py:168-171
            p.n_vel = len(constants.VARRAY)
            p.vel_min, p.vel_res = float(-constants.NVEL), float(constants.VRES)
        self.n_vel = int(p.n_vel)
        sens = r['sensitivity']
        sens = list(sens) if isinstance(sens, (list, tuple)) else [sens]
        if not sensitivity_cut:
            p.sensitivity_mode = 0
        elif len(sens) in (1, 2):
            p.sensitivity_mode = len(sens)
            for i, v in enumerate(sens):
                p.sensitivity[i] = float(v)
        else:
            raise ValueError('3-element sensitivity [dBm, gain, snr] is not supported by the GPU path')
        p.output_model_vars = 1 if output_model_vars else 0
        p.keep_debug = 1 if keep_debug else 0
        p.n_species = len(hydrometeors)
        for i, hy in enumerate(hydrometeors):
            for k, v in hy.abi_fields().items():
                if k == 'melt':
                    for j, x in enumerate(v):
                        p.species[i].melt[j] = x
                else:
                    setattr(p.species[i], k, v)
        self._check(self.lib.radop_set_params(self.h, C.byref(p)))
        self._epoch += 1             # trajectories and gate count belong to the parameters: the next call stages again
        self.params = p
        self.species_order = [hy.kind for hy in hydrometeors]

    def upload_model(self, dic_vars):
        first = next(iter(dic_vars.values()))
        z = np.ascontiguousarray(first.attributes['z-levels'], dtype=np.float32)
        proj = first.attributes['proj_info']
        res = first.attributes['resolution']
        llc = np.asarray((float(proj['Lo1']), float(proj['La1']))).astype('float32')
        urc = np.asarray((float(proj['Lo2']), float(proj['La2']))).astype('float32')
        res32 = np.asarray(res, dtype=np.float32)
        sp = np.asarray([float(proj['Latitude_of_southern_pole']), float(proj['Longitude_of_southern_pole'])],
                        dtype=np.float64)
        nz, ny, nx = z.shape
        self._check(self.lib.radop_upload_grid(self.h, nz, ny, nx, _ptr(z), _ptr(llc), _ptr(urc), _ptr(res32), _ptr(sp)))
        self.uploaded_vars = []
        for name, var in dic_vars.items():
            if name not in VAR_IDS:
                continue
            d = np.ascontiguousarray(var.data, dtype=np.float32)
            if d.shape != z.shape:
                raise ValueError('variable {} does not have the shape of the z-levels'.format(name))
            self._check(self.lib.radop_upload_field(self.h, VAR_IDS[name], _ptr(d)))
            self.uploaded_vars.append(name)
        self.uploaded_vars.sort(key=lambda n: VAR_IDS[n])
        self.grid_shape = (nz, ny, nx)

    def upload_luts(self, lut_sz):
        for h, lut in lut_sz.items():
            if h not in SPECIES_IDS:
                continue
            if h not in self.species_order:
                continue
            vt = np.ascontiguousarray(lut.value_table, dtype=np.float64)
            names = lut.axes_names
            second = 'wc' if h in ('mS', 'mG') else 't'       # wet fraction axis of the melting species
            ie, it, idd = names['e'], names[second], names['d']
            if (ie, it, idd) != (0, 1, 2) or vt.ndim != 4 or vt.shape[3] != 12:
                raise ValueError('lookup table of {} must have axes (e, {}, d, sz[12])'.format(h, second))
            d_axis = np.asarray(lut.axes[idd], dtype=np.float32)
            if d_axis.ndim == 2:                               # [n_wc, nd]: per-gate grids on the device
                d_axis = d_axis[0]
            d_axis = np.ascontiguousarray(d_axis)
            self._check(self.lib.radop_upload_lut(
                self.h, SPECIES_IDS[h], _ptr(vt), vt.shape[0], vt.shape[1], vt.shape[2],
                C.c_float(np.float32(lut.axes_limits[ie][0])), C.c_float(np.float32(lut.axes_step[ie])),
                C.c_float(np.float32(lut.axes_limits[it][0])), C.c_float(np.float32(lut.axes_step[it])), _ptr(d_axis)))

    # ------------------------------------------------------------------------------------
    def _host_buffer(self, key, shape, dtype):
        """Pinned host buffer (through torch when CUDA is there), cached per key."""
        n = int(np.prod(shape))
        ent = self._pinned.get(key)
        if ent is not None and ent[1].size == n and ent[1].dtype == np.dtype(dtype):
            return ent[1].reshape(shape)
        arr = None
        try:
            import torch
            if torch.cuda.is_available():
                t = torch.empty(n, dtype={np.dtype('float32'): torch.float32, np.dtype('float64'): torch.float64}[np.dtype(dtype)],
                                pin_memory=True)
                arr = t.numpy()
                self._pinned[key] = (t, arr)
        except Exception:
            arr = None
        if arr is None:
            arr = np.empty(n, dtype=dtype)
            self._pinned[key] = (None, arr)
        return arr.reshape(shape)

    def make_out(self, scan, want_latlon=True, want_model=False):
        shape = (scan.n_rays, scan.n_gates)
        out = RadopOut()
        arrays = {}
        # observables and mask in ONE pinned block: equally spaced planes let the library copy a chunk of all of them
        # with a single strided device -> host copy (copy_outputs)
        block = self._host_buffer('obs+mask', (len(OUT_NAMES) + 1,) + shape, np.float32)
        for i, name in enumerate(OUT_NAMES):
            arrays[name] = block[i]
            out.obs[i] = block[i].ctypes.data
        arrays['mask'] = block[len(OUT_NAMES)]
        out.mask = arrays['mask'].ctypes.data
        if want_latlon:
            for k in ('lat', 'lon'):
                a = self._host_buffer(k, shape, np.float64)
                arrays[k] = a
                setattr(out, k, a.ctypes.data)
        if want_model:
            for name, slot in self.model_output_slots().items():
                a = self._host_buffer(('model', name), shape, np.float32)
                arrays[name] = a
                out.model[slot] = a.ctypes.data
        if self.params is not None and self.params.doppler_scheme == 3:
            a = self._host_buffer('dspectrum', shape + (self.n_vel,), np.float32)
            arrays['DSPECTRUM'] = a
            out.dspectrum = a.ctypes.data
        return out, arrays

    def _geometry_key(self, scan):
        p = self.params
        return (scan.geometry_key(), p.radar_lat, p.radar_lon, p.n_gates, scan.n_rays)

    def model_output_slots(self):
        """name -> slot of every variable integrate_radials returns (interpolation.py:36-89): the uploaded model
        variables and, with the melting scheme, the re-partitioned masses and wet fractions."""
        slots = {n: VAR_IDS[n] for n in self.uploaded_vars}
        if self.params is not None and self.params.with_melting:
            slots.update(MELT_VAR_SLOTS)
        return slots

    def simulate(self, scan, want_latlon=True, want_model=False):
        """The reference-facing call: host in -> host out.  The gate positions (lat, lon: 16 of the 60 bytes
        per ray gate) depend on the scan geometry only; when the pinned buffers already hold the positions of
        this very ray list, fan and trajectories, they are neither recomputed nor copied again."""
        out, arrays = self.make_out(scan, want_latlon, want_model)
        key = self._geometry_key(scan) if want_latlon else None
        reuse = want_latlon and self.cache_latlon and key == self._latlon_key
        if reuse:
            out.lat = None
            out.lon = None
        self._latlon_key = None if (want_latlon and not reuse) else self._latlon_key
        self._check(self.lib.radop_simulate(self.h, self._scan_arg(scan), C.byref(out)))
        self._staged = (scan, self._epoch)
        if want_latlon:
            self._latlon_key = key
        self.last_d2h_bytes = int(sum(a.nbytes for k, a in arrays.items() if not (reuse and k in ('lat', 'lon'))))
        return arrays

    def _scan_arg(self, scan):
        """The RadopScan to hand to radop_simulate: NULL when the library already has this very scan staged (the Scan
        objects of RadarOperator.build_scan are cached per ray list, so a repeated scan strategy is staged once)."""
        if self.reuse_staged and self._staged is not None and self._staged[0] is scan and self._staged[1] == self._epoch:
            return None
        self._staged = None
        self._keep = scan.as_struct()
        return C.byref(self._keep)

    def host_register(self, array):
        """Page-lock a host array the caller owns (the ranks' shared-memory segment); returns False where it fails."""
        try:
            self._check(self.lib.radop_host_register(self.h, C.c_void_p(array.ctypes.data), C.c_int64(array.nbytes)))
            return True
        except RadopError:
            return False

    def host_unregister(self, array):
        self.lib.radop_host_unregister(self.h, C.c_void_p(array.ctypes.data))

    def simulate_into(self, scan, arrays, rank, world, block, want_latlon=True):
        """radop_simulate of this rank's rays straight into host arrays that hold the WHOLE scan (rows = rays of all
        ranks, dealt in blocks of `block`; radop_set_output_blocks).  arrays: name -> [n_total_rays, ...] for the
        observable names, 'mask', 'lat', 'lon', 'DSPECTRUM' and model variables (any subset)."""
        out = RadopOut()
        row0 = rank * block

        def at(a):
            return a[row0:].ctypes.data if a is not None else None
        for i, name in enumerate(OUT_NAMES):
            if name in arrays:
                out.obs[i] = at(arrays[name])
        out.mask = at(arrays.get('mask'))
        if want_latlon:
            out.lat, out.lon = at(arrays.get('lat')), at(arrays.get('lon'))
        if 'DSPECTRUM' in arrays:
            out.dspectrum = at(arrays['DSPECTRUM'])
        for name, slot in self.model_output_slots().items():
            if name in arrays:
                out.model[slot] = at(arrays[name])
        self._check(self.lib.radop_set_output_blocks(self.h, int(block), int(block * world)))
        try:
            self._check(self.lib.radop_simulate(self.h, self._scan_arg(scan), C.byref(out)))
            self._staged = (scan, self._epoch)
        finally:
            self.lib.radop_set_output_blocks(self.h, 0, 0)
        self._latlon_key = None        # the pinned buffers of make_out do not hold this scan's positions
        self.last_d2h_bytes = int(sum(scan.n_rays * (a.nbytes // max(1, a.shape[0])) for k, a in arrays.items()
                                      if want_latlon or k not in ('lat', 'lon')))

    def stage(self, scan):
        s = scan.as_struct()
        self._staged = None
        self._check(self.lib.radop_stage_scan(self.h, C.byref(s)))
        self._staged = (scan, self._epoch)

    def execute(self):
        self._check(self.lib.radop_execute(self.h))

    def synchronize(self):
        self._check(self.lib.radop_synchronize(self.h))

    def fetch(self, scan, want_latlon=True, want_model=False):
        out, arrays = self.make_out(scan, want_latlon, want_model)
        self._latlon_key = None
        self._check(self.lib.radop_fetch(self.h, C.byref(out)))
        if want_latlon:
            self._latlon_key = self._geometry_key(scan)
        return arrays

    def finish(self):
        """Drain the stream and raise what the run flagged (IndexError when a gate left the model domain)."""
        self._check(self.lib.radop_fetch(self.h, None))

    def device_buffer(self, which):
        """(device pointer, bytes) of a device-resident result of the last execute (BUF_*)."""
        p, n = C.c_void_p(), C.c_int64()
        self._check(self.lib.radop_device_buffer(self.h, C.c_int(which), C.byref(p), C.byref(n)))
        return p.value, n.value

    def enable_timing(self, on=True):
        self._check(self.lib.radop_enable_timing(self.h, 1 if on else 0))

    def stats(self):
        s = RadopStats()
        self._check(self.lib.radop_get_stats(self.h, C.byref(s)))
        d = {k: getattr(s, k) for k, _ in RadopStats._fields_}
        d['pairs_by_species'] = [int(x) for x in s.pairs_by_species]
        d['exec_flop_per_pair'] = [float(x) for x in s.exec_flop_per_pair]
        return d

    def stream_ptr(self):
        p = C.c_void_p()
        self._check(self.lib.radop_get_stream(self.h, C.byref(p)))
        return p.value

    def device_outputs(self):
        p = C.c_void_p()
        nr, ng = C.c_int64(), C.c_int64()
        self._check(self.lib.radop_device_outputs(self.h, C.byref(p), C.byref(nr), C.byref(ng)))
        return p.value, nr.value, ng.value

    def set_chunk(self, subbeam_gates_per_chunk):
        self._check(self.lib.radop_set_chunk(self.h, C.c_int64(int(subbeam_gates_per_chunk))))

    def set_geodesy(self, mode):
        """0: per sub-beam polynomials (default; per-gate solution when they do not converge), 1: per gate."""
        self._check(self.lib.radop_set_geodesy(self.h, C.c_int(int(mode))))

    def geodesy_used_polynomials(self):
        v = C.c_int(0)
        self._check(self.lib.radop_get_geodesy(self.h, C.byref(v)))
        return bool(v.value)

    def set_search(self, mode):
        """Level search of the interpolation: 0 guessed window with bisection as the fallback (default), 1 bisection."""
        self._check(self.lib.radop_set_search(self.h, C.c_int(int(mode))))

    def set_pipeline(self, mode):
        """'fused' / 0 (default): k_gate_fused whenever the configuration allows it; 'staged' / 1: the staged pipeline
        (k_interp_psd -> k_integrate_tab / k_integrate -> k_reduce_subbeams) always."""
        mode = {'fused': 0, 'staged': 1}.get(mode, mode)
        self._check(self.lib.radop_set_pipeline(self.h, C.c_int(int(mode))))

    def set_tuning(self, key, value):
        self._check(self.lib.radop_set_tuning(self.h, C.c_int({'fused_warps': 0}.get(key, key)), C.c_int64(int(value))))

    def used_fused(self):
        v = C.c_int(0)
        self._check(self.lib.radop_get_pipeline(self.h, C.byref(v)))
        return bool(v.value)

    def set_integration(self, mode):
        """How the diameter integral (get_N x lookup_line, doppler_scatter.py:246-268) is evaluated.
        'table' / 0 (default): lambda-transformed tables, verified piece by piece against the direct sum;
        'direct' / 1: sums over the diameters."""
        mode = {'table': 0, 'direct': 1}.get(mode, mode)
        self._check(self.lib.radop_set_integration(self.h, C.c_int(int(mode))))

    def integration_info(self):
        m, n, r, d, ms = C.c_int(0), C.c_int64(0), C.c_int64(0), C.c_int64(0), C.c_float(0)
        self._check(self.lib.radop_get_integration(self.h, C.byref(m), C.byref(n), C.byref(r), C.byref(d), C.byref(ms)))
        why = (C.c_int64 * 4)()
        self._check(self.lib.radop_debug_integration(self.h, C.c_int(0), why, None, C.c_int64(0)))
        return {'mode': 'table' if m.value == 0 else 'direct', 'pieces': int(n.value), 'rejected_pieces': int(r.value),
                'direct_pairs': int(d.value), 'build_ms': float(ms.value),
                'direct_pairs_by_reason': dict(zip(('rejected_piece', 'below_range', 'above_range', 'not_normal'),
                                                   [int(v) for v in why]))}

    def lamtab_piece_status(self, species, n_bins):
        """[n_bins, 176] exponent word of the first column per (bin, piece); INT32_MIN marks a rejected piece."""
        out = np.zeros(n_bins * 176, dtype=np.int32)
        self._check(self.lib.radop_debug_integration(self.h, C.c_int(SPECIES_IDS[species]), None,
                                                     out.ctypes.data_as(C.c_void_p), C.c_int64(out.size)))
        return out.reshape(n_bins, 176)

    def set_overlap_threshold(self, min_output_bytes):
        """simulate() chunks the scan only when the outputs are at least this large (0: always)."""
        self._check(self.lib.radop_set_overlap_threshold(self.h, C.c_int64(int(min_output_bytes))))

    def selftest_division(self, n_samples=1 << 26, seed=1):
        bad = C.c_uint64(0)
        self._check(self.lib.radop_selftest_division(self.h, C.c_uint64(n_samples), C.c_uint64(seed), C.byref(bad)))
        return int(bad.value)

    def measure_fp32_peak(self):
        v = C.c_double()
        self._check(self.lib.radop_measure_fp32_peak(self.h, C.byref(v)))
        return v.value

    def debug_indices(self, scan):
        n = scan.n_subbeam_gates
        nsp = self.params.n_species
        nv = len(self.uploaded_vars)
        d = {'cell': np.zeros((n, 2), np.int32), 'level': np.zeros((n, 4), np.int32), 'status': np.zeros(n, np.int32),
             'lut_e': np.zeros((n, nsp), np.int32), 'lut_t': np.zeros((n, nsp), np.int32),
             'values': np.zeros((n, nv), np.float32), 'rot': np.zeros((n, 2), np.float64)}
        self._check(self.lib.radop_debug_indices(self.h, _ptr(d['cell']), _ptr(d['level']), _ptr(d['status']),
                                                 _ptr(d['lut_e']), _ptr(d['lut_t']), _ptr(d['values']), _ptr(d['rot'])))
        shp = (scan.n_rays, scan.n_sub, scan.n_gates)
        for k in d:
            d[k] = d[k].reshape(shp + d[k].shape[1:])
        d['var_names'] = list(self.uploaded_vars)
        return d

    def get_refl(self, Da, Db, rcs, D, N, step_D, Dmin):
        """Compatibility shim with the data contract of the SWIG function get_refl
        (doppler_scatter.py:715-716, doppler_c.c:11-32): returns refl [len(Da)]."""
        f = lambda a: np.ascontiguousarray(a, dtype=np.float32)
        Da, Db, rcs, D, N, step_D, Dmin = f(Da), f(Db), f(rcs), f(D), f(N), f(step_D), f(Dmin)
        n, nh = Da.shape
        nd = rcs.shape[0]
        out = np.zeros(n, dtype=np.float32)
        self._check(self.lib.radop_get_refl(self.h, n, nh, nd, _ptr(Da), _ptr(Db), _ptr(rcs), _ptr(D), _ptr(N),
                                            _ptr(step_D), _ptr(Dmin), _ptr(out)))
        return out

    def get_refl_noalloc(self, Da, Db, rcs, D, N, step_D, Dmin, out=None):
        """get_refl through radop_get_refl_noalloc: the device side allocates nothing (operands and result are staged in
        a static device array of NOALLOC_SCRATCH_BYTES inside the library; larger problems raise RadopError).  Not in the
        reference.  Pass contiguous float32 arrays and `out` to keep the host side allocation-free as well."""
        f = lambda a: np.ascontiguousarray(a, dtype=np.float32)
        Da, Db, rcs, D, N, step_D, Dmin = f(Da), f(Db), f(rcs), f(D), f(N), f(step_D), f(Dmin)
        n, nh = Da.shape
        nd = rcs.shape[0]
        if out is None:
            out = np.zeros(n, dtype=np.float32)
        self._check(self.lib.radop_get_refl_noalloc(self.h, n, nh, nd, _ptr(Da), _ptr(Db), _ptr(rcs), _ptr(D), _ptr(N),
                                                    _ptr(step_D), _ptr(Dmin), _ptr(out)))
        return out

    def get_all_radar_pts(self, coords, heights_gates, model_data, model_heights, llc, res):
        """Compatibility shim with the data contract of the SWIG function
        (interpolation.py:584-595): returns the interpolated profile."""
        coords = np.ascontiguousarray(coords, dtype=np.float32)
        hg = np.ascontiguousarray(heights_gates, dtype=np.float32)
        data = np.ascontiguousarray(model_data, dtype=np.float32)
        mh = np.ascontiguousarray(model_heights, dtype=np.float32)
        llc = np.ascontiguousarray(llc, dtype=np.float32)
        res = np.ascontiguousarray(res, dtype=np.float32)
        n = coords.shape[0]
        out = np.zeros(n, dtype=np.float32)
        nz, ny, nx = data.shape
        self._check(self.lib.radop_get_all_radar_pts(self.h, n, _ptr(coords), _ptr(hg), _ptr(data), _ptr(mh),
                                                     nz, ny, nx, _ptr(llc), _ptr(res), _ptr(out)))
        return out
