# -*- coding: utf-8 -*-
"""pyart_wrapper.py -- the simulated scan in the field layout of the reference's PyartRadop
(cosmo_pol/radar/pyart_wrapper.py:186-342), assembled straight from the structure-of-arrays
result of the GPU path instead of a list (sweeps) of lists (rays) of ``Radial`` objects
(SURVEY.md section 8f, N2).

``RadopVolume`` holds exactly the dictionaries PyartRadop hands to ``pyart.core.Radar.__init__``
(time, range, fields, metadata, scan_type, latitude, longitude, altitude, sweep_number, sweep_mode,
fixed_angle, sweep_start_ray_index, sweep_stop_ray_index, azimuth, elevation,
instrument_parameters) and works without Py-ART; ``to_pyart()`` builds the ``pyart.core.Radar`` when
that library is installed.  As in the reference: ZH, ZV and ZDR are converted to dB (exact zeros
become NaN first), every field is a masked array with the NaNs masked, per-gate ``Latitude`` /
``Longitude`` fields are added, and ``rangearray`` repeats the range vector for every ray.  The
reference's list of variables to convert spells the spectrum 'DSpectrum' (:256), so ``DSPECTRUM``
stays linear there and here.  Two slips of the reference are not reproduced: ``fixed_angle`` keeps only
the last sweep's angle (:279, :283) and ``time`` has the length of one sweep (:312); both are per
sweep / per ray here, which is what ``pyart.core.Radar`` expects."""
import numpy as np

from ..constants import global_constants as constants

_DB_FIELDS = ('ZDR', 'ZV', 'ZH')
_UNITS = {'ZH': 'dBZ', 'ZV': 'dBZ', 'ZDR': 'dB', 'KDP': 'deg/km', 'PHIDP': 'deg', 'DELTA_HV': 'deg', 'RHOHV': '-',
          'RVEL': 'm/s', 'ATT_H': 'dB/km', 'ATT_V': 'dB/km', 'DSPECTRUM': 'mm6 m-3', 'U': 'm/s', 'V': 'm/s', 'W': 'm/s',
          'T': 'K', 'RHO': 'kg/m3', 'Latitude': 'degrees', 'Longitude': 'degrees', 'rangearray': 'm'}


class RadopVolume(object):
    def __init__(self, scan_type, scan):
        """scan_type: 'ppi' or 'rhi'; scan: the SimulatedScan returned by get_PPI / get_RHI."""
        scan_type = scan_type.lower()
        if scan_type not in ('ppi', 'rhi'):
            raise ValueError("scan_type must be 'ppi' or 'rhi'")
        n_sweeps, n_per, n_gates = scan.mask.shape
        n_rays = n_sweeps * n_per
        self.scan_type = scan_type
        self.nsweeps, self.nrays, self.ngates = n_sweeps, n_rays, n_gates
        fields = {}
        for name, arr in scan.fields.items():
            a = np.array(arr, dtype=np.float64 if arr.dtype == np.float64 else np.float32, copy=True)
            a = a.reshape((n_rays,) + a.shape[2:])
            if name in _DB_FIELDS:
                with np.errstate(divide='ignore', invalid='ignore'):
                    a[a == 0] = np.nan
                    a = 10 * np.log10(a)
            fields[name] = {'data': np.ma.array(a, mask=np.isnan(a))}
        for name, arr in (('Latitude', scan.lats), ('Longitude', scan.lons)):
            a = np.asarray(arr).reshape(n_rays, n_gates)
            fields[name] = {'data': np.ma.array(a, mask=np.isnan(a))}
        ranges = np.asarray(scan.ranges)
        fields['rangearray'] = {'data': np.tile(ranges, (n_rays, 1))}
        for name, f in fields.items():
            if name in _UNITS:
                f['units'] = _UNITS[name]
        self.fields = fields
        start = np.arange(n_sweeps) * n_per
        self.sweep_start_ray_index = {'data': start.astype(float)}
        self.sweep_end_ray_index = {'data': (start + n_per - 1).astype(float)}
        el, az = np.asarray(scan.elevations, dtype=float), np.asarray(scan.azimuths, dtype=float)
        if scan_type == 'ppi':       # one elevation per sweep, all azimuths in every sweep
            self.fixed_angle = {'data': el.copy()}
            self.elevation = {'data': np.repeat(el, n_per)}
            self.azimuth = {'data': np.tile(az, n_sweeps)}
        else:                        # one azimuth per sweep, all elevations in every sweep
            self.fixed_angle = {'data': az.copy()}
            self.elevation = {'data': np.tile(el, n_sweeps)}
            self.azimuth = {'data': np.repeat(az, n_per)}
        self.range = {'data': ranges}
        pt = scan.pos_time
        self.latitude = {'data': np.array(pt['latitude'], dtype=float)}
        self.longitude = {'data': np.array(pt['longitude'], dtype=float)}
        self.altitude = {'data': np.array(pt['altitude'], dtype=float)}
        self.time = {'data': np.zeros((n_rays,)), 'units': 'seconds since ' + str(pt.get('time'))}
        self.sweep_number = {'data': np.arange(0, n_sweeps, dtype=float)}
        self.sweep_mode = {'data': [scan_type] * n_sweeps}
        self.metadata = {}
        varray = getattr(constants, 'VARRAY', None)
        self.instrument_parameters = {'varray': {'data': varray}} if varray is not None else {}

    def get_field(self, sweep_idx, variable):
        """Rays of one sweep, squeezed (pyart_wrapper.py:344-348)."""
        s0 = int(self.sweep_start_ray_index['data'][sweep_idx])
        s1 = int(self.sweep_end_ray_index['data'][sweep_idx]) + 1
        return np.squeeze(self.fields[variable]['data'][s0:s1])

    def to_pyart(self):
        try:
            from pyart import core
        except ImportError:
            raise ImportError('Py-ART is not installed; RadopVolume itself carries the same fields')
        return core.Radar(self.time, self.range, self.fields, self.metadata, self.scan_type, self.latitude,
                          self.longitude, self.altitude, self.sweep_number, self.sweep_mode, self.fixed_angle,
                          {'data': self.sweep_start_ray_index['data'].astype(int)},
                          {'data': self.sweep_end_ray_index['data'].astype(int)}, self.azimuth, self.elevation,
                          instrument_parameters=self.instrument_parameters)


PyartRadop = RadopVolume      # the reference's name (radar_operator.py:447-451)
