# -*- coding: utf-8 -*-
"""global_constants.py -- configuration independent and configuration derived
constants.  Same attribute names as cosmo_pol/constants/global_constants.py:111-198
(``EPS, C, RHO_W, KE, T0, NVEL, VRES, VARRAY, WAVELENGTH, RANGE_RADAR, ...``); the
object is refreshed by ``update()`` whenever ``RadarOperator.config`` is set."""
import numpy as np


class Constant_class(object):
    #: variables censored by cut_at_sensitivity (global_constants.py:117-118)
    SIMULATED_VARIABLES = ['ZH', 'DSPECTRUM', 'RVEL', 'ZV', 'PHIDP', 'ZDR', 'RHOHV', 'KDP']

    def __init__(self, config=None):
        self.EPS = float(np.finfo(np.float64).eps)
        self.C = 299792458.                 # speed of light [m/s]
        self.RHO_W = 1000. / (1000 ** 3)    # water density [kg/mm3]
        self.RHO_I = 916. / (1000 ** 3)     # ice density [kg/mm3]
        self.A = 1.6                        # turbulence constant (Doviak & Zrnic p.409)
        self.RHO_0 = 1.225
        self.KE = 4. / 3.
        self.MAX_MODEL_HEIGHT = 35000
        self.T0 = 273.15
        self.T_K_SQUARED = 283.15
        self.M_AIR = complex(1, 0)
        self._derive(config)

    def _derive(self, config):
        if config is None:
            return
        r = config['radar']
        self.NVEL = self.C / (4 * 1E-6 * r['PRI'] * r['frequency'] * 1E09)
        self.VRES = 2 * self.NVEL / r['FFT_length']
        self.VARRAY = np.arange(-self.NVEL, self.NVEL + self.VRES, self.VRES)
        self.WAVELENGTH = self.C / (r['frequency'] * 1E09) * 1000      # [mm]
        self.PULSE_WIDTH = 2 * r['radial_resolution']
        sens = r.get('sensitivity')
        if isinstance(sens, (list, tuple)) and len(sens) == 3:
            self.RADAR_CONSTANT_DB = (180 - 10 * np.log10(np.pi ** 3 * r['3dB_beamwidth'] ** 2
                                      * r['K_squared'] * self.PULSE_WIDTH / 1000)
                                      - 2 * sens[1]
                                      + 10 * np.log10(1024 * np.log(2) * (self.WAVELENGTH / 1000.) ** 2))
        if r.get('type', 'ground') == 'ground':
            self.RANGE_RADAR = np.arange(r['radial_resolution'] / 2., r['range'], r['radial_resolution'])

    def update(self):
        from ..config import cfg
        self.__init__(cfg.CONFIG)


global_constants = Constant_class()
