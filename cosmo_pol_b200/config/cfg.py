# -*- coding: utf-8 -*-
"""cfg.py -- user configuration: YAML file -> nested dict kept in the module global
``CONFIG`` (same contract as cosmo_pol/config/cfg.py:34-281: ``init``, ``sanity_check``,
``DEFAULTS``, ``VALID_VALUES``).  Invalid optional values fall back to the default with a
message, missing mandatory values raise ValueError."""
import builtins
import copy
import re
from textwrap import dedent

import numpy as np
import yaml

from ..utilities.utilities import Range, TypeList
from .nyquist_fit import nyquist_interpolator

CONFIG = None

DEFAULTS = {
    'radar': {'type': 'ground', 'range': 150000, 'radial_resolution': 500, 'PRI': 700, 'FFT_length': 256,
              'sensitivity': [-5, 10000], '3dB_beamwidth': 1., 'K_squared': 0.93, 'antenna_speed': 0.2,
              'nyquist_velocity': None, 'frequency': 9.41},
    'refraction': {'scheme': 1},
    'integration': {'scheme': 1, 'nv_GH': 9, 'nh_GH': 3, 'n_gaussians': 7, 'antenna_diagram': None,
                    'weight_threshold': 1., 'nr_GH': 7, 'na_GL': 7},
    'doppler': {'scheme': 1, 'turbulence_correction': 0, 'motion_correction': 0},
    'microphysics': {'scheme': '1mom', 'with_melting': 0, 'with_ice_crystals': 1, 'with_attenuation': 1,
                     'scattering': 'tmatrix_masc'},
}

VALID_VALUES = {
    'radar': {'coords': TypeList([float, int], [3]),
              'type': ['ground', 'GPM'],
              'frequency': [2.7, 5.6, 9.41, 9.8, 13.6, 35.6],
              'range': Range(5000, 500000),
              'radial_resolution': Range(25, 5000),
              'PRI': Range(10, 3000),
              'FFT_length': Range(16, 2048),
              'sensitivity': [TypeList([float, int], [3]), TypeList([float, int], [2]), float, int],
              '3dB_beamwidth': Range(0.1, 10.),
              'K_squared': [float, None],
              'nyquist_velocity': [None, str],
              'antenna_speed': Range(1E-6, 10.)},
    'refraction': {'scheme': [1, 2]},
    'integration': {'scheme': [1, 2, 3, 4, 5, 'ml'],     # cfg.py:101 lists [1, 2, 3, 4, 'ml']; 5 exists in interpolation.py:276
                    'nv_GH': range(1, 31, 2), 'nh_GH': range(1, 31, 2),
                    'n_gaussians': range(1, 13, 2),
                    'weight_threshold': Range(0.0001, 1.),
                    'nr_GH': range(1, 31, 2), 'na_GL': range(1, 31, 2)},
    'doppler': {'scheme': [1, 2, 3], 'turbulence_correction': [0, 1], 'motion_correction': [0, 1]},
    'microphysics': {'scheme': ['1mom', '2mom'], 'with_ice_crystals': [0, 1], 'with_melting': [0, 1],
                     'with_attenuation': [0, 1], 'scattering': ['tmatrix_masc', 'tmatrix', 'dda']},
}


def _check_validity(value, valid):
    if isinstance(valid, TypeList):
        return valid == value
    if isinstance(valid, list):
        return any(_check_validity(value, v) for v in valid)
    if isinstance(value, list):
        return all(_check_validity(v, valid) for v in value)
    if type(valid) == builtins.type:
        return type(value) == valid
    if isinstance(valid, (Range, range)):
        return value in valid
    if isinstance(valid, str) and valid.startswith('-reg-'):
        return bool(re.match(valid[5:] + r'\Z', value))
    return valid == value


def init(options_file):
    """Read a YAML options file into the global CONFIG (cfg.py:162-188).  Unreadable
    files give the defaults, with a message."""
    global CONFIG
    CONFIG = None
    if isinstance(options_file, dict):          # convenience: an already parsed configuration
        CONFIG = copy.deepcopy(options_file)
        return CONFIG
    try:
        with open(options_file, 'r') as f:
            CONFIG = yaml.safe_load(f)
        if not isinstance(CONFIG, dict):
            raise ValueError('options file does not contain a mapping')
    except Exception as e:
        CONFIG = copy.deepcopy(DEFAULTS)
        print('Could not find or read {}, using default options...\nThe error was:'.format(options_file))
        print(e)
        return
    return CONFIG


def sanity_check(config):
    """Fill in defaults / reject invalid values (cfg.py:190-281)."""
    for section in VALID_VALUES:
        if section not in config or config[section] is None:
            config[section] = {}
        for key, valid in VALID_VALUES[section].items():
            if key not in config[section]:
                if key in DEFAULTS[section]:
                    config[section][key] = copy.deepcopy(DEFAULTS[section][key])
                else:
                    raise ValueError(dedent('''
                        Mandatory key {:s}/{:s} is missing, please provide a valid value, aborting...
                        '''.format(section, key)))
            if not _check_validity(config[section][key], valid):
                print('Invalid value entered for key: {:s}/{:s}\nThe value must be: {:s}'.format(
                    section, key, str([str(v) for v in valid] if isinstance(valid, list) else valid)))
                if key in DEFAULTS[section]:
                    config[section][key] = copy.deepcopy(DEFAULTS[section][key])
                    print('The default value {:s} was assigned'.format(str(DEFAULTS[section][key])))
                else:
                    raise ValueError('This key is mandatory, please provide a valid value, aborting...')
        for key, val in DEFAULTS[section].items():
            config[section].setdefault(key, copy.deepcopy(val))

    if isinstance(config['radar']['nyquist_velocity'], str):
        config['radar']['nyquist_velocity'] = nyquist_interpolator(config['radar']['nyquist_velocity'])

    # cfg.py:241-262: an antenna diagram file is fitted with a sum of Gaussians, whose parameters integration scheme 2
    # samples (the reference tries this for every configuration, the key being always present; here only when a file is
    # named and no parameters were given)
    integ = config['integration']
    if isinstance(integ.get('antenna_diagram'), str) and integ.get('antenna_params') is None:
        from .antenna_fit import optimize_gaussians
        import numpy as np
        data = np.genfromtxt(integ['antenna_diagram'], delimiter=',')
        integ['antenna_params'] = optimize_gaussians(data[:, 0], 10 * np.log10((10 ** (0.1 * data[:, 1])) ** 2),
                                                     int(integ['n_gaussians']))

    if config['radar']['K_squared'] is None:
        from ..hydrometeors import dielectric
        config['radar']['K_squared'] = dielectric.K_squared(config['radar']['frequency'])
    return config
