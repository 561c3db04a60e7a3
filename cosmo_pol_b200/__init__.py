# -*- coding: utf-8 -*-
"""cosmo_pol_b200 -- B200-native drop-in for the per-gate hot path of wolfidan/cosmo_pol
(polarimetric radar forward operator for COSMO).  See DESIGN.md."""
from .radar_operator import RadarOperator

__all__ = ['RadarOperator']
