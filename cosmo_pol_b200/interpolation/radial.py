# -*- coding: utf-8 -*-
"""Radial -- API-visible result record (cosmo_pol/interpolation/radial.py:16-54)."""


class Radial(object):
    def __init__(self, dic_values, mask, lats_profile, lons_profile, dist_ground_profile, heights_profile,
                 elev_profile=(), quad_pt=(), quad_weight=1):
        self.mask = mask
        self.quad_pt = quad_pt
        self.quad_weight = quad_weight
        self.lats_profile = lats_profile
        self.lons_profile = lons_profile
        self.dist_profile = dist_ground_profile
        self.heights_profile = heights_profile
        self.elev_profile = elev_profile
        self.values = dic_values
        self.has_melting = False
