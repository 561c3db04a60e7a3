from .doppler_scatter import get_radar_observables, cut_at_sensitivity, get_pol_from_sz
