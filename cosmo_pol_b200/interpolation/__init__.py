from .radial import Radial
from .atm_refraction import compute_trajectory_radial
from .quadrature import subbeam_fan
from .interpolation import get_interpolated_radial, integrate_radials
