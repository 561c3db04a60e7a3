"""Physical / derived constants of the radar operator (host side).

Mirrors the *names* of cosmo_pol/constants/ so user code keeps working; the
values that the CUDA kernels need are marshalled into ``RadopParams`` /
``RadopSpecies`` (include/radop.h) by ``cosmo_pol_b200/_abi.py``."""
from .global_constants import global_constants
from . import constants_1mom, constants_2mom
