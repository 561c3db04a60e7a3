from .utilities import (Range, TypeList, vlinspace, nan_cumprod, nan_cumsum, sum_arr, nansum_arr,
                        combine_subradials, get_earth_radius, aliasing)
