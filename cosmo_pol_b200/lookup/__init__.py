from .lut import Lookup_table, save_lut, load_all_lut, load_lut
