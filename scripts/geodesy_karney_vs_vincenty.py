#!/usr/bin/env python
"""How far is the Vincenty stand-in for pyproj.Geod.fwd (oracle geod_fwd, restated in CUDA by the kernels) from
Karney's algorithm (oracle geod_fwd_karney: what pyproj / PROJ actually runs, restated from the 2013 paper)?

For every sub-beam gate of configuration 3 (20 elevations x 360 azimuths x 9 sub-beams x 500 gates = 3.24e7):
the largest coordinate difference, how many float32 casts of the rotated coordinates differ (the casts are what the
reference hands to its C interpolation, interpolation_c.i:16) and how many grid-cell indices differ.
CPU only (numpy); prints one JSON line.  Output committed as profiles/r02_geodesy_karney_vs_vincenty.json."""
import json
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), '..'))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
import bench                      # noqa: E402
import radop_oracle as O          # noqa: E402


def main():
    cfg = bench.checked_config(bench.volume_config())
    from cosmo_pol_b200 import synthetic
    proj = synthetic.grid_projection() if hasattr(synthetic, 'grid_projection') else None
    model = synthetic.model_as_plain(synthetic.make_model(nz=4))          # only the grid geometry is needed
    proj, res = model['proj'], model['resolution']
    llc = np.asarray((float(proj['Lo1']), float(proj['La1']))).astype('float32')
    res32 = np.asarray(res, dtype=np.float32)
    rr = O.derived_constants(cfg)['RANGE_RADAR']
    pos = cfg['radar']['coords']
    pts_hor, pts_ver, weights, keep = O.gauss_hermite_subbeams(cfg)
    az = np.arange(360, dtype=np.float64)
    out = {'sub_beam_gates': 0, 'max_abs_diff_deg': 0.0, 'float32_casts_differ': 0, 'cells_differ': 0,
           'fractions_differ': 0}
    for el in bench.VOLUME_ELEVATIONS:
        refr = [O.ref_4_3(rr, pt + el, pos) for pt in pts_ver]
        for i in range(len(pts_hor)):
            for j in range(len(pts_ver)):
                if not keep[i, j]:
                    continue
                s = np.asarray(refr[j][0], np.float64)[None, :]
                c = []
                for fn in (O.geod_fwd, O.geod_fwd_karney):
                    lon, lat = fn(pos[1], pos[0], (pts_hor[i] + az)[:, None], s)
                    c.append(O.wgs_to_rotated(lat.ravel(), lon.ravel(), proj['Latitude_of_southern_pole'],
                                              proj['Longitude_of_southern_pole']))
                a, b = c
                out['sub_beam_gates'] += a.shape[0]
                out['max_abs_diff_deg'] = max(out['max_abs_diff_deg'], float(np.max(np.abs(a - b))))
                a32, b32 = a.astype(np.float32), b.astype(np.float32)
                out['float32_casts_differ'] += int(np.sum(a32 != b32))
                # grid positions exactly as interpolation_c.c:44-47 forms them (float32)
                pa = [(a32[:, 0] - llc[1]) / res32[1], (a32[:, 1] - llc[0]) / res32[0]]
                pb = [(b32[:, 0] - llc[1]) / res32[1], (b32[:, 1] - llc[0]) / res32[0]]
                out['cells_differ'] += int(np.sum((np.floor(pa[0]) != np.floor(pb[0])) | (np.floor(pa[1]) != np.floor(pb[1]))))
                out['fractions_differ'] += int(np.sum((pa[0] != pb[0]) | (pa[1] != pb[1])))
    out['float32_ulp_deg_at_2deg'] = float(np.spacing(np.float32(2.0)))
    print(json.dumps(out))


if __name__ == '__main__':
    main()
