"""The stand-ins for the absent third-party geodesy (pyproj.Geod.fwd, pycosmo.WGS_to_COSMO)
checked against independent computations."""
import numpy as np
from scipy.integrate import solve_ivp

import radop_oracle as O


def _integrated_geodesic(lat0, lon0, az, dist):
    """Geodesic on the WGS84 ellipsoid integrated numerically (Clairaut form)."""
    a, f = O.WGS84_A, O.WGS84_F
    e2 = f * (2 - f)

    def rhs(s, y):
        phi, lam, alpha = y
        w = np.sqrt(1 - e2 * np.sin(phi) ** 2)
        M = a * (1 - e2) / w ** 3
        N = a / w
        return [np.cos(alpha) / M, np.sin(alpha) / (N * np.cos(phi)), np.sin(alpha) * np.tan(phi) / N]

    sol = solve_ivp(rhs, [0, dist], [np.deg2rad(lat0), np.deg2rad(lon0), np.deg2rad(az)], rtol=1e-13, atol=1e-15,
                    method='DOP853')
    return np.rad2deg(sol.y[1, -1]), np.rad2deg(sol.y[0, -1])


def test_vincenty_direct_against_integrated_geodesic():
    for lat0, lon0 in ((47.28436, 8.51201), (-12.0, 130.0), (70.0, -40.0)):
        for az in (0.0, 37.3, 90.0, 181.0, 269.5, 359.0):
            for d in (150.0, 75000.0, 150000.0, 400000.0):
                lon, lat = O.geod_fwd(lon0, lat0, az, np.array([d]))
                lon_i, lat_i = _integrated_geodesic(lat0, lon0, az, d)
                # 1e-9 deg ~ 0.1 mm: two orders below the float32 rounding of the coordinates
                assert abs(lat[0] - lat_i) < 2e-9 and abs(lon[0] - lon_i) < 3e-9, (lat0, az, d)


def test_karney_direct_against_integrated_geodesic_and_vincenty():
    """geod_fwd_karney restates the algorithm pyproj.Geod.fwd runs (Karney 2013, order 6): held to the numerical
    integration, including lines long enough (3000 km) for the third-order terms of the series to matter, and
    compared with the Vincenty stand-in the kernels restate: the two agree to ~1e-11 degrees (1 micrometre) over
    radar distances (scripts/geodesy_karney_vs_vincenty.py counts what that means for the float32 casts of cfg 3)."""
    worst = 0.0
    for lat0, lon0 in ((47.28436, 8.51201), (-12.0, 130.0), (70.0, -40.0)):
        for az in (12.0, 37.3, 90.0, 181.0, 269.5, 331.0):
            for d in (150.0, 75000.0, 150000.0, 400000.0, 3.0e6):
                lon, lat = O.geod_fwd_karney(lon0, lat0, az, np.array([d]))
                lon_i, lat_i = _integrated_geodesic(lat0, lon0, az, d)
                assert abs(lat[0] - lat_i) < 2e-9 and abs(lon[0] - lon_i) < 6e-9 / np.cos(np.deg2rad(lat_i)), (lat0, az, d)
                if d <= 400000.0:
                    lon_v, lat_v = O.geod_fwd(lon0, lat0, az, np.array([d]))
                    worst = max(worst, abs(lat[0] - lat_v[0]), abs(lon[0] - lon_v[0]))
    assert worst < 1e-10, worst
    # broadcasting over (azimuth, distance) gives the same numbers as one call per azimuth
    az = np.array([3.0, 200.0])[:, None]
    d = np.linspace(150., 150000., 7)[None, :]
    lon2, lat2 = O.geod_fwd_karney(8.5, 47.3, az, d)
    for i in range(2):
        lon1, lat1 = O.geod_fwd_karney(8.5, 47.3, float(az[i, 0]), d[0])
        assert np.array_equal(lon1, lon2[i]) and np.array_equal(lat1, lat2[i])


def test_rotated_pole_is_a_rotation():
    rng = np.random.default_rng(0)
    lat = rng.uniform(40, 52, 100)
    lon = rng.uniform(0, 18, 100)
    r = O.wgs_to_rotated(lat, lon, -43.0, 10.0)
    # a point on the pole meridian keeps rlon = 0 and is shifted by the pole colatitude
    p = O.wgs_to_rotated(np.array([47.0]), np.array([10.0]), -43.0, 10.0)
    assert abs(p[0, 0]) < 1e-12 and abs(p[0, 1]) < 1e-12
    # great-circle distances are preserved
    def gc(a1, o1, a2, o2):
        a1, o1, a2, o2 = map(np.deg2rad, (a1, o1, a2, o2))
        return np.arccos(np.clip(np.sin(a1) * np.sin(a2) + np.cos(a1) * np.cos(a2) * np.cos(o1 - o2), -1, 1))
    d0 = gc(lat[:-1], lon[:-1], lat[1:], lon[1:])
    d1 = gc(r[:-1, 0], r[:-1, 1], r[1:, 0], r[1:, 1])
    assert np.allclose(d0, d1, atol=1e-12)
