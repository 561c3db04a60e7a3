// radop_kernels.cu -- hand-written sm_100a kernels of the radar forward operator hot path.
//
//  K0  k_pack_*            one-time: transpose model fields into a column-major,
//                          variable-interleaved layout; combine + premultiply LUT columns
//  K1  k_interp_psd        per sub-beam gate: geolocation, trilinear interpolation of all
//                          variables (bit-exact restatement of interpolation_c.c), PSD
//                          parameters, LUT bin keys, analytic Doppler moments, histogram
//  K1b k_scan_bins / k_scatter_pairs    counting sort of the active pairs by LUT bin
//  K2t k_build_lamtab / k_integrate_tab   (default) the PSD x scattering-table contraction
//                          through lambda-transformed tables: the sum over the diameters of
//                          one bin is a function of the PSD slope alone, tabulated piecewise
//                          in float64 at table upload and evaluated per pair in ray order
//  K2  k_integrate         (radop_set_integration(1)) the same contraction summed over the
//                          1024 diameters: LUT slice staged in shared memory by TMA bulk
//                          copy, register-tiled blocked-Horner evaluation (FFMA2)
//  K3  k_ice_totals / k_reduce_subbeams / k_ray_epilogue
//                          quadrature + species sums, polarimetric observables,
//                          PHIDP / attenuation prefix scans with warp shuffles
#include "radop_kernels.cuh"
#include <math.h>

// ------------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float ex2f_approx(float x)
{
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok)
                     : "r"(smem_u32(bar)), "r"(parity)
                     : "memory");
    } while (!ok);
}
// 1-D TMA bulk copy global -> shared, completion signalled on an mbarrier
__device__ __forceinline__ void tma_bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ------------------------------------------------------------------------------------------
// K0: one-time layout transforms
// ------------------------------------------------------------------------------------------
// z-levels [nz][ny][nx] -> [ny][nx][nz]
__global__ void k_pack_zlevels(const float *__restrict__ src, float *__restrict__ dst, int nz, int ny, int nx)
{
    size_t n = (size_t)nz * ny * nx;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        size_t col = i % ((size_t)ny * nx);
        size_t l = i / ((size_t)ny * nx);
        dst[col * nz + l] = src[i];
    }
}
// variable [nz][ny][nx] -> fields[ny][nx][nz][NVP] at slot var
__global__ void k_pack_field(const float *__restrict__ src, float *__restrict__ dst, int nz, int ny, int nx, int var)
{
    size_t n = (size_t)nz * ny * nx;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        size_t col = i % ((size_t)ny * nx);
        size_t l = i / ((size_t)ny * nx);
        dst[(col * nz + l) * RADOP_NVP + var] = src[i];
    }
}

// Bounds of the vertical differences of one array [nz][ny][nx] (float32 a[l+1] - a[l], the numerators / divisors of the
// reference's vertical interpolation, interpolation_c.c:84-91, which always pairs vertically adjacent levels):
// out[0] = largest |difference| as a bit pattern (NaN / inf compare above every finite value), out[1] = smallest
// non-zero one, out[2] = 1 if some difference is <= 0 (only meaningful for the z-levels, where the kernel
// subtracts the other way round).  From these the host decides, once per upload, whether EVERY quotient the
// interpolation can ever form takes the shared-reciprocal path exactly (div_shared_rcp), so that the per-gate
// kernels need no per-quotient range checks.
__global__ void k_level_diff_bounds(const float *__restrict__ src, int nz, size_t plane, int descending, unsigned *out)
{
    unsigned hi = 0u, lo = 0xffffffffu, bad = 0u;
    const size_t n = (size_t)(nz - 1) * plane;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float a = src[i], b = src[i + plane];
        const float d = descending ? __fsub_rn(a, b) : __fsub_rn(b, a);
        const unsigned u = __float_as_uint(d) & 0x7fffffffu;
        hi = max(hi, u);
        if (u) lo = min(lo, u);
        if (descending && !(d > 0.f)) bad = 1u;
    }
#pragma unroll
    for (int k = 16; k > 0; k >>= 1) {
        hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, k));
        lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, k));
        bad |= __shfl_xor_sync(0xffffffffu, bad, k);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMax(out, hi);
        atomicMin(out + 1, lo);
        if (bad) atomicOr(out + 2, 1u);
    }
}

// LUT value_table [ne][nt][nd][12] float64 -> combined float32 [ne][nt][nd][ncol]
// columns: 0: b = s0-s1-s2+s3   1: c = s0+s1+s2+s3   2: p = s4+s7   3: q = s6-s5
//          4: kd = s10-s8       5: s9                6: s11         7: spare
// (all the reference derives from the 12 columns is linear in them and enters
//  get_pol_from_sz only through these seven combinations, doppler_scatter.py:516-542;
//  forming s10-s8 in float64 BEFORE the integral removes the KDP cancellation)
// premultiplied by  D^mu * dD  (GEO/GEN modes) or dD (ICE1, whose shape function is
// evaluated in the kernel); column 7 of the 8-column layout is b v(D) (rcs-weighted fall speed,
// Doppler scheme 2); 12-column modes hold 7: v(D) (scheme 1) or b v(D) (scheme 2), 8: 1,
// 9: m(D) (ICE1 only).
__global__ void k_pack_lut(const double *__restrict__ src, float *__restrict__ dst, const float *__restrict__ d_axis,
                           size_t n_rows, int nd, int ncol, int mode, double mu, double dD, double alpha,
                           double beta, double a_m, double b_m, int doppler_scheme)
{
    for (size_t r = blockIdx.x * (size_t)blockDim.x + threadIdx.x; r < n_rows;
         r += (size_t)gridDim.x * blockDim.x) {
        int k = (int)(r % nd);
        const double *s = src + r * 12;
        double D = (double)d_axis[k];
        double pre = (mode == MODE_ICE1) ? dD : (mode == MODE_MELT ? 1.0 : pow(D, mu) * dD);
        float *o = dst + r * ncol;
        o[0] = (float)((s[0] - s[1] - s[2] + s[3]) * pre);
        o[1] = (float)((s[0] + s[1] + s[2] + s[3]) * pre);
        o[2] = (float)((s[4] + s[7]) * pre);
        o[3] = (float)((s[6] - s[5]) * pre);
        o[4] = (float)((s[10] - s[8]) * pre);
        o[5] = (float)(s[9] * pre);
        o[6] = (float)(s[11] * pre);
        // rcs-weighted fall speed (Doppler scheme 2, doppler_scatter.py:286-296): b(D) v(D)
        const double bv = (s[0] - s[1] - s[2] + s[3]) * alpha * pow(D, beta) * pre;
        if (ncol == 8) {
            o[7] = (float)bv;
        } else {
            o[7] = (doppler_scheme == 2) ? (float)bv : (float)(alpha * pow(D, beta) * pre);
            o[8] = (float)pre;
            o[9] = (mode == MODE_ICE1) ? (float)(a_m * pow(D, b_m) * pre) : 0.f;
            o[10] = 0.f;
            o[11] = 0.f;
        }
    }
}
// aux[0][k] = D_k, aux[1][k] = D_k^nu, aux[2][k] = D_k^0.6357
__global__ void k_pack_aux(const float *__restrict__ d_axis, float *__restrict__ aux, int nd, double nu)
{
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < nd) {
        double D = (double)d_axis[k];
        aux[k] = d_axis[k];
        aux[nd + k] = (float)pow(D, nu);
        aux[2 * nd + k] = (float)pow(D, 0.6357);
    }
}

// ------------------------------------------------------------------------------------------
// geometry: WGS84 direct geodesic (Vincenty) + rotated-pole rotation, all float64
// (stand-ins for pyproj.Geod.fwd and pycosmo.WGS_to_COSMO, interpolation.py:526-568)
// ------------------------------------------------------------------------------------------
#define WGS84_A 6378137.0
#define WGS84_F (1.0 / 298.257223563)
#define DEG2RAD 0.017453292519943295
#define RAD2DEG 57.29577951308232

// everything of the direct problem that depends on the azimuth only: one record per
// (ray, sub-beam), filled by k_sub_geometry, read (warp-broadcast) by every gate
struct __align__(16) SubGeo {
    double sinA1, cosA1;     // also the wind projection's sin/cos(phi), proj_vel doppler_scatter.py:30-47
    double sinU1, cosU1;
    double s2s1, c2s1;       // sin / cos of 2 sigma1
    double sinAlpha, B, C;
    double inv_bA;           // 1 / (b A): sigma0 = s * inv_bA
};

__global__ void k_sub_geometry(int n_rays, int n_sub, double lat0_deg, const double *__restrict__ ray_az,
                               const double *__restrict__ sub_daz, SubGeo *__restrict__ geo)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rays * n_sub) return;
    const double f = WGS84_F, a = WGS84_A, b = (1.0 - f) * a;
    double az = ray_az[i / n_sub] + sub_daz[i % n_sub];
    double phi1 = lat0_deg * DEG2RAD, alpha1 = az * DEG2RAD;
    SubGeo k;
    sincos(alpha1, &k.sinA1, &k.cosA1);
    double tanU1 = (1.0 - f) * tan(phi1);
    k.cosU1 = 1.0 / sqrt(1.0 + tanU1 * tanU1);
    k.sinU1 = tanU1 * k.cosU1;
    double sigma1 = atan2(tanU1, k.cosA1);
    sincos(2.0 * sigma1, &k.s2s1, &k.c2s1);
    k.sinAlpha = k.cosU1 * k.sinA1;
    double cosSqAlpha = 1.0 - k.sinAlpha * k.sinAlpha;
    double uSq = cosSqAlpha * (a * a - b * b) / (b * b);
    double A = 1.0 + uSq / 16384.0 * (4096.0 + uSq * (-768.0 + uSq * (320.0 - 175.0 * uSq)));
    k.B = uSq / 1024.0 * (256.0 + uSq * (-128.0 + uSq * (74.0 - 47.0 * uSq)));
    k.C = f / 16.0 * cosSqAlpha * (4.0 + f * (4.0 - 3.0 * cosSqAlpha));
    k.inv_bA = 1.0 / (b * A);
    geo[i] = k;
}

// sin / cos for |x| <= SMALL_ANGLE by Taylor polynomials (truncation < 1e-19 relative): the
// arc length of a radar ray is a few hundredths of a radian, so the argument reduction and
// the slow paths of the library sincos are never needed
#define SMALL_ANGLE 0.06
__device__ __forceinline__ void sincos_small(double x, double &s, double &c)
{
    const double x2 = x * x;
    double ps = -1.0 / 39916800.0;
    ps = fma(ps, x2, 1.0 / 362880.0);
    ps = fma(ps, x2, -1.0 / 5040.0);
    ps = fma(ps, x2, 1.0 / 120.0);
    ps = fma(ps, x2, -1.0 / 6.0);
    s = fma(x * x2, ps, x);
    double pc = 1.0 / 479001600.0;
    pc = fma(pc, x2, -1.0 / 3628800.0);
    pc = fma(pc, x2, 1.0 / 40320.0);
    pc = fma(pc, x2, -1.0 / 720.0);
    pc = fma(pc, x2, 1.0 / 24.0);
    pc = fma(pc, x2, -0.5);
    c = fma(x2, pc, 1.0);
}
// asin / atan for |x| <= SMALL_RATIO by Taylor polynomials (first neglected term < 3e-18 relative)
#define SMALL_RATIO 0.125
__device__ __forceinline__ double asin_small(double z)
{
    const double z2 = z * z;
    double p = 12155.0 / 1245184.0;
    p = fma(p, z2, 6435.0 / 557056.0);
    p = fma(p, z2, 143.0 / 10240.0);
    p = fma(p, z2, 231.0 / 13312.0);
    p = fma(p, z2, 63.0 / 2816.0);
    p = fma(p, z2, 35.0 / 1152.0);
    p = fma(p, z2, 5.0 / 112.0);
    p = fma(p, z2, 3.0 / 40.0);
    p = fma(p, z2, 1.0 / 6.0);
    return fma(z * z2, p, z);
}
__device__ __forceinline__ double atan_small(double t)
{
    const double t2 = t * t;
    double p = -1.0 / 19.0;
    p = fma(p, t2, 1.0 / 17.0);
    p = fma(p, t2, -1.0 / 15.0);
    p = fma(p, t2, 1.0 / 13.0);
    p = fma(p, t2, -1.0 / 11.0);
    p = fma(p, t2, 1.0 / 9.0);
    p = fma(p, t2, -1.0 / 7.0);
    p = fma(p, t2, 1.0 / 5.0);
    p = fma(p, t2, -1.0 / 3.0);
    return fma(t * t2, p, t);
}

// Vincenty direct solution for one gate of a sub-beam: unit vector (cos lat cos lon, ...)
// of the gate is all the rotated-pole transform needs, so the geographic angles themselves
// are only formed for the central sub-beam (want_angles), whose lat / lon are outputs.
// cl0 / sl0: cos / sin of the radar longitude.
__device__ __forceinline__ void geod_forward(const SubGeo &k, double dist, double cl0, double sl0, double lon0_deg,
                                             bool want_angles, double &x, double &y, double &z, double &lon,
                                             double &lat)
{
    const double f = WGS84_F;
    const double sig0 = dist * k.inv_bA;
    double sigma = sig0, sinSigma, cosSigma, c2m;
    const bool small = sig0 < SMALL_ANGLE - 1e-3;
#pragma unroll 1
    for (int it = 0; it < 6; ++it) {      // 5 fixed-point updates, then the evaluation at the result
        if (small) sincos_small(sigma, sinSigma, cosSigma);
        else sincos(sigma, &sinSigma, &cosSigma);
        c2m = k.c2s1 * cosSigma - k.s2s1 * sinSigma;      // cos(2 sigma1 + sigma)
        if (it == 5) break;
        double dS = k.B * sinSigma *
                    (c2m + k.B / 4.0 *
                               (cosSigma * (-1.0 + 2.0 * c2m * c2m) -
                                k.B / 6.0 * c2m * (-3.0 + 4.0 * sinSigma * sinSigma) * (-3.0 + 4.0 * c2m * c2m)));
        sigma = sig0 + dS;
    }
    const double tmp = k.sinU1 * sinSigma - k.cosU1 * cosSigma * k.cosA1;
    const double num = k.sinU1 * cosSigma + k.cosU1 * sinSigma * k.cosA1;
    const double den = (1.0 - f) * sqrt(k.sinAlpha * k.sinAlpha + tmp * tmp);
    const double la = sinSigma * k.sinA1, lb = k.cosU1 * cosSigma - k.sinU1 * sinSigma * k.cosA1;
    const double corr = (1.0 - k.C) * f * k.sinAlpha *
                        (sigma + k.C * sinSigma * (c2m + k.C * cosSigma * (-1.0 + 2.0 * c2m * c2m)));
    // sin / cos of latitude and of (lon0 + lam - corr) without forming the angles
    const double ih1 = rsqrt(num * num + den * den), ih2 = rsqrt(la * la + lb * lb);
    const double sphi = num * ih1, cphi = den * ih1;
    const double slam = la * ih2, clam = lb * ih2;
    double sc, cc;
    sincos_small(corr, sc, cc);            // |corr| < f * sigma
    const double sL = slam * cc - clam * sc, cL = clam * cc + slam * sc;
    const double slon = sl0 * cL + cl0 * sL, clon = cl0 * cL - sl0 * sL;
    x = clon * cphi;
    y = slon * cphi;
    z = sphi;
    if (want_angles) {
        lat = atan2(num, den) * RAD2DEG;
        lon = lon0_deg + (atan2(la, lb) - corr) * RAD2DEG;
    }
}

// geographic unit vector -> rotated-pole (rlat, rlon) in degrees
__device__ __forceinline__ void rotate_pole(const double *__restrict__ R, double x, double y, double z,
                                            double &rlat, double &rlon)
{
    const double xn = R[0] * x + R[1] * y + R[2] * z;
    const double yn = R[3] * x + R[4] * y + R[5] * z;
    const double zn = R[6] * x + R[7] * y + R[8] * z;
    if (xn > 0.0 && fabs(yn) <= SMALL_RATIO * xn && fabs(zn) <= SMALL_RATIO) {
        rlon = atan_small(yn / xn) * RAD2DEG;
        rlat = asin_small(zn) * RAD2DEG;
    } else {
        rlon = atan2(yn, xn) * RAD2DEG;
        rlat = asin(zn) * RAD2DEG;
    }
}

// ------------------------------------------------------------------------------------------
// Geodesy by polynomial: along one sub-beam the azimuth is fixed and the rotated-pole coordinates
// are analytic functions of the ground distance s alone.  The arc a radar ray covers is a few
// hundredths of a radian, so their Chebyshev series on [s_min, s_max] of the scan converges
// geometrically (each coefficient ~ 0.01 x the previous one): the Vincenty solution + rotation is
// evaluated at GEO_NP Chebyshev nodes per (ray, sub-beam) instead of at every gate, the series is
// truncated after GEO_NC terms (the dropped ones are checked to be below 1e-13 degrees, else flag
// bit 2 and the scan is refused) and converted to the power basis, and a gate costs one Horner
// evaluation per coordinate.  The difference to the per-gate solution is ~1e-14 degrees, the
// rounding level of the per-gate solution itself (tests/test_gpu_parity.py::test_geodesy_polynomial).
// ------------------------------------------------------------------------------------------
struct ScanDev {
    const double *ray_az;
    const int *ray_traj;
    const double *sub_daz;
    const int *sub_jv;
    const double *sub_weight;
    const float *traj_s, *traj_h, *traj_e;
    const float *traj_sin, *traj_tan;   // optional (Doppler scheme 3)
    const int *sub_ml;                  // optional (integration scheme 'ml'): 1 = weight only near the melting-layer edges
};

#define GEO_NP 16
#define GEO_NC 12
struct __align__(16) SubPoly {
    double sinA1, cosA1;
    double mid, rhalf;       // x = (s - mid) * rhalf in [-1, 1] over the ground distances of THIS sub-beam's trajectory
    double c[4][GEO_NC];     // rlat, rlon, lat, lon (degrees) as polynomials in x
};

__device__ __forceinline__ double poly_eval(const double2 *__restrict__ c, double x)
{   // coefficients in ascending order, GEO_NC / 2 double2
    double2 t = __ldg(c + GEO_NC / 2 - 1);
    double r = fma(t.y, x, t.x);
#pragma unroll
    for (int k = GEO_NC / 2 - 2; k >= 0; --k) {
        t = __ldg(c + k);
        r = fma(r, x, t.y);
        r = fma(r, x, t.x);
    }
    return r;
}

// block = 128 threads = 8 sub-beams x 16 nodes
__global__ void __launch_bounds__(128)
k_sub_geopoly(const __grid_constant__ DevConfig cfg, int n_rays, ScanDev scan,
              const double *__restrict__ ray_az, const double *__restrict__ sub_daz, SubPoly *__restrict__ out, int *flags)
{
    const int n_sub = cfg.n_sub;
    const double lat0_deg = cfg.radar_lat, lon0_deg = cfg.radar_lon, cl0 = cfg.cos_lon0, sl0 = cfg.sin_lon0;
    const double *rot = cfg.rot;
    __shared__ double s_f[8][4][GEO_NP];
    __shared__ double s_a[8][4][GEO_NP];
    const int sub_l = threadIdx.x >> 4, j = threadIdx.x & 15;
    const int i = blockIdx.x * 8 + sub_l;
    const bool ok = i < n_rays * n_sub;
    const double f = WGS84_F, a = WGS84_A, b = (1.0 - f) * a;
    SubGeo k;
    // The interval is the ground-distance range of the sub-beam's own trajectory row, so that the
    // polynomial (and with it lat / lon to the last bit) does not depend on which other rays share
    // the scan, the shard or the chunk.
    double s_mid = 0.0, s_half = 1.0;
    {
        float lo = 3.0e38f, hi = -3.0e38f;
        if (ok) {
            const float *row = scan.traj_s + ((size_t)scan.ray_traj[i / n_sub] * cfg.n_vert + scan.sub_jv[i % n_sub]) * cfg.n_gates;
            for (int g = j; g < cfg.n_gates; g += GEO_NP) { const float v = __ldg(row + g); lo = fminf(lo, v); hi = fmaxf(hi, v); }
        }
#pragma unroll
        for (int d = 8; d >= 1; d >>= 1) {
            lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, d));
            hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, d));
        }
        if (ok) {
            s_mid = 0.5 * ((double)lo + (double)hi);
            s_half = fmax(0.5 * ((double)hi - (double)lo), 1.0);
        }
    }
    if (ok) {
        double az = ray_az[i / n_sub] + sub_daz[i % n_sub];
        double phi1 = lat0_deg * DEG2RAD, alpha1 = az * DEG2RAD;
        sincos(alpha1, &k.sinA1, &k.cosA1);
        double tanU1 = (1.0 - f) * tan(phi1);
        k.cosU1 = 1.0 / sqrt(1.0 + tanU1 * tanU1);
        k.sinU1 = tanU1 * k.cosU1;
        double sigma1 = atan2(tanU1, k.cosA1);
        sincos(2.0 * sigma1, &k.s2s1, &k.c2s1);
        k.sinAlpha = k.cosU1 * k.sinA1;
        double cosSqAlpha = 1.0 - k.sinAlpha * k.sinAlpha;
        double uSq = cosSqAlpha * (a * a - b * b) / (b * b);
        double A = 1.0 + uSq / 16384.0 * (4096.0 + uSq * (-768.0 + uSq * (320.0 - 175.0 * uSq)));
        k.B = uSq / 1024.0 * (256.0 + uSq * (-128.0 + uSq * (74.0 - 47.0 * uSq)));
        k.C = f / 16.0 * cosSqAlpha * (4.0 + f * (4.0 - 3.0 * cosSqAlpha));
        k.inv_bA = 1.0 / (b * A);
        const double xj = cospi(((double)j + 0.5) / GEO_NP);
        double gx, gy, gz, lon, lat, rlat, rlon;
        geod_forward(k, s_mid + s_half * xj, cl0, sl0, lon0_deg, true, gx, gy, gz, lon, lat);
        rotate_pole(rot, gx, gy, gz, rlat, rlon);
        s_f[sub_l][0][j] = rlat; s_f[sub_l][1][j] = rlon; s_f[sub_l][2][j] = lat; s_f[sub_l][3][j] = lon;
    }
    __syncthreads();
    if (ok) {
        bool bad = false;
#pragma unroll
        for (int v = 0; v < 4; ++v) {
            double acc = 0.0;
            for (int n = 0; n < GEO_NP; ++n) acc += s_f[sub_l][v][n] * cospi((double)j * ((double)n + 0.5) / GEO_NP);
            acc *= (j == 0 ? 1.0 : 2.0) / GEO_NP;
            s_a[sub_l][v][j] = acc;
            if (j >= GEO_NC && fabs(acc) > 1e-13) bad = true;
        }
        if (bad) atomicOr(flags, 4);
    }
    __syncthreads();
    if (ok && j < 4) {
        // Chebyshev -> power basis: T_{n+1} = 2 x T_n - T_{n-1}
        double m[GEO_NC], t0[GEO_NC], t1[GEO_NC];
#pragma unroll
        for (int n = 0; n < GEO_NC; ++n) { m[n] = 0.0; t0[n] = 0.0; t1[n] = 0.0; }
        t0[0] = 1.0; t1[1] = 1.0;
        m[0] = s_a[sub_l][j][0];
        m[1] = s_a[sub_l][j][1];
#pragma unroll
        for (int n = 2; n < GEO_NC; ++n) {
            double t2[GEO_NC];
#pragma unroll
            for (int d = 0; d < GEO_NC; ++d) t2[d] = (d ? 2.0 * t1[d - 1] : 0.0) - t0[d];
            const double an = s_a[sub_l][j][n];
#pragma unroll
            for (int d = 0; d < GEO_NC; ++d) { m[d] = fma(an, t2[d], m[d]); t0[d] = t1[d]; t1[d] = t2[d]; }
        }
#pragma unroll
        for (int d = 0; d < GEO_NC; ++d) out[i].c[j][d] = m[d];
        if (j == 0) { out[i].sinA1 = k.sinA1; out[i].cosA1 = k.cosA1; out[i].mid = s_mid; out[i].rhalf = 1.0 / s_half; }
    }
}

// ------------------------------------------------------------------------------------------
// trilinear interpolation: bit-exact restatement of interpolation_c.c:42-103, 108-164
// on the transposed layout z_t[col][level]
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int column_search(const float *__restrict__ col, int dim, float key)
{
    if (key > __ldg(col)) return -1;
    if (key < __ldg(col + dim - 1)) return -2;
    int hi = 0, lo = dim - 1;
    while (hi < lo) {
        int mid = (hi + lo) / 2;
        float c = __ldg(col + mid);
        if (c == key) return mid;
        if (lo == mid || hi == mid) return hi;
        if (c < key) lo = mid;
        else if (c > key) hi = mid;
    }
    return hi;
}

// Correctly rounded float32 quotient n / den from y = RN(1 / den) (Markstein's theorem:
// q0 = RN(n y) is within one ulp, r = n - den q0 is exact in an FMA, RN(q0 + r y) is the
// correctly rounded quotient).  One reciprocal serves all the variables interpolated in a
// column.  Zero, infinite, NaN and tiny quotients (where the residual could underflow) take
// the IEEE division instead; radop_selftest_division checks the equivalence on the GPU.
__device__ __forceinline__ bool div_den_safe(float den)
{   // the residual n - den q0 is exact when |den| >= 2^-3 and |q| >= 1e-30 (no bits below 2^-149)
    const float ad = fabsf(den);
    return ad >= 0.125f && ad <= 1e15f;
}
__device__ __forceinline__ float div_shared_rcp(float n, float den, float y, bool den_safe)
{
    const float q0 = __fmul_rn(n, y);
    const float r = __fmaf_rn(-den, q0, n);
    float q = __fmaf_rn(r, y, q0);
    const float aq = fabsf(q);
    if (!den_safe || !(aq <= 1e30f) || (aq < 1e-30f && n != 0.f)) q = __fdiv_rn(n, den);
    return q;
}

__global__ void k_selftest_division(unsigned long long n, unsigned long long seed, int mode,
                                    unsigned long long *mismatch)
{
    unsigned long long bad = 0;
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < n;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        // splitmix64
        unsigned long long z = seed + (i + 1) * 0x9E3779B97F4A7C15ull;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z ^= z >> 31;
        unsigned a = (unsigned)z, b = (unsigned)(z >> 32);
        float num, den;
        if (mode == 0) {            // arbitrary bit patterns
            num = __uint_as_float(a);
            den = __uint_as_float(b);
        } else if (mode == 1) {     // level spacings 1 .. 4096 m, field differences of any magnitude
            den = __uint_as_float(0x3f800000u + (b % (12u << 23)));
            num = __uint_as_float(a);
        } else {                    // numerators with few significant bits, zeros, exact multiples
            den = __uint_as_float(0x3f800000u + (b % (12u << 23)));
            unsigned k = a & 0xffu;
            num = (k == 0) ? 0.f : den * (float)((int)(a >> 8 & 0xfff) - 2048) * exp2f((float)((int)(k % 80) - 40));
        }
        float y = __frcp_rn(den);
        float q1 = div_shared_rcp(num, den, y, div_den_safe(den)), q2 = __fdiv_rn(num, den);
        bool same = (__float_as_uint(q1) == __float_as_uint(q2)) || (isnan(q1) && isnan(q2));
        if (!same) ++bad;
    }
    if (bad) atomicAdd(mismatch, bad);
}

// flag[0] |= 1 when some column of z_t is not strictly descending (the fast search below
// is then replaced by the reference's probe sequence)
__global__ void k_check_columns(const float *__restrict__ z_t, int nz, size_t ncol, int *flag)
{
    for (size_t c = blockIdx.x * (size_t)blockDim.x + threadIdx.x; c < ncol; c += (size_t)gridDim.x * blockDim.x) {
        const float *col = z_t + c * nz;
        bool ok = true;
        for (int l = 0; l + 1 < nz; ++l) ok = ok && (col[l] > col[l + 1]);
        if (!ok) atomicOr(flag, 1);
    }
}

struct CellGeom {
    int iy, ix;         // lat index, lon index of the lower-left neighbour
    int y0, y1, x0, x1; // memory-safe (clamped) neighbour indices
    float x, y, dx, dy; // fractional position and complements
    int status;         // 0 ok, 1 above model top, -1 below topography
    int level[4];       // upper level used in each neighbour column
};

// pos0 / pos1 are the float32 grid positions ((coord - llc)/res evaluated in float32).
// strict: every column of z_t is strictly descending (checked at upload).  The reference's
// binary search (interpolation_c.c:108-135) then returns the largest index i <= nz-2 with
// col[i] >= h (an exact match is always probed, the last level never is), which the four
// neighbour columns find together in ceil(log2(nz-1)) rounds of independent loads.
// (top, bottom) of every column of the transposed level array
__global__ void k_pack_topbot(const float *__restrict__ z_t, int nz, size_t n_cols, float2 *__restrict__ out)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_cols; i += (size_t)gridDim.x * blockDim.x)
        out[i] = make_float2(z_t[i * nz], z_t[i * nz + nz - 1]);
}

#define RADOP_NGUESS 1024
__device__ __forceinline__ void locate_cell(const float *__restrict__ z_t, int nz, int ny, int nx, float pos0,
                                            float pos1, float h, bool strict, CellGeom &c,
                                            const uint8_t *__restrict__ guess = nullptr, const float2 *__restrict__ tb = nullptr)
{
    c.iy = (int)floorf(pos0);
    c.ix = (int)floorf(pos1);
    c.x = __fsub_rn(pos0, truncf(pos0));      // fmod(pos, 1.0): exact, keeps the sign of pos
    c.y = __fsub_rn(pos1, truncf(pos1));
    c.dx = __fsub_rn(1.0f, c.x);
    c.dy = __fsub_rn(1.0f, c.y);
    // memory-safe neighbour indices (the reference reads out of bounds on the last row/col)
    const int y0 = min(max(c.iy, 0), ny - 1), y1 = min(max(c.iy + 1, 0), ny - 1);
    const int x0 = min(max(c.ix, 0), nx - 1), x1 = min(max(c.ix + 1, 0), nx - 1);
    c.y0 = y0; c.y1 = y1; c.x0 = x0; c.x1 = x1;
    const float *col[4] = {z_t + ((size_t)y0 * nx + x0) * nz, z_t + ((size_t)y0 * nx + x1) * nz,
                           z_t + ((size_t)y1 * nx + x0) * nz, z_t + ((size_t)y1 * nx + x1) * nz};
    float t0, t1, t2, t3, top0 = 0.f, top1 = 0.f, top2 = 0.f, top3 = 0.f;
    if (tb) {
        const float2 e0 = __ldg(tb + (size_t)y0 * nx + x0), e1 = __ldg(tb + (size_t)y0 * nx + x1);
        const float2 e2 = __ldg(tb + (size_t)y1 * nx + x0), e3 = __ldg(tb + (size_t)y1 * nx + x1);
        top0 = e0.x; t0 = e0.y; top1 = e1.x; t1 = e1.y; top2 = e2.x; t2 = e2.y; top3 = e3.x; t3 = e3.y;
    } else {
        t0 = __ldg(col[0] + nz - 1); t1 = __ldg(col[1] + nz - 1);
        t2 = __ldg(col[2] + nz - 1); t3 = __ldg(col[3] + nz - 1);
    }
    // interp_topo = dx*dy*v0 + x*v2*dy + dx*v1*y + x*y*v3   (left to right, no contraction)
    float topo = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(__fmul_rn(c.dx, c.dy), t0),
                                               __fmul_rn(__fmul_rn(c.x, t2), c.dy)),
                                     __fmul_rn(__fmul_rn(c.dx, t1), c.y)),
                           __fmul_rn(__fmul_rn(c.x, c.y), t3));
    c.level[0] = c.level[1] = c.level[2] = c.level[3] = -1;
    if (!(topo < h)) { c.status = -1; return; }
    c.status = 0;
    if (strict) {
        if (!tb) { top0 = __ldg(col[0]); top1 = __ldg(col[1]); top2 = __ldg(col[2]); top3 = __ldg(col[3]); }
        if (h > top0 || h > top1 || h > top2 || h > top3) { c.status = 1; return; }
        const float bot[4] = {t0, t1, t2, t3};
        int idx[4] = {0, 0, 0, 0};
        unsigned todo = 0xfu;          // columns that still need the bisection
        if (guess) {
            // One round trip instead of six: the level is guessed from the height normalised by the column's
            // own top and bottom (table of a reference column, radop_upload_grid: exact for sigma-like levels,
            // close for hybrid ones), four consecutive levels around the guess are read, and the answer --
            // the largest i <= nz - 2 with col[i] >= h, unique in a strictly descending column -- is taken
            // from them when they bracket h.  A column whose window misses takes the bisection below, so the
            // result is the reference's in every case.
            const float top[4] = {top0, top1, top2, top3};
            int a[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float eta = __fdividef(h - bot[k], top[k] - bot[k]);
                const int t = min(max((int)(eta * (float)RADOP_NGUESS), 0), RADOP_NGUESS - 1);
                a[k] = min(max((int)__ldg(guess + t) - 1, 0), nz - 4);
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float *w = col[k] + a[k];
                // (keeping these four heights for the interpolation instead of loading its two again was measured: the eight
                //  extra live registers spill and the kernel loses 8 %)
                const int cnt = (__ldg(w) >= h) + (__ldg(w + 1) >= h) + (__ldg(w + 2) >= h) + (__ldg(w + 3) >= h);
                const bool ok = (cnt >= 1 || a[k] == 0) && (cnt <= 3 || a[k] + 3 >= nz - 2);
                if (ok) {
                    idx[k] = min(a[k] + max(cnt, 1) - 1, nz - 2);
                    todo &= ~(1u << k);
                }
            }
        }
        if (todo) {
            int step = 1;
            while (2 * step <= nz - 2) step *= 2;
            for (; step > 0; step >>= 1) {
                float v[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) v[k] = (todo & (1u << k)) ? __ldg(col[k] + min(idx[k] + step, nz - 2)) : 0.f;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if ((todo & (1u << k)) && idx[k] + step <= nz - 2 && v[k] >= h) idx[k] += step;
            }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            int i = idx[k];
            if (h < bot[k]) i = nz - 3;
            else if (i == nz - 2) i--;
            c.level[k] = i;
        }
        return;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        int idx = column_search(col[k], nz, h);
        if (idx == -1) { c.status = 1; break; }
        if (idx == -2) idx = nz - 3;
        else if (idx == nz - 2) idx--;
        c.level[k] = idx;
    }
    if (c.status != 0) c.level[0] = c.level[1] = c.level[2] = c.level[3] = -1;
}

// vertical interpolation in one column: v2 - (v2 - v1)/(c1 - c2) * (h - c2)
__device__ __forceinline__ float vlerp(float v1, float v2, float c1, float c2, float h)
{
    return __fsub_rn(v2, __fmul_rn(__fdiv_rn(__fsub_rn(v2, v1), __fsub_rn(c1, c2)), __fsub_rn(h, c2)));
}

// compatibility kernel: same data contract as get_all_radar_pts (reference layout [nz][ny][nx])
__global__ void k_interp_compat(int n, const float *__restrict__ coords, const float *__restrict__ hg,
                                const float *__restrict__ data, const float *__restrict__ heights, int nz, int ny,
                                int nx, float llc0, float llc1, float res0, float res1, float *__restrict__ out)
{
    int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    float pos0 = __fdiv_rn(__fsub_rn(coords[2 * g], llc1), res1);
    float pos1 = __fdiv_rn(__fsub_rn(coords[2 * g + 1], llc0), res0);
    float h = hg[g];
    int iy = (int)floorf(pos0), ix = (int)floorf(pos1);
    float x = fmodf(pos0, 1.0f), y = fmodf(pos1, 1.0f);
    float dx = __fsub_rn(1.0f, x), dy = __fsub_rn(1.0f, y);
    size_t plane = (size_t)ny * nx;
    int y0 = min(max(iy, 0), ny - 1), y1 = min(max(iy + 1, 0), ny - 1);
    int x0 = min(max(ix, 0), nx - 1), x1 = min(max(ix + 1, 0), nx - 1);
    size_t off[4] = {(size_t)y0 * nx + x0, (size_t)y0 * nx + x1, (size_t)y1 * nx + x0, (size_t)y1 * nx + x1};
    float t[4];
    for (int k = 0; k < 4; ++k) t[k] = heights[(size_t)(nz - 1) * plane + off[k]];
    float topo = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(__fmul_rn(dx, dy), t[0]), __fmul_rn(__fmul_rn(x, t[2]), dy)),
                                     __fmul_rn(__fmul_rn(dx, t[1]), y)),
                           __fmul_rn(__fmul_rn(x, y), t[3]));
    if (!(topo < h)) { out[g] = __int_as_float(0x7fc00000); return; }
    float vv[4];
    for (int k = 0; k < 4; ++k) {
        const float *hc = heights + off[k];
        int hi = 0, lo = nz - 1, idx;
        if (h > hc[0]) { out[g] = -9999.0f; return; }
        if (h < hc[(size_t)(nz - 1) * plane]) idx = -2;
        else {
            idx = 0;
            while (hi < lo) {
                int mid = (hi + lo) / 2;
                float c = hc[(size_t)mid * plane];
                if (c == h) { idx = mid; break; }
                if (lo == mid || hi == mid) { idx = hi; break; }
                if (c < h) lo = mid;
                else if (c > h) hi = mid;
                idx = hi;
            }
        }
        if (idx == -2) idx = nz - 3;
        else if (idx == nz - 2) idx--;
        float c1 = hc[(size_t)idx * plane], c2 = hc[(size_t)(idx + 1) * plane];
        float v1 = data[(size_t)idx * plane + off[k]], v2 = data[(size_t)(idx + 1) * plane + off[k]];
        vv[k] = vlerp(v1, v2, c1, c2, h);
    }
    out[g] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(__fmul_rn(dx, dy), vv[0]), __fmul_rn(__fmul_rn(x, vv[2]), dy)),
                                 __fmul_rn(__fmul_rn(dx, vv[1]), y)),
                       __fmul_rn(__fmul_rn(x, y), vv[3]));
}

// compatibility kernel: same data contract and float32 order of operations as get_refl (doppler_c.c:11-32);
// thread = velocity bin, species and diameters in sequence
__global__ void k_refl_compat(int len, int nhydro, int nd, const float *__restrict__ Da, const float *__restrict__ Db,
                              const float *__restrict__ rcs, const float *__restrict__ N, const float *__restrict__ step_D,
                              const float *__restrict__ Dmin, float *__restrict__ refl)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= len) return;
    float r = 0.f;
    for (int j = 0; j < nhydro; ++j) {
        const float st = step_D[j], d0 = Dmin[j];
        // (int) of a float quotient, truncation toward zero as in C
        int ka = (int)__fdiv_rn(__fsub_rn(Da[i * nhydro + j], d0), st);
        int kb = (int)__fdiv_rn(__fsub_rn(Db[i * nhydro + j], d0), st);
        ka = max(ka, 0); kb = min(kb, nd);           // memory safety (undefined behaviour in the reference)
        float sum = 0.f;
        for (int k = ka; k < kb; ++k) sum = __fadd_rn(sum, __fmul_rn(N[k * nhydro + j], rcs[k * nhydro + j]));
        r = __fadd_rn(r, __fmul_rn(sum, st));
    }
    refl[i] = r;
}

// All variables of one gate from the 8 surrounding grid points, in the reference's order of
// operations (interpolation_c.c:84-103, 138-164).  EXACT = false: quotients from the reciprocal
// shared by the variables of a column, no branch; returns true when some quotient needs the
// IEEE division (see div_shared_rcp), in which case the caller repeats with EXACT = true.
template <bool EXACT>
__device__ __forceinline__ bool interp_columns(const DevConfig &cfg, const CellGeom &c, float h_g,
                                               const float *__restrict__ z_t, const float *__restrict__ fields,
                                               float (&val)[RADOP_NVP])
{
    // slots 0..8 are the 1-moment variables: two float4 + one scalar per level;
    // the 2-moment scheme reads all four float4
    // The four neighbour columns run through ONE copy of the body (a rolled loop: the kernel is limited by
    // instruction fetch, and four specialised copies of ~140 instructions each do not fit beside the rest of the
    // gate's code path), in the accumulation order of the reference: columns 0, 2, 1, 3.  -0 is the additive
    // identity of IEEE addition (x + -0 = x for every x, zeros included), so the running sums start from it
    // and every column is added the same way.
    bool rare = false;
#pragma unroll
    for (int j = 0; j < RADOP_NVP; ++j) val[j] = -0.f;
#pragma unroll 1
    for (int oi = 0; oi < 4; ++oi) {
        const int k = (oi == 1) ? 2 : (oi == 2 ? 1 : oi);
        const int yk = (k & 2) ? c.y1 : c.y0, xk = (k & 1) ? c.x1 : c.x0;
        const int lvl = (k == 0) ? c.level[0] : (k == 1 ? c.level[1] : (k == 2 ? c.level[2] : c.level[3]));
        const size_t colidx = (size_t)yk * cfg.nx + xk;
        const float *zc = z_t + colidx * cfg.nz + lvl;
        const float c1 = __ldg(zc), c2 = __ldg(zc + 1);
        const float den = __fsub_rn(c1, c2), dh = __fsub_rn(h_g, c2);
        const float rden = EXACT ? 0.f : __frcp_rn(den);
        if (!EXACT) rare = rare || !div_den_safe(den);
        const float4 *f1 = reinterpret_cast<const float4 *>(fields + (colidx * cfg.nz + lvl) * RADOP_NVP);
        const float4 *f2 = f1 + RADOP_NVP / 4;
        // term = (w0 * v) * w1 in the reference's association: col 1, 2: (w0 * v) * w1 ; col 0, 3: (w0 * w1) * v
        const bool v_mid = (k == 1) || (k == 2);
        const float w0 = (k == 0) ? __fmul_rn(c.dx, c.dy) : (k == 2 ? c.x : (k == 1 ? c.dx : __fmul_rn(c.x, c.y)));
        const float w1 = (k == 2) ? c.dy : c.y;
        auto accum = [&](int j, float a, float b) {
            const float n = __fsub_rn(b, a);
            float q;
            if (EXACT) q = __fdiv_rn(n, den);
            else {
                const float q0 = __fmul_rn(n, rden);
                const float r = __fmaf_rn(-den, q0, n);
                q = __fmaf_rn(r, rden, q0);
                const float aq = fabsf(q);
                rare = rare || !(aq <= 1e30f) || (aq < 1e-30f && n != 0.f);
            }
            float vv = __fsub_rn(b, __fmul_rn(q, dh));
            const float t0 = __fmul_rn(w0, vv);
            const float term = v_mid ? __fmul_rn(t0, w1) : t0;
            val[j] = __fadd_rn(val[j], term);
        };
#pragma unroll
        for (int v4 = 0; v4 < 2; ++v4) {
            float4 a = __ldg(f1 + v4), b = __ldg(f2 + v4);
            accum(4 * v4, a.x, b.x); accum(4 * v4 + 1, a.y, b.y);
            accum(4 * v4 + 2, a.z, b.z); accum(4 * v4 + 3, a.w, b.w);
        }
        if (cfg.two_moment) {
#pragma unroll
            for (int v4 = 2; v4 < 4; ++v4) {
                float4 a = __ldg(f1 + v4), b = __ldg(f2 + v4);
                accum(4 * v4, a.x, b.x); accum(4 * v4 + 1, a.y, b.y);
                accum(4 * v4 + 2, a.z, b.z); accum(4 * v4 + 3, a.w, b.w);
            }
        } else {
            float a = __ldg(reinterpret_cast<const float *>(f1) + 8), b = __ldg(reinterpret_cast<const float *>(f2) + 8);
            accum(8, a, b);
            if (cfg.var_mask & (1u << RADOP_VAR_EDR)) {
                a = __ldg(reinterpret_cast<const float *>(f1) + RADOP_VAR_EDR);
                b = __ldg(reinterpret_cast<const float *>(f2) + RADOP_VAR_EDR);
                accum(RADOP_VAR_EDR, a, b);
            }
        }
    }
    return rare;
}
// ------------------------------------------------------------------------------------------
// K1: per sub-beam gate
// ------------------------------------------------------------------------------------------
struct K1Out {
    int8_t *mask;          // [sbg]
    uint8_t *active;       // [sbg] bit per species
    float *wind;           // [sbg] (U sin phi + V cos phi) cos theta + W sin theta
    float *sinth;          // [sbg]
    float *vsum, *nsum;    // [sbg] analytic fall-speed moments (non-numeric species)
    // Doppler scheme 3 (spectrum): horizontal wind along the sub-beam azimuth, W, RHO, folded elevation
    double *d3_uh;
    float *d3_w, *d3_rho;
    float2 *d3_st;         // float32 (sin, tan) of the folded elevation
    double *d3_std;        // spectral width added by turbulence / antenna motion (or null)
    uint16_t *key;         // [species][sbg]
    float *p0, *p1;        // [species][sbg]
    float *vals;           // [sbg][NVP] interpolated variables (model output / debug) or null
    double *lat, *lon;     // [ray][gate] central sub-beam
    int *hist;             // [total_bins]
    int *flags;            // [0]: out-of-domain
    // debug
    int *dbg_cell, *dbg_level, *dbg_status, *dbg_e, *dbg_t;
    double *dbg_rot;
};

// Logarithm and exponential of the one-moment PSD parameter algebra in float32 accuracy: the parameters end
// up as float32 (lambda, N0) and only have to be good to ~1e-6 (0.01 dB = 2.3e-3), so the double-precision
// library calls (~60 instructions each) are replaced by the single-precision ones.  The two-moment scheme
// keeps the double calls: its KDP has gates where species cancel to 1e-3 of their size.  exp splits its
// argument into a float32 part and the tiny remainder, so the result carries expf's 1 ulp, not the
// rounding of a large argument.
__device__ __forceinline__ double log_sp(float x) { return (double)logf(x); }
__device__ __forceinline__ double exp_sp(double x)
{
    const float hi = (float)x;
    return (double)expf(hi) * (1.0 + (x - (double)hi));
}

// log of the second moment of Field et al. (2005), hydrometeors.py:1277-1299 with n = 3:
// log Q2 = (log QM' - ln10 * A(T)) / B(T)
__device__ __forceinline__ double ice_log_mom2(double T, double logQM)
{
    const double n = 3.0;
    T = T - 273.15;
    double a = 5.065339 - 0.062659 * T - 3.032362 * n + 0.029469 * T * n - 0.000285 * T * T + 0.312550 * n * n +
               0.000204 * T * T * n + 0.003199 * T * n * n - 0.015952 * n * n * n;
    double b = 0.476221 - 0.015896 * T + 0.165977 * n + 0.007468 * T * n - 0.000141 * T * T + 0.060366 * n * n +
               0.000079 * T * T * n + 0.000594 * T * n * n - 0.003577 * n * n * n;
    return (logQM - 2.302585092994046 * a) / b;
}

// quotient n / den from rden = RN(1 / den) (Markstein, see div_shared_rcp) for divisors that are
// scan constants (grid resolution, table axis steps): no branch unless the quotient is odd
__device__ __forceinline__ float div_const(float n, float den, float rden)
{
    const float q0 = __fmul_rn(n, rden);
    const float r = __fmaf_rn(-den, q0, n);
    float q = __fmaf_rn(r, rden, q0);
    const float aq = fabsf(q);
    // the residual is exact when n is not tiny (den is between 1e-6 and 1e6, checked on the host)
    if (!(aq <= 1e30f) || (fabsf(n) < 1e-25f && n != 0.f)) q = __fdiv_rn(n, den);
    return q;
}

__device__ __forceinline__ int lut_bin(float x, float lo, float step, float rstep, int dim)
{   // lut.py:336-341: floor((x - lo)/step) in float32, clamped
    float f = floorf(div_const(__fsub_rn(x, lo), step, rstep));
    int i = (int)f;
    if (f < 0.f) i = 0;
    if (f >= (float)dim) i = dim - 1;
    return i;
}

// one-moment slots that carry the melting re-partition (melting.py:19-90)
#define MELT_SLOT_QMS 9
#define MELT_SLOT_QMG 10
#define MELT_SLOT_FWS 11
#define MELT_SLOT_FWG 12

extern __shared__ int s_hist[];

// register-resident select (a dynamic index would push the value array to local memory)
__device__ __forceinline__ float pick_mass(const float (&v)[RADOP_NVP], int var)
{
    float r = v[RADOP_VAR_QR];
    if (var == RADOP_VAR_QS) r = v[RADOP_VAR_QS];
    if (var == RADOP_VAR_QG) r = v[RADOP_VAR_QG];
    if (var == RADOP_VAR_QI) r = v[RADOP_VAR_QI];
    if (var == RADOP_VAR_QH) r = v[RADOP_VAR_QH];
    return r;
}
__device__ __forceinline__ float pick_number(const float (&v)[RADOP_NVP], int var)
{
    float r = v[RADOP_VAR_QNR];
    if (var == RADOP_VAR_QNS) r = v[RADOP_VAR_QNS];
    if (var == RADOP_VAR_QNG) r = v[RADOP_VAR_QNG];
    if (var == RADOP_VAR_QNI) r = v[RADOP_VAR_QNI];
    if (var == RADOP_VAR_QNH) r = v[RADOP_VAR_QNH];
    return r;
}

// Persistent blocks: each block walks chunks of 128 consecutive sub-beam gates and keeps
// one shared-memory histogram of the LUT bins for its whole life (flushed once).
// GEO_POLY: rotated-pole (and, for the central sub-beam, geographic) coordinates from the per-sub-beam
// polynomials of k_sub_geopoly instead of a Vincenty solution per gate.
// (A phase-synchronous variant -- large blocks whose warps meet at a barrier after geodesy / cell search /
// interpolation / species so that they share the instruction cache -- was measured and is not kept:
// 3.97 ms at 256 / 512 threads and 4.24 ms at 1024 against 3.94 ms, profiles/r01f_k1_experiments.md.)
template <bool GEO_POLY>
__global__ void __launch_bounds__(128, 8)
k_interp_psd(const __grid_constant__ DevConfig cfg, ScanDev scan, const void *__restrict__ geo_v,
             const float *__restrict__ z_t, const float *__restrict__ fields, K1Out o, long long n_sbg, int by_flag)
{
    const int tid = threadIdx.x;
    for (int b = tid; b < cfg.total_bins; b += blockDim.x) s_hist[b] = 0;
    __syncthreads();
    const unsigned lane = tid & 31;
    const int ng = cfg.n_gates;
    const float qnan = __int_as_float(0x7fc00000);
    // both instances are launched; flag bit 2 (a polynomial of k_sub_geopoly did not converge) selects
    // the one that does the work
    if (by_flag && GEO_POLY != !(*(volatile int *)o.flags & 4)) return;

    // Work is handed out per warp in pieces of 128 consecutive sub-beam gates from a global counter
    // (o.flags[3]): warps that start late (an NCCL kernel holding the SM, a slow neighbour) simply
    // take fewer pieces, and the kernel ends when the last piece does.
    const long long n_pieces = (n_sbg + 127) / 128;
    for (;;) {
      long long piece = 0;
      if (lane == 0) piece = (long long)atomicAdd(o.flags + 3, 1);
      piece = __shfl_sync(0xffffffffu, piece, 0);
      if (piece >= n_pieces) break;
#pragma unroll 1
      for (int part = 0; part < 4; ++part) {
        const long long sbg = piece * 128 + part * 32 + (long long)lane;
        uint32_t keys[RADOP_NSPECIES_MAX];
#pragma unroll
        for (int s = 0; s < RADOP_NSPECIES_MAX; ++s) keys[s] = RADOP_KEY_NONE;
        const bool in_range = sbg < n_sbg;
        int g = 0;
        size_t trow = 0;
        float h_g = 0.f, e_g = 0.f, pos0 = 0.f, pos1 = 0.f;
        double sinA1 = 0.0, cosA1 = 0.0, rlat = 0.0, rlon = 0.0;
        CellGeom c;
        c.status = 1;

        if (in_range) {
            // n_sbg < 2^32 (checked by the host): 32-bit index arithmetic
            const unsigned rq = (unsigned)sbg / (unsigned)ng;
            g = (int)((unsigned)sbg - rq * (unsigned)ng);
            const int ray = (int)(rq / (unsigned)cfg.n_sub);
            const int q = (int)(rq - (unsigned)ray * (unsigned)cfg.n_sub);
            trow = ((size_t)scan.ray_traj[ray] * cfg.n_vert + scan.sub_jv[q]) * ng + g;
            const float s_g = __ldg(scan.traj_s + trow);
            h_g = __ldg(scan.traj_h + trow); e_g = __ldg(scan.traj_e + trow);

            // --- geolocation ------------------------------------------------------------------
            const bool center = (q == cfg.center_sub) && o.lat;
            double lon = 0.0, lat = 0.0;
            if (GEO_POLY) {
                const SubPoly *sp = reinterpret_cast<const SubPoly *>(geo_v) + rq;
                const double2 *cp = reinterpret_cast<const double2 *>(sp);
                const double2 t0 = __ldg(cp);
                sinA1 = t0.x; cosA1 = t0.y;
                const double2 iv = __ldg(cp + 1);
                const double x = ((double)s_g - iv.x) * iv.y;
                rlat = poly_eval(cp + 2, x);
                rlon = poly_eval(cp + 2 + GEO_NC / 2, x);
                if (center) {
                    lat = poly_eval(cp + 2 + GEO_NC, x);
                    lon = poly_eval(cp + 2 + 3 * GEO_NC / 2, x);
                }
            } else {
                SubGeo kc;
                const double2 *gp = reinterpret_cast<const double2 *>(reinterpret_cast<const SubGeo *>(geo_v) + rq);
                double2 t0 = __ldg(gp), t1 = __ldg(gp + 1), t2 = __ldg(gp + 2), t3 = __ldg(gp + 3), t4 = __ldg(gp + 4);
                kc.sinA1 = t0.x; kc.cosA1 = t0.y; kc.sinU1 = t1.x; kc.cosU1 = t1.y; kc.s2s1 = t2.x; kc.c2s1 = t2.y;
                kc.sinAlpha = t3.x; kc.B = t3.y; kc.C = t4.x; kc.inv_bA = t4.y;
                sinA1 = kc.sinA1; cosA1 = kc.cosA1;
                double gx, gy, gz;
                geod_forward(kc, (double)s_g, cfg.cos_lon0, cfg.sin_lon0, cfg.radar_lon, center, gx, gy, gz, lon, lat);
                rotate_pole(cfg.rot, gx, gy, gz, rlat, rlon);
            }
            if (center) {
                o.lat[(size_t)ray * ng + g] = lat;
                o.lon[(size_t)ray * ng + g] = lon;
            }
            // domain check on float64 coordinates against float32 corners (interpolation.py:572-575)
            if (rlon < (double)cfg.llc[0] || rlat < (double)cfg.llc[1] || rlon > (double)cfg.urc[0] ||
                rlat > (double)cfg.urc[1])
                atomicOr(o.flags, 1);
            // SWIG typemap casts the coordinates to float32 (interpolation_c.i:16)
            const float crlat = (float)rlat, crlon = (float)rlon;
            pos0 = div_const(__fsub_rn(crlat, cfg.llc[1]), cfg.res[1], cfg.rres[1]);
            pos1 = div_const(__fsub_rn(crlon, cfg.llc[0]), cfg.res[0], cfg.rres[0]);
        }
        if (in_range) {
            locate_cell(z_t, cfg.nz, cfg.ny, cfg.nx, pos0, pos1, h_g, cfg.z_strict != 0, c, cfg.lev_guess);
            o.mask[sbg] = (int8_t)c.status;
            if (o.dbg_cell) {
                o.dbg_cell[2 * sbg] = c.iy;
                o.dbg_cell[2 * sbg + 1] = c.ix;
                for (int k = 0; k < 4; ++k) o.dbg_level[4 * sbg + k] = c.level[k];
                o.dbg_status[sbg] = c.status;
                o.dbg_rot[2 * sbg] = rlat;
                o.dbg_rot[2 * sbg + 1] = rlon;
                for (int s = 0; s < cfg.n_species; ++s) { o.dbg_e[sbg * cfg.n_species + s] = -1; o.dbg_t[sbg * cfg.n_species + s] = -1; }
            }
        }

        float val[RADOP_NVP];
        const bool live = in_range && c.status == 0;
        if (live) {
            // --- interpolate every variable with the same 8 corner weights ----------------
            // fast pass: shared-reciprocal quotients without a branch; a rare flag (zero /
            // tiny / non-finite quotient, odd level spacing) sends the gate through the
            // same code with IEEE divisions
            if (interp_columns<false>(cfg, c, h_g, z_t, fields, val)) interp_columns<true>(cfg, c, h_g, z_t, fields, val);
        }

        if (in_range) {
            uint8_t active = 0;
            float wind = qnan, sinth = 0.f, vsum = 0.f, nsum = 0.f;
            if (live) {
                if (!cfg.two_moment) {
#pragma unroll
                    for (int j = 9; j < RADOP_VAR_EDR; ++j) val[j] = 0.f;
                    if (!(cfg.var_mask & (1u << RADOP_VAR_EDR))) val[RADOP_VAR_EDR] = 0.f;
                }
                if (cfg.with_melting) {
                    // melting.py:19-90 (float32 like the interpolated profiles): where rain coexists
                    // with snow or graupel the rain mass is shared out to wet snow / wet graupel in
                    // proportion to their dry masses, and the three original species are emptied.
                    // Slots 9..12 (unused by the one-moment scheme) carry QmS, QmG, fwet_mS, fwet_mG.
                    const float QR = val[RADOP_VAR_QR], QS = val[RADOP_VAR_QS], QG = val[RADOP_VAR_QG];
                    const float QSG = __fadd_rn(QS, QG);
                    if (QR > 0.f && QSG > 0.f) {
                        if (scan.sub_ml) active |= 0x80;      // mask_ml of melting.py:45 (read and cleared by k_ml_weights)
                        const float qms = __fadd_rn(QS, __fmul_rn(QR, __fdiv_rn(QS, QSG)));
                        const float qmg = __fadd_rn(QG, __fmul_rn(QR, __fdiv_rn(QG, QSG)));
                        val[MELT_SLOT_QMS] = qms;
                        val[MELT_SLOT_QMG] = qmg;
                        val[MELT_SLOT_FWS] = __fdiv_rn(__fdiv_rn(__fmul_rn(QR, QS), QSG), qms);
                        val[MELT_SLOT_FWG] = __fdiv_rn(__fdiv_rn(__fmul_rn(QR, QG), QSG), qmg);
                        if (qms > 0.f || qmg > 0.f) { val[RADOP_VAR_QR] = 0.f; val[RADOP_VAR_QS] = 0.f; val[RADOP_VAR_QG] = 0.f; }
                    }
                }

                // --- incidence elevation folded to [0, 90] (doppler_scatter.py:173-176) ---------
                float e = e_g;
                if (e > 90.f) e = __fsub_rn(180.f, e);
                if (e < 0.f) e = -e;
                const float T = val[RADOP_VAR_T];
                const double Td = (double)T;

                // --- species: PSD parameters (log domain), LUT bin, analytic Doppler moments ----
#pragma unroll 1
                for (int s = 0; s < cfg.n_species; ++s) {
                    const SpeciesDev &S = cfg.sp[s];
                    const bool wet = S.psd == RADOP_PSD_MELTING;
                    const float QMf = wet ? (S.species == RADOP_SP_MS ? val[MELT_SLOT_QMS] : val[MELT_SLOT_QMG])
                                          : pick_mass(val, S.qm_var);
                    if (!(QMf > 0.f)) continue;
                    active |= (uint8_t)(1u << s);
                    if (wet) {
                        // the reference cannot form the Doppler spectrum of a gate with a melting species
                        // (ValueError in get_doppler_spectrum, doppler_scatter.py:686-692): report it
                        if (cfg.doppler_scheme == 3) atomicOr(o.flags, 8);
                        // PSD and integrals per gate in k_integrate_melting; here only the table bin:
                        // lookup_line(e=..., wc=fwet), float64 arithmetic for the float64 fwet profile
                        const float fw = S.species == RADOP_SP_MS ? val[MELT_SLOT_FWS] : val[MELT_SLOT_FWG];
                        const int ie = lut_bin(e, S.e_min, S.e_step, S.e_rstep, S.ne);
                        const double fb = floor(((double)fw - (double)S.t_min) / (double)S.t_step);
                        const int it = fb < 0.0 ? 0 : (fb >= (double)S.nt ? S.nt - 1 : (int)fb);
                        keys[s] = (uint32_t)(S.bin_base + ie * S.nt + it);
                        o.p0[(size_t)s * n_sbg + sbg] = fw;
                        o.p1[(size_t)s * n_sbg + sbg] = QMf;
                        if (o.dbg_e) { o.dbg_e[sbg * cfg.n_species + s] = ie; o.dbg_t[sbg * cfg.n_species + s] = it; }
                        continue;
                    }
                    const double lQM = (S.psd == RADOP_PSD_2MOM) ? log((double)QMf) : log_sp(QMf);
                    double loglam, logn0 = 0.0, n0 = S.N0, ntot_2m = 0.0;
                    bool n0_from_log = false, n0_negative = false;
                    float p0, p1;
                    if (S.psd == RADOP_PSD_1MOM_FIXED_N0) {
                        // lambda = (lambda_factor / QM)^(1/(4+mu)), hydrometeors.py:761-772, 1039-1049
                        loglam = (S.c_l0 - lQM) * S.c_inv;
                    } else if (S.psd == RADOP_PSD_1MOM_SNOW) {
                        // N0 = 13.5 * 5.65e5 * exp(-0.107 (T - 273.15)) / 1000, hydrometeors.py:894-907
                        logn0 = 8.939515416590824 - 0.107 * (Td - 273.15);      // log(13.5 * 5.65e5 / 1000)
                        n0_from_log = true;
                        loglam = (S.c_l0 + logn0 - lQM) * S.c_inv;
                    } else if (S.psd == RADOP_PSD_1MOM_ICE) {
                        // lambda = (Q2 / QM)^(1/(b-2)), hydrometeors.py:1320-1339
                        loglam = (ice_log_mom2(Td, lQM - S.c_l0) - lQM) * S.c_inv - 6.907755278982137;
                    } else {
                        // _Hydrometeor.set_psd, hydrometeors.py:224-255
                        const double QN = (double)pick_number(val, S.qn_var);
                        ntot_2m = QN;
                        const double xm = fmin(fmax((double)QMf / (QN + 2.220446049250313e-16), S.x_min), S.x_max);
                        const double ll0 = S.c_inv * log(S.lambda_factor * xm);
                        // QN can be negative (extrapolation below the lowest usable level): the
                        // reference then carries a negative N0 through; keep the sign outside the log
                        logn0 = log(fabs(S.c_q * QN)) + S.e_q * ll0 - S.c_ka;
                        n0_from_log = true;
                        n0_negative = (S.c_q * QN) < 0.0;
                        loglam = ll0 - S.c_kb;
                    }
                    const bool two = S.psd == RADOP_PSD_2MOM;     // N0 spans 20 orders of magnitude there: full double
                    const double lam = two ? exp(loglam) : exp_sp(loglam);
                    if (n0_from_log) n0 = two ? (n0_negative ? -exp(logn0) : exp(logn0)) : exp_sp(logn0);
                    if (S.psd == RADOP_PSD_1MOM_ICE) { p0 = (float)lam; p1 = QMf; }     // lambda / 1000, QM
                    else { p0 = (float)lam; p1 = (float)n0; }
                    const int ie = lut_bin(e, S.e_min, S.e_step, S.e_rstep, S.ne);
                    const int it = lut_bin(T, S.t_min, S.t_step, S.t_rstep, S.nt);
                    keys[s] = (uint32_t)(S.bin_base + ie * S.nt + it);
                    o.p0[(size_t)s * n_sbg + sbg] = p0;
                    o.p1[(size_t)s * n_sbg + sbg] = p1;
                    if (o.dbg_e) { o.dbg_e[sbg * cfg.n_species + s] = ie; o.dbg_t[sbg * cfg.n_species + s] = it; }
                    if (cfg.doppler_scheme == 1 && !S.numeric_vn) {
                        // _Hydrometeor.integrate_V, hydrometeors.py:178-199
                        const double v = S.c_v * n0 * (two ? exp(S.e_v * loglam) : exp_sp(S.e_v * loglam));
                        const double n = (S.psd == RADOP_PSD_2MOM) ? ntot_2m : S.c_n * n0 * exp_sp(S.e_n * loglam);
                        vsum += (float)v;
                        nsum += (float)n;
                    }
                }
                // --- wind projection (proj_vel, doppler_scatter.py:30-47; theta is the folded e) --
                const float theta = e * 0.017453292519943295f;
                float st, ct;
                sincosf(theta, &st, &ct);
                wind = (float)(((double)val[RADOP_VAR_U] * sinA1 + (double)val[RADOP_VAR_V] * cosA1) * (double)ct +
                               (double)val[RADOP_VAR_W] * (double)st);
                sinth = st;
            } else {
#pragma unroll
                for (int j = 0; j < RADOP_NVP; ++j) val[j] = qnan;
            }
            if (o.d3_uh) {
                o.d3_uh[sbg] = (double)val[RADOP_VAR_U] * sinA1 + (double)val[RADOP_VAR_V] * cosA1;
                o.d3_w[sbg] = val[RADOP_VAR_W];
                o.d3_rho[sbg] = val[RADOP_VAR_RHO];
                if (scan.traj_sin) o.d3_st[sbg] = make_float2(__ldg(scan.traj_sin + trow), __ldg(scan.traj_tan + trow));
                else {
                    float ef = e_g;
                    if (ef > 90.f) ef = __fsub_rn(180.f, ef);
                    if (ef < 0.f) ef = -ef;
                    double sn, cs;
                    sincos((double)__fmul_rn(ef, 0.017453292519943295f), &sn, &cs);
                    o.d3_st[sbg] = make_float2((float)sn, (float)(sn / cs));
                }
                if (o.d3_std) {
                    double sd = 0.0;
                    if (cfg.brd_mode & 1) {      // spectral_width_turb, doppler_scatter.py:727-758
                        const double r = 0.5 * cfg.brd_res + (double)g * cfg.brd_res;
                        const double edr = (double)val[RADOP_VAR_EDR];
                        if (cfg.brd_sigma_r < 0.1 * r * cfg.brd_sigma_theta)
                            sd += pow(r * edr * cfg.brd_turb_far, 1.0 / 3.0);
                        else {
                            const double x = r * cfg.brd_sigma_theta / cfg.brd_sigma_r;
                            const double den = 11.0 / 15.0 + 4.0 / 15.0 * x * x;
                            sd += pow((edr * cfg.brd_turb_near) / (1.0 / (den * sqrt(den))), 1.0 / 3.0);
                        }
                    }
                    if (cfg.brd_mode & 2) {      // spectral_width_motion, :760-777 (float32 cosine)
                        float ef = e_g;
                        if (ef > 90.f) ef = __fsub_rn(180.f, ef);
                        if (ef < 0.f) ef = -ef;
                        sd += (double)__fmul_rn(cfg.brd_motion_num, cosf(__fmul_rn(ef, 0.017453292519943295f))) / cfg.brd_motion_den;
                    }
                    o.d3_std[sbg] = sd;
                }
            }
            o.active[sbg] = active;
            o.wind[sbg] = wind;
            o.sinth[sbg] = sinth;
            o.vsum[sbg] = vsum;
            o.nsum[sbg] = nsum;
            if (o.vals) {
                float4 *dst = reinterpret_cast<float4 *>(o.vals + (size_t)sbg * RADOP_NVP);
#pragma unroll
                for (int j = 0; j < 4; ++j) dst[j] = make_float4(val[4 * j], val[4 * j + 1], val[4 * j + 2], val[4 * j + 3]);
            }
            for (int s = 0; s < cfg.n_species; ++s) o.key[(size_t)s * n_sbg + sbg] = (uint16_t)keys[s];
        }

        // --- warp-aggregated shared-memory histogram of the LUT bins ---------------------------
        for (int s = 0; s < cfg.n_species; ++s) {
            const uint32_t key = keys[s];
            const unsigned peers = __match_any_sync(0xffffffffu, key);
            if (key != RADOP_KEY_NONE && lane == (unsigned)(__ffs(peers) - 1)) atomicAdd(&s_hist[key], __popc(peers));
        }
      }
    }
    __syncthreads();
    for (int b = tid; b < cfg.total_bins; b += blockDim.x) {
        const int cnt = s_hist[b];
        if (cnt) atomicAdd(&o.hist[b], cnt);
    }
}

// ------------------------------------------------------------------------------------------
// K1b: counting sort by bin
// ------------------------------------------------------------------------------------------
// One block.  bin_start = exclusive scan of the counts; tile_start = exclusive scan of
// ceil(count/TILE); chunk_start = exclusive scan of ceil(tiles/CHUNK) (the work units the
// integration kernel hands out dynamically); totals[] = {pairs, tiles, chunks, bins, pairs of species 0..4}
__global__ void k_scan_bins(const __grid_constant__ DevConfig cfg, const int *__restrict__ hist, int *bin_start,
                            int *tile_start, int *chunk_start, int *cursor, long long *totals)
{
    __shared__ int s_tot[3];
    __shared__ int s_warp[3][32];
    const int tid = threadIdx.x, nthr = blockDim.x;
    __shared__ int s_bins;
    __shared__ int s_species[RADOP_NSPECIES_MAX];
    if (tid == 0) { s_tot[0] = s_tot[1] = s_tot[2] = 0; s_bins = 0; }
    if (tid < RADOP_NSPECIES_MAX) s_species[tid] = 0;
    __syncthreads();
    for (int base = 0; base < cfg.total_bins; base += nthr) {
        int b = base + tid;
        int cnt = 0, tiles = 0, chunks = 0;
        if (b < cfg.total_bins) {
            cnt = hist[b];
            tiles = (cnt + RADOP_TILE - 1) / RADOP_TILE;
            chunks = (tiles + RADOP_CHUNK - 1) / RADOP_CHUNK;
        }
        // block-wide inclusive scans (warp shuffle + shared)
        int v0 = cnt, v1 = tiles, v2 = chunks;
        const int lane = tid & 31, wid = tid >> 5;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int t0 = __shfl_up_sync(0xffffffffu, v0, d), t1 = __shfl_up_sync(0xffffffffu, v1, d);
            int t2 = __shfl_up_sync(0xffffffffu, v2, d);
            if (lane >= d) { v0 += t0; v1 += t1; v2 += t2; }
        }
        if (lane == 31) { s_warp[0][wid] = v0; s_warp[1][wid] = v1; s_warp[2][wid] = v2; }
        __syncthreads();
        if (wid == 0) {
            int nw = nthr >> 5;
            int w0 = lane < nw ? s_warp[0][lane] : 0, w1 = lane < nw ? s_warp[1][lane] : 0;
            int w2 = lane < nw ? s_warp[2][lane] : 0;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                int t0 = __shfl_up_sync(0xffffffffu, w0, d), t1 = __shfl_up_sync(0xffffffffu, w1, d);
                int t2 = __shfl_up_sync(0xffffffffu, w2, d);
                if (lane >= d) { w0 += t0; w1 += t1; w2 += t2; }
            }
            s_warp[0][lane] = w0; s_warp[1][lane] = w1; s_warp[2][lane] = w2;
        }
        __syncthreads();
        int off0 = (wid ? s_warp[0][wid - 1] : 0), off1 = (wid ? s_warp[1][wid - 1] : 0);
        int off2 = (wid ? s_warp[2][wid - 1] : 0);
        if (b < cfg.total_bins) {
            int bs = s_tot[0] + off0 + v0 - cnt;
            bin_start[b] = bs;
            cursor[b] = bs;
            tile_start[b] = s_tot[1] + off1 + v1 - tiles;
            chunk_start[b] = s_tot[2] + off2 + v2 - chunks;
            if (cnt) {
                atomicAdd(&s_bins, 1);
                int sp = 0;
                while (sp + 1 < cfg.n_species && b >= cfg.sp[sp + 1].bin_base) ++sp;
                atomicAdd(&s_species[sp], cnt);
            }
        }
        __syncthreads();
        if (tid == nthr - 1) {
            s_tot[0] += off0 + v0;
            s_tot[1] += off1 + v1;
            s_tot[2] += off2 + v2;
        }
        __syncthreads();
    }
    if (tid == 0) {
        bin_start[cfg.total_bins] = s_tot[0];
        tile_start[cfg.total_bins] = s_tot[1];
        chunk_start[cfg.total_bins] = s_tot[2];
        totals[0] = s_tot[0];
        totals[1] = s_tot[1];
        totals[2] = s_tot[2];
        totals[3] = s_bins;
        for (int sp = 0; sp < RADOP_NSPECIES_MAX; ++sp) totals[4 + sp] = s_species[sp];
    }
}

// scatter the active pairs into bin order (positions inside a bin are arbitrary; every
// result is written back to its own (sub-beam gate, species) slot, so the output does not
// depend on them)
__global__ void __launch_bounds__(256)
k_scatter_pairs(const __grid_constant__ DevConfig cfg, const uint16_t *__restrict__ key, const float *__restrict__ p0,
                const float *__restrict__ p1, const double *__restrict__ sub_weight, int *cursor,
                PairRec *__restrict__ rec, long long n_sbg, const float *__restrict__ wsz)
{
    const int tid = threadIdx.x;
    for (int b = tid; b < cfg.total_bins; b += blockDim.x) s_hist[b] = 0;
    __syncthreads();
    long long sbg = (long long)blockIdx.x * blockDim.x + tid;
    const bool in_range = sbg < n_sbg;
    const unsigned lane = tid & 31;
    uint32_t keys[RADOP_NSPECIES_MAX];
    int rank[RADOP_NSPECIES_MAX];
    for (int s = 0; s < cfg.n_species; ++s) {
        uint32_t k = in_range ? (uint32_t)key[(size_t)s * n_sbg + sbg] : RADOP_KEY_NONE;
        keys[s] = k;
        unsigned peers = __match_any_sync(0xffffffffu, k);
        int leader = __ffs(peers) - 1;
        int base = 0;
        if (k != RADOP_KEY_NONE && (int)lane == leader) base = atomicAdd(&s_hist[k], __popc(peers));
        base = __shfl_sync(0xffffffffu, base, leader);
        rank[s] = base + __popc(peers & ((1u << lane) - 1u));
    }
    __syncthreads();
    for (int b = tid; b < cfg.total_bins; b += blockDim.x) {
        int cnt = s_hist[b];
        if (cnt) s_hist[b] = atomicAdd(&cursor[b], cnt);
    }
    __syncthreads();
    if (in_range) {
        int q = (int)((sbg / cfg.n_gates) % cfg.n_sub);
        float w = wsz ? wsz[sbg] : (float)sub_weight[q];
        for (int s = 0; s < cfg.n_species; ++s) {
            if (keys[s] == RADOP_KEY_NONE) continue;
            PairRec r;
            r.sbg = (uint32_t)sbg;
            r.p0 = p0[(size_t)s * n_sbg + sbg];
            r.p1 = p1[(size_t)s * n_sbg + sbg];
            r.w = w;
            rec[s_hist[keys[s]] + rank[s]] = r;
        }
    }
}

// ------------------------------------------------------------------------------------------
// K2: PSD x scattering-table contraction
// ------------------------------------------------------------------------------------------
#define LOG2E 1.4426950408889634f

// Packed FP32 (sm_100 FFMA2 / FMUL2).  Two (sub-beam gate, species) pairs of a thread share one
// 64-bit register pair, the table value enters as a broadcast 32-bit operand.  Every source
// operand is then an (even, odd) register pair or a scalar, which keeps the register-file
// banks balanced: scalar FFMA with three distinct sources measured 65 % of the FMA peak here,
// the packed form 84 % (scripts/probes/probe_ffma.cu), and it needs half the issue slots.
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c)
{
    unsigned long long ra, rb, rc, rd;
    asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rc) : "f"(c.x), "f"(c.y));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    float2 d;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
    return d;
}
__device__ __forceinline__ float2 fmul2(float2 a, float2 b)
{
    unsigned long long ra, rb, rd;
    asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
    float2 d;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
    return d;
}
__device__ __forceinline__ float2 bcast2(float x) { return make_float2(x, x); }

// nu == 1 on a uniform grid: N(D_k) = N0 D_k^mu exp(-lambda D_k) is geometric in k, so the
// integral  sum_k tab[k][c] exp(-lambda D_k) = exp(-lambda D_0) sum_k tab[k][c] rt^k,
// rt = exp(-lambda dD), is ONE Horner recurrence over all nd diameters: one FMA per diameter
// and column, no exponential in the loop, nothing but the running sums in registers.
// The multiplier is the float32 rounding r of rt, except on one row of every RADOP_HBLOCK rows
// where it is rb = r (rt / r)^HBLOCK (formed in float64 by the caller): the product of the
// multipliers over a block equals rt^HBLOCK to float32 rounding, so the error of r does not
// grow with k (it stays below HBLOCK/2 ulp inside a block).
template <int G, int NC>
__device__ __forceinline__ void integrate_geo(const float4 *__restrict__ tab, int nd, const float2 (&r2)[G / 2],
                                              const float2 (&rb2)[G / 2], float2 (&P)[G / 2][NC])
{
    for (int k0 = nd - RADOP_HBLOCK; k0 >= 0; k0 -= RADOP_HBLOCK) {
#pragma unroll
        for (int j = RADOP_HBLOCK - 1; j >= 0; --j) {
            const float4 t0 = tab[(k0 + j) * 2], t1 = tab[(k0 + j) * 2 + 1];
            const float t[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
#pragma unroll
            for (int g = 0; g < G / 2; ++g) {
                const float2 m = (j == RADOP_HBLOCK / 2) ? rb2[g] : r2[g];
#pragma unroll
                for (int c = 0; c < NC; ++c) P[g][c] = ffma2(P[g][c], m, bcast2(t[c]));
            }
        }
    }
}

// general nu: one ex2 per diameter and pair; NC = columns computed, ROW = floats per table row
template <int G, int NC, int ROW>
__device__ __forceinline__ void integrate_gen(const float4 *__restrict__ tab, const float *__restrict__ U, int nd,
                                              const float (&a)[G], float2 (&acc)[G / 2][NC])
{
#pragma unroll 4
    for (int k = 0; k < nd; ++k) {
        const float u = U[k];
        float t[ROW];
#pragma unroll
        for (int v = 0; v < ROW / 4; ++v) {
            const float4 x = tab[k * (ROW / 4) + v];
            t[4 * v] = x.x; t[4 * v + 1] = x.y; t[4 * v + 2] = x.z; t[4 * v + 3] = x.w;
        }
#pragma unroll
        for (int g = 0; g < G / 2; ++g) {
            const float2 E = make_float2(ex2f_approx(a[2 * g] * u), ex2f_approx(a[2 * g + 1] * u));
#pragma unroll
            for (int c = 0; c < NC; ++c) acc[g][c] = ffma2(E, bcast2(t[c]), acc[g][c]);
        }
    }
}

// 1-mom ice: N(D) = N0 * (490.6 exp(-20.78 x) + 17.46 x^0.6357 exp(-3.29 x)), x = lam' D
// (constants_1mom.py:103-104); both exponentials are geometric on the uniform grid and are
// advanced by multiplication, re-anchored with ex2 every RADOP_HBLOCK diameters.
template <int G>
__device__ __forceinline__ void integrate_ice1(const float4 *__restrict__ tab, const float *__restrict__ Dp, int nd,
                                               const float *__restrict__ Dg, float dD, const float (&lam)[G],
                                               float2 (&acc)[G / 2][10])
{
    float a1[G], a2[G], c2[G];
    float2 r1[G / 2], r2[G / 2];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        a1[g] = -20.78f * LOG2E * lam[g];
        a2[g] = -3.290f * LOG2E * lam[g];
        c2[g] = 17.46f * exp2f(0.6357f * log2f(lam[g]));
    }
#pragma unroll
    for (int g = 0; g < G / 2; ++g) {
        r1[g] = make_float2(ex2f_approx(a1[2 * g] * dD), ex2f_approx(a1[2 * g + 1] * dD));
        r2[g] = make_float2(ex2f_approx(a2[2 * g] * dD), ex2f_approx(a2[2 * g + 1] * dD));
    }
    for (int k0 = 0; k0 < nd; k0 += RADOP_HBLOCK) {
        const float d0 = __ldg(Dg + k0);
        float2 E1[G / 2], E2[G / 2];
#pragma unroll
        for (int g = 0; g < G / 2; ++g) {
            E1[g] = make_float2(490.6f * ex2f_approx(a1[2 * g] * d0), 490.6f * ex2f_approx(a1[2 * g + 1] * d0));
            E2[g] = make_float2(c2[2 * g] * ex2f_approx(a2[2 * g] * d0), c2[2 * g + 1] * ex2f_approx(a2[2 * g + 1] * d0));
        }
#pragma unroll
        for (int j = 0; j < RADOP_HBLOCK; ++j) {
            const int k = k0 + j;
            const float dp = Dp[k];
            const float4 t0 = tab[k * 3], t1 = tab[k * 3 + 1], t2 = tab[k * 3 + 2];
            const float t[10] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w, t2.x, t2.y};
#pragma unroll
            for (int g = 0; g < G / 2; ++g) {
                const float2 w = ffma2(bcast2(dp), E2[g], E1[g]);
#pragma unroll
                for (int c = 0; c < 10; ++c) acc[g][c] = ffma2(w, bcast2(t[c]), acc[g][c]);
                E1[g] = fmul2(E1[g], r1[g]);
                E2[g] = fmul2(E2[g], r2[g]);
            }
        }
    }
}

struct K2Args {
    const int *hist, *bin_start, *tile_start, *chunk_start;
    const long long *totals;
    int *counter;       // dynamic work-unit counter (zeroed before the launch)
    const PairRec *rec;
    float *sz_out;      // [n_species][sbg][8] (sz_plane)
    long long n_sbg;
    float2 *ice_vn;     // [sbg] numeric fall-speed moments of the ice species
    float2 *dop_vn;     // [sbg][n_species] rcs-weighted fall-speed moments (Doppler scheme 2)
};

extern __shared__ __align__(128) unsigned char s_k2[];

// sz_out is stored per species, [species][sub-beam gate][8] (32 B rows: columns 0-3 | 4-6, pad): consecutive
// lanes of the table kernel, of the sub-beam reduction and of the beam attenuation touch consecutive rows,
// 1 KB per warp and request, instead of one 128 B line each as with the species interleaved
__device__ __forceinline__ size_t sz_plane(long long n_sbg, int s, int half, size_t sbg)
{
    return ((size_t)s * (size_t)n_sbg + sbg) * 2 + half;
}
__device__ __forceinline__ void store_sz(float *sz_out, long long n_sbg, int s, uint32_t sbg, const float (&v)[7], float pre)
{
    float4 *dst = reinterpret_cast<float4 *>(sz_out);
    dst[sz_plane(n_sbg, s, 0, sbg)] = make_float4(v[0] * pre, v[1] * pre, v[2] * pre, v[3] * pre);
    dst[sz_plane(n_sbg, s, 1, sbg)] = make_float4(v[4] * pre, v[5] * pre, v[6] * pre, 0.f);
}

#define RADOP_NO_PAIR 0xFFFFFFFFu
__device__ __forceinline__ PairRec load_pair(const PairRec *__restrict__ rec, int first, int n_here, int i)
{
    PairRec r;
    if (i < n_here) {
        const float4 x = __ldg(reinterpret_cast<const float4 *>(rec + first + i));
        r.sbg = __float_as_uint(x.x); r.p0 = x.y; r.p1 = x.z; r.w = x.w;
    } else { r.sbg = RADOP_NO_PAIR; r.p0 = 0.f; r.p1 = 0.f; r.w = 0.f; }
    return r;
}

// Trapezoid rule from a rectangle sum: np.trapz (dx = 1) = sum - (first + last) / 2.
// Doppler scheme 2 (doppler_scatter.py:286-296): v = trapz(V N rcs), n = trapz(N rcs), rcs = 2 pi b.
__device__ __forceinline__ float2 doppler2_moments(float sum_b, float sum_bv, float4 row0a, float4 row0b, float4 rowLa,
                                                   float4 rowLb, float w_first, float w_last, float scale)
{
    const float n = sum_b - 0.5f * (row0a.x * w_first + rowLa.x * w_last);
    const float v = sum_bv - 0.5f * (row0b.w * w_first + rowLb.w * w_last);
    return make_float2(v * scale, n * scale);
}

// one tile of a nu == 1 species: 8 pairs per thread, NC = 7 columns (8 with Doppler scheme 2)
template <int NC>
__device__ __forceinline__ void tile_geo(const K2Args &A, const SpeciesDev &S, int s, int nsp, int first, int n_here,
                                         int tid, const float4 *__restrict__ s_tab, int nd)
{
    constexpr int G = 8;
    float2 r2[G / 2], rb2[G / 2], P[G / 2][NC];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        const float lam = load_pair(A.rec, first, n_here, tid + g * RADOP_TILE_THREADS).p0;
        const float a = -LOG2E * lam;
        // rt = exp(-lambda dD) in float64, r = float32(rt), rb = r (rt / r)^HBLOCK
        const double rt = exp2((double)a * S.dD_d);
        const float r = (float)rt;
        const float rb = (float)((double)r + (double)RADOP_HBLOCK * (rt - (double)r));
        if (g & 1) { r2[g / 2].y = r; rb2[g / 2].y = rb; }
        else { r2[g / 2].x = r; rb2[g / 2].x = rb; }
    }
#pragma unroll
    for (int g = 0; g < G / 2; ++g)
#pragma unroll
        for (int c = 0; c < NC; ++c) P[g][c] = make_float2(0.f, 0.f);
    integrate_geo<G, NC>(s_tab, nd, r2, rb2, P);
    const float D0 = __ldg(S.aux);
#pragma unroll
    for (int g = 0; g < G; ++g) {
        const PairRec rec = load_pair(A.rec, first, n_here, tid + g * RADOP_TILE_THREADS);
        if (rec.sbg == RADOP_NO_PAIR) continue;
        float v[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) v[c] = (g & 1) ? P[g / 2][c].y : P[g / 2][c].x;
        const float a = -LOG2E * rec.p0;
        const float e0 = ex2f_approx(a * D0);
        const float v7[7] = {v[0], v[1], v[2], v[3], v[4], v[5], v[6]};
        store_sz(A.sz_out, A.n_sbg, s, rec.sbg, v7, rec.p1 * rec.w * e0);
        if (NC == 8) {
            const float w_last = ex2f_approx(a * (float)((double)(nd - 1) * S.dD_d));
            A.dop_vn[(size_t)rec.sbg * nsp + s] =
                doppler2_moments(v[0], v[NC - 1], s_tab[0], s_tab[1], s_tab[(nd - 1) * 2], s_tab[(nd - 1) * 2 + 1], 1.f, w_last,
                                 6.283185307179586f * rec.p1 * e0 / S.dD_ref);
        }
    }
}

// one tile of a general-nu species
template <int NC>
__device__ __forceinline__ void tile_gen(const K2Args &A, const SpeciesDev &S, int s, int nsp, int first, int n_here,
                                         int tid, const float4 *__restrict__ s_tab, const float *__restrict__ s_aux, int nd)
{
    constexpr int G = 8;
    float a[G];
    float2 acc[G / 2][NC];
#pragma unroll
    for (int g = 0; g < G; ++g) a[g] = -LOG2E * load_pair(A.rec, first, n_here, tid + g * RADOP_TILE_THREADS).p0;
#pragma unroll
    for (int g = 0; g < G / 2; ++g)
#pragma unroll
        for (int c = 0; c < NC; ++c) acc[g][c] = make_float2(0.f, 0.f);
    integrate_gen<G, NC, 8>(s_tab, s_aux, nd, a, acc);
#pragma unroll
    for (int g = 0; g < G; ++g) {
        const PairRec rec = load_pair(A.rec, first, n_here, tid + g * RADOP_TILE_THREADS);
        if (rec.sbg == RADOP_NO_PAIR) continue;
        float v[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) v[c] = (g & 1) ? acc[g / 2][c].y : acc[g / 2][c].x;
        const float v7[7] = {v[0], v[1], v[2], v[3], v[4], v[5], v[6]};
        store_sz(A.sz_out, A.n_sbg, s, rec.sbg, v7, rec.p1 * rec.w);
        if (NC == 8) {
            const float ag = -LOG2E * rec.p0;
            A.dop_vn[(size_t)rec.sbg * nsp + s] =
                doppler2_moments(v[0], v[NC - 1], s_tab[0], s_tab[1], s_tab[(nd - 1) * 2], s_tab[(nd - 1) * 2 + 1],
                                 ex2f_approx(ag * s_aux[0]), ex2f_approx(ag * s_aux[nd - 1]),
                                 6.283185307179586f * rec.p1 / S.dD_ref);
        }
    }
}

// Persistent CTAs pull work units (RADOP_CHUNK consecutive tiles of one LUT bin) from a global
// counter, the most expensive species first, so that the CTAs drain together whatever the
// mix of species; the bin's table slice is staged in shared memory by TMA bulk copies.
// A tile is RADOP_TILE pairs: every thread integrates RADOP_TILE_G of them (in one pass of 8
// or two passes of 4, depending on how many columns the species carries).
__global__ void __launch_bounds__(RADOP_TILE_THREADS)
k_integrate(const __grid_constant__ DevConfig cfg, K2Args A)
{
    __shared__ uint64_t s_bar;
    __shared__ int s_unit;
    float4 *s_tab = reinterpret_cast<float4 *>(s_k2);
    const int tid = threadIdx.x;
    const int total_units = (int)A.totals[2];
    const int nbins = cfg.total_bins;
    const int nsp = cfg.n_species;
    if (tid == 0) mbar_init(&s_bar, 1);
    uint32_t phase = 0;
    int staged_bin = -1;

    for (;;) {
        __syncthreads();         // the previous unit (and its reads of s_unit / the slice) is finished
        if (tid == 0) s_unit = atomicAdd(A.counter, 1);
        __syncthreads();
        const int unit = total_units - 1 - s_unit;      // bins of the costliest species come last
        if (unit < 0) break;
        int b;
        {   // bin with chunk_start[b] <= unit < chunk_start[b + 1]
            int l = 0, r = nbins;
            while (r - l > 1) {
                int m = (l + r) >> 1;
                if (A.chunk_start[m] <= unit) l = m; else r = m;
            }
            b = l;
        }
        int s = 0;
        while (s + 1 < nsp && b >= cfg.sp[s + 1].bin_base) ++s;
        const SpeciesDev &S = cfg.sp[s];
        if (S.mode == MODE_MELT) continue;      // melting species: k_integrate_melting
        const int cnt = A.hist[b];
        const int tiles = (cnt + RADOP_TILE - 1) / RADOP_TILE;
        const int j0 = (unit - A.chunk_start[b]) * RADOP_CHUNK, j1 = min(tiles, j0 + RADOP_CHUNK);

        // ---- stage the LUT slice of this bin (TMA bulk copy, mbarrier completion) --------------
        const int nd = S.nd;
        float *s_aux = reinterpret_cast<float *>(s_k2 + (size_t)nd * S.ncol * 4);
        if (b != staged_bin) {
            if (tid == 0) {
                fence_proxy_async();
                const uint32_t tab_bytes = (uint32_t)nd * S.ncol * 4;
                const bool need_aux = (S.mode != MODE_GEO8);
                mbar_expect_tx(&s_bar, tab_bytes + (need_aux ? (uint32_t)nd * 4 : 0));
                const char *src = reinterpret_cast<const char *>(S.table + (size_t)(b - S.bin_base) * nd * S.ncol);
                for (uint32_t off = 0; off < tab_bytes; off += 16384) {
                    uint32_t n = min(16384u, tab_bytes - off);
                    tma_bulk_g2s(s_k2 + off, src + off, n, &s_bar);
                }
                if (need_aux) {
                    const float *asrc = S.aux + (S.mode == MODE_ICE1 ? 2 * nd : nd);
                    tma_bulk_g2s(s_aux, asrc, (uint32_t)nd * 4, &s_bar);
                }
            }
            mbar_wait(&s_bar, phase);
            phase ^= 1;
            staged_bin = b;
        }

        for (int j = j0; j < j1; ++j) {
            const int first = A.bin_start[b] + j * RADOP_TILE;
            const int n_here = min(RADOP_TILE, cnt - j * RADOP_TILE);
            if (S.mode == MODE_GEO8) {
                if (cfg.doppler_scheme == 2) tile_geo<8>(A, S, s, nsp, first, n_here, tid, s_tab, nd);
                else tile_geo<RADOP_NC>(A, S, s, nsp, first, n_here, tid, s_tab, nd);
            } else if (S.mode == MODE_GEN8) {
                if (cfg.doppler_scheme == 2) tile_gen<8>(A, S, s, nsp, first, n_here, tid, s_tab, s_aux, nd);
                else tile_gen<RADOP_NC>(A, S, s, nsp, first, n_here, tid, s_tab, s_aux, nd);
            } else {
                // 12-column species (numeric fall-speed moments): two passes of 4 pairs per thread
                constexpr int G = 4;
#pragma unroll 1
                for (int pass = 0; pass < RADOP_TILE_G / G; ++pass) {
                    const int i0 = tid + pass * G * RADOP_TILE_THREADS;
                    if (pass * G * RADOP_TILE_THREADS >= n_here) break;
                    if (S.mode == MODE_GEN12) {
                        float a[G];
                        float2 acc[G / 2][9];
#pragma unroll
                        for (int g = 0; g < G; ++g) a[g] = -LOG2E * load_pair(A.rec, first, n_here, i0 + g * RADOP_TILE_THREADS).p0;
#pragma unroll
                        for (int g = 0; g < G / 2; ++g)
#pragma unroll
                            for (int c = 0; c < 9; ++c) acc[g][c] = make_float2(0.f, 0.f);
                        integrate_gen<G, 9, 12>(s_tab, s_aux, nd, a, acc);
#pragma unroll
                        for (int g = 0; g < G; ++g) {
                            const PairRec rec = load_pair(A.rec, first, n_here, i0 + g * RADOP_TILE_THREADS);
                            if (rec.sbg == RADOP_NO_PAIR) continue;
                            float v[9];
#pragma unroll
                            for (int c = 0; c < 9; ++c) v[c] = (g & 1) ? acc[g / 2][c].y : acc[g / 2][c].x;
                            const float n0 = rec.p1;
                            const float v7[7] = {v[0], v[1], v[2], v[3], v[4], v[5], v[6]};
                            store_sz(A.sz_out, A.n_sbg, s, rec.sbg, v7, n0 * rec.w);
                            A.ice_vn[rec.sbg] = make_float2(v[7] * n0, v[8] * n0);
                        }
                    } else {   // MODE_ICE1
                        float lam[G];
                        float2 acc[G / 2][10];
#pragma unroll
                        for (int g = 0; g < G; ++g) {
                            const PairRec rec = load_pair(A.rec, first, n_here, i0 + g * RADOP_TILE_THREADS);
                            lam[g] = (rec.sbg == RADOP_NO_PAIR) ? 1.0f : rec.p0;
                        }
#pragma unroll
                        for (int g = 0; g < G / 2; ++g)
#pragma unroll
                            for (int c = 0; c < 10; ++c) acc[g][c] = make_float2(0.f, 0.f);
                        integrate_ice1<G>(s_tab, s_aux, nd, S.aux, S.dD_geo, lam, acc);
#pragma unroll
                        for (int g = 0; g < G; ++g) {
                            const PairRec rec = load_pair(A.rec, first, n_here, i0 + g * RADOP_TILE_THREADS);
                            if (rec.sbg == RADOP_NO_PAIR) continue;
                            float v[10];
#pragma unroll
                            for (int c = 0; c < 10; ++c) v[c] = (g & 1) ? acc[g / 2][c].y : acc[g / 2][c].x;
                            // N0 renormalised so that the mass integral equals QM (hydrometeors.py:1330-1339)
                            const float n0 = rec.p1 / v[9];
                            const float v7[7] = {v[0], v[1], v[2], v[3], v[4], v[5], v[6]};
                            store_sz(A.sz_out, A.n_sbg, s, rec.sbg, v7, n0 * rec.w);
                            if (cfg.doppler_scheme == 2) {
                                // shape function at the two end diameters (trapezoid correction)
                                const float D0 = __ldg(S.aux), DL = __ldg(S.aux + nd - 1);
                                const float c2 = 17.46f * exp2f(0.6357f * log2f(rec.p0));
                                const float w0 = 490.6f * ex2f_approx(-20.78f * LOG2E * rec.p0 * D0) +
                                                 c2 * s_aux[0] * ex2f_approx(-3.290f * LOG2E * rec.p0 * D0);
                                const float wL = 490.6f * ex2f_approx(-20.78f * LOG2E * rec.p0 * DL) +
                                                 c2 * s_aux[nd - 1] * ex2f_approx(-3.290f * LOG2E * rec.p0 * DL);
                                const float n_ = v[0] - 0.5f * (s_tab[0].x * w0 + s_tab[(nd - 1) * 3].x * wL);
                                const float v_ = v[7] - 0.5f * (s_tab[1].w * w0 + s_tab[(nd - 1) * 3 + 1].w * wL);
                                const float sc = 6.283185307179586f * n0 / S.dD_ref;
                                A.dop_vn[(size_t)rec.sbg * nsp + s] = make_float2(v_ * sc, n_ * sc);
                            } else {
                                A.ice_vn[rec.sbg] = make_float2(v[7] * n0, v[8] * n0);
                            }
                        }
                    }
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// K2t: the contraction through lambda-transformed tables
// ------------------------------------------------------------------------------------------
// For every species the PSD is a ONE-parameter family of shapes: N(D) = N0 * shape(lambda, D)
// (hydrometeors.py:128-152 generalised gamma, :1231-1254 + constants_1mom.py:103-104 Field 2005).
// The diameter integral of one (e, t) bin and one combined column,
//     F_c(lambda) = sum_k tab[k][c] shape(lambda, D_k),
// is therefore a function of lambda alone -- a discrete Laplace-type transform of the table
// column -- and it is analytic in u = log2(lambda).  k_build_lamtab evaluates it in float64 at
// Chebyshev nodes of eighth-octave pieces of u, once per table upload, in the scaled form
//     F_c = 2^{L_c(u)} R_c(u),   L_c = log2 sum_k |tab[k][c]| shape,   R_c = F_c / 2^{L_c} in [-1, 1]
// (envelope and signed ratio: the relative accuracy is that of the float32 sum itself, whatever
// the signs of the column), fits degree-5 polynomials, and VERIFIES every piece at two more
// nodes (x = 0, x = 1, extrema of T_6) against the float64 sum: a piece that misses 2e-7 is
// marked and its pairs are integrated directly, diameter by diameter.  A pair then costs 12 NC
// coefficients (384 B, shared by neighbouring gates: same bin, same piece) and ~100 FMA instead
// of nd x NC.  radop_set_integration(1) selects the direct kernel (k_integrate) instead; both are
// held to the oracle by the same tests and to each other by test_table_vs_direct_integration;
// tests/test_lamtab_math.py restates the construction in numpy against the oracle on the CPU.
#define LT_U0 (-10)              // u = log2(lambda) covered: [LT_U0, LT_U0 + LT_OCT)
#define LT_OCT 22
#define LT_PPO 8                 // pieces per octave
#define LT_NP (LT_OCT * LT_PPO)
#define LT_BAD ((int)0x80000000)
#define LT_TOL 2e-7

__host__ __device__ inline int lt_columns(int mode, int doppler_scheme)
{
    if (mode == MODE_ICE1) return 10;
    if (mode == MODE_GEN12) return 9;
    return doppler_scheme == 2 ? 8 : 7;
}
// floats per (bin, piece): 6 x 2 NC coefficients, NC integer exponents, one mask word, padded to 128 B lines
__host__ __device__ inline int lt_stride(int nc) { return (12 * nc + ((nc + 1 + 3) & ~3) + 31) / 32 * 32; }

// shape(lambda, D_k) / 2^{shift(lambda)} in float64 (the common factor keeps the sum in range)
__device__ __forceinline__ double lt_weight(const SpeciesDev &S, double lam, double lam_p, int k, double y0)
{
    const int nd = S.nd;
    if (S.mode == MODE_GEO8) return exp(-lam * (double)k * S.dD_d);
    if (S.mode == MODE_ICE1) {
        const double D = (double)__ldg(S.aux + k);
        return 490.6 * exp(-lam * (20.78 * D - 3.29 * y0)) +
               17.46 * lam_p * (double)__ldg(S.aux + 2 * nd + k) * exp(-3.29 * lam * (D - y0));
    }
    return exp(-lam * ((double)__ldg(S.aux + nd + k) - y0));
}
__device__ __forceinline__ double lt_y0(const SpeciesDev &S)
{
    return (double)__ldg(S.aux + ((S.mode == MODE_GEO8 || S.mode == MODE_ICE1) ? 0 : S.nd));
}
// log2 of the common factor
__device__ __forceinline__ double lt_shift(const SpeciesDev &S, double lam, double y0)
{
    return -(S.mode == MODE_ICE1 ? 3.29 : 1.0) * lam * y0 * 1.4426950408889634;
}

// one CTA per (bin, piece); warp = node (6 Chebyshev nodes + 2 check nodes), lanes over diameters
__global__ void __launch_bounds__(256)
k_build_lamtab(const __grid_constant__ SpeciesDev S, int nc, int stride, float *__restrict__ out, int *__restrict__ n_bad)
{
    __shared__ double s_v[8][20];
    __shared__ double s_err[20];
    __shared__ int s_flag;
    const int bin = blockIdx.x / LT_NP, piece = blockIdx.x % LT_NP;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) s_flag = 0;
    __syncthreads();
    {
        const double x = warp < 6 ? cospi(((double)warp + 0.5) / 6.0) : (warp == 6 ? 0.0 : 1.0);
        const double u = (double)LT_U0 + ((double)piece + 0.5 * (x + 1.0)) / (double)LT_PPO;
        const double lam = exp2(u), lam_p = pow(lam, 0.6357), y0 = lt_y0(S);
        double F[10], A[10];
#pragma unroll
        for (int c = 0; c < 10; ++c) { F[c] = 0.0; A[c] = 0.0; }
        const float *tab = S.table + (size_t)bin * S.nd * S.ncol;
        for (int k = lane; k < S.nd; k += 32) {
            const double w = lt_weight(S, lam, lam_p, k, y0);
            const float *row = tab + (size_t)k * S.ncol;
#pragma unroll
            for (int c = 0; c < 10; ++c)
                if (c < nc) {
                    const double t = (double)__ldg(row + c);
                    F[c] = fma(w, t, F[c]);
                    A[c] = fma(w, fabs(t), A[c]);
                }
        }
#pragma unroll
        for (int c = 0; c < 10; ++c)
#pragma unroll
            for (int d = 16; d >= 1; d >>= 1) {
                F[c] += __shfl_xor_sync(0xffffffffu, F[c], d);
                A[c] += __shfl_xor_sync(0xffffffffu, A[c], d);
            }
        if (lane == 0) {
            const double sh = lt_shift(S, lam, y0);
#pragma unroll
            for (int c = 0; c < 10; ++c)
                if (c < nc) {
                    double L = 0.0, R = 0.0;
                    if (A[c] > 0.0 && A[c] < 1e300) { L = log2(A[c]) + sh; R = F[c] / A[c]; }
                    else if (A[c] == 0.0) atomicOr(&s_flag, 1 << 1);      // empty column (or total underflow)
                    else atomicOr(&s_flag, 1);
                    if (A[c] > 0.0) atomicOr(&s_flag, 1 << 2);
                    s_v[warp][c] = L;
                    s_v[warp][nc + c] = R;
                }
        }
    }
    __syncthreads();
    float *dst = out + ((size_t)bin * LT_NP + piece) * stride;
    const int f = threadIdx.x;
    double p[6] = {0, 0, 0, 0, 0, 0};
    if (f < 2 * nc) {
        double a[6];
#pragma unroll
        for (int n = 0; n < 6; ++n) {
            double acc = 0.0;
#pragma unroll
            for (int j = 0; j < 6; ++j) acc += s_v[j][f] * cospi((double)n * ((double)j + 0.5) / 6.0);
            a[n] = acc * (n == 0 ? 1.0 : 2.0) / 6.0;
        }
        // Chebyshev -> power basis (degree 5)
        p[0] = a[0] - a[2] + a[4];
        p[1] = a[1] - 3.0 * a[3] + 5.0 * a[5];
        p[2] = 2.0 * a[2] - 8.0 * a[4];
        p[3] = 4.0 * a[3] - 20.0 * a[5];
        p[4] = 8.0 * a[4];
        p[5] = 16.0 * a[5];
        const double e0 = fabs(p[0] - s_v[6][f]);
        const double e1 = fabs(p[0] + p[1] + p[2] + p[3] + p[4] + p[5] - s_v[7][f]);
        // the float32 evaluation carries ~1e-7 of the range the polynomial sweeps inside the piece anyway
        const double range = fmax(1.0, fabs(p[1]) + fabs(p[2]) + fabs(p[3]) + fabs(p[4]) + fabs(p[5]));
        s_err[f] = fmax(e0, e1) * (f < nc ? 0.6931471805599453 : 1.0) / range;
    }
    __syncthreads();
    // a piece is unusable when a node was out of range, when a column is empty at some nodes only
    // (underflow), or when the fit misses the tolerance at the check nodes
    bool bad = (s_flag & 1) != 0;
    for (int c = 0; c < nc; ++c) {
        const double e = s_err[c] + s_err[nc + c];
        if (!(e <= LT_TOL)) bad = true;
    }
    if ((s_flag & 6) == 6) {
        // mixed empty / non-empty: fine when whole columns are empty at all 8 nodes, else underflow
        for (int c = 0; c < nc; ++c) {
            int zero = 0;
            for (int j = 0; j < 8; ++j) zero += (s_v[j][c] == 0.0 && s_v[j][nc + c] == 0.0);
            if (zero != 0 && zero != 8) bad = true;
        }
    }
    // layout per piece: [column][L coefficients 0..5] | [column][R coefficients 0..5] | nc exponent words |
    // mask word: bit c = R_c is the constant +1 or -1 (a column of one sign), bit 16 + c = its sign is minus
    if (f < 2 * nc) {
        if (f < nc) {
            const int L0 = (int)rint(p[0]);
            p[0] -= (double)L0;
            reinterpret_cast<int *>(dst)[12 * nc + f] = (bad && f == 0) ? LT_BAD : L0;
        }
        float *o = (f < nc) ? dst + f * 6 : dst + 6 * nc + (f - nc) * 6;
#pragma unroll
        for (int j = 0; j < 6; ++j) o[j] = (float)p[j];
    }
    if (f == 0) {
        unsigned mask = 0;
        for (int c = 0; c < nc; ++c) {
            bool plus = true, minus = true;
            for (int j = 0; j < 8; ++j) {
                if (fabs(s_v[j][nc + c] - 1.0) > 1e-12) plus = false;
                if (fabs(s_v[j][nc + c] + 1.0) > 1e-12) minus = false;
            }
            if (plus) mask |= 1u << c;
            if (minus) mask |= (1u << c) | (1u << (16 + c));
        }
        reinterpret_cast<unsigned *>(dst)[13 * nc] = mask;
    }
    if (bad && f == 0) atomicAdd(n_bad, 1);
}

// v * 2^e in float32 for any integer e (two exact power-of-two factors; what float32 cannot hold
// flushes to zero / infinity as a float32 sum would)
__device__ __forceinline__ float lt_scale(float v, int e)
{
    e = max(-252, min(252, e));
    const int e1 = e / 2, e2 = e - e1;
    return (v * __int_as_float((e1 + 127) << 23)) * __int_as_float((e2 + 127) << 23);
}

// which piece of the table serves lambda, and where inside it; a non-zero return says why the pair needs
// the direct sum instead (2 below / 3 above the tabulated range, 4 not a positive normal number; 1 = rejected
// piece, found when the piece is read)
__device__ __forceinline__ int lt_locate(float lam, int &piece, float &x)
{
    piece = 0; x = 0.f;
    if (!(lam > 0.f) || !(lam < 3.0e38f)) return 4;
    const int bits = __float_as_int(lam);
    const int e = (bits >> 23) - 127;
    if (e < -126) return 4;
    const float mant = __int_as_float((bits & 0x007fffff) | 0x3f800000);   // [1, 2)
    const float t = (float)LT_PPO * __log2f(mant);                         // [0, LT_PPO)
    const int it = max(0, min(LT_PPO - 1, (int)t));
    const int pc = (e - LT_U0) * LT_PPO + it;
    if (pc < 0) return 2;
    if (pc >= LT_NP) return 3;
    piece = pc;
    x = 2.f * (t - (float)it) - 1.f;
    return 0;
}

// F_c(lambda) = m[c] * 2^ex[c] from the coefficients of one piece.  The envelope polynomials L_c are
// read two columns at a time (12 floats = three 16 B loads); the ratio polynomial R_c is read only for the
// columns that change sign along the diameter axis -- for a column of one sign it is the constant +-1 and
// the piece's mask word says so (in the shipped table families every column but s10 - s8 is of one sign,
// so a pair reads ~1.5 NC + 3 chunks instead of 3 NC + 3; the kernel is bound by the L1 wavefronts of
// exactly these loads, ncu: l1tex__data_pipe_lsu_wavefronts 84 %).
template <int NC>
__device__ __forceinline__ void lt_horner(const float *__restrict__ base, float x, float (&m)[NC], int (&ex)[NC])
{
    const float4 *src = reinterpret_cast<const float4 *>(base);
    int L0[((NC + 1) + 3) & ~3];
#pragma unroll
    for (int i = 0; i < (NC + 1 + 3) / 4; ++i) {
        const float4 v = __ldg(src + 3 * NC + i);
        L0[4 * i] = __float_as_int(v.x); L0[4 * i + 1] = __float_as_int(v.y);
        L0[4 * i + 2] = __float_as_int(v.z); L0[4 * i + 3] = __float_as_int(v.w);
    }
    const unsigned mask = (unsigned)L0[NC];
    float Lv[NC];
#pragma unroll
    for (int k2 = 0; k2 < NC / 2; ++k2) {
        const float4 a = __ldg(src + 3 * k2), b = __ldg(src + 3 * k2 + 1), c = __ldg(src + 3 * k2 + 2);
        float L = b.y, M = c.w;
        L = fmaf(L, x, b.x); M = fmaf(M, x, c.z);
        L = fmaf(L, x, a.w); M = fmaf(M, x, c.y);
        L = fmaf(L, x, a.z); M = fmaf(M, x, c.x);
        L = fmaf(L, x, a.y); M = fmaf(M, x, b.w);
        L = fmaf(L, x, a.x); M = fmaf(M, x, b.z);
        Lv[2 * k2] = L; Lv[2 * k2 + 1] = M;
    }
    if (NC & 1) {
        const float4 a = __ldg(src + 3 * (NC / 2)), b = __ldg(src + 3 * (NC / 2) + 1);
        float L = b.y;
        L = fmaf(L, x, b.x);
        L = fmaf(L, x, a.w);
        L = fmaf(L, x, a.z);
        L = fmaf(L, x, a.y);
        L = fmaf(L, x, a.x);
        Lv[NC - 1] = L;
    }
#pragma unroll
    for (int k = 0; k < NC; ++k) {
        float R;
        if (mask & (1u << k)) {
            R = (mask & (1u << (16 + k))) ? -1.f : 1.f;
        } else {
            const float2 *r = reinterpret_cast<const float2 *>(base + 6 * NC + 6 * k);
            const float2 r0 = __ldg(r), r1 = __ldg(r + 1), r2 = __ldg(r + 2);
            R = r2.y;
            R = fmaf(R, x, r2.x);
            R = fmaf(R, x, r1.y);
            R = fmaf(R, x, r1.x);
            R = fmaf(R, x, r0.y);
            R = fmaf(R, x, r0.x);
        }
        m[k] = ex2f_approx(Lv[k]) * R;
        ex[k] = L0[k];
    }
}

// the same sums diameter by diameter, by a whole warp for ONE pair (pairs outside the tabulated
// range or on a rejected piece): lanes stride over the diameters, shuffle reduction
template <int NC>
__device__ __noinline__ void warp_direct(const SpeciesDev &S, int bin_local, float lam, int lane, float (&dv)[NC])
{
    const int nd = S.nd;
    float acc[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) acc[c] = 0.f;
    const float a = -LOG2E * lam;
    const float y0 = __ldg(S.aux + ((S.mode == MODE_GEO8 || S.mode == MODE_ICE1) ? 0 : nd));
    // everything underflows float32 (the direct kernel returns zeros there as well); NaN runs the loop
    const bool dead = (S.mode == MODE_ICE1 ? 3.29f : 1.f) * a * y0 <= -250.f;
    if (!dead) {
        const float *tab = S.table + (size_t)bin_local * nd * S.ncol;
        const float c2 = (S.mode == MODE_ICE1) ? 17.46f * exp2f(0.6357f * log2f(lam)) : 0.f;
        for (int k = lane; k < nd; k += 32) {
            float w;
            if (S.mode == MODE_GEO8) w = exp2f(a * (float)((double)y0 + (double)k * S.dD_d));
            else if (S.mode == MODE_ICE1) {
                const float D = __ldg(S.aux + k);
                w = 490.6f * exp2f(20.78f * a * D) + c2 * __ldg(S.aux + 2 * nd + k) * exp2f(3.29f * a * D);
            } else w = exp2f(a * __ldg(S.aux + nd + k));
            const float *row = tab + (size_t)k * S.ncol;
#pragma unroll
            for (int c = 0; c < NC; ++c) acc[c] = fmaf(w, __ldg(row + c), acc[c]);
        }
#pragma unroll
        for (int c = 0; c < NC; ++c)
#pragma unroll
            for (int d = 16; d >= 1; d >>= 1) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], d);
    }
#pragma unroll
    for (int c = 0; c < NC; ++c) dv[c] = acc[c];
}

struct K2tArgs {
    const uint16_t *key;     // [species][sbg] global bin id or RADOP_KEY_NONE
    const float *p0, *p1;    // [species][sbg]
    const double *sub_weight;
    const float *wsz;        // integration scheme 'ml': normalised per-gate weights [sbg], else null
    float *sz_out;
    float2 *ice_vn, *dop_vn;
    long long n_sbg;
    unsigned long long *n_fallback;   // [4] pairs integrated directly, by reason (lt_locate; rejected piece)
};

// One species of one warp's 32 sub-beam gates: table evaluation per lane (the coefficient loads of lanes
// that share a piece are broadcasts), then the lanes whose pair the table cannot serve are integrated
// directly one after the other by the whole warp.  Staging the pieces of a warp in shared memory first
// (one coalesced fetch per distinct piece) was measured and is slower: 2.6 ms against 2.0 ms at cfg 3.
// Must be called by all 32 lanes (have = this lane holds an active pair of the species).
template <int NC>
__device__ __forceinline__ void pair_tab(const DevConfig &cfg, const K2tArgs &A, const SpeciesDev &S, int s, bool have,
                                         int bin, uint32_t sbg, float p0, float p1, float w, int lane, float (&acc)[7])
{
    constexpr int STR = (12 * NC + ((NC + 1 + 3) & ~3) + 31) / 32 * 32;      // lt_stride(NC)
    const int nsp = cfg.n_species, nd = S.nd;
    const int bl = bin - S.bin_base;
    float m[NC];
    int ex[NC];
    int piece;
    float x;
    int why = 0;
    if (have) why = lt_locate(p0, piece, x);
    const int id = (have && !why) ? bl * LT_NP + piece : -1;
    // out-of-range lanes read piece 0 of their bin (valid memory), the result is discarded
    lt_horner<NC>(S.lamtab + (size_t)(id >= 0 ? id : bl * LT_NP) * STR, x, m, ex);
    if (id >= 0 && ex[0] == LT_BAD) why = 1;
    unsigned need = __ballot_sync(0xffffffffu, why != 0);
    if (why) atomicAdd(A.n_fallback + (why - 1), 1ull);
    while (need) {
        const int src = __ffs(need) - 1;
        need &= need - 1;
        const int bl_s = __shfl_sync(0xffffffffu, bl, src);
        const float lam_s = __shfl_sync(0xffffffffu, p0, src);
        float ds[NC];
        warp_direct<NC>(S, bl_s, lam_s, lane, ds);
        if (lane == src) {
#pragma unroll
            for (int c = 0; c < NC; ++c) { m[c] = ds[c]; ex[c] = 0; }
        }
    }
    if (!have) return;
    float v7[7];
    if (S.mode == MODE_ICE1) {
        // N0 renormalised so that the mass integral equals QM (hydrometeors.py:1330-1339): n0 = p1 / F_9
        constexpr int C9 = NC > 9 ? 9 : 0, C8 = NC > 8 ? 8 : 0, C7 = NC > 7 ? 7 : 0;
        const float inv9 = 1.f / m[C9];
        const int e9 = ex[C9];
#pragma unroll
        for (int c = 0; c < 7; ++c) v7[c] = lt_scale(m[c] * inv9 * p1 * w, ex[c] - e9);
        if (cfg.sum_species) {
#pragma unroll
            for (int c = 0; c < 7; ++c) acc[c] += v7[c];
        } else store_sz(A.sz_out, A.n_sbg, s, sbg, v7, 1.f);
        if (cfg.doppler_scheme == 2) {
            const float *tab = S.table + (size_t)bl * nd * S.ncol;
            const float D0 = __ldg(S.aux), DL = __ldg(S.aux + nd - 1);
            const float c2 = 17.46f * exp2f(0.6357f * log2f(p0));
            const float w0 = 490.6f * ex2f_approx(-20.78f * LOG2E * p0 * D0) +
                             c2 * __ldg(S.aux + 2 * nd) * ex2f_approx(-3.290f * LOG2E * p0 * D0);
            const float wL = 490.6f * ex2f_approx(-20.78f * LOG2E * p0 * DL) +
                             c2 * __ldg(S.aux + 3 * nd - 1) * ex2f_approx(-3.290f * LOG2E * p0 * DL);
            const float f0 = lt_scale(m[0], ex[0]), f7 = lt_scale(m[C7], ex[C7]);
            const float n0 = p1 / lt_scale(m[C9], e9);
            const float n_ = f0 - 0.5f * (__ldg(tab) * w0 + __ldg(tab + (size_t)(nd - 1) * S.ncol) * wL);
            const float v_ = f7 - 0.5f * (__ldg(tab + 7) * w0 + __ldg(tab + (size_t)(nd - 1) * S.ncol + 7) * wL);
            const float sc = 6.283185307179586f * n0 / S.dD_ref;
            A.dop_vn[(size_t)sbg * nsp + s] = make_float2(v_ * sc, n_ * sc);
        } else {
            A.ice_vn[sbg] = make_float2(lt_scale(m[C7] * inv9 * p1, ex[C7] - e9), lt_scale(m[C8] * inv9 * p1, ex[C8] - e9));
        }
        return;
    }
    const float pre = p1 * w;
#pragma unroll
    for (int c = 0; c < 7; ++c) v7[c] = lt_scale(m[c] * pre, ex[c]);
    if (cfg.sum_species) {
#pragma unroll
        for (int c = 0; c < 7; ++c) acc[c] += v7[c];
    } else store_sz(A.sz_out, A.n_sbg, s, sbg, v7, 1.f);
    if (S.mode == MODE_GEN12) {
        constexpr int C8 = NC > 8 ? 8 : 0, C7 = NC > 7 ? 7 : 0;
        A.ice_vn[sbg] = make_float2(lt_scale(m[C7] * p1, ex[C7]), lt_scale(m[C8] * p1, ex[C8]));
    } else if (NC == 8) {
        // trapezoid end corrections of Doppler scheme 2 (doppler2_moments)
        constexpr int C7 = NC > 7 ? 7 : 0;
        const float *tab = S.table + (size_t)bl * nd * S.ncol;
        const float a = -LOG2E * p0;
        float yF, yL;
        if (S.mode == MODE_GEO8) { yF = __ldg(S.aux); yL = (float)((double)yF + (double)(nd - 1) * S.dD_d); }
        else { yF = __ldg(S.aux + nd); yL = __ldg(S.aux + 2 * nd - 1); }
        const float wF = ex2f_approx(a * yF), wL = ex2f_approx(a * yL);
        const float f0 = lt_scale(m[0], ex[0]), f7 = lt_scale(m[C7], ex[C7]);
        const float n_ = f0 - 0.5f * (__ldg(tab) * wF + __ldg(tab + (size_t)(nd - 1) * S.ncol) * wL);
        const float v_ = f7 - 0.5f * (__ldg(tab + 7) * wF + __ldg(tab + (size_t)(nd - 1) * S.ncol + 7) * wL);
        const float sc = 6.283185307179586f * p1 / S.dD_ref;
        A.dop_vn[(size_t)sbg * nsp + s] = make_float2(v_ * sc, n_ * sc);
    }
}

// Thread per sub-beam gate, in ray order (neighbouring gates share the bin and mostly the piece, so the
// coefficient loads of a warp are broadcasts out of L1; no sort is needed).  Every iteration first issues
// the key loads of all species, then the lambda / N0 loads of the active ones, then evaluates the species
// one after the other.
template <int MINB>
__global__ void __launch_bounds__(256, MINB)
k_integrate_tab(const __grid_constant__ DevConfig cfg, K2tArgs A)
{
    const int ng = cfg.n_gates;
    const int lane = threadIdx.x & 31;
    // warp-uniform trip count: staging and the direct integration of a pair are warp-collective steps
    for (long long base = (long long)blockIdx.x * blockDim.x + (threadIdx.x & ~31); base < A.n_sbg;
         base += (long long)gridDim.x * blockDim.x) {
        const long long sbg = base + lane;
        const bool in_range = sbg < A.n_sbg;
        unsigned long long kp[2] = {~0ull, ~0ull};          // 16-bit keys of up to 8 species
        float P0[RADOP_NSPECIES_MAX], P1[RADOP_NSPECIES_MAX];
#pragma unroll
        for (int s = 0; s < RADOP_NSPECIES_MAX; ++s)
            if (in_range && s < cfg.n_species) {
                const unsigned long long k = (unsigned long long)__ldg(A.key + (size_t)s * A.n_sbg + sbg);
                kp[s >> 2] = (kp[s >> 2] & ~(0xffffull << (16 * (s & 3)))) | (k << (16 * (s & 3)));
            }
#pragma unroll
        for (int s = 0; s < RADOP_NSPECIES_MAX; ++s) {
            P0[s] = 1.f; P1[s] = 0.f;
            if (((kp[s >> 2] >> (16 * (s & 3))) & 0xffffull) != RADOP_KEY_NONE) {
                P0[s] = __ldg(A.p0 + (size_t)s * A.n_sbg + sbg);
                P1[s] = __ldg(A.p1 + (size_t)s * A.n_sbg + sbg);
            }
        }
        const int q = in_range ? (int)(((unsigned)sbg / (unsigned)ng) % (unsigned)cfg.n_sub) : 0;
        const float w = (A.wsz && in_range) ? __ldg(A.wsz + sbg) : (float)__ldg(A.sub_weight + q);
        // sum over the species in the reference's order (doppler_scatter.py:148-158), kept in registers
        float acc[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        bool any = false;
#pragma unroll 1
        for (int s = 0; s < cfg.n_species; ++s) {
            const SpeciesDev &S = cfg.sp[s];
            if (S.mode == MODE_MELT) continue;
            const uint32_t k = (uint32_t)((kp[s >> 2] >> (16 * (s & 3))) & 0xffffull);
            const bool have = k != RADOP_KEY_NONE;
            if (!__any_sync(0xffffffffu, have)) continue;
            any |= have;
            const float p0 = P0[s], p1 = P1[s];
            const int bin = have ? (int)k : S.bin_base;
            switch (S.lt_nc) {
            case 7: pair_tab<7>(cfg, A, S, s, have, bin, (uint32_t)sbg, p0, p1, w, lane, acc); break;
            case 8: pair_tab<8>(cfg, A, S, s, have, bin, (uint32_t)sbg, p0, p1, w, lane, acc); break;
            case 9: pair_tab<9>(cfg, A, S, s, have, bin, (uint32_t)sbg, p0, p1, w, lane, acc); break;
            default: pair_tab<10>(cfg, A, S, s, have, bin, (uint32_t)sbg, p0, p1, w, lane, acc); break;
            }
        }
        if (cfg.sum_species && any) store_sz(A.sz_out, A.n_sbg, 0, (uint32_t)sbg, acc, 1.f);
    }
}

#include "radop_fused.cuh"

// ------------------------------------------------------------------------------------------
// K3: quadrature / species sums and the per-ray epilogue
// ------------------------------------------------------------------------------------------
// IceParticle.integrate_V sums over ALL valid gates of a sub-beam and the total lands on the
// first valid gate (hydrometeors.py:1256-1275 + utilities.py:231-261 zero padding): one warp
// per (ray, sub-beam) computes that total in a fixed order.
__global__ void k_ice_totals(const __grid_constant__ DevConfig cfg, const uint8_t *__restrict__ active,
                             const float2 *__restrict__ ice_vn, float *__restrict__ tot /* [ray*sub][3] */)
{
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    int n_rq = cfg.n_rays * cfg.n_sub;
    if (warp >= n_rq) return;
    const uint8_t bit = (uint8_t)(1u << cfg.ice_numeric);
    size_t base = (size_t)warp * cfg.n_gates;
    float v = 0.f, n = 0.f;
    int first = 0x7fffffff;
    for (int g = lane; g < cfg.n_gates; g += 32) {
        if (active[base + g] & bit) {
            float2 x = ice_vn[base + g];
            v += x.x; n += x.y;
            first = min(first, g);
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        v += __shfl_xor_sync(0xffffffffu, v, d);
        n += __shfl_xor_sync(0xffffffffu, n, d);
        first = min(first, __shfl_xor_sync(0xffffffffu, first, d));
    }
    if (lane == 0) {
        tot[3 * warp] = v;
        tot[3 * warp + 1] = n;
        tot[3 * warp + 2] = __int_as_float(first);
    }
}

// ------------------------------------------------------------------------------------------
// K2m: melting snow / melting graupel (hydrometeors.py:303-500, 1393-1481; doppler_scatter.py:195-268).
// One warp per active (sub-beam gate, melting species) pair, lanes over the diameters.  The PSD has
// no closed form along the diameter axis: every gate has its own diameter grid (limits blended
// with the wet fraction), and
//     N(D) = prop * N_rain(D_r) V_rain(D_r) / V(D) dD_r/dD,   D_r = (m(D) / (pi/6 rho_w))^(1/3),
//     m(D) = fw^2 m_rain(D) + (1 - fw^2) m_dry(D),            V = phi V_rain(D_r) + (1 - phi) V_dry(D),
// with dD_r/dD the reference's forward difference (h = 0.01 mm) and prop = Q / sum(N m dD) so that the
// mass is conserved.  Everything is float64 as in the reference (the pairs of a melting layer are few).
struct K2mArgs {
    const PairRec *rec;
    const uint16_t *key;     // [species][n_sbg]
    const int *bin_start;
    float *sz_out;
    float2 *dop_vn;
    long long n_sbg;
};

__device__ __forceinline__ double pow_small(double x, double b)
{
    if (b == 3.0) return x * x * x;
    if (b == 2.0) return x * x;
    if (b == 1.0) return x;
    if (b == 0.5) return sqrt(x);
    if (b == 0.0) return 1.0;
    return pow(x, b);
}

__global__ void __launch_bounds__(128)
k_integrate_melting(const __grid_constant__ DevConfig cfg, K2mArgs A)
{
    const int lane = threadIdx.x & 31;
    const int warp = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const int n_warps = (int)(((long long)gridDim.x * blockDim.x) >> 5);
    const int nsp = cfg.n_species;
    for (int s = 0; s < nsp; ++s) {
        const SpeciesDev &S = cfg.sp[s];
        if (S.mode != MODE_MELT) continue;
        const int first = A.bin_start[S.bin_base], last = A.bin_start[S.bin_base + S.ne * S.nt];
        const int nd = S.nd;
        const double ar = S.melt[0], br = S.melt[1], alr = S.melt[2], ber = S.melt[3];
        const double as = S.a, bs = S.b, als = S.alpha, bes = S.beta;
        const double inv_vol = 1.0 / (3.141592653589793 / 6.0 * (1000. / (1000. * 1000. * 1000.)));   // 1 / (pi/6 rho_w)
        for (int i = first + warp; i < last; i += n_warps) {
            const PairRec rec = A.rec[i];
            const uint32_t sbg = rec.sbg;
            const double fw = (double)rec.p0, q = (double)rec.p1;
            const int b = (int)A.key[(size_t)s * A.n_sbg + sbg] - S.bin_base;
            const float4 *tab = reinterpret_cast<const float4 *>(S.table + (size_t)b * nd * 8);
            // wet-fraction dependent diameter limits (f_wet setter, hydrometeors.py:333-340), vlinspace
            const double dmax = fw * S.melt[5] + (1.0 - fw) * S.melt[7];
            const double dmin = fw * S.melt[4] + (1.0 - fw) * S.melt[6];
            const double step = (dmax - dmin) / (double)(nd - 1);
            const double dD = (dmin + step) - dmin;
            const double fw2 = fw * fw;
            const double phi = 0.246 * fw + (1.0 - 0.246) * fw * fw * fw * fw * fw * fw * fw;
            // equivalent rain with the mass of the melting species (Rain.set_psd, hydrometeors.py:761-772)
            const double lam = pow(S.lambda_factor / q, 1.0 / (4.0 + S.mu));
            double acc[7] = {0., 0., 0., 0., 0., 0., 0.};
            double sm = 0.0, sv = 0.0, sn = 0.0, edge = 0.0, edge_v = 0.0;
            for (int k = lane; k < nd; k += 32) {
                const double D = dmin + step * (double)k;
                const double D2 = D + 0.01;
                const double m = fw2 * (ar * pow_small(D, br)) + (1.0 - fw2) * (as * pow_small(D, bs));
                const double m2 = fw2 * (ar * pow_small(D2, br)) + (1.0 - fw2) * (as * pow_small(D2, bs));
                const double Dr = cbrt(m * inv_vol), Dr2 = cbrt(m2 * inv_vol);
                const double dDr = (Dr2 - Dr) / 0.01;
                const double Vr = alr * pow_small(Dr, ber);
                const double V = phi * Vr + (1.0 - phi) * (als * pow_small(D, bes));
                const double Nr = S.N0 * pow_small(Dr, S.mu) * exp(-lam * pow_small(Dr, S.nu));
                const double N = Nr * Vr / V * dDr;
                const float4 t0 = __ldg(tab + 2 * k), t1 = __ldg(tab + 2 * k + 1);
                acc[0] += (double)t0.x * N; acc[1] += (double)t0.y * N; acc[2] += (double)t0.z * N; acc[3] += (double)t0.w * N;
                acc[4] += (double)t1.x * N; acc[5] += (double)t1.y * N; acc[6] += (double)t1.z * N;
                sm += N * m;
                if (cfg.doppler_scheme == 2) {
                    const double nb = N * (double)t0.x;       // N rcs / (2 pi)
                    sn += nb; sv += nb * V;
                    if (k == 0 || k == nd - 1) { edge += nb; edge_v += nb * V; }
                } else { sn += N; sv += N * V; }
            }
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) {
#pragma unroll
                for (int c = 0; c < 7; ++c) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], d);
                sm += __shfl_xor_sync(0xffffffffu, sm, d);
                sv += __shfl_xor_sync(0xffffffffu, sv, d);
                sn += __shfl_xor_sync(0xffffffffu, sn, d);
                edge += __shfl_xor_sync(0xffffffffu, edge, d);
                edge_v += __shfl_xor_sync(0xffffffffu, edge_v, d);
            }
            if (lane == 0) {
                const double prop = q / (sm * dD);             // prop_factor = q / integrate_M()
                const double f = prop * dD * (double)rec.w;
                float4 *dst = reinterpret_cast<float4 *>(A.sz_out);
                dst[sz_plane(A.n_sbg, s, 0, sbg)] = make_float4((float)(acc[0] * f), (float)(acc[1] * f), (float)(acc[2] * f), (float)(acc[3] * f));
                dst[sz_plane(A.n_sbg, s, 1, sbg)] = make_float4((float)(acc[4] * f), (float)(acc[5] * f), (float)(acc[6] * f), 0.f);
                if (A.dop_vn) {
                    float2 vn;
                    if (cfg.doppler_scheme == 2) {            // np.trapz with dx = 1 (doppler_scatter.py:286-296)
                        const double c = 6.283185307179586 * prop;
                        vn = make_float2((float)(c * (sv - 0.5 * edge_v)), (float)(c * (sn - 0.5 * edge)));
                    } else                                      // integrate_V, hydrometeors.py:441-460
                        vn = make_float2((float)(sv * prop * dD), (float)(sn * prop * dD));
                    A.dop_vn[(size_t)sbg * nsp + s] = vn;
                }
            }
        }
    }
}

// Integration scheme 'ml' (interpolation.py:423-436; doppler_scatter.py:127-129, 186-189, 259-262): per-gate quadrature
// weights.  One warp per (ray, sub-beam).  A sub-beam flagged sub_ml keeps its weight only around the first and the last
// gate of its own melting layer (mask_ml = QR > 0 and QS + QG > 0, melting.py:45: bit 7 of `active`, set by k_interp_psd):
// weight x scipy.ndimage.gaussian_filter(spikes at the two gates, sigma = 2) -- radius int(4 sigma + 0.5) = 8, normalised
// weights, mode 'reflect'; gk = the nine one-sided filter weights in float64.  Pairs of gates whose weight is zero are
// not evaluated by the reference (valid_data = QM > 0 and weight > 0): their keys are withdrawn from the bin histogram.
struct MlKernel { double gk[9]; };
__global__ void __launch_bounds__(128)
k_ml_weights(const __grid_constant__ DevConfig cfg, ScanDev scan, MlKernel K, uint8_t *__restrict__ active,
             uint16_t *__restrict__ key, int *__restrict__ hist, float *__restrict__ wraw, long long n_sbg)
{
    const int rq = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (rq >= cfg.n_rays * cfg.n_sub) return;
    const int ng = cfg.n_gates, q = rq % cfg.n_sub;
    const size_t base = (size_t)rq * ng;
    const double w = scan.sub_weight[q];
    const bool ml_only = scan.sub_ml && scan.sub_ml[q];
    int first = 0x7fffffff, last = -1;
    if (ml_only) {
        for (int g = lane; g < ng; g += 32)
            if (active[base + g] & 0x80) { first = min(first, g); last = max(last, g); }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            first = min(first, __shfl_xor_sync(0xffffffffu, first, d));
            last = max(last, __shfl_xor_sync(0xffffffffu, last, d));
        }
    }
    for (int g = lane; g < ng; g += 32) {
        double W = w;
        if (ml_only) {
            double G = 0.0;
            if (last >= 0) {
                for (int m = g - 8; m <= g + 8; ++m) {
                    const int r = m < 0 ? -m - 1 : (m >= ng ? 2 * ng - m - 1 : m);      // mode 'reflect'
                    if (r == first || r == last) G += K.gk[abs(g - m)];
                }
            }
            W = w * G;
        }
        const float Wf = (float)W;
        wraw[base + g] = Wf;
        uint8_t act = active[base + g] & 0x7f;
        if (!(Wf > 0.f) && act) {
            for (int s = 0; s < cfg.n_species; ++s) {
                const size_t ki = (size_t)s * n_sbg + base + g;
                const uint16_t k = key[ki];
                if (k != RADOP_KEY_NONE) { atomicSub(hist + k, 1); key[ki] = RADOP_KEY_NONE; }
            }
            act = 0;
        }
        active[base + g] = act;
    }
}

// total weight of a ray gate over the sub-beams (np.sum(all_weights, axis = 0), doppler_scatter.py:128-129) and the
// normalised weights the scattering sums use (:260-261); thread per ray gate, fixed order
__global__ void k_ml_total(const __grid_constant__ DevConfig cfg, const float *__restrict__ wraw, float *__restrict__ wsz)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)cfg.n_rays * cfg.n_gates) return;
    const int g = (int)(idx % cfg.n_gates), ray = (int)(idx / cfg.n_gates);
    double tot = 0.0;
    for (int q = 0; q < cfg.n_sub; ++q) tot += (double)wraw[((size_t)ray * cfg.n_sub + q) * cfg.n_gates + g];
    for (int q = 0; q < cfg.n_sub; ++q) {
        const size_t sbg = ((size_t)ray * cfg.n_sub + q) * cfg.n_gates + g;
        wsz[sbg] = (float)((double)wraw[sbg] / tot);
    }
}

struct K3Args {
    const int8_t *mask;
    const uint8_t *active;
    const float *wind, *sinth, *vsum, *nsum;
    const float *sz_out;
    const float *ice_tot;
    const float2 *dop_vn;
    const float *vals;
    const double *sub_weight;
    const float *ray_nyquist;
    float *comb;        // [ray][gate][8]: b c p q kd s9 s11 rvel
    float *wsum;        // fused pipeline: comb[7] holds sum(rv w) and this the sum of the weights; else null
    const float4 *ice_part;   // fused pipeline with numeric fall-speed moments: [ray][sub][span][2] partial sums (k_gate_fused), else null
    int n_spans;
    const float *wraw, *wsz;   // integration scheme 'ml': per-gate weights [sbg] (raw, and normalised by their sum over the sub-beams)
    float *out;         // [N_OUT+1][ray][gate]
    float *model_out;   // [NVP][ray][gate]
};

// thread per (ray, gate): sums over sub-beams and species in the reference's order
// (doppler_scatter.py:148-158, 259-268, 313-333, 400-401, 472-477)
__global__ void __launch_bounds__(128)
k_reduce_subbeams(const __grid_constant__ DevConfig cfg, K3Args A)
{
    long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    long long n_rg = (long long)cfg.n_rays * cfg.n_gates;
    if (idx >= n_rg) return;
    const int g = (int)(idx % cfg.n_gates);
    const int ray = (int)(idx / cfg.n_gates);
    const int nsp = cfg.n_species;
    const long long n_sbg_all = (long long)cfg.n_rays * cfg.n_sub * cfg.n_gates;
    float comb[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float rsum = 0.f, wsum = 0.f, msum = 0.f;
    float mv[RADOP_NVP];
    float sumw_all = 0.f;
    if (A.model_out) {
#pragma unroll
        for (int v = 0; v < RADOP_NVP; ++v) mv[v] = 0.f;
        for (int q = 0; q < cfg.n_sub; ++q) sumw_all += (float)A.sub_weight[q];
    }
    for (int q = 0; q < cfg.n_sub; ++q) {
        size_t sbg = ((size_t)ray * cfg.n_sub + q) * cfg.n_gates + g;
        int m = A.mask[sbg];
        msum += (float)m;
        if (m != 0) continue;
        float w = A.wraw ? A.wraw[sbg] : (float)A.sub_weight[q];
        uint8_t act = A.active[sbg];
        for (int s = 0; s < (cfg.sum_species ? (act ? 1 : 0) : nsp); ++s) {
            if (!cfg.sum_species && !(act & (1u << s))) continue;
            const float4 *pl = reinterpret_cast<const float4 *>(A.sz_out);
            float4 r0 = __ldg(pl + sz_plane(n_sbg_all, s, 0, sbg)), r1 = __ldg(pl + sz_plane(n_sbg_all, s, 1, sbg));
            comb[0] += r0.x; comb[1] += r0.y; comb[2] += r0.z; comb[3] += r0.w;
            comb[4] += r1.x; comb[5] += r1.y; comb[6] += r1.z;
        }
        if (cfg.doppler_scheme == 1 || cfg.doppler_scheme == 2) {
            float vs, ns;
            if (cfg.doppler_scheme == 1) {
                vs = A.vsum[sbg]; ns = A.nsum[sbg];
                if (cfg.ice_numeric >= 0 && (act & (1u << cfg.ice_numeric))) {
                    const float *t = A.ice_tot + 3 * ((size_t)ray * cfg.n_sub + q);
                    if (__float_as_int(t[2]) == g) { vs += t[0]; ns += t[1]; }
                }
                if (cfg.with_melting) {       // numeric moments of the melting species (hydrometeors.py:441-460)
                    for (int s = 0; s < nsp; ++s) {
                        if (cfg.sp[s].psd != RADOP_PSD_MELTING || !(act & (1u << s))) continue;
                        const float2 x = __ldg(A.dop_vn + sbg * nsp + s);
                        vs += x.x; ns += x.y;
                    }
                }
            } else {
                vs = 0.f; ns = 0.f;
                for (int s = 0; s < nsp; ++s) {
                    if (!(act & (1u << s))) continue;
                    const float2 x = __ldg(A.dop_vn + sbg * nsp + s);
                    vs += x.x; ns += x.y;
                }
            }
            float vh = vs / ns;                              // 0/0 -> NaN where no hydrometeor (Q7)
            float rv = A.wind[sbg] - vh * A.sinth[sbg];
            if (!isnan(rv)) { rsum += rv * w; wsum += w; }
        }
        if (A.model_out) {
            const float4 *vp = reinterpret_cast<const float4 *>(A.vals + sbg * RADOP_NVP);
#pragma unroll
            for (int v4 = 0; v4 < 4; ++v4) {
                float4 x = __ldg(vp + v4);
                float wn = A.wsz ? A.wsz[sbg] : w / sumw_all;
                if (!isnan(x.x)) mv[4 * v4] += x.x * wn;
                if (!isnan(x.y)) mv[4 * v4 + 1] += x.y * wn;
                if (!isnan(x.z)) mv[4 * v4 + 2] += x.z * wn;
                if (!isnan(x.w)) mv[4 * v4 + 3] += x.w * wn;
            }
        }
    }
    float4 *dst = reinterpret_cast<float4 *>(A.comb + idx * 8);
    dst[0] = make_float4(comb[0], comb[1], comb[2], comb[3]);
    dst[1] = make_float4(comb[4], comb[5], comb[6], rsum / wsum);
    float mk = msum / (float)cfg.n_sub;
    if (mk > -1.f && mk <= 0.f) mk = 0.f;
    A.out[(size_t)RADOP_N_OUT * cfg.out_stride + idx] = mk;
    if (A.model_out) {
#pragma unroll
        for (int v = 0; v < RADOP_NVP; ++v)
            // with the melting scheme the one-moment slots 9..12 carry QmS_v, QmG_v, fwet_mS, fwet_mG (melting.py:74-90),
            // which integrate_radials averages like every other entry of the sub-radials' value dictionary
            if ((cfg.var_mask | (cfg.with_melting ? 0x1E00u : 0u)) & (1u << v)) A.model_out[(size_t)v * cfg.out_stride + idx] = mv[v];
    }
}

__device__ __forceinline__ float warp_scan_add(float v, int lane)
{
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        float t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += t;
    }
    return v;
}
__device__ __forceinline__ float warp_scan_mul(float v, int lane)
{
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        float t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v *= t;
    }
    return v;
}

// one warp per ray: observables (get_pol_from_sz, doppler_scatter.py:491-544), inclusive
// scans along the ray with warp shuffles (doppler_scatter.py:406-415; utilities.py:175-199),
// aliasing (utilities.py:142-155) and the sensitivity cut (doppler_scatter.py:804-862)
__global__ void __launch_bounds__(128)
k_ray_epilogue(const __grid_constant__ DevConfig cfg, K3Args A)
{
    const int ray = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (ray >= cfg.n_rays) return;
    const int ng = cfg.n_gates;
    const size_t n_rg = (size_t)cfg.out_stride;      // rays x gates of the whole scan (this launch may be a chunk)
    const float qnan = __int_as_float(0x7fc00000);
    const float PI = 3.14159265358979323846f;
    const float wl = cfg.wavelength;
    const float zfac = (float)((double)wl * wl * wl * wl / (9.869604401089358 * 9.869604401089358 * 3.141592653589793 * (double)cfg.k_squared));
    const float kfac = (float)(1e-3 * (180.0 / 3.141592653589793) * (double)wl);
    const float dr_km = cfg.radial_res / 1000.f;
    float carry_phi = 0.f, carry_h = 1.f, carry_v = 1.f;
    const float nyq = A.ray_nyquist ? A.ray_nyquist[ray] : 0.f;
    if (A.ice_part) {
        // The total of the numeric fall-speed moments of a sub-beam lands on its first gate with ice (IceParticle.integrate_V
        // sums over ALL gates, hydrometeors.py:1256-1275).  The lanes add the spans' partial sums of their sub-beams in span
        // order; lane 0 then replaces the contributions of those gates to the velocity sums in sub-beam order (a fixed
        // order: several sub-beams may share their first icy gate) before the gates are read below.
        int cur = -1;
        float r = 0.f, ws = 0.f;
        for (int q0 = 0; q0 < cfg.n_sub; q0 += 32) {
            const int q = q0 + lane;
            int first = -1;
            float rv_old = 0.f, rv_new = 0.f, w = 0.f;
            if (q < cfg.n_sub) {
                const float4 *p = A.ice_part + ((size_t)ray * cfg.n_sub + q) * A.n_spans * 2;
                float V = 0.f, N = 0.f;
                float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
                for (int sp = 0; sp < A.n_spans; ++sp) {
                    const float4 x = __ldg(p + 2 * sp);
                    V += x.x; N += x.y;
                    if (first < 0 && __float_as_int(x.z) >= 0) { first = __float_as_int(x.z); a0 = x; a1 = __ldg(p + 2 * sp + 1); }
                }
                const float wind = a0.w, sinth = a1.x, vs = a1.y, ns = a1.z;
                w = a1.w;
                rv_old = wind - (vs / ns) * sinth;
                rv_new = wind - ((vs + V) / (ns + N)) * sinth;
            }
            const int nq = min(32, cfg.n_sub - q0);
            for (int j = 0; j < nq; ++j) {
                const int f = __shfl_sync(0xffffffffu, first, j);
                const float o = __shfl_sync(0xffffffffu, rv_old, j), n = __shfl_sync(0xffffffffu, rv_new, j);
                const float wj = __shfl_sync(0xffffffffu, w, j);
                if (lane != 0 || f < 0) continue;
                if (f != cur) {                    // sub-beams of a ray mostly share the gate: keep it in registers
                    if (cur >= 0) { const size_t rg = (size_t)ray * ng + cur; A.comb[rg * 8 + 7] = r; A.wsum[rg] = ws; }
                    cur = f;
                    const size_t rg = (size_t)ray * ng + cur;
                    r = A.comb[rg * 8 + 7]; ws = A.wsum[rg];
                }
                if (!isnan(o)) { r -= o * wj; ws -= wj; }
                if (!isnan(n)) { r += n * wj; ws += wj; }
            }
        }
        if (lane == 0 && cur >= 0) { const size_t rg = (size_t)ray * ng + cur; A.comb[rg * 8 + 7] = r; A.wsum[rg] = ws; }
        __syncwarp();
    }
    for (int g0 = 0; g0 < ng; g0 += 32) {
        int g = g0 + lane;
        bool ok = g < ng;
        float c[8];
        if (ok) {
            const float4 *src = reinterpret_cast<const float4 *>(A.comb + ((size_t)ray * ng + g) * 8);
            float4 a = src[0], b = src[1];
            c[0] = a.x; c[1] = a.y; c[2] = a.z; c[3] = a.w; c[4] = b.x; c[5] = b.y; c[6] = b.z; c[7] = b.w;
            if (A.wsum) c[7] = c[7] / A.wsum[(size_t)ray * ng + g];
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) c[i] = 0.f;
        }
        // exact zeros -> NaN (doppler_scatter.py:401)
#pragma unroll
        for (int i = 0; i < 7; ++i)
            if (c[i] == 0.f) c[i] = qnan;
        float rxh = 2.f * PI * c[0], rxv = 2.f * PI * c[1];
        float zh = zfac * rxh, zv = zfac * rxv;
        float zdr = rxh / rxv;
        float kdp = kfac * c[4];
        float ah = 4.343e-3f * (2.f * wl * c[6]);
        float av = 4.343e-3f * (2.f * wl * c[5]);
        // products of two float32 sums under- / overflow for very weak or strong echoes: float64
        float rho = (float)sqrt(((double)c[2] * c[2] + (double)c[3] * c[3]) / ((double)c[0] * c[1]));
        float dhv = atan2f(-c[3], -c[2]);
        // PHIDP = nan_cumsum(2 KDP) * dr/1000 + DELTA_HV
        float t = isnan(kdp) ? 0.f : 2.f * kdp;
        if (!ok) t = 0.f;
        float sc = warp_scan_add(t, lane) + carry_phi;
        carry_phi = __shfl_sync(0xffffffffu, sc, 31);
        float phidp = sc * dr_km + dhv;
        if (cfg.with_attenuation) {
            float fh = isnan(ah) ? 1.f : exp10f(-0.1f * ah * dr_km);
            float fv = isnan(av) ? 1.f : exp10f(-0.1f * av * dr_km);
            if (!ok) { fh = 1.f; fv = 1.f; }
            float ph = warp_scan_mul(fh, lane) * carry_h;
            float pv = warp_scan_mul(fv, lane) * carry_v;
            carry_h = __shfl_sync(0xffffffffu, ph, 31);
            carry_v = __shfl_sync(0xffffffffu, pv, 31);
            zdr = (zh * ph) / (zv * pv);
        }
        float rvel = c[7];
        if (A.ray_nyquist && cfg.doppler_scheme != 0) {
            float th = (rvel + nyq) / (2.f * nyq) * PI - PI / 2.f;
            float tf = atanf(tanf(th));
            rvel = (tf + PI / 2.f) * (2.f * nyq) / PI - nyq;
        }
        if (cfg.sensitivity_mode && ok) {
            float thr;
            if (cfg.sensitivity_mode == 1) thr = cfg.sens0;
            else thr = (cfg.sens0 - 20.f * log10f(cfg.sens1 / 1000.f)) + 20.f * log10f((cfg.radial_res * (float)g) / 1000.f);
            if (10.f * log10f(zh) < thr) {
                zh = qnan; rvel = qnan; zv = qnan; phidp = qnan; zdr = qnan; rho = qnan; kdp = qnan;
            }
        }
        if (ok) {
            size_t o = (size_t)ray * ng + g;
            A.out[RADOP_OUT_ZH * n_rg + o] = zh;
            A.out[RADOP_OUT_ZDR * n_rg + o] = zdr;
            A.out[RADOP_OUT_ZV * n_rg + o] = zv;
            A.out[RADOP_OUT_KDP * n_rg + o] = kdp;
            A.out[RADOP_OUT_DELTA_HV * n_rg + o] = dhv;
            A.out[RADOP_OUT_PHIDP * n_rg + o] = phidp;
            A.out[RADOP_OUT_RHOHV * n_rg + o] = rho;
            A.out[RADOP_OUT_ATT_H * n_rg + o] = ah;
            A.out[RADOP_OUT_ATT_V * n_rg + o] = av;
            A.out[RADOP_OUT_RVEL * n_rg + o] = rvel;
        }
    }
}

// ------------------------------------------------------------------------------------------
// K4: Doppler spectrum (scheme 3; get_doppler_spectrum doppler_scatter.py:598-725,
//     get_diameter_from_rad_vel :546-595, get_refl doppler_c.c:11-32, attenuation :372-383)
// ------------------------------------------------------------------------------------------
// Per sub-beam gate quantities of get_diameter_from_rad_vel (doppler_scatter.py:571-585) that every (gate, species)
// pair of the gate shares; k_spectrum used to recompute them per pair on all 32 lanes (a fifth of its instructions).
struct __align__(16) SpecPre {
    double c0;        // W + U_h / tan(theta)
    double inv_sin;   // 1 / sin(theta)
    float inv_rho;    // 1 / sqrt(rho / rho_gate0) in float32, as the reference has it
    float p;          // velocity-bin coordinate of "fall speed 0": (c0 sin(theta) - v_min) / v_res; NaN: scan all bins
    float attw;       // quadrature weight x attenuation of the sub-beam at the gate
    uint32_t rg;      // ray gate (ray * n_gates + gate) of the chunk
};

struct K4Args {
    const int8_t *mask;
    const uint8_t *active;
    const float *sz_out;
    const double *sub_weight;
    const double *d3_uh;
    const float *d3_w, *d3_rho;
    const float2 *d3_st;
    const float *ray_nyquist;
    float *attw;          // [sbg] quadrature weight x attenuation factor of the sub-beam at the gate
    struct SpecPre *pre;  // [sbg] what k_spectrum needs per sub-beam gate, prepared once (k_beam_attenuation)
    double *spec_acc;     // [ray][gate][n_vel] float64 accumulator (atomics)
    // broadening (doppler_scatter.py:357-367): spectra are kept per sub-beam until they are filtered
    const double *d3_std; // [sbg] added spectral width, or null
    double *sub_acc;      // [sbg][n_vel] un-attenuated, un-weighted sub-beam spectra, or null
    uint8_t *brd_flag;    // [ray][sub-beam] np.sum(add_specwidth) > 0
    float *spec_out;      // [ray][gate][n_vel]
    float *out;           // [N_OUT+1][ray][gate]
    // work list of the integration kernel (pairs sorted by LUT bin)
    const int *hist, *bin_start, *chunk_start;
    const long long *totals;
    int *counter;
    const PairRec *rec;
};

// One warp per (ray, sub-beam): cumulated one-way attenuation that the reference applies to
// the sub-beam's spectrum.  Quirk mirrored (doppler_scatter.py:298-311): the specific
// attenuations of the VALID gates of a species are added to the beam array aligned at gate 0
// (utilities.nansum_arr zero-pads the shorter array), i.e. gate i receives the value of the
// i-th valid gate; the cumulative sum runs over that compacted sequence.
extern __shared__ float s_k4a[];
__global__ void __launch_bounds__(128)
k_beam_attenuation(const __grid_constant__ DevConfig cfg, K4Args A)
{
    const int warp_in_block = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rq = blockIdx.x * (blockDim.x >> 5) + warp_in_block;
    if (rq >= cfg.n_rays * cfg.n_sub) return;
    const int ng = cfg.n_gates, nsp = cfg.n_species;
    float *tmp = s_k4a + (size_t)warp_in_block * 2 * ng, *total = tmp + ng;
    const int q = rq % cfg.n_sub;
    const float w = (float)A.sub_weight[q];
    const size_t base = (size_t)rq * ng;
    for (int g = lane; g < ng; g += 32) total[g] = 0.f;
    if (cfg.with_attenuation) {
        const float fac = 4.343e-3f * 2.f * cfg.wavelength * (cfg.radial_res / 1000.f);
        for (int s = 0; s < nsp; ++s) {
            int cnt = 0;
            float run = 0.f;
            for (int g0 = 0; g0 < ng; g0 += 32) {
                const int g = g0 + lane;
                const bool valid = g < ng && (A.active[base + g] & (1u << s));
                float x = valid ? fac * (A.sz_out[sz_plane((long long)cfg.n_rays * cfg.n_sub * ng, s, 1, base + g) * 4 + 2] / w) : 0.f;
                const unsigned bal = __ballot_sync(0xffffffffu, valid);
                const int pos = cnt + __popc(bal & ((1u << lane) - 1u));
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const float t = __shfl_up_sync(0xffffffffu, x, d);
                    if (lane >= d) x += t;
                }
                x += run;
                if (valid) tmp[pos] = x;
                run = __shfl_sync(0xffffffffu, x, 31);
                cnt += __popc(bal);
            }
            __syncwarp();
            if (cnt > 0)
                for (int g = lane; g < ng; g += 32) total[g] += tmp[min(g, cnt - 1)];
            __syncwarp();
        }
    }
    const float nanf_ = __int_as_float(0x7fc00000);
    const float rho0 = A.d3_rho[base];
    const uint32_t rg0 = (uint32_t)(rq / cfg.n_sub) * (uint32_t)ng;
    for (int g = lane; g < ng; g += 32) {
        const float aw = w * exp10f(-0.1f * total[g]);
        A.attw[base + g] = aw;
        // np.sin / np.tan of the float32 angle are float32 (doppler_scatter.py:571-575)
        const float2 st = A.d3_st[base + g];
        const double sin_t = (double)st.x, inv_tan = 1.0 / (double)st.y;
        const float rho_corr = sqrtf(__fdiv_rn(A.d3_rho[base + g], rho0));
        SpecPre P;
        P.inv_sin = 1.0 / sin_t;
        P.inv_rho = __fdiv_rn(1.0f, rho_corr);
        P.c0 = (double)A.d3_w[base + g] + A.d3_uh[base + g] * inv_tan;
        P.p = (sin_t > 0.0 && P.inv_rho > 0.f && isfinite(P.c0)) ? (float)((P.c0 * sin_t - cfg.vel_min) * cfg.inv_vel_res) : nanf_;
        P.attw = aw;
        P.rg = rg0 + (uint32_t)g;
        A.pre[base + g] = P;
    }
    if (A.d3_std) {
        // the sub-beam is broadened if np.sum(add_specwidth) > 0: a NaN gate (EDR outside the model
        // domain) switches the broadening of the whole sub-beam off (doppler_scatter.py:366)
        double sum = 0.0;
        for (int g = lane; g < ng; g += 32) sum += A.d3_std[base + g];
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
        if (lane == 0) A.brd_flag[rq] = (sum > 0.0) ? 1 : 0;
    }
}

// Spectrum of every active (sub-beam gate, species) pair, one warp per pair.  The work list
// is the bin-sorted pair list of the integration kernel, so the table column of a bin is staged
// once in shared memory and reused by all its pairs.
//   1. bin edges: fall speed needed at every velocity of VARRAY -> diameter (float64 pow, cast
//      to float32, clipped with the reference's whole-matrix thresholds), float32 index
//      arithmetic of get_refl for the first / last diameter sample of every velocity bin;
//   2. prefix sums (float64) over the nd samples N(D_k) rcs(D_k) dD: lane l owns the samples
//      [l nd/32, (l+1) nd/32), a warp scan joins the lanes;
//   3. reflectivity of a velocity bin = difference of two prefix values, weighted with the
//      quadrature weight and the attenuation of the sub-beam, accumulated over sub-beams and
//      species with float64 atomics.
extern __shared__ __align__(16) unsigned char s_k4[];

// running sums of one lane's samples [k0, k0 + cpl): four independent chains of cpl / 4
// consecutive samples each (latency of the float64 adds), local prefix stored in float32
#define SPEC_SUB 4
// SPEC_WARPS warps per CTA share the three staged table columns (12 KB); each warp owns 5.6 KB (prefix sums, chain bases,
// edge diameters).  Shared memory limits the warps per SM: 16-warp CTAs fit 32 warps where 4-warp CTAs fitted 24.
// ND1024: every table has 1024 diameters (compute_lut_sz.py:69, the only size the reference generates): the index
// arithmetic of the chains (k / cpl, k % cpl, k / sub) becomes shifts and masks instead of integer divisions
template <bool ND1024>
__global__ void __launch_bounds__(32 * SPEC_WARPS)
k_spectrum(const __grid_constant__ DevConfig cfg, K4Args A)
{
    __shared__ int s_unit;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int total_units = (int)A.totals[2];
    const int nbins = cfg.total_bins, nsp = cfg.n_species, ng = cfg.n_gates, nvel = cfg.n_vel;
    int nd_max = 0;
    for (int s = 0; s < nsp; ++s) nd_max = max(nd_max, cfg.sp[s].nd);
    // shared memory carve-up (sample k of lane l = k / cpl sits at (k % cpl) * 32 + l)
    float *s_col0 = reinterpret_cast<float *>(s_k4);            // b D^mu dD
    float *s_auxp = s_col0 + nd_max;                            // D^nu or D^0.6357
    float *s_colm = s_auxp + nd_max;                            // a D^b dD (1-mom ice)
    // s_C: the prefix of sample k of lane l = k / cpl sits at (k % cpl) * 33 + l -- rows of 32 lanes padded to 33 words, so
    // that both the stores of phase 2 (one row, all lanes) and the gathers of phase 3 (neighbouring k of one lane) spread
    // over the banks (the unpadded layout put all samples of a lane into ONE bank: 7e8 conflict cycles per volume)
    const size_t c_words = ((size_t)(nd_max >> 5) * 33 + 1) & ~(size_t)1;
    const size_t per_warp = c_words * 4 + 128 * 8 + (size_t)((nvel + 1) & ~1) * 4;
    unsigned char *wbase = reinterpret_cast<unsigned char *>(s_colm + nd_max) + (size_t)wid * per_warp;
    double *s_base = reinterpret_cast<double *>(wbase);         // [32 lanes][SPEC_SUB] exclusive bases
    float *s_C = reinterpret_cast<float *>(wbase + 128 * 8);    // local inclusive prefix
    float *s_D = s_C + c_words;                                 // diameters at the velocity edges
    const float zfac3 = (float)((double)cfg.wavelength * cfg.wavelength * cfg.wavelength * cfg.wavelength /
                                (306.0196847852814 * (double)cfg.k_squared * (double)cfg.k_squared));   // pi^5
    int staged_bin = -1;

    for (;;) {
        __syncthreads();
        if (tid == 0) s_unit = atomicAdd(A.counter, 1);
        __syncthreads();
        const int unit = total_units - 1 - s_unit;
        if (unit < 0) break;
        int b;
        {
            int l = 0, r = nbins;
            while (r - l > 1) {
                int m = (l + r) >> 1;
                if (A.chunk_start[m] <= unit) l = m; else r = m;
            }
            b = l;
        }
        int s = 0;
        while (s + 1 < nsp && b >= cfg.sp[s + 1].bin_base) ++s;
        const SpeciesDev &S = cfg.sp[s];
        const int nd = ND1024 ? 1024 : S.nd, cpl = ND1024 ? 32 : nd >> 5;   // samples per lane (nd is a multiple of 32)
        const int sub = ND1024 ? 32 / SPEC_SUB : cpl / SPEC_SUB;            // samples per chain (nd is a multiple of 128 here)
        const int cnt = A.hist[b];
        const int tiles = (cnt + RADOP_TILE - 1) / RADOP_TILE;
        const int j0 = (unit - A.chunk_start[b]) * RADOP_CHUNK, j1 = min(tiles, j0 + RADOP_CHUNK);
        if (b != staged_bin) {
            const float *tab = S.table + (size_t)(b - S.bin_base) * nd * S.ncol;
            for (int k = tid; k < nd; k += blockDim.x) {
                const int addr = (k % cpl) * 32 + (k / cpl);
                s_col0[addr] = __ldg(tab + (size_t)k * S.ncol);
                s_auxp[addr] = __ldg(S.aux + (S.mode == MODE_ICE1 ? 2 * nd : nd) + k);
                s_colm[addr] = (S.mode == MODE_ICE1) ? __ldg(tab + (size_t)k * S.ncol + 9) : 0.f;
            }
            staged_bin = b;
            __syncthreads();
        }
        const float spec_scale = (float)((double)S.spec_step / (double)S.dD_ref);
        // diameters at the starts of this lane's four prefix chains: the same for every pair of the unit
        float d_chain[SPEC_SUB];
#pragma unroll
        for (int u = 0; u < SPEC_SUB; ++u) d_chain[u] = __ldg(S.aux + lane * cpl + u * sub);
        for (int jt = j0; jt < j1; ++jt) {
            const int first = A.bin_start[b] + jt * RADOP_TILE;
            const int n_here = min(RADOP_TILE, cnt - jt * RADOP_TILE);
            for (int i = wid; i < n_here; i += SPEC_WARPS) {
                const PairRec rec = A.rec[first + i];
                const uint32_t sbg = rec.sbg;
                const unsigned act = A.active[sbg];
                const SpecPre P = A.pre[sbg];
                const size_t rg = P.rg;
                // ---- 1. diameters at the velocity-bin edges -----------------------------------
                int j_a, j_b;
                {
                    const double inv_sin = P.inv_sin, inv_rho = (double)P.inv_rho, c0 = P.c0;
                    // wh_j = inv_rho (c0 - v_j / sin) falls with j: only the velocities between
                    // "faster than the largest particle" (D = d_max, empty bins) and "wh < 0"
                    // (no bin) can hold power.  Outside the normal case the whole array is done.
                    // Bounds in float32 with a bin of slack on either side (they only have to enclose the bins with power).
                    j_a = 0; j_b = nvel - 1;
                    if (!isnan(P.p)) {
                        // bins per unit of fall speed: rho_corr sin(theta) / v_res = 1 / (inv_rho inv_sin v_res)
                        const float qv = __fdividef(1.000002f * (float)cfg.inv_vel_res, P.inv_rho * (float)inv_sin);
                        const float ja = floorf(P.p - (float)S.wh_hi * qv - 0.02f) - 1.f, jb = floorf(P.p + 0.02f) + 2.f;
                        j_a = (int)fminf(fmaxf(ja, 0.f), (float)(nvel - 1));
                        j_b = (int)fminf(fmaxf(jb, 0.f), (float)(nvel - 1));
                    }
                    for (int j = j_a + lane; j <= j_b; j += 32) {
                        const double v = cfg.vel_min + (double)j * cfg.vel_res;
                        const double wh = inv_rho * (c0 - v * inv_sin);
                        float D = __int_as_float(0x7fc00000);
                        if (wh >= 0.0) {
                            // fall speeds beyond those of the largest / smallest tabulated diameter
                            // clip to the axis ends whatever the power law gives (margin 1e-9 >> rounding)
                            if (wh >= S.wh_hi) D = S.d_max;
                            else if (wh <= S.wh_lo) D = S.d_min;
                            else D = (float)exp2(S.inv_beta * log2(wh * S.inv_alpha));   // pow(): same to 1e-15
                            for (int t = s; t < nsp; ++t) {
                                if (!(act & (1u << t))) continue;
                                if (D >= cfg.sp[t].d_max) D = cfg.sp[t].d_max;
                                if (D <= cfg.sp[t].d_min) D = cfg.sp[t].d_min;
                            }
                        }
                        s_D[j] = D;
                    }
                }
                // ---- 2. prefix sums of N(D_k) rcs(D_k) dD over this lane's samples ------------
                // local running sums in float32: they are stored as float32 anyway, and a chain is 8 samples long
                float loc[SPEC_SUB];
#pragma unroll
                for (int u = 0; u < SPEC_SUB; ++u) loc[u] = 0.f;
                float n0 = rec.p1;
                if (S.mode == MODE_ICE1) {
                    const float lam = rec.p0;
                    const float a1 = -20.78f * LOG2E * lam, a2 = -3.290f * LOG2E * lam;
                    const float c2 = 17.46f * exp2f(0.6357f * log2f(lam));
                    const float r1 = ex2f_approx(a1 * S.dD_geo), r2 = ex2f_approx(a2 * S.dD_geo);
                    float E1[SPEC_SUB], E2[SPEC_SUB], mass = 0.f;
#pragma unroll
                    for (int u = 0; u < SPEC_SUB; ++u) {
                        const float d0 = d_chain[u];
                        E1[u] = 490.6f * ex2f_approx(a1 * d0);
                        E2[u] = c2 * ex2f_approx(a2 * d0);
                    }
                    if (ND1024) {
                        float2 mass2 = make_float2(0.f, 0.f);
                        // as for the geometric shapes below: the table values of two chains and four samples (24 loads) first
#pragma unroll
                        for (int h2 = 0; h2 < SPEC_SUB / 2; ++h2)
#pragma unroll
                            for (int q4 = 0; q4 < 2; ++q4) {
                                float c0v[2][4], axv[2][4], cmv[2][4];
#pragma unroll
                                for (int v = 0; v < 2; ++v)
#pragma unroll
                                    for (int i4 = 0; i4 < 4; ++i4) {
                                        const int ad = ((h2 * 2 + v) * 8 + q4 * 4 + i4) * 32 + lane;
                                        c0v[v][i4] = s_col0[ad]; axv[v][i4] = s_auxp[ad]; cmv[v][i4] = s_colm[ad];
                                    }
                                // the two chains advance as the halves of packed FFMA2 / FMUL2 (the kernel is issue bound)
                                const int u0 = h2 * 2, u1 = u0 + 1;
                                float2 e1 = make_float2(E1[u0], E1[u1]), e2 = make_float2(E2[u0], E2[u1]);
                                float2 lc = make_float2(loc[u0], loc[u1]);
#pragma unroll
                                for (int i4 = 0; i4 < 4; ++i4) {
                                    const int ii = q4 * 4 + i4;
                                    const float2 phi = ffma2(make_float2(axv[0][i4], axv[1][i4]), e2, e1);
                                    mass2 = ffma2(make_float2(cmv[0][i4], cmv[1][i4]), phi, mass2);
                                    lc = ffma2(make_float2(c0v[0][i4], c0v[1][i4]), phi, lc);
                                    s_C[(u0 * 8 + ii) * 33 + lane] = lc.x;
                                    s_C[(u1 * 8 + ii) * 33 + lane] = lc.y;
                                    e1 = fmul2(e1, bcast2(r1)); e2 = fmul2(e2, bcast2(r2));
                                }
                                E1[u0] = e1.x; E1[u1] = e1.y; E2[u0] = e2.x; E2[u1] = e2.y;
                                loc[u0] = lc.x; loc[u1] = lc.y;
                            }
                        mass = mass2.x + mass2.y;
                    } else
                    for (int ii = 0; ii < sub; ++ii) {
#pragma unroll
                        for (int u = 0; u < SPEC_SUB; ++u) {
                            const int ad = (u * sub + ii) * 32 + lane;
                            const float phi = fmaf(s_auxp[ad], E2[u], E1[u]);
                            mass = fmaf(s_colm[ad], phi, mass);
                            loc[u] = fmaf(s_col0[ad], phi, loc[u]);
                            s_C[ad + (u * sub + ii)] = loc[u];
                            E1[u] *= r1; E2[u] *= r2;
                        }
                    }
#pragma unroll
                    for (int d = 16; d > 0; d >>= 1) mass += __shfl_xor_sync(0xffffffffu, mass, d);
                    n0 = rec.p1 / mass;          // hydrometeors.py:1330-1339
                } else if (S.mode == MODE_GEO8) {
                    const float a = -LOG2E * rec.p0;
                    const float r = ex2f_approx(a * S.dD_geo);
                    float E[SPEC_SUB];
#pragma unroll
                    for (int u = 0; u < SPEC_SUB; ++u) E[u] = ex2f_approx(a * d_chain[u]);
                    if (ND1024) {
                        // two chains at a time, their sixteen table values loaded first: the loads are then in flight
                        // together instead of each multiply-add waiting for its own (this line held 17 % of the stall samples)
#pragma unroll
                        for (int h2 = 0; h2 < SPEC_SUB / 2; ++h2) {
                            float cv[2][8];
#pragma unroll
                            for (int v = 0; v < 2; ++v)
#pragma unroll
                                for (int ii = 0; ii < 8; ++ii) cv[v][ii] = s_col0[((h2 * 2 + v) * 8 + ii) * 32 + lane];
                            const int u0 = h2 * 2, u1 = u0 + 1;
                            float2 ev = make_float2(E[u0], E[u1]), lc = make_float2(loc[u0], loc[u1]);
#pragma unroll
                            for (int ii = 0; ii < 8; ++ii) {
                                lc = ffma2(make_float2(cv[0][ii], cv[1][ii]), ev, lc);
                                s_C[(u0 * 8 + ii) * 33 + lane] = lc.x;
                                s_C[(u1 * 8 + ii) * 33 + lane] = lc.y;
                                ev = fmul2(ev, bcast2(r));
                            }
                            loc[u0] = lc.x; loc[u1] = lc.y;
                        }
                    } else
                    for (int ii = 0; ii < sub; ++ii) {
#pragma unroll
                        for (int u = 0; u < SPEC_SUB; ++u) {
                            const int ad = (u * sub + ii) * 32 + lane;
                            loc[u] = fmaf(s_col0[ad], E[u], loc[u]);
                            s_C[ad + (u * sub + ii)] = loc[u];
                            E[u] *= r;
                        }
                    }
                } else {
                    const float a = -LOG2E * rec.p0;
                    for (int ii = 0; ii < sub; ++ii) {
#pragma unroll
                        for (int u = 0; u < SPEC_SUB; ++u) {
                            const int ad = (u * sub + ii) * 32 + lane;
                            loc[u] = fmaf(s_col0[ad], ex2f_approx(a * s_auxp[ad]), loc[u]);
                            s_C[ad + (u * sub + ii)] = loc[u];
                        }
                    }
                }
                double pair_total = 0.0;
                {   // exclusive scan of the lane totals, then the bases of the four chains
                    const double tot = ((double)loc[0] + (double)loc[1]) + ((double)loc[2] + (double)loc[3]);
                    double x = tot;
#pragma unroll
                    for (int d = 1; d < 32; d <<= 1) {
                        const double t = __shfl_up_sync(0xffffffffu, x, d);
                        if (lane >= d) x += t;
                    }
                    const double prev = __shfl_up_sync(0xffffffffu, x, 1);
                    pair_total = __shfl_sync(0xffffffffu, x, 31);
                    double bs = lane ? prev : 0.0;
#pragma unroll
                    // chain c = lane * SPEC_SUB + u lives at u * 32 + lane (lane-major rows: the 8-byte stores of a warp
                    // and the gathers below spread over the banks; [lane][u] cost 6e8 conflict cycles per volume)
                    for (int u = 0; u < SPEC_SUB; ++u) { s_base[u * 32 + lane] = bs; bs += (double)loc[u]; }
                }
                __syncwarp();
                // the next pair of this warp: its per sub-beam gate record and activity mask are random accesses (an L2 round
                // trip each at the head of the pair's dependency chain); asked for now, they are in L1 when it starts
                if (i + SPEC_WARPS < n_here) {
                    const uint32_t sbg_n = A.rec[first + i + SPEC_WARPS].sbg;
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(A.pre + sbg_n));
                    asm volatile("prefetch.global.L1 [%0];" ::"l"(A.active + sbg_n));
                }
                // ---- 3. velocity bins -----------------------------------------------------------
                const double scale = 6.283185307179586 * (double)n0 * (double)zfac3 * (double)spec_scale *
                                     (A.sub_acc ? 1.0 : (double)P.attw);
                double *dst = A.sub_acc ? A.sub_acc + (size_t)sbg * nvel : A.spec_acc + rg * nvel;
                // Second pass only for a pair whose prefix differences ALL came out as zero although it has samples inside
                // the velocity window: its power sits in the far tail of the PSD, below the float32 resolution of the running
                // sums, but it may be the only power of the gate -- and then decides whether RVEL is a number (the reference
                // sums the samples of a bin directly, doppler_c.c:23-30, and keeps 1e-22 of a bin as it keeps 1e+3).
                bool any_power = false, any_samples = false;
                const bool direct = A.sub_acc != nullptr;
                for (int j = j_a + lane; j < j_b; j += 32) {
                    const float Dj = s_D[j];
                    if (isnan(Dj)) continue;
                    // next valid velocity (rows of the filtered list are paired, doppler_scatter.py:591-593)
                    int jn = j + 1;
                    while (jn <= j_b && isnan(s_D[jn])) ++jn;
                    if (jn > j_b) continue;
                    const float Dn = s_D[jn];
                    const float Da = fminf(Dj, Dn), Db = fmaxf(Dj, Dn);
                    const int ia = (int)__fdiv_rn(__fsub_rn(Da, S.d_min), S.spec_step);
                    const int ib = min((int)__fdiv_rn(__fsub_rn(Db, S.d_min), S.spec_step), nd);
                    if (ib <= ia || ia < 0) continue;
                    any_samples = true;
                    const int kb = ib - 1, ka = ia - 1;
                    // chain index = k / sub; position in shared memory (k % cpl) * 32 + k / cpl
                    const int cb = kb / sub, ca = (ka >= 0) ? ka / sub : -1;
                    const double vb = (double)s_C[(kb % cpl) * 33 + kb / cpl];
                    const double va = (ka >= 0) ? (double)s_C[(ka % cpl) * 33 + ka / cpl] : 0.0;
                    double diff;
                    if (ca == cb) diff = vb - va;               // same chain: one running sum
                    else {
                        const double hi = s_base[(cb % SPEC_SUB) * 32 + cb / SPEC_SUB] + vb;
                        const double lo = (ka >= 0) ? s_base[(ca % SPEC_SUB) * 32 + ca / SPEC_SUB] + va : 0.0;
                        diff = hi - lo;
                        if (fabs(diff) < 2e-15 * fabs(hi)) diff = 0.0;      // rounding of the bases
                    }
                    if (direct && !(diff > 1e-6 * (s_base[(cb % SPEC_SUB) * 32 + cb / SPEC_SUB] + vb))) {
                        // Broadening turns a sub-beam spectrum without power into NaN (0 / 0,
                        // doppler_scatter.py:799), so "tiny" and "zero" are not the same thing here:
                        // where the prefix difference has lost the value (far PSD tail), the samples
                        // of the bin are summed directly, as get_refl does (doppler_c.c:23-30)
                        float acc = 0.f;
                        const float lam = rec.p0;
                        const float ea = -LOG2E * lam;
                        const float a1 = -20.78f * LOG2E * lam, a2 = -3.290f * LOG2E * lam;
                        const float c2 = (S.mode == MODE_ICE1) ? 17.46f * exp2f(0.6357f * log2f(lam)) : 0.f;
                        for (int k = ia; k < ib; ++k) {
                            const int ad = (k % cpl) * 32 + k / cpl;
                            float shape;
                            if (S.mode == MODE_ICE1) {
                                const float Dk = __ldg(S.aux + k);
                                shape = fmaf(s_auxp[ad], c2 * exp2f(a2 * Dk), 490.6f * exp2f(a1 * Dk));
                            } else if (S.mode == MODE_GEO8) shape = exp2f(ea * __ldg(S.aux + k));
                            else shape = exp2f(ea * s_auxp[ad]);      // exp2f keeps subnormals (ex2.approx.ftz does not)
                            acc = fmaf(s_col0[ad], shape, acc);
                        }
                        diff = (double)acc;
                    }
                    const double val = diff * scale;
                    if (val != 0.0) { atomicAdd(dst + j, val); any_power = true; }
                }
                // (a pair whose 1024 samples sum to zero in float32 has nothing a float32 spectrum could hold)
                if (!direct && pair_total > 0.0 && __ballot_sync(0xffffffffu, any_samples) && !__any_sync(0xffffffffu, any_power)) {
                    // the second pass, by the whole warp: bin by bin, the lanes stride over the samples of the bin
                    // (a bin at the flat end of the fall-speed law can hold hundreds of samples)
                    const float lam = rec.p0;
                    const float ea = -LOG2E * lam;
                    const float a1 = -20.78f * LOG2E * lam, a2 = -3.290f * LOG2E * lam;
                    const float c2 = (S.mode == MODE_ICE1) ? 17.46f * exp2f(0.6357f * log2f(lam)) : 0.f;
                    for (int j0 = j_a; j0 < j_b; j0 += 32) {
                        const int j = j0 + lane;
                        int ia = 0, ib = 0;
                        if (j < j_b && !isnan(s_D[j])) {
                            int jn = j + 1;
                            while (jn <= j_b && isnan(s_D[jn])) ++jn;
                            if (jn <= j_b) {
                                const float Dj = s_D[j], Dn = s_D[jn];
                                const float Da = fminf(Dj, Dn), Db = fmaxf(Dj, Dn);
                                ia = (int)__fdiv_rn(__fsub_rn(Da, S.d_min), S.spec_step);
                                ib = min((int)__fdiv_rn(__fsub_rn(Db, S.d_min), S.spec_step), nd);
                                if (ia < 0) ib = ia;
                            }
                        }
                        float mine = 0.f;
                        const int nb = min(32, j_b - j0);
                        for (int t = 0; t < nb; ++t) {
                            const int ka = __shfl_sync(0xffffffffu, ia, t), kb = __shfl_sync(0xffffffffu, ib, t);
                            if (kb <= ka) continue;
                            float acc = 0.f;
                            // every lane sums ITS samples of the bin (the cpl consecutive ones it owns in shared memory:
                            // column = lane, no bank conflicts)
                            const int k_hi = min(kb, (lane + 1) * cpl);
                            for (int k = max(ka, lane * cpl); k < k_hi; ++k) {
                                const int ad = (k % cpl) * 32 + k / cpl;
                                float shape;
                                if (S.mode == MODE_ICE1) {
                                    const float Dk = __ldg(S.aux + k);
                                    shape = fmaf(s_auxp[ad], c2 * ex2f_approx(a2 * Dk), 490.6f * ex2f_approx(a1 * Dk));
                                } else if (S.mode == MODE_GEO8) shape = ex2f_approx(ea * __ldg(S.aux + k));
                                else shape = ex2f_approx(ea * s_auxp[ad]);      // flushes below 2^-126: a float32 spectrum cannot hold what that loses
                                acc = fmaf(s_col0[ad], shape, acc);
                            }
#pragma unroll
                            for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
                            if (lane == t) mine = acc;
                        }
                        const double val = (double)mine * scale;
                        if (j < j_b && val != 0.0) atomicAdd(dst + j, val);
                    }
                }
                __syncwarp();
            }
        }
    }
}

// Broadening of the sub-beam spectra (doppler_scatter.py:357-367, 779-801), one warp per sub-beam
// gate: the float32 spectrum of the sub-beam is filtered along the velocity axis with
// scipy.ndimage.gaussian_filter's defaults (sigma = std / v_res; skipped when sigma <= 1e-15 or NaN;
// radius int(4 sigma + 0.5); weights exp(-k^2 / (2 sigma^2)) normalised to 1; mode 'reflect';
// float64 accumulation, float32 result), rescaled to its original power (rows without power become
// 0 / 0 = NaN, as in the reference), then attenuated, weighted and added to the ray spectrum.
#define BRD_WMAX 1024
extern __shared__ __align__(16) unsigned char s_k4b[];
__global__ void __launch_bounds__(128)
k_spectrum_broaden(const __grid_constant__ DevConfig cfg, K4Args A, long long n_sbg)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const long long sbg = (long long)blockIdx.x * (blockDim.x >> 5) + wid;
    if (sbg >= n_sbg) return;
    const int nvel = cfg.n_vel, ng = cfg.n_gates, nvp = (nvel + 1) & ~1;
    unsigned char *wb = s_k4b + (size_t)wid * ((size_t)BRD_WMAX * 8 + (size_t)nvp * 8);
    double *wt = reinterpret_cast<double *>(wb);
    float *row = reinterpret_cast<float *>(wb + (size_t)BRD_WMAX * 8), *flt = row + nvp;
    const long long rq = sbg / ng;
    const int g = (int)(sbg - rq * ng);
    const size_t rg = (size_t)(rq / cfg.n_sub) * ng + g;
    const double *src = A.sub_acc + (size_t)sbg * nvel;
    double *dst = A.spec_acc + rg * nvel;
    float osum = 0.f;
    bool any = false;
    for (int j = lane; j < nvel; j += 32) {
        const float x = (float)src[j];      // get_doppler_spectrum returns float32
        row[j] = x;
        osum += x;
        any = any || (x != 0.f);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) osum += __shfl_xor_sync(0xffffffffu, osum, d);
    any = __any_sync(0xffffffffu, any);
    const bool brd = A.brd_flag[rq] != 0;
    const double attw = (double)A.attw[sbg];
    if (!brd) {
        if (any)
            for (int j = lane; j < nvel; j += 32)
                if (row[j] != 0.f) atomicAdd(dst + j, (double)row[j] * attw);
        return;
    }
    __syncwarp();
    const double sigma = A.d3_std[sbg] / cfg.vel_res;
    const float *res = row;
    if (any && sigma > 1e-15) {
        const double t = 4.0 * sigma + 0.5;
        const int lw = t < 1e9 ? (int)t : 1000000000;
        const double c = -0.5 / (sigma * sigma);
        const bool tab = lw < BRD_WMAX;
        double wsum = 0.0;
        if (tab) {
            for (int k = lane; k <= lw; k += 32) {
                const double w = exp(c * (double)k * (double)k);
                wt[k] = w;
                wsum += k ? 2.0 * w : w;
            }
        } else {
            // weights beyond 39 sigma underflow to 0: the sum is complete long before lw
            for (int k = lane; k <= lw && (double)k <= 40.0 * sigma + 64.0; k += 32) {
                const double w = exp(c * (double)k * (double)k);
                wsum += k ? 2.0 * w : w;
            }
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, d);
        __syncwarp();

        for (int j = lane; j < nvel; j += 32) {
            double acc = (double)row[j];        // weight 1 before the normalisation
            int ia = j, ib = j;                 // reflected indices of j - k and j + k, walked incrementally
            int da = -1, db = 1;
            for (int k = 1; k <= lw; ++k) {
                ia += da; if (ia < 0) { ia = 0; da = 1; } else if (ia >= nvel) { ia = nvel - 1; da = -1; }
                ib += db; if (ib >= nvel) { ib = nvel - 1; db = -1; } else if (ib < 0) { ib = 0; db = 1; }
                double w;
                if (tab) w = wt[k];
                else {
                    if ((double)k > 40.0 * sigma + 64.0) break;
                    w = exp(c * (double)k * (double)k);
                }
                acc += w * ((double)row[ia] + (double)row[ib]);
            }
            flt[j] = (float)(acc / wsum);
        }
        res = flt;
        __syncwarp();
    }
    float csum = 0.f;
    for (int j = lane; j < nvel; j += 32) csum += res[j];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) csum += __shfl_xor_sync(0xffffffffu, csum, d);
    for (int j = lane; j < nvel; j += 32) {
        const float o = __fmul_rn(__fdiv_rn(res[j], csum), osum);   // spectrum / convolved_power * original_power
        if (o != 0.f) atomicAdd(dst + j, (double)o * attw);          // NaN != 0: NaN rows propagate
    }
}

// One warp per (ray, gate): float32 spectrum, RVEL = sum(v S) / sum(S) (doppler_scatter.py:422-429),
// aliasing, sensitivity cut of the spectrum bins (doppler_scatter.py:843-849).
__global__ void __launch_bounds__(128)
k_spectrum_finish(const __grid_constant__ DevConfig cfg, K4Args A)
{
    const long long wg = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const long long n_rg = (long long)cfg.n_rays * cfg.n_gates;
    if (wg >= n_rg) return;
    const int g = (int)(wg % cfg.n_gates), ray = (int)(wg / cfg.n_gates);
    const int nvel = cfg.n_vel;
    const double *src = A.spec_acc + (size_t)wg * nvel;
    float *dst = A.spec_out + (size_t)wg * nvel;
    const float qnan = __int_as_float(0x7fc00000);
    float thr = 0.f;
    if (cfg.sensitivity_mode == 1) thr = cfg.sens0;
    else if (cfg.sensitivity_mode == 2)
        thr = (cfg.sens0 - 20.f * log10f(cfg.sens1 / 1000.f)) + 20.f * log10f((cfg.radial_res * (float)g) / 1000.f);
    double sv = 0.0, ss = 0.0;
    for (int j = lane; j < nvel; j += 32) {
        const double x = src[j];
        sv += (cfg.vel_min + (double)j * cfg.vel_res) * x;
        ss += x;
        float o = (float)x;
        if (cfg.sensitivity_mode && 10.f * log10f(o) < thr) o = qnan;
        dst[j] = o;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        sv += __shfl_xor_sync(0xffffffffu, sv, d);
        ss += __shfl_xor_sync(0xffffffffu, ss, d);
    }
    if (lane == 0) {
        float rvel = (float)(sv / ss);          // 0/0 -> NaN where nothing falls
        if (A.ray_nyquist) {
            const float PI = 3.14159265358979323846f, nyq = A.ray_nyquist[ray];
            const float th = (rvel + nyq) / (2.f * nyq) * PI - PI / 2.f;
            rvel = (atanf(tanf(th)) + PI / 2.f) * (2.f * nyq) / PI - nyq;
        }
        if (cfg.sensitivity_mode && isnan(A.out[(size_t)RADOP_OUT_ZH * cfg.out_stride + wg])) rvel = qnan;
        A.out[(size_t)RADOP_OUT_RVEL * cfg.out_stride + wg] = rvel;
    }
}

// ------------------------------------------------------------------------------------------
// FP32 FMA throughput probe (roofline denominator): 16 independent FFMA chains per thread
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_fma_probe(float *out, int iters, float seed)
{
    float a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = seed + (float)(threadIdx.x + i);
    float m = 0.999f + seed * 1e-9f, c = 1e-3f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], m, c);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += a[i];
    if (s == 123.456f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
