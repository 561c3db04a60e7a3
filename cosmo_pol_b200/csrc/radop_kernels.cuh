// radop_kernels.cuh -- device-side data structures shared by the kernels of libradop.so
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/radop.h"

#define RADOP_NVP 16            // floats per (column, level) in the interleaved field layout
#define RADOP_TILE_THREADS 128  // threads of one PSD-integration CTA
#define RADOP_TILE_G 8          // (sub-beam gate, species) pairs per thread and tile
#define RADOP_TILE (RADOP_TILE_THREADS * RADOP_TILE_G)
#define RADOP_HBLOCK 32         // diameters per Horner block (re-anchoring period)
#ifndef SPEC_WARPS
#define SPEC_WARPS 16           // warps per CTA of the Doppler-spectrum kernel
#endif
#define RADOP_CHUNK 2           // tiles per dynamically scheduled work unit of the integration kernel
#define RADOP_NC 7              // combined columns that carry data (the 8th of a row is padding)
#define RADOP_KEY_NONE 0xFFFFu

// integration modes of the PSD x LUT contraction
enum { MODE_GEO8 = 0,       // nu == 1 on a uniform D grid: blocked Horner, 8 columns
       MODE_GEN8 = 1,       // general nu: one ex2 per diameter, 8 columns
       MODE_ICE1 = 2,       // 1-mom ice (Field 2005 double exponential), 12 columns
       MODE_GEN12 = 3,      // general nu with numeric fall-speed integrals, 12 columns
       MODE_MELT = 4 };     // melting species: raw combined columns, PSD evaluated per gate (k_integrate_melting)

struct SpeciesDev {
    int species, psd, mode, ncol;
    int numeric_vn;
    int ne, nt, nd;
    float e_min, e_step, t_min, t_step;
    float e_rstep, t_rstep;     // RN(1 / step), formed in float64 on the host
    int bin_base;               // first global bin id of this species
    int cost;                   // relative cost of one tile (for work partitioning)
    int qm_var, qn_var;         // RADOP_VAR_* holding mass / number density
    float dD_geo;               // (D[nd-1]-D[0])/(nd-1) computed in double
    float dD_ref;               // D[1]-D[0] in float32 (doppler_scatter.py:224)
    float d_min, d_max;         // first / last diameter of the table axis (float32)
    double inv_beta, inv_alpha;
    double wh_lo, wh_hi;        // fall speeds alpha D^beta of d_min (x (1 - 1e-9)) and d_max (x (1 + 1e-9))
    float spec_step;            // float32(np.linspace(d_min, d_max, nd))[1] - [0]: step_D of the spectrum
                                // (doppler_scatter.py:686-692; differs from dD_ref by ulps)
    double dD_d;                // dD_geo in float64
    const float *table;         // [ne][nt][nd][ncol] combined + premultiplied columns
    const float *aux;           // [3][nd]: D, D^nu, D^0.6357
    const float *lamtab;        // [ne*nt][LT_NP][lt_stride] lambda-transformed table (k_build_lamtab) or null
    int lt_nc, lt_stride;       // columns tabulated, floats per (bin, piece)
    double a, b, alpha, beta, mu, nu, N0, lambda_factor, vel_factor, ntot_factor, x_min, x_max;
    // log-domain constants of the PSD parameters (filled by the host, radop_api.cu prepare()):
    //   log(lambda) = c_inv * (c_l0 [+ log N0(T)] - log QM)          1-mom rain / graupel / snow
    //   log(lambda) = c_inv * log(lambda_factor * x)                   2-mom (before the mm scaling)
    double c_l0, c_inv;
    double c_v, e_v;            // v-moment  = c_v * N0 * lambda^e_v    (_Hydrometeor.integrate_V)
    double c_n, e_n;            // n-moment  = c_n * N0 * lambda^e_n
    double c_q, e_q;            // 2-mom N0  = c_q * QN * lambda^e_q * exp(-c_ka); lambda *= exp(-c_kb)
    double c_ka, c_kb;
    double melt[8];             // RadopSpecies.melt (melting species)
};

struct DevConfig {
    // radar / scan geometry
    int n_rays, n_sub, n_gates, n_vert, center_sub;   // n_rays: rays of this launch (a chunk of the scan)
    long long out_stride;       // rays x gates of the whole scan: stride between output variables
    double radar_lat, radar_lon;
    double cos_lon0, sin_lon0;  // of the radar longitude
    double geo_mid, geo_rhalf;  // polynomial geodesy: x = (s - geo_mid) * geo_rhalf in [-1, 1] over the scan's ground distances
    // model grid
    int nz, ny, nx;
    float llc[2], urc[2], res[2];
    float rres[2];              // RN(1 / res)
    double rot[9];              // rotation matrix geographic -> rotated pole
    uint32_t var_mask;          // uploaded variables
    // microphysics
    int n_species, total_bins;
    int two_moment, doppler_scheme, with_attenuation, sensitivity_mode;
    int z_strict;               // every z-level column is strictly descending (fast column search)
    int div_exact;              // 1: every quotient of the vertical interpolation of this grid and these fields takes the
                                // shared-reciprocal path exactly (bounds taken at upload, k_level_diff_bounds): no range checks
    const uint8_t *lev_guess;   // [1024] level of a reference column against the normalised height, or null
    const float2 *z_tb;         // [ny][nx] (top, bottom) level height of every column: the two ends the search and the topography test
                                // need, in ONE compact array (neighbouring columns share a cache line; in z_t they are nz floats apart)
    int output_model_vars, keep_debug, ice_numeric;   // ice_numeric: species index with numeric v,n or -1
    float wavelength, k_squared, radial_res;
    float sens0, sens1;
    // Doppler spectrum (scheme 3): VARRAY[j] = vel_min + j vel_res, j < n_vel (global_constants.py:168-171)
    int n_vel;
    double vel_min, vel_res, inv_vel_res;
    // broadening of the sub-beam spectra (doppler_scatter.py:727-801): bit 0 turbulence, bit 1 antenna motion
    int with_melting;
    int sum_species;            // sz_out holds ONE row per sub-beam gate, already summed over the species (table mode
                                // without melting species and without Doppler spectra, whose kernels need the rows per species)
    int brd_mode;
    float brd_motion_num;       // float32(wavelength / 100 * antenna_speed)
    double brd_motion_den;      // 2 pi deg2rad(beamwidth)
    double brd_res;             // radial resolution (RANGE_RADAR[g] = res / 2 + g res)
    double brd_sigma_r, brd_sigma_theta, brd_turb_far, brd_turb_near;
    SpeciesDev sp[RADOP_NSPECIES_MAX];
};

// sorted work record of one active (sub-beam gate, species) pair
struct __align__(16) PairRec {
    uint32_t sbg;   // sub-beam gate index
    float p0;       // lambda (GEO/GEN) or lambda/1000 (ICE1)
    float p1;       // N0 (GEO/GEN) or QM (ICE1)
    float w;        // quadrature weight of the sub-beam
};
