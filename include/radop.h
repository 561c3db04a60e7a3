/*
 * radop.h -- C-ABI of libradop.so: the B200 (sm_100a) implementation of the
 * per-gate hot path of wolfidan/cosmo_pol.
 *
 * What it replaces in the reference (paths relative to the reference root):
 *   the body of the closure `worker(event, elev, azimuth)` that get_PPI / get_RHI
 *   map over rays with multiprocess.Pool  (cosmo_pol/radar_operator.py:407-432,
 *   504-530), i.e.
 *     get_interpolated_radial   cosmo_pol/interpolation/interpolation.py:91-496
 *       trilin_interp_radial    cosmo_pol/interpolation/interpolation.py:498-597
 *       get_all_radar_pts       cosmo_pol/interpolation/interpolation_c.c:12-104   (SWIG, C)
 *     get_radar_observables     cosmo_pol/scatter/doppler_scatter.py:49-489
 *       Lookup_table.lookup_line cosmo_pol/lookup/lut.py:309-344
 *       _Hydrometeor.set_psd/get_N/integrate_V  cosmo_pol/hydrometeors/hydrometeors.py:128-255
 *       get_pol_from_sz         cosmo_pol/scatter/doppler_scatter.py:491-544
 *     integrate_radials         cosmo_pol/interpolation/interpolation.py:36-89
 *     cut_at_sensitivity        cosmo_pol/scatter/doppler_scatter.py:804-862
 *   One radop_simulate() call processes the WHOLE ray list of a scan instead of
 *   one pool task per ray.
 *
 * Conventions: plain C, plain pointers and sizes, no torch / CUDA types in any
 * signature.  Every function returns an int status (0 = ok); the text of the
 * last error of a handle is available through radop_last_error().  Nothing
 * aborts or throws across the ABI.  Host buffers are owned by the caller and
 * never retained after a call returns; device memory is owned by the handle.
 * A handle is bound to one GPU and is not thread-safe.
 */
#ifndef RADOP_H
#define RADOP_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RADOP_ABI_VERSION 6

/* ---- status codes ------------------------------------------------------- */
enum {
    RADOP_OK = 0,
    RADOP_ERR_CUDA = 1,          /* a CUDA call failed (see radop_last_error)            */
    RADOP_ERR_ARG = 2,           /* bad argument                                          -> ValueError */
    RADOP_ERR_OUT_OF_DOMAIN = 3, /* a gate falls outside the model domain                 -> IndexError
                                    (cosmo_pol/interpolation/interpolation.py:572-580)   */
    RADOP_ERR_LUT = 4,           /* lookup table missing / wrong shape                    */
    RADOP_ERR_STATE = 5,         /* grid / fields / params not uploaded yet               */
    RADOP_ERR_UNSUPPORTED = 6    /* option not implemented in this build                  */
};

/* ---- model variables (cosmo_pol/radar_operator.py:35-36) ------------------ */
enum {
    /* ids = slots of the interleaved field layout: the eight variables the radar observables of the one-moment
       scheme need come first, so that one (cell, level) of them is two 16-byte loads */
    RADOP_VAR_U = 0, RADOP_VAR_V, RADOP_VAR_W, RADOP_VAR_T,
    RADOP_VAR_QR, RADOP_VAR_QS, RADOP_VAR_QG, RADOP_VAR_QI, RADOP_VAR_RHO,          /* 9 one-moment vars */
    RADOP_VAR_QH, RADOP_VAR_QNR, RADOP_VAR_QNS, RADOP_VAR_QNG, RADOP_VAR_QNI, RADOP_VAR_QNH,
    RADOP_VAR_EDR,               /* eddy dissipation rate, radar_operator.py:253-254 (turbulence broadening) */
    RADOP_NVAR_MAX = 16
};

/* ---- hydrometeor species (order of doppler_scatter.py:99-106) ------------- */
enum { RADOP_SP_R = 0, RADOP_SP_S, RADOP_SP_G, RADOP_SP_H, RADOP_SP_I,
       RADOP_SP_MS, RADOP_SP_MG,          /* melting snow / graupel (with_melting, melting.py:19-90) */
       RADOP_NSPECIES_MAX = 7 };

/* PSD families (cosmo_pol/hydrometeors/hydrometeors.py) */
enum {
    RADOP_PSD_1MOM_FIXED_N0 = 0, /* Rain :761-772, Graupel :1039-1049                     */
    RADOP_PSD_1MOM_SNOW = 1,     /* Snow :894-907 (N0 from temperature)                   */
    RADOP_PSD_1MOM_ICE = 2,      /* IceParticle :1320-1339 (Field et al. 2005 moments)    */
    RADOP_PSD_2MOM = 3,          /* _Hydrometeor.set_psd :224-255, IceParticle :1341-1365 */
    RADOP_PSD_MELTING = 4        /* _MeltingHydrometeor :303-500, MeltingSnow / MeltingGraupel :1393-1481
                                    (one-moment scheme; PSD from the flux of the equivalent rain)  */
};

/* One species: constants of cosmo_pol/constants/constants_{1,2}mom.py */
typedef struct {
    int32_t species;        /* RADOP_SP_*                                                 */
    int32_t psd;            /* RADOP_PSD_*                                                */
    int32_t numeric_vn;     /* 1: fall-speed integrals are numeric (IceParticle.integrate_V,
                               hydrometeors.py:1256-1275, incl. its all-gate np.sum)       */
    int32_t reserved;
    double a, b;            /* m(D) = a D^b                                               */
    double alpha, beta;     /* v(D) = alpha D^beta                                        */
    double mu, nu;          /* N(D) = N0 D^mu exp(-lambda D^nu)                           */
    double N0;              /* fixed intercept (1-mom rain / graupel)                     */
    double lambda_factor, vel_factor, ntot_factor;
    double x_min, x_max;    /* 2-mom mean-mass limits                                     */
    /* RADOP_PSD_MELTING: a, b, alpha, beta above are those of the dry counterpart (snow / graupel);
       mu, nu, N0, lambda_factor those of the equivalent rain; and
       melt[0..3] = rain a, b, alpha, beta; melt[4..5] = rain D_MIN, D_MAX; melt[6..7] = dry D_MIN, D_MAX
       (constants_1mom.py; the wet-fraction dependent limits of hydrometeors.py:333-340)              */
    double melt[8];
} RadopSpecies;

/* Scalars the hot path reads from cfg.CONFIG / global_constants */
typedef struct {
    int32_t abi_version;          /* RADOP_ABI_VERSION                                    */
    int32_t n_gates;              /* len(RANGE_RADAR), global_constants.py:186-190        */
    double radar_lat, radar_lon, radar_alt;   /* CONFIG['radar']['coords']                */
    double wavelength_mm;         /* constants.WAVELENGTH                                 */
    double k_squared;             /* CONFIG['radar']['K_squared']                         */
    double radial_resolution_m;   /* CONFIG['radar']['radial_resolution']                 */
    int32_t two_moment;           /* CONFIG['microphysics']['scheme'] == '2mom'           */
    int32_t with_attenuation;     /* CONFIG['microphysics']['with_attenuation']           */
    int32_t doppler_scheme;       /* CONFIG['doppler']['scheme']: 1 analytic, 2 rcs-weighted,
                                     3 full spectrum (get_doppler_spectrum :598-725)       */
    int32_t sensitivity_mode;     /* 0: no cut; 1: [ZH]; 2: [ZH, range]; cut_at_sensitivity */
    double sensitivity[3];
    int32_t output_model_vars;    /* integrate_radials outputs wanted ('all'/'only_model') */
    int32_t keep_debug;           /* keep per sub-beam-gate indices for radop_debug_*      */
    int32_t n_species;
    int32_t reserved;
    RadopSpecies species[RADOP_NSPECIES_MAX];   /* in summation order                      */
    /* Doppler scheme 3: constants.VARRAY = vel_min + j * vel_res, j < n_vel
       (np.arange(-NVEL, NVEL + VRES, VRES), global_constants.py:168-171)                  */
    int32_t n_vel;
    int32_t reserved2;
    double vel_min, vel_res;
    /* Doppler scheme 3, broadening of the sub-beam spectra (doppler_scatter.py:357-367, 727-801) */
    int32_t spectrum_turbulence;  /* CONFIG['doppler']['turbulence_correction'] (needs RADOP_VAR_EDR) */
    int32_t spectrum_motion;      /* CONFIG['doppler']['motion_correction']                      */
    double antenna_speed;         /* CONFIG['radar']['antenna_speed']                            */
    double beamwidth_deg;         /* CONFIG['radar']['3dB_beamwidth']                            */
    int32_t with_melting;         /* CONFIG['microphysics']['with_melting']: re-partition QR, QS, QG into
                                     the melting species after the interpolation (melting.py:19-90) */
    int32_t reserved3;
} RadopParams;

/*
 * One scan = a list of rays, each with the same fan of sub-beams (quadrature
 * points, interpolation.py:148-166 + 343-354) and a table of refracted
 * trajectories (atm_refraction.py:181-220) shared by all rays of one elevation.
 *   sub-beam q of ray r:  azimuth = ray_az[r] + sub_daz[q]
 *                         trajectory row = (ray_traj[r] * n_vert + sub_jv[q])
 */
typedef struct {
    int32_t n_rays;
    int32_t n_sub;
    int32_t n_traj;
    int32_t n_vert;
    const double *ray_az;       /* [n_rays] degrees                                       */
    const int32_t *ray_traj;    /* [n_rays] trajectory group of the ray                   */
    const float *ray_nyquist;   /* [n_rays] Nyquist velocity or NULL (no aliasing)        */
    const double *sub_daz;      /* [n_sub] azimuth offset (pts_hor)                       */
    const int32_t *sub_jv;      /* [n_sub] vertical quadrature index                      */
    const double *sub_weight;   /* [n_sub] normalised quadrature weight                   */
    const float *traj_s;        /* [n_traj][n_vert][n_gates] ground distance [m]          */
    const float *traj_h;        /* [n_traj][n_vert][n_gates] height [m]                   */
    const float *traj_e;        /* [n_traj][n_vert][n_gates] incidence elevation [deg]    */
    /* Doppler scheme 3 only, optional (NULL: computed on the device): float32 sine and tangent
     * of the elevation folded to [0, 90] deg, as numpy gives them for a float32 angle
     * (get_diameter_from_rad_vel, doppler_scatter.py:571-575), same shape as traj_e         */
    const float *traj_sin;
    const float *traj_tan;
    /* Integration scheme 'ml' (interpolation.py:168-193, 423-436), optional (NULL: every sub-beam keeps its weight at
     * every gate): sub_ml_only[q] = 1 for a sub-beam whose weight only counts near the first and the last
     * melting-layer gate of the sub-beam (Gaussian-smoothed spikes, sigma = 2 gates); needs with_melting. */
    const int32_t *sub_ml_only;   /* [n_sub] */
} RadopScan;

/* radar observables, linear units (doppler_scatter.py:446-460) */
enum {
    RADOP_OUT_ZH = 0, RADOP_OUT_ZDR, RADOP_OUT_ZV, RADOP_OUT_KDP, RADOP_OUT_DELTA_HV,
    RADOP_OUT_PHIDP, RADOP_OUT_RHOHV, RADOP_OUT_ATT_H, RADOP_OUT_ATT_V, RADOP_OUT_RVEL,
    RADOP_N_OUT = 10
};

/* Caller-allocated HOST buffers, each [n_rays][n_gates]; NULL = not wanted. */
typedef struct {
    float *obs[RADOP_N_OUT];
    float *mask;                      /* averaged sub-beam mask, doppler_scatter.py:472-477 */
    double *lat, *lon;                /* central sub-beam gate positions (WGS84)           */
    float *model[RADOP_NVAR_MAX];     /* integrate_radials of each uploaded variable; with with_melting (one-moment
                                         scheme) the slots 9..12 hold QmS_v, QmG_v, fwet_mS, fwet_mG (melting.py:74-90) */
    float *dspectrum;                 /* [n_rays][n_gates][n_vel] DSPECTRUM (Doppler scheme 3) */
} RadopOut;

typedef struct RadopHandle_ RadopHandle;

/* ---- life cycle ----------------------------------------------------------- */
int radop_create(int device, RadopHandle **out);
int radop_destroy(RadopHandle *h);
const char *radop_last_error(const RadopHandle *h);
int radop_abi_version(void);
/* sizeof() of the ABI structs as this library was compiled: a binding checks its own layout against it
 * before the first radop_set_params (which = RADOP_SIZEOF_*; unknown -> -1). */
enum { RADOP_SIZEOF_PARAMS = 0, RADOP_SIZEOF_SPECIES, RADOP_SIZEOF_SCAN, RADOP_SIZEOF_OUT, RADOP_SIZEOF_STATS };
int radop_sizeof(int which);

/* ---- configuration and one-time uploads ----------------------------------- */
int radop_set_params(RadopHandle *h, const RadopParams *p);

/* Model grid: z-levels [nz][ny][nx] float32 (level 0 = model top), lower-left and
 * upper-right corners (Lo, La) and resolution (dlon, dlat) in rotated degrees,
 * rotated south pole (lat, lon).  interpolation.py:547-568. */
int radop_upload_grid(RadopHandle *h, int nz, int ny, int nx, const float *zlevels,
                      const float llc[2], const float urc[2], const float res[2],
                      const double south_pole[2]);
/* One model variable [nz][ny][nx] float32 (var.data, interpolation.py:585). */
int radop_upload_field(RadopHandle *h, int var_id, const float *data);

/* One scattering table (lut.Lookup_table): value_table [ne][nt][nd][12] float64,
 * axis minimum and step as float32 (axes_limits / axes_step), diameters [nd] float32.
 * Melting species: the second axis is the wet fraction 'wc' (t_min, t_step = its minimum and
 * step); their diameters depend on the gate, d_axis is only used for its length. */
int radop_upload_lut(RadopHandle *h, int species, const double *value_table,
                     int ne, int nt, int nd, float e_min, float e_step,
                     float t_min, float t_step, const float *d_axis);

/* ---- the hot path ---------------------------------------------------------- */
/* host in -> device -> host out (the reference-facing call).  scan = NULL: the scan staged by the previous call is
 * used again (same rays, fan and trajectories: its upload and the per sub-beam geometry are skipped -- the case of
 * one scan strategy run over successive model states). */
int radop_simulate(RadopHandle *h, const RadopScan *scan, const RadopOut *out);
/* the same in three steps (used to time the device part alone) */
int radop_stage_scan(RadopHandle *h, const RadopScan *scan);   /* H2D                    */
int radop_execute(RadopHandle *h);                              /* kernels, asynchronous  */
int radop_fetch(RadopHandle *h, const RadopOut *out);           /* sync + D2H             */
int radop_synchronize(RadopHandle *h);
/* The ray list is processed in chunks of about this many sub-beam gates (0 = one chunk; negative = the
 * default: 3e6 where the fused kernel runs, 8e6 for the staged pipeline): work memory is sized for one
 * chunk and, in radop_simulate, the device -> host copies of a chunk overlap the kernels of the next
 * one.  Results do not depend on the chunking. */
int radop_set_chunk(RadopHandle *h, int64_t subbeam_gates_per_chunk);
/* radop_simulate splits the scan into chunks only when the requested outputs are at least this large
 * (default 32 MiB): every chunk costs a pass over the bin table and partial tiles, which only pays when
 * there is a device -> host copy to hide.  0 = always chunk (used by the chunk-invariance tests). */
int radop_set_overlap_threshold(RadopHandle *h, int64_t min_output_bytes);

/* Multi-GPU delivery (stands in for the result list that pool.map returns, radar_operator.py:429-432).
 * The ranks of one job deal blocks of `block_rays` consecutive rays of the scan in turn; every rank hands
 * radop_simulate / radop_fetch output pointers into ONE host array shared by the ranks (each pointer at the
 * first row of the rank: array + rank * block_rays rows) and sets block_pitch_rays = n_ranks * block_rays:
 * local ray i then lands on row (i / block_rays) * block_pitch_rays + i % block_rays, whole blocks as one
 * strided device -> host copy over the rank's own PCIe link.  No device-side collective is involved.
 * block_rays = 0 (default): the caller's arrays hold this handle's rays only, contiguously. */
int radop_set_output_blocks(RadopHandle *h, int block_rays, int block_pitch_rays);
/* Page-lock (cudaHostRegister, portable) / release a host range the caller owns, e.g. the shared-memory
 * segment of the ranks, so that the copies above are asynchronous DMA. */
int radop_host_register(RadopHandle *h, void *ptr, int64_t bytes);
int radop_host_unregister(RadopHandle *h, void *ptr);

/* Device-resident results of the last radop_execute (for NCCL gathers): pointer to
 * float32 [RADOP_N_OUT + 1][n_rays][n_gates] (observables then mask). */
/* Geolocation of the gates (stands in for pyproj.Geod.fwd + pycosmo.WGS_to_COSMO,
 * interpolation.py:526-568).  mode 0 (default): the rotated-pole coordinates of a sub-beam are
 * polynomials in the ground distance, fitted per (ray, sub-beam) to the Vincenty solution at 16
 * Chebyshev nodes; scans longer than 0.05 rad of arc or with a non-converging series use the
 * per-gate solution automatically.  mode 1: per-gate solution always.
 * radop_get_geodesy: *used_polynomials = 1 if the last execute ran on the polynomials. */
int radop_set_geodesy(RadopHandle *h, int mode);
/* Level search in strictly descending z-level columns (interpolation_c.c:108-135 binary_search): mode 0 (default)
 * guesses the level from the height normalised by the column's own top and bottom (table of the grid's middle column,
 * made by radop_upload_grid), reads four consecutive levels around the guess and takes the reference's answer from them
 * when they bracket the height, else bisects; mode 1 bisects always.  The indices are the same in both modes. */
int radop_set_search(RadopHandle *h, int mode);
/* How the PSD x scattering-table contraction (hydrometeors.py:128-152 get_N, doppler_scatter.py:246-268
 * einsum over the diameters) is evaluated.  mode 0 (default): through lambda-transformed tables -- the sum
 * over the diameters of one (e, t) bin is a function of the PSD slope alone; it is tabulated in float64 when a
 * table is first used, every piece verified against the direct sum, and evaluated per pair (k_integrate_tab).
 * mode 1: directly, diameter by diameter (k_integrate).  radop_get_integration reports the pieces tabulated,
 * the pieces rejected by the verification and the pairs of the last execute that were summed directly
 * inside mode 0 (outside the tabulated range or on a rejected piece).  Any output pointer may be NULL. */
int radop_set_integration(RadopHandle *h, int mode);
int radop_get_integration(RadopHandle *h, int *mode, int64_t *n_pieces, int64_t *n_rejected_pieces,
                          int64_t *n_direct_pairs, float *build_ms);
/* Diagnostics of mode 0: why[] = direct pairs of the last execute by reason (rejected piece, slope below the
 * tabulated range, above it, not a positive normal number); piece_status[bin * 176 + piece] of one species'
 * table = exponent word of the first column, INT32_MIN for a rejected piece (n_status >= ne * nt * 176). */
int radop_debug_integration(RadopHandle *h, int species, int64_t why[4], int32_t *piece_status, int64_t n_status);
int radop_get_geodesy(RadopHandle *h, int *used_polynomials);
/* Launch sequence of the hot path.  mode 0 (default): the fused kernel k_gate_fused (geolocation, interpolation,
 * PSD parameters, table contraction, species and sub-beam sums in one pass, nothing per sub-beam gate written to
 * HBM) whenever the configuration allows it -- table integration, no melting species, Doppler scheme 1 or 2, no
 * model-variable outputs, no debug dump, polynomial geodesy -- else the staged pipeline (k_interp_psd ->
 * k_integrate_tab / k_integrate -> k_reduce_subbeams).  mode 1: the staged pipeline always.  Both are held to the
 * oracle and to each other by the GPU tests.  radop_get_pipeline: *used_fused = 1 if the last run was fused. */
int radop_set_pipeline(RadopHandle *h, int mode);
/* Launch-shape knobs (results do not depend on them): RADOP_TUNE_FUSED_WARPS = warps per CTA of the fused kernel,
 * 0 = chosen from the number of sub-beams. */
enum { RADOP_TUNE_FUSED_WARPS = 0 };
int radop_set_tuning(RadopHandle *h, int key, int64_t value);
int radop_get_pipeline(RadopHandle *h, int *used_fused);

int radop_device_outputs(RadopHandle *h, void **dev_ptr, int64_t *n_rays, int64_t *n_gates);
/* Any device-resident result of the last radop_execute, for gathers that must not bounce through the host
 * (one rank per GPU, cosmo_pol_b200/parallel.py): which = RADOP_BUF_*; *dev_ptr / *bytes describe
 *   RADOP_BUF_OBS        float32 [RADOP_N_OUT + 1][n_rays][n_gates]   (same as radop_device_outputs)
 *   RADOP_BUF_LAT / LON  float64 [n_rays][n_gates]                    central sub-beam positions
 *   RADOP_BUF_MODEL      float32 [RADOP_NVAR_MAX][n_rays][n_gates]    integrate_radials (output_model_vars)
 *   RADOP_BUF_DSPECTRUM  float32 [n_rays][n_gates][n_vel]             Doppler scheme 3
 * A buffer the configuration does not produce gives RADOP_ERR_STATE. */
enum { RADOP_BUF_OBS = 0, RADOP_BUF_LAT, RADOP_BUF_LON, RADOP_BUF_MODEL, RADOP_BUF_DSPECTRUM };
int radop_device_buffer(RadopHandle *h, int which, void **dev_ptr, int64_t *bytes);
/* cudaStream_t of the handle (as void*), to record events on. */
int radop_get_stream(RadopHandle *h, void **stream);

/* ---- introspection ---------------------------------------------------------- */
typedef struct {
    int64_t n_subbeam_gates;      /* rays x sub-beams x gates of the last execute         */
    int64_t n_active_pairs;       /* (sub-beam gate, species) pairs with Q > 0            */
    int64_t n_lut_bins;           /* distinct (species, e, t) bins touched                */
    int64_t n_tiles;              /* work tiles of the PSD-integration kernel             */
    int64_t launches;             /* kernels launched by the last execute                 */
    float ms_interp, ms_sort, ms_integrate, ms_reduce, ms_total;  /* CUDA-event times: mean
                                     over the executes since radop_enable_timing(h, 1)     */
    int32_t n_timed;              /* executes in that mean                                  */
    float ms_spectrum;            /* Doppler-spectrum kernels (scheme 3), included in ms_total */
    int64_t pairs_by_species[RADOP_NSPECIES_MAX];   /* active pairs per species (params order)   */
    double exec_flop_per_pair[RADOP_NSPECIES_MAX];  /* FP32 operations the integration kernel
                                     executes per pair of that species (2 per FMA lane)        */
} RadopStats;
int radop_get_stats(RadopHandle *h, RadopStats *s);
int radop_enable_timing(RadopHandle *h, int on);

/* Per sub-beam-gate integer decisions of the last execute (params.keep_debug = 1):
 * cell [n][2] (lat idx, lon idx), level [n][4], status [n] (0 ok, 1 above top,
 * -1 below topography), lut_e / lut_t [n][n_species] (-1 = species inactive),
 * values [n][n_uploaded_vars] interpolated variables; n = rays*sub*gates. NULL = skip. */
int radop_debug_indices(RadopHandle *h, int32_t *cell, int32_t *level, int32_t *status,
                        int32_t *lut_e, int32_t *lut_t, float *values, double *rot_coords);

/* Measured FP32 FMA throughput of this GPU [TFLOP/s] (roofline denominator). */
int radop_measure_fp32_peak(RadopHandle *h, double *tflops);

/* Self-test of the shared-reciprocal float32 division used by the interpolation kernel
 * against the IEEE division (3 x n_samples seeded operand pairs); *mismatches must be 0. */
int radop_selftest_division(RadopHandle *h, uint64_t n_samples, uint64_t seed, uint64_t *mismatches);

/* ---- compatibility shims over the kernels (same data contract as the SWIG
 *      functions they replace; used for A/B tests against the gcc build) -------- */
/* interpolation_c.c:12  get_all_radar_pts(len, coords[n][2], heights[n], data, z, llc, res) */
int radop_get_all_radar_pts(RadopHandle *h, int n, const float *coords_rad_pts,
                            const float *radar_heights, const float *model_data,
                            const float *model_heights, int nz, int ny, int nx,
                            const float *llc_cosmo, const float *res_cosmo, float *output);
/* doppler_c.c:11  get_refl(refl[len], Da[len][nhydro], Db[len][nhydro], rcs[nd][nhydro], D[nd][nhydro],
 *                          N[nd][nhydro], step_D[nhydro], Dmin[nhydro]):
 * refl[i] = sum_j step_D[j] * sum_{k = int((Da - Dmin)/step) .. int((Db - Dmin)/step) - 1} N[k][j] rcs[k][j]
 * in the reference's float32 order of operations (bit-exact against the gcc build for in-range indices;
 * indices outside [0, nd] -- undefined behaviour in the reference -- are clamped). */
int radop_get_refl(RadopHandle *h, int len, int nhydro, int nd, const float *Da, const float *Db,
                   const float *rcs, const float *D, const float *N, const float *step_D,
                   const float *Dmin, float *refl);

/* ---- allocation-free entry point -------------------------------------------------
 * No counterpart in the reference: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid allocation).  Same contract and same bits as radop_get_refl above, but the
 * device copies of the operands and of the result live in one file-scope
 * `static __device__ __align__(256)` array of RADOP_NOALLOC_SCRATCH_BYTES inside libradop.so
 * (every GPU that loads the library holds it until the process exits); the call makes no
 * cudaMalloc / cudaFree and calls no allocating library routine.
 * Limit: 4 * (2*len*nhydro + 2*nd*nhydro + 2*nhydro + len) <= RADOP_NOALLOC_SCRATCH_BYTES,
 * else RADOP_ERR_ARG with the sizes in radop_last_error() (use radop_get_refl then).
 * All calls on a GPU share the array and it keeps the previous call's contents; nothing in it
 * has to start at zero (every word read is first overwritten by this call's copies, every
 * result word is written by the kernel).  The call returns after its stream has drained, so
 * calls from one thread never overlap; calls from several threads or handles on the same
 * GPU that overlap on the device RACE on the array -- serialise them. */
#define RADOP_NOALLOC_SCRATCH_BYTES (4u << 20)
int radop_get_refl_noalloc(RadopHandle *h, int len, int nhydro, int nd, const float *Da, const float *Db,
                           const float *rcs, const float *D, const float *N, const float *step_D,
                           const float *Dmin, float *refl);

#ifdef __cplusplus
}
#endif
#endif /* RADOP_H */
