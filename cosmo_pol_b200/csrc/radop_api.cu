// radop_api.cu -- host side of libradop.so: the C-ABI declared in include/radop.h.
// Owns device memory, the private CUDA stream, the one-time layout transforms and the
// launch sequence of the hot path.  No CPU fallback exists: every entry point needs a GPU.
#include "radop_kernels.cu"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

namespace {

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

struct LutHost {
    bool loaded = false;
    int ne = 0, nt = 0, nd = 0, mode = 0, ncol = 8, cost = 9;
    float e_min = 0, e_step = 1, t_min = 0, t_step = 1, dD_geo = 0, dD_ref = 0;
    double dD_d = 0;
    float d_min = 0, d_max = 0, spec_step = 0;
    RadopSpecies packed_for;   // constants the table was premultiplied with
    int packed_doppler = 0;    // Doppler scheme the table was laid out for
    DevBuf table, aux;
    DevBuf lamtab;             // lambda-transformed table (k_build_lamtab), built on first use
    bool lamtab_built = false;
    int lt_nc = 0, lt_stride = 0;
};

}  // namespace

#define RADOP_TOTALS 16      // slots of the per-chunk totals written by k_scan_bins

struct RadopHandle_ {
    int device = 0;
    cudaStream_t stream = nullptr, copy_stream = nullptr;
    std::vector<cudaEvent_t> chunk_done;
    long long chunk_sbg = -1;        // sub-beam gates per chunk (0: one chunk; -1: the default of the pipeline that runs)
    int n_chunks = 1, n_exec_timed = 0;
    std::string err;
    int n_sm = 148;
    // configuration
    bool have_params = false, have_grid = false;
    RadopParams p;
    int nz = 0, ny = 0, nx = 0;
    float llc[2], urc[2], res[2];
    double rot[9];
    uint32_t var_mask = 0;
    int z_strict = 0;
    // bounds of the vertical differences (k_level_diff_bounds): z-level spacings and, per variable, |f[l+1] - f[l]|
    float z_dmin = 0.f, z_dmax = 0.f;
    bool z_dpos = false;
    float var_nmax[RADOP_NVAR_MAX], var_nmin[RADOP_NVAR_MAX];
    DevBuf z_t, fields, stage, lev_guess, z_tb;
    int search_mode = 0;         // 0: guessed window + bisection when it misses, 1: bisection always
    LutHost lut[RADOP_NSPECIES_MAX];   // indexed by RADOP_SP_*
    // scan
    bool have_scan = false, executed = false;
    bool latlon_valid = false;   // the last run computed the central sub-beam positions
    int n_rays = 0, n_sub = 0, n_traj = 0, n_vert = 0, center_sub = 0;
    bool have_nyquist = false;
    size_t overlap_min_bytes = (size_t)32 << 20;   // radop_simulate chunks the scan only above this output size
    long long n_sbg = 0;
    void *pinned = nullptr;
    size_t pinned_cap = 0;
    DevBuf scan_blob;
    ScanDev scan;
    const float *d_nyquist = nullptr;
    // work buffers
    DevBuf geo_poly;
    int geodesy_mode = 0;        // 0: per sub-beam polynomials (per-gate solution when they do not converge), 1: per gate
    bool geo_poly_ok = false, geo_poly_launched = false, geo_poly_used = false;
    double geo_mid = 0.0, geo_half = 1.0;
    DevBuf geo, mask, active, wind, sinth, vsum, nsum, key, p0, p1, vals, lat, lon, hist, flags;
    DevBuf bin_start, tile_start, cost_start, cursor, totals, rec, sz_out, ice_vn, dop_vn, ice_tot, comb, out, model_out;
    DevBuf d3_uh, d3_w, d3_rho, d3_st, attw, spec_pre, spec_acc, spec_out, d3_std, sub_acc, brd_flag;
    size_t k4_smem = 0;
    DevBuf dbg_cell, dbg_level, dbg_status, dbg_e, dbg_t, dbg_rot;
    DevConfig cfg;
    int k2_blocks = 0;
    size_t k2_smem = 0;
    int integration_mode = 0;    // 0: lambda-transformed tables (k_integrate_tab), 1: direct sums over the diameters (k_integrate)
    int pipeline_mode = 0;       // 0: fused kernel when the configuration allows it, 1: staged pipeline always
    int fused_warps = 0;         // warps per CTA of the fused kernel (0: chosen from the sub-beam count)
    bool last_fused = false;
    int out_block = 0, out_pitch = 0;   // radop_set_output_blocks: host rows of local ray i = (i / block) * pitch + i % block
    bool geo_ready = false;      // geo_poly holds the polynomials of the staged scan (whole ray list)
    bool traj_tab_ready = false; // traj_sc / traj_ie (k_traj_tables) hold the staged trajectories for the current tables
    bool geo_poly_conv = false;  // ... and every series converged
    DevBuf traj_sc, traj_ie, f_wsum, f_ice, f_cnt;
    DevBuf ml_wraw, ml_wsz;      // integration scheme 'ml': per-gate weights of a chunk
    DevBuf lt_stat;              // [0] u64 pairs integrated directly by k_integrate_tab, [1] (int) rejected pieces
    long long lt_pieces = 0, lt_rejected = 0, lt_fallback = 0, lt_why[4] = {0, 0, 0, 0};
    float lt_build_ms = 0.f;
    // timing
    bool timing = false;
    std::vector<cudaEvent_t> ev;   // 6 events per timed execute since radop_enable_timing(1)
    int n_timed = 0;
    RadopStats stats;

    int fail(int code, const std::string &msg)
    {
        err = msg;
        return code;
    }
};

#define CU(h, expr)                                                                                   \
    do {                                                                                              \
        cudaError_t e__ = (expr);                                                                     \
        if (e__ != cudaSuccess) {                                                                     \
            char buf__[512];                                                                          \
            snprintf(buf__, sizeof buf__, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__),   \
                     __FILE__, __LINE__);                                                             \
            return (h)->fail(RADOP_ERR_CUDA, buf__);                                                  \
        }                                                                                             \
    } while (0)

static int species_vars(int sp, int *qm, int *qn)
{
    switch (sp) {
    case RADOP_SP_R: *qm = RADOP_VAR_QR; *qn = RADOP_VAR_QNR; return 0;
    case RADOP_SP_S: *qm = RADOP_VAR_QS; *qn = RADOP_VAR_QNS; return 0;
    case RADOP_SP_G: *qm = RADOP_VAR_QG; *qn = RADOP_VAR_QNG; return 0;
    case RADOP_SP_H: *qm = RADOP_VAR_QH; *qn = RADOP_VAR_QNH; return 0;
    case RADOP_SP_I: *qm = RADOP_VAR_QI; *qn = RADOP_VAR_QNI; return 0;
    // melting species are made from QR, QS and QG after the interpolation (melting.py:19-90)
    case RADOP_SP_MS: *qm = RADOP_VAR_QS; *qn = RADOP_VAR_QNS; return 0;
    case RADOP_SP_MG: *qm = RADOP_VAR_QG; *qn = RADOP_VAR_QNG; return 0;
    }
    return -1;
}

extern "C" {

int radop_abi_version(void) { return RADOP_ABI_VERSION; }

int radop_sizeof(int which)
{
    switch (which) {
    case RADOP_SIZEOF_PARAMS: return (int)sizeof(RadopParams);
    case RADOP_SIZEOF_SPECIES: return (int)sizeof(RadopSpecies);
    case RADOP_SIZEOF_SCAN: return (int)sizeof(RadopScan);
    case RADOP_SIZEOF_OUT: return (int)sizeof(RadopOut);
    case RADOP_SIZEOF_STATS: return (int)sizeof(RadopStats);
    }
    return -1;
}

int radop_create(int device, RadopHandle **out)
{
    if (!out) return RADOP_ERR_ARG;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0 || device < 0 || device >= count) {
        cudaGetLastError();
        return RADOP_ERR_CUDA;   // no GPU: there is no CPU fallback
    }
    RadopHandle *h = new RadopHandle_();
    h->device = device;
    if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete h;
        return RADOP_ERR_CUDA;
    }
    cudaDeviceGetAttribute(&h->n_sm, cudaDevAttrMultiProcessorCount, device);
    memset(&h->stats, 0, sizeof h->stats);
    memset(&h->p, 0, sizeof h->p);
    *out = h;
    return RADOP_OK;
}

int radop_destroy(RadopHandle *h)
{
    if (!h) return RADOP_OK;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    cudaStreamSynchronize(h->copy_stream);
    DevBuf *bufs[] = {&h->lt_stat, &h->ml_wraw, &h->ml_wsz, &h->traj_sc, &h->traj_ie, &h->f_wsum, &h->f_ice, &h->f_cnt, &h->lev_guess, &h->z_tb, &h->geo_poly, &h->geo, &h->z_t, &h->fields, &h->stage, &h->scan_blob, &h->mask, &h->active, &h->wind, &h->sinth,
                      &h->vsum, &h->nsum, &h->key, &h->p0, &h->p1, &h->vals, &h->lat, &h->lon, &h->hist, &h->flags,
                      &h->bin_start, &h->tile_start, &h->cost_start, &h->cursor, &h->totals, &h->rec, &h->sz_out,
                      &h->ice_vn, &h->dop_vn, &h->ice_tot, &h->d3_uh, &h->d3_w, &h->d3_rho, &h->d3_st, &h->attw, &h->spec_pre, &h->spec_acc, &h->spec_out, &h->d3_std, &h->sub_acc, &h->brd_flag, &h->comb, &h->out, &h->model_out, &h->dbg_cell, &h->dbg_level,
                      &h->dbg_status, &h->dbg_e, &h->dbg_t, &h->dbg_rot};
    for (DevBuf *b : bufs) b->release();
    for (auto &l : h->lut) { l.table.release(); l.aux.release(); l.lamtab.release(); }
    if (h->pinned) cudaFreeHost(h->pinned);
    for (auto &e : h->ev) cudaEventDestroy(e);
    for (auto &e : h->chunk_done) cudaEventDestroy(e);
    cudaStreamDestroy(h->copy_stream);
    cudaStreamDestroy(h->stream);
    delete h;
    return RADOP_OK;
}

const char *radop_last_error(const RadopHandle *h) { return h ? h->err.c_str() : "null handle"; }

int radop_set_params(RadopHandle *h, const RadopParams *p)
{
    if (!h || !p) return RADOP_ERR_ARG;
    if (p->abi_version != RADOP_ABI_VERSION) return h->fail(RADOP_ERR_ARG, "RadopParams.abi_version mismatch");
    if (p->n_gates <= 0) return h->fail(RADOP_ERR_ARG, "n_gates must be positive");
    if (p->n_species <= 0 || p->n_species > RADOP_NSPECIES_MAX) return h->fail(RADOP_ERR_ARG, "bad n_species");
    if (p->doppler_scheme < 1 || p->doppler_scheme > 3)
        return h->fail(RADOP_ERR_UNSUPPORTED, "doppler scheme must be 1, 2 or 3");
    if (p->doppler_scheme == 3 && (p->n_vel < 2 || p->n_vel > 4096 || !(p->vel_res > 0.0)))
        return h->fail(RADOP_ERR_ARG, "doppler scheme 3 needs 2 <= n_vel <= 4096 and vel_res > 0");
    if ((p->spectrum_turbulence || p->spectrum_motion) && p->doppler_scheme == 3 &&
        !(p->beamwidth_deg > 0.0 && p->radial_resolution_m > 0.0))
        return h->fail(RADOP_ERR_ARG, "spectrum broadening needs beamwidth_deg > 0 and radial_resolution_m > 0");
    if (p->sensitivity_mode < 0 || p->sensitivity_mode > 2)
        return h->fail(RADOP_ERR_UNSUPPORTED, "sensitivity must be [ZH] or [ZH, range]");
    for (int s = 0; s < p->n_species; ++s) {
        int qm, qn;
        if (species_vars(p->species[s].species, &qm, &qn)) return h->fail(RADOP_ERR_ARG, "unknown species id");
        if (p->species[s].psd < 0 || p->species[s].psd > RADOP_PSD_MELTING) return h->fail(RADOP_ERR_ARG, "unknown psd family");
        const bool wet_species = p->species[s].species == RADOP_SP_MS || p->species[s].species == RADOP_SP_MG;
        if (wet_species != (p->species[s].psd == RADOP_PSD_MELTING))
            return h->fail(RADOP_ERR_ARG, "melting species and RADOP_PSD_MELTING go together");
        if (wet_species && (!p->with_melting || p->two_moment))
            return h->fail(RADOP_ERR_UNSUPPORTED, "melting species need with_melting and the one-moment scheme "
                                                  "(the reference's two-moment call raises ValueError)");
    }
    h->p = *p;
    h->have_params = true;
    h->executed = false;
    h->geo_ready = false;
    h->traj_tab_ready = false;
    return RADOP_OK;
}

int radop_upload_grid(RadopHandle *h, int nz, int ny, int nx, const float *zlevels, const float llc[2],
                      const float urc[2], const float res[2], const double south_pole[2])
{
    if (!h || !zlevels || !llc || !urc || !res || !south_pole) return RADOP_ERR_ARG;
    if (nz < 4 || ny < 2 || nx < 2) return h->fail(RADOP_ERR_ARG, "grid too small (nz >= 4, ny, nx >= 2)");
    CU(h, cudaSetDevice(h->device));
    size_t n = (size_t)nz * ny * nx;
    CU(h, h->z_t.reserve(n * 4));
    CU(h, h->fields.reserve(n * RADOP_NVP * 4));
    CU(h, h->stage.reserve(n * 4));
    CU(h, cudaMemsetAsync(h->fields.p, 0, n * RADOP_NVP * 4, h->stream));
    CU(h, cudaMemcpyAsync(h->stage.p, zlevels, n * 4, cudaMemcpyHostToDevice, h->stream));
    k_pack_zlevels<<<h->n_sm * 8, 256, 0, h->stream>>>(h->stage.as<float>(), h->z_t.as<float>(), nz, ny, nx);
    CU(h, cudaGetLastError());
    CU(h, h->z_tb.reserve((size_t)ny * nx * 8));
    k_pack_topbot<<<h->n_sm * 8, 256, 0, h->stream>>>(h->z_t.as<float>(), nz, (size_t)ny * nx, h->z_tb.as<float2>());
    CU(h, cudaGetLastError());
    {   // strictly descending columns allow the fast column search of k_interp_psd
        CU(h, h->flags.reserve(16));
        CU(h, cudaMemsetAsync(h->flags.p, 0, 16, h->stream));
        k_check_columns<<<h->n_sm * 8, 256, 0, h->stream>>>(h->z_t.as<float>(), nz, (size_t)ny * nx, h->flags.as<int>());
        CU(h, cudaGetLastError());
        int bad = 0;
        CU(h, cudaMemcpyAsync(&bad, h->flags.p, 4, cudaMemcpyDeviceToHost, h->stream));
        CU(h, cudaStreamSynchronize(h->stream));
        h->z_strict = bad ? 0 : 1;
    }
    {   // spacings of vertically adjacent levels: the divisors of the vertical interpolation
        unsigned init[4] = {0u, 0xffffffffu, 0u, 0u}, got[4];
        CU(h, cudaMemcpyAsync(h->flags.p, init, 16, cudaMemcpyHostToDevice, h->stream));
        k_level_diff_bounds<<<h->n_sm * 8, 256, 0, h->stream>>>(h->stage.as<float>(), nz, (size_t)ny * nx, 1, (unsigned *)h->flags.p);
        CU(h, cudaGetLastError());
        CU(h, cudaMemcpyAsync(got, h->flags.p, 16, cudaMemcpyDeviceToHost, h->stream));
        CU(h, cudaStreamSynchronize(h->stream));
        memcpy(&h->z_dmax, &got[0], 4);
        memcpy(&h->z_dmin, &got[1], 4);
        h->z_dpos = got[2] == 0u && got[1] != 0xffffffffu;
    }
    {   // level against normalised height for the column in the middle of the grid (first guess of the search)
        std::vector<uint8_t> g(RADOP_NGUESS, 0);
        const size_t cidx = (size_t)(ny / 2) * nx + nx / 2, plane = (size_t)ny * nx;
        const double top = zlevels[cidx], bot = zlevels[(size_t)(nz - 1) * plane + cidx];
        int i = 0;
        for (int t = RADOP_NGUESS - 1; t >= 0; --t) {        // heights descending, levels ascending
            const double hh = bot + (top - bot) * ((double)t + 0.5) / RADOP_NGUESS;
            while (i + 1 <= nz - 2 && (double)zlevels[(size_t)(i + 1) * plane + cidx] >= hh) ++i;
            g[t] = (uint8_t)std::min(i, 255);
        }
        CU(h, h->lev_guess.reserve(RADOP_NGUESS));
        CU(h, cudaMemcpy(h->lev_guess.p, g.data(), RADOP_NGUESS, cudaMemcpyHostToDevice));
    }
    CU(h, cudaStreamSynchronize(h->stream));
    h->nz = nz; h->ny = ny; h->nx = nx;
    for (int i = 0; i < 2; ++i) { h->llc[i] = llc[i]; h->urc[i] = urc[i]; h->res[i] = res[i]; }
    // rotation geographic -> rotated pole: about z by the pole longitude, about y by 90 + pole latitude
    const double d2r = 0.017453292519943295;
    double th = (90.0 + south_pole[0]) * d2r, ph = south_pole[1] * d2r;
    double ct = cos(th), st = sin(th), cp = cos(ph), sp = sin(ph);
    double R[9] = {ct * cp, ct * sp, st, -sp, cp, 0.0, -st * cp, -st * sp, ct};
    memcpy(h->rot, R, sizeof R);
    h->var_mask = 0;
    h->have_grid = true;
    h->executed = false;
    h->geo_ready = false;
    h->traj_tab_ready = false;
    return RADOP_OK;
}

int radop_upload_field(RadopHandle *h, int var_id, const float *data)
{
    if (!h || !data) return RADOP_ERR_ARG;
    if (!h->have_grid) return h->fail(RADOP_ERR_STATE, "upload the grid before the fields");
    if (var_id < 0 || var_id >= RADOP_NVAR_MAX) return h->fail(RADOP_ERR_ARG, "unknown variable id");
    CU(h, cudaSetDevice(h->device));
    size_t n = (size_t)h->nz * h->ny * h->nx;
    CU(h, cudaMemcpyAsync(h->stage.p, data, n * 4, cudaMemcpyHostToDevice, h->stream));
    k_pack_field<<<h->n_sm * 8, 256, 0, h->stream>>>(h->stage.as<float>(), h->fields.as<float>(), h->nz, h->ny, h->nx, var_id);
    CU(h, cudaGetLastError());
    {   // vertical differences of the variable: the numerators of the vertical interpolation
        unsigned init[4] = {0u, 0xffffffffu, 0u, 0u}, got[4];
        CU(h, h->flags.reserve(16));
        CU(h, cudaMemcpyAsync(h->flags.p, init, 16, cudaMemcpyHostToDevice, h->stream));
        k_level_diff_bounds<<<h->n_sm * 8, 256, 0, h->stream>>>(h->stage.as<float>(), h->nz, (size_t)h->ny * h->nx, 0, (unsigned *)h->flags.p);
        CU(h, cudaGetLastError());
        CU(h, cudaMemcpyAsync(got, h->flags.p, 16, cudaMemcpyDeviceToHost, h->stream));
        CU(h, cudaStreamSynchronize(h->stream));
        memcpy(&h->var_nmax[var_id], &got[0], 4);
        if (got[1] == 0xffffffffu) h->var_nmin[var_id] = 1.0f;        // a constant field: every numerator is zero
        else memcpy(&h->var_nmin[var_id], &got[1], 4);
    }
    CU(h, cudaStreamSynchronize(h->stream));
    h->var_mask |= (1u << var_id);
    h->executed = false;
    return RADOP_OK;
}

int radop_upload_lut(RadopHandle *h, int species, const double *value_table, int ne, int nt, int nd, float e_min,
                     float e_step, float t_min, float t_step, const float *d_axis)
{
    if (!h || !value_table || !d_axis) return RADOP_ERR_ARG;
    if (!h->have_params) return h->fail(RADOP_ERR_STATE, "set the parameters (species constants) before the tables");
    if (species < 0 || species >= RADOP_NSPECIES_MAX) return h->fail(RADOP_ERR_ARG, "unknown species id");
    if (ne <= 0 || nt <= 0 || nd < 2 || (nd % 4) != 0) return h->fail(RADOP_ERR_LUT, "bad table shape (nd must be a multiple of 4)");
    const RadopSpecies *S = nullptr;
    for (int s = 0; s < h->p.n_species; ++s)
        if (h->p.species[s].species == species) S = &h->p.species[s];
    if (!S) return h->fail(RADOP_ERR_ARG, "species is not part of the current parameters");
    CU(h, cudaSetDevice(h->device));
    LutHost &L = h->lut[species];
    h->traj_tab_ready = false;       // elevation bins of the trajectories belong to the table's axis
    // uniformity of the diameter grid (np.linspace in compute_lut_sz.py:398)
    double d0 = d_axis[0], dD = ((double)d_axis[nd - 1] - d0) / (nd - 1), dev = 0;
    for (int k = 0; k < nd; ++k) dev = std::max(dev, fabs((double)d_axis[k] - (d0 + k * dD)));
    bool uniform = dev <= 4e-7 * std::max(1.0, fabs((double)d_axis[nd - 1])) + 1e-12;
    if (S->psd == RADOP_PSD_MELTING) {
        L.mode = MODE_MELT; L.ncol = 8; L.cost = 1;
    } else if (S->psd == RADOP_PSD_1MOM_ICE) {
        if (!uniform || (nd % RADOP_HBLOCK) != 0)
            return h->fail(RADOP_ERR_LUT, "1-mom ice needs a uniform diameter axis with a multiple of 32 bins");
        L.mode = MODE_ICE1; L.ncol = 12; L.cost = 15;
    } else if (S->numeric_vn && h->p.doppler_scheme == 1) {
        L.mode = MODE_GEN12; L.ncol = 12; L.cost = 15;
    } else if (S->nu == 1.0 && uniform && (nd % RADOP_HBLOCK) == 0) {
        L.mode = MODE_GEO8; L.ncol = 8; L.cost = 9;
    } else {
        L.mode = MODE_GEN8; L.ncol = 8; L.cost = 11;
    }
    L.ne = ne; L.nt = nt; L.nd = nd;
    L.e_min = e_min; L.e_step = e_step; L.t_min = t_min; L.t_step = t_step;
    L.dD_geo = (float)dD;
    L.dD_d = dD;
    L.dD_ref = d_axis[1] - d_axis[0];
    L.d_min = d_axis[0]; L.d_max = d_axis[nd - 1];
    {   // D = np.linspace(d_min, d_max, nd) as the reference forms it for the spectrum
        // (doppler_scatter.py:686-692): float32 end points give float32 arithmetic,
        // y = arange(nd) * ((stop - start) / (nd - 1)) + start; step_D = D[1] - D[0]
        volatile float delta = d_axis[nd - 1] - d_axis[0];
        volatile float step = delta / (float)(nd - 1);
        volatile float D1 = step + d_axis[0];
        L.spec_step = D1 - d_axis[0];
    }
    L.packed_for = *S;
    L.packed_doppler = h->p.doppler_scheme;
    size_t rows = (size_t)ne * nt * nd;
    DevBuf raw, dax;
    CU(h, raw.reserve(rows * 12 * 8));
    CU(h, dax.reserve((size_t)nd * 4));
    CU(h, L.table.reserve(rows * L.ncol * 4));
    CU(h, L.aux.reserve((size_t)3 * nd * 4));
    CU(h, cudaMemcpyAsync(raw.p, value_table, rows * 12 * 8, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(dax.p, d_axis, (size_t)nd * 4, cudaMemcpyHostToDevice, h->stream));
    k_pack_lut<<<h->n_sm * 8, 256, 0, h->stream>>>(raw.as<double>(), L.table.as<float>(), dax.as<float>(), rows, nd,
                                                  L.ncol, L.mode, S->mu, (double)L.dD_ref, S->alpha, S->beta, S->a, S->b,
                                                  h->p.doppler_scheme);
    CU(h, cudaGetLastError());
    k_pack_aux<<<(nd + 255) / 256, 256, 0, h->stream>>>(dax.as<float>(), L.aux.as<float>(), nd, S->nu);
    CU(h, cudaGetLastError());
    CU(h, cudaStreamSynchronize(h->stream));
    raw.release();
    dax.release();
    L.loaded = true;
    L.lamtab_built = false;
    h->executed = false;
    return RADOP_OK;
}

// lambda-transformed tables of every species that does not have them yet (table mode only)
static int build_lamtabs(RadopHandle *h)
{
    DevConfig &c = h->cfg;
    CU(h, h->lt_stat.reserve(64));
    bool any = false;
    for (int s = 0; s < c.n_species; ++s) {
        SpeciesDev &d = c.sp[s];
        LutHost &L = h->lut[d.species];
        if (d.mode == MODE_MELT) continue;
        const int nc = lt_columns(d.mode, c.doppler_scheme), stride = lt_stride(nc);
        if (!L.lamtab_built) {
            if (!any) {
                h->lt_pieces = h->lt_rejected = 0;
                for (int i = 0; i < RADOP_NSPECIES_MAX; ++i)
                    if (h->lut[i].lamtab_built) h->lt_pieces += (long long)h->lut[i].ne * h->lut[i].nt * LT_NP;
            }
            cudaEvent_t e0, e1;
            CU(h, cudaEventCreate(&e0)); CU(h, cudaEventCreate(&e1));
            const size_t bins = (size_t)L.ne * L.nt;
            if (bins * LT_NP > 0x7fffffffull) return h->fail(RADOP_ERR_LUT, "too many (bin, piece) combinations for the lambda tables");
            CU(h, L.lamtab.reserve(bins * LT_NP * stride * 4));
            CU(h, cudaMemsetAsync((char *)h->lt_stat.p + 32, 0, 4, h->stream));
            CU(h, cudaEventRecord(e0, h->stream));
            k_build_lamtab<<<(unsigned)(bins * LT_NP), 256, 0, h->stream>>>(d, nc, stride, L.lamtab.as<float>(),
                                                                            (int *)((char *)h->lt_stat.p + 32));
            CU(h, cudaGetLastError());
            CU(h, cudaEventRecord(e1, h->stream));
            int bad = 0;
            CU(h, cudaMemcpyAsync(&bad, (char *)h->lt_stat.p + 32, 4, cudaMemcpyDeviceToHost, h->stream));
            CU(h, cudaStreamSynchronize(h->stream));
            float ms = 0.f;
            cudaEventElapsedTime(&ms, e0, e1);
            cudaEventDestroy(e0); cudaEventDestroy(e1);
            h->lt_build_ms = any ? h->lt_build_ms + ms : ms;
            h->lt_pieces += (long long)bins * LT_NP;
            h->lt_rejected += bad;
            L.lamtab_built = true; L.lt_nc = nc; L.lt_stride = stride;
            any = true;
        }
        d.lamtab = L.lamtab.as<float>(); d.lt_nc = nc; d.lt_stride = stride;
    }
    return RADOP_OK;
}

static int ensure_pinned(RadopHandle *h, size_t bytes)
{
    if (bytes <= h->pinned_cap) return RADOP_OK;
    if (h->pinned) cudaFreeHost(h->pinned);
    h->pinned = nullptr;
    h->pinned_cap = 0;
    CU(h, cudaHostAlloc(&h->pinned, bytes + bytes / 4 + 4096, cudaHostAllocDefault));
    h->pinned_cap = bytes + bytes / 4 + 4096;
    return RADOP_OK;
}

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

int radop_stage_scan(RadopHandle *h, const RadopScan *sc)
{
    if (!h || !sc) return RADOP_ERR_ARG;
    if (!h->have_params) return h->fail(RADOP_ERR_STATE, "parameters not set");
    if (sc->n_rays <= 0 || sc->n_sub <= 0 || sc->n_traj <= 0 || sc->n_vert <= 0)
        return h->fail(RADOP_ERR_ARG, "empty scan");
    if (!sc->ray_az || !sc->ray_traj || !sc->sub_daz || !sc->sub_jv || !sc->sub_weight || !sc->traj_s || !sc->traj_h || !sc->traj_e)
        return h->fail(RADOP_ERR_ARG, "null scan array");
    for (int r = 0; r < sc->n_rays; ++r)
        if (sc->ray_traj[r] < 0 || sc->ray_traj[r] >= sc->n_traj) return h->fail(RADOP_ERR_ARG, "ray_traj out of range");
    for (int q = 0; q < sc->n_sub; ++q)
        if (sc->sub_jv[q] < 0 || sc->sub_jv[q] >= sc->n_vert) return h->fail(RADOP_ERR_ARG, "sub_jv out of range");
    CU(h, cudaSetDevice(h->device));
    const int ng = h->p.n_gates;
    const size_t ntr = (size_t)sc->n_traj * sc->n_vert * ng;
    // one pinned blob, one H2D copy
    size_t off = 0, o_az, o_tr, o_ny, o_daz, o_jv, o_w, o_s, o_h, o_e, o_sin, o_tan, o_ml;
    o_az = off; off = align_up(off + (size_t)sc->n_rays * 8, 256);
    o_tr = off; off = align_up(off + (size_t)sc->n_rays * 4, 256);
    o_ny = off; off = align_up(off + (size_t)sc->n_rays * 4, 256);
    o_daz = off; off = align_up(off + (size_t)sc->n_sub * 8, 256);
    o_jv = off; off = align_up(off + (size_t)sc->n_sub * 4, 256);
    o_w = off; off = align_up(off + (size_t)sc->n_sub * 8, 256);
    o_ml = off; off = align_up(off + (size_t)sc->n_sub * 4, 256);
    o_s = off; off = align_up(off + ntr * 4, 256);
    o_h = off; off = align_up(off + ntr * 4, 256);
    o_e = off; off = align_up(off + ntr * 4, 256);
    const bool have_trig = sc->traj_sin && sc->traj_tan;
    o_sin = off; if (have_trig) off = align_up(off + ntr * 4, 256);
    o_tan = off; if (have_trig) off = align_up(off + ntr * 4, 256);
    int rc = ensure_pinned(h, off);
    if (rc) return rc;
    CU(h, h->scan_blob.reserve(off));
    CU(h, cudaStreamSynchronize(h->stream));   // the pinned blob may still be in flight
    char *P = (char *)h->pinned;
    memcpy(P + o_az, sc->ray_az, (size_t)sc->n_rays * 8);
    memcpy(P + o_tr, sc->ray_traj, (size_t)sc->n_rays * 4);
    if (sc->ray_nyquist) memcpy(P + o_ny, sc->ray_nyquist, (size_t)sc->n_rays * 4);
    memcpy(P + o_daz, sc->sub_daz, (size_t)sc->n_sub * 8);
    memcpy(P + o_jv, sc->sub_jv, (size_t)sc->n_sub * 4);
    memcpy(P + o_w, sc->sub_weight, (size_t)sc->n_sub * 8);
    if (sc->sub_ml_only) memcpy(P + o_ml, sc->sub_ml_only, (size_t)sc->n_sub * 4);
    memcpy(P + o_s, sc->traj_s, ntr * 4);
    {   // ground-distance range of the scan: interval of the geodesy polynomials (k_sub_geopoly)
        float lo = sc->traj_s[0], hi = sc->traj_s[0];
        bool finite = true;
        for (size_t i = 0; i < ntr; ++i) {
            const float v = sc->traj_s[i];
            if (!(v - v == 0.f)) finite = false;
            lo = std::min(lo, v); hi = std::max(hi, v);
        }
        h->geo_mid = 0.5 * ((double)lo + (double)hi);
        h->geo_half = std::max(0.5 * ((double)hi - (double)lo), 1.0);
        // 0.05 rad of arc (~320 km): beyond that the truncated series is not trusted a priori
        h->geo_poly_ok = finite && (h->geo_half * 2.0 <= 0.05 * 6.371e6) && std::fabs(h->geo_mid) <= 0.05 * 6.371e6;
    }
    memcpy(P + o_h, sc->traj_h, ntr * 4);
    memcpy(P + o_e, sc->traj_e, ntr * 4);
    if (have_trig) { memcpy(P + o_sin, sc->traj_sin, ntr * 4); memcpy(P + o_tan, sc->traj_tan, ntr * 4); }
    CU(h, cudaMemcpyAsync(h->scan_blob.p, P, off, cudaMemcpyHostToDevice, h->stream));
    char *D = (char *)h->scan_blob.p;
    h->scan.ray_az = (const double *)(D + o_az);
    h->scan.ray_traj = (const int *)(D + o_tr);
    h->d_nyquist = sc->ray_nyquist ? (const float *)(D + o_ny) : nullptr;
    h->scan.sub_daz = (const double *)(D + o_daz);
    h->scan.sub_jv = (const int *)(D + o_jv);
    h->scan.sub_weight = (const double *)(D + o_w);
    h->scan.sub_ml = sc->sub_ml_only ? (const int *)(D + o_ml) : nullptr;
    h->scan.traj_s = (const float *)(D + o_s);
    h->scan.traj_h = (const float *)(D + o_h);
    h->scan.traj_e = (const float *)(D + o_e);
    h->scan.traj_sin = have_trig ? (const float *)(D + o_sin) : nullptr;
    h->scan.traj_tan = have_trig ? (const float *)(D + o_tan) : nullptr;
    h->n_rays = sc->n_rays; h->n_sub = sc->n_sub; h->n_traj = sc->n_traj; h->n_vert = sc->n_vert;
    h->center_sub = sc->n_sub / 2;          // idx_0 = int(num_beams / 2), doppler_scatter.py:89
    h->n_sbg = (long long)sc->n_rays * sc->n_sub * ng;
    h->have_scan = true;
    h->executed = false;
    h->geo_ready = false;
    h->traj_tab_ready = false;
    return RADOP_OK;
}

static int prepare(RadopHandle *h);

// Per sub-beam polynomials of the rotated-pole coordinates for the WHOLE staged ray list.  They depend on the
// scan geometry, the radar position and the grid's pole only, so they are made once per staged scan (first
// run after radop_stage_scan / radop_set_params / radop_upload_grid) and reused by every execute; the host
// learns here whether every series converged (else the per-gate solution runs).
static int ensure_geometry(RadopHandle *h)
{
    if (h->geo_ready) return RADOP_OK;
    h->geo_poly_conv = false;
    if (h->geodesy_mode == 0 && h->geo_poly_ok) {
        const int n_rq = h->n_rays * h->n_sub;
        CU(h, h->geo_poly.reserve((size_t)n_rq * sizeof(SubPoly)));
        CU(h, h->flags.reserve(16));
        CU(h, cudaMemsetAsync(h->flags.p, 0, 16, h->stream));
        DevConfig c = h->cfg;
        c.n_rays = h->n_rays;
        k_sub_geopoly<<<(n_rq + 7) / 8, 128, 0, h->stream>>>(c, h->n_rays, h->scan, h->scan.ray_az, h->scan.sub_daz,
                                                             h->geo_poly.as<SubPoly>(), h->flags.as<int>());
        CU(h, cudaGetLastError());
        int flag = 0;
        CU(h, cudaMemcpyAsync(&flag, h->flags.p, 4, cudaMemcpyDeviceToHost, h->stream));
        CU(h, cudaStreamSynchronize(h->stream));
        h->geo_poly_conv = !(flag & 4);
    }
    h->geo_ready = true;
    return RADOP_OK;
}

static int prepare(RadopHandle *h)
{
    if (!h->have_params || !h->have_grid || !h->have_scan) return h->fail(RADOP_ERR_STATE, "params, grid and scan are needed before execute");
    const RadopParams &p = h->p;
    DevConfig &c = h->cfg;
    memset(&c, 0, sizeof c);
    c.n_rays = h->n_rays; c.n_sub = h->n_sub; c.n_gates = p.n_gates; c.n_vert = h->n_vert; c.center_sub = h->center_sub;
    c.radar_lat = p.radar_lat; c.radar_lon = p.radar_lon;
    c.cos_lon0 = cos(p.radar_lon * 0.017453292519943295); c.sin_lon0 = sin(p.radar_lon * 0.017453292519943295);
    c.nz = h->nz; c.ny = h->ny; c.nx = h->nx;
    for (int i = 0; i < 2; ++i) {
        c.llc[i] = h->llc[i]; c.urc[i] = h->urc[i]; c.res[i] = h->res[i];
        c.rres[i] = 1.0f / h->res[i];            // IEEE float32 division: correctly rounded reciprocal
        if (!(fabsf(h->res[i]) > 1e-6f && fabsf(h->res[i]) < 1e6f)) return h->fail(RADOP_ERR_ARG, "grid resolution out of range");
    }
    memcpy(c.rot, h->rot, sizeof c.rot);
    c.var_mask = h->var_mask;
    c.z_strict = h->z_strict;
    {   // Is every quotient n / den of the vertical interpolation inside the range where the shared reciprocal gives
        // the IEEE quotient (div_den_safe, div_shared_rcp: 0.125 <= den <= 1e15, 1e-30 <= |q| <= 1e30 or n == 0)?
        // den runs over the spacings of adjacent levels, n over the vertical differences of each uploaded variable.
        bool ok = h->z_dpos && h->z_dmin >= 0.125f && h->z_dmax <= 1e15f;
        for (int v = 0; ok && v < RADOP_NVAR_MAX; ++v)
            if (h->var_mask & (1u << v)) {
                const double nmax = h->var_nmax[v], nmin = h->var_nmin[v];
                ok = std::isfinite(h->var_nmax[v]) && nmax / (double)h->z_dmin <= 1e29 && nmin / (double)h->z_dmax >= 1e-29;
            }
        c.div_exact = ok ? 1 : 0;
    }
    c.lev_guess = (h->search_mode == 0 && h->nz >= 8 && h->nz <= 256) ? h->lev_guess.as<uint8_t>() : nullptr;
    c.z_tb = h->z_tb.as<float2>();
    c.n_species = p.n_species;
    c.two_moment = p.two_moment; c.doppler_scheme = p.doppler_scheme; c.with_attenuation = p.with_attenuation;
    c.sensitivity_mode = p.sensitivity_mode; c.output_model_vars = p.output_model_vars; c.keep_debug = p.keep_debug;
    c.wavelength = (float)p.wavelength_mm; c.k_squared = (float)p.k_squared; c.radial_res = (float)p.radial_resolution_m;
    c.sens0 = (float)p.sensitivity[0]; c.sens1 = (float)p.sensitivity[1];
    c.n_vel = p.n_vel; c.vel_min = p.vel_min; c.vel_res = p.vel_res;
    c.inv_vel_res = p.vel_res > 0 ? 1.0 / p.vel_res : 0.0;
    c.with_melting = p.with_melting ? 1 : 0;
    c.brd_mode = 0;
    if (p.doppler_scheme == 3) c.brd_mode = (p.spectrum_turbulence ? 1 : 0) | (p.spectrum_motion ? 2 : 0);
    if (c.brd_mode) {
        // spectral_width_turb / spectral_width_motion, doppler_scatter.py:727-777
        const double A = 1.6;                              // global_constants.py: turbulence constant
        const double bw = p.beamwidth_deg * 0.017453292519943295;
        c.brd_res = p.radial_resolution_m;
        c.brd_sigma_r = 0.35 * p.radial_resolution_m;
        c.brd_sigma_theta = bw / (4.0 * sqrt(log(2.0)));
        c.brd_turb_far = c.brd_sigma_theta * pow(A, 1.5) / 0.72;
        c.brd_turb_near = c.brd_sigma_r * pow(1.35 * A, 1.5);
        c.brd_motion_num = (float)(p.wavelength_mm / 100.0 * p.antenna_speed);
        c.brd_motion_den = 2.0 * 3.141592653589793 * bw;
        if ((c.brd_mode & 1) && !(h->var_mask & (1u << RADOP_VAR_EDR)))
            return h->fail(RADOP_ERR_STATE, "turbulence broadening needs the EDR variable (RADOP_VAR_EDR)");
    }
    c.ice_numeric = -1;
    uint32_t need = (1u << RADOP_VAR_U) | (1u << RADOP_VAR_V) | (1u << RADOP_VAR_W) | (1u << RADOP_VAR_T);
    int bins = 0, nd_max = 0;
    size_t smem = 0;
    for (int s = 0; s < p.n_species; ++s) {
        const RadopSpecies &S = p.species[s];
        const LutHost &L = h->lut[S.species];
        if (!L.loaded) return h->fail(RADOP_ERR_LUT, "a lookup table is missing for one of the species");
        if (memcmp(&L.packed_for, &S, sizeof S) != 0 || L.packed_doppler != p.doppler_scheme)
            return h->fail(RADOP_ERR_LUT, "species constants changed since the table upload: upload the table again");
        SpeciesDev &d = c.sp[s];
        d.species = S.species; d.psd = S.psd; d.mode = L.mode; d.ncol = L.ncol; d.numeric_vn = S.numeric_vn;
        d.ne = L.ne; d.nt = L.nt; d.nd = L.nd;
        d.e_min = L.e_min; d.e_step = L.e_step; d.t_min = L.t_min; d.t_step = L.t_step;
        d.e_rstep = 1.0f / L.e_step; d.t_rstep = 1.0f / L.t_step;      // correctly rounded float32 reciprocals
        if (!(fabsf(L.e_step) > 1e-6f && fabsf(L.e_step) < 1e6f && fabsf(L.t_step) > 1e-6f && fabsf(L.t_step) < 1e6f))
            return h->fail(RADOP_ERR_LUT, "table axis step out of range");
        d.bin_base = bins; d.cost = L.cost;
        species_vars(S.species, &d.qm_var, &d.qn_var);
        need |= (1u << d.qm_var);
        if (S.psd == RADOP_PSD_2MOM) need |= (1u << d.qn_var);
        if (S.psd == RADOP_PSD_MELTING) need |= (1u << RADOP_VAR_QR) | (1u << RADOP_VAR_QS) | (1u << RADOP_VAR_QG);
        memcpy(d.melt, S.melt, sizeof d.melt);
        d.dD_geo = L.dD_geo; d.dD_ref = L.dD_ref; d.dD_d = L.dD_d;
        d.d_min = L.d_min; d.d_max = L.d_max; d.spec_step = L.spec_step;
        d.inv_beta = 1.0 / S.beta; d.inv_alpha = 1.0 / S.alpha;
        d.wh_lo = S.alpha * pow((double)L.d_min, S.beta) * (1.0 - 1e-9);
        d.wh_hi = S.alpha * pow((double)L.d_max, S.beta) * (1.0 + 1e-9);
        if (!(S.beta > 0.0 && S.alpha > 0.0)) { d.wh_lo = -1e300; d.wh_hi = 1e300; }   // no shortcut
        d.table = L.table.as<float>(); d.aux = L.aux.as<float>();
        d.a = S.a; d.b = S.b; d.alpha = S.alpha; d.beta = S.beta; d.mu = S.mu; d.nu = S.nu; d.N0 = S.N0;
        d.lambda_factor = S.lambda_factor; d.vel_factor = S.vel_factor; d.ntot_factor = S.ntot_factor;
        d.x_min = S.x_min; d.x_max = S.x_max;
        {   // log-domain constants (see SpeciesDev)
            const double ln1000 = 6.907755278982137;
            d.c_l0 = 0; d.c_inv = 0; d.c_q = 0; d.e_q = 0; d.c_ka = 0; d.c_kb = 0;
            if (S.psd == RADOP_PSD_1MOM_FIXED_N0) { d.c_l0 = log(S.lambda_factor); d.c_inv = 1.0 / (4.0 + S.mu); }
            else if (S.psd == RADOP_PSD_1MOM_SNOW) { d.c_l0 = log(S.a * S.lambda_factor); d.c_inv = 1.0 / (S.b + 1.0); }
            else if (S.psd == RADOP_PSD_1MOM_ICE) { d.c_l0 = log(S.b); d.c_inv = 1.0 / (S.b - 2.0); }
            else {
                d.c_inv = -S.nu / S.b;
                d.c_q = S.nu / S.ntot_factor; d.e_q = (S.mu + 1.0) / S.nu;
                d.c_ka = (1.0 + S.mu) * ln1000; d.c_kb = S.nu * ln1000;
            }
            d.c_v = S.vel_factor * S.alpha / S.nu; d.e_v = -(S.beta + S.mu + 1.0) / S.nu;
            d.c_n = S.ntot_factor / S.nu; d.e_n = -(S.mu + 1.0) / S.nu;
        }
        if (S.numeric_vn && p.doppler_scheme == 1) {
            if (c.ice_numeric >= 0) return h->fail(RADOP_ERR_UNSUPPORTED, "only one species may have numeric fall-speed moments");
            c.ice_numeric = s;
        }
        if (p.doppler_scheme == 3 && (L.nd % 128) != 0)
            return h->fail(RADOP_ERR_LUT, "doppler scheme 3 needs a multiple of 128 diameter bins");
        nd_max = std::max(nd_max, L.nd);
        bins += L.ne * L.nt;
        smem = std::max(smem, (size_t)L.nd * L.ncol * 4 + (size_t)L.nd * 4);
    }
    if ((h->var_mask & need) != need) return h->fail(RADOP_ERR_STATE, "a model variable needed by the configuration was not uploaded");
    if (bins >= 0xFFFF) return h->fail(RADOP_ERR_LUT, "too many (species, e, t) bins");
    if (bins * 4 > 200 * 1024) return h->fail(RADOP_ERR_LUT, "bin histogram does not fit shared memory");
    c.total_bins = bins;
    h->k2_smem = smem;
    // 3 staged columns + per warp: local prefix sums (float), 4 bases per lane (double), edge diameters
    h->k4_smem = (size_t)3 * nd_max * 4 + (size_t)SPEC_WARPS * (((((size_t)(nd_max >> 5) * 33 + 1) & ~(size_t)1)) * 4 + 128 * 8 + (size_t)((std::max(p.n_vel, 1) + 1) & ~1) * 4);
    return RADOP_OK;
}

// Device -> host copies of the rays [r0, r0 + nr) of every requested output, on stream st.
// Rows [r0, r0 + nr) of a device array [n_rays][row_bytes] to the caller's host array.  Normally the host array has the
// same layout; with radop_set_output_blocks(block, pitch) the caller's array interleaves the rays of several handles
// (the ranks of a multi-GPU run dealing blocks of `block` rays in turn): local ray i lands on host row
// (i / block) * pitch + i % block, whole blocks as ONE strided copy (cudaMemcpy2DAsync: the copy engine does the striding).
static cudaError_t copy_rows(RadopHandle *h, void *dst_, const void *src_, size_t row_bytes, int r0, int nr, cudaStream_t st)
{
    char *dst = static_cast<char *>(dst_);
    const char *src = static_cast<const char *>(src_);
    const int B = h->out_block;
    if (B <= 0) return cudaMemcpyAsync(dst + (size_t)r0 * row_bytes, src + (size_t)r0 * row_bytes, (size_t)nr * row_bytes, cudaMemcpyDeviceToHost, st);
    const size_t P = (size_t)h->out_pitch;
    int r = r0;
    const int end = r0 + nr;
    cudaError_t e = cudaSuccess;
    if (r % B && r < end) {               // up to the next block boundary
        const int n = std::min(B - r % B, end - r);
        e = cudaMemcpyAsync(dst + ((size_t)(r / B) * P + r % B) * row_bytes, src + (size_t)r * row_bytes, (size_t)n * row_bytes, cudaMemcpyDeviceToHost, st);
        if (e != cudaSuccess) return e;
        r += n;
    }
    const int full = (end - r) / B;
    if (full > 0) {
        e = cudaMemcpy2DAsync(dst + (size_t)(r / B) * P * row_bytes, P * row_bytes, src + (size_t)r * row_bytes, (size_t)B * row_bytes,
                              (size_t)B * row_bytes, (size_t)full, cudaMemcpyDeviceToHost, st);
        if (e != cudaSuccess) return e;
        r += full * B;
    }
    if (r < end)
        e = cudaMemcpyAsync(dst + (size_t)(r / B) * P * row_bytes, src + (size_t)r * row_bytes, (size_t)(end - r) * row_bytes, cudaMemcpyDeviceToHost, st);
    return e;
}

static int copy_outputs(RadopHandle *h, const RadopOut *out, int r0, int nr, cudaStream_t st)
{
    const int ng = h->cfg.n_gates;
    const size_t n_rg = (size_t)h->n_rays * ng, row4 = (size_t)ng * 4, row8 = (size_t)ng * 8;
    // Observables and mask are planes of ONE device array.  When the caller's arrays are equally spaced as well (one host
    // allocation, as the Python binding makes it) and rows are contiguous, a chunk of all planes is ONE strided copy
    // instead of eleven: every copy costs the copy engine ~8 us of latency, 1 ms per volume at twelve chunks.
    bool merged = false;
    if (h->out_block <= 0 && out->mask) {
        bool regular = true;
        for (int v = 0; v < RADOP_N_OUT; ++v) regular = regular && out->obs[v] != nullptr;
        const ptrdiff_t dp = regular ? (const char *)out->obs[1] - (const char *)out->obs[0] : 0;
        regular = regular && dp >= (ptrdiff_t)(n_rg * 4) && dp < (ptrdiff_t)0x7fffffff;
        for (int v = 1; regular && v < RADOP_N_OUT; ++v) regular = (const char *)out->obs[v] - (const char *)out->obs[v - 1] == dp;
        regular = regular && (const char *)out->mask - (const char *)out->obs[RADOP_N_OUT - 1] == dp && n_rg * 4 < 0x7fffffffull;
        if (regular) {
            CU(h, cudaMemcpy2DAsync((char *)out->obs[0] + (size_t)r0 * row4, (size_t)dp, (const char *)h->out.p + (size_t)r0 * row4, n_rg * 4,
                                    (size_t)nr * row4, RADOP_N_OUT + 1, cudaMemcpyDeviceToHost, st));
            merged = true;
        }
    }
    if (!merged) {
        for (int v = 0; v < RADOP_N_OUT; ++v)
            if (out->obs[v]) CU(h, copy_rows(h, out->obs[v], h->out.as<float>() + (size_t)v * n_rg, row4, r0, nr, st));
        if (out->mask) CU(h, copy_rows(h, out->mask, h->out.as<float>() + (size_t)RADOP_N_OUT * n_rg, row4, r0, nr, st));
    }
    if ((out->lat || out->lon) && !h->latlon_valid) return h->fail(RADOP_ERR_STATE, "lat / lon were not computed by the last run");
    if (out->lat) CU(h, copy_rows(h, out->lat, h->lat.as<double>(), row8, r0, nr, st));
    if (out->lon) CU(h, copy_rows(h, out->lon, h->lon.as<double>(), row8, r0, nr, st));
    if (out->dspectrum) {
        if (h->cfg.doppler_scheme != 3) return h->fail(RADOP_ERR_ARG, "DSPECTRUM requested but the Doppler scheme is not 3");
        CU(h, copy_rows(h, out->dspectrum, h->spec_out.as<float>(), row4 * (size_t)h->cfg.n_vel, r0, nr, st));
    }
    for (int v = 0; v < RADOP_NVAR_MAX; ++v)
        if (out->model[v]) {
            if (!h->cfg.output_model_vars || !((h->var_mask | (h->cfg.with_melting ? 0x1E00u : 0u)) & (1u << v)))
                return h->fail(RADOP_ERR_ARG, "model output requested but not computed");
            CU(h, copy_rows(h, out->model[v], h->model_out.as<float>() + (size_t)v * n_rg, row4, r0, nr, st));
        }
    return RADOP_OK;
}

// The launch sequence of the hot path.  The ray list is processed in chunks of about
// h->chunk_sbg sub-beam gates: the work arrays are sized for one chunk, the outputs for the
// whole scan.  With `out` given, the device -> host copies of a chunk run on a second stream
// while the next chunk computes.
static int run_scan(RadopHandle *h, const RadopOut *out)
{
    int rc = prepare(h);
    if (rc) return rc;
    CU(h, cudaSetDevice(h->device));
    if (h->integration_mode == 0) {
        rc = build_lamtabs(h);
        if (rc) return rc;
    }
    DevConfig &c = h->cfg;
    c.sum_species = (h->integration_mode == 0 && !c.with_melting && c.doppler_scheme != 3) ? 1 : 0;
    const int nsp = c.n_species, ng = c.n_gates;
    const int n_rays = h->n_rays;
    const bool ml = h->scan.sub_ml != nullptr;
    if (ml && !c.with_melting)
        // the reference reads subradial.mask_ml, which only melting() sets (interpolation.py:429; AttributeError there)
        return h->fail(RADOP_ERR_ARG, "integration scheme 'ml' needs the melting scheme (with_melting = 1)");
    if (ml && c.doppler_scheme == 3) return h->fail(RADOP_ERR_UNSUPPORTED, "integration scheme 'ml' with Doppler scheme 3 is not implemented");
    rc = ensure_geometry(h);
    if (rc) return rc;
    const bool poly = h->geodesy_mode == 0 && h->geo_poly_ok && h->geo_poly_conv;
    h->geo_poly_launched = poly;
    bool fused = h->pipeline_mode == 0 && h->integration_mode == 0 && !c.with_melting && c.doppler_scheme != 3 &&
                 !c.keep_debug && !c.output_model_vars && poly;
    for (int s = 0; s < nsp; ++s) fused = fused && c.sp[s].ne <= 255 && c.sp[s].mode != MODE_MELT;
    h->last_fused = fused;
    const size_t n_rg = (size_t)n_rays * ng;
    const bool want_vals = c.output_model_vars || c.keep_debug;
    // ---- chunking --------------------------------------------------------------------------
    const long long per_ray = (long long)c.n_sub * ng;
    int n_chunks = 1;
    // without host copies to overlap, chunks only bound the work memory (8x larger chunks)
    // bytes radop_simulate copies back: chunking only pays when there is a copy worth hiding (every chunk
    // costs a pass over the bin table and partial tiles; a dense-quadrature RHI returns 2 MB for 2e7 gates)
    size_t d2h_bytes = 0;
    if (out) {
        for (int v = 0; v < RADOP_N_OUT; ++v) if (out->obs[v]) d2h_bytes += n_rg * 4;
        if (out->mask) d2h_bytes += n_rg * 4;
        if (out->lat) d2h_bytes += n_rg * 8;
        if (out->lon) d2h_bytes += n_rg * 8;
        if (out->dspectrum) d2h_bytes += n_rg * (size_t)c.n_vel * 4;
        for (int v = 0; v < RADOP_NVAR_MAX; ++v) if (out->model[v]) d2h_bytes += n_rg * 4;
    }
    const bool overlap_copies = out && d2h_bytes >= h->overlap_min_bytes;
    // default chunk: 3e6 sub-beam gates for the fused kernel (a chunk costs it little, and many short copies keep the copy
    // engine busy from early on: 4.75 ms end to end at cfg 3 against 5.22 ms with 8e6, scripts/chunk_sweep.py), 8e6 for the
    // staged pipeline (every chunk costs it a pass over the bin table and partial tiles)
    const long long chunk_sbg = h->chunk_sbg >= 0 ? h->chunk_sbg : (fused ? 3000000LL : 8000000LL);
    long long target = overlap_copies ? chunk_sbg : 8 * (h->chunk_sbg >= 0 ? h->chunk_sbg : 8000000LL);
    if (c.brd_mode) {
        // the sub-beam spectra of a chunk are kept in float64: at most ~4 GB of them
        const long long cap = std::max<long long>(per_ray, (4LL << 30) / ((long long)c.n_vel * 8));
        target = target > 0 ? std::min(target, cap) : cap;
    }
    if (!c.keep_debug && target > 0)
        n_chunks = (int)std::min<long long>(n_rays, std::max<long long>(1, (h->n_sbg + target / 2) / target));
    // Chunk boundaries.  With host copies to overlap (radop_simulate) the chunks shrink linearly towards
    // the end: the copy of the last chunk is the only one nothing can hide, so it should be short.
    std::vector<int> chunk_r0(1, 0);
    {
        const int even = (n_rays + n_chunks - 1) / n_chunks;
        if (overlap_copies && n_chunks >= 3) {
            const double unit = (double)n_rays / ((double)n_chunks * n_chunks);     // sum of (2 (n - i) - 1) = n^2
            int r = 0;
            for (int i = 0; i < n_chunks - 1 && r < n_rays; ++i) {
                r = std::min(n_rays, r + std::max(1, (int)(unit * (2 * (n_chunks - i) - 1) + 0.5)));
                if (r < n_rays) chunk_r0.push_back(r);
            }
        } else {
            for (int r = even; r < n_rays; r += even) chunk_r0.push_back(r);
        }
        n_chunks = (int)chunk_r0.size();
        chunk_r0.push_back(n_rays);
    }
    int rays_per_chunk = 1;            // rays of the largest chunk: the work memory is sized for it
    for (int i = 0; i < n_chunks; ++i) rays_per_chunk = std::max(rays_per_chunk, chunk_r0[i + 1] - chunk_r0[i]);
    const long long n = (long long)rays_per_chunk * per_ray;          // sub-beam gates of the largest chunk
    const size_t c_rg = (size_t)rays_per_chunk * ng;
    if (n * nsp >= 0x7fffffffLL || n >= 0xfffffff0LL)
        return h->fail(RADOP_ERR_ARG, "chunk too large (rays x sub-beams x gates x species >= 2^31): lower the chunk size");
    c.geo_mid = h->geo_mid; c.geo_rhalf = 1.0 / h->geo_half;
    const int n_spans = (ng + 31) / 32;
    const long long traj_n = (long long)h->n_traj * h->n_vert * ng;
    if (fused) {
        CU(h, h->traj_sc.reserve((size_t)traj_n * 8));
        CU(h, h->traj_ie.reserve((size_t)traj_n * nsp));
        CU(h, h->f_wsum.reserve(c_rg * 4));
        CU(h, h->f_cnt.reserve(RADOP_NSPECIES_MAX * 8));
        if (c.ice_numeric >= 0 && c.doppler_scheme == 1) CU(h, h->f_ice.reserve((size_t)rays_per_chunk * c.n_sub * n_spans * 32));
    } else {
    if (!poly) CU(h, h->geo.reserve((size_t)rays_per_chunk * c.n_sub * sizeof(SubGeo)));
    CU(h, h->mask.reserve(n));
    CU(h, h->active.reserve(n));
    CU(h, h->wind.reserve(n * 4));
    CU(h, h->sinth.reserve(n * 4));
    CU(h, h->vsum.reserve(n * 4));
    CU(h, h->nsum.reserve(n * 4));
    CU(h, h->key.reserve((size_t)n * nsp * 2));
    CU(h, h->p0.reserve((size_t)n * nsp * 4));
    CU(h, h->p1.reserve((size_t)n * nsp * 4));
    if (want_vals) CU(h, h->vals.reserve((size_t)n * RADOP_NVP * 4));
    if (ml) { CU(h, h->ml_wraw.reserve((size_t)n * 4)); CU(h, h->ml_wsz.reserve((size_t)n * 4)); }
    }
    CU(h, h->lat.reserve(n_rg * 8));
    CU(h, h->lon.reserve(n_rg * 8));
    CU(h, h->hist.reserve((size_t)c.total_bins * 4));
    CU(h, h->flags.reserve(16));
    CU(h, h->bin_start.reserve((size_t)(c.total_bins + 1) * 4));
    CU(h, h->tile_start.reserve((size_t)(c.total_bins + 1) * 4));
    CU(h, h->cost_start.reserve((size_t)(c.total_bins + 1) * 8));
    CU(h, h->cursor.reserve((size_t)c.total_bins * 4));
    CU(h, h->totals.reserve((size_t)n_chunks * RADOP_TOTALS * 8));
    if (!fused) {
        if (h->integration_mode != 0 || c.with_melting || c.doppler_scheme == 3) CU(h, h->rec.reserve((size_t)n * nsp * sizeof(PairRec)));
        CU(h, h->sz_out.reserve((size_t)n * (c.sum_species ? 1 : nsp) * 8 * 4));
        CU(h, h->ice_vn.reserve((size_t)n * 8));
        if (c.doppler_scheme == 2 || c.with_melting) CU(h, h->dop_vn.reserve((size_t)n * nsp * 8));
        CU(h, h->ice_tot.reserve((size_t)rays_per_chunk * c.n_sub * 3 * 4));
    }
    if (c.doppler_scheme == 3) {
        CU(h, h->d3_uh.reserve((size_t)n * 8));
        CU(h, h->d3_w.reserve((size_t)n * 4));
        CU(h, h->d3_rho.reserve((size_t)n * 4));
        CU(h, h->d3_st.reserve((size_t)n * 8));
        CU(h, h->attw.reserve((size_t)n * 4));
        CU(h, h->spec_pre.reserve((size_t)n * sizeof(SpecPre)));
        CU(h, h->spec_acc.reserve(c_rg * c.n_vel * 8));
        CU(h, h->spec_out.reserve(n_rg * c.n_vel * 4));
        if (c.brd_mode) {
            CU(h, h->d3_std.reserve((size_t)n * 8));
            CU(h, h->sub_acc.reserve((size_t)n * c.n_vel * 8));
            CU(h, h->brd_flag.reserve((size_t)rays_per_chunk * c.n_sub));
        }
    }
    CU(h, h->comb.reserve(c_rg * 8 * 4));
    CU(h, h->out.reserve(n_rg * (RADOP_N_OUT + 1) * 4));
    if (c.output_model_vars) CU(h, h->model_out.reserve(n_rg * RADOP_NVP * 4));
    if (c.keep_debug) {
        CU(h, h->dbg_cell.reserve((size_t)n * 2 * 4));
        CU(h, h->dbg_level.reserve((size_t)n * 4 * 4));
        CU(h, h->dbg_status.reserve((size_t)n * 4));
        CU(h, h->dbg_e.reserve((size_t)n * nsp * 4));
        CU(h, h->dbg_t.reserve((size_t)n * nsp * 4));
        CU(h, h->dbg_rot.reserve((size_t)n * 2 * 8));
    }
    static bool attr_done[64] = {false};
    if (!attr_done[h->device & 63]) {
        CU(h, cudaFuncSetAttribute(k_integrate, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(h, cudaFuncSetAttribute(k_interp_psd<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(h, cudaFuncSetAttribute(k_interp_psd<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(h, cudaFuncSetAttribute(k_scatter_pairs, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(h, cudaFuncSetAttribute(k_spectrum<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(h, cudaFuncSetAttribute(k_spectrum<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(h, cudaFuncSetAttribute(k_beam_attenuation, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU(h, cudaFuncSetAttribute(k_spectrum_broaden, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_done[h->device & 63] = true;
    }
    int occ = 1;
    CU(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_integrate, RADOP_TILE_THREADS, h->k2_smem));
    h->k2_blocks = std::max(1, occ) * h->n_sm;
    const size_t hist_smem = (size_t)c.total_bins * 4;
    int occ1 = 1;
    CU(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ1, k_interp_psd<true>, 128, hist_smem));
    const size_t smem_a = (size_t)4 * 2 * ng * 4;
    const size_t smem_b = (size_t)4 * ((size_t)BRD_WMAX * 8 + (size_t)((std::max(c.n_vel, 1) + 1) & ~1) * 8);
    int occ4 = 1;
    if (c.doppler_scheme == 3) {
        if (smem_a > 200 * 1024) return h->fail(RADOP_ERR_ARG, "too many gates per ray for the spectrum attenuation scan");
        if (h->k4_smem > 200 * 1024) return h->fail(RADOP_ERR_ARG, "velocity array / diameter axis too long for the spectrum kernel");
        CU(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ4, k_spectrum<true>, 32 * SPEC_WARPS, h->k4_smem));
    }
    while ((int)h->chunk_done.size() < n_chunks) {
        cudaEvent_t e;
        CU(h, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        h->chunk_done.push_back(e);
    }

    cudaStream_t st = h->stream;
    int launches = 0;
    const bool timed = h->timing && h->n_timed + n_chunks <= 16384;
    if (timed) ++h->n_exec_timed;
    CU(h, cudaMemsetAsync(h->flags.p, 0, 4, st));
    if (h->integration_mode == 0) CU(h, cudaMemsetAsync(h->lt_stat.p, 0, 32, st));
    c.out_stride = (long long)n_rg;

    for (int ch = 0; ch < n_chunks; ++ch) {
        const int r0 = chunk_r0[ch], nr = chunk_r0[ch + 1] - r0;
        c.n_rays = nr;
        const long long nc = (long long)nr * per_ray;
        const size_t cc_rg = (size_t)nr * ng, off_rg = (size_t)r0 * ng;
        cudaEvent_t *ev = nullptr;
        if (timed) {
            while ((int)h->ev.size() < 6 * (h->n_timed + 1)) {
                cudaEvent_t e;
                CU(h, cudaEventCreate(&e));
                h->ev.push_back(e);
            }
            ev = h->ev.data() + 6 * h->n_timed;
            ++h->n_timed;
        }
        if (ev) CU(h, cudaEventRecord(ev[0], st));
        if (!fused) CU(h, cudaMemsetAsync(h->hist.p, 0, (size_t)c.total_bins * 4, st));
        CU(h, cudaMemsetAsync(h->flags.as<int>() + 1, 0, 12, st));
        ScanDev scan = h->scan;
        scan.ray_az += r0;
        scan.ray_traj += r0;
        const float *nyq = h->d_nyquist ? h->d_nyquist + r0 : nullptr;
        long long *totals = h->totals.as<long long>() + RADOP_TOTALS * ch;

        K1Out o;
        memset(&o, 0, sizeof o);
        o.mask = h->mask.as<int8_t>(); o.active = h->active.as<uint8_t>();
        o.wind = h->wind.as<float>(); o.sinth = h->sinth.as<float>(); o.vsum = h->vsum.as<float>(); o.nsum = h->nsum.as<float>();
        o.key = h->key.as<uint16_t>(); o.p0 = h->p0.as<float>(); o.p1 = h->p1.as<float>();
        o.vals = want_vals ? h->vals.as<float>() : nullptr;
        // the positions depend on the scan geometry only: a caller that kept them from an earlier call on the
        // same ray list (radop_simulate with lat = lon = NULL) does not pay for them again
        const bool want_latlon = !out || out->lat || out->lon;
        o.lat = want_latlon ? h->lat.as<double>() + off_rg : nullptr;
        o.lon = want_latlon ? h->lon.as<double>() + off_rg : nullptr;
        h->latlon_valid = want_latlon;
        o.hist = h->hist.as<int>(); o.flags = h->flags.as<int>();
        if (c.doppler_scheme == 3) {
            o.d3_uh = h->d3_uh.as<double>(); o.d3_w = h->d3_w.as<float>(); o.d3_rho = h->d3_rho.as<float>();
            o.d3_st = h->d3_st.as<float2>();
            if (c.brd_mode) o.d3_std = h->d3_std.as<double>();
        }
        if (c.keep_debug) {
            o.dbg_cell = h->dbg_cell.as<int>(); o.dbg_level = h->dbg_level.as<int>(); o.dbg_status = h->dbg_status.as<int>();
            o.dbg_e = h->dbg_e.as<int>(); o.dbg_t = h->dbg_t.as<int>(); o.dbg_rot = h->dbg_rot.as<double>();
        }
        if (fused) {
            // ---- fused pipeline: one kernel from the ray list to the per ray-gate sums -------------------
            FusedArgs fa;
            memset(&fa, 0, sizeof fa);
            fa.poly = h->geo_poly.as<SubPoly>() + (size_t)r0 * c.n_sub;
            fa.traj_sc = h->traj_sc.as<float2>(); fa.traj_ie = h->traj_ie.as<uint8_t>(); fa.traj_n = traj_n;
            fa.comb = h->comb.as<float>(); fa.wsum = h->f_wsum.as<float>();
            fa.mask_out = h->out.as<float>() + (size_t)RADOP_N_OUT * n_rg + off_rg;
            fa.lat = o.lat; fa.lon = o.lon;
            const bool ice = c.ice_numeric >= 0 && c.doppler_scheme == 1;
            fa.ice_part = ice ? h->f_ice.as<float4>() : nullptr;
            fa.flags = h->flags.as<int>();
            fa.n_fallback = h->lt_stat.as<unsigned long long>();
            fa.n_pairs = h->f_cnt.as<unsigned long long>();
            fa.n_spans = n_spans;
            const double log2e = 1.4426950408889634;
            for (int s = 0; s < nsp; ++s) {
                const SpeciesDev &d = c.sp[s];
                FusedSpecies &f = fa.fs[s];
                f.u0 = d.c_l0 * log2e; f.uk = d.c_inv;
                f.na = log2e * (8.939515416590824 + 0.107 * 273.15); f.nb = -0.107 * log2e;
                f.n0 = (float)d.N0;
                f.cv = (float)d.c_v; f.cn = (float)d.c_n; f.ev = (float)d.e_v; f.en = (float)d.e_n;
            }
            if (ch == 0) {
                CU(h, cudaMemsetAsync(h->f_cnt.p, 0, RADOP_NSPECIES_MAX * 8, st));
                if (!h->traj_tab_ready) {      // they depend on the staged trajectories and the tables' elevation axes only
                    k_traj_tables<<<(unsigned)((traj_n + 255) / 256), 256, 0, st>>>(c, h->scan.traj_e, traj_n, h->traj_sc.as<float2>(),
                                                                                   h->traj_ie.as<uint8_t>());
                    CU(h, cudaGetLastError()); ++launches;
                    h->traj_tab_ready = true;
                }
            }
            // warps per CTA = sub-beams of a ray (at most 12), else the count between 6 and 12 that leaves the fewest idle
            int W = h->fused_warps > 0 ? std::min(std::min(h->fused_warps, c.n_sub), FUSED_MAX_WARPS) : 0;
            if (W == 0) {
                // one warp per sub-beam up to 10 sub-beams (3 x 3 quadrature: 9 warps, measured 4.08 ms against 4.46 ms
                // with 3 warps x 3 rounds); beyond, the count between 5 and 10 that leaves the fewest warps idle in the last
                // round over the sub-beams (841: 10 x 85 with 9 idle slots), the larger CTA winning ties
                W = std::min(c.n_sub, FUSED_MAX_WARPS);
                if (c.n_sub > FUSED_MAX_WARPS) {
                    int waste = 1 << 30;
                    for (int w = 5; w <= FUSED_MAX_WARPS; ++w) {
                        const int ws = (c.n_sub + w - 1) / w * w - c.n_sub;
                        if (ws * 64 / c.n_sub <= waste) { waste = ws * 64 / c.n_sub; W = w; }
                    }
                }
            }
            const size_t smemf = (size_t)(FUSED_DBUF ? 2 : 1) * W * FUSED_NPART * 32 * 4;
            typedef void (*FusedKernel)(const DevConfig, const ScanDev, const float *, const float *, const FusedArgs);
            const bool dop2 = c.doppler_scheme == 2;
            FusedKernel fk = c.two_moment ? (dop2 ? (FusedKernel)k_gate_fused<true, true> : (FusedKernel)k_gate_fused<true, false>)
                                          : (dop2 ? (FusedKernel)k_gate_fused<false, true> : (FusedKernel)k_gate_fused<false, false>);
            {   // shared-memory carveout: what the CTAs that the registers allow need for their rows, the rest stays L1
                static int carve_set[64][4] = {{0}};
                int &cur = carve_set[h->device & 63][(c.two_moment ? 2 : 0) + (dop2 ? 1 : 0)];
                cudaFuncAttributes fattr;
                CU(h, cudaFuncGetAttributes(&fattr, fk));
                const int by_regs = 65536 / std::max(1, ((fattr.numRegs + 7) / 8 * 8) * 32 * W), by_threads = 2048 / (32 * W);
                const size_t need = (size_t)std::max(1, std::min(by_regs, by_threads)) * (smemf + fattr.sharedSizeBytes + 1024);
                const int carve = (int)std::min<size_t>(100, (need * 100 + 228 * 1024 - 1) / (228 * 1024) + 1);
                if (carve != cur) {
                    CU(h, cudaFuncSetAttribute(fk, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
                    cur = carve;
                }
            }
            int occf = 1;
            CU(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occf, fk, 32 * W, smemf));
            const long long pieces = (long long)nr * n_spans;
            const unsigned gridf = (unsigned)std::min<long long>(pieces, (long long)std::max(1, occf) * h->n_sm);
            fk<<<gridf, 32 * W, smemf, st>>>(c, scan, h->z_t.as<float>(), h->fields.as<float>(), fa);
            CU(h, cudaGetLastError()); ++launches;
            if (ev) { CU(h, cudaEventRecord(ev[1], st)); CU(h, cudaEventRecord(ev[2], st)); CU(h, cudaEventRecord(ev[3], st)); }
            K3Args a3;
            memset(&a3, 0, sizeof a3);
            a3.sub_weight = scan.sub_weight; a3.ray_nyquist = nyq;
            a3.comb = h->comb.as<float>(); a3.wsum = h->f_wsum.as<float>(); a3.out = h->out.as<float>() + off_rg;
            a3.ice_part = ice ? h->f_ice.as<float4>() : nullptr; a3.n_spans = n_spans;
            k_ray_epilogue<<<(nr * 32 + 127) / 128, 128, 0, st>>>(c, a3);
            CU(h, cudaGetLastError()); ++launches;
            if (ev) { CU(h, cudaEventRecord(ev[4], st)); CU(h, cudaEventRecord(ev[5], st)); }
            if (out) {
                CU(h, cudaEventRecord(h->chunk_done[ch], st));
                CU(h, cudaStreamWaitEvent(h->copy_stream, h->chunk_done[ch], 0));
                rc = copy_outputs(h, out, r0, nr, h->copy_stream);
                if (rc) return rc;
            }
            continue;
        }
        {
            const int n_rq = nr * c.n_sub;
            const long long want = (nc + 127) / 128;
            const unsigned grid1 = (unsigned)std::min<long long>(want, (long long)std::max(1, occ1) * h->n_sm);
            const float *zt = h->z_t.as<float>(), *fl = h->fields.as<float>();
            if (poly) {
                // per sub-beam polynomials of the rotated-pole coordinates, made once per staged scan (ensure_geometry)
                k_interp_psd<true><<<grid1, 128, hist_smem, st>>>(c, scan, (const void *)(h->geo_poly.as<SubPoly>() + (size_t)r0 * c.n_sub),
                                                                  zt, fl, o, nc, 0);
                CU(h, cudaGetLastError()); ++launches;
            } else {
                k_sub_geometry<<<(n_rq + 127) / 128, 128, 0, st>>>(nr, c.n_sub, c.radar_lat, scan.ray_az, scan.sub_daz, h->geo.as<SubGeo>());
                CU(h, cudaGetLastError()); ++launches;
                k_interp_psd<false><<<grid1, 128, hist_smem, st>>>(c, scan, h->geo.p, zt, fl, o, nc, 0);
                CU(h, cudaGetLastError()); ++launches;
            }
        }
        const float *wraw = nullptr, *wsz = nullptr;
        if (ml) {
            MlKernel mk;
            double sum = 0.0;
            for (int k = -8; k <= 8; ++k) sum += exp(-0.5 * k * k / 4.0);
            for (int k = 0; k <= 8; ++k) mk.gk[k] = exp(-0.5 * k * k / 4.0) / sum;
            k_ml_weights<<<(nr * c.n_sub * 32 + 127) / 128, 128, 0, st>>>(c, scan, mk, h->active.as<uint8_t>(), h->key.as<uint16_t>(),
                                                                       h->hist.as<int>(), h->ml_wraw.as<float>(), nc);
            CU(h, cudaGetLastError()); ++launches;
            k_ml_total<<<(unsigned)((cc_rg + 127) / 128), 128, 0, st>>>(c, h->ml_wraw.as<float>(), h->ml_wsz.as<float>());
            CU(h, cudaGetLastError()); ++launches;
            wraw = h->ml_wraw.as<float>(); wsz = h->ml_wsz.as<float>();
        }
        if (ev) CU(h, cudaEventRecord(ev[1], st));

        k_scan_bins<<<1, 1024, 0, st>>>(c, h->hist.as<int>(), h->bin_start.as<int>(), h->tile_start.as<int>(),
                                         h->cost_start.as<int>(), h->cursor.as<int>(), totals);
        CU(h, cudaGetLastError()); ++launches;
        // the sorted work list serves the direct kernel, the melting kernel and the spectrum kernel;
        // the table kernel reads the pairs in ray order straight from K1's outputs
        const bool table_mode = h->integration_mode == 0;
        if (!table_mode || c.with_melting || c.doppler_scheme == 3) {
            k_scatter_pairs<<<(unsigned)((nc + 255) / 256), 256, hist_smem, st>>>(c, h->key.as<uint16_t>(), h->p0.as<float>(), h->p1.as<float>(),
                                                                                   scan.sub_weight, h->cursor.as<int>(), h->rec.as<PairRec>(), nc, wsz);
            CU(h, cudaGetLastError()); ++launches;
        }
        if (ev) CU(h, cudaEventRecord(ev[2], st));

        K2Args a2;
        a2.hist = h->hist.as<int>(); a2.bin_start = h->bin_start.as<int>(); a2.tile_start = h->tile_start.as<int>();
        a2.chunk_start = h->cost_start.as<int>(); a2.totals = totals;
        a2.counter = h->flags.as<int>() + 1;
        a2.rec = h->rec.as<PairRec>(); a2.n_sbg = nc; a2.sz_out = h->sz_out.as<float>(); a2.ice_vn = h->ice_vn.as<float2>(); a2.dop_vn = h->dop_vn.as<float2>();
        if (table_mode) {
            K2tArgs at;
            at.key = h->key.as<uint16_t>(); at.p0 = h->p0.as<float>(); at.p1 = h->p1.as<float>();
            at.sub_weight = scan.sub_weight; at.wsz = wsz; at.sz_out = a2.sz_out; at.ice_vn = a2.ice_vn; at.dop_vn = a2.dop_vn;
            at.n_sbg = nc; at.n_fallback = h->lt_stat.as<unsigned long long>();
            // 4 CTAs of 256 per SM (64 registers), 8 waves of CTAs; measured against 2, 3 and 6 CTAs per SM and
            // against 4 ... all-at-once waves (profiles/README.md)
            const unsigned gridt = (unsigned)std::min<long long>((nc + 255) / 256, (long long)h->n_sm * 32);
            k_integrate_tab<4><<<gridt, 256, 0, st>>>(c, at);
        } else {
            k_integrate<<<h->k2_blocks, RADOP_TILE_THREADS, h->k2_smem, st>>>(c, a2);
        }
        if (c.with_melting) {
            K2mArgs am;
            am.rec = h->rec.as<PairRec>(); am.key = h->key.as<uint16_t>(); am.bin_start = h->bin_start.as<int>();
            am.sz_out = h->sz_out.as<float>(); am.dop_vn = h->dop_vn.as<float2>(); am.n_sbg = nc;
            k_integrate_melting<<<h->n_sm * 8, 128, 0, st>>>(c, am);
            CU(h, cudaGetLastError()); ++launches;
        }
        CU(h, cudaGetLastError()); ++launches;
        if (ev) CU(h, cudaEventRecord(ev[3], st));

        if (c.ice_numeric >= 0 && c.doppler_scheme == 1) {
            int warps = nr * c.n_sub;
            k_ice_totals<<<(warps * 32 + 127) / 128, 128, 0, st>>>(c, h->active.as<uint8_t>(), h->ice_vn.as<float2>(), h->ice_tot.as<float>());
            CU(h, cudaGetLastError()); ++launches;
        }
        K3Args a3;
        a3.mask = h->mask.as<int8_t>(); a3.active = h->active.as<uint8_t>();
        a3.wind = h->wind.as<float>(); a3.sinth = h->sinth.as<float>(); a3.vsum = h->vsum.as<float>(); a3.nsum = h->nsum.as<float>();
        a3.sz_out = h->sz_out.as<float>(); a3.ice_tot = h->ice_tot.as<float>(); a3.dop_vn = h->dop_vn.as<float2>();
        a3.vals = want_vals ? h->vals.as<float>() : nullptr;
        a3.sub_weight = scan.sub_weight; a3.ray_nyquist = nyq;
        a3.comb = h->comb.as<float>(); a3.wsum = nullptr; a3.out = h->out.as<float>() + off_rg;
        a3.ice_part = nullptr; a3.n_spans = 0;
        a3.wraw = wraw; a3.wsz = wsz;
        a3.model_out = c.output_model_vars ? h->model_out.as<float>() + off_rg : nullptr;
        k_reduce_subbeams<<<(unsigned)((cc_rg + 127) / 128), 128, 0, st>>>(c, a3);
        CU(h, cudaGetLastError()); ++launches;
        k_ray_epilogue<<<(nr * 32 + 127) / 128, 128, 0, st>>>(c, a3);
        CU(h, cudaGetLastError()); ++launches;
        if (ev) CU(h, cudaEventRecord(ev[4], st));
        if (c.doppler_scheme == 3) {
            K4Args a4;
            a4.mask = h->mask.as<int8_t>(); a4.active = h->active.as<uint8_t>(); a4.sz_out = h->sz_out.as<float>();
            a4.sub_weight = scan.sub_weight;
            a4.d3_uh = h->d3_uh.as<double>(); a4.d3_w = h->d3_w.as<float>(); a4.d3_rho = h->d3_rho.as<float>();
            a4.d3_st = h->d3_st.as<float2>(); a4.ray_nyquist = nyq;
            a4.attw = h->attw.as<float>(); a4.pre = h->spec_pre.as<SpecPre>(); a4.spec_acc = h->spec_acc.as<double>();
            a4.spec_out = h->spec_out.as<float>() + off_rg * c.n_vel;
            a4.out = h->out.as<float>() + off_rg;
            a4.hist = h->hist.as<int>(); a4.bin_start = h->bin_start.as<int>(); a4.chunk_start = h->cost_start.as<int>();
            a4.totals = totals; a4.counter = h->flags.as<int>() + 2; a4.rec = h->rec.as<PairRec>();
            CU(h, cudaMemsetAsync(h->spec_acc.p, 0, cc_rg * c.n_vel * 8, st));
            a4.d3_std = nullptr; a4.sub_acc = nullptr; a4.brd_flag = nullptr;
            if (c.brd_mode) {
                a4.d3_std = h->d3_std.as<double>(); a4.sub_acc = h->sub_acc.as<double>(); a4.brd_flag = h->brd_flag.as<uint8_t>();
                CU(h, cudaMemsetAsync(h->sub_acc.p, 0, (size_t)nc * c.n_vel * 8, st));
            }
            k_beam_attenuation<<<(nr * c.n_sub + 3) / 4, 128, smem_a, st>>>(c, a4);
            CU(h, cudaGetLastError()); ++launches;
            bool nd1024 = true;
            for (int s = 0; s < nsp; ++s) nd1024 = nd1024 && c.sp[s].nd == 1024;
            if (nd1024) k_spectrum<true><<<std::max(1, occ4) * h->n_sm, 32 * SPEC_WARPS, h->k4_smem, st>>>(c, a4);
            else k_spectrum<false><<<std::max(1, occ4) * h->n_sm, 32 * SPEC_WARPS, h->k4_smem, st>>>(c, a4);
            CU(h, cudaGetLastError()); ++launches;
            if (c.brd_mode) {
                k_spectrum_broaden<<<(unsigned)((nc + 3) / 4), 128, smem_b, st>>>(c, a4, nc);
                CU(h, cudaGetLastError()); ++launches;
            }
            k_spectrum_finish<<<(unsigned)((cc_rg * 32 + 127) / 128), 128, 0, st>>>(c, a4);
            CU(h, cudaGetLastError()); ++launches;
        }
        if (ev) CU(h, cudaEventRecord(ev[5], st));
        if (out) {
            CU(h, cudaEventRecord(h->chunk_done[ch], st));
            CU(h, cudaStreamWaitEvent(h->copy_stream, h->chunk_done[ch], 0));
            rc = copy_outputs(h, out, r0, nr, h->copy_stream);
            if (rc) return rc;
        }
    }
    c.n_rays = n_rays;
    h->n_chunks = n_chunks;
    h->stats.launches = launches;
    h->stats.n_subbeam_gates = h->n_sbg;
    h->executed = true;
    return RADOP_OK;
}

int radop_execute(RadopHandle *h)
{
    if (!h) return RADOP_ERR_ARG;
    return run_scan(h, nullptr);
}

int radop_synchronize(RadopHandle *h)
{
    if (!h) return RADOP_ERR_ARG;
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaStreamSynchronize(h->stream));
    return RADOP_OK;
}

// mean CUDA-event time of each stage over the executes recorded since radop_enable_timing(1)
static void average_timings(RadopHandle *h)
{
    if (!h->n_timed) return;
    double acc[6] = {0, 0, 0, 0, 0, 0};
    for (int i = 0; i < h->n_timed; ++i) {
        cudaEvent_t *e = h->ev.data() + 6 * i;
        float ms[6] = {0, 0, 0, 0, 0, 0};
        cudaEventElapsedTime(&ms[0], e[0], e[1]);
        cudaEventElapsedTime(&ms[1], e[1], e[2]);
        cudaEventElapsedTime(&ms[2], e[2], e[3]);
        cudaEventElapsedTime(&ms[3], e[3], e[4]);
        cudaEventElapsedTime(&ms[4], e[0], e[5]);
        cudaEventElapsedTime(&ms[5], e[4], e[5]);
        for (int k = 0; k < 6; ++k) acc[k] += ms[k];
    }
    const int ne = std::max(1, h->n_exec_timed);     // event sets are per chunk: sum them per execute
    h->stats.ms_interp = (float)(acc[0] / ne);
    h->stats.ms_sort = (float)(acc[1] / ne);
    h->stats.ms_integrate = (float)(acc[2] / ne);
    h->stats.ms_reduce = (float)(acc[3] / ne);
    h->stats.ms_total = (float)(acc[4] / ne);
    h->stats.ms_spectrum = (float)(acc[5] / ne);
    h->stats.n_timed = h->n_exec_timed;
}

static int finish(RadopHandle *h)
{   // after the stream has drained: flags + statistics
    std::vector<long long> tot((size_t)RADOP_TOTALS * std::max(1, h->n_chunks));
    int flags[4];
    CU(h, cudaMemcpy(tot.data(), h->totals.p, tot.size() * 8, cudaMemcpyDeviceToHost));
    CU(h, cudaMemcpy(flags, h->flags.p, sizeof flags, cudaMemcpyDeviceToHost));
    h->stats.n_active_pairs = h->stats.n_tiles = h->stats.n_lut_bins = 0;
    for (int s = 0; s < RADOP_NSPECIES_MAX; ++s) h->stats.pairs_by_species[s] = 0;
    if (h->last_fused) {
        unsigned long long cnt[RADOP_NSPECIES_MAX];
        CU(h, cudaMemcpy(cnt, h->f_cnt.p, sizeof cnt, cudaMemcpyDeviceToHost));
        for (int s = 0; s < RADOP_NSPECIES_MAX; ++s) {
            h->stats.pairs_by_species[s] = (long long)cnt[s];
            h->stats.n_active_pairs += (long long)cnt[s];
        }
    }
    for (int ch = 0; !h->last_fused && ch < std::max(1, h->n_chunks); ++ch) {
        const long long *t = tot.data() + (size_t)RADOP_TOTALS * ch;
        h->stats.n_active_pairs += t[0];
        h->stats.n_tiles += t[1];
        h->stats.n_lut_bins += t[3];      // (species, e, t) bins staged, summed over the chunks
        for (int s = 0; s < RADOP_NSPECIES_MAX; ++s) h->stats.pairs_by_species[s] += t[4 + s];
    }
    // FP32 operations the integration kernel executes per pair (2 per FMA lane): columns x diameters
    for (int s = 0; s < RADOP_NSPECIES_MAX; ++s) h->stats.exec_flop_per_pair[s] = 0;
    for (int s = 0; s < h->cfg.n_species; ++s) {
        const SpeciesDev &S = h->cfg.sp[s];
        const int extra = h->cfg.doppler_scheme == 2 ? 1 : 0;
        double per_d = 0;
        if (S.mode == MODE_GEO8) per_d = 2.0 * (RADOP_NC + extra);
        else if (S.mode == MODE_GEN8) per_d = 2.0 * (RADOP_NC + extra) + 2.0;      // + a D^nu, ex2
        else if (S.mode == MODE_GEN12) per_d = 2.0 * 9 + 2.0;
        else if (S.mode == MODE_MELT) per_d = 0;                                    // float64 kernel, not counted
        else per_d = 2.0 * 10 + 2.0 + 2.0;                                          // ICE1: + shape FMA, two recurrences
        h->stats.exec_flop_per_pair[s] = per_d * S.nd;
    }
    if (h->integration_mode == 0 && h->lt_stat.p) {
        unsigned long long nf[4] = {0, 0, 0, 0};
        CU(h, cudaMemcpy(nf, h->lt_stat.p, 32, cudaMemcpyDeviceToHost));
        h->lt_fallback = (long long)(nf[0] + nf[1] + nf[2] + nf[3]);
        for (int i = 0; i < 4; ++i) h->lt_why[i] = (long long)nf[i];
    }
    average_timings(h);
    h->geo_poly_used = h->geo_poly_launched && !(flags[0] & 4);
    if (flags[0] & 1)
        return h->fail(RADOP_ERR_OUT_OF_DOMAIN, "radar domain is not entirely contained in the model domain");
    if (flags[0] & 8)
        // get_doppler_spectrum fills D[:, j] = np.linspace(d_min, d_max, nbins_D) with the per-gate limits of a
        // melting species, arrays of shape (1,): the assignment raises ValueError (doppler_scatter.py:686-692)
        return h->fail(RADOP_ERR_UNSUPPORTED, "could not broadcast input array from shape (1024,1) into shape (1024,): the Doppler "
                                              "spectrum (scheme 3) of a gate that holds melting snow / graupel -- the reference "
                                              "raises this ValueError in get_doppler_spectrum (doppler_scatter.py:686-692)");
    return RADOP_OK;
}

int radop_fetch(RadopHandle *h, const RadopOut *out)
{
    if (!h) return RADOP_ERR_ARG;
    if (!h->executed) return h->fail(RADOP_ERR_STATE, "nothing executed");
    CU(h, cudaSetDevice(h->device));
    if (out) {
        int rc = copy_outputs(h, out, 0, h->n_rays, h->stream);
        if (rc) return rc;
    }
    CU(h, cudaStreamSynchronize(h->stream));
    return finish(h);
}

int radop_simulate(RadopHandle *h, const RadopScan *scan, const RadopOut *out)
{
    if (!h) return RADOP_ERR_ARG;
    int rc = RADOP_OK;
    if (scan) rc = radop_stage_scan(h, scan);
    else if (h->n_rays <= 0 || h->n_sbg != (long long)h->n_rays * h->n_sub * h->p.n_gates)
        rc = h->fail(RADOP_ERR_STATE, "radop_simulate(scan = NULL) reuses the staged scan: none is staged for these parameters");
    if (rc) return rc;
    rc = run_scan(h, out);          // copies of a chunk overlap the kernels of the next one
    if (rc) {
        // copies into the caller's buffers may already be queued: drain them before the error reaches the caller
        const std::string msg = h->err;
        cudaStreamSynchronize(h->stream);
        cudaStreamSynchronize(h->copy_stream);
        h->err = msg;
        return rc;
    }
    CU(h, cudaStreamSynchronize(h->stream));
    CU(h, cudaStreamSynchronize(h->copy_stream));
    return finish(h);
}

int radop_set_chunk(RadopHandle *h, int64_t subbeam_gates_per_chunk)
{
    if (!h) return RADOP_ERR_ARG;
    h->chunk_sbg = subbeam_gates_per_chunk < 0 ? -1 : subbeam_gates_per_chunk;
    return RADOP_OK;
}

int radop_set_overlap_threshold(RadopHandle *h, int64_t min_output_bytes)
{
    if (!h || min_output_bytes < 0) return RADOP_ERR_ARG;
    h->overlap_min_bytes = (size_t)min_output_bytes;
    return RADOP_OK;
}

int radop_set_geodesy(RadopHandle *h, int mode)
{
    if (!h || (mode != 0 && mode != 1)) return RADOP_ERR_ARG;
    h->geodesy_mode = mode;
    return RADOP_OK;
}

int radop_set_search(RadopHandle *h, int mode)
{
    if (!h || (mode != 0 && mode != 1)) return RADOP_ERR_ARG;
    h->search_mode = mode;
    h->executed = false;
    return RADOP_OK;
}

int radop_set_pipeline(RadopHandle *h, int mode)
{
    if (!h || (mode != 0 && mode != 1)) return RADOP_ERR_ARG;
    h->pipeline_mode = mode;
    h->executed = false;
    return RADOP_OK;
}

int radop_set_tuning(RadopHandle *h, int key, int64_t value)
{
    if (!h) return RADOP_ERR_ARG;
    switch (key) {
    case RADOP_TUNE_FUSED_WARPS:
        if (value < 0 || value > FUSED_MAX_WARPS) return h->fail(RADOP_ERR_ARG, "fused warps per CTA: 0 (automatic) ... 10");
        h->fused_warps = (int)value;
        return RADOP_OK;
    }
    return h->fail(RADOP_ERR_ARG, "unknown tuning key");
}

int radop_get_pipeline(RadopHandle *h, int *used_fused)
{
    if (!h || !used_fused) return RADOP_ERR_ARG;
    *used_fused = h->last_fused ? 1 : 0;
    return RADOP_OK;
}

int radop_set_integration(RadopHandle *h, int mode)
{
    if (!h || (mode != 0 && mode != 1)) return RADOP_ERR_ARG;
    h->integration_mode = mode;
    h->executed = false;
    return RADOP_OK;
}

int radop_get_integration(RadopHandle *h, int *mode, int64_t *n_pieces, int64_t *n_rejected_pieces,
                          int64_t *n_direct_pairs, float *build_ms)
{
    if (!h) return RADOP_ERR_ARG;
    if (mode) *mode = h->integration_mode;
    if (n_pieces) *n_pieces = h->lt_pieces;
    if (n_rejected_pieces) *n_rejected_pieces = h->lt_rejected;
    if (n_direct_pairs) *n_direct_pairs = h->integration_mode == 0 ? h->lt_fallback : 0;
    if (build_ms) *build_ms = h->lt_build_ms;
    return RADOP_OK;
}

int radop_debug_integration(RadopHandle *h, int species, int64_t why[4], int32_t *piece_status, int64_t n_status)
{
    if (!h) return RADOP_ERR_ARG;
    if (why) for (int i = 0; i < 4; ++i) why[i] = h->lt_why[i];
    if (piece_status) {
        if (species < 0 || species >= RADOP_NSPECIES_MAX) return h->fail(RADOP_ERR_ARG, "unknown species id");
        const LutHost &L = h->lut[species];
        if (!L.lamtab_built) return h->fail(RADOP_ERR_STATE, "no lambda table for this species");
        const size_t n = (size_t)L.ne * L.nt * LT_NP;
        if ((size_t)n_status < n) return h->fail(RADOP_ERR_ARG, "piece_status too short");
        CU(h, cudaSetDevice(h->device));
        CU(h, cudaMemcpy2D(piece_status, 4, (const char *)L.lamtab.p + (size_t)12 * L.lt_nc * 4, (size_t)L.lt_stride * 4, 4, n,
                           cudaMemcpyDeviceToHost));
    }
    return RADOP_OK;
}

int radop_get_geodesy(RadopHandle *h, int *used_polynomials)
{
    if (!h || !used_polynomials) return RADOP_ERR_ARG;
    *used_polynomials = h->geo_poly_used ? 1 : 0;
    return RADOP_OK;
}

int radop_device_outputs(RadopHandle *h, void **dev_ptr, int64_t *n_rays, int64_t *n_gates)
{
    if (!h || !dev_ptr) return RADOP_ERR_ARG;
    if (!h->executed) return h->fail(RADOP_ERR_STATE, "nothing executed");
    *dev_ptr = h->out.p;
    if (n_rays) *n_rays = h->cfg.n_rays;
    if (n_gates) *n_gates = h->cfg.n_gates;
    return RADOP_OK;
}

int radop_device_buffer(RadopHandle *h, int which, void **dev_ptr, int64_t *bytes)
{
    if (!h || !dev_ptr) return RADOP_ERR_ARG;
    if (!h->executed) return h->fail(RADOP_ERR_STATE, "nothing executed");
    const size_t n_rg = (size_t)h->n_rays * h->cfg.n_gates;
    void *p = nullptr;
    size_t n = 0;
    switch (which) {
    case RADOP_BUF_OBS: p = h->out.p; n = n_rg * (RADOP_N_OUT + 1) * 4; break;
    case RADOP_BUF_LAT:
    case RADOP_BUF_LON:
        if (!h->latlon_valid) return h->fail(RADOP_ERR_STATE, "lat / lon were not computed by the last run");
        p = which == RADOP_BUF_LAT ? h->lat.p : h->lon.p; n = n_rg * 8; break;
    case RADOP_BUF_MODEL:
        if (!h->cfg.output_model_vars) return h->fail(RADOP_ERR_STATE, "model variables were not integrated (output_model_vars = 0)");
        p = h->model_out.p; n = n_rg * RADOP_NVP * 4; break;
    case RADOP_BUF_DSPECTRUM:
        if (h->cfg.doppler_scheme != 3) return h->fail(RADOP_ERR_STATE, "no Doppler spectrum (scheme 3) in this configuration");
        p = h->spec_out.p; n = n_rg * (size_t)h->cfg.n_vel * 4; break;
    default: return h->fail(RADOP_ERR_ARG, "unknown buffer id");
    }
    *dev_ptr = p;
    if (bytes) *bytes = (int64_t)n;
    return RADOP_OK;
}

int radop_get_stream(RadopHandle *h, void **stream)
{
    if (!h || !stream) return RADOP_ERR_ARG;
    *stream = (void *)h->stream;
    return RADOP_OK;
}

int radop_enable_timing(RadopHandle *h, int on)
{
    if (!h) return RADOP_ERR_ARG;
    h->timing = on != 0;
    h->n_timed = 0;
    h->n_exec_timed = 0;              // a new measurement window
    return RADOP_OK;
}

int radop_get_stats(RadopHandle *h, RadopStats *s)
{
    if (!h || !s) return RADOP_ERR_ARG;
    if (h->executed) {     // drain the stream and refresh counters + timings
        CU(h, cudaSetDevice(h->device));
        CU(h, cudaStreamSynchronize(h->stream));
        int rc = finish(h);
        if (rc != RADOP_OK && rc != RADOP_ERR_OUT_OF_DOMAIN) return rc;
    }
    *s = h->stats;
    return RADOP_OK;
}

int radop_debug_indices(RadopHandle *h, int32_t *cell, int32_t *level, int32_t *status, int32_t *lut_e,
                        int32_t *lut_t, float *values, double *rot_coords)
{
    if (!h) return RADOP_ERR_ARG;
    if (!h->executed || !h->cfg.keep_debug) return h->fail(RADOP_ERR_STATE, "execute with keep_debug = 1 first");
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaStreamSynchronize(h->stream));
    const size_t n = (size_t)h->n_sbg;
    const int nsp = h->cfg.n_species;
    if (cell) CU(h, cudaMemcpy(cell, h->dbg_cell.p, n * 2 * 4, cudaMemcpyDeviceToHost));
    if (level) CU(h, cudaMemcpy(level, h->dbg_level.p, n * 4 * 4, cudaMemcpyDeviceToHost));
    if (status) CU(h, cudaMemcpy(status, h->dbg_status.p, n * 4, cudaMemcpyDeviceToHost));
    if (lut_e) CU(h, cudaMemcpy(lut_e, h->dbg_e.p, n * nsp * 4, cudaMemcpyDeviceToHost));
    if (lut_t) CU(h, cudaMemcpy(lut_t, h->dbg_t.p, n * nsp * 4, cudaMemcpyDeviceToHost));
    if (rot_coords) CU(h, cudaMemcpy(rot_coords, h->dbg_rot.p, n * 2 * 8, cudaMemcpyDeviceToHost));
    if (values) {
        // compact [n][n_uploaded] in ascending variable id
        std::vector<float> tmp(n * RADOP_NVP);
        CU(h, cudaMemcpy(tmp.data(), h->vals.p, n * RADOP_NVP * 4, cudaMemcpyDeviceToHost));
        int nv = 0;
        int ids[RADOP_NVP];
        for (int v = 0; v < RADOP_NVP; ++v)
            if (h->var_mask & (1u << v)) ids[nv++] = v;
        for (size_t i = 0; i < n; ++i)
            for (int j = 0; j < nv; ++j) values[i * nv + j] = tmp[i * RADOP_NVP + ids[j]];
    }
    return RADOP_OK;
}

int radop_measure_fp32_peak(RadopHandle *h, double *tflops)
{
    if (!h || !tflops) return RADOP_ERR_ARG;
    CU(h, cudaSetDevice(h->device));
    const int blocks = h->n_sm * 8, threads = 256, iters = 20000;
    DevBuf sink;
    CU(h, sink.reserve((size_t)blocks * threads * 4));
    cudaEvent_t a, b;
    CU(h, cudaEventCreate(&a));
    CU(h, cudaEventCreate(&b));
    double best = 0;
    for (int rep = 0; rep < 4; ++rep) {
        CU(h, cudaEventRecord(a, h->stream));
        k_fma_probe<<<blocks, threads, 0, h->stream>>>(sink.as<float>(), iters, 0.5f);
        CU(h, cudaGetLastError());
        CU(h, cudaEventRecord(b, h->stream));
        CU(h, cudaEventSynchronize(b));
        float ms = 0;
        cudaEventElapsedTime(&ms, a, b);
        double fl = (double)blocks * threads * 16.0 * 8.0 * iters * 2.0;
        if (rep > 0) best = std::max(best, fl / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    sink.release();
    *tflops = best;
    return RADOP_OK;
}

int radop_selftest_division(RadopHandle *h, uint64_t n_samples, uint64_t seed, uint64_t *mismatches)
{
    if (!h || !mismatches) return RADOP_ERR_ARG;
    CU(h, cudaSetDevice(h->device));
    DevBuf cnt;
    CU(h, cnt.reserve(8));
    CU(h, cudaMemsetAsync(cnt.p, 0, 8, h->stream));
    for (int mode = 0; mode < 3; ++mode) {
        k_selftest_division<<<h->n_sm * 8, 256, 0, h->stream>>>((unsigned long long)n_samples, (unsigned long long)seed + mode, mode,
                                                               cnt.as<unsigned long long>());
        CU(h, cudaGetLastError());
    }
    unsigned long long bad = 0;
    CU(h, cudaMemcpyAsync(&bad, cnt.p, 8, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    cnt.release();
    *mismatches = bad;
    return RADOP_OK;
}

int radop_set_output_blocks(RadopHandle *h, int block_rays, int block_pitch_rays)
{
    if (!h) return RADOP_ERR_ARG;
    if (block_rays < 0 || (block_rays > 0 && block_pitch_rays < block_rays))
        return h->fail(RADOP_ERR_ARG, "output blocks: block_rays >= 0 and block_pitch_rays >= block_rays");
    h->out_block = block_rays;
    h->out_pitch = block_rays > 0 ? block_pitch_rays : 0;
    return RADOP_OK;
}

int radop_host_register(RadopHandle *h, void *ptr, int64_t bytes)
{
    if (!h || !ptr || bytes <= 0) return RADOP_ERR_ARG;
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaHostRegister(ptr, (size_t)bytes, cudaHostRegisterPortable));
    return RADOP_OK;
}

int radop_host_unregister(RadopHandle *h, void *ptr)
{
    if (!h || !ptr) return RADOP_ERR_ARG;
    CU(h, cudaSetDevice(h->device));
    CU(h, cudaHostUnregister(ptr));
    return RADOP_OK;
}

int radop_get_all_radar_pts(RadopHandle *h, int n, const float *coords, const float *heights_gates, const float *data,
                            const float *model_heights, int nz, int ny, int nx, const float *llc, const float *res,
                            float *output)
{
    if (!h || !coords || !heights_gates || !data || !model_heights || !llc || !res || !output || n <= 0)
        return RADOP_ERR_ARG;
    CU(h, cudaSetDevice(h->device));
    DevBuf dc, dh, dd, dz, dout;
    size_t cells = (size_t)nz * ny * nx;
    CU(h, dc.reserve((size_t)n * 8));
    CU(h, dh.reserve((size_t)n * 4));
    CU(h, dd.reserve(cells * 4));
    CU(h, dz.reserve(cells * 4));
    CU(h, dout.reserve((size_t)n * 4));
    CU(h, cudaMemcpyAsync(dc.p, coords, (size_t)n * 8, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(dh.p, heights_gates, (size_t)n * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(dd.p, data, cells * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(dz.p, model_heights, cells * 4, cudaMemcpyHostToDevice, h->stream));
    k_interp_compat<<<(n + 127) / 128, 128, 0, h->stream>>>(n, dc.as<float>(), dh.as<float>(), dd.as<float>(), dz.as<float>(), nz, ny, nx,
                                                           llc[0], llc[1], res[0], res[1], dout.as<float>());
    CU(h, cudaGetLastError());
    CU(h, cudaMemcpyAsync(output, dout.p, (size_t)n * 4, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    dc.release(); dh.release(); dd.release(); dz.release(); dout.release();
    return RADOP_OK;
}

int radop_get_refl(RadopHandle *h, int len, int nhydro, int nd, const float *Da, const float *Db, const float *rcs,
                   const float *D, const float *N, const float *step_D, const float *Dmin, float *refl)
{
    (void)D;      // the reference passes the diameters and never reads them (doppler_c.c:11-32)
    if (!h || !Da || !Db || !rcs || !N || !step_D || !Dmin || !refl || len <= 0 || nhydro <= 0 || nd <= 0)
        return RADOP_ERR_ARG;
    CU(h, cudaSetDevice(h->device));
    DevBuf blob, dout;
    const size_t n_ab = (size_t)len * nhydro, n_tab = (size_t)nd * nhydro;
    CU(h, blob.reserve((2 * n_ab + 2 * n_tab + 2 * (size_t)nhydro) * 4));
    CU(h, dout.reserve((size_t)len * 4));
    float *dDa = blob.as<float>(), *dDb = dDa + n_ab, *drcs = dDb + n_ab, *dN = drcs + n_tab, *dstep = dN + n_tab,
          *dmin = dstep + nhydro;
    CU(h, cudaMemcpyAsync(dDa, Da, n_ab * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(dDb, Db, n_ab * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(drcs, rcs, n_tab * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(dN, N, n_tab * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(dstep, step_D, (size_t)nhydro * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(dmin, Dmin, (size_t)nhydro * 4, cudaMemcpyHostToDevice, h->stream));
    k_refl_compat<<<(len + 127) / 128, 128, 0, h->stream>>>(len, nhydro, nd, dDa, dDb, drcs, dN, dstep, dmin, dout.as<float>());
    CU(h, cudaGetLastError());
    CU(h, cudaMemcpyAsync(refl, dout.p, (size_t)len * 4, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    blob.release(); dout.release();
    return RADOP_OK;
}

}  // extern "C"

// ---- allocation-free variant of radop_get_refl ---------------------------------------------------------------------
// The staged operands and the result of k_refl_compat live in ONE file-scope device array instead of cudaMalloc'ed
// buffers.  Every GPU that loads the library holds these RADOP_NOALLOC_SCRATCH_BYTES until the process exits.  In host
// code the array's name is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress, asked
// once per device with that device current, and is what the copies and the kernel are given.
static __device__ __align__(256) unsigned char g_noalloc_scratch[RADOP_NOALLOC_SCRATCH_BYTES];
static std::mutex g_noalloc_mu;
static void *g_noalloc_ptr[64] = {nullptr};

static cudaError_t noalloc_scratch_address(int device, void **ptr)
{   // the caller has made `device` current
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    void *&slot = g_noalloc_ptr[device & 63];
    if (device >= 64 || !slot) {
        void *q = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&q, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        if (!q) return cudaErrorInvalidSymbol;
        if (device >= 64) { *ptr = q; return cudaSuccess; }
        slot = q;
    }
    *ptr = slot;
    return cudaSuccess;
}

extern "C" int radop_get_refl_noalloc(RadopHandle *h, int len, int nhydro, int nd, const float *Da, const float *Db,
                                      const float *rcs, const float *D, const float *N, const float *step_D,
                                      const float *Dmin, float *refl)
{
    (void)D;
    if (!h) return RADOP_ERR_ARG;
    if (!Da || !Db || !rcs || !N || !step_D || !Dmin || !refl || len <= 0 || nhydro <= 0 || nd <= 0)
        return h->fail(RADOP_ERR_ARG, "radop_get_refl_noalloc: null pointer or non-positive size");
    const size_t n_ab = (size_t)len * nhydro, n_tab = (size_t)nd * nhydro;
    const size_t need = (2 * n_ab + 2 * n_tab + 2 * (size_t)nhydro + (size_t)len) * 4;
    if (need > (size_t)RADOP_NOALLOC_SCRATCH_BYTES) {
        char msg[256];
        snprintf(msg, sizeof msg, "radop_get_refl_noalloc: len=%d nhydro=%d nd=%d need %zu bytes of device scratch, the static "
                 "array holds %u (use radop_get_refl)", len, nhydro, nd, need, (unsigned)RADOP_NOALLOC_SCRATCH_BYTES);
        return h->fail(RADOP_ERR_ARG, msg);
    }
    CU(h, cudaSetDevice(h->device));
    void *base = nullptr;
    CU(h, noalloc_scratch_address(h->device, &base));
    // nothing in the scratch has to start at zero: every word the kernel reads is overwritten by this call's copies first,
    // and the kernel writes all `len` words of the result before they are copied back
    float *dDa = static_cast<float *>(base), *dDb = dDa + n_ab, *drcs = dDb + n_ab, *dN = drcs + n_tab, *dstep = dN + n_tab,
          *dmin = dstep + nhydro, *dout = dmin + nhydro;
    CU(h, cudaMemcpyAsync(dDa, Da, n_ab * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(dDb, Db, n_ab * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(drcs, rcs, n_tab * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(dN, N, n_tab * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(dstep, step_D, (size_t)nhydro * 4, cudaMemcpyHostToDevice, h->stream));
    CU(h, cudaMemcpyAsync(dmin, Dmin, (size_t)nhydro * 4, cudaMemcpyHostToDevice, h->stream));
    k_refl_compat<<<(len + 127) / 128, 128, 0, h->stream>>>(len, nhydro, nd, dDa, dDb, drcs, dN, dstep, dmin, dout);
    CU(h, cudaGetLastError());
    CU(h, cudaMemcpyAsync(refl, dout, (size_t)len * 4, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return RADOP_OK;
}
