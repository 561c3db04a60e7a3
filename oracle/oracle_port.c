/*
 * oracle_port.c -- TEST INFRASTRUCTURE ONLY (CPU oracle; never linked into the product).
 *
 * Plain-C restatement of the two native kernels of the reference:
 *   - trilinear field-to-gate interpolation   cosmo_pol/interpolation/interpolation_c.c:12-164
 *   - Doppler-spectrum bin reflectivity       cosmo_pol/scatter/doppler_c.c:11-32
 *
 * It is validated bit-for-bit against the reference's own sources compiled into
 * oracle/_ref/ (tests/test_oracle_native.py).  Unlike the reference it also
 * reports the integer decisions (grid cell, level per neighbour column, status)
 * so that the CUDA path can be checked bit-exactly on indices.
 *
 * Float semantics follow the reference: every operand is a float, constants such
 * as 1.0 are doubles (the subtraction 1.0-x is exact in double and then rounded
 * once, i.e. identical to a float subtraction).  Compile with -ffp-contract=off.
 */
#include <math.h>
#include <stddef.h>

/* status codes written to status[] */
enum { GATE_OK = 0, GATE_ABOVE_TOP = 1, GATE_BELOW_TOPO = -1 };

/* Descending-column search, interpolation_c.c:108-135.
 * returns -1 (key above arr[0]), -2 (key below arr[dim-1]) or the index of the
 * level just above the key. */
static int column_search(const float *col, ptrdiff_t stride, int dim, float key)
{
    int hi = 0, lo = dim - 1, mid;
    if (key > col[0]) return -1;
    if (key < col[(ptrdiff_t)(dim - 1) * stride]) return -2;
    while (hi < lo) {
        float c;
        mid = (hi + lo) / 2;
        c = col[(ptrdiff_t)mid * stride];
        if (c == key) return mid;
        if (lo == mid || hi == mid) return hi;
        if (c < key) lo = mid;
        else if (c > key) hi = mid;
    }
    return hi;
}

/*
 * coords : [n][2] (rotated lat, rotated lon) float32
 * h      : [n] gate heights float32
 * data   : [nz][ny][nx] float32, heights: [nz][ny][nx] float32 (level 0 = model top)
 * llc    : (Lo1, La1) ; res : (dlon, dlat)
 * out    : [n] interpolated value, NaN below topography, -9999 above model top
 * cell   : [n][2] (lat index, lon index)         (may be NULL)
 * level  : [n][4] upper level per neighbour col   (may be NULL; -1 when unused)
 * status : [n]                                    (may be NULL)
 */
void oracle_interp_gates(int n, const float *coords, const float *h,
                         const float *data, const float *heights,
                         int nz, int ny, int nx,
                         const float *llc, const float *res,
                         float *out, int *cell, int *level, int *status)
{
    const ptrdiff_t plane = (ptrdiff_t)ny * nx;
    for (int g = 0; g < n; ++g) {
        float p0 = (coords[2 * g + 0] - llc[1]) / res[1];
        float p1 = (coords[2 * g + 1] - llc[0]) / res[0];
        int iy = (int)floor(p0);
        int ix = (int)floor(p1);
        float x = (float)fmod(p0, 1.0);
        float y = (float)fmod(p1, 1.0);
        float dx = (float)(1.0 - x);
        float dy = (float)(1.0 - y);
        const int ny_[4] = { iy, iy, iy + 1, iy + 1 };
        const int nx_[4] = { ix, ix + 1, ix, ix + 1 };
        float t[4], topo, vv[4];
        int lv[4] = { -1, -1, -1, -1 };
        int st = GATE_OK;

        if (cell) { cell[2 * g] = iy; cell[2 * g + 1] = ix; }
        for (int k = 0; k < 4; ++k)
            t[k] = heights[(ptrdiff_t)(nz - 1) * plane + (ptrdiff_t)ny_[k] * nx + nx_[k]];
        topo = dx * dy * t[0] + x * t[2] * dy + dx * t[1] * y + x * y * t[3];

        if (!(topo < h[g])) {
            st = GATE_BELOW_TOPO;
            out[g] = (float)NAN;
        } else {
            for (int k = 0; k < 4; ++k) {
                const float *hc = heights + (ptrdiff_t)ny_[k] * nx + nx_[k];
                const float *dc = data + (ptrdiff_t)ny_[k] * nx + nx_[k];
                int idx = column_search(hc, plane, nz, h[g]);
                int l1;
                float c1, c2, v1, v2;
                if (idx == -1) { st = GATE_ABOVE_TOP; break; }
                if (idx == -2) l1 = nz - 3;
                else { if (idx == nz - 2) idx--; l1 = idx; }
                lv[k] = l1;
                c1 = hc[(ptrdiff_t)l1 * plane];
                c2 = hc[(ptrdiff_t)(l1 + 1) * plane];
                v1 = dc[(ptrdiff_t)l1 * plane];
                v2 = dc[(ptrdiff_t)(l1 + 1) * plane];
                vv[k] = v2 - (v2 - v1) / (c1 - c2) * (h[g] - c2);
            }
            if (st == GATE_ABOVE_TOP) out[g] = -9999.0f;
            else out[g] = dx * dy * vv[0] + x * vv[2] * dy + dx * vv[1] * y + x * y * vv[3];
        }
        if (status) status[g] = st;
        if (level) for (int k = 0; k < 4; ++k) level[4 * g + k] = (st == GATE_OK) ? lv[k] : -1;
    }
}

/*
 * Doppler spectrum bin sum, doppler_c.c:11-32.
 * Da, Db : [nbins][nh] diameter bin edges; rcs, N : [nd][nh]; step, dmin : [nh]
 * refl   : [nbins]
 */
void oracle_spectrum_bins(int nbins, int nh, const float *Da, const float *Db,
                          const float *rcs, const float *N,
                          const float *step, const float *dmin, float *refl)
{
    for (int i = 0; i < nbins; ++i) {
        refl[i] = 0.0f;
        for (int j = 0; j < nh; ++j) {
            int ka = (int)((Da[i * nh + j] - dmin[j]) / step[j]);
            int kb = (int)((Db[i * nh + j] - dmin[j]) / step[j]);
            float acc = 0.0f;
            for (int k = ka; k < kb; ++k) acc += N[k * nh + j] * rcs[k * nh + j];
            refl[i] += acc * step[j];
        }
    }
}
