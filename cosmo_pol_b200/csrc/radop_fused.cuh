// radop_fused.cuh -- K1f k_gate_fused: the whole per-gate hot path of the default configuration in ONE kernel.
//
// Included by radop_kernels.cu after the interpolation helpers (locate_cell, div_const, lut_bin) and the
// lambda-table helpers (lt_horner, lt_scale, warp_direct).
//
// What it fuses (reference files as in the kernel file header): geolocation from the per sub-beam polynomials,
// the bit-exact trilinear interpolation of interpolation_c.c:42-103, the PSD parameters of every species
// (hydrometeors.py:224-255, 761-772, 894-907, 1039-1049, 1320-1373), the floor bins of lookup_line (lut.py:336-342),
// the PSD x scattering-table contraction through the lambda-transformed tables (doppler_scatter.py:246-268),
// the sum over the species (:148-158, 400-401), the analytic / rcs-weighted fall-speed moments and the wind
// projection (:276-333), and the quadrature sum over the sub-beams of a ray (:259-268, 313-333, 472-477).
// The older pipeline -- k_interp_psd, k_integrate_tab, k_reduce_subbeams -- wrote keys, slopes, intercepts and one
// 32-byte row of sums per sub-beam gate to HBM and read them back twice (~3 GB per volume); here nothing per
// sub-beam gate leaves the SM.  Melting species, Doppler spectra, the direct integration, model-variable outputs
// and the debug index dump keep the older pipeline (they need the per sub-beam-gate arrays).
//
// Mapping: a work piece is (ray, span of 32 consecutive gates); lanes = gates.  The CTA has W warps; warp w
// takes the sub-beams q = w, w + W, ... of the piece (W = number of sub-beams when there are at most 12), so the
// sub-beams of one gate span run side by side on one SM and find each other's model cells in L1.  Every warp
// adds its sub-beams' sums into a shared-memory row; warp 0 adds the rows in warp order (a fixed order: results
// do not depend on scheduling, chunking or sharding) and writes one 36-byte record per ray gate.
// Pieces are handed out from a global counter, so CTAs that start late (an NCCL kernel on their SM) take fewer.
#pragma once

#ifndef FUSED_MAX_WARPS
#define FUSED_MAX_WARPS 9     // warps per CTA (a build-time experiment knob together with FUSED_MIN_CTAS: registers per thread)
#endif
#ifndef FUSED_MIN_CTAS
#define FUSED_MIN_CTAS 3      // 9 warps x 3 CTAs -> 72 registers (4 CTAs -> 56 registers and more spills: 1.7 % slower)
#endif
#ifndef FUSED_DBUF
#define FUSED_DBUF 1          // 1: two row buffers, ONE CTA barrier per piece, the reducing warp rotates (0: one buffer, two
                              // barriers).  Measured at cfg 3 (profiles/r02_fused_variants.md): 72 registers 3.60 ms with, 3.63 ms
                              // without; 56 registers 3.74 / 3.67 ms.  A barrier-free hand-over through shared-memory flags
                              // was 60 % slower (spinning warps).
#endif
#ifndef FUSED_UNROLL_COLS
#define FUSED_UNROLL_COLS 1   // 1: the four neighbour columns of the interpolation as straight-line code (measured 3.68 ms
                              // against 3.95 ms rolled at 56 registers)
#endif
#define FUSED_NPART 10        // b c p q kd s9 s11 | sum rv w | sum w | sum mask

// constants of one species in the form the fused kernel wants them (filled by the host, prepare_fused())
struct FusedSpecies {
    double u0, uk;            // u = log2(lambda) = uk * (u0 [+ log2 N0(T)] - log2 QM)   (one-moment families)
    double na, nb;            // snow: log2 N0 = na + nb * T
    float n0;                 // fixed intercept
    float cv, cn;             // analytic moments: v = cv n0 2^(ev u), n = cn n0 2^(en u)
    float ev, en;
};

struct FusedArgs {
    const SubPoly *poly;            // [ray][sub] of the rays of this launch
    const float2 *traj_sc;          // [trajectory row][gate] (sin, cos) of the folded incidence elevation
    const uint8_t *traj_ie;         // [species][trajectory row][gate] elevation bin of the species' table
    long long traj_n;               // trajectory rows x gates
    float *comb;                    // [ray][gate][8]: b c p q kd s9 s11 | sum of rv * w
    float *wsum;                    // [ray][gate] sum of the weights behind comb[7]
    float *mask_out;                // [ray][gate] averaged mask
    double *lat, *lon;              // [ray][gate] central sub-beam, or null
    float4 *ice_part;               // [ray][sub][span][2] partial numeric fall-speed moments, or null
    int *flags;                     // [0] bit 0: a gate left the model domain; [3] work counter
    unsigned long long *n_fallback; // [4] pairs summed directly, by reason
    unsigned long long *n_pairs;    // [RADOP_NSPECIES_MAX] active pairs per species
    int n_spans;                    // ceil(n_gates / 32)
    FusedSpecies fs[RADOP_NSPECIES_MAX];
};

// (sin, cos) of the folded elevation and the elevation bins of every species' table for every trajectory row:
// they depend on the scan geometry only (doppler_scatter.py:173-176; lut.py:336-341).
__global__ void k_traj_tables(const __grid_constant__ DevConfig cfg, const float *__restrict__ traj_e, long long n,
                              float2 *__restrict__ sc, uint8_t *__restrict__ ie)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float e = __ldg(traj_e + i);
    if (e > 90.f) e = __fsub_rn(180.f, e);
    if (e < 0.f) e = -e;
    float st, ct;
    sincosf(e * 0.017453292519943295f, &st, &ct);
    sc[i] = make_float2(st, ct);
    for (int s = 0; s < cfg.n_species; ++s) {
        const SpeciesDev &S = cfg.sp[s];
        ie[(size_t)s * n + i] = (uint8_t)lut_bin(e, S.e_min, S.e_step, S.e_rstep, S.ne);
    }
}

// the interpolation of interp_columns for the variables the radar observables need (RHO and EDR are read by
// the Doppler spectrum and the model outputs only): slots 0..7 = U V W T QR QS QG QI, two 16-byte loads per level
// MODE 0: shared-reciprocal quotients with per-quotient range checks (returns true when the gate must be redone with
// IEEE divisions), 1: IEEE divisions, 2: shared-reciprocal quotients WITHOUT checks -- only when the upload-time bounds
// (DevConfig.div_exact) guarantee that no quotient of this grid and these fields can leave the exact range.
template <int MODE, bool TWO_MOM>
__device__ __forceinline__ bool interp_fused(const DevConfig &cfg, const CellGeom &c, float h_g,
                                             const float *__restrict__ z_t, const float *__restrict__ fields,
                                             float (&val)[TWO_MOM ? 15 : 8])
{
    constexpr int NV = TWO_MOM ? 15 : 8;
    bool rare = false;
#pragma unroll
    for (int j = 0; j < NV; ++j) val[j] = -0.f;
#if FUSED_UNROLL_COLS
#pragma unroll
#else
#pragma unroll 1
#endif
    for (int oi = 0; oi < 4; ++oi) {
        const int k = (oi == 1) ? 2 : (oi == 2 ? 1 : oi);        // accumulation order of the reference: 0, 2, 1, 3
        const int yk = (k & 2) ? c.y1 : c.y0, xk = (k & 1) ? c.x1 : c.x0;
        const int lvl = (k == 0) ? c.level[0] : (k == 1 ? c.level[1] : (k == 2 ? c.level[2] : c.level[3]));
        const size_t cell = ((size_t)yk * cfg.nx + xk) * cfg.nz + lvl;
        const float c1 = __ldg(z_t + cell), c2 = __ldg(z_t + cell + 1);
        const float den = __fsub_rn(c1, c2), dh = __fsub_rn(h_g, c2);
        const float rden = MODE == 1 ? 0.f : __frcp_rn(den);
        if (MODE == 0) rare = rare || !div_den_safe(den);
        const float4 *f1 = reinterpret_cast<const float4 *>(fields + cell * RADOP_NVP);
        const float4 *f2 = f1 + RADOP_NVP / 4;
        const bool v_mid = (k == 1) || (k == 2);
        const float w0 = (k == 0) ? __fmul_rn(c.dx, c.dy) : (k == 2 ? c.x : (k == 1 ? c.dx : __fmul_rn(c.x, c.y)));
        const float w1 = (k == 2) ? c.dy : c.y;
        auto accum = [&](int j, float a, float b) {
            const float n = __fsub_rn(b, a);
            float q;
            if (MODE == 1) q = __fdiv_rn(n, den);
            else {
                const float q0 = __fmul_rn(n, rden);
                const float r = __fmaf_rn(-den, q0, n);
                q = __fmaf_rn(r, rden, q0);
                if (MODE == 0) {
                    const float aq = fabsf(q);
                    rare = rare || !(aq <= 1e30f) || (aq < 1e-30f && n != 0.f);
                }
            }
            const float vv = __fsub_rn(b, __fmul_rn(q, dh));
            const float t0 = __fmul_rn(w0, vv);
            val[j] = __fadd_rn(val[j], v_mid ? __fmul_rn(t0, w1) : t0);
        };
        {
            const float4 a = __ldg(f1), b = __ldg(f2);
            accum(0, a.x, b.x); accum(1, a.y, b.y); accum(2, a.z, b.z); accum(3, a.w, b.w);
        }
        {
            const float4 a = __ldg(f1 + 1), b = __ldg(f2 + 1);
            accum(4, a.x, b.x); accum(5, a.y, b.y); accum(6, a.z, b.z); accum(7, a.w, b.w);
        }
        if (TWO_MOM) {               // slot 8 (RHO) and slot 15 (EDR) are not needed
            const float4 a = __ldg(f1 + 2), b = __ldg(f2 + 2), a3 = __ldg(f1 + 3), b3 = __ldg(f2 + 3);
            accum(9, a.y, b.y); accum(10, a.z, b.z); accum(11, a.w, b.w);
            accum(12, a3.x, b3.x); accum(13, a3.y, b3.y); accum(14, a3.z, b3.z);
        }
    }
    return rare;
}

template <int NV>
__device__ __forceinline__ float pick_var(const float (&v)[NV], int var)
{   // register-resident select (a dynamic index would push the array to local memory)
    float r = v[RADOP_VAR_QR];
    if (var == RADOP_VAR_QS) r = v[RADOP_VAR_QS];
    if (var == RADOP_VAR_QG) r = v[RADOP_VAR_QG];
    if (var == RADOP_VAR_QI) r = v[RADOP_VAR_QI];
    if (NV > 8) {
        if (var == RADOP_VAR_QH) r = v[NV > 8 ? RADOP_VAR_QH : 0];
        if (var == RADOP_VAR_QNR) r = v[NV > 8 ? RADOP_VAR_QNR : 0];
        if (var == RADOP_VAR_QNS) r = v[NV > 8 ? RADOP_VAR_QNS : 0];
        if (var == RADOP_VAR_QNG) r = v[NV > 8 ? RADOP_VAR_QNG : 0];
        if (var == RADOP_VAR_QNI) r = v[NV > 8 ? RADOP_VAR_QNI : 0];
        if (var == RADOP_VAR_QNH) r = v[NV > 8 ? RADOP_VAR_QNH : 0];
    }
    return r;
}

// One column of one piece of a lambda-transformed table (layout of k_build_lamtab): mantissa 2^{L(x)} R(x) of
// F_k(lambda); the integer part of the exponent stays in the piece's exponent word.  The sign-mask test is hoisted
// to the warp: a column whose ratio polynomial is the constant +-1 on the pieces of ALL 32 lanes (every column but
// s10 - s8 in the smooth table family) costs no divergent branch.
template <int NC, int K>
__device__ __forceinline__ float lt_col(const float *__restrict__ base, float x, unsigned mask, unsigned all_const)
{
    const float2 *l = reinterpret_cast<const float2 *>(base + 6 * K);
    const float2 l0 = __ldg(l), l1 = __ldg(l + 1), l2 = __ldg(l + 2);
    float L = l2.y;
    L = fmaf(L, x, l2.x);
    L = fmaf(L, x, l1.y);
    L = fmaf(L, x, l1.x);
    L = fmaf(L, x, l0.y);
    L = fmaf(L, x, l0.x);
    // sign bit of the constant ratio moved onto the value: -1 where bit 16 + K of the mask word is set
    float R = __int_as_float(0x3f800000 | ((mask << (15 - K)) & 0x80000000u));
    if (!(all_const & (1u << K))) {               // warp-uniform
        if (!(mask & (1u << K))) {
            const float2 *r = reinterpret_cast<const float2 *>(base + 6 * NC + 6 * K);
            const float2 r0 = __ldg(r), r1 = __ldg(r + 1), r2 = __ldg(r + 2);
            R = r2.y;
            R = fmaf(R, x, r2.x);
            R = fmaf(R, x, r1.y);
            R = fmaf(R, x, r1.x);
            R = fmaf(R, x, r0.y);
            R = fmaf(R, x, r0.x);
        }
    }
    return ex2f_approx(L) * R;
}

// The signed ratio R of column K at x (see lt_col): the constant +-1 from the mask word, or its polynomial.
template <int NC, int K>
__device__ __forceinline__ float lt_ratio(const float *__restrict__ base, float x, unsigned mask, unsigned all_const)
{
    float R = __int_as_float(0x3f800000 | ((mask << (15 - K)) & 0x80000000u));
    if (!(all_const & (1u << K))) {               // warp-uniform
        if (!(mask & (1u << K))) {
            const float2 *r = reinterpret_cast<const float2 *>(base + 6 * NC + 6 * K);
            const float2 r0 = __ldg(r), r1 = __ldg(r + 1), r2 = __ldg(r + 2);
            R = r2.y;
            R = fmaf(R, x, r2.x);
            R = fmaf(R, x, r1.y);
            R = fmaf(R, x, r1.x);
            R = fmaf(R, x, r0.y);
            R = fmaf(R, x, r0.x);
        }
    }
    return R;
}

// Columns K and K + 1 (K even) at once: their twelve envelope coefficients are 48 contiguous, 16-byte aligned bytes,
// read with three 16-byte loads instead of six 8-byte ones -- the kernel is bound by the L1 wavefronts of exactly
// these loads (every distinct piece among the 32 lanes costs one wavefront per load instruction).
template <int NC, int K>
__device__ __forceinline__ float2 lt_col_pair(const float *__restrict__ base, float x, unsigned mask, unsigned all_const)
{
    static_assert((K & 1) == 0 && K + 1 < NC, "pairs start at even columns");
    const float4 *l = reinterpret_cast<const float4 *>(base + 6 * K);
    const float4 a = __ldg(l), b = __ldg(l + 1), c = __ldg(l + 2);
    float L0 = b.y, L1 = c.w;
    L0 = fmaf(L0, x, b.x); L1 = fmaf(L1, x, c.z);
    L0 = fmaf(L0, x, a.w); L1 = fmaf(L1, x, c.y);
    L0 = fmaf(L0, x, a.z); L1 = fmaf(L1, x, c.x);
    L0 = fmaf(L0, x, a.y); L1 = fmaf(L1, x, b.w);
    L0 = fmaf(L0, x, a.x); L1 = fmaf(L1, x, b.z);
    return make_float2(ex2f_approx(L0) * lt_ratio<NC, K>(base, x, mask, all_const),
                       ex2f_approx(L1) * lt_ratio<NC, K + 1>(base, x, mask, all_const));
}

// v * 2^e: one exact multiplication for the exponents the tables really hold (|e| <= 126), else the two-factor form
__device__ __forceinline__ float scale2(float v, int e)
{
    if ((unsigned)(e + 126) > 252u) return lt_scale(v, e);
    return v * __int_as_float((e << 23) + 0x3f800000);
}

// A pair the table cannot serve (slope outside the tabulated range, rejected piece): the sums over the diameters by the
// whole warp (warp_direct), then the arithmetic of the table path with zero exponents.  out[0..6] columns, out[7..8]
// the fall-speed moments.  Cold: a few hundred pairs of 2.7e7.
template <int NC, bool DOP2>
__device__ __noinline__ void pair_direct(const DevConfig &cfg, const SpeciesDev &S, int bl, float lam, float p1, float w,
                                         int lane, float (&out)[9])
{
    float m[NC];
    warp_direct<NC>(S, bl, lam, lane, m);
    const int nd = S.nd;
    out[7] = 0.f; out[8] = 0.f;
    constexpr int C9 = NC > 9 ? 9 : 0, C8 = NC > 8 ? 8 : 0, C7 = NC > 7 ? 7 : 0;
    if (S.mode == MODE_ICE1) {
        const float inv9 = 1.f / m[C9];
#pragma unroll
        for (int c = 0; c < 7; ++c) out[c] = m[c] * inv9 * p1 * w;
        if (!DOP2) { out[7] = m[C7] * inv9 * p1; out[8] = m[C8] * inv9 * p1; }
        else {
            const float *tab = S.table + (size_t)bl * nd * S.ncol;
            const float D0 = __ldg(S.aux), DL = __ldg(S.aux + nd - 1);
            const float c2 = 17.46f * exp2f(0.6357f * log2f(lam));
            const float w0 = 490.6f * ex2f_approx(-20.78f * LOG2E * lam * D0) + c2 * __ldg(S.aux + 2 * nd) * ex2f_approx(-3.290f * LOG2E * lam * D0);
            const float wL = 490.6f * ex2f_approx(-20.78f * LOG2E * lam * DL) + c2 * __ldg(S.aux + 3 * nd - 1) * ex2f_approx(-3.290f * LOG2E * lam * DL);
            const float n_ = m[0] - 0.5f * (__ldg(tab) * w0 + __ldg(tab + (size_t)(nd - 1) * S.ncol) * wL);
            const float v_ = m[C7] - 0.5f * (__ldg(tab + 7) * w0 + __ldg(tab + (size_t)(nd - 1) * S.ncol + 7) * wL);
            const float sc = 6.283185307179586f * (p1 / m[C9]) / S.dD_ref;
            out[7] = v_ * sc; out[8] = n_ * sc;
        }
        return;
    }
#pragma unroll
    for (int c = 0; c < 7; ++c) out[c] = m[c] * (p1 * w);
    if (S.mode == MODE_GEN12) { out[7] = m[C7] * p1; out[8] = m[C8] * p1; }
    else if (DOP2) {
        const float *tab = S.table + (size_t)bl * nd * S.ncol;
        const float a = -LOG2E * lam;
        float yF, yL;
        if (S.mode == MODE_GEO8) { yF = __ldg(S.aux); yL = (float)((double)yF + (double)(nd - 1) * S.dD_d); }
        else { yF = __ldg(S.aux + nd); yL = __ldg(S.aux + 2 * nd - 1); }
        const float wF = ex2f_approx(a * yF), wL = ex2f_approx(a * yL);
        const float n_ = m[0] - 0.5f * (__ldg(tab) * wF + __ldg(tab + (size_t)(nd - 1) * S.ncol) * wL);
        const float v_ = m[C7] - 0.5f * (__ldg(tab + 7) * wF + __ldg(tab + (size_t)(nd - 1) * S.ncol + 7) * wL);
        const float sc = 6.283185307179586f * p1 / S.dD_ref;
        out[7] = v_ * sc; out[8] = n_ * sc;
    }
}

// One species of one warp's 32 gates, from the PSD slope in the log domain (u = log2 lambda) to the seven weighted
// column sums: the arithmetic of pair_tab without the trip through global memory, one column at a time so that no
// array of mantissas and exponents has to stay in registers.  vn: numeric fall-speed moments (ice, Doppler scheme 1)
// or the rcs-weighted ones of scheme 2.  All 32 lanes must call.
template <int NC, bool DOP2>
__device__ __forceinline__ void species_tab(const DevConfig &cfg, const FusedArgs &A, const SpeciesDev &S, bool have,
                                            int bl, double u, float p1, float w, int lane, float (&acc)[7], float2 &vn)
{
    constexpr int STR = (12 * NC + ((NC + 1 + 3) & ~3) + 31) / 32 * 32;      // lt_stride(NC)
    constexpr int C9 = NC > 9 ? 9 : 0, C8 = NC > 8 ? 8 : 0, C7 = NC > 7 ? 7 : 0;
    int piece = 0, why = 0;
    float x = 0.f;
    if (have) {
        // piece and position inside it straight from u (what lt_locate derives from the float32 slope)
        const double t = (u - (double)LT_U0) * (double)LT_PPO;
        if (!(fabs(t) < 1e9)) why = 4;
        else if (t < 0.0) why = 2;
        else if (t >= (double)LT_NP) why = 3;
        else {
            piece = (int)t;
            x = (float)(2.0 * (t - (double)piece) - 1.0);
        }
    }
    // lanes without a servable pair read piece 0 of their bin (valid memory); their result is discarded
    const float *base = S.lamtab + (size_t)(bl * LT_NP + piece) * STR;
    const int *ew = reinterpret_cast<const int *>(base) + 12 * NC;        // NC exponent words, then the mask word
    const int4 e03 = __ldg(reinterpret_cast<const int4 *>(ew));          // exponents of columns 0..3 (word 0 doubles as the reject mark)
    if (have && !why && e03.x == LT_BAD) why = 1;
    const unsigned need = __ballot_sync(0xffffffffu, why != 0);
    const bool tab = have && !why;
    if (need) {
        if (why) atomicAdd(A.n_fallback + (why - 1), 1ull);
        const float lam = exp2f((float)u);
        unsigned todo = need;
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            float out[9];
            pair_direct<NC, DOP2>(cfg, S, __shfl_sync(0xffffffffu, bl, src), __shfl_sync(0xffffffffu, lam, src),
                                  __shfl_sync(0xffffffffu, p1, src), w, lane, out);
            if (lane == src) {
#pragma unroll
                for (int c = 0; c < 7; ++c) acc[c] += out[c];
                vn = make_float2(out[7], out[8]);
            }
        }
    }
    const int4 e47 = __ldg(reinterpret_cast<const int4 *>(ew) + 1);      // columns 4..7; with 7 columns word 7 IS the mask word
    const unsigned mask = NC == 7 ? (unsigned)e47.w : (unsigned)__ldg(ew + NC);
    const unsigned all_const = __reduce_and_sync(0xffffffffu, tab ? mask : 0xffffffffu);
    if (!__any_sync(0xffffffffu, tab)) return;
    float pre;
    int e_off = 0;
    if (S.mode == MODE_ICE1) {
        // N0 renormalised so that the mass integral equals QM (hydrometeors.py:1330-1339): n0 = p1 / F_9
        const float inv9 = 1.f / lt_col<NC, C9>(base, x, mask, all_const);
        e_off = __ldg(ew + C9);
        pre = inv9 * p1;
        if (tab) {
            if (!DOP2) {
                vn = make_float2(scale2(lt_col<NC, C7>(base, x, mask, all_const) * pre, __ldg(ew + C7) - e_off),
                                 scale2(lt_col<NC, C8>(base, x, mask, all_const) * pre, __ldg(ew + C8) - e_off));
            } else {
                const int nd = S.nd;
                const float lam1 = exp2f((float)u);
                const float *tb = S.table + (size_t)bl * nd * S.ncol;
                const float D0 = __ldg(S.aux), DL = __ldg(S.aux + nd - 1);
                const float c2 = 17.46f * exp2f(0.6357f * log2f(lam1));
                const float w0 = 490.6f * ex2f_approx(-20.78f * LOG2E * lam1 * D0) +
                                 c2 * __ldg(S.aux + 2 * nd) * ex2f_approx(-3.290f * LOG2E * lam1 * D0);
                const float wL = 490.6f * ex2f_approx(-20.78f * LOG2E * lam1 * DL) +
                                 c2 * __ldg(S.aux + 3 * nd - 1) * ex2f_approx(-3.290f * LOG2E * lam1 * DL);
                const float f0 = lt_scale(lt_col<NC, 0>(base, x, mask, all_const), __ldg(ew));
                const float f7 = lt_scale(lt_col<NC, C7>(base, x, mask, all_const), __ldg(ew + C7));
                const float n0 = pre / lt_scale(1.f, e_off);
                const float n_ = f0 - 0.5f * (__ldg(tb) * w0 + __ldg(tb + (size_t)(nd - 1) * S.ncol) * wL);
                const float v_ = f7 - 0.5f * (__ldg(tb + 7) * w0 + __ldg(tb + (size_t)(nd - 1) * S.ncol + 7) * wL);
                const float sc = 6.283185307179586f * n0 / S.dD_ref;
                vn = make_float2(v_ * sc, n_ * sc);
            }
        }
        pre *= w;
    } else {
        pre = p1 * w;
        if (S.mode == MODE_GEN12) {
            if (tab) vn = make_float2(scale2(lt_col<NC, C7>(base, x, mask, all_const) * p1, __ldg(ew + C7)),
                                      scale2(lt_col<NC, C8>(base, x, mask, all_const) * p1, __ldg(ew + C8)));
        } else if (DOP2 && NC == 8) {
            // trapezoid end corrections of Doppler scheme 2 (doppler2_moments)
            if (tab) {
                const int nd = S.nd;
                const float *tb = S.table + (size_t)bl * nd * S.ncol;
                const float a = -LOG2E * exp2f((float)u);
                float yF, yL;
                if (S.mode == MODE_GEO8) { yF = __ldg(S.aux); yL = (float)((double)yF + (double)(nd - 1) * S.dD_d); }
                else { yF = __ldg(S.aux + nd); yL = __ldg(S.aux + 2 * nd - 1); }
                const float wF = ex2f_approx(a * yF), wL = ex2f_approx(a * yL);
                const float f0 = lt_scale(lt_col<NC, 0>(base, x, mask, all_const), __ldg(ew));
                const float f7 = lt_scale(lt_col<NC, C7>(base, x, mask, all_const), __ldg(ew + C7));
                const float n_ = f0 - 0.5f * (__ldg(tb) * wF + __ldg(tb + (size_t)(nd - 1) * S.ncol) * wL);
                const float v_ = f7 - 0.5f * (__ldg(tb + 7) * wF + __ldg(tb + (size_t)(nd - 1) * S.ncol + 7) * wL);
                const float sc = 6.283185307179586f * p1 / S.dD_ref;
                vn = make_float2(v_ * sc, n_ * sc);
            }
        }
    }
    if (!tab) pre = 0.f;          // lanes without a table pair add exact zeros
    // the seven columns of get_pol_from_sz, one at a time
    {
        const float2 f = lt_col_pair<NC, 0>(base, x, mask, all_const);
        acc[0] += tab ? scale2(f.x * pre, e03.x - e_off) : 0.f;
        acc[1] += tab ? scale2(f.y * pre, e03.y - e_off) : 0.f;
    }
    {
        const float2 f = lt_col_pair<NC, 2>(base, x, mask, all_const);
        acc[2] += tab ? scale2(f.x * pre, e03.z - e_off) : 0.f;
        acc[3] += tab ? scale2(f.y * pre, e03.w - e_off) : 0.f;
    }
    {
        const float2 f = lt_col_pair<NC, 4>(base, x, mask, all_const);
        acc[4] += tab ? scale2(f.x * pre, e47.x - e_off) : 0.f;
        acc[5] += tab ? scale2(f.y * pre, e47.y - e_off) : 0.f;
    }
    acc[6] += tab ? scale2(lt_col<NC, 6>(base, x, mask, all_const) * pre, e47.z - e_off) : 0.f;
}

// log2 of a positive float32 as a float64: exponent + log2 of the mantissa (MUFU.LG2 on [1, 2): absolute error 2^-22)
__device__ __forceinline__ double log2_split(float x)
{
    int bits = __float_as_int(x);
    int e = (bits >> 23) - 127;
    if (e == -127) {                       // subnormal: renormalise
        x *= 16777216.f;
        bits = __float_as_int(x);
        e = (bits >> 23) - 127 - 24;
    }
    const float mant = __int_as_float((bits & 0x007fffff) | 0x3f800000);
    return (double)e + (double)__log2f(mant);
}

// 2^t for a float64 t with the integer part applied exactly
__device__ __forceinline__ float exp2_split(double t)
{
    const double ip = rint(t);
    return lt_scale(ex2f_approx((float)(t - ip)), (int)fmax(-300.0, fmin(300.0, ip)));
}

template <bool TWO_MOM, bool DOP2>
__global__ void __launch_bounds__(32 * FUSED_MAX_WARPS, FUSED_MIN_CTAS)
k_gate_fused(const __grid_constant__ DevConfig cfg, const ScanDev scan, const float *__restrict__ z_t,
             const float *__restrict__ fields, const __grid_constant__ FusedArgs A)
{
    constexpr int NV = TWO_MOM ? 15 : 8;
    extern __shared__ float s_part_raw[];                  // [FUSED_DBUF ? 2 : 1][W][FUSED_NPART][32]
    __shared__ int s_piece[2][2];
    __shared__ unsigned s_pairs[RADOP_NSPECIES_MAX];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, W = blockDim.x >> 5;
    const int ng = cfg.n_gates, n_sub = cfg.n_sub, n_spans = A.n_spans;
    const int n_pieces = cfg.n_rays * n_spans;
    const float qnan = __int_as_float(0x7fc00000);
    if (threadIdx.x < RADOP_NSPECIES_MAX) s_pairs[threadIdx.x] = 0u;
    auto fetch_piece = [&](int slot) {                     // thread 0: next piece from the global counter
        const int pc = atomicAdd(A.flags + 3, 1);
        const int r = pc / n_spans;
        s_piece[slot][0] = pc < n_pieces ? r : -1;
        s_piece[slot][1] = pc - r * n_spans;
    };
#if FUSED_DBUF
    if (threadIdx.x == 0) fetch_piece(0);
    __syncthreads();
#endif

    for (int k = 0;; ++k) {
#if FUSED_DBUF
        // Piece k uses row buffer k & 1 and piece slot k & 1.  The only barrier of the iteration is the one after the rows
        // are written; the reducing warp (k mod the working warps) adds them up while the others start piece k + 1 in the
        // OTHER buffer, and it has finished before it reaches the next barrier -- after which buffer k & 1 is written again.
        const int buf = k & 1;
        float (*s_part)[FUSED_NPART][32] = reinterpret_cast<float (*)[FUSED_NPART][32]>(s_part_raw) + (size_t)buf * W;
        const int ray = s_piece[buf][0], span = s_piece[buf][1];
        if (ray < 0) break;
        if (threadIdx.x == 0) fetch_piece(buf ^ 1);        // slot last read at the top of iteration k - 1, before its barrier
        const int reducer = k % min(W, n_sub);
#else
        float (*s_part)[FUSED_NPART][32] = reinterpret_cast<float (*)[FUSED_NPART][32]>(s_part_raw);
        if (threadIdx.x == 0) fetch_piece(0);
        __syncthreads();
        const int ray = s_piece[0][0], span = s_piece[0][1];
        if (ray < 0) break;
        const int reducer = 0;
#endif
        const int g = span * 32 + lane;
        const bool in_range = g < ng;

#pragma unroll 1
        for (int q = warp; q < n_sub; q += W) {
            const int rq = ray * n_sub + q;
            const size_t trow = ((size_t)__ldg(scan.ray_traj + ray) * cfg.n_vert + __ldg(scan.sub_jv + q)) * ng + (in_range ? g : 0);
            const float wq = (float)__ldg(scan.sub_weight + q);
            const double2 *cp = reinterpret_cast<const double2 *>(A.poly + rq);
            const double2 az = __ldg(cp);                        // sin, cos of the sub-beam azimuth
            float h_g = 0.f, pos0 = 0.f, pos1 = 0.f;
            CellGeom c;
            c.status = 1;
            if (in_range) {
                const float s_g = __ldg(scan.traj_s + trow);
                h_g = __ldg(scan.traj_h + trow);
                // --- geolocation: rotated-pole coordinates from the sub-beam's polynomials ---------------
                const double2 iv = __ldg(cp + 1);
                const double x = ((double)s_g - iv.x) * iv.y;
                const double rlat = poly_eval(cp + 2, x), rlon = poly_eval(cp + 2 + GEO_NC / 2, x);
                if (q == cfg.center_sub && A.lat) {
                    A.lat[(size_t)ray * ng + g] = poly_eval(cp + 2 + GEO_NC, x);
                    A.lon[(size_t)ray * ng + g] = poly_eval(cp + 2 + 3 * GEO_NC / 2, x);
                }
                // domain check on float64 coordinates against float32 corners (interpolation.py:572-575)
                if (rlon < (double)cfg.llc[0] || rlat < (double)cfg.llc[1] || rlon > (double)cfg.urc[0] ||
                    rlat > (double)cfg.urc[1])
                    atomicOr(A.flags, 1);
                // SWIG typemap casts the coordinates to float32 (interpolation_c.i:16)
                pos0 = div_const(__fsub_rn((float)rlat, cfg.llc[1]), cfg.res[1], cfg.rres[1]);
                pos1 = div_const(__fsub_rn((float)rlon, cfg.llc[0]), cfg.res[0], cfg.rres[0]);
                locate_cell(z_t, cfg.nz, cfg.ny, cfg.nx, pos0, pos1, h_g, cfg.z_strict != 0, c, cfg.lev_guess, cfg.z_tb);
            }
            const float mask_q = in_range ? (float)c.status : 0.f;
            const bool live = in_range && c.status == 0;
            float val[NV];
            if (live) {
                if (cfg.div_exact) interp_fused<2, TWO_MOM>(cfg, c, h_g, z_t, fields, val);
                else if (interp_fused<0, TWO_MOM>(cfg, c, h_g, z_t, fields, val))
                    interp_fused<1, TWO_MOM>(cfg, c, h_g, z_t, fields, val);
            } else {
#pragma unroll
                for (int j = 0; j < NV; ++j) val[j] = 0.f;
            }
            const float T = val[RADOP_VAR_T];
            // --- wind along the sub-beam (proj_vel, doppler_scatter.py:30-47; theta is the folded elevation) ---
            float wind = qnan, sinth = 0.f;
            if (live) {
                const float2 sc = __ldg(A.traj_sc + trow);
                // float32 products of float32 winds: the rounding (1e-7 of 50 m/s) is far below the spread of the
                // float32 sums over the sub-beams that follow
                wind = fmaf(fmaf(val[RADOP_VAR_U], (float)az.x, val[RADOP_VAR_V] * (float)az.y), sc.y, val[RADOP_VAR_W] * sc.x);
                sinth = sc.x;
            }
            // --- species: PSD slope in the log domain, table bin, table evaluation, sums over the species ---
            float acc[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            float vs = 0.f, ns = 0.f;
            bool ice_here = false;
            float2 ice_vn = make_float2(0.f, 0.f);
#pragma unroll 1
            for (int s = 0; s < cfg.n_species; ++s) {
                const SpeciesDev &S = cfg.sp[s];
                const FusedSpecies &F = A.fs[s];
                const float QMf = pick_var<NV>(val, S.qm_var);
                const bool have = live && QMf > 0.f;
                const unsigned who = __ballot_sync(0xffffffffu, have);
                if (!who) continue;
                if (lane == 0) atomicAdd(&s_pairs[s], (unsigned)__popc(who));
                double u = 0.0;
                float p1 = 0.f, n0 = 0.f;
                int bl = 0;
                if (have) {
                    if (S.psd == RADOP_PSD_2MOM) {
                        // _Hydrometeor.set_psd, hydrometeors.py:224-255 (float64 as in k_interp_psd: N0 spans 20 orders
                        // of magnitude and the two-moment KDP has gates where species cancel to 1e-3 of their size)
                        const double QN = (double)pick_var<NV>(val, S.qn_var);
                        const double xm = fmin(fmax((double)QMf / (QN + 2.220446049250313e-16), S.x_min), S.x_max);
                        const double ll0 = S.c_inv * log(S.lambda_factor * xm);
                        const double logn0 = log(fabs(S.c_q * QN)) + S.e_q * ll0 - S.c_ka;
                        const double dn0 = (S.c_q * QN) < 0.0 ? -exp(logn0) : exp(logn0);
                        const double loglam = ll0 - S.c_kb;
                        u = loglam * 1.4426950408889634;
                        n0 = (float)dn0;
                        p1 = n0;
                        if (!DOP2 && !S.numeric_vn) {
                            vs += (float)(S.c_v * dn0 * exp(S.e_v * loglam));
                            ns += (float)QN;
                        }
                    } else {
                        const double lq = log2_split(QMf);
                        if (S.psd == RADOP_PSD_1MOM_FIXED_N0) {
                            // lambda = (lambda_factor / QM)^(1/(4+mu)), hydrometeors.py:761-772, 1039-1049
                            u = F.uk * (F.u0 - lq);
                            n0 = F.n0;
                            p1 = n0;
                        } else if (S.psd == RADOP_PSD_1MOM_SNOW) {
                            // N0 = 13.5 * 5.65e5 * exp(-0.107 (T - 273.15)) / 1000, hydrometeors.py:894-907
                            const double l2n0 = fma(F.nb, (double)T, F.na);
                            u = F.uk * (F.u0 + l2n0 - lq);
                            n0 = exp2_split(l2n0);
                            p1 = n0;
                        } else {
                            // 1-mom ice: lambda = (Q2 / QM)^(1/(b-2)) with the Field et al. (2005) second moment,
                            // hydrometeors.py:1277-1299, 1320-1339; log2 Q2 = (log2 QM' - log2(10) A(T)) / B(T)
                            const double Tc = (double)T - 273.15;
                            const double a = (5.065339 - 3.032362 * 3.0 + 0.312550 * 9.0 - 0.015952 * 27.0) +
                                             Tc * ((-0.062659 + 0.029469 * 3.0 + 0.003199 * 9.0) + Tc * (-0.000285 + 0.000204 * 3.0));
                            const double b = (0.476221 + 0.165977 * 3.0 + 0.060366 * 9.0 - 0.003577 * 27.0) +
                                             Tc * ((-0.015896 + 0.007468 * 3.0 + 0.000594 * 9.0) + Tc * (-0.000141 + 0.000079 * 3.0));
                            double rb = (double)(1.f / (float)b);
                            rb = rb * (2.0 - b * rb);
                            rb = rb * (2.0 - b * rb);
                            const double l2q2 = (lq - F.u0 - 3.321928094887362 * a) * rb;
                            u = F.uk * (l2q2 - lq) - 9.965784284662087;          // lambda / 1000
                            p1 = QMf;
                        }
                        if (!DOP2 && !S.numeric_vn) {
                            // _Hydrometeor.integrate_V, hydrometeors.py:178-199
                            const float fu = (float)u;
                            vs += F.cv * n0 * ex2f_approx(F.ev * fu);
                            ns += F.cn * n0 * ex2f_approx(F.en * fu);
                        }
                    }
                    const int ie = (int)__ldg(A.traj_ie + (size_t)s * A.traj_n + trow);
                    const int it = lut_bin(T, S.t_min, S.t_step, S.t_rstep, S.nt);
                    bl = ie * S.nt + it;
                }
                float2 vn = make_float2(0.f, 0.f);
                // columns tabulated (lt_columns): 7 (8 with the rcs-weighted fall speed of Doppler scheme 2), 10 for 1-mom
                // ice, 9 for a species with numeric fall-speed moments in scheme 1
                constexpr int NC_A = DOP2 ? 8 : 7, NC_B = (TWO_MOM && !DOP2) ? 9 : 10;
                if (S.lt_nc == NC_A) species_tab<NC_A, DOP2>(cfg, A, S, have, bl, u, p1, wq, lane, acc, vn);
                else species_tab<NC_B, DOP2>(cfg, A, S, have, bl, u, p1, wq, lane, acc, vn);
                if (have) {
                    if (DOP2) { vs += vn.x; ns += vn.y; }
                    else if (s == cfg.ice_numeric) { ice_here = true; ice_vn = vn; }
                }
            }
            // --- radial velocity of the sub-beam (doppler_scatter.py:313-333) and this warp's running sums ---
            {
                float rvw = 0.f, ww = 0.f;
                if (live) {
                    const float rv = wind - (vs / ns) * sinth;     // 0/0 -> NaN where no hydrometeor (Q7)
                    if (!isnan(rv)) { rvw = rv * wq; ww = wq; }
                }
                float (*row)[32] = s_part[warp];
                if (q == warp) {
#pragma unroll
                    for (int cc = 0; cc < 7; ++cc) row[cc][lane] = acc[cc];
                    row[7][lane] = rvw; row[8][lane] = ww; row[9][lane] = mask_q;
                } else {
#pragma unroll
                    for (int cc = 0; cc < 7; ++cc) row[cc][lane] += acc[cc];
                    row[7][lane] += rvw; row[8][lane] += ww; row[9][lane] += mask_q;
                }
            }
            // --- numeric fall-speed moments (IceParticle.integrate_V sums over ALL gates of a sub-beam and the
            //     total lands on its first gate with ice, hydrometeors.py:1256-1275): partial sums of this span, folded into that
            //     gate by k_ray_epilogue ---
            if (A.ice_part) {
                const unsigned icy = __ballot_sync(0xffffffffu, ice_here);
                float v = ice_vn.x, n = ice_vn.y;
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) {
                    v += __shfl_xor_sync(0xffffffffu, v, d);
                    n += __shfl_xor_sync(0xffffffffu, n, d);
                }
                const int first = icy ? __ffs(icy) - 1 : 0;
                const float w_f = __shfl_sync(0xffffffffu, wind, first), s_f = __shfl_sync(0xffffffffu, sinth, first);
                const float vs_f = __shfl_sync(0xffffffffu, vs, first), ns_f = __shfl_sync(0xffffffffu, ns, first);
                if (lane == 0) {
                    float4 *dst = A.ice_part + ((size_t)rq * n_spans + span) * 2;
                    dst[0] = make_float4(v, n, __int_as_float(icy ? span * 32 + first : -1), w_f);
                    dst[1] = make_float4(s_f, vs_f, ns_f, wq);
                }
            }
        }
        __syncthreads();
        if (warp == reducer && in_range) {
            float t[FUSED_NPART];
#pragma unroll
            for (int i = 0; i < FUSED_NPART; ++i) t[i] = s_part[0][i][lane];
            const int nw = min(W, n_sub);
            for (int w = 1; w < nw; ++w)
#pragma unroll
                for (int i = 0; i < FUSED_NPART; ++i) t[i] += s_part[w][i][lane];
            const size_t rg = (size_t)ray * ng + g;
            float4 *dst = reinterpret_cast<float4 *>(A.comb + rg * 8);
            dst[0] = make_float4(t[0], t[1], t[2], t[3]);
            dst[1] = make_float4(t[4], t[5], t[6], t[7]);
            A.wsum[rg] = t[8];
            float mk = t[9] / (float)n_sub;
            if (mk > -1.f && mk <= 0.f) mk = 0.f;
            A.mask_out[rg] = mk;
        }
    }
    __syncthreads();
    if (threadIdx.x < cfg.n_species && s_pairs[threadIdx.x]) atomicAdd(A.n_pairs + threadIdx.x, (unsigned long long)s_pairs[threadIdx.x]);
}
