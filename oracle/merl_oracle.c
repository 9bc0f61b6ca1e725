/*
 * merl_oracle.c -- CPU ORACLE for the measured-BRDF hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.  The product (libmerl_b200) never links or calls it.
 *
 * PARITY UNPINNED.  /root/reference holds no plugin source (both submodules are empty
 * gitlinks: /root/reference/.SUBMODULES.json:8-23, /root/reference/README.md:1 is the
 * only description).  This file therefore restates the PUBLIC MERL lookup algorithm
 * (Matusik et al., "BRDFRead.cpp", distributed with the MERL BRDF database) which any
 * MERL plugin must agree with for the file format, the half/difference conversion, the
 * three bin-index formulas, the flat index and the channel scales.
 * Everything the fork could have chosen differently is a named
 * switch or a documented assumption below.  No golden vector exists to check against.
 *
 * Numeric convention ("canonical precision"): query directions arrive
 * as fp32 (Mitsuba's Vector3f), are promoted to double, and every step runs in double
 * exactly as the public reader does.  A long-double twin of the bin computation exists
 * so tests can detect queries whose bin is unstable *in the reference arithmetic
 * itself* (those are excluded from bit-exact comparison and counted).
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off -fopenmp).
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* ------------------------------------------------------------------------------------
 * Table description shared by every entry point (mirrored by ctypes in merl_oracle.py)
 * ---------------------------------------------------------------------------------- */
typedef struct {
    int32_t nh, nd, np;     /* bins in theta_half / theta_diff / phi_diff; MERL = 90/90/180 */
    int32_t theta_h_sqrt;   /* 1: MERL nonlinear theta_half mapping; 0: linear (customized) */
    int32_t ni;             /* sampler: incident-elevation slices, power of two            */
    int32_t clamp_negative; /* App. B2: negative (invalid) cells evaluate to 0             */
    double scale[3];        /* RED/GREEN/BLUE_SCALE = 1/1500, 1.15/1500, 1.66/1500          */
    double spec_weight;     /* sampler: probability of the tabulated half-vector lobe      */
    const double *raw;      /* planar [3][nh*nd*np] exactly as stored in the .binary file   */
    const float *cdf;       /* sampler: [ni][nh+1] row CDFs over theta_half bins (fp32)     */
    const double *dens;     /* optional cache of spec_density(), [ni][nh]; see orc_build_dens */
} orc_table_t;

static inline double dot3(const double a[3], const double b[3]) {
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}

static inline double clamp_unit(double x) { /* acos() guard: the public reader has none and
                                               would hit int(NaN); we define that case. */
    return x > 1.0 ? 1.0 : (x < -1.0 ? -1.0 : x);
}

/* Rodrigues rotation of v about a unit axis (public reader: rotate_vector). */
static void rotate_about(const double v[3], const double axis[3], double angle, double out[3]) {
    const double c = cos(angle), s = sin(angle);
    const double k = dot3(axis, v) * (1.0 - c);
    const double cr[3] = {axis[1] * v[2] - axis[2] * v[1], axis[2] * v[0] - axis[0] * v[2],
                          axis[0] * v[1] - axis[1] * v[0]};
    for (int a = 0; a < 3; ++a) out[a] = v[a] * c + axis[a] * k + cr[a] * s;
}

/* Cartesian (in, out) -> (theta_half, fi_half, theta_diff, fi_diff).
 * Public reader: std_coords_to_half_diff_coords, from the point where it holds the two
 * direction vectors.  Directions are used as given (assumption
 * B9 in DESIGN.md: the plugin hands wi/wo straight to the lookup). */
void orc_half_diff(const double in[3], const double out[3], double hd[4]) {
    double h[3] = {in[0] + out[0], in[1] + out[1], in[2] + out[2]};
    const double len = sqrt(dot3(h, h));
    if (len > 0.0) { h[0] /= len; h[1] /= len; h[2] /= len; }
    const double theta_half = acos(clamp_unit(h[2]));
    const double fi_half = atan2(h[1], h[0]);
    static const double normal[3] = {0.0, 0.0, 1.0}, binormal[3] = {0.0, 1.0, 0.0};
    double tmp[3], d[3];
    rotate_about(in, normal, -fi_half, tmp);
    rotate_about(tmp, binormal, -theta_half, d);
    hd[0] = theta_half;
    hd[1] = fi_half;
    hd[2] = acos(clamp_unit(d[2]));
    hd[3] = atan2(d[1], d[0]);
}

/* Bin-index functions, generalised from the MERL constants
 * 90/90/180 to (nh, nd, np) for the customized_measurment layout (assumption B7). */
int orc_theta_half_index(double theta_half, int nh, int sqrt_map) {
    if (theta_half <= 0.0) return 0;
    double t = theta_half / (M_PI / 2.0) * nh;
    if (sqrt_map) t = sqrt(t * nh);
    int i = (int)t;
    if (i < 0) i = 0;
    if (i >= nh) i = nh - 1;
    return i;
}
int orc_theta_diff_index(double theta_diff, int nd) {
    int i = (int)(theta_diff / (M_PI * 0.5) * nd);
    return i < 0 ? 0 : (i < nd - 1 ? i : nd - 1);
}
int orc_phi_diff_index(double phi_diff, int np) {
    if (phi_diff < 0.0) phi_diff += M_PI;
    int i = (int)(phi_diff / M_PI * (2 * np) / 2);
    return i < 0 ? 0 : (i < np - 1 ? i : np - 1);
}

static inline int flat_index(const orc_table_t *t, int ih, int id, int ip) {
    return ip + id * t->np + ih * t->np * t->nd;
}

static int bins_from_dirs(const orc_table_t *t, const double in[3], const double out[3], int b[3]) {
    double hd[4];
    orc_half_diff(in, out, hd);
    b[0] = orc_theta_half_index(hd[0], t->nh, t->theta_h_sqrt);
    b[1] = orc_theta_diff_index(hd[2], t->nd);
    b[2] = orc_phi_diff_index(hd[3], t->np);
    return flat_index(t, b[0], b[1], b[2]);
}

/* long-double twin, used only to flag queries whose bin is unstable in the reference
 * arithmetic itself (see header). */
static int bins_from_dirs_ld(const orc_table_t *t, const double in_[3], const double out_[3]) {
    typedef long double R;
    const R PI = 3.14159265358979323846264338327950288L;
    R in[3] = {in_[0], in_[1], in_[2]}, out[3] = {out_[0], out_[1], out_[2]};
    R h[3] = {in[0] + out[0], in[1] + out[1], in[2] + out[2]};
    R len = sqrtl(h[0] * h[0] + h[1] * h[1] + h[2] * h[2]);
    if (len > 0) { h[0] /= len; h[1] /= len; h[2] /= len; }
    R hz = h[2] > 1 ? 1 : (h[2] < -1 ? -1 : h[2]);
    R th = acosl(hz), fh = atan2l(h[1], h[0]);
    /* rotate about z by -fh, then about y by -th */
    R c = cosl(-fh), s = sinl(-fh);
    R t0 = in[0] * c - in[1] * s, t1 = in[1] * c + in[0] * s, t2 = in[2];
    c = cosl(-th); s = sinl(-th);
    R d0 = t0 * c + t2 * s, d1 = t1, d2 = t2 * c - t0 * s;
    R dz = d2 > 1 ? 1 : (d2 < -1 ? -1 : d2);
    R td = acosl(dz), fd = atan2l(d1, d0);
    int ih, id, ip;
    if (th <= 0) ih = 0;
    else {
        R x = th / (PI / 2) * t->nh;
        if (t->theta_h_sqrt) x = sqrtl(x * t->nh);
        ih = (int)x; if (ih < 0) ih = 0; if (ih >= t->nh) ih = t->nh - 1;
    }
    id = (int)(td / (PI * 0.5L) * t->nd); id = id < 0 ? 0 : (id < t->nd - 1 ? id : t->nd - 1);
    if (fd < 0) fd += PI;
    ip = (int)(fd / PI * t->np); ip = ip < 0 ? 0 : (ip < t->np - 1 ? ip : t->np - 1);
    return flat_index(t, ih, id, ip);
}

/* Batched bins.  wi/wo are SoA fp32 (x[n], y[n], z[n]) as a Dr.Jit Vector3f would be.
 * Outputs (any may be NULL): ih, id, ip, flat idx, and `stable` = 1 where the long-double
 * twin lands in the same cell. */
void orc_bins(const orc_table_t *t, const float *wix, const float *wiy, const float *wiz,
              const float *wox, const float *woy, const float *woz, int64_t n, int32_t *ih,
              int32_t *id, int32_t *ip, int32_t *idx, uint8_t *stable) {
#pragma omp parallel for schedule(static)
    for (int64_t q = 0; q < n; ++q) {
        const double in[3] = {wix[q], wiy[q], wiz[q]}, out[3] = {wox[q], woy[q], woz[q]};
        int b[3];
        const int f = bins_from_dirs(t, in, out, b);
        if (ih) ih[q] = b[0];
        if (id) id[q] = b[1];
        if (ip) ip[q] = b[2];
        if (idx) idx[q] = f;
        if (stable) stable[q] = (uint8_t)(bins_from_dirs_ld(t, in, out) == f);
    }
}

/* Single-precision twin of bins_from_dirs: every step of SURVEY.md App. A (half vector, acos/atan2,
 * the two Rodrigues rotations, the three index formulas) carried out in `float` with the float libm,
 * as a Mitsuba 3 non-double variant (scalar_rgb, llvm_rgb, cuda_rgb: Float = float) would compute it
 * if the fork's plugin is a straight port of the public reader.  It is NOT
 * what libmerl_b200 matches -- that is the double path above -- it exists to MEASURE how often the two
 * precisions land in different cells, i.e. how much of "bit-exact bins" hinges on assumption B1. */
static void rotate_about_f(const float v[3], const float axis[3], float angle, float out[3]) {
    const float c = cosf(angle), s = sinf(angle);
    const float k = (axis[0] * v[0] + axis[1] * v[1] + axis[2] * v[2]) * (1.0f - c);
    const float cr[3] = {axis[1] * v[2] - axis[2] * v[1], axis[2] * v[0] - axis[0] * v[2],
                         axis[0] * v[1] - axis[1] * v[0]};
    for (int a = 0; a < 3; ++a) out[a] = v[a] * c + axis[a] * k + cr[a] * s;
}
static int bins_from_dirs_f32(const orc_table_t *t, const float in[3], const float out[3]) {
    const float PI_F = (float)M_PI;
    float h[3] = {in[0] + out[0], in[1] + out[1], in[2] + out[2]};
    const float len = sqrtf(h[0] * h[0] + h[1] * h[1] + h[2] * h[2]);
    if (len > 0.0f) { h[0] /= len; h[1] /= len; h[2] /= len; }
    const float hz = h[2] > 1.0f ? 1.0f : (h[2] < -1.0f ? -1.0f : h[2]);
    const float theta_half = acosf(hz), fi_half = atan2f(h[1], h[0]);
    static const float normal[3] = {0.0f, 0.0f, 1.0f}, binormal[3] = {0.0f, 1.0f, 0.0f};
    float tmp[3], d[3];
    rotate_about_f(in, normal, -fi_half, tmp);
    rotate_about_f(tmp, binormal, -theta_half, d);
    const float dz = d[2] > 1.0f ? 1.0f : (d[2] < -1.0f ? -1.0f : d[2]);
    const float theta_diff = acosf(dz);
    float fi_diff = atan2f(d[1], d[0]);
    int ih, id, ip;
    if (theta_half <= 0.0f) ih = 0;
    else {
        float x = theta_half / (PI_F / 2.0f) * (float)t->nh;
        if (t->theta_h_sqrt) x = sqrtf(x * (float)t->nh);
        ih = (int)x; if (ih < 0) ih = 0; if (ih >= t->nh) ih = t->nh - 1;
    }
    id = (int)(theta_diff / (PI_F * 0.5f) * (float)t->nd);
    id = id < 0 ? 0 : (id < t->nd - 1 ? id : t->nd - 1);
    if (fi_diff < 0.0f) fi_diff += PI_F;
    ip = (int)(fi_diff / PI_F * (float)(2 * t->np) / 2.0f);
    ip = ip < 0 ? 0 : (ip < t->np - 1 ? ip : t->np - 1);
    return flat_index(t, ih, id, ip);
}
void orc_bins_f32(const orc_table_t *t, const float *wix, const float *wiy, const float *wiz,
                  const float *wox, const float *woy, const float *woz, int64_t n, int32_t *idx) {
#pragma omp parallel for schedule(static)
    for (int64_t q = 0; q < n; ++q) {
        const float in[3] = {wix[q], wiy[q], wiz[q]}, out[3] = {wox[q], woy[q], woz[q]};
        idx[q] = bins_from_dirs_f32(t, in, out);
    }
}

/* Table read + scale (public reader: lookup_brdf_val's tail). */
static inline void cell_rgb(const orc_table_t *t, int idx, double rgb[3]) {
    const int64_t N = (int64_t)t->nh * t->nd * t->np;
    for (int c = 0; c < 3; ++c) {
        double v = t->raw[idx + c * N] * t->scale[c];
        if (t->clamp_negative && v < 0.0) v = 0.0;
        rgb[c] = v;
    }
}

/* The public reader's user-facing entry: four angles in, RGB out (no cosine factor). */
void orc_lookup_brdf_val(const orc_table_t *t, double theta_in, double fi_in, double theta_out,
                         double fi_out, double rgb[3]) {
    const double in[3] = {cos(fi_in) * sin(theta_in), sin(fi_in) * sin(theta_in), cos(theta_in)};
    const double out[3] = {cos(fi_out) * sin(theta_out), sin(fi_out) * sin(theta_out),
                           cos(theta_out)};
    int b[3];
    const int idx = bins_from_dirs(t, in, out, b);
    const int64_t N = (int64_t)t->nh * t->nd * t->np;
    for (int c = 0; c < 3; ++c) rgb[c] = t->raw[idx + c * N] * t->scale[c];
}

/* ------------------------------------------------------------------------------------
 * Sampler geometry shared by pdf / sample / CDF build (assumption B5):
 * theta_half bin edges theta_i, M_i = 1 - cos(theta_i), solid-angle-uniform inside a bin.
 * ---------------------------------------------------------------------------------- */
static double theta_h_edge(const orc_table_t *t, int i) {
    const double u = (double)i / t->nh;
    return (M_PI / 2.0) * (t->theta_h_sqrt ? u * u : u);
}
static double m_edge(const orc_table_t *t, int i) { /* 1 - cos(theta_i) without cancellation */
    const double s = sin(0.5 * theta_h_edge(t, i));
    return 2.0 * s * s;
}
void orc_m_edges(const orc_table_t *t, double *m) {
    for (int i = 0; i <= t->nh; ++i) m[i] = m_edge(t, i);
}
static inline int incident_slice(const orc_table_t *t, double wiz) {
    int j = (int)(wiz * t->ni);
    return j < 0 ? 0 : (j < t->ni ? j : t->ni - 1);
}
static inline double spec_density(const orc_table_t *t, int j, int ih) {
    /* density of h per solid angle, divided by 4 (the 1/(4 wo.h) Jacobian's constant). */
    if (t->dens) return t->dens[(size_t)j * t->nh + ih];
    const float *row = t->cdf + (size_t)j * (t->nh + 1);
    const double P = (double)row[ih + 1] - (double)row[ih];
    return P / (8.0 * M_PI * (m_edge(t, ih + 1) - m_edge(t, ih)));
}

void orc_build_dens(const orc_table_t *t, double *dens_out) {
    orc_table_t u = *t;
    u.dens = NULL;
    for (int j = 0; j < t->ni; ++j)
        for (int i = 0; i < t->nh; ++i) dens_out[(size_t)j * t->nh + i] = spec_density(&u, j, i);
}

/* eval (with cos theta_o, Mitsuba convention, B3) and pdf of the two-lobe sampler. */
static void eval_pdf_one(const orc_table_t *t, const double wi[3], const double wo[3],
                         double rgb[3], double *pdf, int *idx_out) {
    rgb[0] = rgb[1] = rgb[2] = 0.0;
    if (pdf) *pdf = 0.0;
    if (idx_out) *idx_out = -1;
    if (!(wi[2] > 0.0) || !(wo[2] > 0.0)) return;
    int b[3];
    const int idx = bins_from_dirs(t, wi, wo, b);
    if (idx_out) *idx_out = idx;
    cell_rgb(t, idx, rgb);
    for (int c = 0; c < 3; ++c) rgb[c] *= wo[2];
    if (pdf) {
        double p = (1.0 - t->spec_weight) * wo[2] / M_PI;
        if (t->cdf) {
            const double s[3] = {wi[0] + wo[0], wi[1] + wo[1], wi[2] + wo[2]};
            const double cos_d = dot3(wi, s) / sqrt(dot3(s, s));
            if (cos_d > 0.0)
                p += t->spec_weight * spec_density(t, incident_slice(t, wi[2]), b[0]) / cos_d;
        }
        *pdf = p;
    }
}

/* Batched eval / pdf / eval_pdf (rows A4, A5).  rgb is planar [3][n]; NULL outputs skipped. */
void orc_eval_pdf(const orc_table_t *t, const float *wix, const float *wiy, const float *wiz,
                  const float *wox, const float *woy, const float *woz, int64_t n, double *rgb,
                  double *pdf, int32_t *idx) {
#pragma omp parallel for schedule(static)
    for (int64_t q = 0; q < n; ++q) {
        const double wi[3] = {wix[q], wiy[q], wiz[q]}, wo[3] = {wox[q], woy[q], woz[q]};
        double c[3], p;
        int f;
        eval_pdf_one(t, wi, wo, c, &p, &f);
        if (rgb) { rgb[q] = c[0]; rgb[n + q] = c[1]; rgb[2 * n + q] = c[2]; }
        if (pdf) pdf[q] = p;
        if (idx) idx[q] = f;
    }
}

/* Mitsuba 3's concentric square->disk map followed by the cosine-hemisphere lift. */
static void cosine_hemisphere(double ux, double uy, double wo[3]) {
    const double x = 2.0 * ux - 1.0, y = 2.0 * uy - 1.0;
    double r, phi;
    if (x == 0.0 && y == 0.0) { r = 0.0; phi = 0.0; }
    else if (fabs(x) < fabs(y)) { r = y; phi = 0.5 * M_PI - 0.25 * M_PI * (x / y); }
    else { r = x; phi = 0.25 * M_PI * (y / x); }
    wo[0] = r * cos(phi);
    wo[1] = r * sin(phi);
    const double z2 = 1.0 - wo[0] * wo[0] - wo[1] * wo[1];
    wo[2] = z2 > 0.0 ? sqrt(z2) : 0.0;
}

/* Draw wo (row A6).  lobe: 0 = cosine hemisphere, 1 = tabulated half-vector lobe.
 * The lobe test and the CDF-bin search compare fp32 numbers, so they are exact. */
static int draw_direction(const orc_table_t *t, const double wi[3], float u1, float ux, float uy,
                          double wo[3]) {
    const int lobe = (u1 < (float)t->spec_weight) ? 1 : 0;
    if (!lobe) { cosine_hemisphere(ux, uy, wo); return 0; }
    const int j = incident_slice(t, wi[2]);
    const float *row = t->cdf + (size_t)j * (t->nh + 1);
    int lo = 0, hi = t->nh; /* largest i in [0, nh-1] with row[i] <= ux */
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (row[mid] <= ux) lo = mid; else hi = mid;
    }
    const double width = (double)row[lo + 1] - (double)row[lo];
    const double f = width > 0.0 ? ((double)ux - (double)row[lo]) / width : 0.5;
    const double m0 = m_edge(t, lo), m1 = m_edge(t, lo + 1);
    const double m = m0 + f * (m1 - m0);
    const double ct = 1.0 - m, st = sqrt(m * (2.0 - m));
    const double phi = 2.0 * M_PI * (double)uy;
    const double h[3] = {st * cos(phi), st * sin(phi), ct};
    const double c = dot3(wi, h);
    for (int a = 0; a < 3; ++a) wo[a] = 2.0 * c * h[a] - wi[a];
    return 1;
}

/* Batched sample.  Outputs: wo planar [3][n] (double), pdf[n], weight planar [3][n], lobe[n].
 * pdf/weight are eval_pdf at the *fp32-rounded* wo, because fp32 is what the plugin
 * returns and what every later eval sees. */
void orc_sample(const orc_table_t *t, const float *wix, const float *wiy, const float *wiz,
                const float *u1, const float *u2x, const float *u2y, int64_t n, double *wo_out,
                double *pdf_out, double *weight_out, int32_t *lobe_out) {
#pragma omp parallel for schedule(static)
    for (int64_t q = 0; q < n; ++q) {
        const double wi[3] = {wix[q], wiy[q], wiz[q]};
        double wo[3] = {0, 0, 0}, w[3] = {0, 0, 0}, p = 0.0;
        int lobe = 0;
        if (wi[2] > 0.0) {
            lobe = draw_direction(t, wi, u1[q], u2x[q], u2y[q], wo);
            const double wof[3] = {(float)wo[0], (float)wo[1], (float)wo[2]};
            double f[3];
            eval_pdf_one(t, wi, wof, f, &p, NULL);
            if (p > 0.0) for (int c = 0; c < 3; ++c) w[c] = f[c] / p;
            else { p = 0.0; }
        }
        if (wo_out) { wo_out[q] = wo[0]; wo_out[n + q] = wo[1]; wo_out[2 * n + q] = wo[2]; }
        if (pdf_out) pdf_out[q] = p;
        if (weight_out) { weight_out[q] = w[0]; weight_out[n + q] = w[1]; weight_out[2 * n + q] = w[2]; }
        if (lobe_out) lobe_out[q] = lobe;
    }
}

/* Row CDFs of the half-vector lobe.  For slice j the incident direction
 * is wi = (sqrt(1-c^2), 0, c), c = (j+0.5)/ni, rounded to fp32; bin i's mass is the
 * luminance of f*cos integrated over the bin's half vectors with the reflection Jacobian
 * 4 (wi.h), by an nm x nq midpoint rule; a `floor_frac` share of solid-angle-uniform mass
 * keeps every bin reachable.  cdf_out: [ni][nh+1] fp32, mass_out (optional): [ni][nh] double. */
void orc_build_cdf(const orc_table_t *t, int nm, int nq, double floor_frac, float *cdf_out,
                   double *mass_out) {
    const int nh = t->nh;
    double *medge = (double *)malloc(sizeof(double) * (nh + 1));
    orc_m_edges(t, medge);
#pragma omp parallel for schedule(dynamic, 1)
    for (int j = 0; j < t->ni; ++j) {
        const double c0 = (j + 0.5) / t->ni;
        const double wi[3] = {(float)sqrt(1.0 - c0 * c0), 0.0, (float)c0};
        double *W = (double *)malloc(sizeof(double) * nh);
        double total = 0.0, msum = medge[nh] - medge[0];
        for (int i = 0; i < nh; ++i) {
            double acc = 0.0;
            for (int m = 0; m < nm; ++m) {
                const double mm = medge[i] + (m + 0.5) / nm * (medge[i + 1] - medge[i]);
                const double ct = 1.0 - mm, st = sqrt(mm * (2.0 - mm));
                for (int q = 0; q < nq; ++q) {
                    const double phi = 2.0 * M_PI * (q + 0.5) / nq;
                    const double h[3] = {st * cos(phi), st * sin(phi), ct};
                    const double c = dot3(wi, h);
                    if (!(c > 0.0)) continue;
                    const double wo[3] = {(float)(2.0 * c * h[0] - wi[0]),
                                          (float)(2.0 * c * h[1] - wi[1]),
                                          (float)(2.0 * c * h[2] - wi[2])};
                    if (!(wo[2] > 0.0)) continue;
                    double f[3];
                    eval_pdf_one(t, wi, wo, f, NULL, NULL);
                    acc += (0.2126 * f[0] + 0.7152 * f[1] + 0.0722 * f[2]) * 4.0 * c;
                }
            }
            W[i] = acc / (nm * nq) * 2.0 * M_PI * (medge[i + 1] - medge[i]);
            total += W[i];
        }
        float *row = cdf_out + (size_t)j * (nh + 1);
        double run = 0.0;
        row[0] = 0.0f;
        for (int i = 0; i < nh; ++i) {
            const double uni = (medge[i + 1] - medge[i]) / msum;
            const double P = total > 0.0 ? (1.0 - floor_frac) * W[i] / total + floor_frac * uni : uni;
            if (mass_out) mass_out[(size_t)j * nh + i] = P;
            run += P;
            row[i + 1] = (float)(run > 1.0 ? 1.0 : run);
        }
        row[nh] = 1.0f;
        free(W);
    }
    free(medge);
}

/* Backward (row A10): d loss / d T'[cell][c] where T' is the prescaled, clamped table the
 * device holds and f_c = T'_c * cos(theta_o).  Clamped (invalid) cells get no gradient.
 * grad_table: [nh*nd*np][4] doubles (4th lane unused, mirrors the device float4 layout). */
void orc_backward(const orc_table_t *t, const float *wix, const float *wiy, const float *wiz,
                  const float *wox, const float *woy, const float *woz, const float *gr,
                  const float *gg, const float *gb, int64_t n, double *grad_table) {
    const int64_t N = (int64_t)t->nh * t->nd * t->np;
#pragma omp parallel for schedule(static)
    for (int64_t q = 0; q < n; ++q) {
        const double wi[3] = {wix[q], wiy[q], wiz[q]}, wo[3] = {wox[q], woy[q], woz[q]};
        if (!(wi[2] > 0.0) || !(wo[2] > 0.0)) continue;
        int b[3];
        const int idx = bins_from_dirs(t, wi, wo, b);
        const double g[3] = {gr[q], gg[q], gb[q]};
        for (int c = 0; c < 3; ++c) {
            if (t->clamp_negative && t->raw[idx + c * N] * t->scale[c] < 0.0) continue;
            const double v = g[c] * wo[2];
#pragma omp atomic
            grad_table[(size_t)idx * 4 + c] += v;
        }
    }
}

/* ------------------------------------------------------------------------------------
 * MERL .binary I/O (row A8): int32 dims[3], then 3*n doubles, planar R,G,B.
 * ---------------------------------------------------------------------------------- */
int orc_read_merl(const char *path, int32_t dims[3], double **data_out) {
    FILE *f = fopen(path, "rb");
    if (!f) return -1;
    if (fread(dims, sizeof(int32_t), 3, f) != 3) { fclose(f); return -2; }
    const int64_t n = (int64_t)dims[0] * dims[1] * dims[2];
    if (n <= 0 || n > (1 << 28)) { fclose(f); return -3; }
    double *d = (double *)malloc(sizeof(double) * 3 * n);
    if (fread(d, sizeof(double), 3 * n, f) != (size_t)(3 * n)) { free(d); fclose(f); return -4; }
    fclose(f);
    *data_out = d;
    return 0;
}
int orc_write_merl(const char *path, const int32_t dims[3], const double *data) {
    FILE *f = fopen(path, "wb");
    if (!f) return -1;
    const int64_t n = (int64_t)dims[0] * dims[1] * dims[2];
    int ok = fwrite(dims, sizeof(int32_t), 3, f) == 3 &&
             fwrite(data, sizeof(double), 3 * n, f) == (size_t)(3 * n);
    fclose(f);
    return ok ? 0 : -2;
}
void orc_free(void *p) { free(p); }

int orc_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void orc_set_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n > 0 ? n : 1);
#else
    (void)n;
#endif
}
