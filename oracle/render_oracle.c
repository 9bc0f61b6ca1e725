/*
 * render_oracle.c -- CPU ORACLE of the wavefront path tracer harness.  TEST INFRASTRUCTURE ONLY
 * (same rules as merl_oracle.c: only tests/, smoke() and bench.py's CPU legs may load it).
 *
 * PARITY UNPINNED: the reference holds no source (see merl_oracle.c).  This file restates, as a
 * plain scalar loop in double precision, the algorithm that include/merl_b200.h documents for
 * merl_render(): spheres under a constant environment, pinhole camera, path integrator with
 * emitter sampling + BSDF sampling combined by the power heuristic (the structure of Mitsuba's
 * `path` integrator), Philox4x32-10 streams keyed by (pixel, sample,
 * bounce).  Geometry runs in double; what crosses into the BSDF (wi, wo, the random numbers) is
 * rounded to fp32 first, because fp32 is what a plugin receives.
 *
 * Build: oracle/Makefile -> _build/libmerl_render_oracle.so (includes merl_oracle.c for the BSDF).
 */
#include "merl_oracle.c"

typedef struct {
    int32_t width, height, spp_total, spp_begin, spp_end, max_depth;
    uint64_t seed;
    float cam_pos[3], cam_target[3], cam_up[3], fov_x_deg;
    float env[3];
} orc_render_params_t;

static void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1,
                          uint32_t out[4]) {
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        c1 = (uint32_t)p1; c3 = (uint32_t)p0; c0 = n0; c2 = n2;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
static float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

/* exported so the tests can pin the generator against the published Philox known answers */
void orc_philox(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    philox4x32_10(ctr[0], ctr[1], ctr[2], ctr[3], key[0], key[1], out);
}

#define RAY_EPS 1e-4

static int hit_spheres(const double *sph, int n, const double o[3], const double d[3], double *t_hit) {
    int best = -1;
    double tb = 1e300;
    for (int s = 0; s < n; ++s) {
        const double *c = sph + 4 * s;
        const double l[3] = {o[0] - c[0], o[1] - c[1], o[2] - c[2]};
        const double b = dot3(l, d), cc = dot3(l, l) - c[3] * c[3], disc = b * b - cc;
        if (disc < 0.0) continue;
        const double sq = sqrt(disc);
        double t = -b - sq;
        if (t < RAY_EPS) t = -b + sq;
        if (t >= RAY_EPS && t < tb) { tb = t; best = s; }
    }
    *t_hit = tb;
    return best;
}
static int is_occluded(const double *sph, int n, const double o[3], const double d[3]) {
    double t;
    return hit_spheres(sph, n, o, d, &t) >= 0;
}
static void make_frame(const double n[3], double s[3], double t[3]) {
    const double sign = copysign(1.0, n[2]), a = -1.0 / (sign + n[2]), b = n[0] * n[1] * a;
    s[0] = 1.0 + sign * n[0] * n[0] * a; s[1] = sign * b; s[2] = -sign * n[0];
    t[0] = b; t[1] = sign + n[1] * n[1] * a; t[2] = -n[1];
}
static double power_heuristic(double a, double b) { return a > 0.0 ? a * a / (a * a + b * b) : 0.0; }

/* film: [npix][4] doubles, ADDED to (rgb / spp_total, samples / spp_total), like merl_render. */
void orc_render(const orc_table_t *tables, int ntables, const double *spheres, const int32_t *sphere_mat,
                int nspheres, const orc_render_params_t *p, double *film) {
    (void)ntables;
    const int npix = p->width * p->height;
    const uint32_t k0 = (uint32_t)(p->seed & 0xffffffffu), k1 = (uint32_t)(p->seed >> 32);
    double f[3] = {p->cam_target[0] - p->cam_pos[0], p->cam_target[1] - p->cam_pos[1], p->cam_target[2] - p->cam_pos[2]};
    double len0 = sqrt(dot3(f, f));
    for (int a = 0; a < 3; ++a) f[a] /= len0;
    const double up[3] = {p->cam_up[0], p->cam_up[1], p->cam_up[2]};
    double r[3] = {f[1] * up[2] - f[2] * up[1], f[2] * up[0] - f[0] * up[2], f[0] * up[1] - f[1] * up[0]};
    len0 = sqrt(dot3(r, r));
    for (int a = 0; a < 3; ++a) r[a] /= len0;
    const double u[3] = {r[1] * f[2] - r[2] * f[1], r[2] * f[0] - r[0] * f[2], r[0] * f[1] - r[1] * f[0]};
    const double th = tan(0.5 * p->fov_x_deg * M_PI / 180.0), asp = (double)p->height / p->width;
    const double inv_spp = 1.0 / p->spp_total, pdf_e = 1.0 / (4.0 * M_PI);
#pragma omp parallel for schedule(dynamic, 64)
    for (int pixel = 0; pixel < npix; ++pixel) {
        double acc[4] = {0, 0, 0, 0};
        for (int s = p->spp_begin; s < p->spp_end; ++s) {
            uint32_t rnd[4];
            philox4x32_10((uint32_t)pixel, (uint32_t)s, 0u, 0u, k0, k1, rnd);
            const int x = pixel % p->width, y = pixel / p->width;
            const double sx = ((double)x + u01(rnd[0])) / p->width * 2.0 - 1.0;
            const double sy = 1.0 - ((double)y + u01(rnd[1])) / p->height * 2.0;
            double o[3] = {p->cam_pos[0], p->cam_pos[1], p->cam_pos[2]}, d[3];
            for (int a = 0; a < 3; ++a) d[a] = f[a] + sx * th * r[a] + sy * th * asp * u[a];
            const double dl = sqrt(dot3(d, d));
            for (int a = 0; a < 3; ++a) d[a] /= dl;
            double T[3] = {1, 1, 1}, L[3] = {0, 0, 0}, prev_pdf = 0.0;
            for (int depth = 0; depth <= p->max_depth; ++depth) {
                double t;
                const int hit = hit_spheres(spheres, nspheres, o, d, &t);
                if (hit < 0) {
                    const double w = depth == 0 ? 1.0 : power_heuristic(prev_pdf, pdf_e);
                    for (int c = 0; c < 3; ++c) L[c] += T[c] * p->env[c] * w;
                    break;
                }
                if (depth >= p->max_depth) break;
                const double *sp = spheres + 4 * hit;
                double h[3], n[3], fs[3], ft[3];
                for (int a = 0; a < 3; ++a) { h[a] = o[a] + t * d[a]; n[a] = (h[a] - sp[a]) / sp[3]; }
                const double nl = sqrt(dot3(n, n));
                for (int a = 0; a < 3; ++a) n[a] /= nl;
                make_frame(n, fs, ft);
                const double wi[3] = {(float)-dot3(d, fs), (float)-dot3(d, ft), (float)-dot3(d, n)};
                if (!(wi[2] > 0.0)) break;
                const orc_table_t *tb = tables + sphere_mat[hit];
                /* one Philox call per surface interaction: four uniforms from the high 24 bits of its words
                 * (emitter direction, lobe choice, first BSDF number), a fifth from their unused low bytes */
                uint32_t re[4];
                philox4x32_10((uint32_t)pixel, (uint32_t)s, (uint32_t)(depth + 1), 0u, k0, k1, re);
                const float u5 = (float)(((re[0] & 0xffu) << 16) | ((re[1] & 0xffu) << 8) | (re[2] & 0xffu)) *
                                 (1.0f / 16777216.0f);
                /* emitter sampling: uniform direction, pdf 1/(4 pi) */
                const double ez = 1.0 - 2.0 * u01(re[0]), er = sqrt(fmax(0.0, 1.0 - ez * ez));
                const double ephi = 2.0 * M_PI * u01(re[1]);
                const double e[3] = {er * cos(ephi), er * sin(ephi), ez};
                const double lz = dot3(e, n);
                const double ho[3] = {h[0] + RAY_EPS * n[0], h[1] + RAY_EPS * n[1], h[2] + RAY_EPS * n[2]};
                if (lz > 0.0 && !is_occluded(spheres, nspheres, ho, e)) {
                    const double woe[3] = {(float)dot3(e, fs), (float)dot3(e, ft), (float)lz};
                    double fe[3], pb;
                    eval_pdf_one(tb, wi, woe, fe, &pb, NULL);
                    const double mis = power_heuristic(pdf_e, pb) / pdf_e;
                    for (int c = 0; c < 3; ++c) L[c] += T[c] * fe[c] * p->env[c] * mis;
                }
                /* BSDF sampling */
                double wo[3];
                draw_direction(tb, wi, u01(re[2]), u01(re[3]), u5, wo);
                const double wof[3] = {(float)wo[0], (float)wo[1], (float)wo[2]};
                double fb[3], pdf;
                eval_pdf_one(tb, wi, wof, fb, &pdf, NULL);
                if (!(pdf > 0.0)) break;
                for (int c = 0; c < 3; ++c) T[c] *= fb[c] / pdf;
                if (!(T[0] + T[1] + T[2] > 0.0)) break;
                prev_pdf = pdf;
                for (int a = 0; a < 3; ++a) { d[a] = wof[0] * fs[a] + wof[1] * ft[a] + wof[2] * n[a]; o[a] = ho[a]; }
            }
            for (int c = 0; c < 3; ++c) acc[c] += L[c] * inv_spp;
            acc[3] += inv_spp;
        }
        for (int c = 0; c < 4; ++c) film[(size_t)pixel * 4 + c] += acc[c];
    }
}
