// merl_device.cuh -- device-side half/difference binning for MERL-layout tables (sm_100a).
//
// What it computes is the public MERL lookup's bin triple:
//   ih = floor(nh*sqrt(theta_h/(pi/2)))  id = floor(nd*theta_d/(pi/2))  ip = floor(np*(phi_d mod pi)/pi)
// HOW it computes it is B200-first: no acos/atan2 on the fast path.  Each coordinate is a
// monotone function of a cheap rational quantity of (wi, wo):
//   theta_h : tan^2 = (sx^2+sy^2)/sz^2                       s = wi + wo
//   theta_d : cos^2 = (wi.s)^2/|s|^2
//   phi_d   : direction of (X, Y) = (sz*(wi.s) - wi.z*|s|^2,  |s|*(wi.y*wo.x - wi.x*wo.y))
// so a bin is "candidate from a 0.03-bin-accurate polynomial, then ONE exact comparison
// against a tabulated edge".  Three tiers keep the result bit-exact w.r.t. the double
// reference arithmetic (BASELINE.json: "bin indices ... must be bit-exact"):
//   tier 1  fp32 comparison with a proven error margin          (~99.9 % of queries)
//   tier 2  same comparison in fp64 against fp64 edges          (margin = the double
//           reference's own rounding noise)
//   tier 3  literal double acos/atan2/Rodrigues sequence        (degenerate inputs and
//           queries inside the reference's own noise band; ~1e-9 of queries)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

// Debug build (make EXTRA=-DMERL_DEBUG_BOUNDS): device-side asserts on every computed index before it
// is used -- candidate edges, flat cell index, CDF search result, live-list and stage-slot indices.
// compute-sanitizer is not available on the measurement pool, so tools/sanitize_run.py is run
// against this build instead (profiles/r01_bounds_check.txt).
#ifdef MERL_DEBUG_BOUNDS
#include <assert.h>
#define MERL_CHECK(c) assert(c)
#else
#define MERL_CHECK(c) ((void)0)
#endif

namespace merl {

constexpr int kMaxNH = 256, kMaxND = 256, kMaxNP = 512, kMaxNI = 64;
constexpr int kGuideCells = 128;   // cells of the sampler's per-row guide table (uint8 bin index each)

// Everything a kernel needs about one table; passed by value as a kernel parameter.
struct TableDev {
    const float4* cells;     // [nh*nd*np] prescaled RGB (+ validity bits in .w), L2-resident
    const float* image;      // smem image: thr_h | thr_d | dir_p | dens | cdf | medge (fp32) | guide (uint8)
    const double* thr_h64;   // [nh+1] tan^2(theta_i)   (tier 2)
    const double* thr_d64;   // [nd+1] cos^2(theta_k)
    const double2* dir_p64;  // [np+1] (cos, sin)(phi_k)
    int nh, nd, np, ni;
    int theta_h_sqrt;
    int off_thr_d, off_dir_p, off_dens, off_cdf, off_medge, off_guide;  // float offsets into image
    int image_eval_floats;   // floats needed by eval/pdf/backward kernels (up to dens incl.)
    int image_full_floats;   // floats needed by the sample kernel (everything)
    float kh;                // candidate scale: sqrt map: nh^2*2/pi ; linear: nh*2/pi
    float kd;                // nd*2/pi
    float kp;                // np/pi
    float spec_weight;
    int search_step;         // largest power of two <= nh-1: first stride of the CDF row search
};

// Views into the staged shared-memory image.
struct SmemView {
    const float* thr_h;
    const float* thr_d;
    const float2* dir_p;
    const float* dens;
    const float* cdf;
    const float* medge;
    const uint8_t* guide;   // [ni][kGuideCells]: largest bin whose CDF entry is <= cell / kGuideCells
};

__device__ __forceinline__ SmemView make_view(const TableDev& t, const float* smem) {
    SmemView v;
    v.thr_h = smem;
    v.thr_d = smem + t.off_thr_d;
    v.dir_p = reinterpret_cast<const float2*>(smem + t.off_dir_p);
    v.dens = smem + t.off_dens;
    v.cdf = smem + t.off_cdf;
    v.medge = smem + t.off_medge;
    v.guide = reinterpret_cast<const uint8_t*>(smem + t.off_guide);
    return v;
}

// Cooperative copy of the first `nfloats` of the image into shared memory (16 B vectors).
__device__ __forceinline__ void stage_image(const TableDev& t, float* smem, int nfloats) {
    const float4* src = reinterpret_cast<const float4*>(t.image);
    float4* dst = reinterpret_cast<float4*>(smem);
    const int n4 = (nfloats + 3) >> 2;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) dst[i] = __ldg(src + i);
    __syncthreads();
}

// MUFU-backed approximations (1-2 ulp).  They feed candidates and values with a 1e-5
// tolerance only; every bin decision goes through the margin-checked comparisons below.
__device__ __forceinline__ float sqrt_approx(float x) {
    float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r;
}
__device__ __forceinline__ float rsqrt_approx(float x) {
    float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r;
}
__device__ __forceinline__ float rcp_approx(float x) {
    float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r;
}

// ---- one query (float) or a PAIR of queries (float2) per lane ---------------------------------------
// sm_100 has packed fp32x2 arithmetic (FFMA2 / FMUL2 / FADD2: two IEEE fp32 results per instruction and
// lane, the same rounding as the scalar forms, half the issue slots; measured on B200 at the scalar FMA
// rate and latency, tools/probe/ffma2_probe.cu, profiles/r02_ffma2_probe.txt).  The streaming kernels hold
// four queries per lane, so the multiply/add/fma backbone of the bin math is written once over
// T = float | float2 and the 4-wide loops run it on two pairs.  Everything with an operand modifier or no
// packed form (abs, min/max, saturate, MUFU, compares, selects, float->int) stays per component.  Every
// product and sum is an explicit round-to-nearest operation (never contracted by the compiler), so a query
// gives the same bits whichever width it went through: 4-wide, tail, AoS and redo paths stay bit-identical.
template <class T> __device__ __forceinline__ T p_set(float c);
template <> __device__ __forceinline__ float p_set<float>(float c) { return c; }
template <> __device__ __forceinline__ float2 p_set<float2>(float c) { return make_float2(c, c); }

__device__ __forceinline__ float p_mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float p_add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float p_sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float p_fma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
__device__ __forceinline__ float p_neg(float a) { return -a; }
__device__ __forceinline__ float2 p_mul(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 p_add(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 p_sub(float2 a, float2 b) { return __ffma2_rn(b, make_float2(-1.f, -1.f), a); }   // a - b, one rounding
__device__ __forceinline__ float2 p_fma(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float2 p_neg(float2 a) { return __fmul2_rn(a, make_float2(-1.f, -1.f)); }               // exact

// per-component operations
#define MERL_P1(name, expr)                                                                          \
    __device__ __forceinline__ float name(float a) { return expr(a); }                               \
    __device__ __forceinline__ float2 name(float2 a) { return make_float2(expr(a.x), expr(a.y)); }
MERL_P1(p_abs, fabsf)
MERL_P1(p_sqrt, sqrt_approx)
MERL_P1(p_rsqrt, rsqrt_approx)
MERL_P1(p_rcp, rcp_approx)
#undef MERL_P1
__device__ __forceinline__ float p_max(float a, float c) { return fmaxf(a, c); }
__device__ __forceinline__ float2 p_max(float2 a, float c) { return make_float2(fmaxf(a.x, c), fmaxf(a.y, c)); }
__device__ __forceinline__ float p_mul_sat(float a, float b) { return __saturatef(__fmul_rn(a, b)); }
__device__ __forceinline__ float2 p_mul_sat(float2 a, float2 b) {
    return make_float2(__saturatef(__fmul_rn(a.x, b.x)), __saturatef(__fmul_rn(a.y, b.y)));
}

// ---- candidate polynomials (accuracy requirement: |error| << 0.5 bin) ----
template <class T> __device__ __forceinline__ T acos_poly(T x) {
    return p_fma(p_fma(p_fma(p_set<T>(-0.02087747f), x, p_set<T>(0.07687739f)), x, p_set<T>(-0.21286782f)), x,
                 p_set<T>(1.57075773f));
}
// acos(x), x already clamped to [0,1]: sqrt(1-x)*cubic, max abs. error 3.9e-5 rad.
template <class T> __device__ __forceinline__ T fast_acos01_t(T x) {
    return p_mul(p_sqrt(p_sub(p_set<T>(1.f), x)), acos_poly(x));
}
// acos(x) on [0,1] given 1 - x computed without cancellation (relative error 2.3e-4 as x -> 1).
template <class T> __device__ __forceinline__ T fast_acos01_om_t(T x, T one_minus_x) {
    return p_mul(p_sqrt(p_max(one_minus_x, 0.f)), acos_poly(x));
}
// atan(y/x) for y >= 0, any x: 4-term odd minimax, max rel. error 1.9e-4.
__device__ __forceinline__ float atan_fold(float r, float y, float ax, float x) {
    r = (y > ax) ? __fsub_rn(1.57079633f, r) : r;
    return (x < 0.f) ? __fsub_rn(3.14159265f, r) : r;
}
__device__ __forceinline__ float fast_atan2_pos_t(float y, float x) {
    const float ax = fabsf(x);
    const float mx = fmaxf(ax, y), mn = fminf(ax, y);
    const float a = p_mul(mn, rcp_approx(fmaxf(mx, 1e-37f)));
    const float a2 = p_mul(a, a);
    const float r = p_mul(p_fma(p_fma(p_fma(-0.04508537f, a2, 0.15670808f), a2, -0.32622819f), a2, 0.9998109f), a);
    return atan_fold(r, y, ax, x);
}
__device__ __forceinline__ float2 fast_atan2_pos_t(float2 y, float2 x) {
    typedef float2 T;
    const float2 ax = p_abs(x);
    const float2 mx = make_float2(fmaxf(ax.x, y.x), fmaxf(ax.y, y.y)), mn = make_float2(fminf(ax.x, y.x), fminf(ax.y, y.y));
    const float2 a = p_mul(mn, p_rcp(p_max(mx, 1e-37f)));
    const float2 a2 = p_mul(a, a);
    const float2 r = p_mul(p_fma(p_fma(p_fma(p_set<T>(-0.04508537f), a2, p_set<T>(0.15670808f)), a2, p_set<T>(-0.32622819f)), a2,
                                 p_set<T>(0.9998109f)), a);
    return make_float2(atan_fold(r.x, y.x, ax.x, x.x), atan_fold(r.y, y.y, ax.y, x.y));
}

struct Bins {
    int ih, id, ip;
};

// Quantities of one query (or pair) shared by binning, pdf and the slow tiers.
template <class T> struct HalfDiffT { T sx, sy, sz, rho2, n2, A, inv_s, slen; };
typedef HalfDiffT<float> HalfDiff;

template <class T>
__device__ __forceinline__ HalfDiffT<T> half_diff_setup(T ax, T ay, T az, T bx, T by, T bz) {
    HalfDiffT<T> h;
    h.sx = p_add(ax, bx); h.sy = p_add(ay, by); h.sz = p_add(az, bz);
    h.rho2 = p_fma(h.sx, h.sx, p_mul(h.sy, h.sy));
    h.n2 = p_fma(h.sz, h.sz, h.rho2);
    h.A = p_fma(ax, h.sx, p_fma(ay, h.sy, p_mul(az, h.sz)));
    h.inv_s = p_rsqrt(p_max(h.n2, 1e-37f));
    h.slen = p_mul(h.n2, h.inv_s);
    return h;
}

// Tier 2 + tier 3 (out of line: ~1e-3 of lanes ever get here).  `unc` bit0/1/2 = coordinate
// whose fp32 test was inside its error margin; kh/kd/kp are the tier-1 candidate edges.
__device__ __noinline__ void resolve_bins_slow(const TableDev& t, float ax_, float ay_, float az_,
                                               float bx_, float by_, float bz_, int kh, int kd, int kp,
                                               unsigned unc, Bins& out);

// Tier 1.  Returns the tier-1 bins and, in `unc`, whether any coordinate was decided inside its
// error margin and therefore needs tiers 2-3: nonzero / zero when DETAIL is false (the streaming
// loops), which coordinates (bit 0/1/2 = theta_h / theta_d / phi_d) when DETAIL is true (the redo).
// kc[] receives the candidate edges the slow tiers start from.
//
// Per-component tail: candidate edges from the scaled angles, and -- once the exact-edge comparisons
// dh / dd / dp are formed -- bins and certainty.
__device__ __forceinline__ void bins_candidates(const TableDev& t, float tq, float tdv, float tpv, int& kh, int& kd, int& kp) {
    kh = min(__float2int_rn(t.theta_h_sqrt ? sqrt_approx(tq) : tq), t.nh);
    kd = min(__float2int_rn(tdv), t.nd);
    kp = min(__float2int_rn(tpv), t.np);
    MERL_CHECK(kh >= 0 && kh <= t.nh && kd >= 0 && kd <= t.nd && kp >= 0 && kp <= t.np);
}
template <bool DETAIL>
__device__ __forceinline__ unsigned bins_decide(const TableDev& t, int kh, int kd, int kp, float dh, float dd, float dp,
                                                float rho2, float n2, float mp, float X, float Y, Bins& b) {
    constexpr float EPS_H = 2e-6f, EPS_D = 2e-6f;  // error budgets: DESIGN.md §3
    b.ih = max(min(kh - 1 + (dh >= 0.f), t.nh - 1), 0);  // max(): NaN inputs must not index out of range
    b.id = max(min(kd - 1 + (dd >= 0.f), t.nd - 1), 0);
    b.ip = max(min(kp - 1 + (dp >= 0.f), t.np - 1), 0);
    // Certainty: every comparison must clear its error margin, and for phi_d |(X,Y)| must also be
    // large enough that the same error cannot have moved the candidate edge itself by half a bin.
    // All four conditions are "slack > 0", so one min and one compare decide them together
    // (branch-free: the four queries a thread holds must keep interleaving).  No k == 0 / k == n
    // guards are needed: the edge tables hold 0 / +inf (theta_h) and 4 / -4 (theta_d) there, which
    // puts |d| far outside the margin, and the clamp above makes the bin independent of d.
    // fminf drops NaN operands, so non-finite inputs are caught by the test on n2 instead.
    const float slack_h = fmaf(-EPS_H, rho2, fabsf(dh));
    const float slack_d = fmaf(-EPS_D, n2, fabsf(dd));
    const float slack_p = fabsf(dp) - mp;
    const float slack_c = fmaf(-4.f * (float)t.np, mp, fabsf(X) + Y);
    const bool finite = n2 < __int_as_float(0x7f800000);
    if (DETAIL) {   // which coordinate(s): only the slow tiers need to know
        const unsigned all = finite ? 0u : 7u;
        return all | (unsigned)!(slack_h > 0.f) | ((unsigned)!(slack_d > 0.f) << 1) |
               ((unsigned)(!(slack_p > 0.f) | !(slack_c > 0.f)) << 2);
    }
    const float slack = fminf(fminf(slack_h, slack_d), fminf(slack_p, slack_c));
    return (unsigned)(!(slack > 0.f) | !finite);
}

__device__ __forceinline__ void fold_xy(float& X, float& Y) {
    if (Y < 0.f) { X = -X; Y = -Y; }
}
__device__ __forceinline__ void fold_xy(float2& X, float2& Y) { fold_xy(X.x, Y.x); fold_xy(X.y, Y.y); }

// The multiply/add backbone, for one query or a pair: scaled candidate angles, the folded (X, Y) of
// phi_d, and the operands of the three exact-edge comparisons.
template <class T> struct Tier1 { T tq, tdv, tpv, X, Y, nsz2, nAA, mp; };
template <class T>
__device__ __forceinline__ Tier1<T> tier1_backbone(const TableDev& t, T ax, T ay, T az, T bx, T by, const HalfDiffT<T>& h) {
    constexpr float EPS_P = 1.5e-6f;
    Tier1<T> r;
    // theta_h = acos(sz/|s|) with 1 - cos written as rho^2 / (|s| (|s| + sz)): no cancellation at the
    // small angles where MERL's sqrt-spaced bins are 2e-4 rad wide (sz > 0 for every valid query;
    // an invalid one may produce NaN here, which F2I turns into candidate 0)
    const T ch = p_mul_sat(h.sz, h.inv_s);
    const T th_h = fast_acos01_om_t(ch, p_mul(p_mul(h.rho2, h.inv_s), p_rcp(p_add(h.slen, h.sz))));
    r.tq = p_mul(th_h, p_set<T>(t.kh));
    r.tdv = p_mul(fast_acos01_t(p_mul_sat(h.A, h.inv_s)), p_set<T>(t.kd));
    r.X = p_fma(h.sz, h.A, p_mul(p_neg(az), h.n2));
    r.Y = p_mul(h.slen, p_fma(ay, bx, p_mul(p_neg(ax), by)));
    fold_xy(r.X, r.Y);
    r.tpv = p_mul(fast_atan2_pos_t(r.Y, r.X), p_set<T>(t.kp));
    r.nsz2 = p_mul(p_neg(h.sz), h.sz);          // -(sz^2)
    r.nAA = p_mul(p_neg(h.A), p_abs(h.A));      // -A |A|
    r.mp = p_mul(p_set<T>(EPS_P), p_add(h.n2, h.slen));
    return r;
}

template <bool DETAIL = false>
__device__ __forceinline__ Bins compute_bins_fast(const TableDev& t, const SmemView& v, float ax, float ay,
                                                  float az, float bx, float by, float bz, const HalfDiff& h,
                                                  unsigned& unc, int kc[3]) {
    const Tier1<float> r = tier1_backbone<float>(t, ax, ay, az, bx, by, h);
    int kh, kd, kp;
    bins_candidates(t, r.tq, r.tdv, r.tpv, kh, kd, kp);
    // one exact-edge comparison per coordinate
    const float eh = v.thr_h[kh], ed = v.thr_d[kd];
    const float2 ep = v.dir_p[kp];
    const float dh = p_fma(eh, r.nsz2, h.rho2);                       // >= 0  <=>  theta_h >= edge kh
    const float dd = p_fma(ed, h.n2, r.nAA);                          // >= 0  <=>  theta_d >= edge kd
    const float dp = p_fma(ep.x, r.Y, p_mul(p_neg(ep.y), r.X));       // >= 0  <=>  phi_d   >= edge kp
    Bins b;
    unc = bins_decide<DETAIL>(t, kh, kd, kp, dh, dd, dp, h.rho2, h.n2, r.mp, r.X, r.Y, b);
    kc[0] = kh; kc[1] = kd; kc[2] = kp;
    return b;
}
// The same for a pair of queries (streaming loops only: no DETAIL, no candidates out).
__device__ __forceinline__ void compute_bins_fast2(const TableDev& t, const SmemView& v, float2 ax, float2 ay, float2 az,
                                                   float2 bx, float2 by, const HalfDiffT<float2>& h, unsigned unc[2],
                                                   Bins b[2]) {
    const Tier1<float2> r = tier1_backbone<float2>(t, ax, ay, az, bx, by, h);
    int kh[2], kd[2], kp[2];
    bins_candidates(t, r.tq.x, r.tdv.x, r.tpv.x, kh[0], kd[0], kp[0]);
    bins_candidates(t, r.tq.y, r.tdv.y, r.tpv.y, kh[1], kd[1], kp[1]);
    const float2 eh = make_float2(v.thr_h[kh[0]], v.thr_h[kh[1]]), ed = make_float2(v.thr_d[kd[0]], v.thr_d[kd[1]]);
    const float2 e0 = v.dir_p[kp[0]], e1 = v.dir_p[kp[1]];
    const float2 dh = p_fma(eh, r.nsz2, h.rho2);
    const float2 dd = p_fma(ed, h.n2, r.nAA);
    const float2 dp = p_fma(make_float2(e0.x, e1.x), r.Y, p_mul(p_neg(make_float2(e0.y, e1.y)), r.X));
    unc[0] = bins_decide<false>(t, kh[0], kd[0], kp[0], dh.x, dd.x, dp.x, h.rho2.x, h.n2.x, r.mp.x, r.X.x, r.Y.x, b[0]);
    unc[1] = bins_decide<false>(t, kh[1], kd[1], kp[1], dh.y, dd.y, dp.y, h.rho2.y, h.n2.y, r.mp.y, r.X.y, r.Y.y, b[1]);
}

// All three tiers inline: the exact bins of one query.  Used by the redo path of the streaming
// kernels (uncertain queries only), by the table-build kernels and by the AoS kernels.
__device__ __forceinline__ Bins compute_bins(const TableDev& t, const SmemView& v, float ax, float ay,
                                             float az, float bx, float by, float bz, const HalfDiff& h) {
    unsigned unc;
    int kc[3];
    Bins b = compute_bins_fast<true>(t, v, ax, ay, az, bx, by, bz, h, unc, kc);
    if (unc) resolve_bins_slow(t, ax, ay, az, bx, by, bz, kc[0], kc[1], kc[2], unc, b);
    return b;
}

}  // namespace merl
