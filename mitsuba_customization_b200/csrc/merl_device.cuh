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

// ---- candidate polynomials (accuracy requirement: |error| << 0.5 bin) ----
// atan(y/x) for y >= 0, any x: 4-term odd minimax, max rel. error 1.9e-4.
__device__ __forceinline__ float fast_atan2_pos(float y, float x) {
    const float ax = fabsf(x);
    const float mx = fmaxf(ax, y), mn = fminf(ax, y);
    const float a = mn * rcp_approx(fmaxf(mx, 1e-37f));
    const float a2 = a * a;
    float r = fmaf(fmaf(fmaf(-0.04508537f, a2, 0.15670808f), a2, -0.32622819f), a2, 0.9998109f) * a;
    r = (y > ax) ? 1.57079633f - r : r;
    return (x < 0.f) ? 3.14159265f - r : r;
}
// acos(x) on [0,1]: sqrt(1-x)*cubic, max abs. error 3.9e-5 rad.
__device__ __forceinline__ float fast_acos01(float x) {
    x = fminf(fmaxf(x, 0.f), 1.f);
    const float p = fmaf(fmaf(fmaf(-0.02087747f, x, 0.07687739f), x, -0.21286782f), x, 1.57075773f);
    return sqrt_approx(1.f - x) * p;
}

// acos(x) on [0,1] given 1 - x computed without cancellation (relative error 2.3e-4 as x -> 1).
__device__ __forceinline__ float fast_acos01_om(float x, float one_minus_x) {
    const float p = fmaf(fmaf(fmaf(-0.02087747f, x, 0.07687739f), x, -0.21286782f), x, 1.57075773f);
    return sqrt_approx(fmaxf(one_minus_x, 0.f)) * p;
}

struct Bins {
    int ih, id, ip;
};

// Quantities of one query shared by binning, pdf and the slow tiers.
struct HalfDiff {
    float sx, sy, sz, rho2, n2, A, inv_s, slen;
};

__device__ __forceinline__ HalfDiff half_diff_setup(float ax, float ay, float az, float bx, float by,
                                                    float bz) {
    HalfDiff h;
    h.sx = ax + bx; h.sy = ay + by; h.sz = az + bz;
    h.rho2 = fmaf(h.sx, h.sx, h.sy * h.sy);
    h.n2 = fmaf(h.sz, h.sz, h.rho2);
    h.A = fmaf(ax, h.sx, fmaf(ay, h.sy, az * h.sz));
    h.inv_s = rsqrt_approx(fmaxf(h.n2, 1e-37f));
    h.slen = h.n2 * h.inv_s;
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
template <bool DETAIL = false>
__device__ __forceinline__ Bins compute_bins_fast(const TableDev& t, const SmemView& v, float ax, float ay,
                                                  float az, float bx, float by, float bz, const HalfDiff& h,
                                                  unsigned& unc, int kc[3]) {
    constexpr float EPS_H = 2e-6f, EPS_D = 2e-6f, EPS_P = 1.5e-6f;  // error budgets: DESIGN.md §3
    // candidates
    // theta_h = acos(sz/|s|) with 1 - cos written as rho^2 / (|s| (|s| + sz)): no cancellation at the
    // small angles where MERL's sqrt-spaced bins are 2e-4 rad wide (sz > 0 for every valid query;
    // an invalid one may produce NaN here, which F2I turns into candidate 0)
    const float ch = __saturatef(h.sz * h.inv_s);
    const float th_h = fast_acos01_om(ch, h.rho2 * h.inv_s * rcp_approx(h.slen + h.sz));
    const float tq = th_h * t.kh;
    const int kh = min(__float2int_rn(t.theta_h_sqrt ? sqrt_approx(tq) : tq), t.nh);
    const int kd = min(__float2int_rn(fast_acos01(h.A * h.inv_s) * t.kd), t.nd);
    float X = fmaf(h.sz, h.A, -az * h.n2);
    float Y = h.slen * fmaf(ay, bx, -ax * by);
    if (Y < 0.f) { X = -X; Y = -Y; }
    const int kp = min(__float2int_rn(fast_atan2_pos(Y, X) * t.kp), t.np);
    MERL_CHECK(kh >= 0 && kh <= t.nh && kd >= 0 && kd <= t.nd && kp >= 0 && kp <= t.np);
    // one exact-edge comparison per coordinate
    const float eh = v.thr_h[kh], ed = v.thr_d[kd];
    const float2 ep = v.dir_p[kp];
    const float sz2 = h.sz * h.sz;
    const float dh = fmaf(-eh, sz2, h.rho2);           // >= 0  <=>  theta_h >= edge kh
    const float dd = fmaf(ed, h.n2, -h.A * fabsf(h.A)); // >= 0  <=>  theta_d >= edge kd
    const float dp = fmaf(ep.x, Y, -ep.y * X);         // >= 0  <=>  phi_d   >= edge kp
    Bins b;
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
    const float mp = EPS_P * (h.n2 + h.slen);
    const float slack_h = fmaf(-EPS_H, h.rho2, fabsf(dh));
    const float slack_d = fmaf(-EPS_D, h.n2, fabsf(dd));
    const float slack_p = fabsf(dp) - mp;
    const float slack_c = fmaf(-4.f * (float)t.np, mp, fabsf(X) + Y);
    const bool finite = h.n2 < __int_as_float(0x7f800000);
    if (DETAIL) {   // which coordinate(s): only the slow tiers need to know
        const unsigned all = finite ? 0u : 7u;
        unc = all | (unsigned)!(slack_h > 0.f) | ((unsigned)!(slack_d > 0.f) << 1) |
              ((unsigned)(!(slack_p > 0.f) | !(slack_c > 0.f)) << 2);
    } else {
        const float slack = fminf(fminf(slack_h, slack_d), fminf(slack_p, slack_c));
        unc = (unsigned)(!(slack > 0.f) | !finite);
    }
    kc[0] = kh; kc[1] = kd; kc[2] = kp;
    return b;
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
