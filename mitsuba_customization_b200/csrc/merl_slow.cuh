// merl_slow.cuh -- tiers 2 and 3 of the bin computation (see merl_device.cuh).
// Included exactly once, by merl_b200.cu.
#pragma once
#include "merl_device.cuh"

namespace merl {

// ---- tier 3: the public reader's operation sequence in IEEE double, no FMA contraction ----
// (SURVEY.md App. A.  __dmul_rn/__dadd_rn are never fused, so +,-,*,/,sqrt round exactly as
// gcc -ffp-contract=off does on the host; only acos/atan2/sin/cos can differ in the last ulp.)
__device__ __forceinline__ double dmul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double dadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double dot3_literal(const double a[3], const double b[3]) {
    return dadd(dadd(dmul(a[0], b[0]), dmul(a[1], b[1])), dmul(a[2], b[2]));
}
__device__ __forceinline__ double clamp_unit(double x) { return x > 1.0 ? 1.0 : (x < -1.0 ? -1.0 : x); }

__device__ void rotate_literal(const double v[3], const double axis[3], double angle, double out[3]) {
    const double c = cos(angle), s = sin(angle);
    const double k = dmul(dot3_literal(axis, v), dadd(1.0, -c));
    const double cr[3] = {dadd(dmul(axis[1], v[2]), -dmul(axis[2], v[1])),
                          dadd(dmul(axis[2], v[0]), -dmul(axis[0], v[2])),
                          dadd(dmul(axis[0], v[1]), -dmul(axis[1], v[0]))};
#pragma unroll
    for (int a = 0; a < 3; ++a) out[a] = dadd(dadd(dmul(v[a], c), dmul(axis[a], k)), dmul(cr[a], s));
}

__device__ void literal_bins(const TableDev& t, const double in[3], const double out[3], Bins& b) {
    const double PI = 3.14159265358979323846;
    double h[3] = {dadd(in[0], out[0]), dadd(in[1], out[1]), dadd(in[2], out[2])};
    const double len = __dsqrt_rn(dot3_literal(h, h));
    if (len > 0.0) { h[0] = __ddiv_rn(h[0], len); h[1] = __ddiv_rn(h[1], len); h[2] = __ddiv_rn(h[2], len); }
    const double theta_half = acos(clamp_unit(h[2]));
    const double fi_half = atan2(h[1], h[0]);
    const double normal[3] = {0.0, 0.0, 1.0}, binormal[3] = {0.0, 1.0, 0.0};
    double tmp[3], d[3];
    rotate_literal(in, normal, -fi_half, tmp);
    rotate_literal(tmp, binormal, -theta_half, d);
    const double theta_diff = acos(clamp_unit(d[2]));
    double phi_diff = atan2(d[1], d[0]);
    // index formulas, same operation order as the oracle's orc_*_index
    int ih = 0;
    if (theta_half > 0.0) {
        double x = dmul(__ddiv_rn(theta_half, __ddiv_rn(PI, 2.0)), (double)t.nh);
        if (t.theta_h_sqrt) x = __dsqrt_rn(dmul(x, (double)t.nh));
        ih = (int)x;
    }
    int id = (int)dmul(__ddiv_rn(theta_diff, dmul(PI, 0.5)), (double)t.nd);
    if (phi_diff < 0.0) phi_diff = dadd(phi_diff, PI);
    int ip = (int)__ddiv_rn(dmul(__ddiv_rn(phi_diff, PI), (double)(2 * t.np)), 2.0);
    b.ih = max(min(ih, t.nh - 1), 0);
    b.id = max(min(id, t.nd - 1), 0);
    b.ip = max(min(ip, t.np - 1), 0);
}

// ---- tier 2: the tier-1 comparisons repeated in double against double edges ----
// Margins are the double reference's own rounding noise mapped into each comparison's
// domain: theta_h noise is 1.1e-16 absolute in cos(theta_h), i.e.
// ~2e-16*|s|^4/sz^2 in (rho^2 - T_k sz^2); theta_d and phi_d noise is a few e-16 of |s|^2.
__device__ __noinline__ void resolve_bins_slow(const TableDev& t, float ax_, float ay_, float az_,
                                               float bx_, float by_, float bz_, int kh, int kd, int kp,
                                               unsigned unc, Bins& out) {
    const double ax = ax_, ay = ay_, az = az_, bx = bx_, by = by_, bz = bz_;
    const double sx = ax + bx, sy = ay + by, sz = az + bz;
    const double rho2 = sx * sx + sy * sy, sz2 = sz * sz, n2 = rho2 + sz2;
    const double A = ax * sx + ay * sy + az * sz;
    unsigned still = 0;
    if (unc & 1u) {
        const double dh = rho2 - t.thr_h64[kh] * sz2;
        const double margin = 1e-15 * n2 * n2 / fmax(sz2, 1e-300) + 1e-13 * rho2;
        if (fabs(dh) > margin) out.ih = max(min(kh - 1 + (dh >= 0.0), t.nh - 1), 0);
        else still |= 1u;
    }
    if (unc & 2u) {
        const double dd = t.thr_d64[kd] * n2 - A * fabs(A);
        if (fabs(dd) > 1e-13 * n2) out.id = max(min(kd - 1 + (dd >= 0.0), t.nd - 1), 0);
        else still |= 2u;
    }
    if (unc & 4u) {
        const double slen = sqrt(n2);
        double X = sz * A - az * n2;
        double Y = slen * (ay * bx - ax * by);
        if (Y < 0.0) { X = -X; Y = -Y; }
        // Tier 1's candidate edge is unreliable here (that is one reason to be in tier 2: when
        // |(X,Y)| is small, fp32 cancellation in X moves the direction by many bins), so the
        // candidate is re-derived from the double (X, Y).  phi = 0 and phi = pi are the same
        // fold edge but different bins; atan2 of the folded pair picks the right end.
        const double m = 1e-13 * (n2 + slen);
        if (fabs(X) + Y > m * 4.0 * t.np) {
            kp = max(min((int)rint(atan2(Y, X) * (t.np / 3.14159265358979323846)), t.np), 0);
            const double2 e = t.dir_p64[kp];
            const double dp = e.x * Y - e.y * X;
            if (fabs(dp) > m) out.ip = max(min(kp - 1 + (dp >= 0.0), t.np - 1), 0);
            else still |= 4u;
        } else {
            still |= 4u;
        }
    }
    if (still) {
        const double in[3] = {ax, ay, az}, o[3] = {bx, by, bz};
        Bins lit;
        literal_bins(t, in, o, lit);
        if (still & 1u) out.ih = lit.ih;
        if (still & 2u) out.id = lit.id;
        if (still & 4u) out.ip = lit.ip;
    }
}

}  // namespace merl
