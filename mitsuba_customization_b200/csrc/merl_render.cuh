// merl_render.cuh -- wavefront path tracer harness around the BSDF stage.
//
// The callers either side of the hot path, so that BASELINE configs 2 (one sphere, one table) and
// 5 (many spheres, many tables) can be rendered and checked as images.  Headless: analytic
// spheres, constant environment, pinhole camera, no graphics API.  One wavefront pass:
//   rt_primary_kernel  depth 0: camera ray + first intersection for every path of the pass; escaped
//                      rays go straight to the film, the hits form the first live list
//   rt_trace_kernel    depth >= 1, over the live list: intersect, terminate misses into the film
//   (both)             build the local frame, draw the emitter direction + shadow ray, write the
//                      BSDF-stage inputs (float4 AoS), count the queries per material
//   rt_scan_kernel /   counting sort of the live queries by material, so that each table's queries
//   rt_scatter_kernel  are one contiguous segment (one table: plain stream compaction)
//   BSDF stage         EvalPdfAosOp + SampleAosOp (the kernels of the C ABI's *_aos entry points),
//                      one launch pair per bounce; segmented over tables when there are several
//   rt_shade_kernel    MIS-weighted emitter contribution, throughput update, next ray, next live list
// The CPU oracle restates the same algorithm for the tests.  Included once, by merl_b200.cu.
#pragma once
#include "merl_kernels.cuh"

namespace merl {

// ---- Philox4x32-10: counter (pixel, sample, dimension block, 0), key = seed ------------------
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                              uint32_t k1, uint32_t out[4]) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
__device__ __forceinline__ float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }
// a fifth uniform from the low bytes of three words (the bits u01 discards)
__device__ __forceinline__ float u01_low(uint32_t a, uint32_t b, uint32_t c) {
    return (float)(((a & 0xffu) << 16) | ((b & 0xffu) << 8) | (c & 0xffu)) * (1.0f / 16777216.0f);
}

struct Camera {
    float px, py, pz;     // position
    float fx, fy, fz;     // forward
    float rx, ry, rz;     // right * tan(fov_x/2)
    float ux, uy, uz;     // up * tan(fov_x/2) * height/width
};

struct RenderDev {
    int width, height, spp_total, spp_begin, spp_end, max_depth;
    int nspheres, ntables;
    uint32_t k0, k1;
    float env[3];
    float inv_spp;
    Camera cam;
};

// Path state of one wavefront pass over `cap` paths (device workspace).  From the second bounce on
// the kernels reach it through the live list / the stage slots, i.e. in scattered order, so related
// scalars share a float4 (one 16-byte access, half a sector) instead of one array each.
struct PathState {
    float4* org;      // ray origin, .w = solid-angle pdf of the direction that produced the ray (0 = camera)
    float4* dir;      // ray direction
    float4* thr;      // throughput
    float4* rad;      // accumulated radiance
    float4* nrm;      // shading normal at the current hit
    float4* hit;      // hit point
    int* pixel;       // -1 = dead
    int* sample;
    int* slot;        // indexed by stage slot: the path whose query sits in that slot
    int* mat;         // material at the current hit, -1 when the path has no query this bounce
};

constexpr int kMaxSpheres = 256;
constexpr float kRayEps = 1e-4f;

__device__ __forceinline__ void film_add(float4* film, int pixel, float r, float g, float b, float w) {
    red_add_v4(film + pixel, r, g, b, w);
}

// Two-level intersection.  The host sorts the spheres along a Morton curve and groups kClusterSize
// consecutive ones under an axis-aligned bounding box (merl_render); a ray tests the box first and skips the
// group when it misses it, when the box lies behind the origin, or when it starts beyond the best
// hit so far.  The per-sphere arithmetic is the brute-force loop's (and the oracle's), so the nearest
// hit is the same one; only exact ties in t could resolve to another sphere.
constexpr int kClusterSize = 8;
constexpr int kMaxClusters = kMaxSpheres / kClusterSize;

// ray vs the group's axis-aligned bounding box (slab test): false = no sphere inside can be hit in
// (kRayEps, t_max).  The box is padded on the host, so a ray lying exactly in one of its faces cannot
// touch a sphere and the NaN that 0 * inf produces there (fminf / fmaxf drop it) is harmless.
struct ClusterBox { float4 lo, hi; };
__device__ __forceinline__ bool cluster_may_hit(const ClusterBox c, float ox, float oy, float oz, float idx, float idy,
                                                float idz, float t_max) {
    const float x0 = (c.lo.x - ox) * idx, x1 = (c.hi.x - ox) * idx;
    const float y0 = (c.lo.y - oy) * idy, y1 = (c.hi.y - oy) * idy;
    const float z0 = (c.lo.z - oz) * idz, z1 = (c.hi.z - oz) * idz;
    const float tn = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), kRayEps));
    const float tf = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), t_max));
    return !(tn > tf);
}

// Which groups can the ray touch at all?  One bit per group, found with all lanes of the warp in step
// (the loop has the same trip count for every ray).  The sphere tests then walk each lane's OWN set
// bits: a warp of incoherent rays pays for the longest candidate list among its lanes, not for the
// union of the groups any of its lanes touches (the per-group `continue` of round 1 kept lanes idle
// while a neighbour tested a group they had already rejected: 13 of 32 lanes active in rt_trace_kernel,
// profiles/r02_render_ncu_summary.md).
static_assert(kMaxClusters <= 32, "one mask bit per sphere group");
__device__ __forceinline__ unsigned cluster_mask(const ClusterBox* clus, int n, float ox, float oy, float oz, float dx,
                                                 float dy, float dz) {
    const float idx = 1.f / dx, idy = 1.f / dy, idz = 1.f / dz;
    unsigned mask = 0u;
    for (int k = 0, s0 = 0; s0 < n; ++k, s0 += kClusterSize)
        mask |= cluster_may_hit(clus[k], ox, oy, oz, idx, idy, idz, 3.0e38f) ? (1u << k) : 0u;
    return mask;
}

// nearest hit of (o, d) with the spheres in shared memory; returns sphere index or -1.  Groups are
// visited in index order and a later sphere replaces the best one only when strictly nearer, so the
// result does not depend on which groups the mask prunes.
__device__ __forceinline__ int intersect_spheres(const float4* sph, const ClusterBox* clus, int n, float ox, float oy,
                                                 float oz, float dx, float dy, float dz, float& t_hit) {
    int best = -1;
    float tb = 3.0e38f;
    unsigned mask = cluster_mask(clus, n, ox, oy, oz, dx, dy, dz);
    while (mask) {
        const int s0 = (__ffs(mask) - 1) * kClusterSize;
        mask &= mask - 1u;
        const int s1 = min(s0 + kClusterSize, n);
        for (int s = s0; s < s1; ++s) {
            const float4 c = sph[s];
            const float lx = ox - c.x, ly = oy - c.y, lz = oz - c.z;
            const float b = lx * dx + ly * dy + lz * dz;
            // discriminant from the rejection of l on d (|l - b d|^2), not b^2 - (|l|^2 - r^2): the latter
            // cancels catastrophically in fp32 when the origin is many radii away (Ray Tracing Gems ch. 7)
            const float qx = fmaf(-b, dx, lx), qy = fmaf(-b, dy, ly), qz = fmaf(-b, dz, lz);
            const float disc = c.w * c.w - (qx * qx + qy * qy + qz * qz);
            const float sq = sqrtf(fmaxf(disc, 0.f));
            float t = -b - sq;
            if (t < kRayEps) t = -b + sq;
            const bool ok = disc >= 0.f && t >= kRayEps && t < tb;   // predicated: no branch per sphere
            tb = ok ? t : tb;
            best = ok ? s : best;
        }
    }
    t_hit = tb;
    return best;
}
__device__ __forceinline__ bool occluded(const float4* sph, const ClusterBox* clus, int n, float ox, float oy, float oz,
                                         float dx, float dy, float dz) {
    unsigned mask = cluster_mask(clus, n, ox, oy, oz, dx, dy, dz);
    bool any = false;
    while (mask && !any) {
        const int s0 = (__ffs(mask) - 1) * kClusterSize;
        mask &= mask - 1u;
        const int s1 = min(s0 + kClusterSize, n);
        for (int s = s0; s < s1; ++s) {
            const float4 c = sph[s];
            const float lx = ox - c.x, ly = oy - c.y, lz = oz - c.z;
            const float b = lx * dx + ly * dy + lz * dz;
            const float qx = fmaf(-b, dx, lx), qy = fmaf(-b, dy, ly), qz = fmaf(-b, dz, lz);
            const float disc = c.w * c.w - (qx * qx + qy * qy + qz * qz);
            const float sq = sqrtf(fmaxf(disc, 0.f));
            any |= disc >= 0.f && (-b - sq >= kRayEps || -b + sq >= kRayEps);
        }
    }
    return any;
}

// Duff et al. orthonormal basis (Mitsuba 3's coordinate_system)
__device__ __forceinline__ void frame_from_normal(float nx, float ny, float nz, float s[3], float t[3]) {
    const float sign = copysignf(1.f, nz);
    const float a = -1.f / (sign + nz);
    const float b = nx * ny * a;
    s[0] = 1.f + sign * nx * nx * a; s[1] = sign * b; s[2] = -sign * nx;
    t[0] = b; t[1] = sign + ny * ny * a; t[2] = -ny;
}
__device__ __forceinline__ float power_heuristic(float a, float b) {
    const float a2 = a * a, b2 = b * b;
    return a2 > 0.f ? a2 / (a2 + b2) : 0.f;
}

// ---- ray generation: path i of the pass <-> global sample (first + i) ----------------------------
// Which slice of the render's samples the current wavefront pass covers.  It lives in device memory
// and is advanced by rt_advance_kernel at the end of a pass, so that one captured CUDA graph of a
// pass can be replayed for every pass of a render (merl_render).
struct PassInfo {
    long long first;   // global sample index of the pass's path 0
    int np;            // paths in this pass (<= cap)
    int pad;
};
__global__ void rt_advance_kernel(PassInfo* pass, const long long total, const int cap, int* primary_live_count) {
    const long long first = pass->first + cap;
    const long long left = total - first;
    pass->first = first;
    pass->np = left >= cap ? cap : (left > 0 ? (int)left : 0);
    *primary_live_count = 0;   // the next pass's primary kernel appends its hits from zero
}

// camera ray of path i of the pass <-> global sample (first + i); Philox stream (pixel, sample, 0)
__device__ __forceinline__ void camera_ray(const RenderDev& R, const int64_t gid, int& pixel, int& s, float4& o4,
                                           float4& d4) {
    const int npix = R.width * R.height;
    pixel = (int)(gid % npix);
    s = R.spp_begin + (int)(gid / npix);
    uint32_t rnd[4];
    philox4x32_10((uint32_t)pixel, (uint32_t)s, 0u, 0u, R.k0, R.k1, rnd);
    const int x = pixel % R.width, y = pixel / R.width;
    const float sx = ((float)x + u01(rnd[0])) / (float)R.width * 2.f - 1.f;
    const float sy = 1.f - ((float)y + u01(rnd[1])) / (float)R.height * 2.f;
    const Camera& c = R.cam;
    const float dx = c.fx + sx * c.rx + sy * c.ux, dy = c.fy + sx * c.ry + sy * c.uy, dz = c.fz + sx * c.rz + sy * c.uz;
    const float inv = rsqrtf(dx * dx + dy * dy + dz * dz);
    o4 = make_float4(c.px, c.py, c.pz, 0.f);
    d4 = make_float4(dx * inv, dy * inv, dz * inv, 0.f);
}

// BSDF-stage buffers of one pass (float4 AoS over `cap` slots)
struct StageBuf {
    float4 *wi4, *woe4, *u4;        // inputs: wi (w = material), emitter direction (z < 0: none), random numbers
    float4 *fe4, *wos4, *w4;        // outputs: (f cos, pdf_b) at woe; (wo, pdf) sampled; (weight, lobe)
};

// Appending to a list shared by the whole grid: one global atomic per CTA and iteration instead of
// one per warp (with one counter for the whole grid the per-warp atomics serialise in L2 and the
// kernels spend most of their time waiting for the returned offsets: profiles/r01_tuning.md §5).
// Every thread of the CTA must call it (CTA-uniform control flow); returns the destination index of
// the calling thread's element when `flag` is set.  s_warp: kAppendWarps ints of shared memory.
constexpr int kAppendWarps = 8;   // 256-thread CTAs
__device__ __forceinline__ int cta_append(const bool flag, int* counter, int* s_warp) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned b = __ballot_sync(0xffffffffu, flag);
    if (lane == 0) s_warp[warp] = __popc(b);
    __syncthreads();
    if (warp == 0) {
        const int c = lane < kAppendWarps ? s_warp[lane] : 0;
        int inc = c;
#pragma unroll
        for (int o = 1; o < kAppendWarps; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += v;
        }
        const int total = __shfl_sync(0xffffffffu, inc, kAppendWarps - 1);
        int base = 0;
        if (lane == 0 && total) base = atomicAdd(counter, total);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (lane < kAppendWarps) s_warp[lane] = base + inc - c;
    }
    __syncthreads();
    return s_warp[warp] + __popc(b & ((1u << lane) - 1u));
}

// ---- trace: intersect, terminate misses, prepare the BSDF-stage inputs ---------------------------
// Live paths.  Every per-bounce kernel walks a list of path indices, so it costs in proportion to the
// paths still alive, not to the pass size: at depth 0 the list holds the camera rays that hit
// something (appended by rt_primary_kernel), from depth 1 on the paths that the previous bounce's
// shade kernel continued.
struct LiveList {
    const int* in;
    const int* count_in;
    int* out;              // written by rt_shade_kernel
    int* count_out;        // zeroed by rt_scan_kernel of the same bounce
};

// A bounce's trace step runs in two halves with a shared-memory queue between them, so that each half
// executes with full warps.  FRONT: intersect; a ray that escapes (or has used its interaction budget)
// is finished on the spot, a ray that hits is pushed onto the CTA's queue.  BACK (once 256 hits are
// queued, and for the rest at the end): surface frame, the interaction's Philox draw, the shadow ray,
// the BSDF-stage inputs.  In one piece the second half ran at the hit rate: after a bounce about half of
// the rays escape, and rt_trace_kernel executed with 13 of 32 lanes active (profiles/r02_render_ncu_summary.md).
constexpr int kHitQueue = 512;   // < 256 left over + 256 pushed per iteration
struct HitQueue {
    float4 d[kHitQueue];   // ray direction
    int i[kHitQueue];      // path index
    int hit[kHitQueue];    // sphere
    float t[kHitQueue];
    int n;                 // slot allocator (the CTA-uniform count lives in a register of every thread)
};

// PRIMARY: the camera ray is generated in registers (depth 0) and path state is written only for the
// rays that hit a surface; a camera ray that escapes costs one film update and no path-state traffic.
// Every thread of the CTA calls it (in_range = false: nothing to trace); returns whether it pushed.
template <bool PRIMARY>
__device__ __forceinline__ bool trace_front(const RenderDev& R, const PathState& P, const float4* sph,
                                            const ClusterBox* clus, const int i, const bool in_range, const int depth,
                                            const int64_t first, float4* __restrict__ film, HitQueue& hq) {
    static_assert(kMaxSpheres <= 256, "the sphere index shares a queue word with the sample index");
    bool push = false;
    int hit = -1, pixel_q = 0, smp_q = 0;
    float t = 0.f;
    float4 d4 = make_float4(0.f, 0.f, 1.f, 0.f);
    if (in_range) {
        int pixel, smp;
        float4 o4;
        bool alive = true;
        if (PRIMARY) {
            camera_ray(R, first + i, pixel, smp, o4, d4);
        } else {
            pixel = P.pixel[i];
            if (pixel < 0) { P.mat[i] = -1; alive = false; }
            else { o4 = P.org[i]; d4 = P.dir[i]; }
        }
        if (alive) {
            hit = intersect_spheres(sph, clus, R.nspheres, o4.x, o4.y, o4.z, d4.x, d4.y, d4.z, t);
            if (hit < 0 || depth >= R.max_depth) {
                float lr = 0.f, lg = 0.f, lb = 0.f;
                if (!PRIMARY) { const float4 l4 = P.rad[i]; lr = l4.x; lg = l4.y; lb = l4.z; }
                if (hit < 0) {
                    // escaped: environment radiance, MIS against the emitter sample of the previous bounce
                    const float w = depth == 0 ? 1.f : power_heuristic(o4.w, 0.0795774715f);
                    const float4 t4 = PRIMARY ? make_float4(1.f, 1.f, 1.f, 0.f) : P.thr[i];
                    lr = fmaf(t4.x * R.env[0], w, lr); lg = fmaf(t4.y * R.env[1], w, lg); lb = fmaf(t4.z * R.env[2], w, lb);
                }   // else: interaction budget used up
                film_add(film, pixel, lr * R.inv_spp, lg * R.inv_spp, lb * R.inv_spp, R.inv_spp);
                if (!PRIMARY) { P.pixel[i] = -1; P.mat[i] = -1; }
            } else {
                push = true;
                if (PRIMARY) { pixel_q = pixel; smp_q = smp; }
            }
        }
    }
    const unsigned b = __ballot_sync(0xffffffffu, push);
    if (b) {
        const int lane = threadIdx.x & 31, leader = __ffs(b) - 1;
        int base = 0;
        if (lane == leader) base = atomicAdd(&hq.n, __popc(b));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (push) {
            const int slot = base + __popc(b & ((1u << lane) - 1u));
            MERL_CHECK(slot >= 0 && slot < kHitQueue);
            // PRIMARY: pixel and sample index ride along (d.w, upper bits of the sphere index) so that the
            // second half does not repeat the 64-bit division of the global sample index
            if (PRIMARY) { d4.w = __int_as_float(pixel_q); hit |= smp_q << 8; }
            hq.d[slot] = d4; hq.i[slot] = i; hq.hit[slot] = hit; hq.t[slot] = t;
        }
    }
    return push;
}

// Second half for queue slot `slot` (valid = false: this thread has none).  Returns true when the path
// has a BSDF query this bounce; path_out = its index.
template <bool PRIMARY>
__device__ __forceinline__ bool trace_back(const RenderDev& R, const PathState& P, const StageBuf& S, const float4* sph,
                                           const ClusterBox* clus, const int* smat, const HitQueue& hq, const int slot,
                                           const bool valid, const int depth, const int64_t first,
                                           float4* __restrict__ film, int* hist, int& path_out) {
    if (!valid) return false;
    const float4 d4 = hq.d[slot];
    const int i = hq.i[slot];
    int hit = hq.hit[slot];
    const float t = hq.t[slot];
    path_out = i;
    int pixel, smp;
    float ox, oy, oz;
    if (PRIMARY) {
        pixel = __float_as_int(d4.w);
        smp = hit >> 8;
        hit &= 255;
        ox = R.cam.px; oy = R.cam.py; oz = R.cam.pz;
    } else {
        pixel = P.pixel[i]; smp = P.sample[i];
        const float4 o4 = P.org[i];
        ox = o4.x; oy = o4.y; oz = o4.z;
    }
    const float dx = d4.x, dy = d4.y, dz = d4.z;
    const float4 c = sph[hit];
    const float hx = fmaf(t, dx, ox), hy = fmaf(t, dy, oy), hz = fmaf(t, dz, oz);
    const float ir = 1.f / c.w;
    float nx = (hx - c.x) * ir, ny = (hy - c.y) * ir, nz = (hz - c.z) * ir;
    const float inv = rsqrtf(nx * nx + ny * ny + nz * nz);
    nx *= inv; ny *= inv; nz *= inv;
    float s[3], tt[3];
    frame_from_normal(nx, ny, nz, s, tt);
    const float wix = -(dx * s[0] + dy * s[1] + dz * s[2]);
    const float wiy = -(dx * tt[0] + dy * tt[1] + dz * tt[2]);
    const float wiz = -(dx * nx + dy * ny + dz * nz);
    if (!(wiz > 0.f)) {   // grazing / inside hit: nothing to shade
        float lr = 0.f, lg = 0.f, lb = 0.f;
        if (!PRIMARY) { const float4 l4 = P.rad[i]; lr = l4.x; lg = l4.y; lb = l4.z; }
        film_add(film, pixel, lr * R.inv_spp, lg * R.inv_spp, lb * R.inv_spp, R.inv_spp);
        if (!PRIMARY) { P.pixel[i] = -1; P.mat[i] = -1; }
        return false;
    }
    // ONE Philox call per surface interaction: four 24-bit uniforms from the high bits of its words (emitter
    // direction, lobe choice, first BSDF number) and a fifth from the low bytes u01() does not use
    uint32_t re[4];
    philox4x32_10((uint32_t)pixel, (uint32_t)smp, (uint32_t)(depth + 1), 0u, R.k0, R.k1, re);
    const float u5 = u01_low(re[0], re[1], re[2]);
    // emitter sampling: uniform direction on the sphere (constant environment), pdf 1/(4 pi)
    const float ez = 1.f - 2.f * u01(re[0]);
    const float er = sqrtf(fmaxf(0.f, 1.f - ez * ez));
    float es, ec;
    sincospif(2.f * u01(re[1]), &es, &ec);
    const float ex = er * ec, ey = er * es;
    float4 woe = make_float4(0.f, 0.f, -1.f, 0.f);
    const float lz = ex * nx + ey * ny + ez * nz;
    if (lz > 0.f && !occluded(sph, clus, R.nspheres, fmaf(kRayEps, nx, hx), fmaf(kRayEps, ny, hy), fmaf(kRayEps, nz, hz), ex, ey, ez))
        woe = make_float4(ex * s[0] + ey * s[1] + ez * s[2], ex * tt[0] + ey * tt[1] + ez * tt[2], lz, 0.f);
    const int m = smat[hit];
    {   // this bounce's queries per material, for the counting sort: one shared atomic per group of
        // converged lanes with the same material
        const unsigned peers = __match_any_sync(__activemask(), m);
        if ((threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(&hist[m], __popc(peers));
    }
    if (PRIMARY) {
        P.pixel[i] = pixel; P.sample[i] = smp;
        P.thr[i] = make_float4(1.f, 1.f, 1.f, 0.f);
        P.rad[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    P.nrm[i] = make_float4(nx, ny, nz, 0.f);
    P.hit[i] = make_float4(hx, hy, hz, 0.f);
    P.mat[i] = m;
    S.wi4[i] = make_float4(wix, wiy, wiz, (float)m);
    S.woe4[i] = woe;
    S.u4[i] = make_float4(u01(re[2]), u01(re[3]), u5, 0.f);
    return true;
}

struct SceneSmem {
    float4 sph[kMaxSpheres];
    ClusterBox clus[kMaxClusters];
    int smat[kMaxSpheres];
};
__device__ __forceinline__ void stage_scene(SceneSmem& sc, const int nspheres, const float4* __restrict__ spheres,
                                            const ClusterBox* __restrict__ clusters, const int* __restrict__ sphere_mat,
                                            int* hist, const int ntables) {
    for (int k = threadIdx.x; k < ntables; k += blockDim.x) hist[k] = 0;
    for (int s = threadIdx.x; s < nspheres; s += blockDim.x) { sc.sph[s] = spheres[s]; sc.smat[s] = sphere_mat[s]; }
    for (int k = threadIdx.x; k * kClusterSize < nspheres; k += blockDim.x) sc.clus[k] = clusters[k];
    __syncthreads();
}
// the CTA's per-material query counts -> the global histogram the scan kernel turns into segments
__device__ __forceinline__ void flush_hist(const int* hist, const int ntables, int* __restrict__ counts) {
    __syncthreads();
    for (int k = threadIdx.x; k < ntables; k += blockDim.x)
        if (hist[k]) atomicAdd(&counts[k], hist[k]);
}

// depth 0: camera ray + first intersection in one kernel, one thread per path of the pass; the
// rays that hit are appended to the depth-0 live list (cta_append)
__global__ void __launch_bounds__(256) rt_primary_kernel(const RenderDev R, const PathState P, const StageBuf S,
                                                         const float4* __restrict__ spheres,
                                                         const ClusterBox* __restrict__ clusters,
                                                         const int* __restrict__ sphere_mat,
                                                         const PassInfo* __restrict__ pass, float4* __restrict__ film,
                                                         int* live_out, int* live_count, int* __restrict__ counts) {
    extern __shared__ int hist[];   // R.ntables ints
    __shared__ SceneSmem sc;
    const int np = pass->np;
    const int64_t first = pass->first;
    if (blockIdx.x * blockDim.x >= np) return;
    stage_scene(sc, R.nspheres, spheres, clusters, sphere_mat, hist, R.ntables);
    __shared__ int s_warp[kAppendWarps];
    __shared__ HitQueue hq;
    if (threadIdx.x == 0) hq.n = 0;
    __syncthreads();
    int nq = 0;   // queued hits, the same in every thread
    for (int j0 = blockIdx.x * blockDim.x; j0 < np; j0 += gridDim.x * blockDim.x) {   // CTA-uniform trip count
        const int i = j0 + threadIdx.x;
        const bool pushed = trace_front<true>(R, P, sc.sph, sc.clus, i, i < np, 0, first, film, hq);
        nq += __syncthreads_count(pushed);
        if (nq >= 256) {
            nq -= 256;
            if (threadIdx.x == 0) hq.n = nq;   // (ordered before the next pushes by cta_append's barriers)
            int path = 0;
            const bool q = trace_back<true>(R, P, S, sc.sph, sc.clus, sc.smat, hq, nq + threadIdx.x, true, 0, first, film, hist, path);
            const int dst = cta_append(q, live_count, s_warp);
            MERL_CHECK(!q || (dst >= 0 && dst < np));
            if (q) live_out[dst] = path;
        }
    }
    if (nq > 0) {
        int path = 0;
        const bool q = trace_back<true>(R, P, S, sc.sph, sc.clus, sc.smat, hq, threadIdx.x, (int)threadIdx.x < nq, 0, first, film, hist, path);
        const int dst = cta_append(q, live_count, s_warp);
        MERL_CHECK(!q || (dst >= 0 && dst < np));
        if (q) live_out[dst] = path;
    }
    flush_hist(hist, R.ntables, counts);
}

// depth >= 1: the paths of the live list
__global__ void __launch_bounds__(256) rt_trace_kernel(const RenderDev R, const PathState P, const StageBuf S,
                                                       const float4* __restrict__ spheres,
                                                       const ClusterBox* __restrict__ clusters,
                                                       const int* __restrict__ sphere_mat,
                                                       const PassInfo* __restrict__ pass, const int depth,
                                                       float4* __restrict__ film, const LiveList L,
                                                       int* __restrict__ counts) {
    extern __shared__ int hist[];   // R.ntables ints
    __shared__ SceneSmem sc;
    const int count = *L.count_in;
    if (blockIdx.x * blockDim.x >= count) return;          // whole CTA beyond the live range
    stage_scene(sc, R.nspheres, spheres, clusters, sphere_mat, hist, R.ntables);
    MERL_CHECK(count >= 0 && count <= pass->np);
    __shared__ HitQueue hq;
    if (threadIdx.x == 0) hq.n = 0;
    __syncthreads();
    int nq = 0;
    for (int j0 = blockIdx.x * blockDim.x; j0 < count; j0 += gridDim.x * blockDim.x) {   // CTA-uniform trip count
        const int j = j0 + threadIdx.x;
        const int i = j < count ? L.in[j] : 0;
        MERL_CHECK(i >= 0 && i < pass->np);
        const bool pushed = trace_front<false>(R, P, sc.sph, sc.clus, i, j < count, depth, 0, film, hq);
        nq += __syncthreads_count(pushed);
        if (nq >= 256) {
            nq -= 256;
            if (threadIdx.x == 0) hq.n = nq;
            int path = 0;
            trace_back<false>(R, P, S, sc.sph, sc.clus, sc.smat, hq, nq + threadIdx.x, true, depth, 0, film, hist, path);
            __syncthreads();   // everyone has read its slot and sees the new hq.n before the next pushes
        }
    }
    if (nq > 0) {
        int path = 0;
        trace_back<false>(R, P, S, sc.sph, sc.clus, sc.smat, hq, threadIdx.x, (int)threadIdx.x < nq, depth, 0, film, hist, path);
    }
    flush_hist(hist, R.ntables, counts);
}

// ---- counting sort of the live queries by material (multi-material scenes) -----------------------
// One CTA of 1024 threads (ntables <= 1024), Hillis-Steele scans in shared memory:
//   seg_off[0..ntables]   exclusive scan of counts (cursor = copy; counts cleared for the next bounce)
//   item_off[0..ntables]  exclusive scan of ceil(count / chunk): the BSDF stage's work items, each at
//                         most `chunk` queries of ONE table; item_off[ntables + 1] = chunk.
// chunk adapts to the wavefront: ~target_items items in total, so fat wavefronts give every CTA a
// few items and thin ones (late bounces) still spread over many CTAs instead of one CTA walking
// many tables.
__global__ void __launch_bounds__(1024) rt_scan_kernel(int* counts, const int ntables, int* seg_off, int* cursor,
                                                       int* item_off, const int target_items, int* next_live_count,
                                                       unsigned long long* __restrict__ nquery) {
    __shared__ int warp_tot[32];
    __shared__ int chunk_s;
    const int k = threadIdx.x, lane = k & 31, warp = k >> 5;
    if (k == 0) *next_live_count = 0;   // this bounce's shade kernel appends the continuing paths
    const int c = k < ntables ? counts[k] : 0;
    // inclusive scan over the 1024 threads: shuffles inside a warp, the 32 warp totals scanned by warp 0
    // (two barriers per scan; the Hillis-Steele form it replaces took twenty)
    auto scan = [&](int v, int& total) {
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += u;
        }
        if (lane == 31) warp_tot[warp] = v;
        __syncthreads();
        if (warp == 0) {
            int w = warp_tot[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += u;
            }
            warp_tot[lane] = w;
        }
        __syncthreads();
        total = warp_tot[31];
        const int r = v + (warp ? warp_tot[warp - 1] : 0);
        __syncthreads();   // warp_tot is reused by the next scan
        return r;
    };
    int total;
    const int inc = scan(c, total);
    if (k < ntables) { seg_off[k] = inc - c; cursor[k] = inc - c; counts[k] = 0; }
    if (k == 0) {
        *nquery += (unsigned long long)total;   // statistics: BSDF queries of the whole render
        int ch = (total + target_items - 1) / max(target_items, 1);
        ch = (ch + 31) & ~31;
        chunk_s = min(max(ch, 128), 8192);
    }
    if (k == ntables - 1) seg_off[ntables] = inc;
    __syncthreads();
    const int chunk = chunk_s;
    const int items = (c + chunk - 1) / chunk;
    int itotal;
    const int iinc = scan(items, itotal);
    if (k < ntables) item_off[k] = iinc - items;
    if (k == ntables - 1) { item_off[ntables] = iinc; item_off[ntables + 1] = chunk; }
}
// Live queries -> their material's segment of the stage buffers.  A CTA takes kScatterPer * 256 list
// entries per iteration, counts them per material in shared memory (one shared atomic per warp and
// material: neighbouring pixels mostly share one), reserves each material's range with ONE global
// atomic per CTA, and copies.  Dynamic shared memory: 2 * ntables ints.
constexpr int kScatterPer = 4;
__global__ void __launch_bounds__(256) rt_scatter_kernel(const PathState P, const StageBuf A, const StageBuf B,
                                                         const PassInfo* __restrict__ pass, int* cursor,
                                                         const int ntables, const LiveList L) {
    extern __shared__ int sm_scatter[];
    int* cnt = sm_scatter;
    int* base = sm_scatter + ntables;
    const int count = *L.count_in;
    const int lane = threadIdx.x & 31;
    const int tile = blockDim.x * kScatterPer;
    if (blockIdx.x * tile >= count) return;
    for (int k = threadIdx.x; k < ntables; k += blockDim.x) cnt[k] = 0;
    __syncthreads();
    for (int j0 = blockIdx.x * tile; j0 < count; j0 += gridDim.x * tile) {   // CTA-uniform trip count
        int idx[kScatterPer], mat[kScatterPer], off[kScatterPer];
#pragma unroll
        for (int e = 0; e < kScatterPer; ++e) {
            const int j = j0 + e * blockDim.x + threadIdx.x;
            idx[e] = j < count ? L.in[j] : -1;
        }
#pragma unroll
        for (int e = 0; e < kScatterPer; ++e) mat[e] = idx[e] >= 0 ? P.mat[idx[e]] : -1;
#pragma unroll
        for (int e = 0; e < kScatterPer; ++e) {
            const unsigned peers = __match_any_sync(0xffffffffu, mat[e]);
            const int leader = __ffs(peers) - 1;
            int b = 0;
            if (mat[e] >= 0 && lane == leader) b = atomicAdd(&cnt[mat[e]], __popc(peers));
            b = __shfl_sync(0xffffffffu, b, leader);
            off[e] = b + __popc(peers & ((1u << lane) - 1u));
        }
        __syncthreads();
        for (int k = threadIdx.x; k < ntables; k += blockDim.x) {
            const int c = cnt[k];
            if (c) { base[k] = atomicAdd(&cursor[k], c); cnt[k] = 0; }
        }
        __syncthreads();
#pragma unroll
        for (int e = 0; e < kScatterPer; ++e) {
            if (mat[e] < 0) continue;
            const int i = idx[e], slot = base[mat[e]] + off[e];
            MERL_CHECK(slot >= 0 && slot < pass->np && mat[e] < ntables);
            P.slot[slot] = i;
            B.wi4[slot] = A.wi4[i]; B.woe4[slot] = A.woe4[i]; B.u4[slot] = A.u4[i];
        }
    }
}

// ---- shade: consume the BSDF-stage outputs, continue or finish the path -------------------------------
__device__ __forceinline__ bool shade_one(const RenderDev& R, const PathState& P, const StageBuf& S, const int slot,
                                          const int i, float4* __restrict__ film);

// One thread per BSDF-stage slot (the live queries of this bounce, nq = seg_off[ntables]); paths
// that continue are appended to the next bounce's live list (cta_append).
// by_path: the stage outputs are indexed by path (by-table sharding: they come out of the per-bounce
// exchange in path order, because slot order differs between ranks), else by slot.
__global__ void __launch_bounds__(256) rt_shade_kernel(const RenderDev R, const PathState P, const StageBuf S,
                                                       const int* __restrict__ nq_ptr, float4* __restrict__ film,
                                                       const LiveList L, const int by_path) {
    __shared__ int s_warp[kAppendWarps];
    const int nq = *nq_ptr;
    for (int s0 = blockIdx.x * blockDim.x; s0 < nq; s0 += gridDim.x * blockDim.x) {   // CTA-uniform trip count
        const int slot = s0 + threadIdx.x;
        int i = -1;
        bool go = false;
        if (slot < nq) {
            i = P.slot[slot];
            MERL_CHECK(i >= 0 && P.pixel[i] >= 0 && P.pixel[i] < R.width * R.height);
            go = shade_one(R, P, S, by_path ? i : slot, i, film);
        }
        const int dst = cta_append(go, L.count_out, s_warp);
        MERL_CHECK(!go || (dst >= 0 && dst < nq));
        if (go) L.out[dst] = i;
    }
}

// true when the path continues with a new ray
__device__ __forceinline__ bool shade_one(const RenderDev& R, const PathState& P, const StageBuf& S, const int slot,
                                          const int i, float4* __restrict__ film) {
    const float4 fe = S.fe4[slot], ws = S.wos4[slot], w = S.w4[slot];
    const float4 t4 = P.thr[i], l4 = P.rad[i];
    float tr = t4.x, tg = t4.y, tb = t4.z;
    const float mis = power_heuristic(0.0795774715f, fe.w) * 12.5663706f;   // weight / pdf_e
    const float lr = fmaf(tr * fe.x * R.env[0], mis, l4.x), lg = fmaf(tg * fe.y * R.env[1], mis, l4.y),
                lb = fmaf(tb * fe.z * R.env[2], mis, l4.z);
    tr *= w.x; tg *= w.y; tb *= w.z;
    if (!(ws.w > 0.f) || !(tr + tg + tb > 0.f)) {
        film_add(film, P.pixel[i], lr * R.inv_spp, lg * R.inv_spp, lb * R.inv_spp, R.inv_spp);
        P.pixel[i] = -1; P.mat[i] = -1;
        return false;
    }
    const float4 n4 = P.nrm[i], h4 = P.hit[i];
    const float nx = n4.x, ny = n4.y, nz = n4.z;
    float s[3], tt[3];
    frame_from_normal(nx, ny, nz, s, tt);
    P.dir[i] = make_float4(ws.x * s[0] + ws.y * tt[0] + ws.z * nx, ws.x * s[1] + ws.y * tt[1] + ws.z * ny,
                           ws.x * s[2] + ws.y * tt[2] + ws.z * nz, 0.f);
    P.org[i] = make_float4(fmaf(kRayEps, nx, h4.x), fmaf(kRayEps, ny, h4.y), fmaf(kRayEps, nz, h4.z), ws.w);
    P.thr[i] = make_float4(tr, tg, tb, 0.f);
    P.rad[i] = make_float4(lr, lg, lb, 0.f);
    return true;
}

// ---- by-table sharding: stage outputs of the tables THIS rank owns, from
// their material-sorted slots back to path order, into the exchange buffer (zeroed before; the other
// ranks fill in the rest, one sum per bounce makes every rank's copy complete)
__global__ void __launch_bounds__(256) rt_unsort_kernel(const PathState P, const StageBuf B, const StageBuf O,
                                                        const int* __restrict__ nq_ptr, const int own_world,
                                                        const int own_rank) {
    const int nq = *nq_ptr;
    for (int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < nq; slot += gridDim.x * blockDim.x) {
        const int m = (int)B.wi4[slot].w;
        if (m % own_world != own_rank) continue;
        const int i = P.slot[slot];
        O.fe4[i] = B.fe4[slot]; O.wos4[i] = B.wos4[slot]; O.w4[i] = B.w4[slot];
    }
}

// ---- BSDF stage over several tables: work items (one table, <= chunk queries) are dealt to the CTAs
// in contiguous blocks, so a CTA restages the table image only when its next item is another table --
template <typename Op, int MINB>
__global__ void __launch_bounds__(kThreadsPerCta, MINB) stream_multi_kernel(const TableDev* __restrict__ tabs,
                                                                            const int ntables,
                                                                            const int* __restrict__ seg_off,
                                                                            const int* __restrict__ item_off,
                                                                            const Op op, const int image_bytes,
                                                                            const int own_world, const int own_rank) {
    extern __shared__ float4 smem4[];
    __shared__ TableDev t;
    float* smem = reinterpret_cast<float*>(smem4);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const WarpQueue wq = warp_queue(reinterpret_cast<char*>(smem4) + image_bytes, warp);
    if (lane == 0) *wq.count = 0;
    const int nitems = item_off[ntables], chunk = item_off[ntables + 1];
    const int ipc = (nitems + (int)gridDim.x - 1) / (int)gridDim.x;
    const int i0 = min((int)blockIdx.x * ipc, nitems), i1 = min(i0 + ipc, nitems);
    if (i0 >= i1) return;
    int m = 0;   // table of item i0: last m with item_off[m] <= i0
    { int a = 0, b = ntables; while (b - a > 1) { const int c = (a + b) >> 1; if (item_off[c] <= i0) a = c; else b = c; } m = a; }
    int staged = -1;
    for (int item = i0; item < i1; ++item) {
        while (item_off[m + 1] <= item) ++m;
        if (own_world > 1 && m % own_world != own_rank) continue;   // by-table sharding: another rank's table
        const int a = seg_off[m] + (item - item_off[m]) * chunk, b = min(a + chunk, seg_off[m + 1]);
        if (m != staged) {
            __syncthreads();   // previous table fully processed (queues flushed) before the image changes
            if (threadIdx.x == 0) t = tabs[m];
            __syncthreads();
            stage_image(t, smem, op.image_floats(t));
            staged = m;
        }
        const SmemView v = make_view(t, smem);
        for (int g0 = a + warp * 32; g0 < b; g0 += kWarpsPerCta * 32) {
            const int64_t g = g0 + lane;
            const unsigned q = op.hot(t, v, g, g < b, lane);
            if (q) wq.items[atomicAdd(wq.count, 1)] = g;
            __syncwarp();
            if (*reinterpret_cast<volatile int*>(wq.count) > kQueueCap - 32) flush_queue(op, t, v, wq, lane);
        }
        flush_queue(op, t, v, wq, lane);
    }
}

}  // namespace merl
