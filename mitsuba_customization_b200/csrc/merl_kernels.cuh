// merl_kernels.cuh -- the streaming kernels of libmerl_b200 (sm_100a).
//
// Every batched entry point (bins / eval / pdf / eval_pdf / sample / backward, SoA and AoS) is one
// instantiation of stream_kernel<Op>: a persistent grid-stride kernel sized to the SM count that
//   - stages the per-table edge/density/CDF image in shared memory once per CTA,
//   - streams queries with 16-byte evict-first loads and stores (SoA: four queries per lane),
//   - gathers one float4 per query from the L2-resident table,
//   - runs ONLY tier 1 of the bin computation (merl_device.cuh) in the loop, branch-free, and
//     defers queries whose tier-1 decision was inside its error margin to a warp-private queue
//     in shared memory; Op::redo() re-does those exactly (tiers 2-3) and overwrites their outputs.
// Included exactly once, by merl_b200.cu.
#pragma once
#include "merl_device.cuh"
#include "merl_slow.cuh"

#include <type_traits>

namespace merl {

// ------------------------------------------------------------------------------------------
// streamed SoA access: VEC consecutive queries per thread, evict-first transactions
// ------------------------------------------------------------------------------------------
#ifndef MERL_CTA_THREADS
#define MERL_CTA_THREADS 256
#endif
constexpr int kThreadsPerCta = MERL_CTA_THREADS;

template <int VEC> struct Pack { float v[VEC]; };

template <int VEC> __device__ __forceinline__ Pack<VEC> load_stream(const float* p, int64_t g);
template <> __device__ __forceinline__ Pack<1> load_stream<1>(const float* p, int64_t g) {
    Pack<1> r; r.v[0] = __ldcs(p + g); return r;
}
template <> __device__ __forceinline__ Pack<4> load_stream<4>(const float* p, int64_t g) {
    const float4 f = __ldcs(reinterpret_cast<const float4*>(p) + g);
    Pack<4> r; r.v[0] = f.x; r.v[1] = f.y; r.v[2] = f.z; r.v[3] = f.w; return r;
}
template <> __device__ __forceinline__ Pack<2> load_stream<2>(const float* p, int64_t g) {
    const float2 f = __ldcs(reinterpret_cast<const float2*>(p) + g);
    Pack<2> r; r.v[0] = f.x; r.v[1] = f.y; return r;
}
template <int VEC> __device__ __forceinline__ void store_stream(float* p, int64_t g, const Pack<VEC>& a);
template <> __device__ __forceinline__ void store_stream<2>(float* p, int64_t g, const Pack<2>& a) {
    __stcs(reinterpret_cast<float2*>(p) + g, make_float2(a.v[0], a.v[1]));
}
template <> __device__ __forceinline__ void store_stream<1>(float* p, int64_t g, const Pack<1>& a) {
    __stcs(p + g, a.v[0]);
}
template <> __device__ __forceinline__ void store_stream<4>(float* p, int64_t g, const Pack<4>& a) {
    __stcs(reinterpret_cast<float4*>(p) + g, make_float4(a.v[0], a.v[1], a.v[2], a.v[3]));
}
// L2 prefetch of what a lane will stream Op::kPrefetch iterations from now: moves the HBM latency off
// the load's critical path at no register cost.  Measured on B200 (2^26 queries, profiles/
// r01_tuning.md): eval+pdf 109 -> 117 Gquery/s, eval 125 -> 131 at distance 1; distance 2 is no better;
// the backward kernels lose 5 %.  The sample kernel lost 1 % while it was instruction-bound and gains
// 3 % (64.5 -> 66.7) since the sampler work that cut its instruction count by 15 %.
template <int VEC> __device__ __forceinline__ void prefetch_stream(const float* p, int64_t g) {
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p + g * VEC));
}
// Register-free input staging (ops with kStageVecs > 0): while a lane works on its current group it
// has an LDGSTS (cp.async, L2-only) in flight that lands the NEXT group's kStageVecs x 16 B in the
// lane's own shared-memory slot, so the HBM + L2 latency of the streamed inputs is off the critical
// path without spending registers on a second buffer.  Each lane reads only what it copied itself:
// no barrier, just cp.async.wait_group.  The slot is re-armed after its contents are in registers.
__device__ __forceinline__ void cp_async16(float4* smem_dst, const float* gsrc) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ Pack<4> load_slot(const float4* slot, int k) {
    const float4 f = slot[k * kThreadsPerCta];
    Pack<4> r; r.v[0] = f.x; r.v[1] = f.y; r.v[2] = f.z; r.v[3] = f.w; return r;
}
template <int VEC> __device__ __forceinline__ unsigned load_active(const uint8_t* p, int64_t g);
template <> __device__ __forceinline__ unsigned load_active<1>(const uint8_t* p, int64_t g) {
    return p ? (p[g] != 0) : 1u;
}
template <> __device__ __forceinline__ unsigned load_active<2>(const uint8_t* p, int64_t g) {
    if (!p) return 0x3u;
    const unsigned w = __ldcs(reinterpret_cast<const unsigned short*>(p) + g);
    return ((w & 0xFFu) != 0) | ((w >> 8) != 0) << 1;
}
template <> __device__ __forceinline__ unsigned load_active<4>(const uint8_t* p, int64_t g) {
    if (!p) return 0xFu;
    const uint32_t w = __ldcs(reinterpret_cast<const uint32_t*>(p) + g);
    return ((w & 0xFFu) != 0) | (((w >> 8) & 0xFFu) != 0) << 1 | (((w >> 16) & 0xFFu) != 0) << 2 |
           ((w >> 24) != 0) << 3;
}

// ------------------------------------------------------------------------------------------
// Deferred exactness.  The streaming loop runs tier 1 only (branch-free, no fp64, ~40 fewer
// registers, and the four queries a lane holds interleave).  A query whose tier-1 decision was
// inside its error margin (~1.3e-3 of uniform queries) is appended to the warp's private queue
// in shared memory and re-done exactly when the queue could overflow or the warp runs out of
// work.  The redo overwrites that query's outputs, so results equal the all-tiers computation
// for every input, including adversarial batches in which every query is uncertain (the warp
// then flushes every iteration: slow, still exact).  No scratch memory, no second launch.
// ------------------------------------------------------------------------------------------
constexpr int kWarpsPerCta = kThreadsPerCta / 32;
constexpr int kQueueCap = 256;                      // entries per warp (8 B each)
constexpr int kQueueBytes = kWarpsPerCta * (kQueueCap * 8 + 8);

struct WarpQueue {
    int64_t* items;   // [kQueueCap] absolute query indices
    int* count;
};
__device__ __forceinline__ WarpQueue warp_queue(char* base, int warp) {
    WarpQueue q;
    q.items = reinterpret_cast<int64_t*>(base) + warp * kQueueCap;
    q.count = reinterpret_cast<int*>(base + kWarpsPerCta * kQueueCap * 8) + warp * 2;
    return q;
}

template <typename Op>
__device__ __forceinline__ void flush_queue(const Op& op, const TableDev& t, const SmemView& v, const WarpQueue& wq,
                                            const int lane) {
    __syncwarp();   // orders this warp's streaming stores before the redo stores
    const int n = *reinterpret_cast<volatile int*>(wq.count);
    for (int e = lane; e < n; e += 32) op.redo(t, v, wq.items[e]);
    __syncwarp();
    if (lane == 0) *reinterpret_cast<volatile int*>(wq.count) = 0;
    __syncwarp();
}

// Op interface:
//   static constexpr int kVec       queries per lane and iteration
//   static constexpr int kPrefetch  L2 prefetch distance in iterations (0 = none); void prefetch(g)
//   static constexpr int kStageVecs 16-byte vectors per lane staged through shared memory one iteration
//                                   ahead (0 = none); void stage(slot, g); unsigned hot_staged(...)
//   int image_floats(t)             floats of the table image the op needs in shared memory
//   unsigned hot(t, v, g, in_range, lane)   tier-1 pass over group g (called by every lane of the
//                                   warp, so ops may use warp collectives); returns bit q set
//                                   when query g*kVec+q must be re-done
//   void redo(t, v, i)              exact pass over the single query i
// Optional members of an Op, detected here so that the other ops stay as they are:
//   static constexpr bool kBlocked   true: every CTA walks ONE contiguous range of groups instead of a grid
//                                    stride (coherent batches: what a CTA sees stays local in cell space)
//   static constexpr int kExtraSmem  bytes of CTA-shared scratch after the queues and staging slots;
//                                    void cta_begin(char*) / cta_end(char*) run by all threads around the loop
template <typename Op, typename = void> struct op_blocked : std::false_type {};
template <typename Op> struct op_blocked<Op, std::void_t<decltype(Op::kBlocked)>> : std::bool_constant<Op::kBlocked> {};
template <typename Op, typename = void> struct op_extra_smem : std::integral_constant<int, 0> {};
template <typename Op> struct op_extra_smem<Op, std::void_t<decltype(Op::kExtraSmem)>> : std::integral_constant<int, Op::kExtraSmem> {};

template <typename Op, int MINB>
__global__ void __launch_bounds__(kThreadsPerCta, MINB) stream_kernel(const TableDev t, const Op op,
                                                                      const int64_t ngroups, const int image_bytes) {
    extern __shared__ float4 smem4[];
    float* smem = reinterpret_cast<float*>(smem4);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const WarpQueue wq = warp_queue(reinterpret_cast<char*>(smem4) + image_bytes, warp);
    if (lane == 0) *wq.count = 0;
    constexpr bool kBlk = op_blocked<Op>::value;
    constexpr int kExtra = op_extra_smem<Op>::value;
    char* const extra = reinterpret_cast<char*>(smem4) + image_bytes + kQueueBytes + Op::kStageVecs * 16 * kThreadsPerCta;
    if constexpr (kExtra > 0) op.cta_begin(extra);
    stage_image(t, smem, op.image_floats(t));   // ends with __syncthreads()
    const SmemView v = make_view(t, smem);
    // grid stride: group g of iteration k is blockIdx * blockDim + k * gridDim * blockDim + thread;
    // blocked: CTA b owns groups [b * per_cta, (b + 1) * per_cta), per_cta a multiple of blockDim
    int64_t first, last, stride;
    if constexpr (kBlk) {
        const int64_t per_cta = ((ngroups + gridDim.x - 1) / gridDim.x + kThreadsPerCta - 1) / kThreadsPerCta * kThreadsPerCta;
        first = min((int64_t)blockIdx.x * per_cta, ngroups);
        last = min(first + per_cta, ngroups);
        stride = kThreadsPerCta;
    } else {
        first = blockIdx.x * (int64_t)blockDim.x;
        last = ngroups;
        stride = (int64_t)gridDim.x * blockDim.x;
    }
    // this lane's staging slot: vector k of the group at slot[k * blockDim] (conflict-free)
    float4* const slot = reinterpret_cast<float4*>(reinterpret_cast<char*>(smem4) + image_bytes + kQueueBytes) + threadIdx.x;
    if (Op::kStageVecs > 0) {
        const int64_t gf = first + threadIdx.x;
        if (gf < last) op.stage(slot, gf);
        cp_async_commit();
    }
    // warp-uniform trip count: flush and some ops use warp-level synchronisation
    for (int64_t g0 = first + (threadIdx.x & ~31); g0 < last; g0 += stride) {
        const int64_t g = g0 + lane;
        if (Op::kPrefetch > 0) {
            const int64_t gp = g + stride * Op::kPrefetch;
            if (gp < last) op.prefetch(gp);
        }
        unsigned m;
        if (Op::kStageVecs > 0) {
            cp_async_wait_all();
            m = op.hot_staged(t, v, g, g < last, slot, g + stride < last ? g + stride : (int64_t)-1);
        } else {
            m = op.hot(t, v, g, g < last, lane);
        }
        if (m) {
#pragma unroll
            for (int q = 0; q < Op::kVec; ++q)
                if ((m >> q) & 1u) {
                    const int at = atomicAdd(wq.count, 1);
                    MERL_CHECK(at >= 0 && at < kQueueCap);
                    wq.items[at] = g * Op::kVec + q;
                }
        }
        __syncwarp();
        if (*reinterpret_cast<volatile int*>(wq.count) > kQueueCap - 32 * Op::kVec) flush_queue(op, t, v, wq, lane);
    }
    flush_queue(op, t, v, wq, lane);
    if constexpr (kExtra > 0) op.cta_end(extra);
}

// pdf of the two-lobe sampler given the theta_h bin (row A5; oracle: eval_pdf_one).  Branch-free:
// the density is read for every lane (clamped slice index) and selected.
__device__ __forceinline__ float pdf_from_bins(const TableDev& t, const SmemView& v, float az, float bz,
                                               const HalfDiff& h, int ih) {
    const int j = max(min((int)(az * (float)t.ni), t.ni - 1), 0);
    const float p = (1.f - t.spec_weight) * bz * 0.318309886f;
    const float ps = fmaf(t.spec_weight * v.dens[j * t.nh + ih], h.slen * rcp_approx(h.A), p);
    return h.A > 0.f ? ps : p;
}
__device__ __forceinline__ int flat_index(const TableDev& t, const Bins& b) {
    MERL_CHECK(b.ih >= 0 && b.ih < t.nh && b.id >= 0 && b.id < t.nd && b.ip >= 0 && b.ip < t.np);
    return (b.ih * t.nd + b.id) * t.np + b.ip;
}

// ------------------------------------------------------------------------------------------
// bins / eval / pdf / eval_pdf   (rows A1-A5)
// ------------------------------------------------------------------------------------------
enum QueryMode { MODE_BINS = 0, MODE_EVAL = 1, MODE_PDF = 2, MODE_EVAL_PDF = 3 };

struct QueryIn {
    const float *wix, *wiy, *wiz, *wox, *woy, *woz;
    const uint8_t* active;
};
struct QueryOut {
    float *r, *g, *b, *pdf;
    int32_t* idx;
};

template <int MODE, int VEC>
struct QueryOp {
    static constexpr int kVec = VEC;
#ifndef MERL_PF_QUERY
#define MERL_PF_QUERY 1
#endif
#ifndef MERL_QUERY_STAGE
#define MERL_QUERY_STAGE 1
#endif
    static constexpr bool kEval = (MODE == MODE_EVAL || MODE == MODE_EVAL_PDF);
    // Measured on B200 (profiles/r01_tuning.md section 8): the gather-free modes (pdf, bins) gain 8 %
    // from staging their inputs through shared memory (199 / 206 Gquery/s = 85 / 88 % of HBM) and
    // then want no L2 prefetch; the modes with a table gather lose 6-12 % with staging (the staged
    // copies compete with the latency-critical gather for the load path) and keep the L2 prefetch.
    static constexpr int kStageVecs = (MERL_QUERY_STAGE && VEC == 4 && !kEval) ? 6 : 0;
    static constexpr int kPrefetch = kStageVecs ? 0 : MERL_PF_QUERY;
    static constexpr bool kPdf = (MODE == MODE_PDF || MODE == MODE_EVAL_PDF);
    QueryIn in;
    QueryOut out;

    __device__ __forceinline__ int image_floats(const TableDev& t) const { return kPdf ? t.image_eval_floats : t.off_dens; }
    __device__ __forceinline__ void prefetch(const int64_t g) const {
        prefetch_stream<VEC>(in.wix, g); prefetch_stream<VEC>(in.wiy, g); prefetch_stream<VEC>(in.wiz, g);
        prefetch_stream<VEC>(in.wox, g); prefetch_stream<VEC>(in.woy, g); prefetch_stream<VEC>(in.woz, g);
    }

    __device__ __forceinline__ void stage(float4* slot, const int64_t g) const {
        cp_async16(slot + 0 * kThreadsPerCta, in.wix + 4 * g); cp_async16(slot + 1 * kThreadsPerCta, in.wiy + 4 * g);
        cp_async16(slot + 2 * kThreadsPerCta, in.wiz + 4 * g); cp_async16(slot + 3 * kThreadsPerCta, in.wox + 4 * g);
        cp_async16(slot + 4 * kThreadsPerCta, in.woy + 4 * g); cp_async16(slot + 5 * kThreadsPerCta, in.woz + 4 * g);
    }
    // inputs of group g come from the lane's slot; the slot is then re-armed with group g_next (< 0: none)
    __device__ __forceinline__ unsigned hot_staged(const TableDev& t, const SmemView& v, const int64_t g,
                                                   const bool in_range, float4* slot, const int64_t g_next) const {
        if (!in_range) return 0u;
        const Pack<4> ax = load_slot(slot, 0), ay = load_slot(slot, 1), az = load_slot(slot, 2),
                      bx = load_slot(slot, 3), by = load_slot(slot, 4), bz = load_slot(slot, 5);
        if (g_next >= 0) stage(slot, g_next);
        cp_async_commit();
        return body(t, v, g, reinterpret_cast<const Pack<VEC>&>(ax), reinterpret_cast<const Pack<VEC>&>(ay),
                    reinterpret_cast<const Pack<VEC>&>(az), reinterpret_cast<const Pack<VEC>&>(bx),
                    reinterpret_cast<const Pack<VEC>&>(by), reinterpret_cast<const Pack<VEC>&>(bz));
    }
    __device__ __forceinline__ unsigned hot(const TableDev& t, const SmemView& v, const int64_t g, const bool in_range,
                                            const int) const {
        if (!in_range) return 0u;
        const Pack<VEC> ax = load_stream<VEC>(in.wix, g), ay = load_stream<VEC>(in.wiy, g),
                        az = load_stream<VEC>(in.wiz, g), bx = load_stream<VEC>(in.wox, g),
                        by = load_stream<VEC>(in.woy, g), bz = load_stream<VEC>(in.woz, g);
        return body(t, v, g, ax, ay, az, bx, by, bz);
    }
    __device__ __forceinline__ unsigned body(const TableDev& t, const SmemView& v, const int64_t g, const Pack<VEC>& ax,
                                             const Pack<VEC>& ay, const Pack<VEC>& az, const Pack<VEC>& bx,
                                             const Pack<VEC>& by, const Pack<VEC>& bz) const {
        const unsigned act = load_active<VEC>(in.active, g);
        int idx[VEC];
        Pack<VEC> pdf;
        bool valid[VEC];
        unsigned uncmask = 0;
#pragma unroll
        for (int q = 0; q < VEC; ++q) {
            valid[q] = ((act >> q) & 1u) && az.v[q] > 0.f && bz.v[q] > 0.f;
            const HalfDiff h = half_diff_setup(ax.v[q], ay.v[q], az.v[q], bx.v[q], by.v[q], bz.v[q]);
            unsigned unc;
            int kc[3];
            const Bins b = compute_bins_fast(t, v, ax.v[q], ay.v[q], az.v[q], bx.v[q], by.v[q], bz.v[q], h, unc, kc);
            uncmask |= (unsigned)((unc != 0u) & valid[q]) << q;
            idx[q] = flat_index(t, b);
            if (kPdf) {
                const float pq = pdf_from_bins(t, v, az.v[q], bz.v[q], h, b.ih);
                pdf.v[q] = valid[q] ? pq : 0.f;
            }
        }
        if (kEval) {
            float4 cell[VEC];
#pragma unroll
            for (int q = 0; q < VEC; ++q) cell[q] = __ldg(&t.cells[idx[q]]);
            Pack<VEC> r, gg, bb;
#pragma unroll
            for (int q = 0; q < VEC; ++q) {
                const float c = valid[q] ? bz.v[q] : 0.f;
                r.v[q] = cell[q].x * c; gg.v[q] = cell[q].y * c; bb.v[q] = cell[q].z * c;
            }
            store_stream<VEC>(out.r, g, r);
            store_stream<VEC>(out.g, g, gg);
            store_stream<VEC>(out.b, g, bb);
        }
        if (kPdf) store_stream<VEC>(out.pdf, g, pdf);
        if (MODE == MODE_BINS) {
            if (VEC == 4) {
                reinterpret_cast<int4*>(out.idx)[g] =
                    make_int4(valid[0] ? idx[0] : -1, valid[VEC > 1 ? 1 : 0] ? idx[VEC > 1 ? 1 : 0] : -1,
                              valid[VEC > 2 ? 2 : 0] ? idx[VEC > 2 ? 2 : 0] : -1,
                              valid[VEC > 3 ? 3 : 0] ? idx[VEC > 3 ? 3 : 0] : -1);
            } else {
#pragma unroll
                for (int q = 0; q < VEC; ++q) out.idx[g * VEC + q] = valid[q] ? idx[q] : -1;
            }
        }
        return uncmask;
    }

    // only queries that were valid and uncertain get here
    __device__ __forceinline__ void redo(const TableDev& t, const SmemView& v, const int64_t i) const {
        const float ax = in.wix[i], ay = in.wiy[i], az = in.wiz[i], bx = in.wox[i], by = in.woy[i], bz = in.woz[i];
        const HalfDiff h = half_diff_setup(ax, ay, az, bx, by, bz);
        const Bins b = compute_bins(t, v, ax, ay, az, bx, by, bz, h);
        const int idx = flat_index(t, b);
        if (kEval) {
            const float4 cell = __ldg(&t.cells[idx]);
            out.r[i] = cell.x * bz; out.g[i] = cell.y * bz; out.b[i] = cell.z * bz;
        }
        if (kPdf) out.pdf[i] = pdf_from_bins(t, v, az, bz, h, b.ih);
        if (MODE == MODE_BINS) out.idx[i] = idx;
    }
};

// ------------------------------------------------------------------------------------------
// sample   (row A6; oracle: draw_direction + eval_pdf_one)
// ------------------------------------------------------------------------------------------
// Largest i in [0, nh-1] with row[i] <= u (row[0] = 0).  Per-lane bisection over the smem-resident
// row with compile-time strides (kMaxNH = 256 -> 8 steps, fully unrolled, no branches): the running
// pointer makes each step one LDS with an immediate offset, one bound test, one compare and one
// predicated add.  A probe past the row's end is rejected by the bound test before its value is
// used; it still reads shared memory, at most 128 floats past the row, which stays inside the CTA's
// allocation because the redo queues (kQueueBytes >= 512) follow the table image.
__device__ __forceinline__ int search_lane(const float* row, int nh, int /*step0*/, float u) {
    static_assert(kMaxNH <= 256 && kQueueBytes >= 512, "search_lane: 8 strides, probes may overrun by 128 floats");
    const float* p = row;
    const float* const end = row + nh;
#pragma unroll
    for (int S = 128; S > 0; S >>= 1) {
        const float e = p[S];
        p = (p + S < end && e <= u) ? p + S : p;
    }
    MERL_CHECK(p >= row && p < end);
    return (int)(p - row);
}
// Same result with a guide table (MERL_SAMPLE_GUIDE): guide[g] = largest i with row[i] <= g / kGuideCells, so the
// answer for u in cell g = floor(u kGuideCells) is guide[g] advanced while the next entry is still <= u.  A cell
// covers 1/kGuideCells of the probability mass whatever its number of bins, so the expected number of forward
// steps is (nh + kGuideCells) / kGuideCells < 2; cells that span many low-mass tail bins (the cdf_floor mass of
// high theta_h) take the bisection instead.  Two unrolled steps keep the four queries of a lane interleaved.
__device__ __forceinline__ int search_guided(const float* row, const uint8_t* guide_row, int nh, float u) {
    const int g = max(min((int)(u * (float)kGuideCells), kGuideCells - 1), 0);
    int i = (u == u) ? (int)guide_row[g] : 0;
    const int last = nh - 1;
    i += (i < last && row[i + 1] <= u);
    i += (i < last && row[i + 1] <= u);
    if (i < last && row[i + 1] <= u) i = search_lane(row, nh, 0, u);   // rare: a long run of low-mass bins
    MERL_CHECK(i >= 0 && i < nh);
    return i;
}
// Same result, warp-cooperative: the warp walks its queries; for each, every lane tests a
// strided share of the row (conflict-free LDS) and __ballot_sync/__popc count the entries <= u.
// row[] is non-decreasing, so that count - 1 is the bisection's answer.
__device__ __forceinline__ int search_warp_coop(const float* cdf, int nh, int my_row, float my_u,
                                                bool my_need, unsigned lane) {
    int result = 0;
    unsigned todo = __ballot_sync(0xffffffffu, my_need);
    while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const int row_i = __shfl_sync(0xffffffffu, my_row, src);
        const float u = __shfl_sync(0xffffffffu, my_u, src);
        const float* row = cdf + row_i * (nh + 1);
        int count = 0;
        for (int base = 0; base < nh; base += 32) {   // entries 0..nh-1 (row[nh] = 1 > u always)
            const int e = base + (int)lane;
            const bool le = (e < nh) && (row[e] <= u);
            count += __popc(__ballot_sync(0xffffffffu, le));
        }
        if ((int)lane == src) result = max(count - 1, 0);
    }
    return result;
}

// wo of either lobe with ONE sincospi and ONE sqrt: both lobes are "radius * (cos, sin)(angle)
// plus a height", so the lobe only selects the angle and the radicand (the unselected lobe's
// values are never formed).  Per lobe this is the oracle's draw_direction: cosine lobe = Mitsuba 3's
// concentric square->disk map lifted to the hemisphere, (x, y, z) = (r c, r s, sqrt((1-|r|)(1+|r|)))
// (z^2 = 1 - r^2 factored so that it keeps full relative accuracy at the horizon); tabulated
// lobe h = (st c, st s, 1 - m) with st = sqrt(m (2 - m)), theta_h bin i of the slice's CDF row,
// solid-angle-uniform inside the bin, phi_h = 2 pi uy, and wo = reflect(wi, h).
__device__ __forceinline__ void draw_direction(const SmemView& v, const float* row, int i, bool spec, float ax,
                                               float ay, float az, float ux, float uy, float& x, float& y, float& z) {
    // cosine lobe: concentric square -> disk
    const float sx = fmaf(2.f, ux, -1.f), sy = fmaf(2.f, uy, -1.f);
    const bool q13 = fabsf(sx) < fabsf(sy);
    const float r = q13 ? sy : sx, rp = q13 ? sx : sy;
    float ph = 0.25f * __fdividef(rp, r);            // in units of pi
    if (q13) ph = 0.5f - ph;
    if (sx == 0.f && sy == 0.f) ph = 0.f;
    const float ar = fabsf(r);
    // tabulated lobe: position inside theta_h bin i
    const float c0 = row[i], c1 = row[i + 1];
    const float w = c1 - c0;
    const float f = w > 0.f ? __fdividef(ux - c0, w) : 0.5f;
    const float m0 = v.medge[i], m1 = v.medge[i + 1];
    const float m = fmaf(f, m1 - m0, m0);
    // shared transcendental work
    float sn, cs;
    sincospif(spec ? 2.f * uy : ph, &sn, &cs);
    const float root = sqrt_approx(spec ? m * (2.f - m) : fmaxf(0.f, (1.f - ar) * (1.f + ar)));   // 1 ulp
    const float rad = spec ? root : r;
    const float px = rad * cs, py = rad * sn, pz = spec ? 1.f - m : root;   // cosine: wo; tabulated: h
    const float c2 = 2.f * fmaf(ax, px, fmaf(ay, py, az * pz));
    x = spec ? fmaf(c2, px, -ax) : px;
    y = spec ? fmaf(c2, py, -ay) : py;
    z = spec ? fmaf(c2, pz, -az) : pz;
}

struct SampleIn {
    const float *wix, *wiy, *wiz, *u1, *u2x, *u2y;
    const uint8_t* active;
};
struct SampleOut {
    float *wox, *woy, *woz, *pdf, *wr, *wg, *wb;
    uint32_t* lobe;
};

template <int VEC, int SEARCH>
struct SampleOp {
    static constexpr int kVec = VEC;
#ifndef MERL_PF_SAMPLE
#define MERL_PF_SAMPLE 1
#endif
#ifndef MERL_SAMPLE_STAGE
#define MERL_SAMPLE_STAGE 0
#endif
#ifndef MERL_SAMPLE_GUIDE
#define MERL_SAMPLE_GUIDE 1
#endif
    // input staging (see cp_async16): the first use of every streamed input was the kernel's top stall
    // (profiles/r01_sample_ncu_summary.md: long_scoreboard 4.0 of 9 cycles per issue at the first FMUL / FSETP
    // on loaded data); staged, the inputs of the next group land in the lane's shared-memory slot while the
    // current group's ~1300 instructions run.  Not with the cooperative search (every lane must reach it).
    static constexpr int kStageVecs = (MERL_SAMPLE_STAGE && VEC == 4 && !SEARCH) ? 6 : 0;
    static constexpr int kPrefetch = kStageVecs ? 0 : MERL_PF_SAMPLE;
    SampleIn in;
    SampleOut out;

    __device__ __forceinline__ int image_floats(const TableDev& t) const { return t.image_full_floats; }
    __device__ __forceinline__ void prefetch(const int64_t g) const {
        prefetch_stream<VEC>(in.wix, g); prefetch_stream<VEC>(in.wiy, g); prefetch_stream<VEC>(in.wiz, g);
        prefetch_stream<VEC>(in.u1, g); prefetch_stream<VEC>(in.u2x, g); prefetch_stream<VEC>(in.u2y, g);
    }
    __device__ __forceinline__ void stage(float4* slot, const int64_t g) const {
        cp_async16(slot + 0 * kThreadsPerCta, in.wix + 4 * g); cp_async16(slot + 1 * kThreadsPerCta, in.wiy + 4 * g);
        cp_async16(slot + 2 * kThreadsPerCta, in.wiz + 4 * g); cp_async16(slot + 3 * kThreadsPerCta, in.u1 + 4 * g);
        cp_async16(slot + 4 * kThreadsPerCta, in.u2x + 4 * g); cp_async16(slot + 5 * kThreadsPerCta, in.u2y + 4 * g);
    }
    __device__ __forceinline__ unsigned hot_staged(const TableDev& t, const SmemView& v, const int64_t g, const bool in_range,
                                                   float4* slot, const int64_t g_next) const {
        if (!in_range) return 0u;
        const Pack<4> ax = load_slot(slot, 0), ay = load_slot(slot, 1), az = load_slot(slot, 2), u1 = load_slot(slot, 3),
                      ux = load_slot(slot, 4), uy = load_slot(slot, 5);
        if (g_next >= 0) stage(slot, g_next);
        cp_async_commit();
        typedef const Pack<VEC>& R;
        return body(t, v, g, true, threadIdx.x & 31, load_active<VEC>(in.active, g), reinterpret_cast<R>(ax),
                    reinterpret_cast<R>(ay), reinterpret_cast<R>(az), reinterpret_cast<R>(u1), reinterpret_cast<R>(ux),
                    reinterpret_cast<R>(uy));
    }

    __device__ __forceinline__ unsigned hot(const TableDev& t, const SmemView& v, const int64_t g, const bool in_range,
                                            const int lane) const {
        if (!SEARCH && !in_range) return 0u;
        const int64_t gl = in_range ? g : 0;
        const Pack<VEC> ax = load_stream<VEC>(in.wix, gl), ay = load_stream<VEC>(in.wiy, gl),
                        az = load_stream<VEC>(in.wiz, gl), u1 = load_stream<VEC>(in.u1, gl),
                        ux = load_stream<VEC>(in.u2x, gl), uy = load_stream<VEC>(in.u2y, gl);
        const unsigned act = in_range ? load_active<VEC>(in.active, gl) : 0u;
        return body(t, v, g, in_range, lane, act, ax, ay, az, u1, ux, uy);
    }
    __device__ __forceinline__ unsigned body(const TableDev& t, const SmemView& v, const int64_t g, const bool in_range,
                                             const int lane, const unsigned act, const Pack<VEC>& ax, const Pack<VEC>& ay,
                                             const Pack<VEC>& az, const Pack<VEC>& u1, const Pack<VEC>& ux,
                                             const Pack<VEC>& uy) const {
        Pack<VEC> bx, by, bz, pdf, wr, wg, wb;
        uint32_t lobe[VEC];
        int idx[VEC];
        bool valid[VEC];
        unsigned uncmask = 0;
#pragma unroll
        for (int q = 0; q < VEC; ++q) {
            const bool ok = ((act >> q) & 1u) && az.v[q] > 0.f;
            const bool spec = u1.v[q] < t.spec_weight;
            lobe[q] = (ok && spec) ? 1u : 0u;
            float x, y, z;
            const int j = max(min((int)(az.v[q] * (float)t.ni), t.ni - 1), 0);
            const float* row = v.cdf + j * (t.nh + 1);
            const int i = SEARCH ? search_warp_coop(v.cdf, t.nh, j, ux.v[q], ok && spec, (unsigned)lane)
                          : MERL_SAMPLE_GUIDE ? search_guided(row, v.guide + j * kGuideCells, t.nh, ux.v[q])
                                              : search_lane(row, t.nh, t.search_step, ux.v[q]);
            draw_direction(v, row, i, spec, ax.v[q], ay.v[q], az.v[q], ux.v[q], uy.v[q], x, y, z);
            valid[q] = ok && z > 0.f;
            const HalfDiff h = half_diff_setup(ax.v[q], ay.v[q], az.v[q], x, y, z);
            unsigned unc;
            int kc[3];
            const Bins b = compute_bins_fast(t, v, ax.v[q], ay.v[q], az.v[q], x, y, z, h, unc, kc);
            uncmask |= (unsigned)((unc != 0u) & valid[q]) << q;
            idx[q] = flat_index(t, b);
            const float pq = pdf_from_bins(t, v, az.v[q], z, h, b.ih);
            pdf.v[q] = valid[q] ? pq : 0.f;
            bx.v[q] = ok ? x : 0.f; by.v[q] = ok ? y : 0.f; bz.v[q] = ok ? z : 0.f;
        }
        float4 cell[VEC];
#pragma unroll
        for (int q = 0; q < VEC; ++q) cell[q] = __ldg(&t.cells[idx[q]]);
#pragma unroll
        for (int q = 0; q < VEC; ++q) {
            const bool good = valid[q] && pdf.v[q] > 0.f;
            const float s = good ? __fdividef(bz.v[q], pdf.v[q]) : 0.f;
            wr.v[q] = cell[q].x * s; wg.v[q] = cell[q].y * s; wb.v[q] = cell[q].z * s;
            pdf.v[q] = good ? pdf.v[q] : 0.f;
        }
        if (in_range) {
            store_stream<VEC>(out.wox, g, bx); store_stream<VEC>(out.woy, g, by); store_stream<VEC>(out.woz, g, bz);
            store_stream<VEC>(out.pdf, g, pdf);
            store_stream<VEC>(out.wr, g, wr); store_stream<VEC>(out.wg, g, wg); store_stream<VEC>(out.wb, g, wb);
            if (VEC == 4) {
                reinterpret_cast<uint4*>(out.lobe)[g] = make_uint4(lobe[0], lobe[VEC > 1 ? 1 : 0],
                                                                  lobe[VEC > 2 ? 2 : 0], lobe[VEC > 3 ? 3 : 0]);
            } else if (VEC == 2) {
                reinterpret_cast<uint2*>(out.lobe)[g] = make_uint2(lobe[0], lobe[VEC > 1 ? 1 : 0]);
            } else {
                out.lobe[g] = lobe[0];
            }
        }
        return in_range ? uncmask : 0u;
    }

    // wo and lobe of an uncertain query are already final (they do not depend on the bins);
    // only pdf and weight are re-derived from the exact bins of (wi, stored wo).
    __device__ __forceinline__ void redo(const TableDev& t, const SmemView& v, const int64_t i) const {
        const float ax = in.wix[i], ay = in.wiy[i], az = in.wiz[i];
        const float x = __ldcg(out.wox + i), y = __ldcg(out.woy + i), z = __ldcg(out.woz + i);
        const HalfDiff h = half_diff_setup(ax, ay, az, x, y, z);
        const Bins b = compute_bins(t, v, ax, ay, az, x, y, z, h);
        const float4 cell = __ldg(&t.cells[flat_index(t, b)]);
        float p = pdf_from_bins(t, v, az, z, h, b.ih);
        const bool good = p > 0.f;
        const float s = good ? __fdividef(z, p) : 0.f;
        out.pdf[i] = good ? p : 0.f;
        out.wr[i] = cell.x * s; out.wg[i] = cell.y * s; out.wb[i] = cell.z * s;
    }
};

// ------------------------------------------------------------------------------------------
// backward   (row A10): grad_cells[idx] += (g_r, g_g, g_b) * cos(theta_o), one 16-byte RED
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void red_add_v4(float4* addr, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
}

// GATHER = 1: each query reads its cell's validity bits (one more L2 sector per query) and masks
// its own contribution.  GATHER = 0: contributions are added unmasked and backward_mask_kernel
// zeroes the channels of invalid cells afterwards in one pass over the table (47 MB of traffic,
// cheaper than the per-query gather once the batch has more queries than the table has cells).
// CTA = 1 (MERL_BWD_CTA_AGGREGATE, coherent batches): blocked work assignment (a CTA walks one contiguous range
// of the batch, so what it sees stays local in cell space), and after the lane- and warp-level folds the
// contribution goes into the warp's PRIVATE accumulator table in shared memory (kWarpTableSlots cells, direct
// mapped: key = cell index, value = float4 sum) instead of straight to L2.  A warp owns its table, and after the
// warp fold the contributing lanes hold distinct cells, so the table needs no atomics: lanes that map to the
// same slot in one step are told apart with one more __match_any_sync (the lowest takes the slot, the others
// go to L2), and a slot that holds another cell is EVICTED (its sum goes to L2 with one vector RED) and taken
// over, which keeps the table on the cells the warp is currently visiting.  Flushed when the warp is done.
// (A first version used one CTA-wide table with shared-memory float atomics: those are CAS loops on sm_100a
// -- ATOMS.CAST.SPIN -- and it lost 15-30 % on every workload, profiles/r02_coherent_bwd.txt.)
constexpr int kWarpTableSlots = 128;
constexpr int kWarpTableLog2 = 7;
constexpr int kCtaTableBytes = kWarpsPerCta * kWarpTableSlots * (4 + 16);

template <int VEC, int AGG, int GATHER, int CTA = 0>
struct BackwardOp {
    static constexpr int kVec = VEC;
    static constexpr bool kBlocked = CTA != 0;
    static constexpr int kExtraSmem = CTA ? kCtaTableBytes : 0;
    static_assert(!CTA || AGG, "the warp tables need the warp fold (distinct cells per step)");
    __device__ __forceinline__ int* warp_keys() const {
        extern __shared__ float4 smem4[];
        return reinterpret_cast<int*>(reinterpret_cast<char*>(smem4) + extra_off) + (threadIdx.x >> 5) * kWarpTableSlots;
    }
    __device__ __forceinline__ float4* warp_vals() const {
        extern __shared__ float4 smem4[];
        return reinterpret_cast<float4*>(reinterpret_cast<char*>(smem4) + extra_off + kWarpsPerCta * kWarpTableSlots * 4) +
               (threadIdx.x >> 5) * kWarpTableSlots;
    }
    __device__ __forceinline__ void cta_begin(char* extra) const {
        int* keys = reinterpret_cast<int*>(extra);
        for (int i = threadIdx.x; i < kWarpsPerCta * kWarpTableSlots; i += blockDim.x) keys[i] = -1;
        // (stage_image's barrier, which follows, orders this before the first use)
    }
    __device__ __forceinline__ void cta_end(char*) const {
        __syncwarp();
        const int* keys = warp_keys();
        const float4* vals = warp_vals();
        for (int i = threadIdx.x & 31; i < kWarpTableSlots; i += 32) {
            const int key = keys[i];
            if (key >= 0) { const float4 a = vals[i]; red_add_v4(grad + key, a.x, a.y, a.z, a.w); }
        }
    }
    // One step of the warp (all 32 lanes call it): lanes with `em` hold DISTINCT cells.
    __device__ __forceinline__ void table_add(const bool em, const int idx, const float v0, const float v1, const float v2,
                                              const float v3, const int lane) const {
        const unsigned slot = ((unsigned)idx * 2654435761u) >> (32 - kWarpTableLog2);
        const unsigned peers = __match_any_sync(0xffffffffu, em ? (int)slot : -1 - lane);
        const bool win = em && lane == __ffs(peers) - 1;
        if (em && !win) red_add_v4(grad + idx, v0, v1, v2, v3);      // another cell of this step took the slot
        if (win) {
            int* keys = warp_keys();
            float4* vals = warp_vals();
            const int key = keys[slot];
            float4 a = vals[slot];
            if (key == idx) {
                a.x += v0; a.y += v1; a.z += v2; a.w += v3;
            } else {
                if (key >= 0) red_add_v4(grad + key, a.x, a.y, a.z, a.w);   // evict
                keys[slot] = idx;
                a = make_float4(v0, v1, v2, v3);
            }
            vals[slot] = a;
        }
        __syncwarp();
    }
#ifndef MERL_PF_BACKWARD
#define MERL_PF_BACKWARD 0
#endif
    static constexpr int kPrefetch = MERL_PF_BACKWARD;
#ifndef MERL_BWD_STAGE
#define MERL_BWD_STAGE 1
#endif
    // input staging (see cp_async16): only the big-batch path, which has no per-query gather
    static constexpr int kStageVecs = (MERL_BWD_STAGE && VEC == 4 && !GATHER) ? 9 : 0;
    QueryIn q;
    const float *gr, *gg, *gb;
    float4* grad;
    int extra_off;   // byte offset of the CTA table in dynamic shared memory (set by the host launcher)

    __device__ __forceinline__ int image_floats(const TableDev& t) const { return t.off_dens; }
    __device__ __forceinline__ void prefetch(const int64_t g) const {
        prefetch_stream<VEC>(q.wix, g); prefetch_stream<VEC>(q.wiy, g); prefetch_stream<VEC>(q.wiz, g);
        prefetch_stream<VEC>(q.wox, g); prefetch_stream<VEC>(q.woy, g); prefetch_stream<VEC>(q.woz, g);
        prefetch_stream<VEC>(gr, g); prefetch_stream<VEC>(gg, g); prefetch_stream<VEC>(gb, g);
    }

    __device__ __forceinline__ void stage(float4* slot, const int64_t g) const {
        cp_async16(slot + 0 * kThreadsPerCta, q.wix + 4 * g); cp_async16(slot + 1 * kThreadsPerCta, q.wiy + 4 * g);
        cp_async16(slot + 2 * kThreadsPerCta, q.wiz + 4 * g); cp_async16(slot + 3 * kThreadsPerCta, q.wox + 4 * g);
        cp_async16(slot + 4 * kThreadsPerCta, q.woy + 4 * g); cp_async16(slot + 5 * kThreadsPerCta, q.woz + 4 * g);
        cp_async16(slot + 6 * kThreadsPerCta, gr + 4 * g); cp_async16(slot + 7 * kThreadsPerCta, gg + 4 * g);
        cp_async16(slot + 8 * kThreadsPerCta, gb + 4 * g);
    }
    // (an out-of-range lane still takes part, for the warp collectives of AGG; its slot holds stale
    // data, which is harmless because its `act` is 0)
    __device__ __forceinline__ unsigned hot_staged(const TableDev& t, const SmemView& v, const int64_t g,
                                                   const bool in_range, float4* slot, const int64_t g_next) const {
        if (!AGG && !in_range) return 0u;
        const Pack<4> ax = load_slot(slot, 0), ay = load_slot(slot, 1), az = load_slot(slot, 2), bx = load_slot(slot, 3),
                      by = load_slot(slot, 4), bz = load_slot(slot, 5), cr = load_slot(slot, 6), cg = load_slot(slot, 7),
                      cb = load_slot(slot, 8);
        if (g_next >= 0) stage(slot, g_next);
        cp_async_commit();
        const unsigned act = in_range ? load_active<VEC>(q.active, g) : 0u;
        typedef const Pack<VEC>& R;
        return body(t, v, g, in_range, threadIdx.x & 31, act, reinterpret_cast<R>(ax), reinterpret_cast<R>(ay),
                    reinterpret_cast<R>(az), reinterpret_cast<R>(bx), reinterpret_cast<R>(by), reinterpret_cast<R>(bz),
                    reinterpret_cast<R>(cr), reinterpret_cast<R>(cg), reinterpret_cast<R>(cb));
    }
    __device__ __forceinline__ unsigned hot(const TableDev& t, const SmemView& v, const int64_t g, const bool in_range,
                                            const int lane) const {
        if (!AGG && !in_range) return 0u;
        const int64_t gl = in_range ? g : 0;
        const Pack<VEC> ax = load_stream<VEC>(q.wix, gl), ay = load_stream<VEC>(q.wiy, gl),
                        az = load_stream<VEC>(q.wiz, gl), bx = load_stream<VEC>(q.wox, gl),
                        by = load_stream<VEC>(q.woy, gl), bz = load_stream<VEC>(q.woz, gl),
                        cr = load_stream<VEC>(gr, gl), cg = load_stream<VEC>(gg, gl), cb = load_stream<VEC>(gb, gl);
        const unsigned act = in_range ? load_active<VEC>(q.active, gl) : 0u;
        return body(t, v, g, in_range, lane, act, ax, ay, az, bx, by, bz, cr, cg, cb);
    }
    __device__ __forceinline__ unsigned body(const TableDev& t, const SmemView& v, const int64_t g, const bool in_range,
                                             const int lane, const unsigned act, const Pack<VEC>& ax, const Pack<VEC>& ay,
                                             const Pack<VEC>& az, const Pack<VEC>& bx, const Pack<VEC>& by,
                                             const Pack<VEC>& bz, const Pack<VEC>& cr, const Pack<VEC>& cg,
                                             const Pack<VEC>& cb) const {
        int idx[VEC];
        bool live[VEC];
        unsigned uncmask = 0;
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
            const bool valid = ((act >> k) & 1u) && az.v[k] > 0.f && bz.v[k] > 0.f;
            const HalfDiff h = half_diff_setup(ax.v[k], ay.v[k], az.v[k], bx.v[k], by.v[k], bz.v[k]);
            unsigned unc;
            int kc[3];
            const Bins b = compute_bins_fast(t, v, ax.v[k], ay.v[k], az.v[k], bx.v[k], by.v[k], bz.v[k], h, unc, kc);
            uncmask |= (unsigned)((unc != 0u) & valid) << k;
            live[k] = valid & (unc == 0u);   // an uncertain query adds nothing here: redo() does its RED
            idx[k] = flat_index(t, b);
        }
        float bits[VEC];
#pragma unroll
        for (int k = 0; k < VEC; ++k)
            bits[k] = GATHER ? __ldg(&reinterpret_cast<const float*>(t.cells)[4 * (int64_t)idx[k] + 3]) : 7.f;
        float a0[VEC], a1[VEC], a2[VEC], a3[VEC];
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
            const int m = (int)bits[k];
            const float c = live[k] ? bz.v[k] : 0.f;
            a0[k] = (m & 1) ? cr.v[k] * c : 0.f; a1[k] = (m & 2) ? cg.v[k] * c : 0.f;
            a2[k] = (m & 4) ? cb.v[k] * c : 0.f; a3[k] = live[k] ? 1.f : 0.f;
        }
        if (AGG && VEC > 1) {
            // a lane's own consecutive queries first (registers only): query k folds into the first earlier
            // query of the lane with the same cell
#pragma unroll
            for (int k = VEC - 1; k > 0; --k) {
#pragma unroll
                for (int j = 0; j < k; ++j) {
                    if (live[k] && live[j] && idx[k] == idx[j]) {
                        a0[j] += a0[k]; a1[j] += a1[k]; a2[j] += a2[k]; a3[j] += a3[k];
                        live[k] = false;
                    }
                }
            }
        }
#pragma unroll
        for (int k = 0; k < VEC; ++k) {
            float v0 = a0[k], v1 = a1[k], v2 = a2[k], v3 = a3[k];
            bool em = live[k];
            if (AGG) {
                // lanes of this warp that hit the same cell fold into the lowest such lane
                const int key = em ? idx[k] : -1 - lane;
                const unsigned peers = __match_any_sync(0xffffffffu, key);
                if (__any_sync(0xffffffffu, peers & (peers - 1))) {
                    if (peers == 0xffffffffu) {      // whole warp on one cell: butterfly
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) {
                            v0 += __shfl_xor_sync(0xffffffffu, v0, o); v1 += __shfl_xor_sync(0xffffffffu, v1, o);
                            v2 += __shfl_xor_sync(0xffffffffu, v2, o); v3 += __shfl_xor_sync(0xffffffffu, v3, o);
                        }
                        em = em && lane == 0;
                    } else {
                        const bool leader = lane == __ffs(peers) - 1;
                        unsigned rest = leader ? (peers & ~(1u << lane)) : 0u;
                        const float m0 = v0, m1 = v1, m2 = v2, m3 = v3;
                        while (__any_sync(0xffffffffu, rest != 0u)) {
                            const int src = rest ? __ffs(rest) - 1 : lane;
                            const float a0 = __shfl_sync(0xffffffffu, m0, src), a1 = __shfl_sync(0xffffffffu, m1, src),
                                        a2 = __shfl_sync(0xffffffffu, m2, src), a3 = __shfl_sync(0xffffffffu, m3, src);
                            if (rest) { v0 += a0; v1 += a1; v2 += a2; v3 += a3; rest &= rest - 1; }
                        }
                        em = em && leader;
                    }
                }
            }
            if (CTA) table_add(em, idx[k], v0, v1, v2, v3, lane);
            else if (em) red_add_v4(grad + idx[k], v0, v1, v2, v3);
        }
        return in_range ? uncmask : 0u;
    }

    __device__ __forceinline__ void redo(const TableDev& t, const SmemView& v, const int64_t i) const {
        const float ax = q.wix[i], ay = q.wiy[i], az = q.wiz[i], bx = q.wox[i], by = q.woy[i], bz = q.woz[i];
        const HalfDiff h = half_diff_setup(ax, ay, az, bx, by, bz);
        const Bins b = compute_bins(t, v, ax, ay, az, bx, by, bz, h);
        const int idx = flat_index(t, b);
        const int m = GATHER ? (int)__ldg(&reinterpret_cast<const float*>(t.cells)[4 * (int64_t)idx + 3]) : 7;
        red_add_v4(grad + idx, (m & 1) ? gr[i] * bz : 0.f, (m & 2) ? gg[i] * bz : 0.f, (m & 4) ? gb[i] * bz : 0.f, 1.f);
    }
};

// grad[c].rgb *= valid bits of cell c  (invalid channels never receive gradient; see BackwardOp)
__global__ void backward_mask_kernel(const float4* __restrict__ cells, float4* __restrict__ grad, const int64_t ncell) {
    for (int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; c < ncell; c += (int64_t)gridDim.x * blockDim.x) {
        const int m = (int)__ldg(&reinterpret_cast<const float*>(cells)[4 * c + 3]);
        if (m == 7) continue;
        float4 g = grad[c];
        if (!(m & 1)) g.x = 0.f;
        if (!(m & 2)) g.y = 0.f;
        if (!(m & 4)) g.z = 0.f;
        grad[c] = g;
    }
}

// ------------------------------------------------------------------------------------------
// AoS variants for the Mitsuba 0.6 wavefront BSDF stage (one float4 per query and buffer)
// ------------------------------------------------------------------------------------------
struct EvalPdfAosOp {
    static constexpr int kVec = 1;
    static constexpr int kPrefetch = 0;
    static constexpr int kStageVecs = 0;
    __device__ __forceinline__ void stage(float4*, int64_t) const {}
    __device__ __forceinline__ unsigned hot_staged(const TableDev&, const SmemView&, int64_t, bool, float4*, int64_t) const { return 0u; }
    const float4 *wi4, *wo4;
    float4* out4;

    __device__ __forceinline__ int image_floats(const TableDev& t) const { return t.image_eval_floats; }
    __device__ __forceinline__ void prefetch(const int64_t g) const {
        prefetch_stream<4>(reinterpret_cast<const float*>(wi4), g); prefetch_stream<4>(reinterpret_cast<const float*>(wo4), g);
    }

    __device__ __forceinline__ unsigned hot(const TableDev& t, const SmemView& v, const int64_t g, const bool in_range,
                                            const int) const {
        if (!in_range) return 0u;
        const float4 a = __ldcs(wi4 + g), b = __ldcs(wo4 + g);
        const bool valid = a.z > 0.f && b.z > 0.f;
        const HalfDiff h = half_diff_setup(a.x, a.y, a.z, b.x, b.y, b.z);
        unsigned unc;
        int kc[3];
        const Bins bn = compute_bins_fast(t, v, a.x, a.y, a.z, b.x, b.y, b.z, h, unc, kc);
        const float4 cell = __ldg(&t.cells[flat_index(t, bn)]);
        const float c = valid ? b.z : 0.f;
        const float pq = pdf_from_bins(t, v, a.z, b.z, h, bn.ih);
        __stcs(out4 + g, make_float4(cell.x * c, cell.y * c, cell.z * c, valid ? pq : 0.f));
        return (unsigned)((unc != 0u) & valid);
    }
    __device__ __forceinline__ void redo(const TableDev& t, const SmemView& v, const int64_t i) const {
        const float4 a = wi4[i], b = wo4[i];
        const HalfDiff h = half_diff_setup(a.x, a.y, a.z, b.x, b.y, b.z);
        const Bins bn = compute_bins(t, v, a.x, a.y, a.z, b.x, b.y, b.z, h);
        const float4 cell = __ldg(&t.cells[flat_index(t, bn)]);
        out4[i] = make_float4(cell.x * b.z, cell.y * b.z, cell.z * b.z, pdf_from_bins(t, v, a.z, b.z, h, bn.ih));
    }
};

struct SampleAosOp {
    static constexpr int kVec = 1;
    static constexpr int kPrefetch = 0;
    static constexpr int kStageVecs = 0;
    __device__ __forceinline__ void stage(float4*, int64_t) const {}
    __device__ __forceinline__ unsigned hot_staged(const TableDev&, const SmemView&, int64_t, bool, float4*, int64_t) const { return 0u; }
    const float4 *wi4, *u4;
    float4 *wo4, *w4;

    __device__ __forceinline__ int image_floats(const TableDev& t) const { return t.image_full_floats; }
    __device__ __forceinline__ void prefetch(const int64_t g) const {
        prefetch_stream<4>(reinterpret_cast<const float*>(wi4), g); prefetch_stream<4>(reinterpret_cast<const float*>(u4), g);
    }

    __device__ __forceinline__ unsigned hot(const TableDev& t, const SmemView& v, const int64_t g, const bool in_range,
                                            const int) const {
        if (!in_range) return 0u;
        const float4 a = __ldcs(wi4 + g), u = __ldcs(u4 + g);
        const bool ok = a.z > 0.f;
        const bool spec = u.x < t.spec_weight;
        float x, y, z;
        const int j = max(min((int)(a.z * (float)t.ni), t.ni - 1), 0);
        const float* row = v.cdf + j * (t.nh + 1);
        const int i = MERL_SAMPLE_GUIDE ? search_guided(row, v.guide + j * kGuideCells, t.nh, u.y)
                                        : search_lane(row, t.nh, t.search_step, u.y);
        draw_direction(v, row, i, spec, a.x, a.y, a.z, u.y, u.z, x, y, z);
        const bool valid = ok && z > 0.f;
        const HalfDiff h = half_diff_setup(a.x, a.y, a.z, x, y, z);
        unsigned unc;
        int kc[3];
        const Bins bn = compute_bins_fast(t, v, a.x, a.y, a.z, x, y, z, h, unc, kc);
        const float4 cell = __ldg(&t.cells[flat_index(t, bn)]);
        const float pq = pdf_from_bins(t, v, a.z, z, h, bn.ih);
        const bool good = valid && pq > 0.f;
        const float s = good ? __fdividef(z, pq) : 0.f;
        __stcs(wo4 + g, make_float4(ok ? x : 0.f, ok ? y : 0.f, ok ? z : 0.f, good ? pq : 0.f));
        __stcs(w4 + g, make_float4(cell.x * s, cell.y * s, cell.z * s, (ok && spec) ? 1.f : 0.f));
        return (unsigned)((unc != 0u) & valid);
    }
    __device__ __forceinline__ void redo(const TableDev& t, const SmemView& v, const int64_t i) const {
        const float4 a = wi4[i];
        const float4 o = __ldcg(wo4 + i);
        const float lobe = __ldcg(w4 + i).w;
        const HalfDiff h = half_diff_setup(a.x, a.y, a.z, o.x, o.y, o.z);
        const Bins bn = compute_bins(t, v, a.x, a.y, a.z, o.x, o.y, o.z, h);
        const float4 cell = __ldg(&t.cells[flat_index(t, bn)]);
        const float p = pdf_from_bins(t, v, a.z, o.z, h, bn.ih);
        const bool good = p > 0.f;
        const float s = good ? __fdividef(o.z, p) : 0.f;
        wo4[i] = make_float4(o.x, o.y, o.z, good ? p : 0.f);
        w4[i] = make_float4(cell.x * s, cell.y * s, cell.z * s, lobe);
    }
};

}  // namespace merl
