// merl_b200.cu -- libmerl_b200: C ABI (include/merl_b200.h) + sm_100a kernels.
//
// Kernels:  repack_kernel, cdf_build_kernel, dens_kernel (table build, one-off; here)
//                          stream_kernel<Op> with Op = QueryOp / SampleOp / BackwardOp /
//                          EvalPdfAosOp / SampleAosOp (merl_kernels.cuh)
#include <nvtx3/nvToolsExt.h>
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <unistd.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <utility>
#include <atomic>
#include <mutex>
#include <string>
#include <vector>

// NCCL's device-side API (2.28+): only its host-visible types are used here -- the symmetric window and
// device communicator are opened on the host and reduced to raw unicast / multicast pointers that
// merl_exchange.cuh's own kernel takes (transport MERL_XCHG_NVLS).  Optional at build time.
#if defined(__has_include)
#if __has_include(<nccl.h>) && __has_include(<nccl_device/impl/comm__types.h>)
#include <nccl.h>
#include <nccl_device/impl/comm__types.h>
#define MERL_HAVE_NCCL_DEVICE 1
#endif
#endif
#ifndef MERL_HAVE_NCCL_DEVICE
#define MERL_HAVE_NCCL_DEVICE 0
#endif

#include "../../include/merl_b200.h"
#include "merl_kernels.cuh"
#include "merl_render.cuh"
#include "merl_exchange.cuh"


using namespace merl;

// resident CTAs per SM the register allocation of the 4-wide sample / backward ops is capped for
#ifndef MERL_QUERY_MINB
#define MERL_QUERY_MINB (3 * kMinbScale)
#endif
#ifndef MERL_SAMPLE_MINB
#define MERL_SAMPLE_MINB (3 * kMinbScale)
#endif
#ifndef MERL_BWD_MINB
#define MERL_BWD_MINB (3 * kMinbScale)
#endif

// ------------------------------------------------------------------------------------------
// errors, counters, profiler ranges
// ------------------------------------------------------------------------------------------
// One NVTX range per public call: named spans in an nsys / ncu timeline of the
// host renderer.  NVTX 3 is header-only and a no-op unless a tool injects itself into the process.
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
    NvtxRange(const NvtxRange&) = delete;
    NvtxRange& operator=(const NvtxRange&) = delete;
};
#define MERL_RANGE(name) NvtxRange nvtx_range_(name)

static thread_local std::string g_err;
static std::atomic<int64_t> g_launches{0};
static std::atomic<int> g_search_mode{0};

static int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}
#define CUDA_TRY(x)                                                                          \
    do {                                                                                     \
        cudaError_t e_ = (x);                                                                \
        if (e_ != cudaSuccess)                                                               \
            return fail(MERL_ERR_CUDA, "%s: %s (%s:%d)", #x, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// No exception crosses the C boundary (include/merl_b200.h): the bodies that allocate host memory run
// inside this guard, which turns bad_alloc / length_error / anything else into a status code.
template <typename F> static int guarded(const char* what, F&& body) {
    try {
        return body();
    } catch (const std::bad_alloc&) {
        return fail(MERL_ERR_NOMEM, "%s: out of host memory", what);
    } catch (const std::exception& e) {
        return fail(MERL_ERR_NOMEM, "%s: %s", what, e.what());
    } catch (...) {
        return fail(MERL_ERR_NOMEM, "%s: unknown C++ exception", what);
    }
}

extern "C" const char* merl_last_error(void) { return g_err.c_str(); }
extern "C" int merl_abi_version(void) { return MERL_B200_ABI_VERSION; }
extern "C" int64_t merl_launch_count(void) { return g_launches.load(); }
extern "C" int merl_set_sample_search(int mode) {
    if (mode != 0 && mode != 1) return fail(MERL_ERR_ARG, "sample search mode must be 0 or 1");
    g_search_mode.store(mode);
    return MERL_OK;
}

// ------------------------------------------------------------------------------------------
// host-side table object
// ------------------------------------------------------------------------------------------
struct merl_table {
    TableDev dev{};
    merl_table_params_t params{};
    int device = 0;
    int num_sms = 0;
    int64_t ncell = 0;
    float4* d_cells = nullptr;
    float* d_image = nullptr;
    double* d_edges64 = nullptr;   // dir_p64 | thr_h64 | thr_d64 | medge64
    double* d_medge64 = nullptr;   // view into d_edges64
    size_t persist_bytes = 0;      // persisting-L2 window actually granted (0 = none)
    // host-buffer pipeline state (lazy)
    std::mutex host_mu;
    void* pipe = nullptr;
};

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
        else prev = -1;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

extern "C" void merl_table_params_default(merl_table_params_t* p) {
    if (!p) return;
    memset(p, 0, sizeof *p);
    p->nh = 90; p->nd = 90; p->np = 180;
    p->theta_h_sqrt = 1;
    p->ni = 32;
    p->clamp_negative = 1;
    p->cdf_nm = 2; p->cdf_nq = 64;
    p->l2_persist = 1;
    p->cdf_floor = 0.01f;
    p->spec_weight = 0.5f;
    p->scale[0] = 1.0 / 1500.0; p->scale[1] = 1.15 / 1500.0; p->scale[2] = 1.66 / 1500.0;
}

// ------------------------------------------------------------------------------------------
// table-build kernels
// ------------------------------------------------------------------------------------------
// Planar double RGB (file layout) -> channel-interleaved float4, prescaled by the channel
// scales, negatives clamped.  .w keeps one validity bit per channel (raw >= 0) as a float.
__global__ void repack_kernel(const double* __restrict__ raw, int64_t ncell, double s0, double s1,
                              double s2, int clamp_negative, float4* __restrict__ cells) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < ncell;
         i += (int64_t)gridDim.x * blockDim.x) {
        double v0 = raw[i] * s0, v1 = raw[i + ncell] * s1, v2 = raw[i + 2 * ncell] * s2;
        int bits = (v0 >= 0.0) | ((v1 >= 0.0) << 1) | ((v2 >= 0.0) << 2);
        if (clamp_negative) { v0 = fmax(v0, 0.0); v1 = fmax(v1, 0.0); v2 = fmax(v2, 0.0); }
        else bits = 7;
        cells[i] = make_float4((float)v0, (float)v1, (float)v2, (float)bits);
    }
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Row CDFs of the tabulated half-vector lobe (the oracle's orc_build_cdf states
// the same quadrature).  One CTA per incident slice j; warp w integrates bins w, w+W, ...
__global__ void __launch_bounds__(256) cdf_build_kernel(TableDev t, const double* __restrict__ medge,
                                                        int nm, int nq, double floor_frac,
                                                        float* __restrict__ cdf_out) {
    extern __shared__ float4 smem4[];
    float* smem = reinterpret_cast<float*>(smem4);
    stage_image(t, smem, t.off_dens);
    const SmemView v = make_view(t, smem);
    double* W = reinterpret_cast<double*>(smem + ((t.off_dens + 3) & ~3));
    const int j = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    const double c0 = (j + 0.5) / t.ni;
    const float ax = (float)sqrt(1.0 - c0 * c0), ay = 0.f, az = (float)c0;
    const double PI = 3.14159265358979323846;
    for (int i = warp; i < t.nh; i += nwarp) {
        double acc = 0.0;
        for (int s = lane; s < nm * nq; s += 32) {
            const int m = s / nq, q = s - m * nq;
            const double mm = medge[i] + (m + 0.5) / nm * (medge[i + 1] - medge[i]);
            const double ct = 1.0 - mm, st = sqrt(mm * (2.0 - mm));
            double sp, cp;
            sincos(2.0 * PI * (q + 0.5) / nq, &sp, &cp);
            const double hx = st * cp, hy = st * sp, hz = ct;
            const double c = (double)ax * hx + (double)az * hz;
            if (!(c > 0.0)) continue;
            const float bx = (float)(2.0 * c * hx - ax), by = (float)(2.0 * c * hy - ay),
                        bz = (float)(2.0 * c * hz - az);
            if (!(bz > 0.f)) continue;
            const HalfDiff h = half_diff_setup(ax, ay, az, bx, by, bz);
            const Bins b = compute_bins(t, v, ax, ay, az, bx, by, bz, h);
            const float4 cell = __ldg(&t.cells[(b.ih * t.nd + b.id) * t.np + b.ip]);
            const double lum = 0.2126 * (double)(cell.x * bz) + 0.7152 * (double)(cell.y * bz) +
                               0.0722 * (double)(cell.z * bz);
            acc += lum * 4.0 * c;
        }
        acc = warp_sum(acc);
        if (lane == 0) W[i] = acc / (nm * nq) * 2.0 * PI * (medge[i + 1] - medge[i]);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double total = 0.0;
        for (int i = 0; i < t.nh; ++i) total += W[i];
        const double msum = medge[t.nh] - medge[0];
        float* row = cdf_out + (size_t)j * (t.nh + 1);
        double run = 0.0;
        row[0] = 0.f;
        for (int i = 0; i < t.nh; ++i) {
            const double uni = (medge[i + 1] - medge[i]) / msum;
            const double P = total > 0.0 ? (1.0 - floor_frac) * W[i] / total + floor_frac * uni : uni;
            run += P;
            row[i + 1] = (float)(run > 1.0 ? 1.0 : run);
        }
        row[t.nh] = 1.f;
    }
}

// dens[j][i] = P_ji / (8 pi (M_{i+1}-M_i)) with P_ji the fp32 CDF's own increment, so that
// pdf() integrates to exactly what sample() draws.
__global__ void dens_kernel(const float* __restrict__ cdf, const double* __restrict__ medge, int ni,
                            int nh, float* __restrict__ dens) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= ni * nh) return;
    const int j = k / nh, i = k - j * nh;
    const double P = (double)cdf[j * (nh + 1) + i + 1] - (double)cdf[j * (nh + 1) + i];
    dens[k] = (float)(P / (8.0 * 3.14159265358979323846 * (medge[i + 1] - medge[i])));
}

// guide[j][g] = largest i in [0, nh-1] with cdf[j][i] <= g / kGuideCells (search_guided starts there)
__global__ void guide_kernel(const float* __restrict__ cdf, int nh, uint8_t* __restrict__ guide) {
    const float* row = cdf + (size_t)blockIdx.x * (nh + 1);
    const float u = (float)threadIdx.x / (float)kGuideCells;
    int i = 0;
    while (i < nh - 1 && row[i + 1] <= u) ++i;
    guide[blockIdx.x * kGuideCells + threadIdx.x] = (uint8_t)i;
}

// ------------------------------------------------------------------------------------------
// host: launch plumbing
// ------------------------------------------------------------------------------------------
static constexpr int kThreads = kThreadsPerCta;
static constexpr int kMinbScale = 256 / kThreadsPerCta;   // MINB values below are per 256-thread CTA
static constexpr size_t kMaxSmemBytes = 100 * 1024;   // per-table image (edges, densities, CDF rows)

template <typename K> static int blocks_per_sm(K kernel, size_t smem) {
    // Cached per (device, kernel instantiation, smem size): cudaFuncSetAttribute and the occupancy
    // figure both apply to the CURRENT device only, and one process may hold tables on several.
    struct Key { int dev; const void* fn; size_t smem; };
    static std::mutex mu;
    static std::vector<std::pair<Key, int>> cache;
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lk(mu);
    for (auto& e : cache)
        if (e.first.dev == dev && e.first.fn == (const void*)kernel && e.first.smem == smem) return e.second;
    if (smem > 48 * 1024) cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int nb = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, kThreads, smem) != cudaSuccess || nb < 1) nb = 1;
    cudaGetLastError();
    cache.push_back({Key{dev, (const void*)kernel, smem}, nb});
    return nb;
}

// Persistent launch: grid = min(work, SMs x resident CTAs), with the table's persisting-L2
// access window attached to the launch itself (no mutation of the caller's stream).
// The window normally covers the table's cells.  A launch that never reads them (the unmasked scatter of
// merl_backward) sets tl_window_base to its own L2-resident working set -- the gradient table -- for the
// duration of the launch call; nullptr = the cells.
static thread_local const void* tl_window_base = nullptr;
struct WindowOverride {
    explicit WindowOverride(const void* p) { tl_window_base = p; }
    ~WindowOverride() { tl_window_base = nullptr; }
};
static bool pin_gradient_window() {
    static const bool on = [] { const char* e = getenv("MERL_BWD_PIN_GRAD"); return !e || atoi(e) != 0; }();
    return on;
}

template <typename K, typename... Args>
static int launch_persistent(const merl_table* t, K kernel, size_t smem, int64_t items, cudaStream_t stream,
                             Args... args) {
    if (items <= 0) return MERL_OK;
    const int64_t want = (items + kThreads - 1) / kThreads;
    const int64_t cap = (int64_t)t->num_sms * blocks_per_sm(kernel, smem);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)(want < cap ? want : cap));
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    if (t->persist_bytes) {
        attr[0].id = cudaLaunchAttributeAccessPolicyWindow;
        attr[0].val.accessPolicyWindow.base_ptr = tl_window_base ? const_cast<void*>(tl_window_base) : (void*)t->d_cells;
        attr[0].val.accessPolicyWindow.num_bytes = t->persist_bytes;
        attr[0].val.accessPolicyWindow.hitRatio = 1.0f;
        attr[0].val.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        attr[0].val.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
    }
    CUDA_TRY(cudaLaunchKernelEx(&cfg, kernel, args...));
    g_launches.fetch_add(1);
    return MERL_OK;
}

static inline bool aligned(const void* p, size_t a) { return p == nullptr || ((uintptr_t)p % a) == 0; }

// Shared-memory layout of the streaming kernels: [table image, 128-byte rounded][warp queues].
template <typename Op, int MINB>
static int launch_op(const merl_table* t, const Op& op, int image_floats, int64_t ngroups, cudaStream_t s) {
    const int img = (int)(((size_t)image_floats * sizeof(float) + 127) & ~(size_t)127);
    const size_t stage = (size_t)Op::kStageVecs * 16 * kThreads;
    return launch_persistent(t, stream_kernel<Op, MINB>, (size_t)img + kQueueBytes + stage + op_extra_smem<Op>::value, ngroups, s,
                             t->dev, op, ngroups, img);
}
// byte offset of an op's CTA scratch (Op::kExtraSmem) in the kernel's dynamic shared memory
template <typename Op> static int extra_smem_offset(int image_floats) {
    const int img = (int)(((size_t)image_floats * sizeof(float) + 127) & ~(size_t)127);
    return img + kQueueBytes + Op::kStageVecs * 16 * kThreads;
}
static inline int eval_floats(const merl_table* t, bool pdf) { return pdf ? t->dev.image_eval_floats : t->dev.off_dens; }

static QueryIn offset_in(QueryIn q, int64_t o) {
    q.wix += o; q.wiy += o; q.wiz += o; q.wox += o; q.woy += o; q.woz += o;
    if (q.active) q.active += o;
    return q;
}
static QueryOut offset_out(QueryOut q, int64_t o) {
    if (q.r) q.r += o; if (q.g) q.g += o; if (q.b) q.b += o; if (q.pdf) q.pdf += o; if (q.idx) q.idx += o;
    return q;
}

// Four queries per lane when every buffer allows 16-byte transactions; the (< 4 query) tail and
// unaligned buffers take the one-query-per-lane instantiation of the same op.
template <int MODE>
static int run_query_mode(const merl_table* t, const QueryIn& in, int64_t n, const QueryOut& out, cudaStream_t s) {
    const int fl = eval_floats(t, MODE == MODE_PDF || MODE == MODE_EVAL_PDF);
    const bool vec = aligned(in.wix, 16) && aligned(in.wiy, 16) && aligned(in.wiz, 16) && aligned(in.wox, 16) &&
                     aligned(in.woy, 16) && aligned(in.woz, 16) && aligned(in.active, 4) && aligned(out.r, 16) &&
                     aligned(out.g, 16) && aligned(out.b, 16) && aligned(out.pdf, 16) && aligned(out.idx, 16);
    int64_t done = 0;
    if (vec && n >= 4) {
        const int rc = launch_op<QueryOp<MODE, 4>, MERL_QUERY_MINB>(t, QueryOp<MODE, 4>{in, out}, fl, n / 4, s);
        if (rc) return rc;
        done = (n / 4) * 4;
    }
    if (done < n)
        return launch_op<QueryOp<MODE, 1>, 4 * kMinbScale>(t, QueryOp<MODE, 1>{offset_in(in, done), offset_out(out, done)}, fl,
                                              n - done, s);
    return MERL_OK;
}

static int check_query_args(const merl_table* t, const float* a, const float* b, const float* c, const float* d,
                            const float* e, const float* f, int64_t n) {
    if (!t) return fail(MERL_ERR_ARG, "table handle is NULL");
    if (n < 0) return fail(MERL_ERR_ARG, "n must be >= 0");
    if (n > 0 && (!a || !b || !c || !d || !e || !f)) return fail(MERL_ERR_ARG, "query pointer is NULL");
    return MERL_OK;
}

// ------------------------------------------------------------------------------------------
// host: table creation (rows A3, A8)
// ------------------------------------------------------------------------------------------
static bool is_pow2(int x) { return x > 0 && (x & (x - 1)) == 0; }
static int round4(int x) { return (x + 3) & ~3; }

static int upload_sampler(merl_table* t, const float* cdf_host, cudaStream_t s) {
    // cdf_host == NULL: build on the device from the table; otherwise install the given rows.
    const TableDev& d = t->dev;
    float* d_cdf = t->d_image + d.off_cdf;
    if (cdf_host) {
        CUDA_TRY(cudaMemcpyAsync(d_cdf, cdf_host, sizeof(float) * d.ni * (d.nh + 1), cudaMemcpyHostToDevice, s));
    } else {
        const size_t smem = (size_t)round4(d.off_dens) * sizeof(float) + sizeof(double) * d.nh;
        cudaFuncSetAttribute(cdf_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxSmemBytes);
        cdf_build_kernel<<<d.ni, kThreads, smem, s>>>(d, t->d_medge64, t->params.cdf_nm, t->params.cdf_nq,
                                                     (double)t->params.cdf_floor, d_cdf);
        CUDA_TRY(cudaGetLastError());
        g_launches.fetch_add(1);
    }
    const int nd = d.ni * d.nh;
    dens_kernel<<<(nd + 255) / 256, 256, 0, s>>>(d_cdf, t->d_medge64, d.ni, d.nh, t->d_image + d.off_dens);
    CUDA_TRY(cudaGetLastError());
    guide_kernel<<<d.ni, kGuideCells, 0, s>>>(d_cdf, d.nh, reinterpret_cast<uint8_t*>(t->d_image + d.off_guide));
    CUDA_TRY(cudaGetLastError());
    g_launches.fetch_add(2);
    return MERL_OK;
}

static void free_table(merl_table* t) {
    if (!t) return;
    DeviceGuard g(t->device);
    cudaFree(t->d_cells);
    cudaFree(t->d_image);
    cudaFree(t->d_edges64);
    delete t;
}

static void free_pipe(merl_table* t);

// Either `planar` (file-layout doubles: prescale + repack on the device, CDF quadrature on the device)
// or `cells_host` + `cdf_host` (a packed sidecar: the float4 table and the CDF rows as they were on
// the device when it was written; uploaded as they are).
static int build_table(const double* planar, const merl_table_params_t* pin, int device, merl_table_t** out,
                       const float* cells_host = nullptr, const float* cdf_host = nullptr) {
    if ((!planar && !(cells_host && cdf_host)) || !out) return fail(MERL_ERR_ARG, "NULL argument");
    merl_table_params_t p;
    if (pin) p = *pin; else merl_table_params_default(&p);
    if (p.nh < 2 || p.nh > kMaxNH || p.nd < 2 || p.nd > kMaxND || p.np < 2 || p.np > kMaxNP)
        return fail(MERL_ERR_UNSUPPORTED, "resolution %dx%dx%d outside [2..%d]x[2..%d]x[2..%d]", p.nh, p.nd, p.np,
                    kMaxNH, kMaxND, kMaxNP);
    if (!is_pow2(p.ni) || p.ni > kMaxNI) return fail(MERL_ERR_ARG, "ni must be a power of two <= %d", kMaxNI);
    if (!(p.spec_weight >= 0.f && p.spec_weight <= 1.f)) return fail(MERL_ERR_ARG, "spec_weight outside [0,1]");
    if (p.cdf_nm < 1 || p.cdf_nq < 1) return fail(MERL_ERR_ARG, "cdf quadrature orders must be >= 1");
    int ndev = 0;
    CUDA_TRY(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(MERL_ERR_ARG, "device %d not in [0,%d)", device, ndev);
    DeviceGuard guard(device);

    merl_table* t = new merl_table();
    t->params = p;
    t->device = device;
    t->ncell = (int64_t)p.nh * p.nd * p.np;
    cudaDeviceGetAttribute(&t->num_sms, cudaDevAttrMultiProcessorCount, device);
    TableDev& d = t->dev;
    d.nh = p.nh; d.nd = p.nd; d.np = p.np; d.ni = p.ni; d.theta_h_sqrt = p.theta_h_sqrt;
    d.off_thr_d = round4(p.nh + 1);
    d.off_dir_p = d.off_thr_d + round4(p.nd + 1);
    d.off_dens = d.off_dir_p + round4(2 * (p.np + 1));
    d.off_cdf = d.off_dens + round4(p.ni * p.nh);
    d.off_medge = d.off_cdf + round4(p.ni * (p.nh + 1));
    d.image_eval_floats = d.off_dens + p.ni * p.nh;
    d.off_guide = d.off_medge + round4(p.nh + 1);
    d.image_full_floats = d.off_guide + p.ni * kGuideCells / 4;
    if ((size_t)round4(d.image_full_floats) * sizeof(float) > kMaxSmemBytes) {
        const int need = d.image_full_floats;
        delete t;
        return fail(MERL_ERR_UNSUPPORTED, "sampler tables (%d floats) exceed the %zu-byte shared-memory image", need,
                    kMaxSmemBytes);
    }
    const double PI = 3.14159265358979323846;
    d.kh = (float)(p.theta_h_sqrt ? (double)p.nh * p.nh * 2.0 / PI : p.nh * 2.0 / PI);
    d.kd = (float)(p.nd * 2.0 / PI);
    d.kp = (float)(p.np / PI);
    d.spec_weight = p.spec_weight;
    d.search_step = 1;
    while (d.search_step * 2 <= p.nh - 1) d.search_step *= 2;

    // bin edges in double (host libm), then rounded to fp32 for tier 1
    const int n64 = (p.nh + 1) + (p.nd + 1) + 2 * (p.np + 1) + (p.nh + 1);
    std::vector<double> e64(n64);
    double* dir_p = e64.data();                 // first: double2 loads need 16-byte alignment
    double* thr_h = dir_p + 2 * (p.np + 1);
    double* thr_d = thr_h + p.nh + 1;
    double* medge = thr_d + p.nd + 1;
    for (int i = 0; i <= p.nh; ++i) {
        const double u = (double)i / p.nh;
        const double th = (PI / 2.0) * (p.theta_h_sqrt ? u * u : u);
        const double tn = tan(th);
        thr_h[i] = (i == 0) ? 0.0 : (i == p.nh ? INFINITY : tn * tn);
        const double sh = sin(0.5 * th);
        medge[i] = 2.0 * sh * sh;
    }
    for (int k = 0; k <= p.nd; ++k) {
        const double c = cos((double)k / p.nd * (PI / 2.0));
        thr_d[k] = (k == 0) ? 4.0 : (k == p.nd ? -4.0 : c * c);
    }
    for (int k = 0; k <= p.np; ++k) {
        const double a = (double)k / p.np * PI;
        dir_p[2 * k] = (k == 0) ? 1.0 : (k == p.np ? -1.0 : cos(a));
        dir_p[2 * k + 1] = (k == 0 || k == p.np) ? 0.0 : sin(a);
    }
    std::vector<float> img(round4(d.image_full_floats), 0.f);
    for (int i = 0; i <= p.nh; ++i) { img[i] = (float)thr_h[i]; img[d.off_medge + i] = (float)medge[i]; }
    for (int k = 0; k <= p.nd; ++k) img[d.off_thr_d + k] = (float)thr_d[k];
    for (int k = 0; k < 2 * (p.np + 1); ++k) img[d.off_dir_p + k] = (float)dir_p[k];

    cudaStream_t s = nullptr;
    double* d_raw = nullptr;
    int rc = MERL_OK;
    auto bail = [&](int code) { cudaFree(d_raw); free_table(t); return code; };
#define TRY_OR_BAIL(x)                                                                            \
    do {                                                                                          \
        cudaError_t e_ = (x);                                                                     \
        if (e_ != cudaSuccess)                                                                    \
            return bail(fail(MERL_ERR_CUDA, "%s: %s (%s:%d)", #x, cudaGetErrorString(e_), __FILE__, __LINE__)); \
    } while (0)
    TRY_OR_BAIL(cudaMalloc(&t->d_cells, sizeof(float4) * t->ncell));
    TRY_OR_BAIL(cudaMalloc(&t->d_image, sizeof(float) * img.size()));
    TRY_OR_BAIL(cudaMalloc(&t->d_edges64, sizeof(double) * n64));
    if (planar) {
        TRY_OR_BAIL(cudaMalloc(&d_raw, sizeof(double) * 3 * t->ncell));
        TRY_OR_BAIL(cudaMemcpyAsync(d_raw, planar, sizeof(double) * 3 * t->ncell, cudaMemcpyHostToDevice, s));
    } else {
        TRY_OR_BAIL(cudaMemcpyAsync(t->d_cells, cells_host, sizeof(float4) * t->ncell, cudaMemcpyHostToDevice, s));
    }
    TRY_OR_BAIL(cudaMemcpyAsync(t->d_image, img.data(), sizeof(float) * img.size(), cudaMemcpyHostToDevice, s));
    TRY_OR_BAIL(cudaMemcpyAsync(t->d_edges64, e64.data(), sizeof(double) * n64, cudaMemcpyHostToDevice, s));
    d.cells = t->d_cells;
    d.image = t->d_image;
    d.dir_p64 = reinterpret_cast<const double2*>(t->d_edges64);
    d.thr_h64 = t->d_edges64 + 2 * (p.np + 1);
    d.thr_d64 = d.thr_h64 + p.nh + 1;
    t->d_medge64 = const_cast<double*>(d.thr_d64) + p.nd + 1;
    if (planar) {
        repack_kernel<<<t->num_sms * 4, 256, 0, s>>>(d_raw, t->ncell, p.scale[0], p.scale[1], p.scale[2],
                                                     p.clamp_negative, t->d_cells);
        TRY_OR_BAIL(cudaGetLastError());
        g_launches.fetch_add(1);
    }
    rc = upload_sampler(t, planar ? nullptr : cdf_host, s);
    if (rc) return bail(rc);
    TRY_OR_BAIL(cudaStreamSynchronize(s));
    cudaFree(d_raw);
    d_raw = nullptr;

    // L2 persistence (BASELINE.json: "pinned in L2 with a cudaAccessPolicyWindow")
    if (p.l2_persist) {
        int max_persist = 0, max_window = 0;
        cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, device);
        cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, device);
        size_t want = sizeof(float4) * (size_t)t->ncell;
        size_t cur = 0;
        cudaDeviceGetLimit(&cur, cudaLimitPersistingL2CacheSize);
        if (max_persist > 0 && max_window > 0) {
            const size_t lim = want < (size_t)max_persist ? want : (size_t)max_persist;
            if (cur < lim) cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, lim);
            t->persist_bytes = want < (size_t)max_window ? want : (size_t)max_window;
        }
        cudaGetLastError();  // persistence is an optimisation; never fail creation on it
    }
#undef TRY_OR_BAIL
    *out = t;
    return MERL_OK;
}

static int merl_table_create_from_host_impl(const double* planar_rgb, const merl_table_params_t* p, int device,
                                           merl_table_t** out) {
    MERL_RANGE("merl_table_create_from_host");
    return build_table(planar_rgb, p, device, out);
}
extern "C" int merl_table_create_from_host(const double* planar_rgb, const merl_table_params_t* p, int device,
                                           merl_table_t** out) {
    return guarded("merl_table_create_from_host", [&]() -> int { return merl_table_create_from_host_impl(planar_rgb, p, device, out); });
}

static int merl_table_create_from_file_impl(const char* path, const merl_table_params_t* pin, int device,
                                           merl_table_t** out) {
    MERL_RANGE("merl_table_create_from_file");
    if (!path || !out) return fail(MERL_ERR_ARG, "NULL argument");
    FILE* f = fopen(path, "rb");
    if (!f) return fail(MERL_ERR_IO, "cannot open '%s'", path);
    int32_t dims[3];
    if (fread(dims, sizeof(int32_t), 3, f) != 3) { fclose(f); return fail(MERL_ERR_IO, "'%s': short header", path); }
    merl_table_params_t p;
    if (pin) p = *pin; else merl_table_params_default(&p);
    // a wrong or corrupt file must fail here, before anything is sized from its header
    if (dims[0] < 1 || dims[1] < 1 || dims[2] < 1 || dims[0] > (1 << 20) || dims[1] > (1 << 20) || dims[2] > (1 << 20)) {
        fclose(f);
        return fail(MERL_ERR_IO, "'%s': implausible header %d x %d x %d", path, dims[0], dims[1], dims[2]);
    }
    if (p.nh == 0 && p.nd == 0 && p.np == 0) { p.nh = dims[0]; p.nd = dims[1]; p.np = dims[2]; }
    if (p.nh < 2 || p.nh > kMaxNH || p.nd < 2 || p.nd > kMaxND || p.np < 2 || p.np > kMaxNP) {
        fclose(f);
        return fail(MERL_ERR_UNSUPPORTED, "'%s': resolution %dx%dx%d outside [2..%d]x[2..%d]x[2..%d]", path, p.nh, p.nd,
                    p.np, kMaxNH, kMaxND, kMaxNP);
    }
    const int64_t n = (int64_t)dims[0] * dims[1] * dims[2];
    // the public reader's check: the product of the header must equal the expected cell count
    if (n != (int64_t)p.nh * p.nd * p.np) {
        fclose(f);
        return fail(MERL_ERR_IO, "'%s': header %dx%dx%d does not match the expected %dx%dx%d cells", path, dims[0],
                    dims[1], dims[2], p.nh, p.nd, p.np);
    }
    {   // the payload must be there before it is allocated for
        fseek(f, 0, SEEK_END);
        const long long have = (long long)ftell(f), need = 12 + 24 * (long long)n;
        fseek(f, 12, SEEK_SET);
        if (have < need) {
            fclose(f);
            return fail(MERL_ERR_IO, "'%s': %lld bytes, a %d x %d x %d table needs %lld", path, have, dims[0], dims[1],
                        dims[2], need);
        }
    }
    std::vector<double> raw((size_t)3 * n);
    const size_t got = fread(raw.data(), sizeof(double), raw.size(), f);
    fclose(f);
    if (got != raw.size()) return fail(MERL_ERR_IO, "'%s': expected %zu doubles, read %zu", path, raw.size(), got);
    return build_table(raw.data(), &p, device, out);
}
extern "C" int merl_table_create_from_file(const char* path, const merl_table_params_t* pin, int device,
                                           merl_table_t** out) {
    return guarded("merl_table_create_from_file", [&]() -> int { return merl_table_create_from_file_impl(path, pin, device, out); });
}

extern "C" int merl_table_destroy(merl_table_t* t) {
    if (!t) return MERL_OK;
    free_pipe(t);
    free_table(t);
    return MERL_OK;
}

extern "C" int merl_table_info(const merl_table_t* t, merl_table_info_t* o) {
    if (!t || !o) return fail(MERL_ERR_ARG, "NULL argument");
    memset(o, 0, sizeof *o);
    o->nh = t->dev.nh; o->nd = t->dev.nd; o->np = t->dev.np; o->ni = t->dev.ni;
    o->theta_h_sqrt = t->dev.theta_h_sqrt; o->device = t->device;
    o->ncell = t->ncell;
    o->table_bytes = t->ncell * (int64_t)sizeof(float4);
    o->d_cells = t->d_cells;
    o->d_cdf = t->d_image + t->dev.off_cdf;
    o->d_dens = t->d_image + t->dev.off_dens;
    o->spec_weight = t->dev.spec_weight;
    o->smem_eval_bytes = (int32_t)(t->dev.image_eval_floats * sizeof(float)) + kQueueBytes;
    o->smem_sample_bytes = (int32_t)(t->dev.image_full_floats * sizeof(float)) + kQueueBytes;
    return MERL_OK;
}

extern "C" int merl_table_set_cdf(merl_table_t* t, const float* cdf_host) {
    if (!t || !cdf_host) return fail(MERL_ERR_ARG, "NULL argument");
    DeviceGuard g(t->device);
    const int rc = upload_sampler(t, cdf_host, nullptr);
    if (rc) return rc;
    CUDA_TRY(cudaStreamSynchronize(nullptr));
    return MERL_OK;
}

// The float4 table is the differentiable parameter (merl_table_info_t.d_cells); an optimiser updates
// it in place on the device.  This re-derives what depends on it -- the sampler's CDF rows and the
// density table -- asynchronously on `stream` (Mitsuba 3: parameters_changed({"data"})).
extern "C" int merl_table_refresh_sampler(merl_table_t* t, merl_stream_t stream) {
    MERL_RANGE("merl_table_refresh_sampler");
    if (!t) return fail(MERL_ERR_ARG, "NULL argument");
    DeviceGuard g(t->device);
    return upload_sampler(t, nullptr, (cudaStream_t)stream);
}

// The optimiser's copy of the parameter -> the table the kernels read (Dr.Jit replaces parameter
// arrays rather than writing into them), then what depends on it.
extern "C" int merl_table_update_cells(merl_table_t* t, const void* cells4_device, merl_stream_t stream) {
    MERL_RANGE("merl_table_update_cells");
    if (!t || !cells4_device) return fail(MERL_ERR_ARG, "NULL argument");
    DeviceGuard g(t->device);
    if (cells4_device != (const void*)t->d_cells)
        CUDA_TRY(cudaMemcpyAsync(t->d_cells, cells4_device, sizeof(float4) * t->ncell, cudaMemcpyDeviceToDevice,
                                 (cudaStream_t)stream));
    return upload_sampler(t, nullptr, (cudaStream_t)stream);
}

extern "C" int merl_table_download(const merl_table_t* t, float* cells4, float* cdf, float* dens) {
    if (!t) return fail(MERL_ERR_ARG, "NULL argument");
    DeviceGuard g(t->device);
    const TableDev& d = t->dev;
    if (cells4) CUDA_TRY(cudaMemcpy(cells4, t->d_cells, sizeof(float4) * t->ncell, cudaMemcpyDeviceToHost));
    if (cdf) CUDA_TRY(cudaMemcpy(cdf, t->d_image + d.off_cdf, sizeof(float) * d.ni * (d.nh + 1), cudaMemcpyDeviceToHost));
    if (dens) CUDA_TRY(cudaMemcpy(dens, t->d_image + d.off_dens, sizeof(float) * d.ni * d.nh, cudaMemcpyDeviceToHost));
    return MERL_OK;
}

// ------------------------------------------------------------------------------------------
// table files either side of the loader: a report on a .binary file, and a
// packed sidecar that holds the device-format table so a scene with many tables starts fast
// ------------------------------------------------------------------------------------------
static int merl_file_report_impl(const char* path, merl_file_report_t* o) {
    if (!path || !o) return fail(MERL_ERR_ARG, "NULL argument");
    memset(o, 0, sizeof *o);
    FILE* f = fopen(path, "rb");
    if (!f) return fail(MERL_ERR_IO, "cannot open '%s'", path);
    if (fread(o->dims, sizeof(int32_t), 3, f) != 3) { fclose(f); return fail(MERL_ERR_IO, "'%s': short header", path); }
    fseek(f, 0, SEEK_END);
    o->file_bytes = (int64_t)ftell(f);
    fseek(f, 12, SEEK_SET);
    const int64_t n = (int64_t)o->dims[0] * o->dims[1] * o->dims[2];
    o->ncell = n;
    o->is_merl_layout = (n == 90 * 90 * 180);
    if (o->dims[0] <= 0 || o->dims[1] <= 0 || o->dims[2] <= 0 || n > ((int64_t)1 << 31)) {
        fclose(f);
        return fail(MERL_ERR_IO, "'%s': implausible header %d x %d x %d", path, o->dims[0], o->dims[1], o->dims[2]);
    }
    o->expected_bytes = 12 + 3 * n * 8;
    if (o->file_bytes != o->expected_bytes) {
        fclose(f);
        return fail(MERL_ERR_IO, "'%s': %lld bytes, a %d x %d x %d table needs %lld", path, (long long)o->file_bytes,
                    o->dims[0], o->dims[1], o->dims[2], (long long)o->expected_bytes);
    }
    std::vector<double> buf((size_t)1 << 16);
    for (int c = 0; c < 3; ++c) {
        double mn = INFINITY, mx = -INFINITY;
        int64_t left = n;
        while (left > 0) {
            const size_t want = (size_t)(left < (int64_t)buf.size() ? left : (int64_t)buf.size());
            if (fread(buf.data(), 8, want, f) != want) { fclose(f); return fail(MERL_ERR_IO, "'%s': read error", path); }
            for (size_t i = 0; i < want; ++i) {
                const double v = buf[i];
                if (!(v == v) || v == INFINITY || v == -INFINITY) { o->nonfinite[c]++; continue; }
                if (v < 0.0) { o->negative[c]++; continue; }       // MERL marks below-horizon cells with negatives
                mn = v < mn ? v : mn;
                mx = v > mx ? v : mx;
            }
            left -= (int64_t)want;
        }
        o->min_valid[c] = mn;
        o->max_valid[c] = mx;
    }
    fclose(f);
    return MERL_OK;
}
extern "C" int merl_file_report(const char* path, merl_file_report_t* o) {
    return guarded("merl_file_report", [&]() -> int { return merl_file_report_impl(path, o); });
}

namespace {
struct PackedHeader {
    char magic[8];               // "MERLB200"
    int32_t version, header_bytes;
    merl_table_params_t params;
    int64_t ncell, cdf_floats;
    uint64_t checksum;           // FNV-1a over the payload
};
constexpr int32_t kPackedVersion = 1;
uint64_t fnv1a(const void* p, size_t n, uint64_t h) {
    const unsigned char* c = static_cast<const unsigned char*>(p);
    for (size_t i = 0; i < n; ++i) { h ^= c[i]; h *= 1099511628211ull; }
    return h;
}
}  // namespace

static int merl_table_save_packed_impl(const merl_table_t* t, const char* path) {
    if (!t || !path) return fail(MERL_ERR_ARG, "NULL argument");
    const TableDev& d = t->dev;
    std::vector<float> cells((size_t)4 * t->ncell), cdf((size_t)d.ni * (d.nh + 1));
    const int rc = merl_table_download(t, cells.data(), cdf.data(), nullptr);
    if (rc) return rc;
    PackedHeader h{};
    memcpy(h.magic, "MERLB200", 8);
    h.version = kPackedVersion;
    h.header_bytes = (int32_t)sizeof h;
    h.params = t->params;
    h.ncell = t->ncell;
    h.cdf_floats = (int64_t)cdf.size();
    h.checksum = fnv1a(cdf.data(), cdf.size() * 4, fnv1a(cells.data(), cells.size() * 4, 1469598103934665603ull));
    FILE* f = fopen(path, "wb");
    if (!f) return fail(MERL_ERR_IO, "cannot create '%s'", path);
    const bool ok = fwrite(&h, sizeof h, 1, f) == 1 && fwrite(cells.data(), 4, cells.size(), f) == cells.size() &&
                    fwrite(cdf.data(), 4, cdf.size(), f) == cdf.size();
    if (fclose(f) != 0 || !ok) return fail(MERL_ERR_IO, "'%s': write error", path);
    return MERL_OK;
}
extern "C" int merl_table_save_packed(const merl_table_t* t, const char* path) {
    return guarded("merl_table_save_packed", [&]() -> int { return merl_table_save_packed_impl(t, path); });
}

static int merl_table_create_from_packed_impl(const char* path, int device, merl_table_t** out) {
    MERL_RANGE("merl_table_create_from_packed");
    if (!path || !out) return fail(MERL_ERR_ARG, "NULL argument");
    FILE* f = fopen(path, "rb");
    if (!f) return fail(MERL_ERR_IO, "cannot open '%s'", path);
    PackedHeader h{};
    if (fread(&h, sizeof h, 1, f) != 1 || memcmp(h.magic, "MERLB200", 8) != 0) {
        fclose(f);
        return fail(MERL_ERR_IO, "'%s' is not a packed libmerl_b200 table", path);
    }
    const merl_table_params_t& p = h.params;
    if (h.version != kPackedVersion || h.header_bytes != (int32_t)sizeof h || p.nh < 2 || p.nh > kMaxNH || p.nd < 2 ||
        p.nd > kMaxND || p.np < 2 || p.np > kMaxNP || p.ni < 1 || p.ni > kMaxNI ||
        h.ncell != (int64_t)p.nh * p.nd * p.np || h.cdf_floats != (int64_t)p.ni * (p.nh + 1)) {
        fclose(f);
        return fail(MERL_ERR_IO, "'%s': unsupported version or inconsistent header", path);
    }
    std::vector<float> cells((size_t)4 * h.ncell), cdf((size_t)h.cdf_floats);
    const bool ok = fread(cells.data(), 4, cells.size(), f) == cells.size() && fread(cdf.data(), 4, cdf.size(), f) == cdf.size();
    fclose(f);
    if (!ok) return fail(MERL_ERR_IO, "'%s' is truncated", path);
    if (fnv1a(cdf.data(), cdf.size() * 4, fnv1a(cells.data(), cells.size() * 4, 1469598103934665603ull)) != h.checksum)
        return fail(MERL_ERR_IO, "'%s': checksum mismatch", path);
    return build_table(nullptr, &p, device, out, cells.data(), cdf.data());
}
extern "C" int merl_table_create_from_packed(const char* path, int device, merl_table_t** out) {
    return guarded("merl_table_create_from_packed", [&]() -> int { return merl_table_create_from_packed_impl(path, device, out); });
}

// ------------------------------------------------------------------------------------------
// C ABI: batched queries on device pointers
// ------------------------------------------------------------------------------------------
extern "C" int merl_bins(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz,
                         const float* wox, const float* woy, const float* woz, const uint8_t* active, int64_t n,
                         int32_t* idx_out, merl_stream_t stream) {
    MERL_RANGE("merl_bins");
    int rc = check_query_args(t, wix, wiy, wiz, wox, woy, woz, n);
    if (rc) return rc;
    if (n > 0 && !idx_out) return fail(MERL_ERR_ARG, "idx_out is NULL");
    DeviceGuard g(t->device);
    const QueryIn in{wix, wiy, wiz, wox, woy, woz, active};
    const QueryOut out{nullptr, nullptr, nullptr, nullptr, idx_out};
    return run_query_mode<MODE_BINS>(t, in, n, out, (cudaStream_t)stream);
}

extern "C" int merl_eval(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz,
                         const float* wox, const float* woy, const float* woz, const uint8_t* active, int64_t n,
                         float* r, float* g, float* b, merl_stream_t stream) {
    MERL_RANGE("merl_eval");
    int rc = check_query_args(t, wix, wiy, wiz, wox, woy, woz, n);
    if (rc) return rc;
    if (n > 0 && (!r || !g || !b)) return fail(MERL_ERR_ARG, "output pointer is NULL");
    DeviceGuard gd(t->device);
    const QueryIn in{wix, wiy, wiz, wox, woy, woz, active};
    const QueryOut out{r, g, b, nullptr, nullptr};
    return run_query_mode<MODE_EVAL>(t, in, n, out, (cudaStream_t)stream);
}

extern "C" int merl_pdf(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz,
                        const float* wox, const float* woy, const float* woz, const uint8_t* active, int64_t n,
                        float* pdf, merl_stream_t stream) {
    MERL_RANGE("merl_pdf");
    int rc = check_query_args(t, wix, wiy, wiz, wox, woy, woz, n);
    if (rc) return rc;
    if (n > 0 && !pdf) return fail(MERL_ERR_ARG, "output pointer is NULL");
    DeviceGuard gd(t->device);
    const QueryIn in{wix, wiy, wiz, wox, woy, woz, active};
    const QueryOut out{nullptr, nullptr, nullptr, pdf, nullptr};
    return run_query_mode<MODE_PDF>(t, in, n, out, (cudaStream_t)stream);
}

extern "C" int merl_eval_pdf(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz,
                             const float* wox, const float* woy, const float* woz, const uint8_t* active,
                             int64_t n, float* r, float* g, float* b, float* pdf, merl_stream_t stream) {
    MERL_RANGE("merl_eval_pdf");
    int rc = check_query_args(t, wix, wiy, wiz, wox, woy, woz, n);
    if (rc) return rc;
    if (n > 0 && (!r || !g || !b || !pdf)) return fail(MERL_ERR_ARG, "output pointer is NULL");
    DeviceGuard gd(t->device);
    const QueryIn in{wix, wiy, wiz, wox, woy, woz, active};
    const QueryOut out{r, g, b, pdf, nullptr};
    return run_query_mode<MODE_EVAL_PDF>(t, in, n, out, (cudaStream_t)stream);
}

template <int SEARCH>
static int run_sample(const merl_table* t, SampleIn in, int64_t n, SampleOut out, cudaStream_t s) {
    const int fl = t->dev.image_full_floats;
    const bool vec = aligned(in.wix, 16) && aligned(in.wiy, 16) && aligned(in.wiz, 16) && aligned(in.u1, 16) &&
                     aligned(in.u2x, 16) && aligned(in.u2y, 16) && aligned(in.active, 4) && aligned(out.wox, 16) &&
                     aligned(out.woy, 16) && aligned(out.woz, 16) && aligned(out.pdf, 16) && aligned(out.wr, 16) &&
                     aligned(out.wg, 16) && aligned(out.wb, 16) && aligned(out.lobe, 16);
    int64_t done = 0;
#ifndef MERL_SAMPLE_VEC
#define MERL_SAMPLE_VEC 4
#endif
    constexpr int SV = MERL_SAMPLE_VEC;
    if (vec && n >= SV) {
        const int rc = launch_op<SampleOp<SV, SEARCH>, MERL_SAMPLE_MINB>(t, SampleOp<SV, SEARCH>{in, out}, fl, n / SV, s);
        if (rc) return rc;
        done = (n / SV) * SV;
    }
    if (done < n) {
        in.wix += done; in.wiy += done; in.wiz += done; in.u1 += done; in.u2x += done; in.u2y += done;
        if (in.active) in.active += done;
        out.wox += done; out.woy += done; out.woz += done; out.pdf += done;
        out.wr += done; out.wg += done; out.wb += done; out.lobe += done;
        return launch_op<SampleOp<1, SEARCH>, 3 * kMinbScale>(t, SampleOp<1, SEARCH>{in, out}, fl, n - done, s);
    }
    return MERL_OK;
}

extern "C" int merl_sample(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz,
                           const float* u1, const float* u2x, const float* u2y, const uint8_t* active, int64_t n,
                           float* wox, float* woy, float* woz, float* pdf, float* wr, float* wg, float* wb,
                           uint32_t* lobe, merl_stream_t stream) {
    MERL_RANGE("merl_sample");
    int rc = check_query_args(t, wix, wiy, wiz, u1, u2x, u2y, n);
    if (rc) return rc;
    if (n > 0 && (!wox || !woy || !woz || !pdf || !wr || !wg || !wb || !lobe))
        return fail(MERL_ERR_ARG, "output pointer is NULL");
    DeviceGuard gd(t->device);
    const SampleIn in{wix, wiy, wiz, u1, u2x, u2y, active};
    const SampleOut out{wox, woy, woz, pdf, wr, wg, wb, lobe};
    return g_search_mode.load() ? run_sample<1>(t, in, n, out, (cudaStream_t)stream)
                                : run_sample<0>(t, in, n, out, (cudaStream_t)stream);
}

struct BackwardIn {
    QueryIn q;
    const float *gr, *gg, *gb;
};

template <int AGG, int GATHER, int CTA = 0>
static int run_backward(const merl_table* t, BackwardIn in, int64_t n, float4* grad, cudaStream_t s, bool mask_pass) {
    const int fl = t->dev.off_dens;
    const QueryIn& q = in.q;
    const bool vec = aligned(q.wix, 16) && aligned(q.wiy, 16) && aligned(q.wiz, 16) && aligned(q.wox, 16) &&
                     aligned(q.woy, 16) && aligned(q.woz, 16) && aligned(q.active, 4) && aligned(in.gr, 16) &&
                     aligned(in.gg, 16) && aligned(in.gb, 16);
    int64_t done = 0;
    // GATHER == 0 never reads the cells: the persisting window goes to the gradient table (same size) instead
    WindowOverride win(!GATHER && pin_gradient_window() ? (const void*)grad : nullptr);
    if (vec && n >= 4) {
        using Op = BackwardOp<4, AGG, GATHER, CTA>;
        const int rc = launch_op<Op, MERL_BWD_MINB>(t, Op{in.q, in.gr, in.gg, in.gb, grad, extra_smem_offset<Op>(fl)}, fl, n / 4, s);
        if (rc) return rc;
        done = (n / 4) * 4;
    }
    if (done < n) {
        using Op = BackwardOp<1, AGG, GATHER, 0>;
        in.q = offset_in(in.q, done);
        in.gr += done; in.gg += done; in.gb += done;
        const int rc = launch_op<Op, 3 * kMinbScale>(t, Op{in.q, in.gr, in.gg, in.gb, grad, 0}, fl, n - done, s);
        if (rc) return rc;
    }
    if (!GATHER && n > 0 && mask_pass) {
        backward_mask_kernel<<<t->num_sms * 8, 256, 0, s>>>(t->d_cells, grad, t->ncell);
        CUDA_TRY(cudaGetLastError());
        g_launches.fetch_add(1);
    }
    return MERL_OK;
}

extern "C" int merl_backward(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz,
                             const float* wox, const float* woy, const float* woz, const float* gr,
                             const float* gg, const float* gb, const uint8_t* active, int64_t n,
                             float* grad_cells4, uint32_t flags, merl_stream_t stream) {
    MERL_RANGE("merl_backward");
    int rc = check_query_args(t, wix, wiy, wiz, wox, woy, woz, n);
    if (rc) return rc;
    if (n > 0 && (!gr || !gg || !gb || !grad_cells4)) return fail(MERL_ERR_ARG, "gradient pointer is NULL");
    if (!aligned(grad_cells4, 16)) return fail(MERL_ERR_ARG, "grad_cells4 must be 16-byte aligned");
    DeviceGuard gd(t->device);
    BackwardIn in{{wix, wiy, wiz, wox, woy, woz, active}, gr, gg, gb};
    float4* grad = reinterpret_cast<float4*>(grad_cells4);
    // big batches: unmasked scatter + one mask pass over the table; small ones: per-query validity gather
    bool gather = n < t->ncell || !t->params.clamp_negative;
    cudaStream_t s = (cudaStream_t)stream;
    // MERL_BWD_DEFER_MASK: every query scatters unmasked whatever the batch size and the caller applies the
    // validity mask once per exchange (merl_peerbuf_allreduce with the table, or merl_grad_apply_mask)
    const bool defer = (flags & MERL_BWD_DEFER_MASK) != 0;
    if (defer && t->params.clamp_negative) gather = false;
    const bool mp = !defer;
    // coherent batches: lane fold + warp fold + CTA-wide shared-memory accumulator, blocked work assignment
    if (flags & MERL_BWD_CTA_AGGREGATE) return gather ? run_backward<1, 1, 1>(t, in, n, grad, s, mp) : run_backward<1, 0, 1>(t, in, n, grad, s, mp);
    if (flags & MERL_BWD_WARP_AGGREGATE) return gather ? run_backward<1, 1>(t, in, n, grad, s, mp) : run_backward<1, 0>(t, in, n, grad, s, mp);
    return gather ? run_backward<0, 1>(t, in, n, grad, s, mp) : run_backward<0, 0>(t, in, n, grad, s, mp);
}

// hit counts out of a gradient table
static __global__ void grad_clear_counts_kernel(float4* __restrict__ grad, const int64_t ncell) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < ncell; i += (int64_t)gridDim.x * blockDim.x)
        grad[i].w = 0.f;
}
extern "C" int merl_grad_clear_counts(const merl_table_t* t, float* grad_cells4, merl_stream_t stream) {
    if (!t || !grad_cells4) return fail(MERL_ERR_ARG, "NULL argument");
    if ((uintptr_t)grad_cells4 & 15) return fail(MERL_ERR_ARG, "grad_cells4 must be 16-byte aligned");
    DeviceGuard g(t->device);
    grad_clear_counts_kernel<<<t->num_sms * 4, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<float4*>(grad_cells4), t->ncell);
    g_launches.fetch_add(1);
    CUDA_TRY(cudaGetLastError());
    return MERL_OK;
}

extern "C" int merl_eval_pdf_aos(const merl_table_t* t, const float* wi4, const float* wo4, int64_t n, float* out4,
                                 merl_stream_t stream) {
    MERL_RANGE("merl_eval_pdf_aos");
    if (!t || n < 0 || (n > 0 && (!wi4 || !wo4 || !out4))) return fail(MERL_ERR_ARG, "bad argument");
    if (!aligned(wi4, 16) || !aligned(wo4, 16) || !aligned(out4, 16))
        return fail(MERL_ERR_ARG, "AoS buffers must be 16-byte aligned");
    DeviceGuard gd(t->device);
    const EvalPdfAosOp op{reinterpret_cast<const float4*>(wi4), reinterpret_cast<const float4*>(wo4),
                          reinterpret_cast<float4*>(out4)};
    return launch_op<EvalPdfAosOp, 4 * kMinbScale>(t, op, t->dev.image_eval_floats, n, (cudaStream_t)stream);
}

extern "C" int merl_sample_aos(const merl_table_t* t, const float* wi4, const float* u4, int64_t n, float* wo4_out,
                               float* weight4_out, merl_stream_t stream) {
    MERL_RANGE("merl_sample_aos");
    if (!t || n < 0 || (n > 0 && (!wi4 || !u4 || !wo4_out || !weight4_out))) return fail(MERL_ERR_ARG, "bad argument");
    if (!aligned(wi4, 16) || !aligned(u4, 16) || !aligned(wo4_out, 16) || !aligned(weight4_out, 16))
        return fail(MERL_ERR_ARG, "AoS buffers must be 16-byte aligned");
    DeviceGuard gd(t->device);
    const SampleAosOp op{reinterpret_cast<const float4*>(wi4), reinterpret_cast<const float4*>(u4),
                         reinterpret_cast<float4*>(wo4_out), reinterpret_cast<float4*>(weight4_out)};
    return launch_op<SampleAosOp, 3 * kMinbScale>(t, op, t->dev.image_full_floats, n, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------
// host-buffer entry points: chunked H2D -> kernel -> D2H over three internal streams, so the
// two copy engines and the SMs overlap (PCIe is full duplex).  Used by scalar-variant
// plugins and by bench.py's end-to-end leg.
// ------------------------------------------------------------------------------------------
struct HostPipe {
    static constexpr int kSlots = 3;
    static constexpr int64_t kChunk = 1 << 21;
    int64_t cap = 0;          // floats per array the slots are sized for (<= kChunk)
    cudaStream_t stream[kSlots]{};
    float* d_in[kSlots]{};    // 6 arrays of cap floats
    float* d_out[kSlots]{};   // 8 arrays of cap floats
};

static void release_pipe(HostPipe* p) {
    if (!p) return;
    for (int s = 0; s < HostPipe::kSlots; ++s) {
        if (p->stream[s]) { cudaStreamSynchronize(p->stream[s]); cudaStreamDestroy(p->stream[s]); }
        cudaFree(p->d_in[s]);
        cudaFree(p->d_out[s]);
    }
    delete p;
}

// Slots hold min(n, kChunk) queries (rounded up to a power of two >= 1024): a scalar-variant host
// that asks for one query at a time pins 3 x 14 x 4 KB per table, not 3 x 14 x 8 MB; a later, larger
// call grows the pipe.  Partial allocation failures release what was made.
static int get_pipe(merl_table* t, int64_t n, HostPipe** out) {
    int64_t want = 1024;
    while (want < n && want < HostPipe::kChunk) want <<= 1;
    HostPipe* p = static_cast<HostPipe*>(t->pipe);
    if (p && p->cap >= want) { *out = p; return MERL_OK; }
    if (p) { release_pipe(p); t->pipe = nullptr; }
    p = new HostPipe();
    p->cap = want;
    for (int s = 0; s < HostPipe::kSlots; ++s) {
        cudaError_t e = cudaStreamCreateWithFlags(&p->stream[s], cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaMalloc(&p->d_in[s], sizeof(float) * 6 * want);
        if (e == cudaSuccess) e = cudaMalloc(&p->d_out[s], sizeof(float) * 8 * want);
        if (e != cudaSuccess) {
            release_pipe(p);
            return fail(e == cudaErrorMemoryAllocation ? MERL_ERR_NOMEM : MERL_ERR_CUDA, "host pipeline buffers: %s",
                        cudaGetErrorString(e));
        }
    }
    t->pipe = p;
    *out = p;
    return MERL_OK;
}

static void free_pipe(merl_table* t) {
    if (!t || !t->pipe) return;
    DeviceGuard g(t->device);
    release_pipe(static_cast<HostPipe*>(t->pipe));
    t->pipe = nullptr;
}

static int merl_eval_pdf_host_impl(const merl_table_t* tc, const float* wix, const float* wiy, const float* wiz,
                                  const float* wox, const float* woy, const float* woz, int64_t n, float* r,
                                  float* g, float* b, float* pdf) {
    MERL_RANGE("merl_eval_pdf_host");
    int rc = check_query_args(tc, wix, wiy, wiz, wox, woy, woz, n);
    if (rc) return rc;
    if (n > 0 && (!r || !g || !b || !pdf)) return fail(MERL_ERR_ARG, "output pointer is NULL");
    merl_table* t = const_cast<merl_table*>(tc);
    DeviceGuard gd(t->device);
    std::lock_guard<std::mutex> lk(t->host_mu);
    HostPipe* p;
    if ((rc = get_pipe(t, n, &p))) return rc;
    const int64_t C = p->cap;
    const float* src[6] = {wix, wiy, wiz, wox, woy, woz};
    float* dst[4] = {r, g, b, pdf};
    int slot = 0;
    for (int64_t o = 0; o < n; o += C, slot = (slot + 1) % HostPipe::kSlots) {
        const int64_t m = (n - o < C) ? n - o : C;
        cudaStream_t s = p->stream[slot];
        float* di = p->d_in[slot];
        float* dq = p->d_out[slot];
        for (int a = 0; a < 6; ++a)
            CUDA_TRY(cudaMemcpyAsync(di + a * C, src[a] + o, sizeof(float) * m, cudaMemcpyHostToDevice, s));
        const QueryIn in{di, di + C, di + 2 * C, di + 3 * C, di + 4 * C, di + 5 * C, nullptr};
        const QueryOut out{dq, dq + C, dq + 2 * C, dq + 3 * C, nullptr};
        if ((rc = run_query_mode<MODE_EVAL_PDF>(t, in, m, out, s))) return rc;
        for (int a = 0; a < 4; ++a)
            CUDA_TRY(cudaMemcpyAsync(dst[a] + o, dq + a * C, sizeof(float) * m, cudaMemcpyDeviceToHost, s));
    }
    for (int s = 0; s < HostPipe::kSlots; ++s) CUDA_TRY(cudaStreamSynchronize(p->stream[s]));
    return MERL_OK;
}
extern "C" int merl_eval_pdf_host(const merl_table_t* tc, const float* wix, const float* wiy, const float* wiz,
                                  const float* wox, const float* woy, const float* woz, int64_t n, float* r,
                                  float* g, float* b, float* pdf) {
    return guarded("merl_eval_pdf_host", [&]() -> int { return merl_eval_pdf_host_impl(tc, wix, wiy, wiz, wox, woy, woz, n, r, g, b, pdf); });
}

static int merl_sample_host_impl(const merl_table_t* tc, const float* wix, const float* wiy, const float* wiz,
                                const float* u1, const float* u2x, const float* u2y, int64_t n, float* wox,
                                float* woy, float* woz, float* pdf, float* wr, float* wg, float* wb, uint32_t* lobe) {
    MERL_RANGE("merl_sample_host");
    int rc = check_query_args(tc, wix, wiy, wiz, u1, u2x, u2y, n);
    if (rc) return rc;
    if (n > 0 && (!wox || !woy || !woz || !pdf || !wr || !wg || !wb || !lobe))
        return fail(MERL_ERR_ARG, "output pointer is NULL");
    merl_table* t = const_cast<merl_table*>(tc);
    DeviceGuard gd(t->device);
    std::lock_guard<std::mutex> lk(t->host_mu);
    HostPipe* p;
    if ((rc = get_pipe(t, n, &p))) return rc;
    const int64_t C = p->cap;
    const float* src[6] = {wix, wiy, wiz, u1, u2x, u2y};
    float* dst[8] = {wox, woy, woz, pdf, wr, wg, wb, reinterpret_cast<float*>(lobe)};
    int slot = 0;
    for (int64_t o = 0; o < n; o += C, slot = (slot + 1) % HostPipe::kSlots) {
        const int64_t m = (n - o < C) ? n - o : C;
        cudaStream_t s = p->stream[slot];
        float* di = p->d_in[slot];
        float* dq = p->d_out[slot];
        for (int a = 0; a < 6; ++a)
            CUDA_TRY(cudaMemcpyAsync(di + a * C, src[a] + o, sizeof(float) * m, cudaMemcpyHostToDevice, s));
        const SampleIn in{di, di + C, di + 2 * C, di + 3 * C, di + 4 * C, di + 5 * C, nullptr};
        const SampleOut out{dq, dq + C, dq + 2 * C, dq + 3 * C, dq + 4 * C, dq + 5 * C, dq + 6 * C,
                            reinterpret_cast<uint32_t*>(dq + 7 * C)};
        rc = g_search_mode.load() ? run_sample<1>(t, in, m, out, s) : run_sample<0>(t, in, m, out, s);
        if (rc) return rc;
        for (int a = 0; a < 8; ++a)
            CUDA_TRY(cudaMemcpyAsync(dst[a] + o, dq + a * C, sizeof(float) * m, cudaMemcpyDeviceToHost, s));
    }
    for (int s = 0; s < HostPipe::kSlots; ++s) CUDA_TRY(cudaStreamSynchronize(p->stream[s]));
    return MERL_OK;
}
extern "C" int merl_sample_host(const merl_table_t* tc, const float* wix, const float* wiy, const float* wiz,
                                const float* u1, const float* u2x, const float* u2y, int64_t n, float* wox,
                                float* woy, float* woz, float* pdf, float* wr, float* wg, float* wb, uint32_t* lobe) {
    return guarded("merl_sample_host", [&]() -> int { return merl_sample_host_impl(tc, wix, wiy, wiz, u1, u2x, u2y, n, wox, woy, woz, pdf, wr, wg, wb, lobe); });
}

// ------------------------------------------------------------------------------------------
// merl_eval_pdf_host_static: the host-buffer eval+pdf without any allocation.
//
// merl_eval_pdf_host above stages the caller's host arrays through device buffers that it
// cudaMallocs on first use (and regrows for a larger call).  Some hosts of this library must not
// trigger that: a renderer whose own pool (Dr.Jit's allocator, a torch caching allocator, a 100-table
// scene) has already claimed the device memory, so a late cudaMalloc fails or evicts; a frame loop
// that cannot take the device-wide synchronisation cudaMalloc/cudaFree imply; a scalar-variant plugin
// called from many worker threads where the first query of each new table would stall the others.
// For them the staging area is a fixed array in the library's own image, below.  It costs
// kStaticStageBytes of device memory on every GPU this library is loaded on, from load until exit,
// whether or not the entry point is ever called.
//
// g_static_stage is a HOST-SIDE STAND-IN in host code: its address there is not a device address.
// static_stage_for_current_device() resolves the real one with cudaGetSymbolAddress, once per device,
// with that device current, and only that pointer is handed to copies and kernels.
// ------------------------------------------------------------------------------------------
static constexpr int64_t kStaticStageQueries = MERL_STATIC_STAGE_MAX_QUERIES;
static constexpr int kStaticStageArrays = 10;   // wi.xyz, wo.xyz in; r, g, b, pdf out
static constexpr size_t kStaticStageBytes = sizeof(float) * kStaticStageArrays * kStaticStageQueries;
static __device__ __align__(256) float g_static_stage[kStaticStageArrays * kStaticStageQueries];
static_assert(kStaticStageBytes <= (4u << 20), "the resident staging array is meant to stay small");

namespace {
constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mu[kStaticMaxDevices];         // one call at a time per device: they share the array
float* g_static_ptr[kStaticMaxDevices] = {};      // device address of g_static_stage, per device
std::mutex g_static_setup_mu;

// The caller has made `device` current.  First call per device: look the symbol up (no allocation;
// the array was placed when the module was loaded onto that device).
int static_stage_for_current_device(int device, float** out) {
    if (device < 0 || device >= kStaticMaxDevices) return fail(MERL_ERR_UNSUPPORTED, "device %d: static staging supports devices 0..%d", device, kStaticMaxDevices - 1);
    std::lock_guard<std::mutex> lk(g_static_setup_mu);
    if (!g_static_ptr[device]) {
        float* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_static_stage);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return fail(MERL_ERR_CUDA, "cudaGetSymbolAddress(static staging array) on device %d: %s", device, cudaGetErrorString(e));
        }
        g_static_ptr[device] = p;
    }
    *out = g_static_ptr[device];
    return MERL_OK;
}
}  // namespace

static int merl_eval_pdf_host_static_impl(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz,
                                         const float* wox, const float* woy, const float* woz, int64_t n, float* r,
                                         float* g, float* b, float* pdf, cudaStream_t s) {
    MERL_RANGE("merl_eval_pdf_host_static");
    int rc = check_query_args(t, wix, wiy, wiz, wox, woy, woz, n);
    if (rc) return rc;
    if (n > kStaticStageQueries)
        return fail(MERL_ERR_ARG, "merl_eval_pdf_host_static: n = %lld exceeds MERL_STATIC_STAGE_MAX_QUERIES = %lld "
                    "(split the batch, or use merl_eval_pdf_host, which allocates)", (long long)n, (long long)kStaticStageQueries);
    if (n > 0 && (!r || !g || !b || !pdf)) return fail(MERL_ERR_ARG, "output pointer is NULL");
    if (n == 0) return MERL_OK;
    DeviceGuard gd(t->device);
    float* stage = nullptr;
    if ((rc = static_stage_for_current_device(t->device, &stage))) return rc;
    // The array is shared by every call on this device and still holds the previous call's queries and
    // results.  Nothing in it has to start at zero: elements [0, n) of the six input arrays are
    // overwritten by the copies below before the kernel reads them, the kernel writes elements [0, n)
    // of the four output arrays for every lane (inactive and below-horizon lanes store zeros), and
    // nothing past n is read or copied back.
    std::lock_guard<std::mutex> lk(g_static_mu[t->device]);
    const int64_t C = kStaticStageQueries;
    const float* src[6] = {wix, wiy, wiz, wox, woy, woz};
    float* dst[4] = {r, g, b, pdf};
    for (int a = 0; a < 6; ++a)
        CUDA_TRY(cudaMemcpyAsync(stage + a * C, src[a], sizeof(float) * n, cudaMemcpyHostToDevice, s));
    float* dq = stage + 6 * C;
    const QueryIn in{stage, stage + C, stage + 2 * C, stage + 3 * C, stage + 4 * C, stage + 5 * C, nullptr};
    const QueryOut out{dq, dq + C, dq + 2 * C, dq + 3 * C, nullptr};
    if ((rc = run_query_mode<MODE_EVAL_PDF>(t, in, n, out, s))) return rc;
    for (int a = 0; a < 4; ++a)
        CUDA_TRY(cudaMemcpyAsync(dst[a], dq + a * C, sizeof(float) * n, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));   // results are in the host buffers on return; the lock covers the GPU work
    return MERL_OK;
}
extern "C" int merl_eval_pdf_host_static(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz,
                                         const float* wox, const float* woy, const float* woz, int64_t n, float* r,
                                         float* g, float* b, float* pdf, merl_stream_t stream) {
    return guarded("merl_eval_pdf_host_static", [&]() -> int {
        return merl_eval_pdf_host_static_impl(t, wix, wiy, wiz, wox, woy, woz, n, r, g, b, pdf, (cudaStream_t)stream);
    });
}

// ------------------------------------------------------------------------------------------
// multi-GPU exchange step: NCCL bound at run time (no link-time dependency)
// ------------------------------------------------------------------------------------------
namespace {
struct NcclId { char bytes[MERL_COMM_ID_BYTES]; };   // ncclUniqueId (passed by value)
struct NcclApi {
    // the four entry points used, with NCCL's stable C signatures (nccl.h 2.x)
    int (*GetUniqueId)(void* id) = nullptr;
    int (*CommInitRank)(void** comm, int nranks, NcclId id, int rank) = nullptr;
    int (*CommDestroy)(void* comm) = nullptr;
    int (*AllReduce)(const void* send, void* recv, size_t count, int dtype, int op, void* comm, cudaStream_t s) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    // optional (present in NCCL >= 2.19 / 2.27): gather of the IPC handles, registered buffers
    int (*AllGather)(const void* send, void* recv, size_t count, int dtype, void* comm, cudaStream_t s) = nullptr;
    int (*MemAlloc)(void** ptr, size_t size) = nullptr;
    int (*MemFree)(void* ptr) = nullptr;
    int (*CommRegister)(void* comm, void* buf, size_t size, void** handle) = nullptr;
    int (*CommDeregister)(void* comm, void* handle) = nullptr;
    int (*CommWindowRegister)(void* comm, void* buf, size_t size, void** win, int flags) = nullptr;
    int (*CommWindowDeregister)(void* comm, void* win) = nullptr;
    int (*GetVersion)(int* v) = nullptr;
    int (*DevCommCreate)(void* comm, const void* requirements, void* dev_comm_out) = nullptr;   // NCCL >= 2.28
    int (*DevCommDestroy)(void* comm, const void* dev_comm) = nullptr;
    bool ok = false;
};
NcclApi* nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);   // the host's copy, if loaded
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) return;
        *(void**)&api.GetUniqueId = dlsym(h, "ncclGetUniqueId");
        *(void**)&api.CommInitRank = dlsym(h, "ncclCommInitRank");
        *(void**)&api.CommDestroy = dlsym(h, "ncclCommDestroy");
        *(void**)&api.AllReduce = dlsym(h, "ncclAllReduce");
        *(void**)&api.GetErrorString = dlsym(h, "ncclGetErrorString");
        *(void**)&api.AllGather = dlsym(h, "ncclAllGather");
        *(void**)&api.MemAlloc = dlsym(h, "ncclMemAlloc");
        *(void**)&api.MemFree = dlsym(h, "ncclMemFree");
        *(void**)&api.CommRegister = dlsym(h, "ncclCommRegister");
        *(void**)&api.CommDeregister = dlsym(h, "ncclCommDeregister");
        *(void**)&api.CommWindowRegister = dlsym(h, "ncclCommWindowRegister");
        *(void**)&api.CommWindowDeregister = dlsym(h, "ncclCommWindowDeregister");
        *(void**)&api.GetVersion = dlsym(h, "ncclGetVersion");
        *(void**)&api.DevCommCreate = dlsym(h, "ncclDevCommCreate");
        *(void**)&api.DevCommDestroy = dlsym(h, "ncclDevCommDestroy");
        api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllReduce;
    });
    return &api;
}
int nccl_fail(const char* what, int rc) {
    NcclApi* a = nccl_api();
    return fail(MERL_ERR_CUDA, "%s: NCCL error %d (%s)", what, rc, a->GetErrorString ? a->GetErrorString(rc) : "?");
}
}  // namespace

struct merl_comm {
    void* comm = nullptr;
    int device = 0, nranks = 1, rank = 0;
#if MERL_HAVE_NCCL_DEVICE
    // device communicator with a multicast mapping of the local (NVLink) team, made on first use by an
    // MERL_XCHG_NVLS buffer (collective: every rank creates its NVLS buffers in the same order)
    ncclDevComm dev_comm{};
    int dev_comm_state = 0;   // 0 not tried, 1 ready, -1 unavailable
#endif
};

extern "C" int merl_comm_unique_id(void* id_out) {
    if (!id_out) return fail(MERL_ERR_ARG, "NULL argument");
    NcclApi* a = nccl_api();
    if (!a->ok) return fail(MERL_ERR_UNSUPPORTED, "libnccl.so.2 not found");
    const int rc = a->GetUniqueId(id_out);
    return rc ? nccl_fail("ncclGetUniqueId", rc) : MERL_OK;
}

static int merl_comm_create_impl(const void* id, int nranks, int rank, int device, merl_comm_t** out) {
    if (!id || !out || nranks < 1 || rank < 0 || rank >= nranks) return fail(MERL_ERR_ARG, "bad argument");
    NcclApi* a = nccl_api();
    if (!a->ok) return fail(MERL_ERR_UNSUPPORTED, "libnccl.so.2 not found");
    int ndev = 0;
    CUDA_TRY(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(MERL_ERR_ARG, "device %d not in [0,%d)", device, ndev);
    DeviceGuard g(device);
    merl_comm* c = new merl_comm();
    c->device = device; c->nranks = nranks; c->rank = rank;
    NcclId idv;
    memcpy(&idv, id, sizeof idv);
    const int rc = a->CommInitRank(&c->comm, nranks, idv, rank);
    if (rc) { delete c; return nccl_fail("ncclCommInitRank", rc); }
    *out = c;
    return MERL_OK;
}
extern "C" int merl_comm_create(const void* id, int nranks, int rank, int device, merl_comm_t** out) {
    return guarded("merl_comm_create", [&]() -> int { return merl_comm_create_impl(id, nranks, rank, device, out); });
}

extern "C" int merl_comm_destroy(merl_comm_t* c) {
    if (!c) return MERL_OK;
    NcclApi* a = nccl_api();
    DeviceGuard g(c->device);
#if MERL_HAVE_NCCL_DEVICE
    if (c->dev_comm_state == 1 && a->DevCommDestroy) a->DevCommDestroy(c->comm, &c->dev_comm);
#endif
    const int rc = (a->ok && c->comm) ? a->CommDestroy(c->comm) : 0;
    delete c;
    return rc ? nccl_fail("ncclCommDestroy", rc) : MERL_OK;
}

extern "C" int merl_allreduce_sum_f32(merl_comm_t* c, float* buf, int64_t count, merl_stream_t stream) {
    MERL_RANGE("merl_allreduce_sum_f32");
    if (!c || count < 0 || (count > 0 && !buf)) return fail(MERL_ERR_ARG, "bad argument");
    if (count == 0) return MERL_OK;
    NcclApi* a = nccl_api();
    DeviceGuard g(c->device);
    const int rc = a->AllReduce(buf, buf, (size_t)count, /*ncclFloat32*/ 7, /*ncclSum*/ 0, c->comm, (cudaStream_t)stream);
    return rc ? nccl_fail("ncclAllReduce", rc) : MERL_OK;
}

// ------------------------------------------------------------------------------------------
// exchange buffers: gradient tables / films that the ranks of a communicator sum through NVLink
// peer memory with ONE kernel (merl_exchange.cuh), or -- transport MERL_XCHG_NCCL -- through
// ncclAllReduce on a buffer from ncclMemAlloc that is registered with the communicator (symmetric
// window when NCCL has them, else ncclCommRegister), which makes NCCL's zero-copy / NVLS path eligible
// ------------------------------------------------------------------------------------------
typedef void (*XchgKernel)(const PeerView, const int64_t, const float4*, uint32_t*, int*);
static XchgKernel xchg_kernel_for(int nranks) {
    switch (nranks) {
        // ranks x unroll = 16 loads of 16 B in flight per thread (a remote load takes ~3 us under load; on 2
        // GPUs unroll 2 left the kernel latency-bound: 67 us for 23.3 MB, profiles/r02_v2_bench_n2.json)
        case 1: return peer_allreduce_kernel<1, 4>;   // nothing to sum; the mask pass still runs
        case 2: return peer_allreduce_kernel<2, 8>;
        case 4: return peer_allreduce_kernel<4, 4>;
        case 8: return peer_allreduce_kernel<8, 2>;
        default: return peer_allreduce_kernel<0, 1>;
    }
}

struct merl_peerbuf {
    merl_comm* comm = nullptr;
    int transport = MERL_XCHG_PEER;
    int64_t count = 0;             // floats (multiple of 4)
    float* d_buf = nullptr;
    uint32_t* d_flags = nullptr;   // PEER: arrive words | epoch words
    int* h_status = nullptr;       // PEER: pinned, mapped; set by a barrier that timed out
    int* d_status = nullptr;
    PeerView view{};
    bool opened[kMaxRanks] = {};
    int ctas_per_sm = 2, num_sms = 0;
    void* nccl_handle = nullptr;   // NCCL: ncclCommRegister handle
    void* nccl_window = nullptr;   // NCCL: symmetric window
    bool nccl_alloc = false;       // d_buf came from ncclMemAlloc
    NvlsView nvls{};               // NVLS: multicast / local addresses inside the symmetric window
    // 32 CTAs x 256 threads x 4 = 32 Ki 16-byte requests in flight per GPU is where the switch path peaks; more
    // CTAs or deeper unrolling only queue up in front of the link (profiles/r02_nvls_sweep_n8.txt: 23.3 MB on 8
    // GPUs in 71.7 us at 32 x 4 against 84.9 us at 148 x 4; NCCL's own kernel uses 3 CTAs x 512 threads).
    // MERL_NVLS_GRID / MERL_NVLS_UNROLL override at buffer creation (tuning).
    int nvls_grid_cap = 32, nvls_unroll = 4;
    const char* registration = "none";
};

namespace {
struct PeerHello {
    cudaIpcMemHandle_t data, flags;
    int32_t num_sms, pid;
    int64_t count;
};

// every rank contributes `bytes` bytes; out = nranks x bytes in rank order (host memory)
int comm_allgather_bytes(merl_comm* c, const void* mine, size_t bytes, void* out) {
    NcclApi* a = nccl_api();
    if (c->nranks == 1) { memcpy(out, mine, bytes); return MERL_OK; }
    if (!a->AllGather) return fail(MERL_ERR_UNSUPPORTED, "this NCCL has no ncclAllGather symbol");
    unsigned char* d = nullptr;
    CUDA_TRY(cudaMalloc(&d, bytes * (size_t)(c->nranks + 1)));
    cudaError_t e = cudaMemcpy(d, mine, bytes, cudaMemcpyHostToDevice);
    int rc = 0;
    if (e == cudaSuccess) rc = a->AllGather(d, d + bytes, bytes, /*ncclUint8*/ 1, c->comm, nullptr);
    if (e == cudaSuccess && !rc) e = cudaStreamSynchronize(nullptr);
    if (e == cudaSuccess && !rc) e = cudaMemcpy(out, d + bytes, bytes * (size_t)c->nranks, cudaMemcpyDeviceToHost);
    cudaFree(d);
    if (rc) return nccl_fail("ncclAllGather", rc);
    if (e != cudaSuccess) return fail(MERL_ERR_CUDA, "gather of the peer handles: %s", cudaGetErrorString(e));
    return MERL_OK;
}
// all ranks have reached this point (and everything they enqueued on the NULL stream before has run)
int comm_barrier(merl_comm* c) {
    int32_t one = 1;
    std::vector<int32_t> all((size_t)c->nranks);
    return comm_allgather_bytes(c, &one, sizeof one, all.data());
}

void peerbuf_release(merl_peerbuf* b) {
    if (!b) return;
    NcclApi* a = nccl_api();
    for (int p = 0; p < kMaxRanks; ++p)
        if (b->opened[p]) {
            cudaIpcCloseMemHandle(b->view.data[p]);
            cudaIpcCloseMemHandle(b->view.flags[p]);
        }
    if (b->nccl_window && a->CommWindowDeregister) a->CommWindowDeregister(b->comm->comm, b->nccl_window);
    if (b->nccl_handle && a->CommDeregister) a->CommDeregister(b->comm->comm, b->nccl_handle);
    if (b->d_buf) { if (b->nccl_alloc) a->MemFree(b->d_buf); else cudaFree(b->d_buf); }
    cudaFree(b->d_flags);
    if (b->h_status) cudaFreeHost(b->h_status);
    cudaGetLastError();
    delete b;
}
}  // namespace

static int merl_peerbuf_create_impl(merl_comm_t* c, int64_t count, int transport, merl_peerbuf_t** out) {
    MERL_RANGE("merl_peerbuf_create");
    if (!c || !out || count < 1) return fail(MERL_ERR_ARG, "bad argument");
    if (transport != MERL_XCHG_PEER && transport != MERL_XCHG_NCCL && transport != MERL_XCHG_NVLS)
        return fail(MERL_ERR_ARG, "unknown transport %d", transport);
    if (c->nranks > kMaxRanks) return fail(MERL_ERR_UNSUPPORTED, "at most %d ranks", kMaxRanks);
    NcclApi* a = nccl_api();
    DeviceGuard g(c->device);
    merl_peerbuf* b = new merl_peerbuf();
    b->comm = c;
    b->transport = transport;
    b->count = (count + 3) & ~(int64_t)3;
    cudaDeviceGetAttribute(&b->num_sms, cudaDevAttrMultiProcessorCount, c->device);
    const size_t bytes = sizeof(float) * (size_t)b->count;
    auto bail = [&](int code) { peerbuf_release(b); return code; };
#define PB_TRY(x)                                                                                   \
    do {                                                                                            \
        cudaError_t e_ = (x);                                                                       \
        if (e_ != cudaSuccess)                                                                      \
            return bail(fail(e_ == cudaErrorMemoryAllocation ? MERL_ERR_NOMEM : MERL_ERR_CUDA, "%s: %s (%s:%d)", #x, \
                             cudaGetErrorString(e_), __FILE__, __LINE__));                          \
    } while (0)
    if (transport == MERL_XCHG_NVLS) {
#if MERL_HAVE_NCCL_DEVICE
        // symmetric window with a multicast mapping; this library's own kernel does the exchange through it
        if (!(a->MemAlloc && a->MemFree && a->CommWindowRegister && a->CommWindowDeregister && a->DevCommCreate))
            return bail(fail(MERL_ERR_UNSUPPORTED, "the loaded NCCL has no symmetric windows / device communicators (needs 2.28+)"));
        int v = 0;
        if (a->GetVersion) a->GetVersion(&v);
        // EXACT match, patch level included: the multicast address below is assembled from NCCL's device-side
        // structures (ncclWindow_vidmem, ncclDevComm) as the headers of the build lay them out, and those are
        // private to a release.  With any other library a wrong pointer would make multimem.st write into
        // arbitrary peer memory without an error, so the host falls back to another transport instead.
        if (v != NCCL_VERSION_CODE)
            return bail(fail(MERL_ERR_UNSUPPORTED, "built against NCCL %d headers, running NCCL %d: the device structures the NVLS "
                             "transport reads are private to a release, an exact match is required", NCCL_VERSION_CODE, v));
        const size_t data_bytes = (bytes + 4095) & ~(size_t)4095, flag_bytes = sizeof(uint32_t) * 2 * kXchgMaxCtas;
        if (a->MemAlloc((void**)&b->d_buf, data_bytes + flag_bytes) != 0 || !b->d_buf) {
            b->d_buf = nullptr;
            return bail(fail(MERL_ERR_NOMEM, "ncclMemAlloc of %zu bytes failed", data_bytes + flag_bytes));
        }
        b->nccl_alloc = true;
        PB_TRY(cudaMemset(b->d_buf, 0, data_bytes + flag_bytes));
        PB_TRY(cudaHostAlloc((void**)&b->h_status, sizeof(int), cudaHostAllocMapped));
        *b->h_status = 0;
        PB_TRY(cudaHostGetDevicePointer((void**)&b->d_status, b->h_status, 0));
        PB_TRY(cudaDeviceSynchronize());
        if (a->CommWindowRegister(c->comm, b->d_buf, data_bytes + flag_bytes, &b->nccl_window, /*NCCL_WIN_COLL_SYMMETRIC*/ 1) != 0 ||
            !b->nccl_window) {
            b->nccl_window = nullptr;
            return bail(fail(MERL_ERR_UNSUPPORTED, "ncclCommWindowRegister(symmetric) failed"));
        }
        if (c->dev_comm_state == 0) {
            ncclDevCommRequirements req{};
            req.lsaMultimem = true;
            c->dev_comm_state = a->DevCommCreate(c->comm, &req, &c->dev_comm) == 0 ? 1 : -1;
        }
        if (c->dev_comm_state != 1 || !c->dev_comm.lsaMultimem.mcBasePtr || c->dev_comm.lsaSize != c->nranks)
            return bail(fail(MERL_ERR_UNSUPPORTED, "no multicast mapping over all %d ranks (NVSwitch multicast unavailable, or "
                             "ranks on several nodes)", c->nranks));
        ncclWindow_vidmem wv;
        PB_TRY(cudaMemcpy(&wv, b->nccl_window, sizeof wv, cudaMemcpyDeviceToHost));
        char* mc = static_cast<char*>(c->dev_comm.lsaMultimem.mcBasePtr) + ((size_t)wv.mcOffset4K << 12);
        b->nvls.mc_data = reinterpret_cast<float4*>(mc);
        b->nvls.mc_flags = reinterpret_cast<uint32_t*>(mc + data_bytes);
        b->nvls.my_flags = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(b->d_buf) + data_bytes);
        b->nvls.epoch_words = b->nvls.my_flags + kXchgMaxCtas;
        b->nvls.nranks = c->nranks;
        b->nvls.rank = c->rank;
        b->registration = "ncclMemAlloc + symmetric window, own multimem kernel (NVLS)";
        {
            int occ = 0;
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, nvls_allreduce_kernel<4>, kXchgThreads, 0) != cudaSuccess || occ < 1) occ = 1;
            cudaGetLastError();
            int want = 1;   // CTAs per SM: the switch, not the SMs, does the adds
            if (const char* e = getenv("MERL_NVLS_CTAS_PER_SM")) want = std::max(1, atoi(e));
            b->ctas_per_sm = std::min(want, occ);
            if (const char* e = getenv("MERL_NVLS_GRID")) b->nvls_grid_cap = std::max(1, atoi(e));
            if (const char* e = getenv("MERL_NVLS_UNROLL")) b->nvls_unroll = atoi(e);
        }
        {
            const int rc = comm_barrier(c);   // every rank has zeroed its flags before anyone arrives
            if (rc) return bail(rc);
        }
        *out = b;
        return MERL_OK;
#else
        return bail(fail(MERL_ERR_UNSUPPORTED, "built without NCCL's device headers (nccl_device/): no MERL_XCHG_NVLS"));
#endif
    }
    if (transport == MERL_XCHG_NCCL) {
        // registered buffer: NCCL can then reduce in place out of user memory (no staging copies) and,
        // on NVSwitch systems, use the in-switch reduction with fewer SMs
        if (a->MemAlloc && a->MemFree && a->MemAlloc((void**)&b->d_buf, bytes) == 0 && b->d_buf) {
            b->nccl_alloc = true;
            if (a->CommWindowRegister && a->CommWindowDeregister &&
                a->CommWindowRegister(c->comm, b->d_buf, bytes, &b->nccl_window, /*NCCL_WIN_COLL_SYMMETRIC*/ 1) == 0 &&
                b->nccl_window)
                b->registration = "ncclMemAlloc + ncclCommWindowRegister(symmetric)";
            else if (a->CommRegister && a->CommDeregister && a->CommRegister(c->comm, b->d_buf, bytes, &b->nccl_handle) == 0)
                b->registration = "ncclMemAlloc + ncclCommRegister";
            else
                b->registration = "ncclMemAlloc, unregistered";
            b->nccl_window = b->nccl_window ? b->nccl_window : nullptr;
        } else {
            b->d_buf = nullptr;
            PB_TRY(cudaMalloc(&b->d_buf, bytes));
            b->registration = "cudaMalloc, unregistered (this NCCL has no ncclMemAlloc)";
        }
        cudaGetLastError();
        PB_TRY(cudaMemset(b->d_buf, 0, bytes));
        PB_TRY(cudaDeviceSynchronize());
        *out = b;
        return MERL_OK;
    }
    // peer transport: legacy CUDA IPC needs cudaMalloc memory (not a pool, not a VMM mapping)
    PB_TRY(cudaMalloc(&b->d_buf, bytes));
    PB_TRY(cudaMalloc(&b->d_flags, sizeof(uint32_t) * kXchgFlagWords));
    PB_TRY(cudaMemset(b->d_buf, 0, bytes));
    PB_TRY(cudaMemset(b->d_flags, 0, sizeof(uint32_t) * kXchgFlagWords));
    PB_TRY(cudaHostAlloc((void**)&b->h_status, sizeof(int), cudaHostAllocMapped));
    *b->h_status = 0;
    PB_TRY(cudaHostGetDevicePointer((void**)&b->d_status, b->h_status, 0));
    PB_TRY(cudaDeviceSynchronize());
    b->view.nranks = c->nranks;
    b->view.rank = c->rank;
    b->view.data[c->rank] = reinterpret_cast<float4*>(b->d_buf);
    b->view.flags[c->rank] = b->d_flags;
    b->registration = "CUDA IPC peer mapping";
    {   // all CTAs of an exchange must be resident at once on every rank (they spin on each other)
        int occ = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, xchg_kernel_for(c->nranks), kXchgThreads, 0) != cudaSuccess || occ < 1) occ = 1;
        cudaGetLastError();
        b->ctas_per_sm = std::min(2, occ);
    }
    if (c->nranks > 1) {
        PeerHello me{};
        PB_TRY(cudaIpcGetMemHandle(&me.data, b->d_buf));
        PB_TRY(cudaIpcGetMemHandle(&me.flags, b->d_flags));
        me.num_sms = b->num_sms;
        me.pid = (int32_t)getpid();
        me.count = b->count;
        std::vector<PeerHello> all((size_t)c->nranks);
        int rc = comm_allgather_bytes(c, &me, sizeof me, all.data());
        if (rc) return bail(rc);
        for (int p = 0; p < c->nranks; ++p) {
            if (all[p].count != b->count) return bail(fail(MERL_ERR_ARG, "rank %d asked for %lld floats, this rank for %lld", p,
                                                          (long long)all[p].count, (long long)b->count));
            b->num_sms = std::min(b->num_sms, (int)all[p].num_sms);
            if (p == c->rank) continue;
            if (all[p].pid == me.pid)
                return bail(fail(MERL_ERR_UNSUPPORTED, "peer exchange needs one process per GPU (rank %d shares this process)", p));
            void *pd = nullptr, *pf = nullptr;
            cudaError_t e = cudaIpcOpenMemHandle(&pd, all[p].data, cudaIpcMemLazyEnablePeerAccess);
            if (e == cudaSuccess) e = cudaIpcOpenMemHandle(&pf, all[p].flags, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) {
                if (pd) cudaIpcCloseMemHandle(pd);
                return bail(fail(MERL_ERR_UNSUPPORTED, "cannot map rank %d's buffer (no NVLink / PCIe peer access?): %s", p,
                                 cudaGetErrorString(e)));
            }
            b->view.data[p] = reinterpret_cast<float4*>(pd);
            b->view.flags[p] = reinterpret_cast<uint32_t*>(pf);
            b->opened[p] = true;
        }
        if ((rc = comm_barrier(c))) return bail(rc);   // everyone has mapped everyone before the first exchange
    }
#undef PB_TRY
    *out = b;
    return MERL_OK;
}
extern "C" int merl_peerbuf_create(merl_comm_t* c, int64_t count, int transport, merl_peerbuf_t** out) {
    return guarded("merl_peerbuf_create", [&]() -> int { return merl_peerbuf_create_impl(c, count, transport, out); });
}

extern "C" int merl_peerbuf_destroy(merl_peerbuf_t* b) {
    if (!b) return MERL_OK;
    DeviceGuard g(b->comm->device);
    cudaDeviceSynchronize();
    // unmap the peers, then meet, then free: no rank frees memory another rank still has mapped
    for (int p = 0; p < kMaxRanks; ++p)
        if (b->opened[p]) {
            cudaIpcCloseMemHandle(b->view.data[p]);
            cudaIpcCloseMemHandle(b->view.flags[p]);
            b->opened[p] = false;
        }
    int rc = MERL_OK;
    if (b->transport != MERL_XCHG_NCCL && b->comm->nranks > 1) rc = comm_barrier(b->comm);
    peerbuf_release(b);
    return rc;
}

extern "C" int merl_peerbuf_info(const merl_peerbuf_t* b, merl_peerbuf_info_t* o) {
    if (!b || !o) return fail(MERL_ERR_ARG, "NULL argument");
    memset(o, 0, sizeof *o);
    o->data = b->d_buf;
    o->count = b->count;
    o->transport = b->transport;
    o->nranks = b->comm->nranks;
    o->rank = b->comm->rank;
    o->status = b->h_status ? *reinterpret_cast<volatile int*>(b->h_status) : 0;
    snprintf(o->registration, sizeof o->registration, "%s", b->registration);
    return MERL_OK;
}

// grid of the exchange kernel: the same on every rank (min SM count over the ranks), all CTAs co-resident
static int xchg_grid(const merl_peerbuf* b, int64_t n4) {
    const int64_t per = (n4 + b->comm->nranks - 1) / b->comm->nranks;
    int64_t want = (per + kXchgThreads - 1) / kXchgThreads;
    const int64_t cap = std::min<int64_t>((int64_t)b->num_sms * b->ctas_per_sm, kXchgMaxCtas);
    if (want > cap) want = cap;
    return (int)(want < 1 ? 1 : want);
}

extern "C" int merl_peerbuf_allreduce(merl_peerbuf_t* b, int64_t count, const merl_table_t* mask_table, merl_stream_t stream) {
    MERL_RANGE("merl_peerbuf_allreduce");
    if (!b || count < 0 || count > b->count || (count & 3)) return fail(MERL_ERR_ARG, "count must be a multiple of 4 in [0, capacity]");
    if (mask_table && (mask_table->device != b->comm->device || 4 * mask_table->ncell != count))
        return fail(MERL_ERR_ARG, "mask table: other device, or count != 4 * ncell");
    if (count == 0) return MERL_OK;
    DeviceGuard g(b->comm->device);
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t n4 = count / 4;
    if (b->transport == MERL_XCHG_NCCL) {
        if (b->comm->nranks > 1) {
            NcclApi* a = nccl_api();
            const int rc = a->AllReduce(b->d_buf, b->d_buf, (size_t)count, /*ncclFloat32*/ 7, /*ncclSum*/ 0, b->comm->comm, s);
            if (rc) return nccl_fail("ncclAllReduce", rc);
        }
        if (mask_table) {
            backward_mask_kernel<<<mask_table->num_sms * 8, 256, 0, s>>>(mask_table->d_cells, reinterpret_cast<float4*>(b->d_buf), n4);
            CUDA_TRY(cudaGetLastError());
            g_launches.fetch_add(1);
        }
        return MERL_OK;
    }
    if (*reinterpret_cast<volatile int*>(b->h_status))
        return fail(MERL_ERR_CUDA, "an earlier peer exchange timed out waiting for another rank; the buffer is unusable");
    const float4* mask = mask_table ? mask_table->d_cells : nullptr;
    if (b->transport == MERL_XCHG_NVLS) {
        if (b->comm->nranks == 1) {
            if (mask) { backward_mask_kernel<<<mask_table->num_sms * 8, 256, 0, s>>>(mask, reinterpret_cast<float4*>(b->d_buf), n4); g_launches.fetch_add(1); }
        } else {
            int grid = xchg_grid(b, n4);
            if (b->nvls_grid_cap > 0) grid = std::min(grid, b->nvls_grid_cap);
            if (b->nvls_unroll == 8) nvls_allreduce_kernel<8><<<grid, kXchgThreads, 0, s>>>(b->nvls, n4, mask, b->d_status);
            else if (b->nvls_unroll == 2) nvls_allreduce_kernel<2><<<grid, kXchgThreads, 0, s>>>(b->nvls, n4, mask, b->d_status);
            else nvls_allreduce_kernel<4><<<grid, kXchgThreads, 0, s>>>(b->nvls, n4, mask, b->d_status);
            g_launches.fetch_add(1);
        }
        CUDA_TRY(cudaGetLastError());
        return MERL_OK;
    }
    uint32_t* epoch = b->d_flags + kXchgMaxCtas * kMaxRanks;
    const int grid = xchg_grid(b, n4);
    xchg_kernel_for(b->comm->nranks)<<<grid, kXchgThreads, 0, s>>>(b->view, n4, mask, epoch, b->d_status);
    CUDA_TRY(cudaGetLastError());
    g_launches.fetch_add(1);
    return MERL_OK;
}

extern "C" int merl_grad_apply_mask(const merl_table_t* t, float* grad_cells4, merl_stream_t stream) {
    if (!t || !grad_cells4) return fail(MERL_ERR_ARG, "NULL argument");
    if ((uintptr_t)grad_cells4 & 15) return fail(MERL_ERR_ARG, "grad_cells4 must be 16-byte aligned");
    DeviceGuard g(t->device);
    backward_mask_kernel<<<t->num_sms * 8, 256, 0, (cudaStream_t)stream>>>(t->d_cells, reinterpret_cast<float4*>(grad_cells4), t->ncell);
    CUDA_TRY(cudaGetLastError());
    g_launches.fetch_add(1);
    return MERL_OK;
}

// ------------------------------------------------------------------------------------------
// wavefront path tracer harness (merl_render.cuh): host driver
// ------------------------------------------------------------------------------------------
static void normalize3(float v[3]) {
    const float l = sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    if (l > 0.f) { v[0] /= l; v[1] /= l; v[2] /= l; }
}
static void cross3(const float a[3], const float b[3], float o[3]) {
    o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}

// Side stream + cached pass graph of merl_render, one per device.  Instantiating a graph costs far
// more than launching it, so the executable graph of the last call is kept and re-targeted with
// cudaGraphExecUpdate (same topology, new pointers / parameters); only a topology change (another
// max_depth) instantiates again.  A replaced executable must outlive its last launch; instead of
// synchronising, it is destroyed by a later call once its event has completed.
struct RenderSide {
    std::mutex mu;
    cudaStream_t stream = nullptr;
    cudaEvent_t fork = nullptr, join = nullptr;
    cudaGraphExec_t exec = nullptr;
    struct Retired { cudaGraphExec_t exec; cudaEvent_t done; };
    std::vector<Retired> retired;
    void reap() {
        size_t keep = 0;
        for (size_t k = 0; k < retired.size(); ++k) {
            if (cudaEventQuery(retired[k].done) == cudaSuccess) {
                cudaGraphExecDestroy(retired[k].exec);
                cudaEventDestroy(retired[k].done);
            } else {
                retired[keep++] = retired[k];
            }
        }
        retired.resize(keep);
        cudaGetLastError();   // cudaErrorNotReady from the query is not an error
    }
    // make `exec` an executable of `graph`: update in place when the topology allows it
    cudaError_t adopt(cudaGraph_t graph) {
        if (exec) {
            cudaGraphExecUpdateResultInfo info;
            if (cudaGraphExecUpdate(exec, graph, &info) == cudaSuccess) return cudaSuccess;
            cudaGetLastError();
            cudaEvent_t done = nullptr;
            if (cudaEventCreateWithFlags(&done, cudaEventDisableTiming) == cudaSuccess &&
                cudaEventRecord(done, stream) == cudaSuccess)
                retired.push_back({exec, done});
            exec = nullptr;   // (if the event could not be made the old executable is leaked, not freed early)
        }
        return cudaGraphInstantiate(&exec, graph, 0);
    }
};
static std::atomic<int> g_render_graphs{1};
// merl_render's workspace comes from a private stream-ordered pool that keeps its memory between
// calls (release threshold = max): the default pool hands freed memory back to the OS at the next
// synchronisation, which made every call pay milliseconds for a fresh 0.5-4 GB mapping.
static cudaMemPool_t render_pool(int device) {
    static std::mutex mu;
    static cudaMemPool_t pools[64] = {};
    if (device < 0 || device >= 64) return nullptr;
    std::lock_guard<std::mutex> lk(mu);
    if (!pools[device]) {
        cudaMemPoolProps props{};
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = device;
        cudaMemPool_t pool = nullptr;
        if (cudaMemPoolCreate(&pool, &props) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        uint64_t keep = UINT64_MAX;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        pools[device] = pool;
    }
    return pools[device];
}
static RenderSide* render_side(int device) {
    static std::mutex mu;
    static RenderSide* sides[64] = {};
    if (device < 0 || device >= 64) { fail(MERL_ERR_ARG, "device %d out of range", device); return nullptr; }
    std::lock_guard<std::mutex> lk(mu);
    if (!sides[device]) {
        RenderSide* r = new RenderSide();
        if (cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking) != cudaSuccess ||
            cudaEventCreateWithFlags(&r->fork, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&r->join, cudaEventDisableTiming) != cudaSuccess) {
            fail(MERL_ERR_CUDA, "cannot create the render side stream: %s", cudaGetErrorString(cudaGetLastError()));
            delete r;
            return nullptr;
        }
        sides[device] = r;
    }
    return sides[device];
}
extern "C" int merl_set_render_graphs(int on) {
    g_render_graphs.store(on ? 1 : 0);
    return MERL_OK;
}

static int merl_render_impl(const merl_table_t* const* tables, int ntables, const merl_sphere_t* spheres, int nspheres,
                           const merl_render_params_t* p, float* film4, merl_render_stats_t* stats_out,
                           merl_stream_t stream_) {
    MERL_RANGE("merl_render");
    if (!tables || !spheres || !p || !film4) return fail(MERL_ERR_ARG, "NULL argument");
    if (ntables < 1 || ntables > 1024) return fail(MERL_ERR_ARG, "ntables must be in [1, 1024]");
    if (nspheres < 1 || nspheres > kMaxSpheres) return fail(MERL_ERR_ARG, "nspheres must be in [1, %d]", kMaxSpheres);
    if (p->width < 1 || p->height < 1 || p->max_depth < 1 || p->spp_total < 1 || p->spp_begin < 0 ||
        p->spp_end < p->spp_begin || p->spp_end > p->spp_total)
        return fail(MERL_ERR_ARG, "bad render parameters");
    if (p->spp_total > (1 << 22)) return fail(MERL_ERR_ARG, "spp_total must not exceed 2^22 (the sample index shares a queue word)");
    if (!aligned(film4, 16)) return fail(MERL_ERR_ARG, "film4 must be 16-byte aligned");
    // by-table sharding: this rank evaluates tables m with m % own_world == own_rank only
    const int own_world = p->table_shard_world > 1 ? p->table_shard_world : 1;
    const int own_rank = own_world > 1 ? p->table_shard_rank : 0;
    if (own_world > 1 && (own_rank < 0 || own_rank >= own_world || !p->stage_exchange))
        return fail(MERL_ERR_ARG, "by-table sharding needs table_shard_rank in [0, world) and a stage_exchange buffer");
    const merl_table* t0 = nullptr;
    for (int k = 0; k < ntables; ++k) {
        const bool mine = k % own_world == own_rank;
        if (!tables[k]) {
            if (mine) return fail(MERL_ERR_ARG, "table %d is NULL", k);
            continue;
        }
        if (!t0) t0 = tables[k];
        if (tables[k]->device != t0->device) return fail(MERL_ERR_ARG, "all tables must live on one device");
    }
    if (!t0) return fail(MERL_ERR_ARG, "this rank owns no table");
    for (int k = 0; k < nspheres; ++k)
        if (spheres[k].material < 0 || spheres[k].material >= ntables || !(spheres[k].radius > 0.f))
            return fail(MERL_ERR_ARG, "sphere %d: bad material or radius", k);
    DeviceGuard guard(t0->device);
    cudaStream_t s = (cudaStream_t)stream_;

    RenderDev R{};
    R.width = p->width; R.height = p->height; R.spp_total = p->spp_total; R.spp_begin = p->spp_begin;
    R.spp_end = p->spp_end; R.max_depth = p->max_depth; R.nspheres = nspheres; R.ntables = ntables;
    R.k0 = (uint32_t)(p->seed & 0xffffffffu); R.k1 = (uint32_t)(p->seed >> 32);
    for (int c = 0; c < 3; ++c) R.env[c] = p->env[c];
    R.inv_spp = 1.0f / (float)p->spp_total;
    {
        float f[3] = {p->cam_target[0] - p->cam_pos[0], p->cam_target[1] - p->cam_pos[1], p->cam_target[2] - p->cam_pos[2]};
        normalize3(f);
        float up[3] = {p->cam_up[0], p->cam_up[1], p->cam_up[2]}, r[3], u[3];
        cross3(f, up, r); normalize3(r); cross3(r, f, u);
        const float th = tanf(0.5f * p->fov_x_deg * 3.14159265358979f / 180.f), asp = (float)p->height / (float)p->width;
        R.cam = Camera{p->cam_pos[0], p->cam_pos[1], p->cam_pos[2], f[0], f[1], f[2], r[0] * th, r[1] * th, r[2] * th,
                       u[0] * th * asp, u[1] * th * asp, u[2] * th * asp};
    }
    const int64_t npix = (int64_t)p->width * p->height;
    const int64_t total = npix * (p->spp_end - p->spp_begin);
    if (total == 0) { if (stats_out) memset(stats_out, 0, sizeof *stats_out); return MERL_OK; }
    // Paths per pass.  Default: as many as fit comfortably, up to 64 Mi (264 B of workspace per path = 17.7 GB of
    // a B200's 180 GB, kept in the pool): the late bounces of a pass hold a few percent of the paths but cost
    // six launches at their latency floor each, so fewer, fatter passes are faster -- config 5 (268 M paths):
    // 4873 / 5515 / 5850 / 6019 spp Mpix/s at 8 / 16 / 32 / 64 Mi paths per pass (profiles/r02_render_wavefront.txt).
    // The default stays below a quarter of the device memory that is free right now.
    // (decided once per device and process: cudaMemGetInfo costs ~50 us, a tenth of a small frame)
    int64_t cap = p->wavefront;
    if (cap <= 0) {
        static std::atomic<int64_t> decided[64];
        const int di = t0->device & 63;
        cap = decided[di].load();
        if (cap == 0) {
            cap = (int64_t)1 << 26;
            size_t free_b = 0, total_b = 0;
            if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) { cudaGetLastError(); free_b = (size_t)8 << 30; }
            while (cap > (1 << 20) && (size_t)cap * 264 > free_b / 4) cap >>= 1;
            decided[di].store(cap);
        }
    }
    if (cap > total) cap = total;
    cap = (cap + 255) & ~(int64_t)255;
    if (own_world > 1) {
        merl_peerbuf* xb = p->stage_exchange;
        if (xb->comm->device != t0->device || xb->comm->nranks != own_world || xb->comm->rank != own_rank)
            return fail(MERL_ERR_ARG, "stage_exchange belongs to another device / communicator shape");
        if (xb->count < 12 * cap)
            return fail(MERL_ERR_ARG, "stage_exchange holds %lld floats, a pass of %lld paths needs %lld", (long long)xb->count,
                        (long long)cap, (long long)(12 * cap));
    }

    // one stream-ordered workspace: path state, stage buffers, sort scratch, scene
    const size_t nf4 = 6, ni = 4;   // path state: six float4 and four int arrays
    char* ws = nullptr;
    for (;;) {
        size_t bytes = cap * (nf4 * 16 + ni * 4) + cap * 16 * 9 + 512;
        bytes += (size_t)nspheres * (16 + 4) + 256 + (size_t)(4 * ntables + 4) * 4 + 256 + sizeof(TableDev) * ntables + 256 + 64;
        bytes += (size_t)cap * 4 * 2 + 512 + 256 + (size_t)kMaxClusters * 32 + 256 + 256;   // live lists, their counters, clusters, pass info
        cudaMemPool_t pool = render_pool(t0->device);
        const cudaError_t e = pool ? cudaMallocFromPoolAsync((void**)&ws, bytes, pool, s) : cudaMallocAsync(&ws, bytes, s);
        if (e == cudaSuccess) break;
        cudaGetLastError();
        // the default pass size was chosen from the memory free at the first call: if the device has filled up
        // since, halve it (an explicit p->wavefront is the caller's decision and fails loudly instead)
        if (e == cudaErrorMemoryAllocation && p->wavefront <= 0 && own_world == 1 && cap > (1 << 20)) {
            cap = ((cap / 2) + 255) & ~(int64_t)255;
            continue;
        }
        return fail(e == cudaErrorMemoryAllocation ? MERL_ERR_NOMEM : MERL_ERR_CUDA, "render workspace of %zu bytes: %s", bytes,
                    cudaGetErrorString(e));
    }
    char* cur = ws;
    auto take = [&](size_t b) { char* r = cur; cur += (b + 255) & ~(size_t)255; return r; };
    float4* f4 = reinterpret_cast<float4*>(take(cap * 16 * 9));
    StageBuf A{f4, f4 + cap, f4 + 2 * cap, f4 + 3 * cap, f4 + 4 * cap, f4 + 5 * cap};
    StageBuf B{f4 + 6 * cap, f4 + 7 * cap, f4 + 8 * cap, A.fe4, A.wos4, A.w4};
    float4* ps = reinterpret_cast<float4*>(take(cap * nf4 * 16));
    int* in = reinterpret_cast<int*>(take(cap * ni * 4));
    PathState P{ps, ps + cap, ps + 2 * cap, ps + 3 * cap, ps + 4 * cap, ps + 5 * cap, in, in + cap, in + 2 * cap, in + 3 * cap};
    float4* d_sph = reinterpret_cast<float4*>(take((size_t)nspheres * 16));
    int* d_smat = reinterpret_cast<int*>(take((size_t)nspheres * 4));
    int* d_counts = reinterpret_cast<int*>(take((size_t)(4 * ntables + 4) * 4));
    int *d_seg = d_counts + ntables, *d_cursor = d_seg + ntables + 1, *d_items = d_cursor + ntables;
    TableDev* d_tabs = reinterpret_cast<TableDev*>(take(sizeof(TableDev) * ntables));
    unsigned long long* d_nq = reinterpret_cast<unsigned long long*>(take(8));
    int* d_live[2] = {reinterpret_cast<int*>(take((size_t)cap * 4)), reinterpret_cast<int*>(take((size_t)cap * 4))};
    int* d_nlive = reinterpret_cast<int*>(take(2 * 4));
    ClusterBox* d_clus = reinterpret_cast<ClusterBox*>(take((size_t)kMaxClusters * sizeof(ClusterBox)));
    PassInfo* d_pass = reinterpret_cast<PassInfo*>(take(sizeof(PassInfo)));

    // Spheres along a Morton curve, kClusterSize consecutive ones under one bounding sphere
    // (merl_render.cuh: intersect_spheres).  The device only ever uses the sorted order.
    std::vector<float4> hs(nspheres);
    std::vector<ClusterBox> hc((nspheres + kClusterSize - 1) / kClusterSize);
    std::vector<int> hm(nspheres);
    {
        float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f};
        for (int k = 0; k < nspheres; ++k) {
            const float c[3] = {spheres[k].cx, spheres[k].cy, spheres[k].cz};
            for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], c[a]); hi[a] = std::max(hi[a], c[a]); }
        }
        auto spread = [](uint32_t v) {   // 10 bits -> every third bit
            v = (v | (v << 16)) & 0x030000FFu; v = (v | (v << 8)) & 0x0300F00Fu;
            v = (v | (v << 4)) & 0x030C30C3u; v = (v | (v << 2)) & 0x09249249u;
            return v;
        };
        std::vector<std::pair<uint32_t, int>> order(nspheres);
        for (int k = 0; k < nspheres; ++k) {
            const float c[3] = {spheres[k].cx, spheres[k].cy, spheres[k].cz};
            uint32_t code = 0;
            for (int a = 0; a < 3; ++a) {
                const float ext = hi[a] - lo[a];
                const uint32_t q = ext > 0.f ? (uint32_t)std::min(1023.f, (c[a] - lo[a]) / ext * 1023.f) : 0u;
                code |= spread(q) << a;
            }
            order[k] = {code, k};
        }
        std::sort(order.begin(), order.end());
        for (int k = 0; k < nspheres; ++k) {
            const merl_sphere_t& sp = spheres[order[k].second];
            hs[k] = make_float4(sp.cx, sp.cy, sp.cz, sp.radius);
            hm[k] = sp.material;
        }
        for (size_t g = 0; g < hc.size(); ++g) {
            const int s0 = (int)g * kClusterSize, s1 = std::min(s0 + kClusterSize, nspheres);
            float lo3[3] = {3e38f, 3e38f, 3e38f}, hi3[3] = {-3e38f, -3e38f, -3e38f};
            for (int k = s0; k < s1; ++k) {
                const float c[3] = {hs[k].x, hs[k].y, hs[k].z};
                const float pad = hs[k].w * (1.f + 1e-4f) + 1e-5f;   // conservative
                for (int a = 0; a < 3; ++a) { lo3[a] = std::min(lo3[a], c[a] - pad); hi3[a] = std::max(hi3[a], c[a] + pad); }
            }
            hc[g].lo = make_float4(lo3[0], lo3[1], lo3[2], 0.f);
            hc[g].hi = make_float4(hi3[0], hi3[1], hi3[2], 0.f);
        }
    }
    std::vector<TableDev> ht(ntables);
    int img_floats_eval = 0, img_floats_full = 0;
    for (int k = 0; k < ntables; ++k) {
        if (!tables[k]) { ht[k] = TableDev{}; continue; }   // another rank's table (by-table sharding)
        ht[k] = tables[k]->dev;
        img_floats_eval = std::max(img_floats_eval, ht[k].image_eval_floats);
        img_floats_full = std::max(img_floats_full, ht[k].image_full_floats);
    }
    const int img_e = (int)(((size_t)img_floats_eval * 4 + 127) & ~(size_t)127);
    const int img_f = (int)(((size_t)img_floats_full * 4 + 127) & ~(size_t)127);
    const int grid_e = t0->num_sms * blocks_per_sm(stream_multi_kernel<EvalPdfAosOp, 4 * kMinbScale>, (size_t)img_e + kQueueBytes);
    const int grid_f = t0->num_sms * blocks_per_sm(stream_multi_kernel<SampleAosOp, 3 * kMinbScale>, (size_t)img_f + kQueueBytes);
    int rc = MERL_OK;
    int64_t launches = 0;
    RenderSide* forked = nullptr;   // set once work may have been enqueued on the side stream
#define RT_TRY(x)                                                                                  \
    do {                                                                                           \
        cudaError_t e_ = (x);                                                                      \
        if (e_ != cudaSuccess) {                                                                   \
            rc = fail(MERL_ERR_CUDA, "%s: %s (%s:%d)", #x, cudaGetErrorString(e_), __FILE__, __LINE__); \
            goto done;                                                                             \
        }                                                                                          \
    } while (0)
    {
    RT_TRY(cudaMemcpyAsync(d_sph, hs.data(), sizeof(float4) * nspheres, cudaMemcpyHostToDevice, s));
    RT_TRY(cudaMemcpyAsync(d_smat, hm.data(), sizeof(int) * nspheres, cudaMemcpyHostToDevice, s));
    RT_TRY(cudaMemcpyAsync(d_clus, hc.data(), sizeof(ClusterBox) * hc.size(), cudaMemcpyHostToDevice, s));
    RT_TRY(cudaMemcpyAsync(d_tabs, ht.data(), sizeof(TableDev) * ntables, cudaMemcpyHostToDevice, s));
    RT_TRY(cudaMemsetAsync(d_counts, 0, sizeof(int) * ntables, s));
    RT_TRY(cudaMemsetAsync(d_nq, 0, 8, s));
    RT_TRY(cudaMemsetAsync(d_nlive, 0, 8, s));
    float4* film = reinterpret_cast<float4*>(film4);
    const PassInfo pass0{0, (int)std::min<int64_t>(cap, total), 0};
    RT_TRY(cudaMemcpyAsync(d_pass, &pass0, sizeof pass0, cudaMemcpyHostToDevice, s));
    const int64_t npasses = (total + cap - 1) / cap;
    const int nb = (int)((cap + 255) / 256);
    // One wavefront pass = 2 + 6 max_depth launches, most of them over a handful of live paths at
    // the later bounces: launch-latency bound when enqueued one by one.  The pass is recorded once
    // as a CUDA graph (it reads its sample range from d_pass, advanced by its last node) and
    // replayed npasses times.  Capture needs a non-legacy stream, so the work runs on a side
    // stream that is forked from and joined back into the caller's stream with events: the call
    // stays asynchronous and ordered on `stream`.
    bool xchg_failed = false;
    const int64_t launches_before = g_launches.load();
    auto enqueue_pass = [&](cudaStream_t q) {
        // depth 0: camera rays and their first intersection in one kernel over the whole pass; its
        // hits form the depth-0 live list.  Every other kernel walks a live list (grid-stride over
        // a grid that covers a few waves; CTAs beyond the live count exit at once).
        const int g = std::min(nb, t0->num_sms * 16);
#ifndef MERL_PRIMARY_PERSISTENT
#define MERL_PRIMARY_PERSISTENT 1
#endif
        // the camera-ray kernel on the same few-waves grid as the others: the scene is staged and the
        // per-material histogram flushed once per CTA, i.e. ~14x less often than with one CTA per 256 paths
        rt_primary_kernel<<<MERL_PRIMARY_PERSISTENT ? g : nb, 256, sizeof(int) * ntables, q>>>(R, P, A, d_sph, d_clus, d_smat, d_pass, film, d_live[0], d_nlive, d_counts);
        for (int depth = 0; depth <= p->max_depth; ++depth) {
            const int cur = depth & 1, nxt = cur ^ 1;
            const LiveList L{d_live[cur], d_nlive + cur, d_live[nxt], d_nlive + nxt};
            if (depth) rt_trace_kernel<<<g, 256, sizeof(int) * ntables, q>>>(R, P, A, d_sph, d_clus, d_smat, d_pass, depth, film, L, d_counts);
            if (depth == p->max_depth) break;   // the last trace only collects escaped paths
            // live queries -> contiguous, material-sorted segments (with one table this is plain
            // stream compaction: the BSDF stage then touches live queries only)
            rt_scan_kernel<<<1, 1024, 0, q>>>(d_counts, ntables, d_seg, d_cursor, d_items, 4 * std::min(grid_e, grid_f),
                                              d_nlive + nxt, d_nq);
            rt_scatter_kernel<<<g, 256, 2 * sizeof(int) * ntables, q>>>(P, A, B, d_pass, d_cursor, ntables, L);
            stream_multi_kernel<EvalPdfAosOp, 4 * kMinbScale><<<grid_e, kThreads, (size_t)img_e + kQueueBytes, q>>>(
                d_tabs, ntables, d_seg, d_items, EvalPdfAosOp{B.wi4, B.woe4, B.fe4}, img_e, own_world, own_rank);
            stream_multi_kernel<SampleAosOp, 3 * kMinbScale><<<grid_f, kThreads, (size_t)img_f + kQueueBytes, q>>>(
                d_tabs, ntables, d_seg, d_items, SampleAosOp{B.wi4, B.u4, B.wos4, B.w4}, img_f, own_world, own_rank);
            if (own_world > 1) {
                // this rank's share of the stage outputs -> path order in the exchange buffer, summed over
                // the ranks (every entry has exactly one non-zero contributor), shaded from there
                float4* x4 = reinterpret_cast<float4*>(p->stage_exchange->d_buf);
                const StageBuf O{nullptr, nullptr, nullptr, x4, x4 + cap, x4 + 2 * cap};
                cudaMemsetAsync(x4, 0, (size_t)cap * 48, q);
                rt_unsort_kernel<<<g, 256, 0, q>>>(P, B, O, d_seg + ntables, own_world, own_rank);
                if (merl_peerbuf_allreduce(p->stage_exchange, 12 * cap, nullptr, q) != MERL_OK) xchg_failed = true;
                rt_shade_kernel<<<g, 256, 0, q>>>(R, P, O, d_seg + ntables, film, L, 1);
            } else {
                rt_shade_kernel<<<g, 256, 0, q>>>(R, P, B, d_seg + ntables, film, L, 0);
            }
        }
        rt_advance_kernel<<<1, 1, 0, q>>>(d_pass, (long long)total, (int)cap, d_nlive);
    };
    const int64_t per_pass = 2 + (own_world > 1 ? 8 : 6) * (int64_t)p->max_depth;
    // (a caller that is itself capturing `stream` gets the plain launches: they become part of its graph)
    cudaStreamCaptureStatus cap_status = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &cap_status) != cudaSuccess) { cudaGetLastError(); cap_status = cudaStreamCaptureStatusNone; }
    if (npasses >= 2 && g_render_graphs.load() && cap_status == cudaStreamCaptureStatusNone) {
        RenderSide* side = render_side(t0->device);
        if (!side) { rc = MERL_ERR_CUDA; goto done; }
        std::lock_guard<std::mutex> lk(side->mu);
        side->reap();
        RT_TRY(cudaEventRecord(side->fork, s));
        RT_TRY(cudaStreamWaitEvent(side->stream, side->fork, 0));
        forked = side;
        cudaGraph_t graph = nullptr;
        RT_TRY(cudaStreamBeginCapture(side->stream, cudaStreamCaptureModeThreadLocal));
        enqueue_pass(side->stream);
        {
            const cudaError_t ec = cudaStreamEndCapture(side->stream, &graph);
            if (ec != cudaSuccess || !graph) {
                rc = fail(MERL_ERR_CUDA, "graph capture of a render pass failed: %s", cudaGetErrorString(ec));
                goto done;
            }
            const cudaError_t ei = side->adopt(graph);
            cudaGraphDestroy(graph);
            if (ei != cudaSuccess) {
                rc = fail(MERL_ERR_CUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(ei));
                goto done;
            }
        }
        for (int64_t k = 0; k < npasses; ++k) RT_TRY(cudaGraphLaunch(side->exec, side->stream));
        RT_TRY(cudaEventRecord(side->join, side->stream));
        RT_TRY(cudaStreamWaitEvent(s, side->join, 0));
        forked = nullptr;   // joined
    } else {
        for (int64_t k = 0; k < npasses; ++k) enqueue_pass(s);
        RT_TRY(cudaGetLastError());
    }
    if (xchg_failed) { rc = MERL_ERR_CUDA; goto done; }   // (message set by merl_peerbuf_allreduce)
    launches = npasses * per_pass;
    g_launches.fetch_add(launches - (g_launches.load() - launches_before));   // (exchanges enqueued above counted themselves)
    if (stats_out) {
        unsigned long long nq = 0;
        RT_TRY(cudaMemcpyAsync(&nq, d_nq, 8, cudaMemcpyDeviceToHost, s));
        RT_TRY(cudaStreamSynchronize(s));
        stats_out->paths = total; stats_out->bsdf_queries = (int64_t)nq; stats_out->launches = launches;
        // by-table mode: the frame's per-bounce exchanges have run by now; one that timed out left the stage
        // outputs of its pass unsummed (the kernel skips the data phase), so the film is not to be used
        if (own_world > 1 && p->stage_exchange->h_status && *reinterpret_cast<volatile int*>(p->stage_exchange->h_status)) {
            rc = fail(MERL_ERR_CUDA, "a stage exchange of this frame timed out waiting for another rank; the film is incomplete");
            goto done;
        }
    }
    }
done:
#undef RT_TRY
    if (forked) {
        // error after the fork: graph launches may still be queued on the side stream and they use the
        // workspace, so `s` must wait for them before the stream-ordered free below
        if (cudaEventRecord(forked->join, forked->stream) != cudaSuccess ||
            cudaStreamWaitEvent(s, forked->join, 0) != cudaSuccess)
            cudaStreamSynchronize(forked->stream);
        cudaGetLastError();
    }
    cudaFreeAsync(ws, s);
    return rc;
}
extern "C" int merl_render(const merl_table_t* const* tables, int ntables, const merl_sphere_t* spheres, int nspheres,
                           const merl_render_params_t* p, float* film4, merl_render_stats_t* stats_out,
                           merl_stream_t stream_) {
    return guarded("merl_render", [&]() -> int { return merl_render_impl(tables, ntables, spheres, nspheres, p, film4, stats_out, stream_); });
}
