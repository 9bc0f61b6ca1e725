// fake_device.cpp -- TEST DOUBLE for the CPU-only tests of the plugin hosts (tests/test_plugins.py,
// -m "not gpu").  It stands in for BOTH libraries the hosts link -- the CUDA runtime and
// libmerl_b200 -- so that the host plumbing (argument order, broadcast, masks, result wrapping,
// the CustomOp, parameters_changed, serialisation, the wavefront stage's buffer handling) can
// execute on a machine without a GPU.
//
// This is NOT an implementation of the BSDF and not a fallback: nothing here computes a bin, a
// lookup, a pdf or a sample.  "Device memory" is malloc, and every merl_* query entry point
// writes a POSITION-CODED function of its arguments (documented per function below) so a test can
// tell which pointer went where.  Only the binaries under plugins/_build/cpu/ link it; the product
// library and the GPU test binaries never see it.
#include <cuda_runtime_api.h>
#include <merl_b200.h>

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

// ---- CUDA runtime subset used by plugins/mock and plugins/mitsuba06/wavefront_stage.h -------------
extern "C" {
cudaError_t cudaMalloc(void** p, size_t n) { *p = std::malloc(n ? n : 1); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
cudaError_t cudaFree(void* p) { std::free(p); return cudaSuccess; }
cudaError_t cudaMallocHost(void** p, size_t n) { return cudaMalloc(p, n); }
cudaError_t cudaFreeHost(void* p) { std::free(p); return cudaSuccess; }
cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) { std::memmove(d, s, n); return cudaSuccess; }
cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t) { std::memmove(d, s, n); return cudaSuccess; }
cudaError_t cudaMemset(void* d, int v, size_t n) { std::memset(d, v, n); return cudaSuccess; }
cudaError_t cudaDeviceSynchronize(void) { return cudaSuccess; }
cudaError_t cudaSetDevice(int) { return cudaSuccess; }
cudaError_t cudaStreamCreate(cudaStream_t* s) { *s = reinterpret_cast<cudaStream_t>(0x1); return cudaSuccess; }
cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, unsigned) { return cudaStreamCreate(s); }
cudaError_t cudaStreamDestroy(cudaStream_t) { return cudaSuccess; }
cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
const char* cudaGetErrorString(cudaError_t) { return "fake device error"; }
}

// ---- libmerl_b200 subset ---------------------------------------------------------------------
struct merl_table {
    merl_table_params_t p;
    int64_t ncell;
    float* cells;      // float4[ncell]: cells[4k + c] = 1 + (k % 1024) + 0.25 c, .w = 7
    int refreshes;
};
static thread_local std::string g_err;
static std::atomic<int64_t> g_calls{0};
static int fail(int code, const std::string& m) { g_err = m; return code; }
static bool on(const uint8_t* active, int64_t i) { return !active || active[i]; }

extern "C" {
const char* merl_last_error(void) { return g_err.c_str(); }
int64_t merl_launch_count(void) { return g_calls.load(); }

void merl_table_params_default(merl_table_params_t* p) {
    std::memset(p, 0, sizeof *p);
    p->nh = 90; p->nd = 90; p->np = 180; p->theta_h_sqrt = 1; p->ni = 16; p->clamp_negative = 1;
    p->cdf_nm = 4; p->cdf_nq = 64; p->l2_persist = 1; p->cdf_floor = 0.05f; p->spec_weight = 0.5f;
    p->scale[0] = 1.0 / 1500.0; p->scale[1] = 1.15 / 1500.0; p->scale[2] = 1.66 / 1500.0;
}
static int create(const merl_table_params_t* p, merl_table_t** out) {
    if (p->ni < 1 || p->ni > 64 || (p->ni & (p->ni - 1))) return fail(MERL_ERR_ARG, "ni must be a power of two <= 64");
    if (p->spec_weight < 0.f || p->spec_weight > 1.f) return fail(MERL_ERR_ARG, "spec_weight must be in [0, 1]");
    merl_table* t = new merl_table{*p, (int64_t) p->nh * p->nd * p->np, nullptr, 0};
    t->cells = static_cast<float*>(std::malloc(sizeof(float) * 4 * t->ncell));
    for (int64_t k = 0; k < t->ncell; ++k) {
        for (int c = 0; c < 3; ++c) t->cells[4 * k + c] = 1.f + (float) (k % 1024) + 0.25f * c;
        t->cells[4 * k + 3] = 7.f;
    }
    *out = t;
    return MERL_OK;
}
// reads only the header: resolution from the file when the parameters leave it open
int merl_table_create_from_file(const char* path, const merl_table_params_t* p, int, merl_table_t** out) {
    if (!path || !out) return fail(MERL_ERR_ARG, "NULL argument");
    FILE* f = std::fopen(path, "rb");
    if (!f) return fail(MERL_ERR_IO, std::string("cannot open '") + path + "'");
    int32_t d[3] = {0, 0, 0};
    const size_t got = std::fread(d, 4, 3, f);
    std::fclose(f);
    if (got != 3 || d[0] <= 0 || d[1] <= 0 || d[2] <= 0) return fail(MERL_ERR_IO, std::string("'") + path + "': bad header");
    merl_table_params_t q;
    if (p) q = *p; else merl_table_params_default(&q);
    if (q.nh == 0 && q.nd == 0 && q.np == 0) { q.nh = d[0]; q.nd = d[1]; q.np = d[2]; }
    if ((int64_t) q.nh * q.nd * q.np != (int64_t) d[0] * d[1] * d[2]) return fail(MERL_ERR_IO, "resolution does not match the file");
    return create(&q, out);
}
int merl_table_create_from_host(const double* rgb, const merl_table_params_t* p, int, merl_table_t** out) {
    if (!rgb || !p || !out) return fail(MERL_ERR_ARG, "NULL argument");
    return create(p, out);
}
int merl_table_destroy(merl_table_t* t) { if (t) { std::free(t->cells); delete t; } return MERL_OK; }
int merl_table_info(const merl_table_t* t, merl_table_info_t* o) {
    std::memset(o, 0, sizeof *o);
    o->nh = t->p.nh; o->nd = t->p.nd; o->np = t->p.np; o->ni = t->p.ni; o->theta_h_sqrt = t->p.theta_h_sqrt;
    o->ncell = t->ncell; o->table_bytes = t->ncell * 16; o->d_cells = t->cells; o->spec_weight = t->p.spec_weight;
    return MERL_OK;
}
int merl_table_download(const merl_table_t* t, float* cells4, float*, float*) {
    if (cells4) std::memcpy(cells4, t->cells, sizeof(float) * 4 * t->ncell);
    return MERL_OK;
}
int merl_table_refresh_sampler(merl_table_t* t, merl_stream_t) { ++t->refreshes; return MERL_OK; }
int merl_table_update_cells(merl_table_t* t, const void* cells4, merl_stream_t) {
    if (!t || !cells4) return fail(MERL_ERR_ARG, "NULL argument");
    if (cells4 != t->cells) std::memcpy(t->cells, cells4, sizeof(float) * 4 * t->ncell);
    ++t->refreshes;
    return MERL_OK;
}

// eval: (r, g, b)[i] = cells[i % ncell].(r, g, b) * wo.z[i];  pdf[i] = wi.x + 2 wi.y + 4 wi.z + 8 wo.x + 16 wo.y + 32 wo.z
static float code_pdf(float wix, float wiy, float wiz, float wox, float woy, float woz) {
    return wix + 2.f * wiy + 4.f * wiz + 8.f * wox + 16.f * woy + 32.f * woz;
}
int merl_eval_pdf(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz, const float* wox,
                  const float* woy, const float* woz, const uint8_t* active, int64_t n, float* r, float* g, float* b,
                  float* pdf, merl_stream_t) {
    ++g_calls;
    for (int64_t i = 0; i < n; ++i) {
        const bool a = on(active, i);
        const float* c = t->cells + 4 * (i % t->ncell);
        if (r) { r[i] = a ? c[0] * woz[i] : 0.f; g[i] = a ? c[1] * woz[i] : 0.f; b[i] = a ? c[2] * woz[i] : 0.f; }
        if (pdf) pdf[i] = a ? code_pdf(wix[i], wiy[i], wiz[i], wox[i], woy[i], woz[i]) : 0.f;
    }
    return MERL_OK;
}
int merl_eval(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz, const float* wox,
              const float* woy, const float* woz, const uint8_t* active, int64_t n, float* r, float* g, float* b,
              merl_stream_t s) {
    return merl_eval_pdf(t, wix, wiy, wiz, wox, woy, woz, active, n, r, g, b, nullptr, s);
}
int merl_pdf(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz, const float* wox,
             const float* woy, const float* woz, const uint8_t* active, int64_t n, float* pdf, merl_stream_t s) {
    return merl_eval_pdf(t, wix, wiy, wiz, wox, woy, woz, active, n, nullptr, nullptr, nullptr, pdf, s);
}
// sample: wo = (u1, u2x, u2y); pdf = wi.x + 2 wi.y + 4 wi.z; weight = 2 (u1, u2x, u2y); lobe = i & 1
int merl_sample(const merl_table_t*, const float* wix, const float* wiy, const float* wiz, const float* u1,
                const float* u2x, const float* u2y, const uint8_t* active, int64_t n, float* wox, float* woy,
                float* woz, float* pdf, float* wr, float* wg, float* wb, uint32_t* lobe, merl_stream_t) {
    ++g_calls;
    for (int64_t i = 0; i < n; ++i) {
        const float a = on(active, i) ? 1.f : 0.f;
        wox[i] = a * u1[i]; woy[i] = a * u2x[i]; woz[i] = a * u2y[i];
        pdf[i] = a * (wix[i] + 2.f * wiy[i] + 4.f * wiz[i]);
        wr[i] = a * 2.f * u1[i]; wg[i] = a * 2.f * u2x[i]; wb[i] = a * 2.f * u2y[i];
        lobe[i] = a != 0.f ? (uint32_t) (i & 1) : 0u;
    }
    return MERL_OK;
}
// backward: grad[i % ncell] += (gr, gg, gb, 1/wo.z)[i] * wo.z[i]
int merl_backward(const merl_table_t* t, const float*, const float*, const float*, const float*, const float*,
                  const float* woz, const float* gr, const float* gg, const float* gb, const uint8_t* active,
                  int64_t n, float* grad, uint32_t, merl_stream_t) {
    ++g_calls;
    for (int64_t i = 0; i < n; ++i) {
        if (!on(active, i)) continue;
        float* c = grad + 4 * (i % t->ncell);
        c[0] += gr[i] * woz[i]; c[1] += gg[i] * woz[i]; c[2] += gb[i] * woz[i]; c[3] += 1.f;
    }
    return MERL_OK;
}
int merl_grad_clear_counts(const merl_table_t* t, float* grad, merl_stream_t) {
    ++g_calls;
    for (int64_t k = 0; k < t->ncell; ++k) grad[4 * k + 3] = 0.f;
    return MERL_OK;
}
// AoS: out4 = (cells[i % ncell].rgb * wo4.z, code_pdf);  sample: wo4_out = (u4.xyz, wi code), weight4 = (2 u4.xyz, i & 1)
int merl_eval_pdf_aos(const merl_table_t* t, const float* wi4, const float* wo4, int64_t n, float* out4, merl_stream_t) {
    ++g_calls;
    for (int64_t i = 0; i < n; ++i) {
        const float* c = t->cells + 4 * (i % t->ncell);
        const float *a = wi4 + 4 * i, *b = wo4 + 4 * i;
        out4[4 * i + 0] = c[0] * b[2]; out4[4 * i + 1] = c[1] * b[2]; out4[4 * i + 2] = c[2] * b[2];
        out4[4 * i + 3] = code_pdf(a[0], a[1], a[2], b[0], b[1], b[2]);
    }
    return MERL_OK;
}
int merl_sample_aos(const merl_table_t*, const float* wi4, const float* u4, int64_t n, float* wo4, float* w4, merl_stream_t) {
    ++g_calls;
    for (int64_t i = 0; i < n; ++i) {
        const float *a = wi4 + 4 * i, *u = u4 + 4 * i;
        wo4[4 * i + 0] = u[0]; wo4[4 * i + 1] = u[1]; wo4[4 * i + 2] = u[2]; wo4[4 * i + 3] = a[0] + 2.f * a[1] + 4.f * a[2];
        w4[4 * i + 0] = 2.f * u[0]; w4[4 * i + 1] = 2.f * u[1]; w4[4 * i + 2] = 2.f * u[2]; w4[4 * i + 3] = (float) (i & 1);
    }
    return MERL_OK;
}
int merl_eval_pdf_host(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz, const float* wox,
                       const float* woy, const float* woz, int64_t n, float* r, float* g, float* b, float* pdf) {
    return merl_eval_pdf(t, wix, wiy, wiz, wox, woy, woz, nullptr, n, r, g, b, pdf, nullptr);
}
int merl_sample_host(const merl_table_t* t, const float* wix, const float* wiy, const float* wiz, const float* u1,
                     const float* u2x, const float* u2y, int64_t n, float* wox, float* woy, float* woz, float* pdf,
                     float* wr, float* wg, float* wb, uint32_t* lobe) {
    return merl_sample(t, wix, wiy, wiz, u1, u2x, u2y, nullptr, n, wox, woy, woz, pdf, wr, wg, wb, lobe, nullptr);
}
}
