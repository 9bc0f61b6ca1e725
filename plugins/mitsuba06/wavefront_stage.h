// wavefront_stage.h -- the batched BSDF stage for Mitsuba 0.6 hosts.
//
// Upstream 0.6 calls BSDF::eval/pdf/sample once per path vertex from scalar integrators; it has no
// wavefront stage (BASELINE.json asks for one).  An integrator that wants the GPU path gathers one
// bounce's queries here and issues ONE call: records are packed as float4 AoS into pinned host
// memory, copied to the device, run through merl_eval_pdf_aos / merl_sample_aos and copied back.
// All of it is asynchronous on the stage's stream until fetch().  The stage computes nothing.
#pragma once
#include <cuda_runtime_api.h>
#include <merl_b200.h>

#include <stdexcept>
#include <string>

namespace merl_host {

class WavefrontStage {
public:
    struct Float4 { float x, y, z, w; };

    explicit WavefrontStage(size_t capacity, int device = 0) : m_cap(capacity), m_device(device) {
        ok(cudaSetDevice(device), "cudaSetDevice");
        ok(cudaStreamCreateWithFlags(&m_stream, cudaStreamNonBlocking), "cudaStreamCreate");
        const size_t bytes = capacity * sizeof(Float4);
        for (int k = 0; k < 4; ++k) {
            ok(cudaMallocHost((void**) &m_host[k], bytes), "cudaMallocHost");
            ok(cudaMalloc((void**) &m_dev[k], bytes), "cudaMalloc");
        }
    }
    WavefrontStage(const WavefrontStage&) = delete;
    WavefrontStage& operator=(const WavefrontStage&) = delete;
    ~WavefrontStage() {
        for (int k = 0; k < 4; ++k) { cudaFreeHost(m_host[k]); cudaFree(m_dev[k]); }
        if (m_stream) cudaStreamDestroy(m_stream);
    }

    size_t capacity() const { return m_cap; }
    size_t size() const { return m_n; }
    void clear() { m_n = 0; }

    // eval/pdf query: (wi, wo) in the local frame.  Returns the slot to read the result from.
    size_t pushEval(float wix, float wiy, float wiz, float wox, float woy, float woz) {
        const size_t i = slot();
        m_host[0][i] = { wix, wiy, wiz, 0.f };
        m_host[1][i] = { wox, woy, woz, 0.f };
        return i;
    }
    // sample query: wi and three uniform numbers (lobe choice, 2-D sample).
    size_t pushSample(float wix, float wiy, float wiz, float u1, float u2x, float u2y) {
        const size_t i = slot();
        m_host[0][i] = { wix, wiy, wiz, 0.f };
        m_host[1][i] = { u1, u2x, u2y, 0.f };
        return i;
    }

    // out slot i: (r, g, b, pdf), eval including cos(theta_o).
    void launchEvalPdf(const merl_table_t* table) {
        upload(2);
        call(merl_eval_pdf_aos(table, (const float*) m_dev[0], (const float*) m_dev[1], (int64_t) m_n,
                               (float*) m_dev[2], m_stream), "merl_eval_pdf_aos");
        download(2, 1);
    }
    // out slot i: wo() = (wo.xyz, pdf), weight() = (weight.rgb, lobe).
    void launchSample(const merl_table_t* table) {
        upload(2);
        call(merl_sample_aos(table, (const float*) m_dev[0], (const float*) m_dev[1], (int64_t) m_n,
                             (float*) m_dev[2], (float*) m_dev[3], m_stream), "merl_sample_aos");
        download(2, 2);
    }
    // Blocks until the last launch*'s results are in host memory.
    void fetch() { ok(cudaStreamSynchronize(m_stream), "cudaStreamSynchronize"); }

    const Float4& result(size_t i) const { return m_host[2][i]; }
    const Float4& wo(size_t i) const { return m_host[2][i]; }
    const Float4& weight(size_t i) const { return m_host[3][i]; }

private:
    size_t slot() {
        if (m_n >= m_cap) throw std::runtime_error("merl wavefront stage: capacity exceeded");
        return m_n++;
    }
    void upload(int nbuf) {
        ok(cudaSetDevice(m_device), "cudaSetDevice");
        for (int k = 0; k < nbuf; ++k)
            ok(cudaMemcpyAsync(m_dev[k], m_host[k], m_n * sizeof(Float4), cudaMemcpyHostToDevice, m_stream), "H2D");
    }
    void download(int first, int nbuf) {
        for (int k = first; k < first + nbuf; ++k)
            ok(cudaMemcpyAsync(m_host[k], m_dev[k], m_n * sizeof(Float4), cudaMemcpyDeviceToHost, m_stream), "D2H");
    }
    static void ok(cudaError_t e, const char* what) {
        if (e != cudaSuccess) throw std::runtime_error(std::string("merl wavefront stage: ") + what + ": " + cudaGetErrorString(e));
    }
    static void call(int rc, const char* what) {
        if (rc != MERL_OK) throw std::runtime_error(std::string("merl wavefront stage: ") + what + ": " + merl_last_error());
    }

    size_t m_cap, m_n = 0;
    int m_device;
    cudaStream_t m_stream = nullptr;
    Float4* m_host[4] = {};
    Float4* m_dev[4] = {};
};

}  // namespace merl_host
