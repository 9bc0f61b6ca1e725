// measured_bsdf.h -- Mitsuba 0.6 host side shared by the `merl` and `customized_measurment` BSDF
// plugins, on libmerl_b200.
//
// Replaces: the reference's mitsuba/src/bsdfs/{merl,customized_measurment}.cpp (named by
// /root/reference/README.md:1; sources absent, /root/reference/.SUBMODULES.json:8-14), i.e. the
// upstream virtuals BSDF::eval / pdf / sample(bRec, sample) / sample(bRec, pdf, sample) on a
// BSDFSamplingRecord.  Semantics kept: bRec.wi / bRec.wo in the local frame;
// zero unless measure == ESolidAngle, the glossy-reflection component is selected by
// bRec.typeMask / bRec.component, and both directions are in the upper hemisphere; eval includes
// cos(theta_o); sample() fills bRec.wo, eta = 1, sampledComponent = 0, sampledType =
// EGlossyReflection and returns eval/pdf; errors by Log(EError, ...), which throws.
//
// The scalar virtuals forward ONE query to the *_host entry points (correct and PCIe-latency
// bound: for parity runs and for integrators that have not been ported).  The throughput path is
// the wavefront stage (wavefront_stage.h): evalBatch()/sampleBatch() below.  No CPU code path.
#pragma once
#include <mitsuba/core/fresolver.h>
#include <mitsuba/core/fstream.h>
#include <mitsuba/core/properties.h>
#include <mitsuba/render/bsdf.h>

#include <fstream>

#include "../common/merl_host.h"
#include "wavefront_stage.h"

MTS_NAMESPACE_BEGIN

class MeasuredBSDF : public BSDF {
public:
    MeasuredBSDF(const Properties& props, const char* plugin, const merl_table_params_t& params)
        : BSDF(props), m_plugin(plugin), m_params(params) {
        m_params.spec_weight = props.getFloat("specWeight", m_params.spec_weight);
        m_params.ni = props.getInteger("incidentSlices", m_params.ni);
        m_device = props.getInteger("device", 0);
        m_filename = Thread::getThread()->getFileResolver()->resolve(props.getString("filename"));
        readFile();
    }

    // Network rendering: the table travels in the stream (mtssrv nodes have no access to the file).
    MeasuredBSDF(Stream* stream, InstanceManager* manager, const char* plugin) : BSDF(stream, manager), m_plugin(plugin) {
        m_filename = stream->readString();
        stream->read(&m_params, sizeof m_params);
        m_device = stream->readInt();
        m_raw.resize(3 * cellCount());
        stream->readDoubleArray(m_raw.data(), m_raw.size());
    }

    void serialize(Stream* stream, InstanceManager* manager) const override {
        BSDF::serialize(stream, manager);
        stream->writeString(m_filename);
        stream->write(&m_params, sizeof m_params);
        stream->writeInt(m_device);
        stream->writeDoubleArray(m_raw.data(), m_raw.size());
    }

    void configure() override {
        check(merl_table_create_from_host(m_raw.data(), &m_params, m_device, m_table.put()), "building the table");
        m_info = m_table.info();
        m_components.clear();
        m_components.push_back(EGlossyReflection | EFrontSide);
        m_usesRayDifferentials = false;
        BSDF::configure();
    }

    // ---- scalar virtuals -------------------------------------------------------------------
    Spectrum eval(const BSDFSamplingRecord& bRec, EMeasure measure) const override {
        Spectrum value; Float pdf;
        return evalPdfOne(bRec, measure, value, pdf) ? value : Spectrum(0.0f);
    }
    Float pdf(const BSDFSamplingRecord& bRec, EMeasure measure) const override {
        Spectrum value; Float pdf;
        return evalPdfOne(bRec, measure, value, pdf) ? pdf : 0.0f;
    }
    Spectrum sample(BSDFSamplingRecord& bRec, Float& pdf, const Point2& sample_) const override {
        pdf = 0.0f;
        if (!enabled(bRec) || Frame::cosTheta(bRec.wi) <= 0) return Spectrum(0.0f);
        float u1, u2x;
        splitSample(sample_.x, u1, u2x);
        const float u2y = sample_.y;
        warnScalarOnce();
        float o[3], w[3];
        uint32_t lobe;
        check(merl_sample_host(m_table.get(), &bRec.wi.x, &bRec.wi.y, &bRec.wi.z, &u1, &u2x, &u2y, 1, &o[0], &o[1],
                               &o[2], &pdf, &w[0], &w[1], &w[2], &lobe), "sample");
        if (pdf <= 0) return Spectrum(0.0f);
        bRec.wo = Vector(o[0], o[1], o[2]);
        bRec.eta = 1.0f;
        bRec.sampledComponent = 0;
        bRec.sampledType = EGlossyReflection;
        Spectrum weight;
        weight.fromLinearRGB(w[0], w[1], w[2]);
        return weight;
    }
    Spectrum sample(BSDFSamplingRecord& bRec, const Point2& sample_) const override {
        Float pdf;
        return sample(bRec, pdf, sample_);
    }

    // ---- wavefront stage -------------------------------------------------------------------
    // 0.6 hands the plugin two random numbers; the lobe choice reuses sample.x (the usual 0.6
    // idiom), rescaled so that the 2-D sample stays uniform.
    void splitSample(float sx, float& u1, float& u2x) const {
        const float w = m_info.spec_weight;
        u1 = sx;
        u2x = sx < w ? sx / w : (sx - w) / (1.0f - w);
    }
    // Queue one record; results come from stage.result(slot) after evalBatch() + stage.fetch().
    size_t push(merl_host::WavefrontStage& stage, const BSDFSamplingRecord& bRec) const {
        const bool on = enabled(bRec);
        // a disabled record is queued below the horizon, which evaluates to zero
        return stage.pushEval(bRec.wi.x, bRec.wi.y, on ? bRec.wi.z : -1.0f, bRec.wo.x, bRec.wo.y, bRec.wo.z);
    }
    size_t pushSample(merl_host::WavefrontStage& stage, const BSDFSamplingRecord& bRec, const Point2& sample_) const {
        float u1, u2x;
        splitSample(sample_.x, u1, u2x);
        return stage.pushSample(bRec.wi.x, bRec.wi.y, enabled(bRec) ? bRec.wi.z : -1.0f, u1, u2x, sample_.y);
    }
    void evalBatch(merl_host::WavefrontStage& stage) const { stage.launchEvalPdf(m_table.get()); }
    void sampleBatch(merl_host::WavefrontStage& stage) const { stage.launchSample(m_table.get()); }

    std::string toString() const override {
        std::ostringstream oss;
        oss << getClassName() << "[" << std::endl
            << "  filename = \"" << m_filename << "\"," << std::endl
            << "  resolution = " << m_info.nh << "x" << m_info.nd << "x" << m_info.np << "," << std::endl
            << "  specWeight = " << m_info.spec_weight << std::endl
            << "]";
        return oss.str();
    }
    const merl_table_t* table() const { return m_table.get(); }

protected:
    void check(int rc, const char* what) const {
        merl_host::check(rc, m_plugin, what, [](const std::string& m) { Log(EError, "%s", m.c_str()); });
    }
    bool enabled(const BSDFSamplingRecord& bRec) const {
        return (bRec.typeMask & EGlossyReflection) && (bRec.component == -1 || bRec.component == 0);
    }
    // The scalar virtuals are one host->GPU->host round trip per query (no CPU implementation exists in this
    // plugin): correct, and ~1e4-1e5 queries/s.  Said once per process; the batched path is WavefrontStage.
    void warnScalarOnce() const {
        static bool warned = false;
        if (!warned) {
            warned = true;
            Log(EWarn, "%s: BSDF::eval/pdf/sample run ONE PCIe round trip per call on this plugin (no CPU path); "
                       "an integrator that gathers a bounce's hits should use WavefrontStage (wavefront_stage.h), "
                       "which serves whole batches at ~1e11 queries/s.", m_plugin);
        }
    }
    size_t cellCount() const { return (size_t) m_params.nh * (size_t) m_params.nd * (size_t) m_params.np; }

    bool evalPdfOne(const BSDFSamplingRecord& bRec, EMeasure measure, Spectrum& value, Float& pdf) const {
        if (measure != ESolidAngle || !enabled(bRec) || Frame::cosTheta(bRec.wi) <= 0 || Frame::cosTheta(bRec.wo) <= 0)
            return false;
        warnScalarOnce();
        float c[3];
        check(merl_eval_pdf_host(m_table.get(), &bRec.wi.x, &bRec.wi.y, &bRec.wi.z, &bRec.wo.x, &bRec.wo.y,
                                 &bRec.wo.z, 1, &c[0], &c[1], &c[2], &pdf), "eval");
        value.fromLinearRGB(c[0], c[1], c[2]);
        return true;
    }

    // MERL .binary layout: int32 dims[3], then planar R, G, B doubles.  Kept
    // on the host only so that serialize() can ship it; all validation beyond the header is the
    // library's (merl_table_create_from_host).
    void readFile() {
        std::ifstream f(m_filename.c_str(), std::ios::binary);
        int32_t dims[3];
        if (!f.read(reinterpret_cast<char*>(dims), 12)) Log(EError, "%s: cannot read \"%s\"", m_plugin, m_filename.c_str());
        if (m_params.nh == 0) { m_params.nh = dims[0]; m_params.nd = dims[1]; m_params.np = dims[2]; }
        if (dims[0] <= 0 || dims[1] <= 0 || dims[2] <= 0 || (size_t) dims[0] * (size_t) dims[1] * (size_t) dims[2] != cellCount())
            Log(EError, "%s: \"%s\" has dimensions %dx%dx%d, expected %dx%dx%d", m_plugin, m_filename.c_str(), dims[0],
                dims[1], dims[2], m_params.nh, m_params.nd, m_params.np);
        m_raw.resize(3 * cellCount());
        if (!f.read(reinterpret_cast<char*>(m_raw.data()), (std::streamsize) (m_raw.size() * 8)))
            Log(EError, "%s: \"%s\" is truncated", m_plugin, m_filename.c_str());
    }

    const char* m_plugin;
    merl_table_params_t m_params;
    std::string m_filename;
    int m_device = 0;
    std::vector<double> m_raw;
    merl_host::Table m_table;
    merl_table_info_t m_info{};
};

MTS_NAMESPACE_END
