// measured_bsdf.h -- Mitsuba 3 host side shared by the `merl` and `customized_measurment` BSDF
// plugins: a BSDF subclass whose eval / pdf / eval_pdf / sample hand device pointers of Dr.Jit
// CUDA arrays to libmerl_b200 (include/merl_b200.h) and wrap the results without a copy.
//
// Replaces: the reference's mitsuba3/src/bsdfs/{merl,customized_measurment}.cpp (named by
// /root/reference/README.md:1; sources absent, /root/reference/.SUBMODULES.json:15-21), i.e. the
// upstream virtuals `BSDF::sample/eval/pdf/eval_pdf(const BSDFContext&, const
// SurfaceInteraction3f&, ...)`.  Semantics: wi = si.wi and wo in the local
// shading frame, eval includes cos(theta_o), zero unless wi.z > 0 and wo.z > 0, one component
// (GlossyReflection | FrontSide), errors by Throw().
//
// Variants: cuda_* -> device pointers straight through.  llvm_* -> the whole packet through the
// *_host entry points (host pointers of the evaluated arrays, n = packet width; the library stages the
// copies in chunks over three streams), inactive lanes zeroed with dr::select afterwards.  scalar_* ->
// the *_host entry points with n = 1: correct, but ONE PCIe round trip per query (~1e4-1e5 queries/s);
// for parity runs only -- the plugin logs that once.  There is no CPU implementation in this plugin.
//
// JIT variants need WAVEFRONT mode.  These virtuals call an external kernel, so their arguments must
// be evaluated arrays (dr::eval + .data()).  Under Dr.Jit's default flags Mitsuba records virtual
// function calls and loops symbolically: the arguments are placeholders then and evaluating one
// raises deep inside Dr.Jit.  The plugin checks for that first and says what to do:
//     dr.set_flag(dr.JitFlag.VCallRecord, False); dr.set_flag(dr.JitFlag.LoopRecord, False)   # Dr.Jit 0.4
//     dr.set_flag(dr.JitFlag.SymbolicCalls, False); dr.set_flag(dr.JitFlag.SymbolicLoops, False)   # 1.x
#pragma once
#include <mitsuba/core/fresolver.h>
#include <mitsuba/core/properties.h>
#include <mitsuba/render/bsdf.h>

#include <drjit/custom.h>

#include <algorithm>

#include "../common/merl_host.h"

// "Dr.Jit is recording symbolically right now": JitFlag::Recording in drjit-core 0.4.x, renamed
// SymbolicScope in 1.x (both: set only while a call / loop is being captured).
#if defined(DRJIT_VERSION_MAJOR) && DRJIT_VERSION_MAJOR >= 1
#  define MERL_JIT_RECORDING_FLAG JitFlag::SymbolicScope
#else
#  define MERL_JIT_RECORDING_FLAG JitFlag::Recording
#endif

NAMESPACE_BEGIN(mitsuba)

template <typename Float, typename Spectrum> struct MerlEvalOp;   // merl_eval_op.h (end of this file)

template <typename Float, typename Spectrum> class MeasuredBSDF : public BSDF<Float, Spectrum> {
public:
    MI_IMPORT_BASE(BSDF, m_flags, m_components)
    MI_IMPORT_TYPES()

    // `params` arrives filled by the concrete plugin (MERL: fixed layout; customized: from props).
    MeasuredBSDF(const Properties& props, const char* plugin, merl_table_params_t params)
        : Base(props), m_plugin(plugin) {
        // sampler knobs common to both plugins (assumption B5: the fork's own are unknown)
        params.spec_weight = props.get<float>("spec_weight", params.spec_weight);
        params.ni = props.get<int>("ni", params.ni);
        params.l2_persist = props.get<bool>("l2_persist", params.l2_persist != 0) ? 1 : 0;
        m_device = props.get<int>("device", 0);
        m_filename = Thread::thread()->file_resolver()->resolve(props.string("filename"));
        check(merl_table_create_from_file(m_filename.c_str(), &params, m_device, m_table.put()),
              "loading the table");
        m_info = m_table.info();
        // the differentiable parameter: the library's float4 table seen as a flat Float (no copy)
        if constexpr (dr::is_cuda_v<Float>)
            m_data = dr::map<Float>(const_cast<void*>(m_info.d_cells), (size_t) (4 * m_info.ncell));
        m_flags = BSDFFlags::GlossyReflection | BSDFFlags::FrontSide;
        m_components.push_back(m_flags);
    }

    // ---- BSDF interface --------------------------------------------------------------------
    Spectrum eval(const BSDFContext& ctx, const SurfaceInteraction3f& si, const Vector3f& wo_,
                  Mask active_) const override {
        MI_MASKED_FUNCTION(ProfilerPhase::BSDFEvaluate, active_);
        if (!ctx.is_enabled(BSDFFlags::GlossyReflection)) return Spectrum(0.f);
        if constexpr (dr::is_cuda_v<Float>) {
            require_wavefront();
            // *_ad_* variants: when `data` carries gradients the evaluation enters the AD graph as
            // one custom operation whose adjoint is the scatter kernel (merl_eval_op.h)
            if constexpr (dr::is_diff_v<Float>) {
                if (dr::grad_enabled(m_data))
                    return dr::custom<MerlEvalOp<Float, Spectrum>>(this, m_data, si.wi, wo_, active_);
            }
            return eval_primal(si.wi, wo_, active_);
        } else {
            return eval_pdf_packet(si.wi, wo_, active_).first;
        }
    }

    // The kernel call itself, on detached operands (cuda variants only).
    Spectrum eval_primal(const Vector3f& wi_, const Vector3f& wo_, Mask active_) const {
        if constexpr (dr::is_cuda_v<Float>) {
            require_wavefront();
            Vector3f wi = wi_, wo = wo_;
            Mask active = active_;
            const size_t n = materialise(wi, wo, active);
            Float r = dr::empty<Float>(n), g = dr::empty<Float>(n), b = dr::empty<Float>(n);
            check(merl_eval(m_table.get(), wi.x().data(), wi.y().data(), wi.z().data(), wo.x().data(),
                            wo.y().data(), wo.z().data(), mask_ptr(active), (int64_t) n, r.data(), g.data(),
                            b.data(), jit_cuda_stream()),
                  "eval");
            return Spectrum(r, g, b);
        } else {
            Throw("%s: eval_primal is the cuda variants' path", m_plugin);
        }
    }

    Float pdf(const BSDFContext& ctx, const SurfaceInteraction3f& si, const Vector3f& wo_,
              Mask active_) const override {
        MI_MASKED_FUNCTION(ProfilerPhase::BSDFEvaluate, active_);
        if (!ctx.is_enabled(BSDFFlags::GlossyReflection)) return Float(0.f);
        if constexpr (dr::is_cuda_v<Float>) {
            require_wavefront();
            Vector3f wi = si.wi, wo = wo_;
            Mask active = active_;
            const size_t n = materialise(wi, wo, active);
            Float p = dr::empty<Float>(n);
            check(merl_pdf(m_table.get(), wi.x().data(), wi.y().data(), wi.z().data(), wo.x().data(),
                           wo.y().data(), wo.z().data(), mask_ptr(active), (int64_t) n, p.data(),
                           jit_cuda_stream()),
                  "pdf");
            return p;
        } else {
            return eval_pdf_packet(si.wi, wo_, active_).second;
        }
    }

    std::pair<Spectrum, Float> eval_pdf(const BSDFContext& ctx, const SurfaceInteraction3f& si,
                                        const Vector3f& wo_, Mask active_) const override {
        MI_MASKED_FUNCTION(ProfilerPhase::BSDFEvaluate, active_);
        if (!ctx.is_enabled(BSDFFlags::GlossyReflection)) return { Spectrum(0.f), Float(0.f) };
        if constexpr (dr::is_cuda_v<Float>) {
            require_wavefront();
            Vector3f wi = si.wi, wo = wo_;
            Mask active = active_;
            const size_t n = materialise(wi, wo, active);
            Float r = dr::empty<Float>(n), g = dr::empty<Float>(n), b = dr::empty<Float>(n),
                  p = dr::empty<Float>(n);
            check(merl_eval_pdf(m_table.get(), wi.x().data(), wi.y().data(), wi.z().data(), wo.x().data(),
                                wo.y().data(), wo.z().data(), mask_ptr(active), (int64_t) n, r.data(),
                                g.data(), b.data(), p.data(), jit_cuda_stream()),
                  "eval_pdf");
            return { Spectrum(r, g, b), p };
        } else {
            return eval_pdf_packet(si.wi, wo_, active_);
        }
    }

    std::pair<BSDFSample3f, Spectrum> sample(const BSDFContext& ctx, const SurfaceInteraction3f& si,
                                             Float sample1_, const Point2f& sample2_,
                                             Mask active_) const override {
        MI_MASKED_FUNCTION(ProfilerPhase::BSDFSample, active_);
        BSDFSample3f bs;
        if (!ctx.is_enabled(BSDFFlags::GlossyReflection)) {
            bs.wo = Vector3f(0.f); bs.pdf = Float(0.f); bs.eta = Float(1.f);
            bs.sampled_type = UInt32(0u); bs.sampled_component = UInt32(0u);
            return { bs, Spectrum(0.f) };
        }
        if constexpr (dr::is_jit_v<Float>) {
            require_wavefront();
            Vector3f wi = si.wi;
            Float u1 = sample1_;
            Point2f u2 = sample2_;
            Mask active = active_;
            const size_t n = dr::width(wi.x(), wi.y(), wi.z(), u1, u2.x(), u2.y(), active);
            dr::resize(wi.x(), n); dr::resize(wi.y(), n); dr::resize(wi.z(), n);
            dr::resize(u1, n); dr::resize(u2.x(), n); dr::resize(u2.y(), n); dr::resize(active, n);
            dr::eval(wi, u1, u2, active);                 // materialise: the external call breaks the megakernel
            // llvm_*: dr::eval() only queues the kernels on Dr.Jit's thread pool; the host-buffer entry point
            // reads the packets' memory right away, so wait for them (CUDA variants are ordered by the stream)
            if constexpr (dr::is_llvm_v<Float>) dr::sync_thread();
            Float ox = dr::empty<Float>(n), oy = dr::empty<Float>(n), oz = dr::empty<Float>(n),
                  p = dr::empty<Float>(n), wr = dr::empty<Float>(n), wg = dr::empty<Float>(n),
                  wb = dr::empty<Float>(n);
            UInt32 lobe = dr::empty<UInt32>(n);
            if constexpr (dr::is_cuda_v<Float>) {
                check(merl_sample(m_table.get(), wi.x().data(), wi.y().data(), wi.z().data(), u1.data(),
                                  u2.x().data(), u2.y().data(), mask_ptr(active), (int64_t) n, ox.data(), oy.data(),
                                  oz.data(), p.data(), wr.data(), wg.data(), wb.data(), lobe.data(),
                                  jit_cuda_stream()),
                      "sample");
            } else {
                // llvm_*: the packet's host memory through the host-buffer entry point, all n lanes in one
                // call; lanes that were masked out are zeroed afterwards (the entry point takes no mask)
                check(merl_sample_host(m_table.get(), wi.x().data(), wi.y().data(), wi.z().data(), u1.data(),
                                       u2.x().data(), u2.y().data(), (int64_t) n, ox.data(), oy.data(), oz.data(),
                                       p.data(), wr.data(), wg.data(), wb.data(), lobe.data()),
                      "sample");
                const Float zero(0.f);
                ox = dr::select(active, ox, zero); oy = dr::select(active, oy, zero); oz = dr::select(active, oz, zero);
                p = dr::select(active, p, zero);
                wr = dr::select(active, wr, zero); wg = dr::select(active, wg, zero); wb = dr::select(active, wb, zero);
                lobe = dr::select(active, lobe, UInt32(0u));
            }
            bs.wo = Vector3f(ox, oy, oz);
            bs.pdf = p;
            bs.eta = Float(1.f);
            bs.sampled_type = UInt32(+BSDFFlags::GlossyReflection);
            bs.sampled_component = UInt32(0u);            // one component; `lobe` is internal to the sampler
            m_last_lobe = lobe;
            return { bs, Spectrum(wr, wg, wb) };
        } else {
            warn_scalar_once();
            const float wi[3] = { (float) si.wi.x(), (float) si.wi.y(), (float) si.wi.z() };
            const float u[3] = { (float) sample1_, (float) sample2_.x(), (float) sample2_.y() };
            float o[3] = {}, p = 0.f, w[3] = {};
            uint32_t lobe = 0;
            if ((bool) active_)
                check(merl_sample_host(m_table.get(), &wi[0], &wi[1], &wi[2], &u[0], &u[1], &u[2], 1, &o[0],
                                       &o[1], &o[2], &p, &w[0], &w[1], &w[2], &lobe),
                      "sample");
            bs.wo = Vector3f(o[0], o[1], o[2]);
            bs.pdf = p;
            bs.eta = 1.f;
            bs.sampled_type = +BSDFFlags::GlossyReflection;
            bs.sampled_component = 0u;
            m_last_lobe = lobe;
            return { bs, Spectrum(w[0], w[1], w[2]) };
        }
    }

    // ---- differentiable parameter ----------------------------------------------------------
    // `data` is the prescaled float4 table on the device, exposed as a flat Float of 4 * ncell
    // entries (r, g, b, validity bits per cell).  In *_ad_* variants Dr.Jit would derive the
    // adjoint of a dr::gather itself; with an external kernel it is wired by hand: eval() enters
    // the AD graph through MerlEvalOp, whose backward() calls backward_table().
    void traverse(TraversalCallback* cb) override {
        if constexpr (dr::is_cuda_v<Float>)
            cb->put_parameter("data", m_data, +ParamFlags::Differentiable);
        else
            cb->put_parameter("data", m_info.d_cells, +ParamFlags::NonDifferentiable);
    }

    // An optimiser step replaces m_data by a new array (Dr.Jit arrays are immutable values): its
    // contents go back into the table the kernels read, then the sampler is rebuilt from it.
    void parameters_changed(const std::vector<std::string>& keys = {}) override {
        if (!keys.empty() && std::find(keys.begin(), keys.end(), "data") == keys.end()) return;
        if constexpr (dr::is_cuda_v<Float>) {
            if (dr::width(m_data) != (size_t) (4 * m_info.ncell))
                Throw("%s: parameter \"data\" must keep its %zu entries (got %zu)", m_plugin,
                      (size_t) (4 * m_info.ncell), dr::width(m_data));
            dr::eval(m_data);
            check(merl_table_update_cells(m_table.get(), dr::detach(m_data).data(), jit_cuda_stream()),
                  "updating the table");
        } else {
            check(merl_table_refresh_sampler(m_table.get(), jit_cuda_stream()), "rebuilding the sampler");
        }
    }

    const Float& data() const { return m_data; }
    Float& data() { return m_data; }

    // grad_cells4: device float4[ncell] (zeroed by the caller) += d eval / d data . grad_out.
    void backward_table(const SurfaceInteraction3f& si, const Vector3f& wo_, const Spectrum& grad_out_,
                        Mask active_, float* grad_cells4, bool warp_aggregate = false) const {
        backward_table(si.wi, wo_, grad_out_, active_, grad_cells4, warp_aggregate);
    }
    void backward_table(const Vector3f& wi_, const Vector3f& wo_, const Spectrum& grad_out_, Mask active_,
                        float* grad_cells4, bool warp_aggregate = false) const {
        if constexpr (dr::is_cuda_v<Float>) {
            require_wavefront();
            Vector3f wi = wi_, wo = wo_;
            Spectrum go = grad_out_;
            Mask active = active_;
            const size_t n = materialise(wi, wo, active);
            dr::resize(go.x(), n); dr::resize(go.y(), n); dr::resize(go.z(), n);
            dr::eval(go);
            check(merl_backward(m_table.get(), wi.x().data(), wi.y().data(), wi.z().data(), wo.x().data(),
                                wo.y().data(), wo.z().data(), go.x().data(), go.y().data(), go.z().data(),
                                mask_ptr(active), (int64_t) n, grad_cells4,
                                warp_aggregate ? MERL_BWD_WARP_AGGREGATE : 0u, jit_cuda_stream()),
                  "backward");
        } else {
            Throw("%s: the table gradient needs a cuda_ad_* variant", m_plugin);
        }
    }

    // backward_table() leaves the number of queries per cell in the .w lane; a gradient of `data`
    // must be zero there (that lane of the table holds validity bits).
    void clear_hit_counts(float* grad_cells4) const {
        check(merl_grad_clear_counts(m_table.get(), grad_cells4, jit_cuda_stream()), "clearing the hit counts");
    }

    std::string to_string() const override {
        std::ostringstream oss;
        oss << this->class_name() << "[" << std::endl
            << "  filename = \"" << m_filename << "\"," << std::endl
            << "  resolution = " << m_info.nh << "x" << m_info.nd << "x" << m_info.np << "," << std::endl
            << "  table_bytes = " << m_info.table_bytes << "," << std::endl
            << "  spec_weight = " << m_info.spec_weight << std::endl
            << "]";
        return oss.str();
    }

    const merl_table_info_t& table_info() const { return m_info; }
    const merl_table_t* table() const { return m_table.get(); }
    const UInt32& last_sampled_lobe() const { return m_last_lobe; }   // diagnostic (parity tests)

protected:
    void check(int rc, const char* what) const {
        merl_host::check(rc, m_plugin, what, [](const std::string& m) { Throw("%s", m.c_str()); });
    }

    // Broadcast width-1 operands, then force evaluation so that .data() is a valid device pointer.
    static size_t materialise(Vector3f& wi, Vector3f& wo, Mask& active) {
        const size_t n = dr::width(wi.x(), wi.y(), wi.z(), wo.x(), wo.y(), wo.z(), active);
        dr::resize(wi.x(), n); dr::resize(wi.y(), n); dr::resize(wi.z(), n);
        dr::resize(wo.x(), n); dr::resize(wo.y(), n); dr::resize(wo.z(), n);
        dr::resize(active, n);
        dr::eval(wi, wo, active);
        // llvm_*: evaluation is asynchronous on Dr.Jit's thread pool and the packets are handed to the library
        // as plain host memory next; CUDA variants need no wait (same stream as the library's launches)
        if constexpr (dr::is_llvm_v<Float>) dr::sync_thread();
        return n;
    }
    // Dr.Jit CUDA masks store one byte per lane, which is what the ABI's `active` expects.
    static const uint8_t* mask_ptr(const Mask& m) { return reinterpret_cast<const uint8_t*>(m.data()); }

    // JIT variants: refuse to run inside a symbolic recording (see the header comment).
    void require_wavefront() const {
        if constexpr (dr::is_jit_v<Float>) {
            if (jit_flag(MERL_JIT_RECORDING_FLAG))
                Throw("%s: this BSDF calls an external CUDA kernel and needs evaluated arrays, but Dr.Jit is recording "
                      "this call symbolically.  Render in wavefront mode: switch off JitFlag::VCallRecord and "
                      "JitFlag::LoopRecord (Dr.Jit 1.x: SymbolicCalls and SymbolicLoops) before rendering.", m_plugin);
        }
    }
    // scalar_* variants pay one PCIe round trip per query: say so once per process.
    void warn_scalar_once() const {
        static bool warned = false;
        if (!warned) {
            warned = true;
            Log(Warn, "%s: scalar variant -- every eval/pdf/sample call is one host->GPU->host round trip "
                      "(~1e4-1e5 queries/s).  Use a cuda_* variant (device arrays, ~1e11 queries/s) or an llvm_* "
                      "variant (whole packets per call); scalar variants are for parity checks.", m_plugin);
        }
    }

    // non-CUDA variants.  llvm_*: the whole packet in one call of the host-buffer entry point;
    // scalar_*: one query.
    std::pair<Spectrum, Float> eval_pdf_packet(const Vector3f& wi_, const Vector3f& wo_, Mask active) const {
        if constexpr (dr::is_llvm_v<Float>) {
            require_wavefront();
            Vector3f wi = wi_, wo = wo_;
            const size_t n = materialise(wi, wo, active);
            Float r = dr::empty<Float>(n), g = dr::empty<Float>(n), b = dr::empty<Float>(n), p = dr::empty<Float>(n);
            check(merl_eval_pdf_host(m_table.get(), wi.x().data(), wi.y().data(), wi.z().data(), wo.x().data(),
                                     wo.y().data(), wo.z().data(), (int64_t) n, r.data(), g.data(), b.data(), p.data()),
                  "eval_pdf");
            const Float zero(0.f);
            return { Spectrum(dr::select(active, r, zero), dr::select(active, g, zero), dr::select(active, b, zero)),
                     dr::select(active, p, zero) };
        } else {
        warn_scalar_once();
        const float wi[3] = { (float) wi_.x(), (float) wi_.y(), (float) wi_.z() };
        const float wo[3] = { (float) wo_.x(), (float) wo_.y(), (float) wo_.z() };
        float c[3] = {}, p = 0.f;
        if ((bool) active)
            check(merl_eval_pdf_host(m_table.get(), &wi[0], &wi[1], &wi[2], &wo[0], &wo[1], &wo[2], 1, &c[0],
                                     &c[1], &c[2], &p),
                  "eval_pdf");
        return { Spectrum(c[0], c[1], c[2]), Float(p) };
        }
    }

    const char* m_plugin;
    std::string m_filename;
    int m_device = 0;
    merl_host::Table m_table;
    merl_table_info_t m_info{};
    Float m_data;                    // cuda variants: view of the table (or the optimiser's replacement)
    mutable UInt32 m_last_lobe;
};

NAMESPACE_END(mitsuba)

#include "merl_eval_op.h"
