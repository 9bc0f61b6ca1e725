// merl_eval_op.h -- the table lookup as ONE node of Dr.Jit's AD graph (cuda_ad_* variants).
//
// Replaces: what Dr.Jit derives by itself in the reference plugin, where the lookup is a
// dr::gather on the table array and its adjoint a dr::scatter_reduce(ReduceOp::Add) (SURVEY.md
// row A10).  Here the lookup is an external kernel, so the pair is declared by hand:
//   eval()      MeasuredBSDF::eval_primal  -> merl_eval      (radiance, cos(theta_o) included)
//   backward()  MeasuredBSDF::backward_table -> merl_backward (scatter-add into float4[ncell])
// Inputs: the plugin instance (not differentiable), `data` = the flat float4 table (the only
// differentiable input), wi, wo, active.  d eval / d wi and d wo are zero almost everywhere
// (nearest-bin lookup; only the cos(theta_o) factor depends on wo, and upstream's gather-based plugin
// propagates that through ordinary arithmetic outside the gather): this operation is the gather's
// stand-in, so like the gather it propagates to `data` only.
// Interface: Dr.Jit 0.4.x `dr::CustomOp<DiffType, Output, Input...>` (the fork's
// Dr.Jit version is unknown); compiled and run here against plugins/mock/mitsuba3/drjit/custom.h.
#pragma once
#include <drjit/custom.h>

NAMESPACE_BEGIN(mitsuba)

template <typename Float, typename Spectrum>
struct MerlEvalOp : dr::CustomOp<Float, Spectrum, const MeasuredBSDF<Float, Spectrum>*, Float,
                                 Vector<Float, 3>, Vector<Float, 3>, dr::mask_t<Float>> {
    using Plugin = MeasuredBSDF<Float, Spectrum>;
    using Vector3f = Vector<Float, 3>;
    using Mask = dr::mask_t<Float>;
    using Base = dr::CustomOp<Float, Spectrum, const Plugin*, Float, Vector3f, Vector3f, Mask>;

    Spectrum eval(const Plugin* const& bsdf, const Float& /* data: read by the kernel through the table handle */,
                  const Vector3f& wi, const Vector3f& wo, const Mask& active) override {
        return bsdf->eval_primal(wi, wo, active);
    }

    void backward() override {
        const Plugin* bsdf = Base::template value_in<0>();
        const size_t entries = dr::width(Base::template value_in<1>());
        Float grad = dr::zeros<Float>(entries);                       // float4[ncell], flat
        dr::eval(grad);
        bsdf->backward_table(Base::template value_in<2>(), Base::template value_in<3>(), Base::grad_out(),
                             Base::template value_in<4>(), grad.data());
        bsdf->clear_hit_counts(grad.data());     // .w of a cell holds validity bits, not a parameter
        Base::template set_grad_in<1>(grad);
    }

    const char* name() const override { return "merl_eval"; }
};

NAMESPACE_END(mitsuba)
