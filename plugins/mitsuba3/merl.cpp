// merl.cpp -- `<bsdf type="merl">` for Mitsuba 3, hosted on libmerl_b200.
//
//   <bsdf type="merl">
//       <string name="filename" value="gold-metallic-paint.binary"/>
//   </bsdf>
//
// Replaces the reference's mitsuba3/src/bsdfs/merl.cpp (named by /root/reference/README.md:1;
// source absent).  The MERL layout is fixed: 90 x 90 x 180 cells, planar doubles, channel
// scales 1/1500, 1.15/1500, 1.66/1500 (public BRDFRead.cpp).
#include "measured_bsdf.h"

NAMESPACE_BEGIN(mitsuba)

template <typename Float, typename Spectrum> class MERL final : public MeasuredBSDF<Float, Spectrum> {
public:
    using Base = MeasuredBSDF<Float, Spectrum>;
    explicit MERL(const Properties& props) : Base(props, "merl", merl_params()) {}
    MI_DECLARE_CLASS()
private:
    static merl_table_params_t merl_params() {
        merl_table_params_t p;
        merl_table_params_default(&p);   // 90/90/180, sqrt theta_h bins, MERL scales, negatives -> 0
        return p;
    }
};

template <typename Float, typename Spectrum> const char* MERL<Float, Spectrum>::class_name() const { return "MERL"; }

MI_IMPLEMENT_CLASS_VARIANT(MERL, BSDF)
MI_EXPORT_PLUGIN(MERL, "MERL measured BRDF")
NAMESPACE_END(mitsuba)
