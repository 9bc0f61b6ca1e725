// customized_measurment.cpp -- `<bsdf type="customized_measurment">` for Mitsuba 3, hosted on
// libmerl_b200 (plugin name spelled as in /root/reference/README.md:1).
//
// The fork's format for this plugin is unknown.  Assumption
// built here: a MERL-layout file (3 x int32 header, planar doubles) whose resolution comes from
// the header unless `resolution_theta_h/_theta_d/_phi_d` are given, per-channel `scale_r/g/b`
// or one `scale` (default: the MERL channel scales), `theta_h_mapping` = "sqrt" (MERL's nonlinear bins) or "linear", `clamp_negative`.
//
//   <bsdf type="customized_measurment">
//       <string name="filename" value="sample_17.binary"/>
//       <string name="theta_h_mapping" value="linear"/>
//       <float name="scale_r" value="1"/> ...
//   </bsdf>
#include "measured_bsdf.h"

NAMESPACE_BEGIN(mitsuba)

template <typename Float, typename Spectrum>
class CustomizedMeasurment final : public MeasuredBSDF<Float, Spectrum> {
public:
    using Base = MeasuredBSDF<Float, Spectrum>;
    explicit CustomizedMeasurment(const Properties& props)
        : Base(props, "customized_measurment", params_from(props)) {}
    MI_DECLARE_CLASS()
private:
    static merl_table_params_t params_from(const Properties& props) {
        merl_table_params_t p;
        merl_table_params_default(&p);
        p.nh = props.get<int>("resolution_theta_h", 0);     // 0 = take the file header's
        p.nd = props.get<int>("resolution_theta_d", 0);
        p.np = props.get<int>("resolution_phi_d", 0);
        if ((p.nh == 0) != (p.nd == 0) || (p.nd == 0) != (p.np == 0))
            Throw("customized_measurment: give all three resolution_* parameters or none");
        if (props.has_property("scale"))                      // one factor for all three channels
            p.scale[0] = p.scale[1] = p.scale[2] = props.get<double>("scale", 1.0);
        p.scale[0] = props.get<double>("scale_r", p.scale[0]);  // default: the MERL channel scales
        p.scale[1] = props.get<double>("scale_g", p.scale[1]);
        p.scale[2] = props.get<double>("scale_b", p.scale[2]);
        const std::string mapping = props.string("theta_h_mapping", "sqrt");
        if (mapping != "sqrt" && mapping != "linear")
            Throw("customized_measurment: theta_h_mapping must be \"sqrt\" or \"linear\"");
        p.theta_h_sqrt = mapping == "sqrt";
        p.clamp_negative = props.get<bool>("clamp_negative", true) ? 1 : 0;
        return p;
    }
};

template <typename Float, typename Spectrum>
const char* CustomizedMeasurment<Float, Spectrum>::class_name() const { return "CustomizedMeasurment"; }

MI_IMPLEMENT_CLASS_VARIANT(CustomizedMeasurment, BSDF)
MI_EXPORT_PLUGIN(CustomizedMeasurment, "Customized measured BRDF")
NAMESPACE_END(mitsuba)
