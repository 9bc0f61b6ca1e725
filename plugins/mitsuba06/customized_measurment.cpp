// customized_measurment.cpp -- `<bsdf type="customized_measurment">` for Mitsuba 0.6, hosted on
// libmerl_b200 (spelling per /root/reference/README.md:1).  Format assumption B7 (SURVEY.md
// Appendix B): MERL-layout file, resolution from the header unless `thetaHRes/thetaDRes/phiDRes`
// are given, `scale` or `scaleR/G/B`, `thetaHMapping` = "sqrt" | "linear", `clampNegative`.
#include "measured_bsdf.h"

MTS_NAMESPACE_BEGIN

class CustomizedMeasurment : public MeasuredBSDF {
public:
    CustomizedMeasurment(const Properties& props) : MeasuredBSDF(props, "customized_measurment", paramsFrom(props)) {}
    CustomizedMeasurment(Stream* stream, InstanceManager* manager)
        : MeasuredBSDF(stream, manager, "customized_measurment") { configure(); }
    MTS_DECLARE_CLASS()
private:
    static merl_table_params_t paramsFrom(const Properties& props) {
        merl_table_params_t p;
        merl_table_params_default(&p);
        p.nh = props.getInteger("thetaHRes", 0);            // 0 = take the file header's
        p.nd = props.getInteger("thetaDRes", 0);
        p.np = props.getInteger("phiDRes", 0);
        if ((p.nh == 0) != (p.nd == 0) || (p.nd == 0) != (p.np == 0))
            Log(EError, "customized_measurment: give thetaHRes, thetaDRes and phiDRes together or not at all");
        if (props.hasProperty("scale")) p.scale[0] = p.scale[1] = p.scale[2] = props.getFloat("scale", 1.0f);
        p.scale[0] = props.hasProperty("scaleR") ? props.getFloat("scaleR", 1.0f) : p.scale[0];
        p.scale[1] = props.hasProperty("scaleG") ? props.getFloat("scaleG", 1.0f) : p.scale[1];
        p.scale[2] = props.hasProperty("scaleB") ? props.getFloat("scaleB", 1.0f) : p.scale[2];
        const std::string mapping = props.getString("thetaHMapping", "sqrt");
        if (mapping != "sqrt" && mapping != "linear")
            Log(EError, "customized_measurment: thetaHMapping must be \"sqrt\" or \"linear\"");
        p.theta_h_sqrt = mapping == "sqrt";
        p.clamp_negative = props.getBoolean("clampNegative", true) ? 1 : 0;
        return p;
    }
};

MTS_IMPLEMENT_CLASS_S(CustomizedMeasurment, false, BSDF)
MTS_EXPORT_PLUGIN(CustomizedMeasurment, "Customized measured BRDF")
MTS_NAMESPACE_END
