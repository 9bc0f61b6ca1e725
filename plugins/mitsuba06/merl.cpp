// merl.cpp -- `<bsdf type="merl">` for Mitsuba 0.6, hosted on libmerl_b200.
//
//   <bsdf type="merl"> <string name="filename" value="gold-metallic-paint.binary"/> </bsdf>
//
// Replaces the reference's mitsuba/src/bsdfs/merl.cpp (named by /root/reference/README.md:1;
// source absent).  Fixed MERL layout: 90 x 90 x 180, scales 1/1500, 1.15/1500, 1.66/1500.
#include "measured_bsdf.h"

MTS_NAMESPACE_BEGIN

class MERL : public MeasuredBSDF {
public:
    MERL(const Properties& props) : MeasuredBSDF(props, "merl", merlParams()) {}
    MERL(Stream* stream, InstanceManager* manager) : MeasuredBSDF(stream, manager, "merl") { configure(); }
    MTS_DECLARE_CLASS()
private:
    static merl_table_params_t merlParams() {
        merl_table_params_t p;
        merl_table_params_default(&p);
        return p;
    }
};

MTS_IMPLEMENT_CLASS_S(MERL, false, BSDF)
MTS_EXPORT_PLUGIN(MERL, "MERL measured BRDF")
MTS_NAMESPACE_END
