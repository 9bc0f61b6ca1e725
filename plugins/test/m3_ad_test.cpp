// m3_ad_test.cpp -- the differentiable side of the Mitsuba 3 host (plugins/mitsuba3/merl_eval_op.h,
// MeasuredBSDF::traverse / parameters_changed), driven the way Dr.Jit's AD layer and an optimiser
// would, against the mock headers and the real libmerl_b200:
//   1. eval through dr::custom<MerlEvalOp>() == the plain eval, bit for bit;
//   2. the recorded operation's backward() with grad_out = (u1, u2x, u2y) -> gradient of `data`
//      (flat float4[ncell]); its .w lane must be zero; a direct backward_table() into a second
//      buffer gives the same rgb lanes (and the hit counts in .w);
//   3. an "optimiser step": traverse() exposes `data`; it is REPLACED by a new array holding
//      2 x the cells (validity lane untouched), parameters_changed({"data"}) -> eval doubles exactly.
//
//   m3_ad_test <plugin> <table.binary> <queries.bin> <out.bin> [key=value ...]
//   out.bin: float32 rgb_custom[3][n], grad[4 * ncell], grad_direct[4 * ncell], rgb_after[3][n]
#include "m3_test_common.h"

#include <cmath>

struct Collect : TraversalCallback {
    std::map<std::string, std::pair<void*, uint32_t>> seen;
    void put_parameter_impl(const std::string& name, void* ptr, uint32_t flags) override { seen[name] = { ptr, flags }; }
};

int main(int argc, char** argv) {
    if (argc < 5) { std::cerr << "usage: m3_ad_test plugin table queries out [key=value ...]\n"; return 2; }
    try {
        const std::string kind = argv[1];
        const Properties props = parse_props(argv[2], argc, argv, 5);
        const Queries Q(argv[3]);
        const int64_t n = Q.n;
        std::ofstream fo(argv[4], std::ios::binary);
        using Plugin = MeasuredBSDF<FloatC, SpectrumC>;
        using Op = MerlEvalOp<FloatC, SpectrumC>;

        auto owner = make<FloatC, SpectrumC>(kind, props);
        Plugin* bsdf = static_cast<Plugin*>(owner.get());
        const size_t ncell = (size_t) bsdf->table_info().ncell;
        BSDFContext ctx;
        SurfaceInteraction<FloatC, SpectrumC> si;
        Vector<FloatC, 3> wo;
        for (int c = 0; c < 3; ++c) {
            si.wi.v[c] = dr::load<FloatC>(Q.wi.data() + c * n, n);
            wo.v[c] = dr::load<FloatC>(Q.wo.data() + c * n, n);
        }
        auto active = dr::load<dr::CUDAArray<bool>>(Q.act.data(), n);

        // 1. forward through the custom operation
        const SpectrumC plain = bsdf->eval(ctx, si, wo, active);
        const SpectrumC viaop = dr::custom<Op>((const Plugin*) bsdf, bsdf->data(), si.wi, wo, active);
        dr::sync_thread();
        std::vector<float> rgb0[3];
        for (int c = 0; c < 3; ++c) {
            rgb0[c] = to_host(plain.v[c]);
            const auto h = to_host(viaop.v[c]);
            if (std::memcmp(h.data(), rgb0[c].data(), sizeof(float) * n) != 0) Throw("custom-op eval differs from eval (channel %d)", c);
            write_vec(fo, h);
        }

        // 2. backward of the recorded operation
        if (dr::mock_tape().size() != 1) Throw("expected one recorded operation, found %zu", dr::mock_tape().size());
        auto op = std::dynamic_pointer_cast<Op>(dr::mock_tape().back());
        if (!op || std::string(op->name()) != "merl_eval") Throw("recorded operation is not merl_eval");
        const SpectrumC go(dr::load<FloatC>(Q.u1.data(), n), dr::load<FloatC>(Q.u2.data(), n),
                           dr::load<FloatC>(Q.u2.data() + n, n));
        op->mock_set_grad_out(go);
        op->backward();
        const FloatC& grad = op->template mock_grad_in<1>();
        if (dr::width(grad) != 4 * ncell) Throw("gradient of data has %zu entries, expected %zu", dr::width(grad), 4 * ncell);
        FloatC direct = dr::zeros<FloatC>(4 * ncell);
        bsdf->backward_table(si, wo, go, active, direct.data());
        dr::sync_thread();
        const auto hg = to_host(grad), hd = to_host(direct);
        double hits = 0;
        for (size_t k = 0; k < ncell; ++k) {
            if (hg[4 * k + 3] != 0.f) Throw("gradient of the validity lane is not zero at cell %zu", k);
            hits += hd[4 * k + 3];
            for (int c = 0; c < 3; ++c) {
                const float a = hg[4 * k + c], b = hd[4 * k + c];
                if (std::fabs(a - b) > 1e-6f + 1e-5f * std::fabs(b)) Throw("gradient mismatch at cell %zu channel %d: %g vs %g", k, c, a, b);
            }
        }
        write_vec(fo, hg);
        write_vec(fo, hd);

        // 3. optimiser step: replace `data` by 2 x cells, parameters_changed, eval again
        Collect cb;
        bsdf->traverse(&cb);
        if (!cb.seen.count("data") || cb.seen["data"].second != +ParamFlags::Differentiable)
            Throw("traverse() does not expose a differentiable \"data\"");
        FloatC* param = static_cast<FloatC*>(cb.seen["data"].first);
        auto cells = to_host(*param);
        if (cells.size() != 4 * ncell) Throw("parameter data has %zu entries, expected %zu", cells.size(), 4 * ncell);
        for (size_t k = 0; k < ncell; ++k)
            for (int c = 0; c < 3; ++c) cells[4 * k + c] *= 2.f;
        *param = dr::load<FloatC>(cells.data(), cells.size());     // a NEW array, as Dr.Jit's optimisers produce
        bsdf->parameters_changed({ "data" });
        const SpectrumC after = bsdf->eval(ctx, si, wo, active);
        dr::sync_thread();
        for (int c = 0; c < 3; ++c) {
            const auto h = to_host(after.v[c]);
            for (int64_t i = 0; i < n; ++i)
                if (h[i] != 2.f * rgb0[c][i]) Throw("eval after the step is not 2 x eval at query %lld channel %d: %g vs %g", (long long) i, c, h[i], rgb0[c][i]);
            write_vec(fo, h);
        }
        // a wrong-sized replacement is refused
        *param = dr::zeros<FloatC>(8);
        bool refused = false;
        try { bsdf->parameters_changed({ "data" }); } catch (const std::exception&) { refused = true; }
        if (!refused) Throw("parameters_changed accepted a data array of the wrong size");
        std::cout << "m3_ad_test ok: " << n << " queries, " << (long long) hits << " hits, " << merl_launch_count()
                  << " kernel launches" << std::endl;
        return 0;
    } catch (const std::exception& e) {
        std::cerr << "m3_ad_test: " << e.what() << std::endl;
        return 1;
    }
}
