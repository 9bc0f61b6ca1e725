// m3_host_test.cpp -- drives plugins/mitsuba3/{merl,customized_measurment}.cpp the way a Mitsuba 3
// integrator would (BSDFContext, SurfaceInteraction3f, wo, Mask), built against the mock headers
// under plugins/mock/mitsuba3 and the real libmerl_b200.  tests/test_gpu_plugins.py feeds it seeded
// queries and compares what it writes with the CPU oracle.
//
//   m3_host_test <plugin: merl|customized_measurment> <table.binary> <queries.bin> <out.bin> [key=value ...]
//   queries.bin: int64 n, then float32 wi[3][n], wo[3][n], u1[n], u2[2][n], uint8 active[n]
//   out.bin    : float32 rgb[3][n], pdf[n], eval_only[3][n], pdf_only[n],
//                s_wo[3][n], s_pdf[n], s_weight[3][n], uint32 s_lobe[n]      (cuda variant, all n)
//                then the same block for the first min(n, 64) queries through the scalar variant
#include "m3_test_common.h"

int main(int argc, char** argv) {
    if (argc < 5) { std::cerr << "usage: m3_host_test plugin table queries out [key=value ...]\n"; return 2; }
    try {
        const std::string kind = argv[1];
        const Properties props = parse_props(argv[2], argc, argv, 5);
        const Queries Q(argv[3]);
        const int64_t n = Q.n;
        const auto &wi = Q.wi, &wo = Q.wo, &u1 = Q.u1, &u2 = Q.u2;
        const auto& act = Q.act;
        std::ofstream fo(argv[4], std::ios::binary);

        BSDFContext ctx;
        {   // ---- cuda_rgb-like variant: whole batch, device arrays ------------------------------
            auto bsdf = make<FloatC, SpectrumC>(kind, props);
            std::cout << bsdf->to_string() << std::endl;
            SurfaceInteraction<FloatC, SpectrumC> si;
            Vector<FloatC, 3> wod;
            for (int c = 0; c < 3; ++c) {
                si.wi.v[c] = dr::load<FloatC>(wi.data() + c * n, n);
                wod.v[c] = dr::load<FloatC>(wo.data() + c * n, n);
            }
            auto active = dr::load<dr::CUDAArray<bool>>(act.data(), n);
            auto [rgb, pdf] = bsdf->eval_pdf(ctx, si, wod, active);
            auto rgb_only = bsdf->eval(ctx, si, wod, active);
            auto pdf_only = bsdf->pdf(ctx, si, wod, active);
            Vector<FloatC, 2> s2(dr::load<FloatC>(u2.data(), n), dr::load<FloatC>(u2.data() + n, n));
            auto [bs, weight] = bsdf->sample(ctx, si, dr::load<FloatC>(u1.data(), n), s2, active);
            dr::sync_thread();
            for (int c = 0; c < 3; ++c) write_vec(fo, to_host(rgb.v[c]));
            write_vec(fo, to_host(pdf));
            for (int c = 0; c < 3; ++c) write_vec(fo, to_host(rgb_only.v[c]));
            write_vec(fo, to_host(pdf_only));
            for (int c = 0; c < 3; ++c) write_vec(fo, to_host(bs.wo.v[c]));
            write_vec(fo, to_host(bs.pdf));
            for (int c = 0; c < 3; ++c) write_vec(fo, to_host(weight.v[c]));
            std::vector<uint32_t> lobe(n);
            dr::store(lobe.data(), static_cast<MeasuredBSDF<FloatC, SpectrumC>*>(bsdf.get())->last_sampled_lobe());
            write_vec(fo, lobe);
            // a disabled component yields zeros without touching the device
            BSDFContext off;
            off.type_mask = +BSDFFlags::DiffuseReflection;
            auto z = bsdf->eval(off, si, wod, active);
            if (dr::width(z.x()) != 1) Throw("disabled component should return a literal zero");
        }
        {   // ---- scalar_rgb-like variant: one query per call ------------------------------------
            auto bsdf = make<float, SpectrumS>(kind, props);
            const int64_t m = n < 64 ? n : 64;
            std::vector<float> out(15 * m);
            std::vector<uint32_t> lobe(m);
            for (int64_t i = 0; i < m; ++i) {
                SurfaceInteraction<float, SpectrumS> si;
                si.wi = Vector<float, 3>(wi[i], wi[n + i], wi[2 * n + i]);
                const Vector<float, 3> w(wo[i], wo[n + i], wo[2 * n + i]);
                auto [rgb, pdf] = bsdf->eval_pdf(ctx, si, w, act[i] != 0);
                auto rgb_only = bsdf->eval(ctx, si, w, act[i] != 0);
                const float pdf_only = bsdf->pdf(ctx, si, w, act[i] != 0);
                auto [bs, weight] = bsdf->sample(ctx, si, u1[i], Vector<float, 2>(u2[i], u2[n + i]), act[i] != 0);
                const float row[15] = { rgb.x(), rgb.y(), rgb.z(), pdf, rgb_only.x(), rgb_only.y(), rgb_only.z(),
                                        pdf_only, bs.wo.x(), bs.wo.y(), bs.wo.z(), bs.pdf, weight.x(), weight.y(),
                                        weight.z() };
                for (int k = 0; k < 15; ++k) out[k * m + i] = row[k];
                lobe[i] = static_cast<MeasuredBSDF<float, SpectrumS>*>(bsdf.get())->last_sampled_lobe();
            }
            write_vec(fo, out);
            write_vec(fo, lobe);
        }
        {   // ---- llvm_rgb-like variant: whole packet in host memory, one call per virtual --------
            using FloatL = dr::LLVMArray<float>;
            using SpectrumL = Vector<FloatL, 3>;
            auto bsdf = make<FloatL, SpectrumL>(kind, props);
            SurfaceInteraction<FloatL, SpectrumL> si;
            Vector<FloatL, 3> wod;
            for (int c = 0; c < 3; ++c) {
                si.wi.v[c] = dr::load<FloatL>(wi.data() + c * n, n);
                wod.v[c] = dr::load<FloatL>(wo.data() + c * n, n);
            }
            auto active = dr::load<dr::LLVMArray<bool>>(act.data(), n);
            const int64_t calls0 = merl_launch_count();
            const int syncs0 = dr::mock_sync_thread_calls();
            auto [rgb, pdf] = bsdf->eval_pdf(ctx, si, wod, active);
            auto rgb_only = bsdf->eval(ctx, si, wod, active);
            auto pdf_only = bsdf->pdf(ctx, si, wod, active);
            Vector<FloatL, 2> s2(dr::load<FloatL>(u2.data(), n), dr::load<FloatL>(u2.data() + n, n));
            auto [bs, weight] = bsdf->sample(ctx, si, dr::load<FloatL>(u1.data(), n), s2, active);
            // every virtual waits for Dr.Jit's asynchronous LLVM evaluation before handing host pointers over
            // (the mock's data() raises otherwise; this counts the waits)
            if (dr::mock_sync_thread_calls() - syncs0 < 4) Throw("llvm variant: fewer than one dr::sync_thread() per virtual");
            std::ofstream fl(std::string(argv[4]) + ".llvm", std::ios::binary);
            auto host = [](const FloatL& a) { std::vector<float> h(a.size()); dr::store(h.data(), a); return h; };
            for (int c = 0; c < 3; ++c) write_vec(fl, host(rgb.v[c]));
            write_vec(fl, host(pdf));
            for (int c = 0; c < 3; ++c) write_vec(fl, host(rgb_only.v[c]));
            write_vec(fl, host(pdf_only));
            for (int c = 0; c < 3; ++c) write_vec(fl, host(bs.wo.v[c]));
            write_vec(fl, host(bs.pdf));
            for (int c = 0; c < 3; ++c) write_vec(fl, host(weight.v[c]));
            std::vector<uint32_t> lobe(n);
            dr::store(lobe.data(), static_cast<MeasuredBSDF<FloatL, SpectrumL>*>(bsdf.get())->last_sampled_lobe());
            write_vec(fl, lobe);
            std::cout << "llvm variant: " << n << " lanes per call, " << (merl_launch_count() - calls0)
                      << " library launches for 4 virtuals" << std::endl;
            // JIT variants refuse to run inside a symbolic recording, with a message that says what to do
            jit_set_flag(JitFlag::Recording, 1);
            bool refused = false;
            try { bsdf->eval(ctx, si, wod, active); } catch (const std::exception& e) {
                refused = std::string(e.what()).find("wavefront") != std::string::npos;
            }
            jit_set_flag(JitFlag::Recording, 0);
            if (!refused) Throw("the llvm variant did not refuse a symbolic recording");
        }
        {   // the same guard on the cuda variant (before any data() / eval() can raise inside Dr.Jit)
            auto bsdf = make<FloatC, SpectrumC>(kind, props);
            SurfaceInteraction<FloatC, SpectrumC> si;
            Vector<FloatC, 3> wod;
            for (int c = 0; c < 3; ++c) { si.wi.v[c] = dr::load<FloatC>(wi.data() + c * n, n); wod.v[c] = dr::load<FloatC>(wo.data() + c * n, n); }
            jit_set_flag(JitFlag::Recording, 1);
            int refused = 0;
            auto expect = [&](auto&& call) {
                try { call(); } catch (const std::exception& e) {
                    refused += std::string(e.what()).find("wavefront") != std::string::npos;
                }
            };
            expect([&] { bsdf->eval(ctx, si, wod, true); });
            expect([&] { bsdf->pdf(ctx, si, wod, true); });
            expect([&] { bsdf->eval_pdf(ctx, si, wod, true); });
            expect([&] { bsdf->sample(ctx, si, FloatC(0.5f), Vector<FloatC, 2>(FloatC(0.5f), FloatC(0.5f)), true); });
            jit_set_flag(JitFlag::Recording, 0);
            if (refused != 4) Throw("the cuda variant refused %d of 4 virtuals inside a recording", refused);
            std::cout << "recording guard ok" << std::endl;
        }
        {
            size_t warns = 0;
            for (const auto& l : mock_log()) warns += l.find("[WARN]") == 0 && l.find("scalar variant") != std::string::npos;
            std::cout << "scalar warning lines: " << warns << std::endl;
        }
        std::cout << "m3_host_test ok: " << n << " queries, " << merl_launch_count() << " kernel launches" << std::endl;
        return 0;
    } catch (const std::exception& e) {
        std::cerr << "m3_host_test: " << e.what() << std::endl;
        return 1;
    }
}
