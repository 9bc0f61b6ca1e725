// m06_host_test.cpp -- drives plugins/mitsuba06/{merl,customized_measurment}.cpp the way a Mitsuba
// 0.6 integrator would (BSDFSamplingRecord, EMeasure, Point2 sample), built against the mock
// headers under plugins/mock/mitsuba06 and the real libmerl_b200.  tests/test_gpu_plugins.py feeds
// it seeded queries and compares what it writes with the CPU oracle.
//
//   m06_host_test <plugin> <table.binary> <queries.bin> <out.bin> [key=value ...]
//   queries.bin: int64 n, float32 wi[3][n], wo[3][n], u1[n] (unused), u2[2][n], uint8 active[n]
//   out.bin    : wavefront stage, all n:   float32 evalpdf[n][4], s_wo[n][4], s_weight[n][4]
//                scalar virtuals, first m = min(n, 64): float32 rows[m][15]
//                  = eval rgb, pdf, sample wo, sample pdf, sample weight rgb, sampledType, sampledComponent,
//                    eval(EDiscrete) r, eval of the unserialised clone r
#include "../mitsuba06/merl.cpp"
#include "../mitsuba06/customized_measurment.cpp"

#include <cstdlib>
#include <iostream>
#include <memory>

using namespace mitsuba;

template <typename T> static std::vector<T> read_vec(std::ifstream& f, size_t n) {
    std::vector<T> v(n);
    f.read(reinterpret_cast<char*>(v.data()), (std::streamsize) (n * sizeof(T)));
    if (!f) throw std::runtime_error("short read");
    return v;
}

static std::unique_ptr<MeasuredBSDF> make(const std::string& kind, const Properties& p) {
    if (kind == "merl") return std::unique_ptr<MeasuredBSDF>(new MERL(p));
    if (kind == "customized_measurment") return std::unique_ptr<MeasuredBSDF>(new CustomizedMeasurment(p));
    Log(EError, "unknown plugin \"%s\"", kind.c_str());
    return nullptr;
}

int main(int argc, char** argv) {
    if (argc < 5) { std::cerr << "usage: m06_host_test plugin table queries out [key=value ...]\n"; return 2; }
    try {
        const std::string kind = argv[1];
        Properties props;
        props.setString("filename", argv[2]);
        for (int i = 5; i < argc; ++i) {
            const std::string kv = argv[i];
            const size_t eq = kv.find('=');
            if (eq == std::string::npos) Log(EError, "bad parameter \"%s\"", kv.c_str());
            const std::string k = kv.substr(0, eq), v = kv.substr(eq + 1);
            char* end = nullptr;
            const double d = std::strtod(v.c_str(), &end);
            if (end && *end == 0 && !v.empty()) props.setFloat(k, d); else props.setString(k, v);
        }
        std::ifstream fq(argv[3], std::ios::binary);
        int64_t n = 0;
        fq.read(reinterpret_cast<char*>(&n), 8);
        if (!fq || n < 0) Log(EError, "cannot read %s", argv[3]);
        const auto wi = read_vec<float>(fq, 3 * n), wo = read_vec<float>(fq, 3 * n), u1 = read_vec<float>(fq, n),
                   u2 = read_vec<float>(fq, 2 * n);
        const auto act = read_vec<uint8_t>(fq, n);
        (void) u1;
        std::ofstream fo(argv[4], std::ios::binary);

        auto bsdf = make(kind, props);
        bsdf->configure();                      // the scene loader's job upstream
        std::cout << bsdf->toString() << std::endl;
        Intersection its;

        {   // ---- wavefront stage: one launch per bounce-sized batch -----------------------------
            merl_host::WavefrontStage stage((size_t) (n ? n : 1));
            for (int64_t i = 0; i < n; ++i) {
                BSDFSamplingRecord bRec(its, Vector(wi[i], wi[n + i], wi[2 * n + i]), Vector(wo[i], wo[n + i], wo[2 * n + i]));
                if (!act[i]) bRec.typeMask = BSDF::EDiffuseReflection;      // component switched off
                bsdf->push(stage, bRec);
            }
            bsdf->evalBatch(stage);
            stage.fetch();
            for (int64_t i = 0; i < n; ++i) fo.write(reinterpret_cast<const char*>(&stage.result(i)), 16);
            stage.clear();
            for (int64_t i = 0; i < n; ++i) {
                BSDFSamplingRecord bRec(its, Vector(wi[i], wi[n + i], wi[2 * n + i]));
                if (!act[i]) bRec.typeMask = BSDF::EDiffuseReflection;
                bsdf->pushSample(stage, bRec, Point2(u2[i], u2[n + i]));
            }
            bsdf->sampleBatch(stage);
            stage.fetch();
            for (int64_t i = 0; i < n; ++i) fo.write(reinterpret_cast<const char*>(&stage.wo(i)), 16);
            for (int64_t i = 0; i < n; ++i) fo.write(reinterpret_cast<const char*>(&stage.weight(i)), 16);
        }
        {   // ---- scalar virtuals, and a clone rebuilt from serialize() --------------------------
            Stream stream;
            InstanceManager manager;
            bsdf->serialize(&stream, &manager);
            stream.seek(0);
            std::unique_ptr<MeasuredBSDF> clone(kind == "merl" ? (MeasuredBSDF*) new MERL(&stream, &manager)
                                                               : (MeasuredBSDF*) new CustomizedMeasurment(&stream, &manager));
            const int64_t m = n < 64 ? n : 64;
            for (int64_t i = 0; i < m; ++i) {
                BSDFSamplingRecord bRec(its, Vector(wi[i], wi[n + i], wi[2 * n + i]), Vector(wo[i], wo[n + i], wo[2 * n + i]));
                if (!act[i]) bRec.typeMask = BSDF::EDiffuseReflection;
                float row[15] = {};
                bsdf->eval(bRec, ESolidAngle).toLinearRGB(row[0], row[1], row[2]);
                row[3] = bsdf->pdf(bRec, ESolidAngle);
                BSDFSamplingRecord sRec(its, bRec.wi);
                sRec.typeMask = bRec.typeMask;
                Float pdf;
                const Spectrum w = bsdf->sample(sRec, pdf, Point2(u2[i], u2[n + i]));
                row[4] = sRec.wo.x; row[5] = sRec.wo.y; row[6] = sRec.wo.z; row[7] = pdf;
                w.toLinearRGB(row[8], row[9], row[10]);
                row[11] = (float) sRec.sampledType; row[12] = (float) sRec.sampledComponent;
                row[13] = bsdf->eval(bRec, EDiscrete)[0];
                row[14] = clone->eval(bRec, ESolidAngle)[0];
                fo.write(reinterpret_cast<const char*>(row), sizeof row);
            }
        }
        std::cout << "m06_host_test ok: " << n << " queries, " << merl_launch_count() << " kernel launches" << std::endl;
        return 0;
    } catch (const std::exception& e) {
        std::cerr << "m06_host_test: " << e.what() << std::endl;
        return 1;
    }
}
