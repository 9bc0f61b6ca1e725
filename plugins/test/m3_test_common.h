// m3_test_common.h -- what the Mitsuba 3 host test drivers share: the plugin sources (compiled
// into the driver, as a unity build), file helpers, plugin construction from key=value arguments.
#pragma once
#include "../mitsuba3/merl.cpp"
#include "../mitsuba3/customized_measurment.cpp"

#include <cstdlib>
#include <fstream>
#include <iostream>

using namespace mitsuba;
using FloatC = dr::CUDAArray<float>;
using SpectrumC = Vector<FloatC, 3>;
using SpectrumS = Vector<float, 3>;

template <typename T> static std::vector<T> read_vec(std::ifstream& f, size_t n) {
    std::vector<T> v(n);
    f.read(reinterpret_cast<char*>(v.data()), (std::streamsize) (n * sizeof(T)));
    if (!f) throw std::runtime_error("short read");
    return v;
}
template <typename T> static void write_vec(std::ofstream& f, const std::vector<T>& v) {
    f.write(reinterpret_cast<const char*>(v.data()), (std::streamsize) (v.size() * sizeof(T)));
}
static std::vector<float> to_host(const FloatC& a) {
    std::vector<float> h(a.size());
    dr::store(h.data(), a);
    return h;
}

template <typename F, typename S> static std::unique_ptr<BSDF<F, S>> make(const std::string& kind, const Properties& p) {
    if (kind == "merl") return std::make_unique<MERL<F, S>>(p);
    if (kind == "customized_measurment") return std::make_unique<CustomizedMeasurment<F, S>>(p);
    Throw("unknown plugin \"%s\"", kind.c_str());
}


static Properties parse_props(const char* table, int argc, char** argv, int first) {
    Properties props;
    props.set_string("filename", table);
    for (int i = first; i < argc; ++i) {
        const std::string kv = argv[i];
        const size_t eq = kv.find('=');
        if (eq == std::string::npos) Throw("bad parameter \"%s\"", kv.c_str());
        const std::string k = kv.substr(0, eq), v = kv.substr(eq + 1);
        char* end = nullptr;
        const double d = std::strtod(v.c_str(), &end);
        if (end && *end == 0 && !v.empty()) props.set_float(k, d); else props.set_string(k, v);
    }
    return props;
}

// queries.bin: int64 n, then float32 wi[3][n], wo[3][n], u1[n], u2[2][n], uint8 active[n]
struct Queries {
    int64_t n = 0;
    std::vector<float> wi, wo, u1, u2;
    std::vector<uint8_t> act;
    explicit Queries(const char* path) {
        std::ifstream fq(path, std::ios::binary);
        fq.read(reinterpret_cast<char*>(&n), 8);
        if (!fq || n < 0) Throw("cannot read %s", path);
        wi = read_vec<float>(fq, 3 * n); wo = read_vec<float>(fq, 3 * n); u1 = read_vec<float>(fq, n);
        u2 = read_vec<float>(fq, 2 * n);
        act = read_vec<uint8_t>(fq, n);
    }
};
