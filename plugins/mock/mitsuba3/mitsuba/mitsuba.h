// MOCK of the Mitsuba 3 names the merl / customized_measurment plugin hosts touch (upstream
// include/mitsuba/{core,render}; SURVEY.md §8b "Host -> plugin, Mitsuba 3").  Signatures and
// member names follow upstream Mitsuba 3.x; bodies are the minimum that lets the plugin sources
// under plugins/mitsuba3/ compile and run against libmerl_b200 in this repository.
#pragma once
#include <drjit/array.h>

#include <cstdarg>
#include <cstdio>
#include <map>
#include <sstream>
#include <string>
#include <utility>
#include <vector>

namespace dr = drjit;

#define NAMESPACE_BEGIN(x) namespace x {
#define NAMESPACE_END(x) }
#define MI_VARIANT template <typename Float, typename Spectrum>
#define MI_USING_1(a) using Base::a;
#define MI_USING_2(a, b) using Base::a; using Base::b;
#define MI_USING_3(a, b, c) using Base::a; using Base::b; using Base::c;
#define MI_USING_PICK(_1, _2, _3, NAME, ...) NAME
#define MI_IMPORT_BASE(Name, ...)                                                                   \
    using Base = mitsuba::Name<Float, Spectrum>;                                                    \
    MI_USING_PICK(__VA_ARGS__, MI_USING_3, MI_USING_2, MI_USING_1)(__VA_ARGS__)
#define MI_IMPORT_TYPES(...)                                                                        \
    using Mask = dr::mask_t<Float>;                                                                 \
    using UInt32 = dr::uint32_array_t<Float>;                                                       \
    using Vector3f = mitsuba::Vector<Float, 3>;                                                     \
    using Point2f = mitsuba::Vector<Float, 2>;                                                      \
    using Color3f = mitsuba::Vector<Float, 3>;                                                      \
    using SurfaceInteraction3f = mitsuba::SurfaceInteraction<Float, Spectrum>;                      \
    using BSDFSample3f = mitsuba::BSDFSample3<Float, Spectrum>;
#define MI_DECLARE_CLASS() const char* class_name() const override;
#define MI_IMPLEMENT_CLASS_VARIANT(Name, Parent, ...)
#define MI_EXPORT_PLUGIN(Name, Descr)
#define MI_MASKED_FUNCTION(...)
#define MI_INLINE inline

NAMESPACE_BEGIN(mitsuba)

[[noreturn]] inline void Throw(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    throw std::runtime_error(buf);
}

// upstream: Log(LogLevel, fmt, ...) routes through the thread's Logger; the mock keeps the lines
enum LogLevel { Trace, Debug, Info, Warn, Error };
inline std::vector<std::string>& mock_log() { static std::vector<std::string> l; return l; }
inline void Log(LogLevel level, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    mock_log().push_back(std::string(level == Warn ? "[WARN] " : "[LOG] ") + buf);
    std::fprintf(stderr, "%s\n", mock_log().back().c_str());
}

template <typename T> class ref {                                   // intrusive in upstream; plain here
public:
    ref(T* p = nullptr) : m_p(p) {}
    T* operator->() const { return m_p.get(); }
    T* get() const { return m_p.get(); }
    explicit operator bool() const { return (bool) m_p; }
private:
    std::shared_ptr<T> m_p;
};

class Object {
public:
    virtual ~Object() = default;
    virtual std::string to_string() const { return "Object[]"; }
};

template <typename Value, size_t N> struct Vector {
    Value v[N];
    Vector() = default;
    Vector(const Value& a, const Value& b) : v{ a, b } {}
    Vector(const Value& a, const Value& b, const Value& c) : v{ a, b, c } {}
    Vector(float s) { for (auto& e : v) e = Value(s); }
    Value& x() { return v[0]; } const Value& x() const { return v[0]; }
    Value& y() { return v[1]; } const Value& y() const { return v[1]; }
    Value& z() { return v[2]; } const Value& z() const { return v[2]; }
    Value& r() { return v[0]; } const Value& r() const { return v[0]; }
    Value& g() { return v[1]; } const Value& g() const { return v[1]; }
    Value& b() { return v[2]; } const Value& b() const { return v[2]; }
};

enum class BSDFFlags : uint32_t {
    Empty = 0x0, Null = 0x1, DiffuseReflection = 0x2, DiffuseTransmission = 0x4, GlossyReflection = 0x8,
    GlossyTransmission = 0x10, DeltaReflection = 0x20, DeltaTransmission = 0x40, FrontSide = 0x10000,
    BackSide = 0x20000, All = 0x1ff
};
constexpr uint32_t operator|(BSDFFlags a, BSDFFlags b) { return (uint32_t) a | (uint32_t) b; }
constexpr uint32_t operator+(BSDFFlags a) { return (uint32_t) a; }
enum class TransportMode : uint32_t { Radiance, Importance };
enum class ParamFlags : uint32_t { Differentiable = 0x0, NonDifferentiable = 0x1, Discontinuous = 0x2 };
constexpr uint32_t operator+(ParamFlags a) { return (uint32_t) a; }

struct BSDFContext {
    TransportMode mode = TransportMode::Radiance;
    uint32_t type_mask = (uint32_t) 0x1ff;
    uint32_t component = (uint32_t) -1;
    bool is_enabled(BSDFFlags type, uint32_t comp = 0) const {
        return (component == (uint32_t) -1 || component == comp) && (type_mask & (uint32_t) type) != 0;
    }
};

template <typename Float, typename Spectrum> struct SurfaceInteraction { Vector<Float, 3> wi; };

template <typename Float, typename Spectrum> struct BSDFSample3 {
    Vector<Float, 3> wo;
    Float pdf;
    Float eta;
    dr::uint32_array_t<Float> sampled_type;
    dr::uint32_array_t<Float> sampled_component;
};

class Properties {
public:
    void set_string(const std::string& k, const std::string& v) { m_s[k] = v; }
    void set_float(const std::string& k, double v) { m_f[k] = v; }
    bool has_property(const std::string& k) const { return m_s.count(k) || m_f.count(k); }
    std::string string(const std::string& k) const {
        auto it = m_s.find(k);
        if (it == m_s.end()) Throw("Property \"%s\" has not been specified!", k.c_str());
        return it->second;
    }
    std::string string(const std::string& k, const std::string& d) const { return m_s.count(k) ? m_s.at(k) : d; }
    template <typename T> T get(const std::string& k, T d) const { return m_f.count(k) ? (T) m_f.at(k) : d; }
    std::string id() const { return string("id", "unnamed"); }
private:
    std::map<std::string, std::string> m_s;
    std::map<std::string, double> m_f;
};

struct FileResolver {                                               // upstream: Thread::thread()->file_resolver()
    std::string resolve(const std::string& p) const { return p; }
};
struct Thread {
    static Thread* thread() { static Thread t; return &t; }
    const FileResolver* file_resolver() const { static FileResolver f; return &f; }
};

struct TraversalCallback {
    virtual ~TraversalCallback() = default;
    virtual void put_parameter_impl(const std::string& name, void* ptr, uint32_t flags) = 0;
    template <typename T> void put_parameter(const std::string& name, T& v, uint32_t flags) {
        put_parameter_impl(name, &v, flags);
    }
};

template <typename Float_, typename Spectrum_> class BSDF : public Object {
public:
    using Float = Float_;
    using Spectrum = Spectrum_;
    MI_IMPORT_TYPES()
    explicit BSDF(const Properties& props) : m_id(props.id()) {}
    virtual std::pair<BSDFSample3f, Spectrum> sample(const BSDFContext& ctx, const SurfaceInteraction3f& si,
                                                     Float sample1, const Point2f& sample2,
                                                     Mask active = true) const = 0;
    virtual Spectrum eval(const BSDFContext& ctx, const SurfaceInteraction3f& si, const Vector3f& wo,
                          Mask active = true) const = 0;
    virtual Float pdf(const BSDFContext& ctx, const SurfaceInteraction3f& si, const Vector3f& wo,
                      Mask active = true) const = 0;
    virtual std::pair<Spectrum, Float> eval_pdf(const BSDFContext& ctx, const SurfaceInteraction3f& si,
                                                const Vector3f& wo, Mask active = true) const {
        return { eval(ctx, si, wo, active), pdf(ctx, si, wo, active) };
    }
    virtual void traverse(TraversalCallback*) {}
    virtual void parameters_changed(const std::vector<std::string>& = {}) {}
    virtual const char* class_name() const = 0;
    uint32_t flags() const { return m_flags; }
    size_t component_count() const { return m_components.size(); }
protected:
    uint32_t m_flags = 0;
    std::vector<uint32_t> m_components;
    std::string m_id;
};

NAMESPACE_END(mitsuba)
