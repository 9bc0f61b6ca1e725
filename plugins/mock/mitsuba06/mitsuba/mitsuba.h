// MOCK of the Mitsuba 0.6 names the merl / customized_measurment plugin hosts touch (upstream
// include/mitsuba/{core,render}; SURVEY.md §8b "Host -> plugin, Mitsuba 0.6").  Names and
// signatures follow upstream 0.6 (single-precision, RGB build: SPECTRUM_SAMPLES = 3); bodies are
// the minimum that lets plugins/mitsuba06/*.cpp compile and run against libmerl_b200 here.
#pragma once
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#define MTS_NAMESPACE_BEGIN namespace mitsuba {
#define MTS_NAMESPACE_END }
#define MTS_DECLARE_CLASS() virtual const char* getClassName() const;
#define MTS_IMPLEMENT_CLASS_S(Name, abstract, Parent) const char* Name::getClassName() const { return #Name; }
#define MTS_EXPORT_PLUGIN(Name, Descr)

MTS_NAMESPACE_BEGIN

typedef float Float;
enum ELogLevel { ETrace, EDebug, EInfo, EWarn, EError };
inline void Log(ELogLevel lvl, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (lvl == EError) throw std::runtime_error(buf);               // upstream: EError throws
    fprintf(stderr, "%s\n", buf);
}
#define SLog Log

struct Vector { Float x = 0, y = 0, z = 0; Vector() {} Vector(Float a, Float b, Float c) : x(a), y(b), z(c) {} };
struct Point2 { Float x = 0, y = 0; Point2() {} Point2(Float a, Float b) : x(a), y(b) {} };
struct Frame { static Float cosTheta(const Vector& v) { return v.z; } };

struct Spectrum {                                                   // RGB build
    Float s[3];
    Spectrum(Float v = 0) { s[0] = s[1] = s[2] = v; }
    void fromLinearRGB(Float r, Float g, Float b) { s[0] = r; s[1] = g; s[2] = b; }
    void toLinearRGB(Float& r, Float& g, Float& b) const { r = s[0]; g = s[1]; b = s[2]; }
    Float& operator[](int i) { return s[i]; }
    const Float& operator[](int i) const { return s[i]; }
    bool isZero() const { return s[0] == 0 && s[1] == 0 && s[2] == 0; }
};

enum EMeasure { EInvalidMeasure = 0, ESolidAngle = 1, ELength, EArea, EDiscrete };
enum ETransportMode { ERadiance = 0, EImportance = 1 };

struct Intersection {};

class Stream {                                                      // in-memory stand-in for MemoryStream
public:
    void write(const void* p, size_t n) { const char* c = (const char*) p; m_buf.insert(m_buf.end(), c, c + n); }
    void read(void* p, size_t n) {
        if (m_pos + n > m_buf.size()) Log(EError, "Stream: read past the end");
        memcpy(p, m_buf.data() + m_pos, n);
        m_pos += n;
    }
    void writeInt(int v) { write(&v, 4); }
    int readInt() { int v; read(&v, 4); return v; }
    void writeFloat(Float v) { write(&v, 4); }
    Float readFloat() { Float v; read(&v, 4); return v; }
    void writeDouble(double v) { write(&v, 8); }
    double readDouble() { double v; read(&v, 8); return v; }
    void writeBool(bool v) { char c = v; write(&c, 1); }
    bool readBool() { char c; read(&c, 1); return c != 0; }
    void writeString(const std::string& s) { writeInt((int) s.size()); write(s.data(), s.size()); }
    std::string readString() { std::string s((size_t) readInt(), '\0'); read(&s[0], s.size()); return s; }
    void writeDoubleArray(const double* p, size_t n) { write(p, n * 8); }
    void readDoubleArray(double* p, size_t n) { read(p, n * 8); }
    void seek(size_t p) { m_pos = p; }
    size_t getSize() const { return m_buf.size(); }
private:
    std::vector<char> m_buf;
    size_t m_pos = 0;
};
class InstanceManager {};

class Properties {
public:
    void setString(const std::string& k, const std::string& v) { m_s[k] = v; }
    void setFloat(const std::string& k, double v) { m_f[k] = v; }
    bool hasProperty(const std::string& k) const { return m_s.count(k) || m_f.count(k); }
    std::string getString(const std::string& k) const {
        if (!m_s.count(k)) Log(EError, "Property \"%s\" missing", k.c_str());
        return m_s.at(k);
    }
    std::string getString(const std::string& k, const std::string& d) const { return m_s.count(k) ? m_s.at(k) : d; }
    Float getFloat(const std::string& k, Float d) const { return m_f.count(k) ? (Float) m_f.at(k) : d; }
    int getInteger(const std::string& k, int d) const { return m_f.count(k) ? (int) m_f.at(k) : d; }
    bool getBoolean(const std::string& k, bool d) const { return m_f.count(k) ? m_f.at(k) != 0 : d; }
private:
    std::map<std::string, std::string> m_s;
    std::map<std::string, double> m_f;
};

struct FileResolver { std::string resolve(const std::string& p) const { return p; } };
struct Thread {
    static Thread* getThread() { static Thread t; return &t; }
    const FileResolver* getFileResolver() const { static FileResolver f; return &f; }
};

class BSDF;
struct BSDFSamplingRecord {
    const Intersection& its;
    Vector wi, wo;
    Float eta = 1;
    ETransportMode mode = ERadiance;
    unsigned int typeMask = 0xFFFFFFFFu;
    int component = -1;
    unsigned int sampledType = 0;
    int sampledComponent = -1;
    BSDFSamplingRecord(const Intersection& i, const Vector& wi_, const Vector& wo_, ETransportMode m = ERadiance)
        : its(i), wi(wi_), wo(wo_), mode(m) {}
    BSDFSamplingRecord(const Intersection& i, const Vector& wi_, ETransportMode m = ERadiance)
        : its(i), wi(wi_), mode(m) {}
};

class ConfigurableObject {
public:
    virtual ~ConfigurableObject() {}
    virtual void configure() {}
    virtual void serialize(Stream*, InstanceManager*) const {}
    virtual std::string toString() const { return "ConfigurableObject[]"; }
};

class BSDF : public ConfigurableObject {
public:
    enum EBSDFType {
        ENull = 0x00001, EDiffuseReflection = 0x00002, EDiffuseTransmission = 0x00004, EGlossyReflection = 0x00008,
        EGlossyTransmission = 0x00010, EDeltaReflection = 0x00020, EDeltaTransmission = 0x00040,
        EAnisotropic = 0x01000, ESpatiallyVarying = 0x02000, ENonSymmetric = 0x04000, EFrontSide = 0x10000,
        EBackSide = 0x20000, EUsesSampler = 0x40000
    };
    explicit BSDF(const Properties&) {}
    BSDF(Stream*, InstanceManager*) {}
    void configure() override {
        m_combinedType = 0;
        for (unsigned c : m_components) m_combinedType |= c;
    }
    virtual Spectrum eval(const BSDFSamplingRecord& bRec, EMeasure measure = ESolidAngle) const = 0;
    virtual Float pdf(const BSDFSamplingRecord& bRec, EMeasure measure = ESolidAngle) const = 0;
    virtual Spectrum sample(BSDFSamplingRecord& bRec, const Point2& sample) const = 0;
    virtual Spectrum sample(BSDFSamplingRecord& bRec, Float& pdf, const Point2& sample) const = 0;
    virtual const char* getClassName() const = 0;
    unsigned int getType() const { return m_combinedType; }
    int getComponentCount() const { return (int) m_components.size(); }
protected:
    std::vector<unsigned int> m_components;
    unsigned int m_combinedType = 0;
    bool m_usesRayDifferentials = false;
};

MTS_NAMESPACE_END
