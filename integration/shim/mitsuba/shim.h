// SYNTAX-CHECK SHIM — not Mitsuba.  The reference's headers all include Enoki, which is missing from the reference archive
// (ext/enoki is an empty submodule), so the plugin sources under integration/src cannot be compiled against them here.  This file
// declares, with the reference's names and signatures (cited), just the part of the API those sources use, so that
// `g++ -fsyntax-only` type-checks them -- including one explicit instantiation per plugin (MTS_EXPORT_PLUGIN below).
// tests/test_integration_sources.py runs that check.  Nothing here is linked into anything.
#pragma once
#include <cstdint>
#include <cstdio>
#include <map>
#include <stdexcept>
#include <string>
#include <typeinfo>
#include <utility>
#include <vector>

#define NAMESPACE_BEGIN(name) namespace name {
#define NAMESPACE_END(name) }
#define MTS_C 299792458.f                 /* include/mitsuba/core/spectrum.h:32-33 */
#define MTS_WAVELENGTH_MIN 3090644.f      /* spectrum.h:15-27 (the 94 GHz alternative) */
#define MTS_WAVELENGTH_MAX 3294422.f
#define MTS_CIE_MIN 360.f
#define MTS_CIE_MAX 830.f
#define MTS_P 1.f

// FOR_EACH over up to 12 macro arguments
#define SHIM_CAT(a, b) a##b
#define SHIM_NARGS_(_1, _2, _3, _4, _5, _6, _7, _8, _9, _10, _11, _12, N, ...) N
#define SHIM_NARGS(...) SHIM_NARGS_(__VA_ARGS__, 12, 11, 10, 9, 8, 7, 6, 5, 4, 3, 2, 1)
#define SHIM_FE_1(M, a) M(a)
#define SHIM_FE_2(M, a, ...) M(a) SHIM_FE_1(M, __VA_ARGS__)
#define SHIM_FE_3(M, a, ...) M(a) SHIM_FE_2(M, __VA_ARGS__)
#define SHIM_FE_4(M, a, ...) M(a) SHIM_FE_3(M, __VA_ARGS__)
#define SHIM_FE_5(M, a, ...) M(a) SHIM_FE_4(M, __VA_ARGS__)
#define SHIM_FE_6(M, a, ...) M(a) SHIM_FE_5(M, __VA_ARGS__)
#define SHIM_FE_7(M, a, ...) M(a) SHIM_FE_6(M, __VA_ARGS__)
#define SHIM_FE_8(M, a, ...) M(a) SHIM_FE_7(M, __VA_ARGS__)
#define SHIM_FE_9(M, a, ...) M(a) SHIM_FE_8(M, __VA_ARGS__)
#define SHIM_FE_10(M, a, ...) M(a) SHIM_FE_9(M, __VA_ARGS__)
#define SHIM_FE_11(M, a, ...) M(a) SHIM_FE_10(M, __VA_ARGS__)
#define SHIM_FE_12(M, a, ...) M(a) SHIM_FE_11(M, __VA_ARGS__)
#define SHIM_FE_N(N, M, ...) SHIM_CAT(SHIM_FE_, N)(M, __VA_ARGS__)
#define SHIM_FOR_EACH(M, ...) SHIM_FE_N(SHIM_NARGS(__VA_ARGS__), M, __VA_ARGS__)

// include/mitsuba/core/config.h / class.h / render/fwd.h
#define SHIM_USING_MEMBER(m) using Base::m;
#define MTS_IMPORT_BASE(Name, ...) using Base = Name<Float, Spectrum>; SHIM_FOR_EACH(SHIM_USING_MEMBER, __VA_ARGS__)
#define SHIM_IMPORT_TYPE(T) using T = ::mitsuba::T<Float, Spectrum>;
#define MTS_IMPORT_TYPES(...) SHIM_FOR_EACH(SHIM_IMPORT_TYPE, __VA_ARGS__)
#define MTS_DECLARE_CLASS() public: const ::mitsuba::Class *class_() const override { return nullptr; }
#define MTS_IMPLEMENT_CLASS_VARIANT(Name, Parent)
#define MTS_EXPORT_PLUGIN(Name, descr) template class Name<float, ::mitsuba::ShimSpectrum>;

namespace mitsuba {

struct ShimSpectrum { float v[1]; };
template <typename T> constexpr size_t array_size_v = 1;

template <typename... A> [[noreturn]] void Throw(const char *fmt, A...) { throw std::runtime_error(fmt); }              // core/logger.h:134-138
[[noreturn]] inline void NotImplementedError(const char *name) { throw std::logic_error(name); }
namespace tfm { template <typename... A> std::string format(const char *fmt, A...) { return fmt; } }

struct Class { std::string m_name; const std::string &name() const { return m_name; } };
class TraversalCallback;
class Object {                                                                                                             // core/object.h
public:
    virtual const Class *class_() const = 0;
    virtual void traverse(TraversalCallback *) {}
    virtual std::string to_string() const { return ""; }
    virtual ~Object() = default;
};
template <typename T> class ref {
public:
    ref(T *p = nullptr) : m_p(p) {}
    T *get() const { return m_p; }
    T *operator->() const { return m_p; }
    operator T *() const { return m_p; }
private:
    T *m_p;
};
class TraversalCallback {                                                                                                  // core/object.h:272-287
public:
    template <typename T> void put_parameter(const std::string &name, T &t) { put_parameter_impl(name, typeid(T), &t); }
    virtual void put_object(const std::string &name, Object *obj) = 0;
    virtual ~TraversalCallback() = default;
protected:
    virtual void put_parameter_impl(const std::string &name, const std::type_info &type, void *ptr) = 0;
};
class Properties {                                                                                                         // core/properties.h
public:
    int int_(const std::string &, int d) const { return d; }
    long long long_(const std::string &, long long d) const { return d; }
    bool bool_(const std::string &, bool d) const { return d; }
    float float_(const std::string &, float d) const { return d; }
};
struct ScalarVector3f { float a[3]; float x() const { return a[0]; } float y() const { return a[1]; } float z() const { return a[2]; } };
struct ScalarVector2i { int a[2]; int x() const { return a[0]; } int y() const { return a[1]; } };
struct ScalarVector2f { float a[2]; float x() const { return a[0]; } float y() const { return a[1]; } };
using ScalarPoint2i = ScalarVector2i;
inline int hprod(const ScalarVector2i &v) { return v.x() * v.y(); }
struct ScalarMatrix4f { float m[4][4]; float operator()(int r, int c) const { return m[r][c]; } };
struct ScalarTransform4f { ScalarMatrix4f matrix; };                                                                       // core/transform.h
template <typename T> struct DynamicBuffer { const T *data() const { return nullptr; } size_t size() const { return 0; } T *data() { return nullptr; } };

#define SHIM_TEMPLATE template <typename Float, typename Spectrum>
SHIM_TEMPLATE class Texture : public Object {};
SHIM_TEMPLATE class Medium : public Object {};
SHIM_TEMPLATE class BSDF : public Object {};
SHIM_TEMPLATE class ReconstructionFilter : public Object { public: float radius() const { return 0.5f; } };              // core/rfilter.h
SHIM_TEMPLATE class Sampler : public Object { public: size_t sample_count() const { return 4; } size_t seed() const { return 0; } };   // render/sampler.h
SHIM_TEMPLATE class Shape : public Object {                                                                                // render/shape.h
public:
    bool is_mesh() const { return false; }
    const BSDF<Float, Spectrum> *bsdf() const { return nullptr; }
};
SHIM_TEMPLATE class Mesh : public Shape<Float, Spectrum> {                                                                 // render/mesh.h
public:
    uint32_t vertex_count() const { return 0; } uint32_t face_count() const { return 0; }
    bool has_vertex_normals() const { return false; }
    const DynamicBuffer<float> &vertex_positions_buffer() const { return m_f; }
    const DynamicBuffer<float> &vertex_normals_buffer() const { return m_f; }
    const DynamicBuffer<uint32_t> &faces_buffer() const { return m_u; }
private:
    DynamicBuffer<float> m_f; DynamicBuffer<uint32_t> m_u;
};
SHIM_TEMPLATE class Endpoint : public Object { public: const Shape<Float, Spectrum> *shape() const { return nullptr; } };  // render/endpoint.h
SHIM_TEMPLATE class Emitter : public Endpoint<Float, Spectrum> {};
SHIM_TEMPLATE class Transmitter : public Endpoint<Float, Spectrum> {};
SHIM_TEMPLATE class SignalBlock : public Object {                                                                          // render/signalblock.h
public:
    SignalBlock(const ScalarVector2i &, size_t, const ReconstructionFilter<Float, Spectrum> * = nullptr, bool = true, bool = true, bool = true) {}
    void set_offset(const ScalarPoint2i &) {} void clear() {}
    DynamicBuffer<float> &data() { return m_d; }
    const Class *class_() const override { return nullptr; }
private:
    DynamicBuffer<float> m_d;
};
SHIM_TEMPLATE class ImageBlock : public SignalBlock<Float, Spectrum> { public: using SignalBlock<Float, Spectrum>::SignalBlock; };   // render/imageblock.h
SHIM_TEMPLATE class ADC : public Object {                                                                                  // render/adc.h:26-97
public:
    virtual void prepare(const std::vector<std::string> &channels) = 0;
    virtual void put(const SignalBlock<Float, Spectrum> *block) = 0;
    const ScalarVector2i &size() const { return m_i; } const ScalarVector2i &window_size() const { return m_i; }
    const ScalarPoint2i &window_offset() const { return m_i; } const ScalarVector2f &bandwidth() const { return m_f; }
    const ReconstructionFilter<Float, Spectrum> *reconstruction_filter() const { return nullptr; }
private:
    ScalarVector2i m_i; ScalarVector2f m_f;
};
SHIM_TEMPLATE class Film : public Object {                                                                                 // render/film.h
public:
    virtual void prepare(const std::vector<std::string> &channels) = 0;
    virtual void put(const ImageBlock<Float, Spectrum> *block) = 0;
    const ScalarVector2i &size() const { return m_i; } const ScalarVector2i &crop_size() const { return m_i; } const ScalarPoint2i &crop_offset() const { return m_i; }
    const ReconstructionFilter<Float, Spectrum> *reconstruction_filter() const { return nullptr; }
private:
    ScalarVector2i m_i;
};
SHIM_TEMPLATE class Sensor : public Endpoint<Float, Spectrum> {                                                            // render/sensor.h
public:
    float shutter_open() const { return 0.f; } float shutter_open_time() const { return 0.f; }
    Film<Float, Spectrum> *film() const { return nullptr; } Sampler<Float, Spectrum> *sampler() const { return nullptr; }
};
SHIM_TEMPLATE class Receiver : public Endpoint<Float, Spectrum> {                                                          // render/receiver.h:61-112
public:
    float adc_sampling_start() const { return 0.f; } float adc_sampling_time() const { return 0.f; }
    std::string receive_type() const { return "raw"; }
    ADC<Float, Spectrum> *adc() const { return nullptr; } Sampler<Float, Spectrum> *sampler() const { return nullptr; }
};
SHIM_TEMPLATE class Scene : public Object {                                                                                // render/scene.h:150-200
public:
    const std::vector<ref<Shape<Float, Spectrum>>> &shapes() const { return m_shapes; }
    const std::vector<ref<Emitter<Float, Spectrum>>> &emitters() const { return m_emitters; }
    const std::vector<ref<Transmitter<Float, Spectrum>>> &transmitters() const { return m_transmitters; }
private:
    std::vector<ref<Shape<Float, Spectrum>>> m_shapes; std::vector<ref<Emitter<Float, Spectrum>>> m_emitters; std::vector<ref<Transmitter<Float, Spectrum>>> m_transmitters;
};
SHIM_TEMPLATE class SamplingIntegrator : public Object {                                                                   // render/integrator.h:43-53, 116-171
public:
    using Scene = ::mitsuba::Scene<Float, Spectrum>; using Sensor = ::mitsuba::Sensor<Float, Spectrum>; using Receiver = ::mitsuba::Receiver<Float, Spectrum>;
    explicit SamplingIntegrator(const Properties &) {}
    virtual bool render(Scene *scene, Sensor *sensor) = 0;
    virtual bool receive(Scene *scene, Receiver *receiver) = 0;
    virtual void cancel() { m_stop = true; }
    virtual std::vector<std::string> aov_names() const { return {}; }
    bool should_stop() const { return m_stop; }
protected:
    int m_max_depth = -1, m_rr_depth = 5;
    bool m_stop = false;
};

}  // namespace mitsuba
