// beifong_b200 — the Mitsuba-side half of the drop-in: everything the three integrator plugins of this
// directory share.  A loaded mitsuba::Scene is FLATTENED into the calls of include/beifong_b200/abi.h:
//
//   Scene::shapes()        -> bf_scene_add_mesh (Mesh buffers, mesh.cpp:789-805 hands the same two to Embree) / bf_scene_add_rect
//                             (`to_world` through Shape::traverse, shape.cpp:483-485) + bf_scene_set_shape_velocity (`velocity`)
//   Shape::bsdf()          -> bsdf_id(): diffuse / twosided / roughconductor / conductor -> bf_bsdf_desc (+ their spectra)
//   Scene::transmitters()  -> endpoint_desc(): areatransmitter / wignertransmitter / phasedtransmitter
//   Scene::emitters()      -> endpoint_desc(): spot / area                    (gen 1 / 2 scenes)
//   Receiver / Sensor      -> endpoint_desc(): omnidirectional / wignerreceiver / phasedreceiver ; fluxmeter / perspective
//   ADC / Film             -> bf_adc_desc / the film size
//
// The reference's endpoint plugins keep their signal parameters in private members without accessors
// (wignertransmitter.cpp:601-612, wignerreceiver.cpp:303-312, phasedtransmitter.cpp, phasedreceiver.cpp) and RoughConductor
// keeps `distribution` / `sample_visible` private (roughconductor.cpp:425-430).  Mitsuba's own mechanism for exposing plugin state
// is Object::traverse; integration/patches/expose_parameters.md lists the put_parameter lines those plugins gain (no behaviour
// changes).  With them in place every value below is read through a TraversalCallback -- no XML re-parsing, no friend access.
//
// Written against /root/reference/include/mitsuba/{core,render}; not compiled here (ext/enoki is missing from the reference
// archive), syntax-checked against integration/shim/ by tests/test_integration_sources.py.
#pragma once

#include <mitsuba/core/properties.h>
#include <mitsuba/core/transform.h>
#include <mitsuba/render/adc.h>
#include <mitsuba/render/bsdf.h>
#include <mitsuba/render/emitter.h>
#include <mitsuba/render/film.h>
#include <mitsuba/render/mesh.h>
#include <mitsuba/render/receiver.h>
#include <mitsuba/render/scene.h>
#include <mitsuba/render/sensor.h>
#include <mitsuba/render/texture.h>
#include <mitsuba/render/transmitter.h>

#include <cstring>
#include <map>
#include <string>
#include <typeinfo>
#include <unordered_map>
#include <vector>

#include <beifong_b200/abi.h>

NAMESPACE_BEGIN(mitsuba)
NAMESPACE_BEGIN(b200)

/// Collects what an object reports through Object::traverse: scalar parameters by name, child objects by name.
class ParamCollector : public TraversalCallback {
public:
    std::map<std::string, double> numbers;
    std::map<std::string, std::string> strings;
    std::map<std::string, ScalarTransform4f> transforms;
    std::map<std::string, std::vector<float>> vectors;
    std::map<std::string, Object *> objects;

    explicit ParamCollector(Object *o) { if (o) o->traverse(this); }
    void put_object(const std::string &name, Object *obj) override { objects[name] = obj; }

    double number(const std::string &name, double dflt) const { auto it = numbers.find(name); return it == numbers.end() ? dflt : it->second; }
    bool flag(const std::string &name, bool dflt) const { return number(name, dflt ? 1.0 : 0.0) != 0.0; }
    std::string string(const std::string &name, const std::string &dflt) const { auto it = strings.find(name); return it == strings.end() ? dflt : it->second; }
    bool has(const std::string &name) const { return numbers.count(name) || strings.count(name) || transforms.count(name) || vectors.count(name); }

protected:
    void put_parameter_impl(const std::string &name, const std::type_info &type, void *ptr) override {
        if (type == typeid(float)) numbers[name] = *(const float *) ptr;
        else if (type == typeid(double)) numbers[name] = *(const double *) ptr;
        else if (type == typeid(int)) numbers[name] = *(const int *) ptr;
        else if (type == typeid(uint32_t)) numbers[name] = *(const uint32_t *) ptr;
        else if (type == typeid(bool)) numbers[name] = *(const bool *) ptr ? 1.0 : 0.0;
        else if (type == typeid(std::string)) strings[name] = *(const std::string *) ptr;
        else if (type == typeid(ScalarTransform4f)) transforms[name] = *(const ScalarTransform4f *) ptr;
        else if (type == typeid(ScalarVector3f)) { const ScalarVector3f &v = *(const ScalarVector3f *) ptr; vectors[name] = {v.x(), v.y(), v.z()}; }
        else if (type == typeid(std::vector<float>)) vectors[name] = *(const std::vector<float> *) ptr;
        // anything else (dynamic buffers, textures' bitmaps) is not needed on this path
    }
};

inline void to_row_major(const ScalarTransform4f &t, float out[16]) {
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c)
            out[4 * r + c] = t.matrix(r, c);
}

/// Scene -> bf_scene.  One instance per (integrator, scene); `handle()` stays valid until destruction.
template <typename Float, typename Spectrum> class Flattener {
public:
    MTS_IMPORT_TYPES(Scene, Shape, Mesh, BSDF, Texture, Emitter, Transmitter, Sensor, Receiver, ADC, Film, ReconstructionFilter)

    Flattener() = default;
    ~Flattener() { if (m_handle) bf_scene_destroy(m_handle); }
    Flattener(const Flattener &) = delete;
    Flattener &operator=(const Flattener &) = delete;

    bf_scene *handle() const { return m_handle; }
    bool ready() const { return m_committed; }
    void check(int rc, const char *what) const {
        if (rc == BF_OK) return;
        const char *msg = bf_last_error(m_handle);
        if (rc == BF_ERR_UNSUPPORTED) NotImplementedError(msg && *msg ? msg : what);          // -> Python NotImplementedError
        Throw("beifong_b200: %s: %s", what, msg && *msg ? msg : "error");                      // -> std::runtime_error -> RuntimeError
    }

    /// Shapes, BSDFs and transmitters / emitters of `scene`.  Call once, then set_receiver / set_sensor, then commit().
    void add_scene(const Scene *scene, bool use_emitters) {
        bf_constants c;
        std::memset(&c, 0, sizeof(c));
        c.c = (float) MTS_C; c.lambda_min_nm = (float) MTS_WAVELENGTH_MIN; c.lambda_max_nm = (float) MTS_WAVELENGTH_MAX;   // spectrum.h:15-40
        c.cie_min_nm = (float) MTS_CIE_MIN; c.cie_max_nm = (float) MTS_CIE_MAX; c.P = (float) MTS_P;
        c.spectrum_width = (uint32_t) array_size_v<Spectrum>;
        if (bf_scene_create(&c, &m_handle) != BF_OK) Throw("beifong_b200: %s", bf_last_error(nullptr));
        for (auto &shape : scene->shapes()) {
            ParamCollector sp(shape.get());
            const uint32_t bsdf = bsdf_id(const_cast<BSDF *>(shape->bsdf()));
            uint32_t id = BF_NONE;
            if (shape->is_mesh()) {
                const Mesh *m = static_cast<const Mesh *>(shape.get());
                check(bf_scene_add_mesh(m_handle, (const float *) m->vertex_positions_buffer().data(), (uint32_t) m->vertex_count(),
                                        (const uint32_t *) m->faces_buffer().data(), (uint32_t) m->face_count(),
                                        m->has_vertex_normals() ? (const float *) m->vertex_normals_buffer().data() : nullptr, bsdf, &id),
                      "bf_scene_add_mesh");
            } else if (shape->class_()->name() == "Rectangle") {
                float m16[16];
                to_row_major(sp.transforms.at("to_world"), m16);                               // includes flip_normals (rectangle.cpp:76-81)
                check(bf_scene_add_rect(m_handle, m16, bsdf, &id), "bf_scene_add_rect");
            } else {
                NotImplementedError(("shape plugin \"" + shape->class_()->name() + "\" (the B200 path takes meshes and rectangles)").c_str());
            }
            if (sp.transforms.count("velocity")) {                                             // shape.cpp:39-42, :485
                float v16[16];
                to_row_major(sp.transforms.at("velocity"), v16);
                check(bf_scene_set_shape_velocity(m_handle, id, v16), "bf_scene_set_shape_velocity");
            }
            m_shape_ids[shape.get()] = id;
        }
        if (use_emitters) {
            for (auto &e : scene->emitters()) {
                bf_endpoint_desc d = endpoint_desc(e.get(), e->class_()->name(), /*is_tx*/ true);
                check(bf_scene_add_transmitter(m_handle, shape_id_of(e->shape()), &d, nullptr), "bf_scene_add_transmitter");
            }
        } else {
            for (auto &t : scene->transmitters()) {
                bf_endpoint_desc d = endpoint_desc(t.get(), t->class_()->name(), /*is_tx*/ true);
                check(bf_scene_add_transmitter(m_handle, shape_id_of(t->shape()), &d, nullptr), "bf_scene_add_transmitter");
            }
        }
    }

    /// gen 3: the receiver and its ADC (receiver.cpp:16-62, adc.cpp:7-76)
    void set_receiver(const Receiver *rx) {
        bf_endpoint_desc d = endpoint_desc(const_cast<Receiver *>(rx), rx->class_()->name(), /*is_tx*/ false);
        d.adc_sampling_start = (float) rx->adc_sampling_start();
        d.adc_sampling_time = (float) rx->adc_sampling_time();
        const std::string rt = rx->receive_type();
        d.receive_type = rt == "raw" ? BF_RECV_RAW : rt == "raw_resample" ? BF_RECV_RAW_RESAMPLE : rt == "mix_resample" ? BF_RECV_MIX_RESAMPLE : BF_RECV_MIXER;
        check(bf_scene_set_receiver(m_handle, shape_id_of(rx->shape()), &d), "bf_scene_set_receiver");
        const ADC *adc = rx->adc();
        bf_adc_desc a;
        std::memset(&a, 0, sizeof(a));
        a.t_bins = (uint32_t) adc->size().x(); a.f_bins = (uint32_t) adc->size().y();
        a.window_t_bins = (uint32_t) adc->window_size().x(); a.window_f_bins = (uint32_t) adc->window_size().y();
        a.window_offset_t = (uint32_t) adc->window_offset().x(); a.window_offset_f = (uint32_t) adc->window_offset().y();
        a.t_bandwidth = adc->bandwidth().x(); a.f_bandwidth = adc->bandwidth().y();
        filter_desc(adc->reconstruction_filter(), a.filter_radius, a.filter_kind);
        check(bf_scene_set_adc(m_handle, &a), "bf_scene_set_adc");
    }

    /// gen 1 / 2: the sensor and its film (sensor.cpp, film.cpp); the film must be 1 x 1 box-filtered bins, as the radar scripts use it
    void set_sensor(const Sensor *sensor) {
        bf_endpoint_desc d = endpoint_desc(const_cast<Sensor *>(sensor), sensor->class_()->name(), /*is_tx*/ false);
        d.shutter_open = (float) sensor->shutter_open(); d.shutter_open_time = (float) sensor->shutter_open_time();
        d.film_width = (uint32_t) sensor->film()->size().x(); d.film_height = (uint32_t) sensor->film()->size().y();
        check(bf_scene_set_receiver(m_handle, shape_id_of(sensor->shape()), &d), "bf_scene_set_receiver");
    }

    void commit(const bf_integrator_desc &integ, int device) {
        check(bf_scene_set_integrator(m_handle, &integ), "bf_scene_set_integrator");
        check(bf_scene_commit(m_handle, device), "bf_scene_commit");                           // SAH build + upload; fails loudly without a GPU
        m_committed = true;
    }

    /// A shape whose `to_world` changed since the last call (moving antennas / targets between frames): re-pose it in place.
    void update_shape(const Shape *shape) {
        ParamCollector sp(const_cast<Shape *>(shape));
        float m16[16];
        to_row_major(sp.transforms.at("to_world"), m16);
        check(bf_scene_update_transform(m_handle, shape_id_of(shape), m16), "bf_scene_update_transform");
    }

    uint32_t shape_id_of(const Shape *shape) const {
        if (!shape) return BF_NONE;
        auto it = m_shape_ids.find(shape);
        if (it == m_shape_ids.end()) Throw("beifong_b200: endpoint attached to a shape that is not part of the scene");
        return it->second;
    }

    // ------------------------------------------------------------------------------------------------------------------
    /// Texture plugins `uniform` / `regular` / `irregular` (src/spectra) -> spectrum id.  `d65`, `srgb*`, bitmaps are not radar inputs.
    uint32_t spectrum_id(Object *tex_obj, float dflt) {
        uint32_t id = BF_NONE;
        if (!tex_obj) {
            check(bf_scene_add_spectrum(m_handle, BF_SPEC_UNIFORM, 1, nullptr, &dflt, &id), "bf_scene_add_spectrum");
            return id;
        }
        auto cached = m_spectrum_ids.find(tex_obj);
        if (cached != m_spectrum_ids.end()) return cached->second;
        ParamCollector tp(tex_obj);
        const std::string cls = tex_obj->class_()->name();
        if (cls == "UniformSpectrum") {
            float v = (float) tp.number("value", dflt);                                          // uniform.cpp:42-44
            check(bf_scene_add_spectrum(m_handle, BF_SPEC_UNIFORM, 1, nullptr, &v, &id), "bf_scene_add_spectrum");
        } else if (cls == "RegularSpectrum") {
            const std::vector<float> &values = tp.vectors.at("values");                         // regular.cpp:87-91 ("range", "values")
            const std::vector<float> &range = tp.vectors.at("range");
            std::vector<float> lam(values.size());
            for (size_t i = 0; i < lam.size(); ++i) lam[i] = range[0] + (range[1] - range[0]) * (float) i / (float) (lam.size() - 1);
            check(bf_scene_add_spectrum(m_handle, BF_SPEC_REGULAR, (uint32_t) values.size(), lam.data(), values.data(), &id), "bf_scene_add_spectrum");
        } else if (cls == "IrregularSpectrum") {
            const std::vector<float> &values = tp.vectors.at("values"), &lam = tp.vectors.at("wavelengths");   // irregular.cpp:98-102
            check(bf_scene_add_spectrum(m_handle, BF_SPEC_IRREGULAR, (uint32_t) values.size(), lam.data(), values.data(), &id), "bf_scene_add_spectrum");
        } else {
            NotImplementedError(("texture plugin \"" + cls + "\" on the B200 radar path").c_str());
        }
        m_spectrum_ids[tex_obj] = id;
        return id;
    }

    /// diffuse.cpp / twosided.cpp / roughconductor.cpp / conductor.cpp -> bf_bsdf_desc.  A shape without a BSDF gets Mitsuba's default
    /// `diffuse` with reflectance 0.5 (shape.cpp:52-60).
    uint32_t bsdf_id(BSDF *bsdf) {
        auto cached = m_bsdf_ids.find(bsdf);
        if (cached != m_bsdf_ids.end()) return cached->second;
        bf_bsdf_desc d;
        std::memset(&d, 0, sizeof(d));
        d.back_bsdf = d.reflectance = d.eta = d.k = d.specular_reflectance = BF_NONE;
        d.alpha_u = d.alpha_v = 0.1f; d.sample_visible = 1; d.distribution = BF_MICROFACET_BECKMANN;
        BSDF *inner = bsdf;
        if (bsdf && bsdf->class_()->name() == "TwoSidedBRDF") {                                // twosided.cpp:94-181
            ParamCollector tp(bsdf);
            inner = static_cast<BSDF *>(tp.objects.at("brdf_0"));
            BSDF *back = static_cast<BSDF *>(tp.objects.at("brdf_1"));
            d.two_sided = 1;
            if (back != inner) d.back_bsdf = bsdf_id(back);
        }
        fill_one_sided(inner, d);
        uint32_t id = BF_NONE;
        check(bf_scene_add_bsdf(m_handle, &d, &id), "bf_scene_add_bsdf");
        m_bsdf_ids[bsdf] = id;
        return id;
    }

    /// One endpoint plugin -> bf_endpoint_desc.  `cls` is the class name the plugin registers (MTS_IMPLEMENT_CLASS_VARIANT).
    bf_endpoint_desc endpoint_desc(Object *ep, const std::string &cls, bool is_tx) {
        bf_endpoint_desc d;
        std::memset(&d, 0, sizeof(d));
        const float eye[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
        std::memcpy(d.to_world, eye, 64); std::memcpy(d.sample_to_camera, eye, 64);
        d.film_width = d.film_height = 1;
        ParamCollector p(ep);
        auto tex = [&](const char *name, float dflt) { auto it = p.objects.find(name); return spectrum_id(it == p.objects.end() ? nullptr : it->second, dflt); };
        if (is_tx) {
            if (cls == "AreaTransmitter") { d.kind = BF_TX_AREA; d.texture = tex("radiance", 1.f); }                       // areatransmitter.cpp:52-63
            else if (cls == "WignerTransmitter") { d.kind = BF_TX_WIGNER; d.texture = tex("antenna_texture", 1.f); signal_block(p, d.signal); }   // wignertransmitter.cpp:59-102
            else if (cls == "PhasedTransmitter") { d.kind = BF_TX_PHASED; d.texture = tex("antenna_texture", 1.f); signal_block(p, d.signal); array_block(p, d.array); }   // phasedtransmitter.cpp:53-167
            else if (cls == "SpotLight") {                                                                                 // spot.cpp:72-95
                d.kind = BF_TX_SPOT; d.texture = tex("intensity", 1.f);
                d.cutoff_angle_deg = (float) p.number("cutoff_angle", 20.0); d.beam_width_deg = (float) p.number("beam_width", d.cutoff_angle_deg * 3.0 / 4.0);
                if (p.transforms.count("to_world")) to_row_major(p.transforms.at("to_world"), d.to_world);
            }
            else if (cls == "AreaLight") { d.kind = BF_TX_AREALIGHT; d.texture = tex("radiance", 1.f); }                    // area.cpp
            else NotImplementedError(("transmitter / emitter plugin \"" + cls + "\" on the B200 radar path").c_str());
        } else {
            if (cls == "Omnidirectional") { d.kind = BF_RX_OMNIDIRECTIONAL; d.texture = tex("antenna_texture", 1.f); }      // omnidirectional.cpp:71-107
            else if (cls == "Wignerreceiver") { d.kind = BF_RX_WIGNER; d.texture = tex("antenna_texture", 1.f); signal_block(p, d.signal); }    // wignerreceiver.cpp:59-112
            else if (cls == "Phasedreceiver") { d.kind = BF_RX_PHASED; d.texture = tex("antenna_texture", 1.f); signal_block(p, d.signal); array_block(p, d.array); }   // phasedreceiver.cpp:50-175
            else if (cls == "FluxMeter") { d.kind = BF_RX_FLUXMETER; }                                                     // fluxmeter.cpp:63-84
            else if (cls == "PerspectiveCamera") {                                                                         // perspective.cpp:101-164
                d.kind = BF_RX_PERSPECTIVE;
                d.near_clip = (float) p.number("near_clip", 1e-2); d.far_clip = (float) p.number("far_clip", 1e4);
                if (p.transforms.count("to_world")) to_row_major(p.transforms.at("to_world"), d.to_world);
                if (p.transforms.count("sample_to_camera")) to_row_major(p.transforms.at("sample_to_camera"), d.sample_to_camera);
            }
            else NotImplementedError(("receiver / sensor plugin \"" + cls + "\" on the B200 radar path").c_str());
        }
        return d;
    }

private:
    /// signaltype, amplitude, crf / prf, chirp_len / pulse_len, freq_centre, freq_sweep / freq_ext, phase, sig_is_delta, gain,
    /// resample_freq: the members wignertransmitter.cpp:601-612 lists, under the names the patch exposes them with
    static void signal_block(const ParamCollector &p, bf_signal_desc &s) {
        const std::string type = p.string("signaltype", "cw");
        s.signaltype = type == "cw" ? BF_SIG_CW : type == "linfmcw" ? BF_SIG_LINFMCW : type == "pulse" ? BF_SIG_PULSE : BF_SIG_OTHER;
        s.sig_is_delta = p.flag("sig_is_delta", false); s.resample_freq = p.flag("resample_freq", false);
        s.amplitude = (float) p.number("sig_amplitude", 1.0); s.repfreq = (float) p.number("sig_repfreq", 1.0); s.t_ext = (float) p.number("sig_t_ext", 1.0);
        s.f_centre = (float) p.number("sig_f_centre", 1.0); s.f_ext = (float) p.number("sig_f_ext", 1.0); s.phi0 = (float) p.number("sig_phi0", 0.0);
        s.gain = (float) p.number("gain", 1.0);
    }
    /// n_elems, steering_vector, array_loc, elem_dims, elem_spacing, elem_axis (phasedtransmitter.cpp:104-117)
    static void array_block(const ParamCollector &p, bf_array_desc &a) {
        a.n_elems = (uint32_t) p.number("n_elems", 1.0);
        auto v3 = [&](const char *name, float out[3], float dflt) { auto it = p.vectors.find(name); for (int i = 0; i < 3; ++i) out[i] = it == p.vectors.end() ? dflt : it->second[i]; };
        v3("steering_vector", a.steering_vector, 0.f); v3("elem_dims", a.elem_dims, 1.f); v3("elem_spacing", a.elem_spacing, 0.f); v3("elem_axis", a.elem_axis, 0.f);
        const float eye[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
        std::memcpy(a.array_loc, eye, 64);
        if (p.transforms.count("array_loc")) to_row_major(p.transforms.at("array_loc"), a.array_loc);
    }
    void fill_one_sided(BSDF *b, bf_bsdf_desc &d) {
        if (!b) { d.kind = BF_BSDF_DIFFUSE; d.reflectance = spectrum_id(nullptr, 0.5f); return; }
        ParamCollector p(b);
        const std::string cls = b->class_()->name();
        auto obj = [&](const char *name) -> Object * { auto it = p.objects.find(name); return it == p.objects.end() ? nullptr : it->second; };
        if (cls == "SmoothDiffuse") {                                                           // diffuse.cpp:68-76
            d.kind = BF_BSDF_DIFFUSE; d.reflectance = spectrum_id(obj("reflectance"), 0.5f);
        } else if (cls == "RoughConductor" || cls == "SmoothConductor") {                       // roughconductor.cpp:156-194, conductor.cpp
            d.kind = cls == "RoughConductor" ? BF_BSDF_ROUGHCONDUCTOR : BF_BSDF_CONDUCTOR;
            d.eta = spectrum_id(obj("eta"), 0.f); d.k = spectrum_id(obj("k"), 1.f);
            if (obj("specular_reflectance")) d.specular_reflectance = spectrum_id(obj("specular_reflectance"), 1.f);
            if (cls == "RoughConductor") {
                // alpha / alpha_u / alpha_v are textures in the reference (roughconductor.cpp:396-401); radar scenes give constants
                ParamCollector au(obj("alpha_u") ? obj("alpha_u") : obj("alpha")), av(obj("alpha_v") ? obj("alpha_v") : obj("alpha"));
                d.alpha_u = (float) au.number("value", 0.1); d.alpha_v = (float) av.number("value", 0.1);
                d.distribution = p.string("distribution", "beckmann") == "ggx" ? BF_MICROFACET_GGX : BF_MICROFACET_BECKMANN;
                d.sample_visible = p.flag("sample_visible", true);
            }
        } else {
            NotImplementedError(("bsdf plugin \"" + cls + "\" on the B200 radar path").c_str());
        }
    }
    static void filter_desc(const ReconstructionFilter *f, float &radius, uint32_t &kind) {
        const std::string cls = f ? f->class_()->name() : "BoxFilter";
        radius = f ? (float) f->radius() : 0.5f;
        kind = cls == "BoxFilter" ? 0u : cls == "GaussianFilter" ? 1u : cls == "TentFilter" ? 2u : cls == "MitchellNetravaliFilter" ? 3u
             : cls == "CatmullRomFilter" ? 4u : cls == "LanczosSincFilter" ? 5u : 0u;
    }

    bf_scene *m_handle = nullptr;
    bool m_committed = false;
    std::unordered_map<const Shape *, uint32_t> m_shape_ids;
    std::unordered_map<const Object *, uint32_t> m_bsdf_ids, m_spectrum_ids;
};

NAMESPACE_END(b200)
NAMESPACE_END(mitsuba)
