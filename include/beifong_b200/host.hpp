// beifong_b200 — C++ host layer over the C ABI (include/beifong_b200/abi.h).
//
// The reference is compiled C++ (Mitsuba 2 plugins); this header mirrors, for the radar path only, the
// interface a caller of the reference sees, with the same names, argument meaning and error behaviour:
//
//   reference                                                   here
//   ---------------------------------------------------------   ------------------------------------------
//   Throw(...) -> std::runtime_error   (core/logger.h:134-138)  beifong::Error
//   NotImplementedError("name")                                 beifong::NotImplementedError
//   Scene (librender/scene.cpp:22-120): shapes, endpoints       beifong::Scene::add_* / set_*
//   Receiver::adc(), adc_sampling_start(), receive_type()       beifong::Receiver (receiver.h:61-112)
//   ADC::size(), bandwidth(), prepare(), bitmap(raw)            beifong::ADC      (adc.h:26-97, hdradc.cpp:219-233)
//   Integrator::receive(scene, receiver) -> bool, cancel(),     beifong::Integrator (integrator.h:43-53,116-158;
//   aov_names(); sampler sample_count / seed                    integrator.cpp:315-768, phase.cpp:140-145)
//
// Header only, C++17, no CUDA or torch types; link with libbeifong_b200.so.  There is no CPU fallback behind it:
// Scene::commit() throws when no CUDA device is present.  A Mitsuba-side plugin is this layer with
// the reference's own Scene / Receiver / ADC objects on top.
#pragma once
#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "abi.h"

namespace beifong {

struct Error : std::runtime_error { using std::runtime_error::runtime_error; };
struct NotImplementedError : Error { using Error::Error; };

using Matrix4f = float[16];   // row major, include/mitsuba/core/transform.h

// Physical constants of a variant build (spectrum.h:15-40): the 94 GHz set used by the radar scripts, or the defaults.
inline bf_constants constants(float c, float lambda_min_nm, float lambda_max_nm, uint32_t spectrum_width = 1) {
    bf_constants k;
    std::memset(&k, 0, sizeof(k));
    k.c = c; k.lambda_min_nm = lambda_min_nm; k.lambda_max_nm = lambda_max_nm; k.cie_min_nm = 360.f; k.cie_max_nm = 830.f;
    k.P = 1.f; k.spectrum_width = spectrum_width;
    return k;
}

class Scene;

// ADC (src/librender/adc.cpp, src/adcs/hdradc.cpp): geometry of the accumulation buffer + access to it.
class ADC {
public:
    ADC() { std::memset(&m_desc, 0, sizeof(m_desc)); m_desc.t_bins = m_desc.f_bins = 1024; m_desc.window_t_bins = m_desc.window_f_bins = 1024;
            m_desc.t_bandwidth = 3.81e-6f; m_desc.f_bandwidth = 250e6f; m_desc.filter_radius = 0.5f; }
    ADC(uint32_t t_bins, uint32_t f_bins, float t_bandwidth, float f_bandwidth) : ADC() {
        m_desc.t_bins = m_desc.window_t_bins = t_bins; m_desc.f_bins = m_desc.window_f_bins = f_bins;
        m_desc.t_bandwidth = t_bandwidth; m_desc.f_bandwidth = f_bandwidth;
    }
    void set_window(uint32_t t_bins, uint32_t f_bins, uint32_t offset_t, uint32_t offset_f) {
        m_desc.window_t_bins = t_bins; m_desc.window_f_bins = f_bins; m_desc.window_offset_t = offset_t; m_desc.window_offset_f = offset_f;
    }
    // <rfilter>: "box" (default here; the reference's scripts all use it), "gaussian" (radius = 4 stddev, gaussian.cpp:37-45),
    // "tent" (radius 1, tent.cpp:30-34); the wide ones take the n x n splat of SignalBlock::put (signalblock.cpp:113-161)
    void set_reconstruction_filter(const std::string &type, float stddev = 0.5f, int lobes = 3) {
        if (type == "box") { m_desc.filter_kind = 0; m_desc.filter_radius = 0.5f; }
        else if (type == "gaussian") { m_desc.filter_kind = 1; m_desc.filter_radius = 4.f * stddev; }
        else if (type == "tent") { m_desc.filter_kind = 2; m_desc.filter_radius = 1.f; }
        else if (type == "mitchell") { m_desc.filter_kind = 3; m_desc.filter_radius = 2.f; }       // B = C = 1/3
        else if (type == "catmullrom") { m_desc.filter_kind = 4; m_desc.filter_radius = 2.f; }
        else if (type == "lanczos") { m_desc.filter_kind = 5; m_desc.filter_radius = (float) lobes; }
        else throw NotImplementedError("rfilter plugin \"" + type + "\"");
    }
    uint32_t t_bins() const { return m_desc.t_bins; }
    uint32_t f_bins() const { return m_desc.f_bins; }
    float t_bandwidth() const { return m_desc.t_bandwidth; }
    float f_bandwidth() const { return m_desc.f_bandwidth; }
    const bf_adc_desc &desc() const { return m_desc; }
    // ADC::prepare(channels): the channel names Integrator::receive hands over ("Y", "A", "W" + aov_names())
    void prepare(const std::vector<std::string> &channels) { m_channels = channels; }
    const std::vector<std::string> &channels() const { return m_channels; }
    // ADC::bitmap(raw = true): the un-normalised fp32 buffer [f][t][channel] (hdradc.cpp:219-233)
    // ADC::bitmap(raw = false): every channel times 1.f / weight, weight channel dropped (hdradc.cpp:235-275), developed on
    // the device from the profile of the last receive(); fp32 components here, bitmap_f16() for the reference's default.
    const std::vector<float> &bitmap(bool raw = true) {
        if (raw) return m_storage;
        if (!m_source) throw Error("ADC::bitmap(raw=false): nothing received yet");
        const size_t n = (size_t) m_w * m_h * (m_channels.size() - 1);
        m_developed.resize(n);
        if (int rc = bf_profile_develop(m_source, BF_COMPONENT_FLOAT32, m_developed.data(), n * sizeof(float)))
            throw_rc(rc, "bf_profile_develop");
        return m_developed;
    }
    std::vector<uint16_t> bitmap_f16() {
        if (!m_source) throw Error("ADC::bitmap(raw=false): nothing received yet");
        std::vector<uint16_t> out((size_t) m_w * m_h * (m_channels.size() - 1));
        if (int rc = bf_profile_develop(m_source, BF_COMPONENT_FLOAT16, out.data(), out.size() * sizeof(uint16_t)))
            throw_rc(rc, "bf_profile_develop");
        return out;
    }
    uint32_t width() const { return m_w; }
    uint32_t height() const { return m_h; }

private:
    friend class Integrator;
    bf_adc_desc m_desc;
    std::vector<std::string> m_channels;
    std::vector<float> m_storage, m_developed;
    uint32_t m_w = 0, m_h = 0;
    bf_scene *m_source = nullptr;      // scene of the last receive() (must outlive the ADC's developed reads)
    void throw_rc(int rc, const char *what) const {
        const char *msg = bf_last_error(m_source);
        std::string text = std::string(what) + ": " + (msg && *msg ? msg : "error");
        if (rc == BF_ERR_UNSUPPORTED) throw NotImplementedError(text);
        throw Error(text);
    }
};

// Signal block shared by the wigner / phased endpoints (wignertransmitter.cpp:59-102): same property names.
struct Signal {
    std::string signaltype = "linfmcw";     // "cw" | "linfmcw" | "pulse"
    float amplitude = 1.f, repfreq = 5e3f, length = 200e-6f, freq_centre = 94e9f, freq_ext = 6e9f, phase = 0.f, gain = 1.f;
    bool sig_is_delta = true, resample_freq = false;
    bf_signal_desc desc() const {
        bf_signal_desc d;
        std::memset(&d, 0, sizeof(d));
        if (signaltype == "cw") d.signaltype = BF_SIG_CW;
        else if (signaltype == "linfmcw") d.signaltype = BF_SIG_LINFMCW;
        else if (signaltype == "pulse") d.signaltype = BF_SIG_PULSE;
        else d.signaltype = BF_SIG_OTHER;
        d.sig_is_delta = sig_is_delta; d.resample_freq = resample_freq; d.amplitude = amplitude; d.repfreq = repfreq; d.t_ext = length;
        d.f_centre = freq_centre; d.f_ext = freq_ext; d.phi0 = phase; d.gain = gain;
        return d;
    }
};

// Phased array block of `phasedtransmitter` / `phasedreceiver` (phasedtransmitter.cpp:104-117): same property names.
struct Array {
    uint32_t n_elems = 1;
    float steering_vector[3] = {0.f, 0.f, 0.f};
    Matrix4f array_loc = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    float elem_dims[3] = {1.f, 1.f, 1.f}, elem_spacing[3] = {0.f, 0.f, 0.f}, elem_axis[3] = {0.f, 1.f, 0.f};
    bf_array_desc desc() const {
        bf_array_desc d;
        std::memset(&d, 0, sizeof(d));
        d.n_elems = n_elems;
        std::memcpy(d.steering_vector, steering_vector, 12); std::memcpy(d.array_loc, array_loc, 64);
        std::memcpy(d.elem_dims, elem_dims, 12); std::memcpy(d.elem_spacing, elem_spacing, 12); std::memcpy(d.elem_axis, elem_axis, 12);
        return d;
    }
};

// Receiver (src/librender/receiver.cpp:16-62 + src/receivers/*.cpp): owns its ADC like the reference's does.
class Receiver {
public:
    // type: "omnidirectional" | "wignerreceiver" | "phasedreceiver" (the plugin names of src/receivers); receive_type as receiver.cpp:44-52
    Receiver(const std::string &type, uint32_t shape_id, const ADC &adc, const Signal &signal = Signal(), const std::string &receive_type = "raw",
             float adc_sampling_start = 0.f, float adc_sampling_end = 0.f, uint32_t antenna_texture = 0, const Array *array = nullptr)
        : m_shape(shape_id), m_adc(adc), m_receive_type(receive_type) {
        std::memset(&m_desc, 0, sizeof(m_desc));
        if (type == "omnidirectional") m_desc.kind = BF_RX_OMNIDIRECTIONAL;
        else if (type == "wignerreceiver") m_desc.kind = BF_RX_WIGNER;
        else if (type == "phasedreceiver") {
            if (!array) throw Error("phasedreceiver: the array parameters (n_elems, steering_vector, array_loc, elem_dims, elem_spacing, elem_axis) are required");
            m_desc.kind = BF_RX_PHASED; m_desc.array = array->desc();
        }
        else throw NotImplementedError("receiver plugin \"" + type + "\"");
        if (receive_type == "raw") m_desc.receive_type = BF_RECV_RAW;
        else if (receive_type == "raw_resample") m_desc.receive_type = BF_RECV_RAW_RESAMPLE;
        else if (receive_type == "mix_resample") m_desc.receive_type = BF_RECV_MIX_RESAMPLE;
        else if (receive_type == "mixer") m_desc.receive_type = BF_RECV_MIXER;
        else throw Error("Receiver: unknown receive_type \"" + receive_type + "\"");
        if (adc_sampling_end - adc_sampling_start < 0.f)   // receiver.cpp:23-25
            throw Error("ADC sampling time must be less than or equal to the adc sampling end time!");
        m_desc.adc_sampling_start = adc_sampling_start; m_desc.adc_sampling_time = adc_sampling_end - adc_sampling_start;
        m_desc.signal = signal.desc(); m_desc.texture = antenna_texture;
        const float eye[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
        std::memcpy(m_desc.to_world, eye, 64); std::memcpy(m_desc.sample_to_camera, eye, 64);
        m_desc.film_width = m_desc.film_height = 1;
    }
    float adc_sampling_start() const { return m_desc.adc_sampling_start; }
    float adc_sampling_time() const { return m_desc.adc_sampling_time; }
    const std::string &receive_type() const { return m_receive_type; }
    ADC *adc() { return &m_adc; }
    uint32_t shape() const { return m_shape; }
    const bf_endpoint_desc &desc() const { return m_desc; }

private:
    bf_endpoint_desc m_desc;
    uint32_t m_shape;
    ADC m_adc;
    std::string m_receive_type;
};

// Scene (src/librender/scene.cpp:22-120): what the plugin tree flattens to.  Shapes, BSDFs and spectra are referred to
// by the ids the add_* calls return.
class Scene {
public:
    explicit Scene(const bf_constants &k) { check(bf_scene_create(&k, &m_h), "bf_scene_create"); }
    ~Scene() { if (m_h) bf_scene_destroy(m_h); }
    Scene(const Scene &) = delete;
    Scene &operator=(const Scene &) = delete;

    uint32_t add_uniform_spectrum(float value) {
        uint32_t id; check(bf_scene_add_spectrum(m_h, BF_SPEC_UNIFORM, 1, nullptr, &value, &id), "bf_scene_add_spectrum"); return id;
    }
    // <bsdf type="diffuse"> optionally inside <bsdf type="twosided">
    uint32_t add_diffuse(float reflectance, bool two_sided) {
        bf_bsdf_desc d = bsdf_defaults(); d.kind = BF_BSDF_DIFFUSE; d.two_sided = two_sided; d.reflectance = add_uniform_spectrum(reflectance);
        uint32_t id; check(bf_scene_add_bsdf(m_h, &d, &id), "bf_scene_add_bsdf"); return id;
    }
    // <bsdf type="roughconductor"> (Beckmann, sample_visible): eta / k spectra by id
    uint32_t add_roughconductor(uint32_t eta, uint32_t k, float alpha, bool two_sided) {
        bf_bsdf_desc d = bsdf_defaults(); d.kind = BF_BSDF_ROUGHCONDUCTOR; d.two_sided = two_sided; d.eta = eta; d.k = k; d.alpha_u = d.alpha_v = alpha;
        uint32_t id; check(bf_scene_add_bsdf(m_h, &d, &id), "bf_scene_add_bsdf"); return id;
    }
    // <shape type="rectangle"> with its to_world
    uint32_t add_rectangle(const Matrix4f to_world, uint32_t bsdf) {
        uint32_t id; check(bf_scene_add_rect(m_h, to_world, bsdf, &id), "bf_scene_add_rect"); return id;
    }
    // <shape type="ply|obj">: world-space vertex positions (3V), faces (3F), optional vertex normals (3V)
    uint32_t add_mesh(const std::vector<float> &positions, const std::vector<uint32_t> &faces, const std::vector<float> &normals, uint32_t bsdf) {
        uint32_t id;
        check(bf_scene_add_mesh(m_h, positions.data(), (uint32_t) (positions.size() / 3), faces.data(), (uint32_t) (faces.size() / 3),
                                normals.empty() ? nullptr : normals.data(), bsdf, &id), "bf_scene_add_mesh");
        return id;
    }
    // <transmitter type="wignertransmitter|phasedtransmitter|areatransmitter"> attached to a rectangle
    uint32_t add_transmitter(const std::string &type, uint32_t shape_id, const Signal &signal, uint32_t antenna_texture, const Array *array = nullptr) {
        bf_endpoint_desc d;
        std::memset(&d, 0, sizeof(d));
        if (type == "wignertransmitter") d.kind = BF_TX_WIGNER;
        else if (type == "areatransmitter") d.kind = BF_TX_AREA;
        else if (type == "phasedtransmitter") {
            if (!array) throw Error("phasedtransmitter: the array parameters are required");
            d.kind = BF_TX_PHASED; d.array = array->desc();
        }
        else throw NotImplementedError("transmitter plugin \"" + type + "\"");
        d.signal = signal.desc(); d.texture = antenna_texture;
        uint32_t id; check(bf_scene_add_transmitter(m_h, shape_id, &d, &id), "bf_scene_add_transmitter"); return id;
    }
    void set_receiver(const Receiver &rx) {
        check(bf_scene_set_receiver(m_h, rx.shape(), &rx.desc()), "bf_scene_set_receiver");
    }
    // Moving geometry between frames (animated_trans_rad.py rebuilds the scene; here the rectangle is re-posed in place)
    void update_rectangle(uint32_t shape_id, const Matrix4f to_world) { check(bf_scene_update_rect(m_h, shape_id, to_world), "bf_scene_update_rect"); }
    // <transform name="velocity"> of a shape (shape.cpp:39-42); read by Shape::doppler when the integrator's doppler flag is set
    void set_shape_velocity(uint32_t shape_id, const Matrix4f velocity) { check(bf_scene_set_shape_velocity(m_h, shape_id, velocity), "bf_scene_set_shape_velocity"); }
    // A mesh moved rigidly from the pose it was added in: goes into the build before commit(), a device refit (no rebuild) after it
    void update_transform(uint32_t shape_id, const Matrix4f motion) { check(bf_scene_update_transform(m_h, shape_id, motion), "bf_scene_update_transform"); }
    // SAH build + upload; throws when there is no CUDA device (no CPU fallback)
    void commit(int device = 0) { check(bf_scene_commit(m_h, device), "bf_scene_commit"); m_committed = true; }
    bool committed() const { return m_committed; }
    // every later call of this scene is enqueued on the application's stream (a cudaStream_t; nullptr: the library's own)
    void set_stream(void *cuda_stream) { check(bf_scene_set_stream(m_h, cuda_stream), "bf_scene_set_stream"); }
    bf_scene *handle() { return m_h; }

    void check(int rc, const char *what) const {
        if (rc == BF_OK) return;
        const char *msg = bf_last_error(m_h);
        std::string text = std::string(what) + ": " + (msg && *msg ? msg : "error");
        if (rc == BF_ERR_UNSUPPORTED) throw NotImplementedError(text);
        throw Error(text);
    }

private:
    static bf_bsdf_desc bsdf_defaults() {
        bf_bsdf_desc d;
        std::memset(&d, 0, sizeof(d));
        d.back_bsdf = BF_NONE; d.specular_reflectance = BF_NONE; d.distribution = BF_MICROFACET_BECKMANN; d.sample_visible = 1;
        d.alpha_u = d.alpha_v = 0.1f;
        return d;
    }
    bf_scene *m_h = nullptr;
    bool m_committed = false;
};

// Integrator: `phase` wrapping `pathtimefrequency` (src/integrators/phase.cpp:63-145, pathtimefrequency.cpp:103-477)
// behind SamplingIntegrator::receive (src/librender/integrator.cpp:315-768).
class Integrator {
public:
    // XML parameters of the reference, same names and defaults (integrator.cpp:1713-1728, phase.cpp:78)
    explicit Integrator(int max_depth = -1, int rr_depth = 5, uint32_t bins = 1) : m_stop(false) {
        if (max_depth < 0 && max_depth != -1) throw Error("\"max_depth\" must be set to -1 (infinite) or a value >= 0");   // integrator.cpp:1717-1719
        if (rr_depth <= 0) throw Error("\"rr_depth\" must be set to a value greater than zero!");                             // :1722-1723
        std::memset(&m_desc, 0, sizeof(m_desc));
        m_desc.gen = BF_GEN_PHASE; m_desc.max_depth = max_depth; m_desc.rr_depth = rr_depth; m_desc.phase_bins = bins;
    }
    // phase.cpp:140-145
    std::vector<std::string> aov_names() const {
        std::vector<std::string> names;
        for (uint32_t i = 0; i < m_desc.phase_bins; ++i) names.push_back("phase.bin" + std::to_string(i));
        return names;
    }
    const bf_integrator_desc &desc() const { return m_desc; }
    // Restores the Doppler statement the reference has commented out (pathtimefrequency.cpp:252-256); off by default.  Before prepare().
    void set_doppler(bool on) { m_desc.doppler = on ? 1u : 0u; }

    // Configure the scene for this integrator and receiver (the part of Scene construction that depends on them) and
    // commit it.  Call once; receive() may then be called any number of times.
    void prepare(Scene &scene, Receiver &receiver, int device = 0) {
        scene.set_receiver(receiver);
        scene.check(bf_scene_set_adc(scene.handle(), &receiver.adc()->desc()), "bf_scene_set_adc");
        scene.check(bf_scene_set_integrator(scene.handle(), &m_desc), "bf_scene_set_integrator");
        scene.commit(device);
    }

    // SamplingIntegrator::receive(scene, receiver): total_spp = sample_count samples with sampler seed `seed`, accumulated into
    // receiver->adc().  Returns false when cancel() stopped it (integrator.cpp:767).  Samples are issued in slices so
    // that cancel() takes effect between slices, like should_stop() polled per sample (integrator.h:168-171).
    bool receive(Scene &scene, Receiver &receiver, uint64_t sample_count, uint64_t seed = 0, uint64_t slice = 1ull << 28) {
        if (!scene.committed()) throw Error("Integrator::receive: call prepare() first");
        m_stop = false;
        ADC *adc = receiver.adc();
        std::vector<std::string> channels = {"Y", "A", "W"};               // integrator.cpp:484-489
        for (const std::string &n : aov_names()) channels.push_back(n);
        adc->prepare(channels);
        uint64_t done = 0;
        bool first = true;
        while (done < sample_count && !m_stop) {
            const uint64_t n = std::min<uint64_t>(slice, sample_count - done);
            scene.check(bf_receive(scene.handle(), seed, done, n, first ? 0 : 1), "bf_receive");
            scene.check(bf_synchronize(scene.handle()), "bf_synchronize");
            done += n; first = false;
        }
        uint32_t w = 0, h = 0, c = 0;
        scene.check(bf_profile_size(scene.handle(), &w, &h, &c), "bf_profile_size");
        adc->m_w = w; adc->m_h = h; adc->m_source = scene.handle();
        adc->m_storage.resize((size_t) w * h * c);
        scene.check(bf_profile_read(scene.handle(), adc->m_storage.data(), adc->m_storage.size()), "bf_profile_read");   // adc->put(block)
        m_samples_done = done;
        return !m_stop;
    }
    // Many receive() calls as one submission (the frames of a sequence, the receivers of a scene: scene.h:197 holds a vector of them):
    // item i re-poses rectangles and / or the receiver (bf_batch_item), then integrates `sample_count` samples into profile i.
    // Returns [item][f][t][channel]; the receiver's ADC is not touched.
    std::vector<float> receive_batch(Scene &scene, const std::vector<bf_batch_item> &items, uint64_t sample_count) {
        if (!scene.committed()) throw Error("Integrator::receive_batch: call prepare() first");
        uint32_t w = 0, h = 0, c = 0;
        scene.check(bf_profile_size(scene.handle(), &w, &h, &c), "bf_profile_size");
        std::vector<float> out((size_t) items.size() * w * h * c);
        scene.check(bf_receive_batch(scene.handle(), (uint32_t) items.size(), items.data(), sample_count, out.data(), out.size()), "bf_receive_batch");
        return out;
    }
    void cancel() { m_stop = true; }                 // integrator.cpp:48-51: only sets m_stop
    bool should_stop() const { return m_stop; }
    uint64_t samples_done() const { return m_samples_done; }

private:
    bf_integrator_desc m_desc;
    std::atomic<bool> m_stop;
    uint64_t m_samples_done = 0;
};

}  // namespace beifong
