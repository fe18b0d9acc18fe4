// <integrator type="b200phase"> — drop-in for <integrator type="phase"><integrator type="pathtimefrequency"/></integrator>
// (src/integrators/phase.cpp:63-145 around pathtimefrequency.cpp:103-477) on a B200: the receive() sample loop
// (src/librender/integrator.cpp:315-768, CPU branch :484-668) runs in libbeifong_b200.so; everything else of Mitsuba stays.
//
//   scene file   same XML; only the integrator's type changes, same parameters: max_depth (-1), rr_depth (5), bins (1)
//   Python       scene.integrator().receive(scene, scene.receivers()[0]); receiver.adc().bitmap(raw=True)   -- unchanged
//
// Interfaces implemented: include/mitsuba/render/integrator.h:43-44 (render / receive), :53 (cancel via the base's m_stop),
// :150 (aov_names).  The result goes back through the receiver's own ADC: ADC::prepare / ADC::put (adc.h:26-57), so
// HDRADC::bitmap / develop (hdradc.cpp:219-295) and the scripts that read them work as before.
#include <mitsuba/render/integrator.h>
#include <mitsuba/render/signalblock.h>

#include "b200_flatten.h"

NAMESPACE_BEGIN(mitsuba)

template <typename Float, typename Spectrum>
class B200PhaseIntegrator final : public SamplingIntegrator<Float, Spectrum> {
public:
    MTS_IMPORT_BASE(SamplingIntegrator, m_max_depth, m_rr_depth, should_stop)
    MTS_IMPORT_TYPES(Scene, Sensor, Receiver, ADC, SignalBlock, Sampler)

    B200PhaseIntegrator(const Properties &props) : Base(props) {
        m_bins = props.int_("bins", 1);                          // phase.cpp:78
        m_doppler = props.bool_("doppler", false);               // restores pathtimefrequency.cpp:252-256 (commented out upstream)
        m_device = props.int_("device", 0);
        m_slice = (uint64_t) props.long_("samples_per_slice", 1ll << 28);   // cancel() takes effect between slices
    }

    /// Radar scenes are received, not rendered (integrator.cpp:352-355 throws the same way for the opposite case)
    bool render(Scene *, Sensor *) override {
        Throw("b200phase: this integrator implements receive(); use b200range / b200time for film-based scenes");
    }

    /// Replaces SamplingIntegrator::receive (integrator.cpp:315-768)
    bool receive(Scene *scene, Receiver *receiver) override {
        this->m_stop = false;
        if (!m_flat.ready() || m_scene != scene || m_receiver != receiver) {
            m_flat.~Flattener(); new (&m_flat) b200::Flattener<Float, Spectrum>();
            m_flat.add_scene(scene, /*use_emitters*/ false);
            m_flat.set_receiver(receiver);
            bf_integrator_desc i;
            std::memset(&i, 0, sizeof(i));
            i.gen = BF_GEN_PHASE; i.max_depth = m_max_depth; i.rr_depth = m_rr_depth; i.phase_bins = (uint32_t) m_bins; i.doppler = m_doppler ? 1u : 0u;
            m_flat.commit(i, m_device);
            m_scene = scene; m_receiver = receiver;
        }
        ADC *adc = receiver->adc();
        std::vector<std::string> channels = { "Y", "A", "W" };                 // integrator.cpp:359-370
        for (const std::string &n : aov_names()) channels.push_back(n);
        adc->prepare(channels);
        Sampler *sampler = receiver->sampler();
        const uint64_t total = (uint64_t) sampler->sample_count();             // integrator.cpp:347: total_spp
        uint64_t done = 0;
        while (done < total && !should_stop()) {
            const uint64_t n = std::min<uint64_t>(m_slice, total - done);
            m_flat.check(bf_receive(m_flat.handle(), (uint64_t) sampler->seed(), done, n, done ? 1 : 0), "bf_receive");
            m_flat.check(bf_synchronize(m_flat.handle()), "bf_synchronize");
            done += n;
        }
        // adc->put(block): the accumulated window as ONE SignalBlock (integrator.cpp:664, signalblock.cpp:49-77)
        ref<SignalBlock> block = new SignalBlock(adc->window_size(), channels.size(), adc->reconstruction_filter(),
                                                 /*warn_negative*/ false, /*warn_invalid*/ false, /*border*/ false);
        block->set_offset(adc->window_offset());
        block->clear();
        m_flat.check(bf_profile_read(m_flat.handle(), (float *) block->data().data(), block->data().size()), "bf_profile_read");   // [f][t][channel] fp32
        adc->put(block);
        return !should_stop();
    }

    std::vector<std::string> aov_names() const override {                      // phase.cpp:140-145
        std::vector<std::string> r;
        for (int i = 0; i < m_bins; ++i) r.push_back("phase.bin" + std::to_string(i));
        return r;
    }

    std::string to_string() const override {
        return tfm::format("B200PhaseIntegrator[\n  max_depth = %i,\n  rr_depth = %i,\n  bins = %i,\n  doppler = %i\n]", m_max_depth, m_rr_depth, m_bins, (int) m_doppler);
    }

    MTS_DECLARE_CLASS()
private:
    b200::Flattener<Float, Spectrum> m_flat;
    const Scene *m_scene = nullptr;
    const Receiver *m_receiver = nullptr;
    int m_bins, m_device;
    bool m_doppler;
    uint64_t m_slice;
};

MTS_IMPLEMENT_CLASS_VARIANT(B200PhaseIntegrator, SamplingIntegrator)
MTS_EXPORT_PLUGIN(B200PhaseIntegrator, "B200 phase / path-time-frequency integrator")
NAMESPACE_END(mitsuba)
