// <integrator type="b200time"> — drop-in for time(pathtime) (src/integrators/time.cpp:86-206 around pathtime.cpp:114-302) on a B200.
// Same parameters as the reference's wrapper (50 bins of 0.5 ns are hard-coded: time.cpp:118) + max_depth, rr_depth of the inner path tracer.
#include "b200film.h"

NAMESPACE_BEGIN(mitsuba)

template <typename Float, typename Spectrum>
class B200TimeIntegrator final : public SamplingIntegrator<Float, Spectrum> {
public:
    MTS_IMPORT_BASE(SamplingIntegrator, m_max_depth, m_rr_depth, should_stop)
    MTS_IMPORT_TYPES(Scene, Sensor, Receiver, Film, ImageBlock, Sampler)

    B200TimeIntegrator(const Properties &props) : Base(props) {
        m_dr = 0.f; m_bins = 50;                                                          // time.cpp:118
        m_device = props.int_("device", 0);
    }

    bool render(Scene *scene, Sensor *sensor) override {
        this->m_stop = false;
        bf_integrator_desc i;
        std::memset(&i, 0, sizeof(i));
        i.gen = BF_GEN_TIME; i.max_depth = m_max_depth; i.rr_depth = m_rr_depth; i.range_bins = (uint32_t) m_bins; i.dr = m_dr;
        return b200::render_through_b200<Float, Spectrum>(this, m_flat, scene, sensor, i, m_device, aov_names());
    }
    bool receive(Scene *, Receiver *) override { Throw("b200time: this integrator implements render(); use b200phase for receive()"); }

    std::vector<std::string> aov_names() const override {
        std::vector<std::string> r;
        for (int i = 0; i < m_bins; ++i) for (const char *c : { ".X", ".Y", ".Z" }) r.push_back("bin" + std::to_string(i) + c);   // time.cpp:200-206
        return r;
    }

    MTS_DECLARE_CLASS()
private:
    b200::Flattener<Float, Spectrum> m_flat;
    float m_dr; int m_bins, m_device;
};

MTS_IMPLEMENT_CLASS_VARIANT(B200TimeIntegrator, SamplingIntegrator)
MTS_EXPORT_PLUGIN(B200TimeIntegrator, "B200 time integrator")
NAMESPACE_END(mitsuba)
