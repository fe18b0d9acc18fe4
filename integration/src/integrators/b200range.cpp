// <integrator type="b200range"> — drop-in for range(pathlength) (src/integrators/range.cpp:63-221 around pathlength.cpp:114-342) on a B200.
// Same parameters as the reference's wrapper (dr, bins: range.cpp:74-75) + max_depth, rr_depth of the inner path tracer.
#include "b200film.h"

NAMESPACE_BEGIN(mitsuba)

template <typename Float, typename Spectrum>
class B200RangeIntegrator final : public SamplingIntegrator<Float, Spectrum> {
public:
    MTS_IMPORT_BASE(SamplingIntegrator, m_max_depth, m_rr_depth, should_stop)
    MTS_IMPORT_TYPES(Scene, Sensor, Receiver, Film, ImageBlock, Sampler)

    B200RangeIntegrator(const Properties &props) : Base(props) {
        m_dr = props.float_("dr", 1.f); m_bins = props.int_("bins", 1);                  // range.cpp:74-75
        m_device = props.int_("device", 0);
    }

    bool render(Scene *scene, Sensor *sensor) override {
        this->m_stop = false;
        bf_integrator_desc i;
        std::memset(&i, 0, sizeof(i));
        i.gen = BF_GEN_RANGE; i.max_depth = m_max_depth; i.rr_depth = m_rr_depth; i.range_bins = (uint32_t) m_bins; i.dr = m_dr;
        return b200::render_through_b200<Float, Spectrum>(this, m_flat, scene, sensor, i, m_device, aov_names());
    }
    bool receive(Scene *, Receiver *) override { Throw("b200range: this integrator implements render(); use b200phase for receive()"); }

    std::vector<std::string> aov_names() const override {
        std::vector<std::string> r;
        for (int i = 0; i < m_bins; ++i) r.push_back("bin" + std::to_string(i));          // range.cpp:216-221 (mono / spectral: one channel per bin)
        return r;
    }

    MTS_DECLARE_CLASS()
private:
    b200::Flattener<Float, Spectrum> m_flat;
    float m_dr; int m_bins, m_device;
};

MTS_IMPLEMENT_CLASS_VARIANT(B200RangeIntegrator, SamplingIntegrator)
MTS_EXPORT_PLUGIN(B200RangeIntegrator, "B200 range integrator")
NAMESPACE_END(mitsuba)
