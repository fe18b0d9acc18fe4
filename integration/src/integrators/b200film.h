// Shared body of b200range.cpp / b200time.cpp: the gen-1 / gen-2 integrators of the reference are wrappers that write their bins
// into AOV channels of a 1 x 1 film (range.cpp:141-161, time.cpp:118-153) and are driven by render() (integrator.cpp:53-166).
#pragma once
#include <mitsuba/render/film.h>
#include <mitsuba/render/imageblock.h>
#include <mitsuba/render/integrator.h>

#include "b200_flatten.h"

NAMESPACE_BEGIN(mitsuba)
NAMESPACE_BEGIN(b200)

/// render(): flatten once, run spp * pixels samples, hand the accumulated film to Film::put as one ImageBlock (integrator.cpp:150)
template <typename Float, typename Spectrum, typename Integrator>
bool render_through_b200(Integrator *self, Flattener<Float, Spectrum> &flat, const typename Integrator::Scene *scene,
                         typename Integrator::Sensor *sensor, const bf_integrator_desc &desc, int device,
                         const std::vector<std::string> &aov_names) {
    using Film = typename Integrator::Film;
    using ImageBlock = typename Integrator::ImageBlock;
    if (!flat.ready()) {
        flat.add_scene(scene, /*use_emitters*/ true);
        flat.set_sensor(sensor);
        flat.commit(desc, device);
    }
    Film *film = sensor->film();
    std::vector<std::string> channels = { "X", "Y", "Z", "A", "W" };           // integrator.cpp:88-92
    for (const std::string &n : aov_names) channels.push_back(n);
    film->prepare(channels);
    const uint64_t spp = (uint64_t) sensor->sampler()->sample_count();
    const uint64_t total = spp * (uint64_t) hprod(film->crop_size());
    flat.check(bf_receive(flat.handle(), (uint64_t) sensor->sampler()->seed(), 0, total, 0), "bf_receive");
    ref<ImageBlock> block = new ImageBlock(film->crop_size(), channels.size(), film->reconstruction_filter(), false, false, false);
    block->set_offset(film->crop_offset());
    block->clear();
    flat.check(bf_profile_read(flat.handle(), (float *) block->data().data(), block->data().size()), "bf_profile_read");
    film->put(block);
    return !self->should_stop();
}

NAMESPACE_END(b200)
NAMESPACE_END(mitsuba)
