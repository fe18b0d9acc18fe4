// Test program for the C++ host layer (include/beifong_b200/host.hpp): a 94 GHz FMCW radar scene built through the classes
// that mirror the reference's Scene / Receiver / ADC / Integrator, received on the GPU, checked against the CPU oracle.
//
//   host_radar --host-only            error behaviour of the host layer only (runs without a GPU)
//   host_radar <path/to/liboracle.so> full run: receive() on cuda:0 vs the oracle on the exported scene, then cancel()
//
// The oracle is loaded with dlopen here, as the checker; nothing in the host layer knows about it.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <dlfcn.h>
#include <thread>

#include "beifong_b200/host.hpp"

using namespace beifong;

static int failures = 0;
#define EXPECT(cond) do { if (!(cond)) { std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); ++failures; } } while (0)

template <class E, class F> static bool throws(F &&f, const char *needle = nullptr) {
    try { f(); } catch (const E &e) { return !needle || std::string(e.what()).find(needle) != std::string::npos; } catch (...) { return false; }
    return false;
}

static const float C0 = 299792458.f;

static void build(Scene &scene, uint32_t &rx_shape) {
    const uint32_t default_bsdf = scene.add_diffuse(0.5f, false);
    const uint32_t ground = scene.add_diffuse(0.5f, true);
    const uint32_t eta = scene.add_uniform_spectrum(0.f), k = scene.add_uniform_spectrum(1.f);
    const uint32_t metal = scene.add_roughconductor(eta, k, 0.1f, true);
    // 5 mm x 5 mm apertures at (0, 0, 0.3) facing +x: translate * rotate_y(90) * scale(0.0025, 0.0025, 1)
    const float rx_m[16] = {0, 0, 1, 0,  0, 0.0025f, 0, 0,  -0.0025f, 0, 0, 0.3f,  0, 0, 0, 1};
    const float tx_m[16] = {0, 0, 1, 0,  0, 0.0025f, 0, 0.01f,  -0.0025f, 0, 0, 0.3f,  0, 0, 0, 1};
    const float gnd_m[16] = {20, 0, 0, 0,  0, 20, 0, 0,  0, 0, 1, 0,  0, 0, 0, 1};
    rx_shape = scene.add_rectangle(rx_m, default_bsdf);
    const uint32_t tx_shape = scene.add_rectangle(tx_m, default_bsdf);
    scene.add_rectangle(gnd_m, ground);
    // a 4 m x 3 m metal plate 8 m ahead, slightly tilted, as a 2 x 2-cell grid (8 triangles)
    std::vector<float> pos; std::vector<uint32_t> idx;
    for (int j = 0; j < 3; ++j) for (int i = 0; i < 3; ++i) { pos.push_back(8.f + 0.1f * i); pos.push_back(-2.f + 2.f * i); pos.push_back(0.2f + 1.5f * j); }
    for (int j = 0; j < 2; ++j) for (int i = 0; i < 2; ++i) {
        uint32_t a = j * 3 + i, b = a + 1, c = a + 4, d = a + 3;
        idx.insert(idx.end(), {a, b, c, a, c, d});
    }
    scene.add_mesh(pos, idx, {}, metal);
    Signal tx_sig; tx_sig.repfreq = 1.f / 200e-6f; tx_sig.resample_freq = true;
    scene.add_transmitter("wignertransmitter", tx_shape, tx_sig, scene.add_uniform_spectrum(1.f));
}

int main(int argc, char **argv) {
    const bool host_only = argc > 1 && std::string(argv[1]) == "--host-only";
    // ---- error behaviour (the reference's Throw / NotImplementedError sites) ----
    EXPECT((throws<NotImplementedError>([] { ADC a; Receiver r("irradiancemeter", 0, a); }, "irradiancemeter")));
    EXPECT((throws<Error>([] { ADC a; Receiver r("wignerreceiver", 0, a, Signal(), "raw", 2e-6f, 1e-6f); }, "ADC sampling time")));
    EXPECT((throws<Error>([] { ADC a; Receiver r("wignerreceiver", 0, a, Signal(), "sideways"); }, "receive_type")));
    EXPECT((throws<Error>([] { Integrator i(4, 0); }, "rr_depth")));
    EXPECT((throws<Error>([] { Integrator i(-3); }, "max_depth")));
    EXPECT((throws<Error>([] { ADC a; a.bitmap(false); }, "nothing received")));
    EXPECT((throws<NotImplementedError>([] { ADC a; a.set_reconstruction_filter("sinc"); }, "sinc")));
    { ADC a; a.set_reconstruction_filter("gaussian"); EXPECT(a.desc().filter_kind == 1 && a.desc().filter_radius == 2.f); }
    {
        Scene s(constants(C0, C0 / 97e9f * 1e9f, C0 / 91e9f * 1e9f));
        EXPECT((throws<Error>([&] { const float m[16] = {1,0,0,0, 0,1,0,0, 0,0,1,0, 0,0,0,1}; s.add_rectangle(m, 99); }, "bsdf id out of range")));
        EXPECT((throws<NotImplementedError>([&] { s.add_transmitter("laser", 0, Signal(), 0); }, "laser")));
        Integrator integ(4);
        EXPECT((integ.aov_names() == std::vector<std::string>{"phase.bin0"}));
        ADC adc(1, 1024, 200e-6f, 1.f);
        Receiver rx("wignerreceiver", 0, adc);
        EXPECT((throws<Error>([&] { integ.receive(s, rx, 16); }, "prepare")));
    }
    if (host_only) {
        // without a device commit must fail loudly: there is no CPU fallback
        Scene s(constants(C0, C0 / 97e9f * 1e9f, C0 / 91e9f * 1e9f));
        uint32_t rx_shape; build(s, rx_shape);
        const float sweep_rate = 6e9f / 200e-6f;
        ADC adc(1, 1024, 200e-6f, 1024 * 0.03f * 2.f / C0 * sweep_rate);
        Receiver rx("wignerreceiver", rx_shape, adc, Signal(), "mix_resample", 20e-6f, 180e-6f, 0);
        Integrator integ(4);
        bool committed = true;
        try { integ.prepare(s, rx); } catch (const Error &e) { committed = false; EXPECT(std::string(e.what()).find("no CUDA device") != std::string::npos); }
        std::printf("host-only: %d failure(s)%s\n", failures, committed ? " (a device was present: commit succeeded)" : "");
        return failures ? 1 : 0;
    }

    // ---- full run ----
    Scene scene(constants(C0, C0 / 97e9f * 1e9f, C0 / 91e9f * 1e9f));
    uint32_t rx_shape; build(scene, rx_shape);
    const float sweep_rate = 6e9f / 200e-6f;
    ADC adc(1, 1024, 200e-6f, 1024 * 0.03f * 2.f / C0 * sweep_rate);
    Signal rx_sig; rx_sig.repfreq = 1.f / 200e-6f;
    Receiver rx("wignerreceiver", rx_shape, adc, rx_sig, "mix_resample", 20e-6f, 180e-6f, scene.add_uniform_spectrum(1.f));
    Integrator integ(4, 5, 1);
    integ.prepare(scene, rx, 0);
    const uint64_t n = 1 << 18;
    EXPECT(integ.receive(scene, rx, n, 7));
    const std::vector<float> &g = rx.adc()->bitmap(true);
    EXPECT(rx.adc()->channels().size() == 4 && rx.adc()->channels()[3] == "phase.bin0");
    EXPECT(g.size() == 1024 * 4);

    // the oracle on the exported scene
    void *lib = argc > 1 ? dlopen(argv[1], RTLD_NOW) : nullptr;
    if (!lib) { std::printf("cannot load the oracle (%s)\n", argc > 1 ? dlerror() : "no path given"); return 2; }
    auto orc_import = (void *(*)(const void *, size_t)) dlsym(lib, "orc_scene_import");
    auto orc_accel = (int (*)(void *, int, const char *)) dlsym(lib, "orc_scene_set_accel");
    auto orc_receive = (int (*)(void *, uint64_t, uint64_t, uint64_t, float *, int, int, int, uint64_t *)) dlsym(lib, "orc_receive");
    size_t need = 0;
    bf_scene_export(scene.handle(), nullptr, 0, &need);
    std::vector<uint8_t> blob(need);
    scene.check(bf_scene_export(scene.handle(), blob.data(), blob.size(), &need), "bf_scene_export");
    void *orc = orc_import(blob.data(), blob.size());
    EXPECT(orc != nullptr);
    orc_accel(orc, 1, nullptr);
    std::vector<float> c(g.size(), 0.f);
    uint64_t queries = 0;
    orc_receive(orc, 7, 0, n, c.data(), 0, 0, 8, &queries);
    double num = 0, den = 0, wg = 0, wc = 0;
    for (size_t i = 0; i < g.size(); i += 4) { num += (double) (g[i] - c[i]) * (g[i] - c[i]); den += (double) c[i] * c[i]; wg += g[i + 2]; wc += c[i + 2]; }
    const double rel = std::sqrt(num / std::max(den, 1e-300));
    std::printf("receive: %llu samples, weight channel %.0f vs %.0f, range-profile rel L2 vs oracle %.3e\n", (unsigned long long) n, wg, wc, rel);
    EXPECT(wg == wc);
    EXPECT(den > 0 && rel <= 1e-3);

    // bitmap(raw = false): channels Y, A, phase.bin0 = raw * (1.f / W)
    {
        const std::vector<float> &d = rx.adc()->bitmap(false);
        EXPECT(d.size() == 1024 * 3);
        bool same = true;
        for (size_t b = 0; b < 1024; ++b) {
            const float inv = 1.f / g[4 * b + 2];
            const float want[3] = {g[4 * b] * inv, g[4 * b + 1] * inv, g[4 * b + 3] * inv};
            for (int k = 0; k < 3; ++k) {
                const float got = d[3 * b + k];
                if (std::isnan(want[k]) ? !std::isnan(got) : std::memcmp(&want[k], &got, 4) != 0) same = false;
            }
        }
        EXPECT(same);
        EXPECT(rx.adc()->bitmap_f16().size() == 1024 * 3);
    }

    // cancel() from another thread stops receive() between slices; it reports "cancelled" like the reference
    std::thread killer([&] { std::this_thread::sleep_for(std::chrono::milliseconds(30)); integ.cancel(); });
    const bool finished = integ.receive(scene, rx, 1ull << 36, 1, 1ull << 22);
    killer.join();
    EXPECT(!finished && integ.samples_done() > 0 && integ.samples_done() < (1ull << 36));
    std::printf("cancel: stopped after %llu samples\n", (unsigned long long) integ.samples_done());
    std::printf("%d failure(s)\n", failures);
    return failures ? 1 : 0;
}
