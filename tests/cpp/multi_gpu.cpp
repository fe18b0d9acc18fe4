// Two GPUs from ONE C++ process, no torch anywhere: the torch-free multi-GPU path of the C ABI (bf_multi_unique_id / bf_multi_init /
// bf_multi_reduce).  One thread per device: build the scene through include/beifong_b200/host.hpp, commit on device r,
// join the NCCL communicator, receive this rank's slice of the sample range, reduce onto rank 0.  Rank 0 then runs the WHOLE range
// alone and compares: count channels identical, range profile within 1e-3 relative L2 (fp32 summation order differs).
//   multi_gpu            exits 0 with "SKIP" when fewer than two devices are visible
#include <cmath>
#include <cstdio>
#include <thread>
#include <vector>
#include <cuda_runtime_api.h>

#include "beifong_b200/host.hpp"

using namespace beifong;
static const float C0 = 299792458.f;

struct Rig {
    Scene scene;
    ADC adc;
    uint32_t rx_shape = 0;
    Rig() : scene(constants(C0, C0 / 97e9f * 1e9f, C0 / 91e9f * 1e9f)), adc(1, 1024, 200e-6f, 1024 * 0.03f * 2.f / C0 * (6e9f / 200e-6f)) {
        const uint32_t dflt = scene.add_diffuse(0.5f, false), ground = scene.add_diffuse(0.5f, true);
        const uint32_t metal = scene.add_roughconductor(scene.add_uniform_spectrum(0.f), scene.add_uniform_spectrum(1.f), 0.1f, true);
        const float rx_m[16] = {0, 0, 1, 0,  0, 0.0025f, 0, 0,  -0.0025f, 0, 0, 0.3f,  0, 0, 0, 1};
        const float tx_m[16] = {0, 0, 1, 0,  0, 0.0025f, 0, 0.01f,  -0.0025f, 0, 0, 0.3f,  0, 0, 0, 1};
        const float gnd_m[16] = {20, 0, 0, 0,  0, 20, 0, 0,  0, 0, 1, 0,  0, 0, 0, 1};
        rx_shape = scene.add_rectangle(rx_m, dflt);
        const uint32_t tx_shape = scene.add_rectangle(tx_m, dflt);
        scene.add_rectangle(gnd_m, ground);
        std::vector<float> pos; std::vector<uint32_t> idx;
        for (int j = 0; j < 9; ++j) for (int i = 0; i < 9; ++i) { pos.push_back(8.f + 0.05f * i); pos.push_back(-2.f + 0.5f * i); pos.push_back(0.2f + 0.4f * j); }
        for (int j = 0; j < 8; ++j) for (int i = 0; i < 8; ++i) {
            uint32_t a = j * 9 + i, b = a + 1, c = a + 10, d = a + 9;
            idx.insert(idx.end(), {a, b, c, a, c, d});
        }
        scene.add_mesh(pos, idx, {}, metal);
        Signal tx_sig; tx_sig.repfreq = 1.f / 200e-6f; tx_sig.resample_freq = true;
        scene.add_transmitter("wignertransmitter", tx_shape, tx_sig, scene.add_uniform_spectrum(1.f));
    }
    void prepare(int device) {
        Signal rx_sig; rx_sig.repfreq = 1.f / 200e-6f;
        Receiver rx("wignerreceiver", rx_shape, adc, rx_sig, "mix_resample", 20e-6f, 180e-6f, scene.add_uniform_spectrum(1.f));
        Integrator integ(4, 5, 1);
        integ.prepare(scene, rx, device);
    }
};

int main() {
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev < 2) { std::printf("SKIP: %d CUDA device(s) visible, two needed\n", n_dev); return 0; }
    const int world = 2;
    const uint64_t n_total = (1ull << 22) + 12345, seed = 11;
    unsigned char id[BF_MULTI_ID_BYTES];
    if (bf_multi_unique_id(id, sizeof(id)) != BF_OK) { std::printf("FAILED bf_multi_unique_id: %s\n", bf_last_error(nullptr)); return 1; }
    std::vector<float> reduced;
    int failures = 0;
    std::vector<std::thread> threads;
    for (int r = 0; r < world; ++r)
        threads.emplace_back([&, r] {
            try {
                Rig rig;
                rig.prepare(r);
                bf_scene *s = rig.scene.handle();
                rig.scene.check(bf_multi_init(s, r, world, id, sizeof(id)), "bf_multi_init");
                const uint64_t base = n_total / world, first = r * base, n = r == world - 1 ? n_total - first : base;
                rig.scene.check(bf_receive(s, seed, first, n, 0), "bf_receive");
                rig.scene.check(bf_multi_reduce(s, 0), "bf_multi_reduce");
                rig.scene.check(bf_synchronize(s), "bf_synchronize");
                if (r == 0) {
                    uint32_t w, h, c; rig.scene.check(bf_profile_size(s, &w, &h, &c), "bf_profile_size");
                    reduced.resize((size_t) w * h * c);
                    rig.scene.check(bf_profile_read(s, reduced.data(), reduced.size()), "bf_profile_read");
                }
            } catch (const std::exception &e) { std::printf("FAILED rank %d: %s\n", r, e.what()); ++failures; }
        });
    for (std::thread &t : threads) t.join();
    if (failures) return 1;
    // the whole range on one device
    Rig rig; rig.prepare(0);
    bf_scene *s = rig.scene.handle();
    rig.scene.check(bf_receive(s, seed, 0, n_total, 0), "bf_receive");
    std::vector<float> whole(reduced.size());
    rig.scene.check(bf_profile_read(s, whole.data(), whole.size()), "bf_profile_read");
    double num = 0, den = 0; bool counts = true; double w_sum = 0;
    for (size_t i = 0; i < whole.size(); i += 4) {
        num += (double) (reduced[i] - whole[i]) * (reduced[i] - whole[i]); den += (double) whole[i] * whole[i];
        counts &= reduced[i + 1] == whole[i + 1] && reduced[i + 2] == whole[i + 2];
        w_sum += reduced[i + 2];
    }
    const double rel = std::sqrt(num / std::max(den, 1e-300));
    std::printf("2 GPUs, %llu samples: weight sum %.0f, counts %s, range-profile rel L2 reduced vs single %.3e\n", (unsigned long long) n_total, w_sum,
                counts ? "identical" : "DIFFER", rel);
    if (!counts || !(den > 0) || !(rel <= 1e-3) || w_sum < 0.9 * (double) n_total) { std::printf("FAILED\n"); return 1; }
    std::printf("0 failure(s)\n");
    return 0;
}
