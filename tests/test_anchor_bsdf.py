"""Independent anchor for the car's material class on the path (row a18): a plate of `roughconductor` (Beckmann, visible-normal
sampling, complex index eta + ik) -- BSDF value, its pdf inside the MIS weight, conductor Fresnel term -- evaluated per sample by
tests/plate_path_bsdf.py from the reference's text; oracle (not gpu) and device (gpu) must reproduce it."""
import math

import numpy as np
import pytest

from beifong_b200 import _native as N
from beifong_b200.meshes import Transform as T
from beifong_b200.scene import FlatScene, RADAR_94GHZ
from oracle_lib import OracleScene
from plate_path import PlateScene
from plate_path_bsdf import evaluate_sample

ALPHA, ETA, K = 0.3, 0.2, 3.0


def build(sc, two_sided=False, ggx=False):
    s = FlatScene(spectrum_width=1, **RADAR_94GHZ)
    dflt = s.add_diffuse(0.5, two_sided=False)
    rough = s.add_bsdf(N.BF_BSDF_ROUGHCONDUCTOR, two_sided=two_sided, eta=s.add_uniform(ETA), k=s.add_uniform(K), alpha_u=ALPHA, alpha_v=ALPHA,
                       distribution=N.BF_MICROFACET_GGX if ggx else N.BF_MICROFACET_BECKMANN)
    floor_m = s.add_diffuse(sc.rho, two_sided=False)
    rx = s.add_rect((T.scale([sc.a, sc.a, 1.0])).matrix32(), dflt)
    tx = s.add_rect((T.translate([0.0, sc.tx_y, 0.0]) * T.scale([sc.a, sc.a, 1.0])).matrix32(), dflt)
    s.add_rect((T.translate([0.0, 0.0, sc.R]) * T.rotate([1, 0, 0], 180.0) * T.scale([sc.plate_half, sc.plate_half, 1.0])).matrix32(), rough)
    s.add_rect((T.translate([0.0, 0.0, sc.floor_z]) * T.scale([sc.floor_half, sc.floor_half, 1.0])).matrix32(), floor_m)
    tex = s.add_uniform(1.0)
    sig = dict(amplitude=1.0, repfreq=sc.repfreq, t_ext=sc.t_ext, f_centre=sc.f_c, f_ext=sc.f_ext, gain=1.0)
    s.add_transmitter(tx, FlatScene.endpoint(N.BF_TX_WIGNER, texture=tex, signal=FlatScene.signal(N.BF_SIG_LINFMCW, sig_is_delta=True, resample_freq=True, **sig)))
    s.set_receiver(rx, FlatScene.endpoint(N.BF_RX_WIGNER, texture=tex, signal=FlatScene.signal(N.BF_SIG_LINFMCW, sig_is_delta=True, resample_freq=False, **sig),
                                          receive_type=N.BF_RECV_MIX_RESAMPLE, adc_sampling_start=sc.adc_start, adc_sampling_time=sc.adc_time))
    s.set_adc(t_bins=sc.t_bins, f_bins=sc.f_bins, t_bandwidth=sc.t_bw, f_bandwidth=sc.f_bw)
    s.set_integrator(N.BF_GEN_PHASE, max_depth=2, rr_depth=5, phase_bins=1)
    return s


def check_records(sc, rec, seed, n, ggx=False):
    ev = [evaluate_sample(sc, seed, i, ALPHA, ETA, K, ggx) for i in range(n)]
    scale = np.median([abs(e["value"]) for e in ev if e["value"] != 0.0])
    valid = nonzero = 0
    for i, e in enumerate(ev):
        r = rec[i]
        assert int(r["valid"]) == e["valid"], (i, r, e)
        assert r["tf_t"] == pytest.approx(e["tf_t"], rel=2e-6, abs=1e-6)
        if not e["valid"]:
            assert r["value"] == 0.0 and not math.isfinite(r["first_t"])
            continue
        valid += 1
        assert r["first_t"] == pytest.approx(e["first_t"], rel=3e-6)
        assert r["tf_f"] == pytest.approx(e["tf_f"], abs=0.02 + 1e-5 * e["tf_f"])
        nonzero += e["value"] != 0.0
        # the Beckmann lobe spans 30 orders of magnitude over the plate and is cut at D cos > 1e-20: gate the far tail absolutely
        assert abs(r["value"] - e["value"]) <= 2e-3 * abs(e["value"]) + 1e-6 * scale, (i, float(r["value"]), e["value"], scale)
    return valid, nonzero


@pytest.mark.parametrize("ggx", [False, True])
@pytest.mark.parametrize("two_sided", [False, True])
def test_oracle_reproduces_the_independent_roughconductor_evaluation(two_sided, ggx):
    sc = PlateScene()
    s = build(sc, two_sided, ggx)
    orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
    rec = orc.sample_records(3000, seed=29)
    valid, nonzero = check_records(sc, rec, 29, 3000, ggx)
    assert valid > 0.15 * 3000 and nonzero > 0.5 * valid
    s.close(); orc.close()


@pytest.mark.gpu
@pytest.mark.parametrize("ggx", [False, True])
@pytest.mark.parametrize("pipeline", ["wavefront", "mega"])
def test_device_reproduces_the_independent_roughconductor_evaluation(native, pipeline, ggx):
    sc = PlateScene()
    s = build(sc, True, ggx).commit()
    s.set_tunable("mega_max", (1 << 28) if pipeline == "mega" else -1)
    rec = s.sample_records(3000, seed=29)
    valid, nonzero = check_records(sc, rec, 29, 3000, ggx)
    assert valid > 0.15 * 3000 and nonzero > 0.5 * valid
    s.close()
