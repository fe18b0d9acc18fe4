"""Independent anchor for the endpoint / signal rows DESIGN.md section 2 listed as "parity unpinned": the phased endpoints and
the `pulse` / `cw` / non-delta `linfmcw` signals.  tests/plate_path_endpoints.py evaluates single path samples of the plate scene
in numpy straight from the reference's text (not from oracle.cpp); the CPU oracle (not gpu) and the CUDA path (gpu, wavefront
and k_mega pipelines, through the C ABI) must reproduce it sample by sample."""
import math

import numpy as np
import pytest

from beifong_b200 import _native as N
from beifong_b200.meshes import Transform as T
from beifong_b200.scene import FlatScene, RADAR_94GHZ
from oracle_lib import OracleScene
from plate_path import PlateScene
from plate_path_endpoints import Signal, Variant, evaluate_sample

KIND = {"linfmcw": N.BF_SIG_LINFMCW, "pulse": N.BF_SIG_PULSE, "cw": N.BF_SIG_CW}


def make_variant(name):
    """(PlateScene, Variant).  The non-delta cases use a 2 MHz frequency extent and a 150 us pulse: the PSD's sinc argument stays
    below 1e3 rad, where one fp32 ulp of it is 6e-5 rad, so a sample's value is decided by the fp32 value of its frequency (244
    distinct values across the extent) and not by the last bit of a path time."""
    if name.startswith("phased"):
        sc = PlateScene()
        n = int(name[len("phased"):])
        return sc, Variant(sc, endpoints="phased", n_elems=n, steer=(0.0, 0.2, 0.0) if n == 2 else (0.1, -0.15, 0.0))
    if name == "two_tx":
        sc = PlateScene()
        return sc, Variant(sc, second_tx=dict(y=-0.02, gain=0.25))
    sc = PlateScene(f_ext=2e6, t_ext=150e-6, adc_start=5e-6, adc_time=140e-6, f_bw=3.0e6)
    kind = {"pulse": "pulse", "cw": "cw", "fmcw_nodelta": "linfmcw"}[name]
    base = dict(repfreq=sc.repfreq, t_ext=sc.t_ext, f_c=sc.f_c, f_ext=sc.f_ext, amplitude=1.5, gain=0.5)
    return sc, Variant(sc, rx_signal=Signal(kind, False, False, **base), tx_signal=Signal(kind, False, False, **base))


def abi_signal(sg):
    return FlatScene.signal(KIND[sg.kind], sig_is_delta=sg.delta, resample_freq=sg.resample, amplitude=float(sg.A), repfreq=float(sg.repfreq),
                            t_ext=float(sg.T), f_centre=float(sg.f_c), f_ext=float(sg.f_ext), gain=float(sg.gain))


def build(sc, var, spectrum_width=1):
    s = FlatScene(spectrum_width=spectrum_width, **RADAR_94GHZ)
    dflt = s.add_diffuse(0.5, two_sided=False)
    plate_m = s.add_diffuse(sc.rho, two_sided=False)
    rx = s.add_rect((T.scale([sc.a, sc.a, 1.0])).matrix32(), dflt)
    tx = s.add_rect((T.translate([0.0, sc.tx_y, 0.0]) * T.scale([sc.a, sc.a, 1.0])).matrix32(), dflt)
    s.add_rect((T.translate([0.0, 0.0, sc.R]) * T.rotate([1, 0, 0], 180.0) * T.scale([sc.plate_half, sc.plate_half, 1.0])).matrix32(), plate_m)
    s.add_rect((T.translate([0.0, 0.0, sc.floor_z]) * T.scale([sc.floor_half, sc.floor_half, 1.0])).matrix32(), plate_m)
    tex = s.add_uniform(1.0)
    tx_kind, rx_kind, tx_arr, rx_arr = N.BF_TX_WIGNER, N.BF_RX_WIGNER, None, None
    if var.endpoints == "phased":
        def arr(pose):
            return FlatScene.array(var.n_elems, var.steer, pose.matrix32(), (2.0 * sc.a, var.pitch, 1.0), (var.pitch,) * 3, (0, 1, 0))
        tx_kind, rx_kind = N.BF_TX_PHASED, N.BF_RX_PHASED
        tx_arr, rx_arr = arr(T.translate([0.0, sc.tx_y, 0.0])), arr(T.translate([0.0, 0.0, 0.0]))
    s.add_transmitter(tx, FlatScene.endpoint(tx_kind, texture=tex, signal=abi_signal(var.tx_signal), array=tx_arr))
    if var.second_tx is not None:
        import copy
        sg2 = copy.copy(var.tx_signal)
        sg2.gain = np.float32(float(var.tx_signal.gain) * var.second_tx["gain"])
        tx2 = s.add_rect((T.translate([0.0, var.second_tx["y"], 0.0]) * T.scale([sc.a, sc.a, 1.0])).matrix32(), dflt)
        s.add_transmitter(tx2, FlatScene.endpoint(tx_kind, texture=tex, signal=abi_signal(sg2)))
    s.set_receiver(rx, FlatScene.endpoint(rx_kind, texture=tex, signal=abi_signal(var.rx_signal), receive_type=N.BF_RECV_MIX_RESAMPLE,
                                          adc_sampling_start=sc.adc_start, adc_sampling_time=sc.adc_time, array=rx_arr))
    s.set_adc(t_bins=sc.t_bins, f_bins=sc.f_bins, t_bandwidth=sc.t_bw, f_bandwidth=sc.f_bw)
    s.set_integrator(N.BF_GEN_PHASE, max_depth=2, rr_depth=5, phase_bins=1)
    return s


def check_records(sc, var, rec, seed, n, value_scale=1.0):
    """Every modelled sample against the numpy evaluation; -> (checked, valid, nonzero, share of values within 2e-3)."""
    ev = [evaluate_sample(sc, var, seed, i) for i in range(n)]
    for e in ev:
        if e is not None:
            e["value"] *= value_scale
    scale = np.median([abs(e["value"]) for e in ev if e is not None and e["value"] != 0.0])
    checked = valid = nonzero = close = 0
    for i, e in enumerate(ev):
        if e is None:
            continue
        r = rec[i]
        checked += 1
        assert int(r["valid"]) == e["valid"], (i, r, e)
        assert r["tf_t"] == pytest.approx(e["tf_t"], rel=2e-6, abs=1e-6)
        if not e["valid"]:
            assert r["value"] == 0.0 and not math.isfinite(r["first_t"])
            continue
        valid += 1
        assert r["first_t"] == pytest.approx(e["first_t"], rel=3e-6)
        assert int(r["depth"]) == e["depth"]
        assert r["tf_f"] == pytest.approx(e["tf_f"], abs=0.02 + 1e-5 * e["tf_f"]), (i, r["tf_f"], e["tf_f"])
        if e["value"] != 0.0:
            nonzero += 1
        # the sum over n^2 complex terms / the PSD's sinc pass through zero: gate the absolute error against the typical sample there
        ok = abs(r["value"] - e["value"]) <= 2e-3 * abs(e["value"]) + 2e-4 * scale
        close += bool(ok)
        assert ok, (i, float(r["value"]), e["value"], scale)
    return checked, valid, nonzero, close / max(valid, 1)


VARIANTS = ["phased2", "phased3", "pulse", "cw", "fmcw_nodelta", "two_tx"]


@pytest.mark.parametrize("name", VARIANTS)
def test_oracle_reproduces_the_independent_endpoint_evaluation(name):
    sc, var = make_variant(name)
    s = build(sc, var)
    orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
    seed, n = 21, 3000
    rec = orc.sample_records(n, seed=seed)
    checked, valid, nonzero, _ = check_records(sc, var, rec, seed, n)
    assert checked >= 0.99 * n and valid > 0.15 * n and nonzero > 0.8 * valid
    s.close(); orc.close()


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", ["wavefront", "mega"])
@pytest.mark.parametrize("name", VARIANTS)
def test_device_reproduces_the_independent_endpoint_evaluation(native, name, pipeline):
    sc, var = make_variant(name)
    s = build(sc, var).commit()
    s.set_tunable("mega_max", (1 << 28) if pipeline == "mega" else -1)
    seed, n = 21, 3000
    rec = s.sample_records(n, seed=seed)
    checked, valid, nonzero, _ = check_records(sc, var, rec, seed, n)
    assert checked >= 0.99 * n and valid > 0.15 * n and nonzero > 0.8 * valid
    s.close()


def test_oracle_phased_endpoints_spectral_variant():
    """Four wavelengths: sample_wigner reads wavelength[0] and a delta signal gives all four the same frequency, so hsum(...) is four
    times the scalar evaluation (phasedreceiver.cpp:264-345 returns a Spectrum broadcast from real(W))."""
    sc, var = make_variant("phased2")
    s = build(sc, var, spectrum_width=4)
    orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
    rec = orc.sample_records(2000, seed=23)
    checked, valid, nonzero, _ = check_records(sc, var, rec, 23, 2000, value_scale=4.0)
    assert checked >= 0.99 * 2000 and valid > 0.15 * 2000 and nonzero > 0.8 * valid
    s.close(); orc.close()


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", ["wavefront", "mega"])
def test_device_phased_endpoints_spectral_variant(native, pipeline):
    sc, var = make_variant("phased2")
    s = build(sc, var, spectrum_width=4).commit()
    s.set_tunable("mega_max", (1 << 28) if pipeline == "mega" else -1)
    rec = s.sample_records(2000, seed=23)
    checked, valid, nonzero, _ = check_records(sc, var, rec, 23, 2000, value_scale=4.0)
    assert checked >= 0.99 * 2000 and valid > 0.15 * 2000 and nonzero > 0.8 * valid
    s.close()
