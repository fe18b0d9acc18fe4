"""Independent anchor for the radar rows the reference pins with no golden vector:
tests/plate_path.py evaluates single path samples of a three-rectangle scene in numpy float64 straight from the reference's text;
here the CPU oracle (not gpu) and the CUDA path (gpu) must reproduce it sample by sample -- first-hit distance, valid flag, both bin
coordinates, the value hsum(|ray_weight| L) and the phase handed to the PhaseIntegrator -- and the binned profile must be the
histogram of those values."""
import math

import numpy as np
import pytest

from beifong_b200 import _native as N
from beifong_b200.meshes import Transform as T
from beifong_b200.scene import FlatScene, RADAR_94GHZ
from oracle_lib import OracleScene
from plate_path import PlateScene, evaluate_sample
from record_binning import profile_from_records


def build(sc, plate_velocity=None, max_depth=2, mesh_plate=None, spectrum_width=1):
    s = FlatScene(spectrum_width=spectrum_width, **RADAR_94GHZ)
    dflt = s.add_diffuse(0.5, two_sided=False)
    plate_m = s.add_diffuse(sc.rho, two_sided=False)
    rx = s.add_rect((T.scale([sc.a, sc.a, 1.0])).matrix32(), dflt)
    tx = s.add_rect((T.translate([0.0, sc.tx_y, 0.0]) * T.scale([sc.a, sc.a, 1.0])).matrix32(), dflt)
    if mesh_plate:
        # the same plate as a triangle mesh facing -z (an n x n grid of quads; "smooth": with vertex normals, all (0, 0, -1)): the
        # first hit then goes through the BVH, Moeller-Trumbore and Mesh::compute_surface_interaction (mesh.cpp:452-548) instead of
        # the analytic rectangle; without UVs the tangent comes from coordinate_system(n), which for n = (0, 0, -1) gives the
        # rectangle's own shading frame ((1, 0, 0), (0, -1, 0)), so the independent evaluation applies unchanged
        g = 6
        xs = np.linspace(-sc.plate_half, sc.plate_half, g + 1)
        pos = np.array([[x, y, sc.R] for y in xs for x in xs], np.float32)
        faces = []
        for j in range(g):
            for i in range(g):
                v = j * (g + 1) + i
                faces += [[v, v + g + 1, v + 1], [v + 1, v + g + 1, v + g + 2]]       # wound so that the normal is -z
        nrm = np.tile(np.array([[0.0, 0.0, -1.0]], np.float32), (len(pos), 1)) if mesh_plate == "smooth" else None
        plate = s.add_mesh(pos, np.array(faces, np.uint32), plate_m, normals=nrm)
    else:
        plate = s.add_rect((T.translate([0.0, 0.0, sc.R]) * T.rotate([1, 0, 0], 180.0) * T.scale([sc.plate_half, sc.plate_half, 1.0])).matrix32(), plate_m)
    if plate_velocity is not None:
        s.set_shape_velocity(plate, T.translate(list(plate_velocity)).matrix32())
    s.add_rect((T.translate([0.0, 0.0, sc.floor_z]) * T.scale([sc.floor_half, sc.floor_half, 1.0])).matrix32(), plate_m)
    tex = s.add_uniform(1.0)
    sig = dict(amplitude=1.0, repfreq=sc.repfreq, t_ext=sc.t_ext, f_centre=sc.f_c, f_ext=sc.f_ext, gain=1.0)
    s.add_transmitter(tx, FlatScene.endpoint(N.BF_TX_WIGNER, texture=tex, signal=FlatScene.signal(N.BF_SIG_LINFMCW, sig_is_delta=True, resample_freq=True, **sig)))
    s.set_receiver(rx, FlatScene.endpoint(N.BF_RX_WIGNER, texture=tex, signal=FlatScene.signal(N.BF_SIG_LINFMCW, sig_is_delta=True, resample_freq=False, **sig),
                                          receive_type=N.BF_RECV_MIX_RESAMPLE, adc_sampling_start=sc.adc_start, adc_sampling_time=sc.adc_time))
    s.set_adc(t_bins=sc.t_bins, f_bins=sc.f_bins, t_bandwidth=sc.t_bw, f_bandwidth=sc.f_bw)
    s.set_integrator(N.BF_GEN_PHASE, max_depth=max_depth, rr_depth=5, phase_bins=1, doppler=plate_velocity is not None)
    return s


def check_records(sc, rec, seed, idxs, plate_velocity=None, max_depth=2, value_scale=1.0):
    """Every modelled sample of `idxs` against the numpy evaluation."""
    n_checked = n_valid = n_phase = 0
    worst = 0.0
    for i in idxs:
        e = evaluate_sample(sc, seed, int(i), plate_velocity, max_depth)
        if e is None:
            continue
        e["value"] *= value_scale
        r = rec[i]
        n_checked += 1
        assert int(r["valid"]) == e["valid"], (i, r, e)
        assert r["tf_t"] == pytest.approx(e["tf_t"], rel=2e-6, abs=1e-6), (i, r["tf_t"], e["tf_t"])
        if not e["valid"]:
            assert r["value"] == 0.0 and not math.isfinite(r["first_t"])
            continue
        n_valid += 1
        assert r["first_t"] == pytest.approx(e["first_t"], rel=3e-6)
        assert int(r["depth"]) == e["depth"]
        if plate_velocity is not None:      # the Doppler-shifted wavelength itself, to fp32 rounding
            assert r["wavelength0"] == pytest.approx(e["wavelength0"], rel=3e-7), (i, r["wavelength0"], e["wavelength0"])
        # beat-frequency bin coordinate: a difference of two ~1e11 Hz values; the fp32 roundings of the two wavelengths are
        # reproduced by the evaluator, so it agrees to a small fraction of a bin
        assert r["tf_f"] == pytest.approx(e["tf_f"], abs=(0.02 if plate_velocity is None else 1.5) + 1e-5 * e["tf_f"]), (i, r["tf_f"], e["tf_f"])
        scale = max(abs(e["value"]), 1e-12)
        worst = max(worst, abs(r["value"] - e["value"]) / scale)
        assert r["value"] == pytest.approx(e["value"], rel=2e-3, abs=1e-9), (i, r["value"], e["value"])
        if math.isfinite(e["phase"]):
            n_phase += 1
            assert r["phase"] == pytest.approx(e["phase"], rel=1e-5)      # t2 of a grazing continuation ray carries the fp32 error of its direction
        else:
            assert r["phase"] == -math.inf
    return n_checked, n_valid, n_phase, worst


def test_oracle_reproduces_the_independent_path_evaluation():
    sc = PlateScene()
    s = build(sc)
    orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
    seed, n = 9, 4000
    rec = orc.sample_records(n, seed=seed)
    checked, valid, with_phase, worst = check_records(sc, rec, seed, range(n))
    assert checked >= 0.99 * n and valid > 0.15 * n and with_phase > 0.9 * valid
    # median agreement is at fp32 rounding level, far inside the gate of the individual samples
    vals = np.array([(rec[i]["value"], (evaluate_sample(sc, seed, i) or {"value": np.nan})["value"]) for i in range(400)])
    ok = np.isfinite(vals[:, 1]) & (vals[:, 1] != 0)
    assert np.median(np.abs(vals[ok, 0] - vals[ok, 1]) / np.abs(vals[ok, 1])) <= 2e-5
    # SignalBlock::put: the oracle's binned profile is the histogram of its per-sample records (numpy restatement of the box splat)
    prof = orc.receive(n, seed=seed, threads=1)
    mine, _, _ = profile_from_records(rec, prof.shape[1], prof.shape[0], 1)
    assert np.array_equal(mine[..., 1:3], prof[..., 1:3].astype(np.float64))
    assert np.linalg.norm(mine[..., 0] - prof[..., 0]) <= 1e-5 * np.linalg.norm(mine[..., 0])
    # the plate at R = 5 m and the sweep rate put the energy where the closed form says: f_beat = slope * (2 R / c) up to the
    # offset of the transmitter, i.e. bin f_beat * f_bins / f_bw of the frequency axis
    k = sc.f_ext / sc.t_ext * (2.0 * sc.R / sc.c) * sc.f_bins / sc.f_bw
    energy = np.abs(prof[..., 0]).sum(axis=1)
    assert abs(int(np.argmax(energy)) - int(k)) <= 1
    assert energy[int(k) - 1:int(k) + 40].sum() >= 0.98 * energy.sum()       # oblique paths are longer: the tail lies above bin k
    s.close(); orc.close()


VELOCITY = (0.0, 0.0, 30.0)      # 30 m/s along the plate normal: a shift of ~2e-7 relative, ~19 kHz at 94 GHz, ~1.6 frequency bins


def test_oracle_doppler_flag_against_the_independent_evaluation():
    """Shape::doppler (shape.cpp:375-389) behind the integrator's doppler flag: the shifted wavelength and beat-frequency coordinate of
    every sample against the numpy evaluation; with the flag off (the default, what the reference executes) the velocity is inert."""
    sc = PlateScene()
    s = build(sc, VELOCITY)
    orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
    seed, n = 9, 1500
    rec = orc.sample_records(n, seed=seed)
    checked, valid, _, _ = check_records(sc, rec, seed, range(n), VELOCITY)
    assert checked >= 0.99 * n and valid > 0.15 * n
    off = build(sc)
    off.set_shape_velocity(2, T.translate(list(VELOCITY)).matrix32())       # velocity present, flag off
    rec_off = OracleScene(off.export(), accel=OracleScene.BRUTE).sample_records(n, seed=seed)
    base = OracleScene(build(sc).export(), accel=OracleScene.BRUTE).sample_records(n, seed=seed)
    assert rec_off.tobytes() == base.tobytes()
    hit = rec["valid"] == 1
    assert np.median(rec["tf_f"][hit] - base["tf_f"][hit]) > 0.5             # the approaching plate raises the beat frequency
    s.close(); off.close(); orc.close()


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", ["wavefront", "mega"])
def test_device_doppler_flag(native, pipeline):
    sc = PlateScene()
    s = build(sc, VELOCITY).commit()
    s.set_tunable("mega_max", (1 << 28) if pipeline == "mega" else -1)
    seed, n = 9, 1500
    rec = s.sample_records(n, seed=seed)
    checked, valid, _, _ = check_records(sc, rec, seed, range(n), VELOCITY)
    assert checked >= 0.99 * n and valid > 0.15 * n
    orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
    c = orc.sample_records(n, seed=seed)
    assert np.array_equal(rec["wavelength0"].view(np.uint32)[rec["valid"] == 1], c["wavelength0"].view(np.uint32)[c["valid"] == 1])
    s.close(); orc.close()


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", ["wavefront", "mega"])
def test_device_reproduces_the_independent_path_evaluation(native, pipeline):
    sc = PlateScene()
    s = build(sc).commit()
    s.set_tunable("mega_max", (1 << 28) if pipeline == "mega" else -1)
    seed, n = 9, 4000
    rec = s.sample_records(n, seed=seed)
    checked, valid, with_phase, worst = check_records(sc, rec, seed, range(n))
    assert checked >= 0.99 * n and valid > 0.15 * n and with_phase > 0.9 * valid
    s.receive(n, seed=seed)
    prof = s.profile()
    mine, _, _ = profile_from_records(rec, prof.shape[1], prof.shape[0], 1)
    assert np.array_equal(mine[..., 1:3], prof[..., 1:3].astype(np.float64))
    assert np.linalg.norm(mine[..., 0] - prof[..., 0]) <= 1e-5 * np.linalg.norm(mine[..., 0])
    s.close()


def test_oracle_three_segment_path_against_the_independent_evaluation():
    """max_depth 3: a second loop iteration at the floor.  Its emitter sampling contributes nothing (the transmitter is seen from
    behind) but resamples the interaction's wavelengths at the later time, so the sample's beat-frequency coordinate moves on by the
    sweep rate times the extra delay; the time is carried over three segments and the phase comes from the third."""
    sc = PlateScene()
    s = build(sc, max_depth=3)
    orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
    seed, n = 13, 3000
    rec = orc.sample_records(n, seed=seed)
    checked, valid, with_phase, _ = check_records(sc, rec, seed, range(n), max_depth=3)
    assert checked >= 0.99 * n and valid > 0.15 * n and with_phase > 0.05 * valid       # the third segment meets the plate for ~1 in 10
    two = OracleScene(build(sc).export(), accel=OracleScene.BRUTE).sample_records(n, seed=seed)
    deeper = (rec["valid"] == 1) & (rec["depth"] == 3)
    assert deeper.sum() > 0.1 * n
    assert np.array_equal(rec["value"][deeper], two["value"][deeper])                    # same radiance ...
    assert np.median(rec["tf_f"][deeper] - two["tf_f"][deeper]) > 1.0                    # ... binned at a later beat frequency
    s.close(); orc.close()


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", ["wavefront", "mega"])
def test_device_three_segment_path(native, pipeline):
    sc = PlateScene()
    s = build(sc, max_depth=3).commit()
    s.set_tunable("mega_max", (1 << 28) if pipeline == "mega" else -1)
    seed, n = 13, 3000
    rec = s.sample_records(n, seed=seed)
    checked, valid, with_phase, _ = check_records(sc, rec, seed, range(n), max_depth=3)
    assert checked >= 0.99 * n and valid > 0.15 * n and with_phase > 0.05 * valid
    s.close()


@pytest.mark.parametrize("mesh_plate", ["flat", "smooth"])
def test_oracle_mesh_plate_against_the_independent_evaluation(mesh_plate):
    """The plate as a triangle mesh: BVH / Moeller-Trumbore / Mesh::compute_surface_interaction in place of the analytic rectangle."""
    sc = PlateScene()
    s = build(sc, max_depth=3, mesh_plate=mesh_plate)
    for accel in (OracleScene.BRUTE, OracleScene.BVH):
        orc = OracleScene(s.export(), accel=accel)
        rec = orc.sample_records(2000, seed=17)
        checked, valid, with_phase, _ = check_records(sc, rec, 17, range(2000), max_depth=3)
        assert checked >= 0.99 * 2000 and valid > 0.15 * 2000 and with_phase > 0.05 * valid
        assert (rec["first_prim"][rec["valid"] == 1] < 72).all() and len(np.unique(rec["first_prim"][rec["valid"] == 1])) > 20
        orc.close()
    s.close()


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", ["wavefront", "mega"])
@pytest.mark.parametrize("mesh_plate", ["flat", "smooth"])
def test_device_mesh_plate(native, mesh_plate, pipeline):
    sc = PlateScene()
    s = build(sc, max_depth=3, mesh_plate=mesh_plate).commit()
    s.set_tunable("mega_max", (1 << 28) if pipeline == "mega" else -1)
    rec = s.sample_records(2000, seed=17)
    checked, valid, with_phase, _ = check_records(sc, rec, 17, range(2000), max_depth=3)
    assert checked >= 0.99 * 2000 and valid > 0.15 * 2000 and with_phase > 0.05 * valid
    s.close()


def test_oracle_spectral_variant_against_the_independent_evaluation():
    """`*_spectral` variants carry four wavelengths; a delta signal gives all four the same frequency (wignerreceiver.cpp:163-175
    assigns a Float to a Wavelength), so every per-wavelength quantity is the scalar one and aovs[0] = hsum(...) is four times it."""
    sc = PlateScene()
    s = build(sc, max_depth=3, spectrum_width=4)
    orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
    rec = orc.sample_records(2000, seed=19)
    checked, valid, _, _ = check_records(sc, rec, 19, range(2000), max_depth=3, value_scale=4.0)
    assert checked >= 0.99 * 2000 and valid > 0.15 * 2000
    s.close(); orc.close()


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", ["wavefront", "mega"])
def test_device_spectral_variant(native, pipeline):
    sc = PlateScene()
    s = build(sc, max_depth=3, spectrum_width=4).commit()
    s.set_tunable("mega_max", (1 << 28) if pipeline == "mega" else -1)
    rec = s.sample_records(2000, seed=19)
    checked, valid, _, _ = check_records(sc, rec, 19, range(2000), max_depth=3, value_scale=4.0)
    assert checked >= 0.99 * 2000 and valid > 0.15 * 2000
    s.close()


def test_oracle_phase_bins_and_window_offsets_against_numpy_binning():
    """PhaseIntegrator with 8 phase bins (phase.cpp:117-124) and a cropped ADC window with offsets (adc.cpp:29-47): the oracle's
    binned profile against the numpy restatement of the binning (tests/record_binning.py) applied to its own per-sample records."""
    from record_binning import phase_channel
    sc = PlateScene()
    s = build(sc, max_depth=3)
    s.set_integrator(N.BF_GEN_PHASE, max_depth=3, rr_depth=5, phase_bins=8)
    s.set_adc(t_bins=sc.t_bins, f_bins=sc.f_bins, t_bandwidth=sc.t_bw, f_bandwidth=sc.f_bw, window_f_bins=96, window_offset_f=40, window_t_bins=3, window_offset_t=1)
    orc = OracleScene(s.export(), accel=OracleScene.BRUTE)
    n = 20000
    rec = orc.sample_records(n, seed=31)
    prof = orc.receive(n, seed=31, threads=1)
    assert prof.shape == (96, 3, 3 + 8)
    mine, pix, ch = profile_from_records(rec, 3, 96, 8, off_t=1, off_f=40)
    assert np.array_equal(mine[..., 1:3], prof[..., 1:3].astype(np.float64))           # valid and weight counts, bin by bin
    assert np.linalg.norm(mine[..., 0] - prof[..., 0]) <= 1e-5 * np.linalg.norm(mine[..., 0])
    # the phase AOVs hold the integrator's UNWEIGHTED hsum(L) (phase.cpp:117-124 runs inside sample(), before receive_sample applies
    # |ray_weight|), in the channel of the sample's phase bin: L comes from the independent evaluation, bin and channel from the records
    aov = np.zeros((96 * 3, 8))
    unmodelled = np.zeros(96 * 3, bool)
    for i in range(n):
        if pix[i] < 0 or not rec["valid"][i]:
            continue
        e = evaluate_sample(sc, 31, i, None, 3)
        if e is None:
            unmodelled[pix[i]] = True
        elif ch[i] >= 0:
            aov[pix[i], ch[i]] += e["radiance"]
    got = prof[..., 3:].reshape(-1, 8).astype(np.float64)
    assert (np.abs(got).sum(axis=0) > 0).sum() >= 6                                    # spread over the bins, not piled into one
    clean = ~unmodelled
    assert clean.mean() > 0.5 and np.abs(aov[clean]).sum() > 0
    assert np.linalg.norm(got[clean] - aov[clean]) <= 2e-3 * np.linalg.norm(aov[clean])
    assert 0 < (pix >= 0).sum() < n                                                    # the window really crops
    s.close(); orc.close()
