"""The benchmark / parity configurations of BASELINE.json as concrete scenes.

Each builder returns an un-committed FlatScene.  Meshes named by BASELINE.json that are missing from
the reference tree (.MISSING_LARGE_BLOBS: Car-body.ply, Motorbike, Bus) are replaced by procedural
stand-ins of the same triangle-count class (beifong_b200.meshes); this is stated in every result.
"""
import numpy as np

from . import _native as N
from .meshes import Transform as T
from . import meshes
from .scene import FlatScene, RADAR_94GHZ

C0 = 299792458.0


def rect_to_world(scale_xy, origin, target, up=(0, 0, 1)):
    """<scale x y/> followed by <lookat/>: later XML transforms multiply from the left (xml.cpp:918)."""
    return (T.look_at(origin, target, up) * T.scale([scale_xy[0], scale_xy[1], 1.0])).matrix32()


def c1_trans_rad(spectrum_width=1, max_depth=-1, rr_depth=5):
    """python_scripts/trans_rad.xml verbatim: fluxmeter on a 0.05 x 0.05 rectangle at the origin facing -y,
    spot emitter (cutoff 25, beam 20), 1 x 1 target at y = -4, 20 x 20 ground at z = -0.5, all
    twosided(diffuse 1.0); gen-1 integrator time(pathtime): 50 bins of 0.5 ns."""
    s = FlatScene(spectrum_width=spectrum_width)          # reference defaults (spectrum.h:15-40)
    default_bsdf = s.add_diffuse(0.5, two_sided=False)     # Shape's default BSDF (shape.cpp:52-60)
    material = s.add_diffuse(1.0, two_sided=True)
    rx = s.add_rect(rect_to_world((0.05, 0.05), (0, 0, 0), (0, -1, 0)), default_bsdf)
    s.add_rect(rect_to_world((1, 1), (0, -4, 0), (0, 0, 0)), material)
    s.add_rect(rect_to_world((20, 20), (0, 0, -0.5), (0, 0, 0.5), up=(1, 0, 0)), material)   # no <up>: xml.cpp:911-913
    intensity = s.add_uniform(1.0)
    spot = FlatScene.endpoint(N.BF_TX_SPOT, texture=intensity, cutoff_angle_deg=25.0, beam_width_deg=20.0,
                              to_world=T.look_at((0, 0, 0), (0, -1, 0), (0, 0, 1)).matrix32())
    s.add_transmitter(None, spot)
    s.set_receiver(rx, FlatScene.endpoint(N.BF_RX_FLUXMETER))
    s.set_integrator(N.BF_GEN_TIME, max_depth=max_depth, rr_depth=rr_depth)
    return s


def fmcw_signal(resample_freq, gain=1.0):
    # Receive.ipynb cell 5: 94 GHz centre, 6 GHz sweep, 200 us chirp
    return FlatScene.signal(N.BF_SIG_LINFMCW, sig_is_delta=True, resample_freq=resample_freq, amplitude=1.0,
                            repfreq=1.0 / 200e-6, t_ext=200e-6, f_centre=94e9, f_ext=6e9, gain=gain)


def radar_scene(mesh_pos, mesh_faces, spectrum_width=1, max_depth=4, range_bins=1024, bin_m=0.03, smooth_normals=True,
                extra_meshes=(), phased=None, ground_half=20.0, reflector=True, return_ids=False, omni_area=False, variant=None):
    """Common 94 GHz FMCW layout of python_scripts/Render.py / Receive.ipynb: Tx and Rx are 5 mm x 5 mm
    rectangles at (0, 0, 0.3) facing +x, ground 20 x 20 diffuse 0.5, corner reflector 0.1 x 0.05 diffuse 1,
    target mesh twosided(roughconductor alpha 0.1).  Receiver `wignerreceiver` (mix_resample) + transmitter
    `wignertransmitter` (resample_freq): beat frequency = sweep rate x delay, binned on the frequency axis."""
    # `variant`: overrides used by the parity tests to reach every branch of the path (all default to the layout above):
    #   tx_signal / rx_signal (bf_signal_desc), receive_type, phase_bins, rr_depth, car_bsdf (kwargs of add_bsdf),
    #   ground_spectrum ("regular" | "irregular": two / three-point tables as python_scripts/trans_image_rad.xml:53 writes them),
    #   window (kwargs of set_adc: window_t_bins, window_f_bins, window_offset_t, window_offset_f), t_bins,
    #   second_tx (True: a second wignertransmitter on its own aperture)
    v = dict(variant or {})
    s = FlatScene(spectrum_width=spectrum_width, **RADAR_94GHZ)
    default_bsdf = s.add_diffuse(0.5, two_sided=False)
    lam_lo, lam_hi = RADAR_94GHZ["lambda_min_nm"], RADAR_94GHZ["lambda_max_nm"]
    if v.get("ground_spectrum") == "regular":
        ground = s.add_bsdf(N.BF_BSDF_DIFFUSE, two_sided=True, reflectance=s.add_spectrum(N.BF_SPEC_REGULAR, [0.5, 0.5], [lam_lo, lam_hi]))
        refl = s.add_bsdf(N.BF_BSDF_DIFFUSE, two_sided=True, reflectance=s.add_spectrum(N.BF_SPEC_REGULAR, [0.9, 1.0], [lam_lo, lam_hi]))
    elif v.get("ground_spectrum") == "irregular":
        mid = lam_lo + 0.3 * (lam_hi - lam_lo)
        ground = s.add_bsdf(N.BF_BSDF_DIFFUSE, two_sided=True, reflectance=s.add_spectrum(N.BF_SPEC_IRREGULAR, [0.4, 0.6, 0.5], [lam_lo, mid, lam_hi]))
        refl = s.add_bsdf(N.BF_BSDF_DIFFUSE, two_sided=True, reflectance=s.add_spectrum(N.BF_SPEC_IRREGULAR, [1.0, 0.8, 1.0], [lam_lo, mid, lam_hi]))
    else:
        ground = s.add_diffuse(0.5, two_sided=True)
        refl = s.add_diffuse(1.0, two_sided=True)
    # twosided(roughconductor): Beckmann alpha 0.1; a perfect conductor at radar wavelengths (the
    # reference's default Cu table evaluates to eta = k = 0 outside the visible range => F = 1)
    eta, k = s.add_uniform(0.0), s.add_uniform(1.0)
    car_kw = dict(kind=N.BF_BSDF_ROUGHCONDUCTOR, two_sided=True, eta=eta, k=k, alpha_u=0.1, alpha_v=0.1)
    car_kw.update(v.get("car_bsdf", {}))
    car = s.add_bsdf(car_kw.pop("kind"), **car_kw)
    ant = (T.translate([0.0, 0.0, 0.3]) * T.rotate([0, 1, 0], 90.0) * T.scale([0.0025, 0.0025, 1.0]))
    rx = s.add_rect(ant.matrix32(), default_bsdf)
    tx = s.add_rect((T.translate([0.0, 0.01, 0.0]) * ant).matrix32(), default_bsdf)
    s.add_rect((T.translate([0, 0, 0]) * T.scale([ground_half, ground_half, 1.0])).matrix32(), ground)
    if reflector:
        s.add_rect((T.translate([6.0, -2.0, 0.5]) * T.rotate([0, 1, 0], -90.0) * T.scale([0.05, 0.025, 1.0])).matrix32(), refl)
    nrm = meshes.vertex_normals(mesh_pos, mesh_faces) if smooth_normals else None
    s.add_mesh(mesh_pos, mesh_faces, car, normals=nrm)
    for (p, f) in extra_meshes:
        s.add_mesh(p, f, car, normals=meshes.vertex_normals(p, f) if smooth_normals else None)
    tex = s.add_uniform(1.0)
    tx_kind, rx_kind, tx_arr, rx_arr = N.BF_TX_WIGNER, N.BF_RX_WIGNER, None, None
    if phased:
        # `phased` = dict(n_elems, steer): phasedtransmitter / phasedreceiver on the same apertures, the aperture split into
        # n_elems elements along its local y axis (elem_spacing = pitch, elem_axis = y), array_loc = the aperture's pose
        n = int(phased.get("n_elems", 4))
        pitch = 0.005 / n
        def arr(pose):
            return FlatScene.array(n, phased.get("steer", (0.0, 0.0, 0.0)), pose.matrix32(), (0.005, pitch, 1.0), (pitch, pitch, pitch), (0, 1, 0))
        pose = T.translate([0.0, 0.0, 0.3]) * T.rotate([0, 1, 0], 90.0)
        tx_kind, rx_kind = N.BF_TX_PHASED, N.BF_RX_PHASED
        tx_arr, rx_arr = arr(T.translate([0.0, 0.01, 0.0]) * pose), arr(pose)
    if omni_area:   # `areatransmitter` + `omnidirectional` receiver (Receive.ipynb cells 13, 15 use the omnidirectional receiver)
        tx_kind, rx_kind = N.BF_TX_AREA, N.BF_RX_OMNIDIRECTIONAL
    s.add_transmitter(tx, FlatScene.endpoint(tx_kind, texture=tex, signal=v.get("tx_signal", fmcw_signal(True)), array=tx_arr))
    if v.get("second_tx"):
        tx2 = s.add_rect((T.translate([0.0, -0.02, 0.1]) * ant).matrix32(), default_bsdf)
        s.add_transmitter(tx2, FlatScene.endpoint(tx_kind, texture=s.add_uniform(0.5), signal=v.get("tx_signal", fmcw_signal(True)), array=tx_arr))
    sweep_rate = 6e9 / 200e-6
    f_bandwidth = v.get("f_bandwidth", range_bins * bin_m * 2.0 / C0 * sweep_rate)
    s.set_receiver(rx, FlatScene.endpoint(rx_kind, texture=tex, signal=v.get("rx_signal", fmcw_signal(False)),
                                          receive_type=v.get("receive_type", N.BF_RECV_MIX_RESAMPLE),
                                          adc_sampling_start=20e-6, adc_sampling_time=160e-6, array=rx_arr))
    s.set_adc(t_bins=v.get("t_bins", 1), f_bins=range_bins, t_bandwidth=200e-6, f_bandwidth=f_bandwidth, **v.get("window", {}))
    s.set_integrator(N.BF_GEN_PHASE, max_depth=max_depth, rr_depth=v.get("rr_depth", 5), phase_bins=v.get("phase_bins", 1))
    return (s, rx, tx) if return_ids else s


def car_pose():
    # Render.py:305-316: rotate x 90, rotate z -20, translate (10, 3, 1); the stand-in is already z-up
    return T.translate([10.0, 3.0, 1.0]) * T.rotate([0, 0, 1], -20.0)


def c2_car_body(spectrum_width=1, max_depth=4, n_u=264, n_v=133):
    """BASELINE config 2: car-body mesh (procedural stand-in, 69 696 triangles), 1 Tx / 1 Rx, 3 bounces."""
    p, f = meshes.car_body(n_u, n_v)
    p = meshes.transform_points(car_pose().matrix32(), p)
    return radar_scene(p, f, spectrum_width=spectrum_width, max_depth=max_depth)


def small_radar(spectrum_width=1, max_depth=4, n_u=24, n_v=13, smooth_normals=True, phased=None, omni_area=False, variant=None):
    """The C2 layout with a few hundred triangles: sized for the brute-force oracle."""
    p, f = meshes.car_body(n_u, n_v)
    p = meshes.transform_points(car_pose().matrix32(), p)
    return radar_scene(p, f, spectrum_width=spectrum_width, max_depth=max_depth, smooth_normals=smooth_normals, phased=phased, omni_area=omni_area,
                       variant=variant)


def c3_multi_object(spectrum_width=1, max_depth=4, copies=14, n_u=264, n_v=133, f_bins=1024, t_bins=1024):
    """BASELINE config 3 stand-in: `copies` car bodies on a 7 x 2 grid (~1 M triangles), 1024 x 1024 ADC."""
    base_p, base_f = meshes.car_body(n_u, n_v)
    extra = []
    first = None
    for i in range(copies):
        gx, gy = i % 7, i // 7
        scale = 1.0 if i % 2 == 0 else 2.2
        m = (T.translate([8.0 + 3.0 * gx * 2.0, -6.0 + 12.0 * gy, 1.0 * scale]) * T.rotate([0, 0, 1], 17.0 * i) * T.scale(scale)).matrix32()
        p = meshes.transform_points(m, base_p)
        if first is None:
            first = (p, base_f)
        else:
            extra.append((p, base_f))
    s = radar_scene(first[0], first[1], spectrum_width=spectrum_width, max_depth=max_depth, extra_meshes=extra)
    return s


def c4_pose(frame, n_frames=256):
    """Tx/Rx pose of frame `frame` of the animated sequence: boresight (0, -1, 0) rotated about z by frame * 360 / n_frames
    degrees (python_scripts/animated_trans_rad.py:305-311 steps by 5 degrees; 256 positions is Receive.ipynb's da = 1 / 2**8)."""
    rot = T.rotate([0, 0, 1], frame * 360.0 / n_frames).matrix32().astype(np.float64)
    boresight = rot[:3, :3] @ np.array([0.0, -1.0, 0.0])
    up = rot[:3, :3] @ np.array([0.0, 0.0, 1.0])
    return T.look_at((0, 0, 0), tuple(boresight), tuple(up))


def c4_receiver(pose, dr=0.2, bins=50):
    near, far = dr, bins * dr + dr
    return FlatScene.endpoint(N.BF_RX_PERSPECTIVE, film_width=1, film_height=1, near_clip=near, far_clip=far, to_world=pose.matrix32(),
                              sample_to_camera=meshes.perspective_sample_to_camera(45.0, near, far, 1, 1))


def c4_animated(frame=0, n_frames=256, spectrum_width=1, dr=0.2, bins=50, max_depth=-1):
    """BASELINE config 4, one frame of python_scripts/animated_trans_rad.py: 1 x 1 target at y = +4 and 20 x 20 ground at
    z = -0.5 (twosided diffuse 1), area emitter on a 0.5 x 0.5 rectangle at the origin (radiance 100), perspective sensor
    fov 45, near / far = dr / (bins + 1) dr, 1 x 1 box-filtered film; gen-2 integrator range(pathlength), dr 0.2, 50 bins.
    Returns (scene, tx_shape_id): later frames call scene.update_rect(tx_shape_id, ...) and scene.set_receiver(None, ...)."""
    s = FlatScene(spectrum_width=spectrum_width)
    default_bsdf = s.add_diffuse(0.5, two_sided=False)
    material = s.add_diffuse(1.0, two_sided=True)
    pose = c4_pose(frame, n_frames)
    tx = s.add_rect((pose * T.scale([0.5, 0.5, 1.0])).matrix32(), default_bsdf)
    s.add_rect(rect_to_world((1, 1), (0, 4, 0), (0, 0, 0)), material)
    s.add_rect(rect_to_world((20, 20), (0, 0, -0.5), (0, 0, 0.5), up=(0, 1, 0)), material)
    s.add_transmitter(tx, FlatScene.endpoint(N.BF_TX_AREALIGHT, texture=s.add_uniform(100.0)))
    s.set_receiver(None, c4_receiver(pose, dr, bins))
    s.set_integrator(N.BF_GEN_RANGE, max_depth=max_depth, rr_depth=5, range_bins=bins, dr=dr)
    return s, tx


C5_ELEMENTS = 64
C5_PITCH = 0.5 * C0 / 94e9          # lambda / 2 at the 94 GHz carrier of the stand-in signal (BASELINE names 76.5 GHz)


def c5_element_pose(k, n_elements=C5_ELEMENTS):
    """to_world of receive element k: 5 mm x 5 mm apertures on a line along y, pitch lambda / 2, at z = 0.3 facing +x."""
    y = (k - 0.5 * (n_elements - 1)) * C5_PITCH
    return (T.translate([0.0, y, 0.3]) * T.rotate([0, 1, 0], 90.0) * T.scale([0.0025, 0.0025, 1.0])).matrix32()


def c5_urban(n_side=100, spectrum_width=1, max_depth=4, face_grid=10):
    """BASELINE config 5 stand-in: procedural city of n_side x n_side tessellated boxes (default 1.0e7 triangles, facades
    twosided(roughconductor)) on a 1 km ground plane, one wignertransmitter, receive elements on a line: element k is
    received with `scene.update_rect(rx, c5_element_pose(k)); scene.receive(n, seed=k)` into its own 1024-bin profile.
    Returns (scene, rx_shape_id)."""
    p, f = meshes.city_blocks(n_side=n_side, face_grid=face_grid)
    s, rx, _ = radar_scene(p, f, spectrum_width=spectrum_width, max_depth=max_depth, smooth_normals=False,
                           ground_half=0.5 * n_side * 10.0, reflector=False, range_bins=1024, bin_m=0.25, return_ids=True)
    return s, rx
