"""An INDEPENDENT evaluation of one radar path sample on a three-rectangle scene (TEST INFRASTRUCTURE).

VERDICT r01 weak #1: the oracle (oracle/oracle.cpp) and the kernels are two copies of one transcription of the reference, and the
radar-specific arithmetic -- `receive_sample` binning, `pathtimefrequency` time / phase bookkeeping, the wigner endpoints,
`SignalBlock::put`, `PhaseIntegrator` -- is pinned by no golden vector of the reference.  This module is a third
implementation of exactly those rows, written from the reference's text (file:line below), NOT from oracle.cpp, in plain
Python / numpy float64 with float32 rounding only where the reference stores a float (sample time, frequencies, wavelengths:
the beat frequency is a difference of two ~1e11 Hz values, so those roundings decide the bin).  For sample index i it draws the
same random numbers (PCG32 seeded with (tea64(seed, i), tea64(i, seed)), src/librender/sampler.cpp:89-92) and walks the one
path the integrator can take on this scene:

    receiver rectangle at the origin facing +z  ->  diffuse plate at z = R facing -z  ->  NEE to the transmitter rectangle
                                                 ->  BSDF-sampled continuation ray  ->  floor at z = -1 (max_depth 2 ends there)

and returns what `bf_sample_record` exposes: first-hit distance, valid, the fractional bin coordinates tf, the value
hsum(|ray_weight| * L), and the phase the PhaseIntegrator sees.
"""
import math

import numpy as np

F = np.float32
MASK64 = (1 << 64) - 1
PCG_MULT = 0x5851f42d4c957f2d
RAY_EPS = 1500.0 * (2.0 ** -24)          # math::RayEpsilon = Epsilon * 1500, Epsilon = 2^-24 (include/mitsuba/core/math.h:34-38)


def tea64(v0, v1, rounds=4):
    """sample_tea_64 instantiated with 64-bit arguments (include/mitsuba/core/random.h:104-116; sampler.cpp:89-92 passes UInt64)."""
    s = 0
    for _ in range(rounds):
        s = (s + 0x9e3779b9) & MASK64
        v0 = (v0 + ((((v1 << 4) & MASK64) + 0xa341316c) ^ ((v1 + s) & MASK64) ^ (((v1 >> 5) + 0xc8013ea4) & MASK64))) & MASK64
        v1 = (v1 + ((((v0 << 4) & MASK64) + 0xad90777d) ^ ((v0 + s) & MASK64) ^ (((v0 >> 5) + 0x7e95761e) & MASK64))) & MASK64
    return (v0 + ((v1 << 32) & MASK64)) & MASK64


class PCG32:
    """The public pcg32 generator (O'Neill; enoki/random.h wraps it): seed(initstate, initseq), next_float32 = (u >> 9 | 0x3f800000) - 1."""

    def __init__(self, initstate, initseq):
        self.state, self.inc = 0, ((initseq << 1) | 1) & MASK64
        self.next_u32()
        self.state = (self.state + initstate) & MASK64
        self.next_u32()

    def next_u32(self):
        old = self.state
        self.state = (old * PCG_MULT + self.inc) & MASK64
        xorshifted = (((old >> 18) ^ old) >> 27) & 0xffffffff
        rot = old >> 59
        return ((xorshifted >> rot) | (xorshifted << ((-rot) & 31))) & 0xffffffff

    def next_float(self):
        return float(np.array([(self.next_u32() >> 9) | 0x3f800000], np.uint32).view(np.float32)[0] - F(1.0))


def tri(x):
    return 1.0 - 2.0 * abs(x) if abs(x) < 0.5 else 0.0          # math.h:80-82


def sinc(x):
    return math.sin(x) / x if abs(x) > 2.0 ** -24 else 1.0       # math.h:76-78


def cosine_hemisphere(u, v):
    """warp::square_to_cosine_hemisphere over square_to_uniform_disk_concentric (include/mitsuba/core/warp.h:54-90, 325-349)."""
    x, y = 2.0 * u - 1.0, 2.0 * v - 1.0
    if x == 0.0 and y == 0.0:
        r, phi = 0.0, 0.0
    elif abs(x) < abs(y):
        r, phi = y, 0.5 * math.pi - 0.25 * math.pi * x / y
    else:
        r, phi = x, 0.25 * math.pi * y / x
    px, py = r * math.cos(phi), r * math.sin(phi)
    return np.array([px, py, math.sqrt(max(0.0, 1.0 - px * px - py * py))])


class PlateScene:
    """Geometry and signal parameters; the same numbers build the FlatScene in tests/test_gpu_anchor.py."""

    def __init__(self, R=5.0, a=0.0025, tx_y=0.01, plate_half=3.0, floor_half=30.0, rho=0.7, floor_z=-1.0,
                 f_c=94e9, f_ext=6e9, t_ext=200e-6, repfreq=5000.0, adc_start=20e-6, adc_time=160e-6,
                 t_bins=4, f_bins=256, t_bw=200e-6, f_bw=3.0e6, c=299792458.0):
        self.__dict__.update(locals())
        del self.__dict__["self"]
        self.c32 = float(F(c))               # MTS_C = math::CVac<Float> is a FLOAT constant (spectrum.h:32-33): 299792448
        self.lam_min = F(c / 97e9 * 1e9)     # RADAR_94GHZ constants of beifong_b200/scene.py (spectrum.h:17,25 alternative)
        self.lam_max = F(c / 91e9 * 1e9)

    # wignerreceiver.cpp:163-175 == wignertransmitter.cpp sample_delta_frequency, in the fp32 operations the reference performs
    def delta_frequency(self, time):
        period = F(1.0) / F(self.repfreq)
        t = F(abs(F(time)))                   # fmodulo (math.h:92-104) leaves |a| < b unchanged; a < 0 gives b - |a|
        assert t < period
        if F(time) < 0:
            t = period - t
        ti = F(0) + F(self.t_ext) / F(2)
        return F(F(self.f_c) + (F(self.f_ext) / F(self.t_ext)) * (t - ti))

    def wavelength_nm(self, f):
        return F(F(self.c) * (F(1.0) / f) * F(1e9))         # MTS_C*rcp(frequencies)*1e9

    def wigner_gain(self, local_xy, dvec, lam_nm, wid):
        """Rectangle::sample_wigner (src/shapes/rectangle.cpp:132-220) for an axis-aligned aperture: r_hat = local / 2,
        nu_hat = (d expressed in the aperture frame) / (lambda * 1e-9)."""
        rx, ry = local_xy[0] / 2.0, local_xy[1] / 2.0
        lam_m = float(lam_nm) * 1e-9
        nux, nuy = dvec[0] / lam_m, dvec[1] / lam_m
        return 4.0 * tri(rx) * tri(ry) * sinc(2.0 * math.pi * nux * wid * tri(rx)) * sinc(2.0 * math.pi * nuy * wid * tri(ry))


def evaluate_sample(sc, seed, idx, plate_velocity=None, max_depth=2):
    """-> dict(first_t, valid, tf_t, tf_f, value, phase, depth, wavelength0) of sample `idx`; None when the path leaves the modelled cases.
    plate_velocity = (vx, vy, vz): the integrator's doppler flag is set and the plate's `velocity` transform is that translation."""
    rng = PCG32(tea64(seed, idx), tea64(idx, seed))
    # ---- receive_sample head: src/librender/integrator.cpp:1544-1569 ----
    pos = (rng.next_float(), rng.next_float())
    aperture = (rng.next_float(), rng.next_float())
    time = F(F(sc.adc_start) + F(rng.next_float()) * F(sc.adc_time))
    rng.next_float()                                             # wavelength_sample: unused by a delta signal
    # ---- Wignerreceiver::sample_ray_differential: src/receivers/wignerreceiver.cpp:208-269 ----
    f_rx_f = sc.delta_frequency(time)
    lam = sc.wavelength_nm(f_rx_f)
    a = sc.a
    local = (2.0 * pos[0] - 1.0, 2.0 * pos[1] - 1.0)
    o = np.array([a * local[0], a * local[1], 0.0])              # Rectangle::sample_position, rectangle.cpp:111-125
    area = (2.0 * a) * (2.0 * a)                                 # |dp_du x dp_dv|, dp_du = to_world * (2, 0, 0)
    dloc = cosine_hemisphere(*aperture)
    gain_rx = sc.wigner_gain(local, dloc, lam, 2.0 * a)          # ds.d is the LOCAL direction here (wignerreceiver.cpp:243-246)
    weight = 1.0 * 1.0 * (gain_rx * (1.0 / area)) * (area * math.pi)
    d = dloc                                                     # Frame3f((0, 0, 1)) is the identity (vector.h:116-135)
    f_rx = sc.c32 * (1.0 / (float(lam) * 1e-9))                    # integrator.cpp:1592, double because of the 1e-9 literal
    tf_t = float(time - F(sc.adc_start))
    out = dict(first_t=math.inf, valid=0, value=0.0, phase=0.0, depth=1, tf_t=tf_t * sc.t_bins / sc.t_bw, tf_f=0.0)
    # ---- first hit: plate at z = R (Rectangle::ray_intersect, rectangle.cpp:229-263) ----
    if d[2] <= 0.0:
        return out
    t1 = (sc.R - o[2]) / d[2]
    p1 = o + t1 * d
    if abs(p1[0]) > sc.plate_half or abs(p1[1]) > sc.plate_half:
        return out                                                # passes the plate: nothing else lies above the antennas
    out["first_t"], out["valid"] = t1, 1
    # ---- Ray::update_state(-t) (include/mitsuba/core/ray.h:89-93) ----
    denom = float((sc.lam_max - sc.lam_min) / F(2)) * 1e-9
    ray_time = F(time + F(-t1) / F(sc.c))
    # ---- NEE: Scene::sample_transmitter_direction (scene.cpp:249-299) -> WignerTransmitter::sample_direction (wignertransmitter.cpp:373-534)
    s2 = (rng.next_float(), rng.next_float())
    tloc = (2.0 * s2[0] - 1.0, 2.0 * s2[1] - 1.0)
    p_tx = np.array([a * tloc[0], sc.tx_y + a * tloc[1], 0.0])
    dv = p_tx - p1
    dist2 = float(dv @ dv); dist = math.sqrt(dist2); dv = dv / dist
    n_tx = np.array([0.0, 0.0, 1.0])
    dp = abs(float(dv @ n_tx))
    pdf = (1.0 / area) * (dist2 / dp if dp != 0.0 else 0.0)      # Shape::sample_direction, shape.cpp:323-342
    active = float(dv @ n_tx) < 0.0 and pdf != 0.0
    geom = 1.0 / pdf if pdf != 0.0 else 0.0
    ds_time = F(ray_time + F(-dist) / F(sc.c))                    # wignertransmitter.cpp:422-425
    f_tx = sc.delta_frequency(ds_time)                            # resample_freq: the interaction's wavelengths are overwritten (:436-441)
    lam2 = sc.wavelength_nm(f_tx)
    gain_tx = sc.wigner_gain(tloc, -dv, lam2, 2.0 * a)
    geom *= gain_tx
    pdf *= gain_tx
    tx_val = (1.0 * 1.0 * geom * ((1.0 / area) * 2.0 * math.pi)) if active else 0.0
    # diffuse plate facing -z: cos_theta_o = n . wo, eval = rho / pi * cos_theta_o, pdf = cos_theta_o / pi (src/bsdfs/diffuse.cpp:78-135)
    n_p = np.array([0.0, 0.0, -1.0])
    cos_i, cos_o = float(n_p @ (-d)), float(n_p @ dv)
    res = 0.0
    if active and cos_i > 0.0 and cos_o > 0.0:
        bsdf_val, bsdf_pdf = sc.rho / math.pi * cos_o, cos_o / math.pi
        mis = (pdf * pdf) / (pdf * pdf + bsdf_pdf * bsdf_pdf) if pdf * pdf > 0.0 else 0.0      # pathtimefrequency.cpp:472-477
        res = mis * 1.0 * bsdf_val * tx_val
    # ---- BSDF sampling + continuation ray (pathtimefrequency.cpp:346-371) ----
    rng.next_float()
    s3 = (rng.next_float(), rng.next_float())
    wo = cosine_hemisphere(*s3)
    d2 = np.array([wo[0], -wo[1], -wo[2]])                        # plate shading frame: s = (1, 0, 0), t = n x s = (0, -1, 0), n = (0, 0, -1)
    phase = -math.inf
    out["depth"] = 2
    if d2[2] < 0.0:
        t0 = (0.0 - p1[2]) / d2[2]                                # the antenna plane z = 0: a hit on an aperture is not modelled
        q = p1 + t0 * d2
        if (abs(q[0]) <= a and abs(q[1]) <= a) or (abs(q[0]) <= a and abs(q[1] - sc.tx_y) <= a):
            return None
        t2 = (sc.floor_z - p1[2]) / d2[2]
        p2 = p1 + t2 * d2
        if abs(p2[0]) <= sc.floor_half and abs(p2[1]) <= sc.floor_half:
            phase = 2.0 * math.pi * (-t2) / denom                 # spawn_ray starts a fresh ray (phase 0); ray_.phase += ray.phase
            if max_depth >= 3:
                # ---- second loop iteration, at the floor (pathtimefrequency.cpp:153-345 once more) ----
                # spawn_ray carries the interaction's time and (resampled) wavelengths; update_state(-t2) advances the time
                ray_time2 = F(ray_time + F(-t2) / F(sc.c))
                # NEE: the transmitter faces +z and is seen from below, so `dot(ds.d, ds.n) < 0` fails and the contribution is
                # masked out -- but sample_direction has already overwritten the interaction's wavelengths with the delta
                # frequency at ds.time (wignertransmitter.cpp:436-441, unconditional in a scalar variant)
                s4 = (rng.next_float(), rng.next_float())
                tl = (2.0 * s4[0] - 1.0, 2.0 * s4[1] - 1.0)
                p_tx2 = np.array([a * tl[0], sc.tx_y + a * tl[1], 0.0])
                dist_b = math.sqrt(float((p_tx2 - p2) @ (p_tx2 - p2)))
                ds_time2 = F(ray_time2 + F(-dist_b) / F(sc.c))
                lam2 = sc.wavelength_nm(sc.delta_frequency(ds_time2))
                # BSDF sampling on the floor (frame = world axes), third segment
                rng.next_float()
                s5 = (rng.next_float(), rng.next_float())
                d3 = cosine_hemisphere(*s5)
                out["depth"] = 3
                phase = -math.inf
                if d3[2] > 0.0:
                    tz = (0.0 - p2[2]) / d3[2]
                    qz = p2 + tz * d3
                    if (abs(qz[0]) <= a and abs(qz[1]) <= a) or (abs(qz[0]) <= a and abs(qz[1] - sc.tx_y) <= a):
                        return None                               # reaches an aperture from behind: not modelled
                    t3 = (sc.R - p2[2]) / d3[2]
                    p3 = p2 + t3 * d3
                    if abs(p3[0]) <= sc.plate_half and abs(p3[1]) <= sc.plate_half:
                        phase = 2.0 * math.pi * (-t3) / denom
    if plate_velocity is not None:
        # Shape::doppler (src/librender/shape.cpp:375-389) at the statement of pathtimefrequency.cpp:254-256 (commented out in the
        # reference, behind a flag here): 2 * dot(si.wi, velocity * Point3f(si.to_local(si.p))) / MTS_C * si.wavelengths, with the hit
        # position expressed in the plate's shading frame (s, t, n) = ((1, 0, 0), (0, -1, 0), (0, 0, -1)), as the reference writes it
        p_loc = np.array([p1[0], -p1[1], -p1[2]]) + np.asarray(plate_velocity, float)
        wi = np.array([-d[0], d[1], d[2]])
        fshift = F(2.0 * float(wi @ p_loc) / sc.c32)
        lam2 = F(lam2 + fshift * lam2)
    out["wavelength0"] = float(lam2)
    # ---- receive_sample tail: integrator.cpp:1588-1665 ----
    f_end = sc.c32 * (1.0 / (float(lam2) * 1e-9))
    out["tf_f"] = abs(f_end - float(F(f_rx))) * sc.f_bins / sc.f_bw      # f_rx is stored in a fp32 Wavelength (:1592)
    out["value"] = abs(weight) * res
    out["radiance"] = res                                         # what the PhaseIntegrator writes to its AOV: hsum(L) before the ray weight (phase.cpp:117-124)
    out["phase"] = phase
    return out
