"""INDEPENDENT evaluation of one `render_sample` of the gen-1 / gen-2 integrators on a plate scene (TEST INFRASTRUCTURE).

The companion of tests/plate_path.py for the image-side entry point: `SamplingIntegrator::render_sample`
(src/librender/integrator.cpp:255-310) with a `perspective` sensor (src/sensors/perspective.cpp:95-206,
include/mitsuba/render/sensor.h:196-231, include/mitsuba/core/transform.h:203-220), an `area` emitter
(src/emitters/area.cpp:64-72, 125-139), a diffuse plate, and the wrapped integrators

    gen 2:  range(pathlength)   src/integrators/pathlength.cpp:98-233 + range.cpp:78-118
    gen 1:  time(pathtime)      src/integrators/pathtime.cpp + time.cpp (same walk; the scalar is a travel time)

written in plain Python / numpy float64 from the reference's TEXT, not from oracle/oracle.cpp.  Scene: camera at the origin looking
down +z, emitter rectangle beside it in the plane z = 0 facing +z, diffuse plate at z = R facing -z, floor at z = -1; `max_depth 2`,
so a sample is: camera ray -> plate -> emitter sampling (NEE) -> BSDF-sampled ray -> floor (or nothing), and the integrator returns
(radiance, valid, scalar) with the reference's own bookkeeping of the scalar (the first-hit distance is added once after the first
intersection and AGAIN inside the emitter-sampling block, pathlength.cpp:122, :188).
"""
import math

import numpy as np

from plate_path import F, PCG32, cosine_hemisphere, tea64


class RenderScene:
    def __init__(self, R=4.0, plate_half=3.0, rho=0.8, floor_z=-1.0, floor_half=30.0, em_x=0.8, em_half=0.25, radiance=100.0,
                 fov=45.0, near=0.2, far=10.2, dr=0.2, bins=50):
        self.__dict__.update(locals())
        del self.__dict__["self"]


def evaluate_render_sample(sc, seed, idx):
    """-> dict(pos, first_t, valid, value, scalar, depth[, on_emitter])."""
    rng = PCG32(tea64(seed, idx), tea64(idx, seed))
    # ---- render_sample head (integrator.cpp:259-274): pixel (0, 0) of a 1 x 1 film, no aperture sample, shutter closed ----
    pos = (rng.next_float(), rng.next_float())
    rng.next_float()                                              # wavelength_sample: a monochromatic ray weighs 1
    # ---- PerspectiveCamera::sample_ray_differential: near_p = sample_to_camera * (x, y, 0) ----
    # camera_to_sample = scale(-1/2, -1/2, 1) translate(-1, -1, 0) perspective(fov, near, far) for a square uncropped film
    # (sensor.h:224-230), so a sample (sx, sy, 0) comes from the point ((1 - 2 sx) near tan, (1 - 2 sy) near tan, near)
    tan = math.tan(math.radians(sc.fov * 0.5))
    near_p = np.array([(1.0 - 2.0 * pos[0]) * sc.near * tan, (1.0 - 2.0 * pos[1]) * sc.near * tan, sc.near])
    d = near_p / np.linalg.norm(near_p)
    mint, maxt = sc.near / d[2], sc.far / d[2]
    o = np.zeros(3)
    out = dict(pos=pos, first_t=math.inf, valid=0, value=0.0, scalar=0.0, depth=1)
    # ---- first intersection: the plate at z = R ----
    t1 = sc.R / d[2]
    p1 = o + t1 * d
    if not (mint <= t1 <= maxt) or abs(p1[0]) > sc.plate_half or abs(p1[1]) > sc.plate_half:
        return out
    out["first_t"], out["valid"] = t1, 1
    scalar = t1                                                   # pathlength += si.t after the first intersection (:122)
    # depth 1: the plate is no emitter; max_depth 2 > 1: emitter sampling
    s2 = (rng.next_float(), rng.next_float())
    a = sc.em_half
    p_e = np.array([sc.em_x + a * (2.0 * s2[0] - 1.0), a * (2.0 * s2[1] - 1.0), 0.0])     # Rectangle::sample_position
    area = (2.0 * a) ** 2
    dv = p_e - p1
    dist2 = float(dv @ dv); dist = math.sqrt(dist2); dv = dv / dist
    dp = abs(dv[2])                                               # emitter normal (0, 0, 1)
    pdf = (1.0 / area) * (dist2 / dp) if dp != 0.0 else 0.0       # Shape::sample_direction, shape.cpp:323-342
    active_e = dv[2] < 0.0 and pdf != 0.0                         # area.cpp:135
    emitter_val = sc.radiance / pdf if active_e else 0.0          # nothing stands between the plate and the emitter
    cos_i, cos_o = d[2], -dv[2]                                   # plate normal (0, 0, -1): wi = -d, wo = dv
    res = 0.0
    if active_e and cos_i > 0.0 and cos_o > 0.0:
        bsdf_val, bsdf_pdf = sc.rho / math.pi * cos_o, cos_o / math.pi
        mis = (pdf * pdf) / (pdf * pdf + bsdf_pdf * bsdf_pdf)
        res = mis * bsdf_val * emitter_val
    scalar += t1                                                  # ... and again inside the emitter-sampling block (:188)
    # BSDF sampling, continuation ray (pathlength.cpp:192-205)
    rng.next_float()
    s3 = (rng.next_float(), rng.next_float())
    wo = cosine_hemisphere(*s3)
    d2 = np.array([wo[0], -wo[1], -wo[2]])                        # plate frame (s, t, n) = ((1, 0, 0), (0, -1, 0), (0, 0, -1))
    out["depth"] = 2
    if d2[2] < 0.0:
        t0 = -p1[2] / d2[2]
        q = p1 + t0 * d2
        if abs(q[0] - sc.em_x) <= a and abs(q[1]) <= a:
            # The BSDF-sampled ray lands on the emitter (pathlength.cpp:207-226, then the top of the loop at depth 2, :104-109):
            # emission weighted against emitter sampling, throughput = rho (cosine-sampled diffuse lobe), and the segment
            # length added twice -- once for the new intersection, once more where the emission is evaluated.
            bs_pdf = wo[2] / math.pi
            em_pdf = (1.0 / area) * (t0 * t0) / abs(d2[2])        # Shape::pdf_direction for the point hit
            w_em = (bs_pdf * bs_pdf) / (bs_pdf * bs_pdf + em_pdf * em_pdf)
            out["value"], out["scalar"], out["on_emitter"] = res + w_em * sc.rho * sc.radiance, scalar + 2.0 * t0, True
            return out
        t2 = (sc.floor_z - p1[2]) / d2[2]
        p2 = p1 + t2 * d2
        if abs(p2[0]) <= sc.floor_half and abs(p2[1]) <= sc.floor_half:
            scalar += t2                                          # pathlength += si.t of the new intersection (:225)
    out["value"], out["scalar"] = res, scalar
    return out


def range_bin(sc, scalar):
    """range.cpp:102-105: bin i takes ranges >= i dr and < i dr + dr, compared in fp32."""
    for i in range(sc.bins):
        lo = F(F(i) * F(sc.dr))
        hi = F(lo + F(sc.dr))
        if F(scalar) >= lo and F(scalar) < hi:
            return i
    return -1


class FluxScene:
    """gen 1: `fluxmeter` on a rectangle at the origin facing +z (src/sensors/fluxmeter.cpp:44-84), `spot` emitter beside it looking
    down +z (src/emitters/spot.cpp:69-151), diffuse plate at z = R, floor at z = floor_z; time(pathtime), 50 bins of 0.5 ns."""

    def __init__(self, R=1.5, plate_half=2.0, rho=0.8, floor_z=-0.5, floor_half=30.0, a=0.05, spot=(0.3, 0.0, 0.02),
                 intensity=2.0, cutoff_deg=60.0, beam_deg=35.0):
        self.__dict__.update(locals())
        del self.__dict__["self"]


def evaluate_flux_sample(sc, seed, idx):
    """-> dict(pos, first_t, valid, value, scalar, depth) of one render_sample; None when the BSDF-sampled ray lands on the sensor's rectangle."""
    rng = PCG32(tea64(seed, idx), tea64(idx, seed))
    pos = (rng.next_float(), rng.next_float())                    # integrator.cpp:259
    aperture = (rng.next_float(), rng.next_float())               # needs_aperture_sample(): endpoint.h:241 default true
    rng.next_float()                                              # wavelength_sample
    # FluxMeter::sample_ray_differential: position on the shape, cosine-weighted direction about its normal, weight pi
    o = np.array([sc.a * (2.0 * pos[0] - 1.0), sc.a * (2.0 * pos[1] - 1.0), 0.0])
    d = cosine_hemisphere(*aperture)                              # Frame3f((0, 0, 1)) is the identity
    weight = math.pi
    out = dict(pos=pos, first_t=math.inf, valid=0, value=0.0, radiance=0.0, scalar=0.0, depth=1)
    if d[2] <= 0.0:
        return out
    t1 = sc.R / d[2]
    p1 = o + t1 * d
    if abs(p1[0]) > sc.plate_half or abs(p1[1]) > sc.plate_half:
        return out                                                # nothing else lies above the sensor
    out["first_t"], out["valid"] = t1, 1
    scalar = float(F(F(t1) / F(3.0e8)))                           # pathtime = si.t / 3.0e8 (pathtime.cpp:121)
    # emitter sampling: SpotLight::sample_direction (delta position: pdf 1, mis 1), falloff_curve
    rng.next_float(); rng.next_float()                            # the 2-D sample is drawn and ignored
    ps = np.asarray(sc.spot, float)
    dv = ps - p1
    dist = math.sqrt(float(dv @ dv)); dv = dv / dist
    local_d = -dv                                                 # trafo.inverse() * -ds.d, the spot's frame is the world's
    cos_theta = local_d[2] / np.linalg.norm(local_d)
    cutoff, beam = math.radians(sc.cutoff_deg), math.radians(sc.beam_deg)
    if cos_theta <= math.cos(cutoff):
        fall = 0.0
    elif cos_theta >= math.cos(beam):
        fall = sc.intensity
    else:
        fall = sc.intensity * ((cutoff - math.acos(cos_theta)) * (1.0 / (cutoff - beam)))
    emitter_val = fall * (1.0 / dist) ** 2
    cos_i, cos_o = d[2], -dv[2]
    res = 0.0
    if cos_i > 0.0 and cos_o > 0.0:
        res = (sc.rho / math.pi * cos_o) * emitter_val
    rng.next_float()
    s3 = (rng.next_float(), rng.next_float())
    wo = cosine_hemisphere(*s3)
    d2 = np.array([wo[0], -wo[1], -wo[2]])
    out["depth"] = 2
    if d2[2] < 0.0:
        t0 = -p1[2] / d2[2]
        q = p1 + t0 * d2
        if abs(q[0]) <= sc.a and abs(q[1]) <= sc.a:
            return None
        t2 = (sc.floor_z - p1[2]) / d2[2]
        p2 = p1 + t2 * d2
        if abs(p2[0]) <= sc.floor_half and abs(p2[1]) <= sc.floor_half:
            scalar = float(F(F(scalar) + F(F(t2) / F(3.0e8))))    # pathtime += si.t / 3.0e8 (:209)
    out["value"], out["radiance"], out["scalar"] = weight * res, res, scalar     # aovs[0] = ray_weight * L; the wrappers bin the unweighted L
    return out
