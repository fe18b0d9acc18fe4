"""INDEPENDENT evaluation of the endpoint / signal variants the reference pins with no golden vector (TEST INFRASTRUCTURE).

tests/plate_path.py walks one radar path sample of the three-rectangle plate scene for the wigner endpoints with a delta
linear-FMCW signal.  This module does the same walk, again in plain Python / numpy written from the reference's TEXT
(file:line below) and not from oracle/oracle.cpp, for the rows DESIGN.md section 2 still listed as "parity unpinned":

  * `phasedreceiver` + `phasedtransmitter` (src/receivers/phasedreceiver.cpp:116-164, 264-345;
    src/transmitters/phasedtransmitter.cpp:104-164, 270-291, 402-489): virtual-element tables, steering phases, the complex sum
    of `sample_wigner`, the receiver's `1 - cos^4` factor, the transmitter's `|pdf * W|`;
  * `pulse`, `cw` and non-delta `linfmcw` signals with `sig_is_delta = false`, `resample_freq = false`
    (wignerreceiver.cpp:112-194, wignertransmitter.cpp:106-146, 423-441): uniform frequency sampling, `eval_signal` / `wchirp`
    (include/mitsuba/core/math.h:79-128) at the receiver and at the transmitter, the frequency-extent factor of the ray weight.

The chirp PSD `sinc(2 pi (f - f_i) T tri)` turns on the fp32 value of the sampled frequency (ulp 8192 Hz at 94 GHz) and on the
wavelength -> frequency round trip at the transmitter, so those operations are carried out in fp32 exactly where the reference
stores a Float / Wavelength and in double where a double literal (`1e-9`) promotes the expression.
"""
import cmath
import math

import numpy as np

from plate_path import F, PCG32, cosine_hemisphere, sinc, tea64, tri

TWO_PI32 = F(6.28318530717958647692)


class Signal:
    """The signal block of an endpoint plugin (wignertransmitter.cpp:59-102 and its three siblings: same fields everywhere)."""

    def __init__(self, kind, delta, resample, amplitude=1.0, repfreq=5000.0, t_ext=200e-6, f_c=94e9, f_ext=6e9, gain=1.0):
        self.kind, self.delta, self.resample = kind, delta, resample
        self.A, self.repfreq, self.T, self.f_c, self.f_ext, self.gain = F(amplitude), F(repfreq), F(t_ext), F(f_c), F(f_ext), F(gain)

    def fold(self, time):
        """math::fmodulo(time, rcp(repfreq)) for |time| < period (math.h:104-119)."""
        period = F(1.0) / self.repfreq
        t = F(abs(F(time)))
        assert t - period < F(2.0 ** -24)
        return period - t if F(time) < 0 else t

    def delta_frequency(self, time):
        """sample_delta_frequency (wignertransmitter.cpp:152-168 / phasedtransmitter.cpp:214-230): weight 1."""
        if self.kind == "linfmcw":
            t = self.fold(time)
            ti = F(0) + self.T / F(2)
            return F(self.f_c + (self.f_ext / self.T) * (t - ti))
        return self.f_c                                        # cw

    def uniform_frequency(self, u):
        """sample_frequency with sig_is_delta = false: sample_shifted<Wavelength> of a width-1 spectrum is the sample itself."""
        return F(F(u) * self.f_ext + (self.f_c - self.f_ext / F(2)))

    def eval(self, time, f):
        """eval_signal(time, frequency).first, every operand a Float (wignertransmitter.cpp:106-146)."""
        if self.kind in ("cw", "other"):
            return float(self.A * self.A)
        t = self.fold(time)
        ti = F(0) + self.T / F(2)
        fi = F(self.f_c + (self.f_ext / self.T) * (t - ti)) if self.kind == "linfmcw" else self.f_c
        x = F(t - ti)
        if not abs(F(x / self.T)) < F(0.5):                     # math::rect
            return 0.0
        df = F(F(f) - fi)                                      # exact: both are fp32 values of ~1e11
        tr = F(1.0) - F(2.0) * abs(F(x / self.T)) if abs(F(x / self.T)) < F(0.5) else F(0.0)
        arg = F(F(F(TWO_PI32 * df) * self.T) * tr)             # TwoPi*f*w*tri(t/w), left to right
        s = math.sin(float(arg)) / float(arg) if abs(arg) > F(2.0 ** -24) else 1.0
        return float(F(F(F(F(2) * self.A * self.A) * self.T) * tr)) * s


def wavelength_nm(c, f):
    """MTS_C * rcp(frequencies) * 1e9 stored in a Wavelength."""
    return F(F(c) * (F(1.0) / F(f)) * F(1e9))


def frequency_of(c32, lam_nm):
    """MTS_C * rcp(wavelengths * 1e-9): double because of the literal; callers round it where the reference stores a Float."""
    return c32 * (1.0 / (float(lam_nm) * 1e-9))


class PhasedArray:
    """Constructor tables of phasedtransmitter.cpp:104-164 (= phasedreceiver.cpp:116-164) for an array whose `array_loc` is a
    translation: elements along `axis` with pitch `spacing`, each `wid` wide, n^2 virtual elements."""

    def __init__(self, n, steer, origin, wid, spacing, axis, lam_min_nm, lam_max_nm):
        self.n, self.wid, self.origin = n, np.asarray(wid, float), np.asarray(origin, float)
        sv = np.sin(np.asarray(steer, float))
        sp, ax = np.asarray(spacing, float), np.asarray(axis, float)
        locs = []
        for i in range(n):
            k = (i - n / 2.0 + 0.5) if n % 2 == 0 else (i - (n - 1.0) / 2.0)
            locs.append(-sp * ax * k)
        k0 = float(F(1.0 / ((float(lam_max_nm) - float(lam_min_nm)) * 1e-9 / 2)))      # static_cast<ScalarFloat>(rcp(...))
        self.centres, self.r_dash, self.psi_dash = [], [], []
        for i in range(n):
            for j in range(n):
                self.centres.append((locs[i] + locs[j]) / 2)
                self.r_dash.append(locs[i] - locs[j])
                self.psi_dash.append(cmath.exp(1j * k0 * float((-self.r_dash[-1]) @ sv)))

    def wigner(self, p_world, d, lam_nm):
        """sample_wigner (phasedtransmitter.cpp:270-291): d is whatever the caller left in ds.d."""
        lam_m = float(lam_nm) * 1e-9
        W = 0j
        for c, rd, psi in zip(self.centres, self.r_dash, self.psi_dash):
            # velem_to_object = inverse(array_loc * translate(r_v) * scale(wid.x/2, wid.y/2, wid.z)); r_hat = that * p / 2
            q = (np.asarray(p_world, float) - self.origin - c)
            rx, ry = q[0] / (self.wid[0] / 2) / 2, q[1] / (self.wid[1] / 2) / 2
            if abs(rx) <= 0.5 and abs(ry) <= 0.5:
                nu = np.asarray(d, float) / lam_m            # from_frame of an axis-aligned frame is the identity (transform.h:284-295)
                w = 4 * self.wid[0] * self.wid[1] * tri(rx) * tri(ry) * \
                    sinc(2 * math.pi * nu[0] * self.wid[0] * tri(rx)) * sinc(2 * math.pi * nu[1] * self.wid[1] * tri(ry))
                W += w * cmath.exp(2j * math.pi * float(nu @ rd)) * psi
        return W.real


def rect_wigner(local_xy, dvec, lam_nm, wid):
    """Rectangle::sample_wigner for an axis-aligned aperture (src/shapes/rectangle.cpp:132-220), as plate_path.wigner_gain."""
    rx, ry = local_xy[0] / 2.0, local_xy[1] / 2.0
    lam_m = float(lam_nm) * 1e-9
    nux, nuy = dvec[0] / lam_m, dvec[1] / lam_m
    return 4.0 * tri(rx) * tri(ry) * sinc(2.0 * math.pi * nux * wid * tri(rx)) * sinc(2.0 * math.pi * nuy * wid * tri(ry))


class Variant:
    """What differs from plate_path.evaluate_sample: the endpoint plugins and their signal blocks."""

    def __init__(self, sc, endpoints="wigner", rx_signal=None, tx_signal=None, n_elems=2, steer=(0.0, 0.2, 0.0), second_tx=None):
        self.endpoints = endpoints
        self.second_tx = second_tx      # dict(y=..., gain=...): a second wignertransmitter on its own aperture (Scene::sample_transmitter_direction picks one)
        base = dict(repfreq=sc.repfreq, t_ext=sc.t_ext, f_c=sc.f_c, f_ext=sc.f_ext)
        self.rx_signal = rx_signal or Signal("linfmcw", True, False, **base)
        self.tx_signal = tx_signal or Signal("linfmcw", True, True, **base)
        if endpoints == "phased":
            w = 2.0 * sc.a
            self.pitch = w / n_elems
            self.n_elems, self.steer = n_elems, steer
            geo = dict(wid=(w, self.pitch, 1.0), spacing=(self.pitch,) * 3, axis=(0, 1, 0), lam_min_nm=sc.lam_min, lam_max_nm=sc.lam_max)
            self.rx_array = PhasedArray(n_elems, steer, (0.0, 0.0, 0.0), **geo)
            self.tx_array = PhasedArray(n_elems, steer, (0.0, sc.tx_y, 0.0), **geo)


def evaluate_sample(sc, var, seed, idx):
    """-> dict(first_t, valid, tf_t, tf_f, value, depth) of sample `idx`, None when the path leaves the modelled cases."""
    rng = PCG32(tea64(seed, idx), tea64(idx, seed))
    # ---- receive_sample head: src/librender/integrator.cpp:1544-1569 ----
    pos = (rng.next_float(), rng.next_float())
    aperture = (rng.next_float(), rng.next_float())
    time = F(F(sc.adc_start) + F(rng.next_float()) * F(sc.adc_time))
    wavelength_sample = rng.next_float()
    # ---- Receiver::sample_ray_differential (wignerreceiver.cpp:208-269 / phasedreceiver.cpp:264-345): mix_resample ----
    rs = var.rx_signal
    if rs.delta:
        f_sampled, power_rx = rs.delta_frequency(time), 1.0
    else:
        f_sampled = rs.uniform_frequency(wavelength_sample)
        power_rx = rs.eval(time, f_sampled)
    lam = wavelength_nm(sc.c, f_sampled)
    a = sc.a
    local = (2.0 * pos[0] - 1.0, 2.0 * pos[1] - 1.0)
    o = np.array([a * local[0], a * local[1], 0.0])              # Rectangle::sample_position, rectangle.cpp:111-125
    area = (2.0 * a) * (2.0 * a)
    d = cosine_hemisphere(*aperture)                             # the receiver rectangle's frame is the identity
    if var.endpoints == "phased":
        cos_n = d[2]                                             # dot(ds.d, ds.n), ds.n = (0, 0, 1)
        geom_rx = var.rx_array.wigner(o, d, lam) * (1.0 / area) * (1.0 - cos_n ** 4)
    else:
        geom_rx = rect_wigner(local, d, lam, 2.0 * a) * (1.0 / area)
    extents = area * math.pi
    if not rs.delta:
        extents *= sc.c32 * (1.0 / float(rs.f_ext)) * 1e9         # extents *= MTS_C*rcp(m_sig_f_ext)*1e9
    weight = power_rx * float(rs.gain) * geom_rx * extents
    f_rx = F(frequency_of(sc.c32, lam))                          # integrator.cpp:1592, stored in a Wavelength
    tf_t = float(time - F(sc.adc_start))
    out = dict(first_t=math.inf, valid=0, value=0.0, depth=1, tf_t=tf_t * sc.t_bins / sc.t_bw,
               tf_f=abs(frequency_of(sc.c32, lam) - float(f_rx)) * sc.f_bins / sc.f_bw)
    # ---- first hit: the plate at z = R ----
    if d[2] <= 0.0:
        return out
    t1 = (sc.R - o[2]) / d[2]
    p1 = o + t1 * d
    if abs(p1[0]) > sc.plate_half or abs(p1[1]) > sc.plate_half:
        return out
    out["first_t"], out["valid"] = t1, 1
    ray_time = F(time + F(-t1) / F(sc.c))                         # Ray::update_state, ray.h:89-93
    # ---- NEE: Scene::sample_transmitter_direction -> Transmitter::sample_direction ----
    s2 = (rng.next_float(), rng.next_float())
    tx_y, tx_gain, pick = sc.tx_y, 1.0, 1.0
    if var.second_tx is not None:
        # scene.cpp:262-274: index = min(uint(sample.x * n), n - 1); the sample is re-stretched, pdf *= 1 / n, value *= n
        index = min(int(float(F(F(s2[0]) * F(2.0)))), 1)
        s2 = (float(F(F(F(s2[0]) - F(index) * F(0.5)) * F(2.0))), s2[1])
        pick = 0.5
        if index == 1:
            tx_y, tx_gain = var.second_tx["y"], var.second_tx["gain"]
    tloc = (2.0 * s2[0] - 1.0, 2.0 * s2[1] - 1.0)
    p_tx = np.array([a * tloc[0], tx_y + a * tloc[1], 0.0])
    dv = p_tx - p1
    dist2 = float(dv @ dv); dist = math.sqrt(dist2); dv = dv / dist
    dp = abs(dv[2])
    pdf = (1.0 / area) * (dist2 / dp if dp != 0.0 else 0.0)      # Shape::sample_direction, shape.cpp:323-342
    active = dv[2] < 0.0 and pdf != 0.0
    geom = 1.0 / pdf if pdf != 0.0 else 0.0                      # antenna_texture = 1
    ds_time = F(ray_time + F(-dist) / F(sc.c))                    # ds.time += -ds.dist / MTS_C
    ts = var.tx_signal
    lam2 = lam
    if ts.resample:
        lam2 = wavelength_nm(sc.c, ts.delta_frequency(ds_time))   # it.wavelengths overwritten; weight 1
        power_tx = 1.0
    else:
        power_tx = ts.eval(ds_time, F(frequency_of(sc.c32, lam)))  # eval_signal(ds.time, MTS_C*rcp(it.wavelengths*1e-9))
    if var.endpoints == "phased":
        wgain = var.tx_array.wigner(p_tx, -dv, lam2)              # phasedtransmitter.cpp:470-478
        tx_val = power_tx * float(ts.gain) * geom * wgain
        pdf = abs(pdf * wgain)
    else:
        wgain = rect_wigner(tloc, -dv, lam2, 2.0 * a)             # wignertransmitter.cpp:443-452
        tx_val = power_tx * float(ts.gain) * tx_gain * geom * wgain * ((1.0 / area) * 2.0 * math.pi)
        pdf = pdf * wgain
    pdf *= pick                                                   # ds.pdf *= transmitter_pdf; spec *= rcp(transmitter_pdf)
    tx_val *= 1.0 / pick
    if not active:
        tx_val = 0.0
    # diffuse plate facing -z (src/bsdfs/diffuse.cpp:78-135) and the MIS weight of pathtimefrequency.cpp:276-300, 472-477
    cos_i, cos_o = d[2], -dv[2]
    res = 0.0
    if active and pdf != 0.0 and cos_i > 0.0 and cos_o > 0.0:
        bsdf_val, bsdf_pdf = sc.rho / math.pi * cos_o, cos_o / math.pi
        mis = (pdf * pdf) / (pdf * pdf + bsdf_pdf * bsdf_pdf)
        res = mis * bsdf_val * tx_val
    # ---- BSDF sampling + continuation ray: only its geometry matters here (max_depth 2 ends the path at the next hit) ----
    rng.next_float()
    s3 = (rng.next_float(), rng.next_float())
    wo = cosine_hemisphere(*s3)
    d2 = np.array([wo[0], -wo[1], -wo[2]])
    out["depth"] = 2
    if d2[2] < 0.0:
        t0 = (0.0 - p1[2]) / d2[2]
        q = p1 + t0 * d2
        ys = [0.0, sc.tx_y] + ([var.second_tx["y"]] if var.second_tx is not None else [])
        if any(abs(q[0]) <= a and abs(q[1] - y0) <= a for y0 in ys):
            return None                                            # lands on an aperture: not modelled
    # ---- receive_sample tail: integrator.cpp:1588-1665 ----
    out["tf_f"] = abs(frequency_of(sc.c32, lam2) - float(f_rx)) * sc.f_bins / sc.f_bw
    out["value"] = abs(weight) * res
    return out
