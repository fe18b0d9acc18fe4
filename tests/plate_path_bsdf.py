"""INDEPENDENT evaluation of the emitter-sampling term on a rough conducting plate (TEST INFRASTRUCTURE).

The plate of tests/plate_path.py with the car's material class: `roughconductor` (src/bsdfs/roughconductor.cpp:283-394) over the
Beckmann `MicrofacetDistribution` (include/mitsuba/render/microfacet.h:184-223, 331-355) and `fresnel_conductor`
(include/mitsuba/render/fresnel.h:92-116), written from the reference's text in numpy float64.  What is evaluated per sample is what
the first loop iteration adds to the result -- mis * bsdf->eval * transmitter_val with the BSDF's pdf in the MIS weight -- together
with the receiver side of plate_path.py; the BSDF-sampled continuation ray (visible-normal sampling, pinned by the reference's own
tables, test_microfacet.py) is not followed, so phase and depth are not compared for this variant.
"""
import math

import numpy as np

from plate_path import F, PCG32, cosine_hemisphere, tea64


def microfacet_d(m, alpha, ggx=False):
    """MicrofacetDistribution::eval (microfacet.h:184-202), isotropic."""
    c = m[2]
    c2 = c * c
    if ggx:
        r = 1.0 / (math.pi * alpha * alpha * ((m[0] / alpha) ** 2 + (m[1] / alpha) ** 2 + m[2] ** 2) ** 2)
    else:
        r = math.exp(-((m[0] / alpha) ** 2 + (m[1] / alpha) ** 2) / c2) / (math.pi * alpha * alpha * c2 * c2)
    return r if r * c > 1e-20 else 0.0


def smith_g1(v, m, alpha, ggx=False):
    """microfacet.h:331-355: the rational Beckmann fit, or the closed GGX form."""
    xy = (alpha * v[0]) ** 2 + (alpha * v[1]) ** 2
    if float(v @ m) * v[2] <= 0.0:
        return 0.0
    if xy == 0.0:
        return 1.0
    t2 = xy / (v[2] * v[2])
    if ggx:
        return 2.0 / (1.0 + math.sqrt(1.0 + t2))
    a = 1.0 / math.sqrt(t2)
    return 1.0 if a >= 1.6 else (3.535 * a + 2.181 * a * a) / (1.0 + 2.276 * a + 2.577 * a * a)


def fresnel_conductor(cos_i, eta, k):
    c2 = cos_i * cos_i
    s2 = 1.0 - c2
    s4 = s2 * s2
    t1 = eta * eta - k * k - s2
    a2pb2 = math.sqrt(max(t1 * t1 + 4.0 * k * k * eta * eta, 0.0))
    a = math.sqrt(max(0.5 * (a2pb2 + t1), 0.0))
    term1, term2 = a2pb2 + c2, 2.0 * cos_i * a
    rs = (term1 - term2) / (term1 + term2)
    term3, term4 = a2pb2 * c2 + s4, term2 * s2
    rp = rs * (term3 - term4) / (term3 + term4)
    return 0.5 * (rs + rp)


def evaluate_sample(sc, seed, idx, alpha, eta, k, ggx=False):
    """-> dict(first_t, valid, tf_t, tf_f, value) of sample `idx` on a plate of roughconductor(alpha, eta + ik)."""
    rng = PCG32(tea64(seed, idx), tea64(idx, seed))
    pos = (rng.next_float(), rng.next_float())
    aperture = (rng.next_float(), rng.next_float())
    time = F(F(sc.adc_start) + F(rng.next_float()) * F(sc.adc_time))
    rng.next_float()
    lam = sc.wavelength_nm(sc.delta_frequency(time))
    a = sc.a
    local = (2.0 * pos[0] - 1.0, 2.0 * pos[1] - 1.0)
    o = np.array([a * local[0], a * local[1], 0.0])
    area = (2.0 * a) * (2.0 * a)
    d = cosine_hemisphere(*aperture)
    weight = sc.wigner_gain(local, d, lam, 2.0 * a) * (1.0 / area) * (area * math.pi)
    f_rx = sc.c32 * (1.0 / (float(lam) * 1e-9))
    out = dict(first_t=math.inf, valid=0, value=0.0, tf_t=float(time - F(sc.adc_start)) * sc.t_bins / sc.t_bw, tf_f=0.0)
    if d[2] <= 0.0:
        return out
    t1 = sc.R / d[2]
    p1 = o + t1 * d
    if abs(p1[0]) > sc.plate_half or abs(p1[1]) > sc.plate_half:
        return out
    out["first_t"], out["valid"] = t1, 1
    ray_time = F(time + F(-t1) / F(sc.c))
    s2 = (rng.next_float(), rng.next_float())
    tloc = (2.0 * s2[0] - 1.0, 2.0 * s2[1] - 1.0)
    p_tx = np.array([a * tloc[0], sc.tx_y + a * tloc[1], 0.0])
    dv = p_tx - p1
    dist2 = float(dv @ dv); dist = math.sqrt(dist2); dv = dv / dist
    dp = abs(dv[2])
    pdf = (1.0 / area) * (dist2 / dp if dp != 0.0 else 0.0)
    active = dv[2] < 0.0 and pdf != 0.0
    geom = 1.0 / pdf if pdf != 0.0 else 0.0
    ds_time = F(ray_time + F(-dist) / F(sc.c))
    lam2 = sc.wavelength_nm(sc.delta_frequency(ds_time))
    gain_tx = sc.wigner_gain(tloc, -dv, lam2, 2.0 * a)
    tx_val = (geom * gain_tx * ((1.0 / area) * 2.0 * math.pi)) if active else 0.0
    pdf *= gain_tx
    # ---- RoughConductor::eval / pdf in the plate's shading frame (s, t, n) = ((1, 0, 0), (0, -1, 0), (0, 0, -1)) ----
    wi = np.array([-d[0], d[1], d[2]])
    wo = np.array([dv[0], -dv[1], -dv[2]])
    res = 0.0
    if active and pdf != 0.0 and wi[2] > 0.0 and wo[2] > 0.0:
        H = (wo + wi) / np.linalg.norm(wo + wi)
        D = microfacet_d(H, alpha, ggx)
        if D != 0.0:
            G = smith_g1(wi, H, alpha, ggx) * smith_g1(wo, H, alpha, ggx)
            bsdf_val = fresnel_conductor(float(wi @ H), eta, k) * (D * G / (4.0 * wi[2]))
            bsdf_pdf = D * smith_g1(wi, H, alpha, ggx) / (4.0 * wi[2]) if (float(wi @ H) > 0.0 and float(wo @ H) > 0.0) else 0.0
            mis = (pdf * pdf) / (pdf * pdf + bsdf_pdf * bsdf_pdf)
            res = mis * bsdf_val * tx_val
    out["tf_f"] = abs(sc.c32 * (1.0 / (float(lam2) * 1e-9)) - float(F(f_rx))) * sc.f_bins / sc.f_bw
    out["value"] = abs(weight) * res
    return out
