"""Loader for the reference's XML scene format (src/libcore/xml.cpp) into a FlatScene, so that the same scene
files (python_scripts/trans_rad.xml, trans_image_rad.xml, ...) drive the B200 path unchanged.

Supported subset = what the radar path needs: <default>/$param substitution
(xml.cpp:132-150) with overrides (the CLI's -Dkey=value), <integrator> nesting (time(pathtime),
range(pathlength), phase(pathtimefrequency)), <shape type=rectangle|ply|obj> with <transform name="to_world">
(scale, translate, rotate, lookat, matrix; later entries multiply from the left, xml.cpp:918), nested or
referenced <bsdf> (diffuse, twosided, roughconductor, conductor), <emitter|transmitter>, <sensor|receiver> with
<film|adc> and <sampler>, <spectrum> / <float> reflectance values, <ref id>, <include filename>, <alias id as>, <shape type=serialized>.  Anything else raises
NotImplementedError, as the reference's plugin loader does for an unknown plugin.
"""
import os
import re
import xml.etree.ElementTree as ET

import numpy as np

from . import _native as N
from . import meshes
from .meshes import Transform as T
from .scene import FlatScene, BeifongError, ULTRASOUND_DEFAULTS

CIE_Y_NORMALIZATION = 1.0 / 106.7502593994140625      # include/mitsuba/core/spectrum.h:176

SIGNAL_TYPES = {"cw": N.BF_SIG_CW, "linfmcw": N.BF_SIG_LINFMCW, "pulse": N.BF_SIG_PULSE}
RECEIVE_TYPES = {"raw": N.BF_RECV_RAW, "raw_resample": N.BF_RECV_RAW_RESAMPLE, "mix_resample": N.BF_RECV_MIX_RESAMPLE,
                 "mixer": N.BF_RECV_MIXER}


class LoadedScene:
    """What load_file returns: the FlatScene plus the sampler parameters the sample loop reads."""

    def __init__(self, scene, sample_count, seed, shape_ids):
        self.scene, self.sample_count, self.seed, self.shape_ids = scene, sample_count, seed, shape_ids


def _floats(text):
    return [float(x) for x in re.split(r"[\s,]+", text.strip()) if x]


def _substitute(node, params):
    """$name substitution in every attribute (xml.cpp:132-150); an unknown parameter is an error."""
    for el in node.iter():
        for k, v in list(el.attrib.items()):
            if "$" in v:
                def rep(m):
                    if m.group(1) not in params:
                        raise BeifongError(f"xml: undefined parameter ${m.group(1)}")
                    return str(params[m.group(1)])
                el.set(k, re.sub(r"\$(\w+)", rep, v))


def _transform(el):
    m = T()
    for op in el:
        g = lambda k, d=0.0: float(op.get(k, d))
        if op.tag == "scale":
            if "value" in op.attrib:
                v = _floats(op.get("value")); s = v * 3 if len(v) == 1 else v
            else:
                s = [g("x", 1.0), g("y", 1.0), g("z", 1.0)]
            t = T.scale(s)
        elif op.tag == "translate":
            t = T.translate(_floats(op.get("value")) if "value" in op.attrib else [g("x"), g("y"), g("z")])
        elif op.tag == "rotate":
            axis = _floats(op.get("value")) if "value" in op.attrib else [g("x"), g("y"), g("z")]
            t = T.rotate(axis, g("angle"))
        elif op.tag == "lookat":
            o, tg = np.array(_floats(op.get("origin"))), np.array(_floats(op.get("target")))
            up = np.array(_floats(op.get("up"))) if op.get("up") else np.zeros(3)
            if not np.any(up):                              # xml.cpp:911-913: first vector of coordinate_system(dir)
                n = (tg - o) / np.linalg.norm(tg - o)
                sign = 1.0 if n[2] >= 0 else -1.0
                a = -1.0 / (sign + n[2]); b = n[0] * n[1] * a
                up = np.array([1.0 + sign * n[0] * n[0] * a, sign * b, -sign * n[0]])
            t = T.look_at(o, tg, up)
        elif op.tag == "matrix":
            v = _floats(op.get("value"))
            if len(v) == 9:
                mm = np.eye(4); mm[:3, :3] = np.array(v).reshape(3, 3)
            else:
                mm = np.array(v, dtype=np.float64).reshape(4, 4)
            t = T(mm)
        else:
            raise NotImplementedError(f"xml: transform operation <{op.tag}>")
        m = t * m
    return m


class _Loader:
    def __init__(self, root, base_dir, spectrum_width, consts):
        self.root, self.base_dir, self.width = root, base_dir, spectrum_width
        self.s = FlatScene(spectrum_width=spectrum_width, **consts)
        self.named_bsdf_nodes = {el.get("id"): el for el in root if el.tag == "bsdf" and el.get("id")}
        self.bsdf_ids = {}
        self.default_bsdf = None
        self.shape_ids = {}
        self.sample_count, self.seed = 4, 0               # sampler.cpp:11-19
        self.have_rx = False
        self.film = None

    # ---- properties ----------------------------------------------------------------------------------
    @staticmethod
    def props(el):
        p = {}
        for c in el:
            if c.tag in ("float", "integer", "boolean", "string") and c.get("name"):
                v = c.get("value")
                p[c.get("name")] = (float(v) if c.tag == "float" else int(v) if c.tag == "integer" else
                                    (v.strip().lower() == "true") if c.tag == "boolean" else v)
            elif c.tag in ("vector", "point") and c.get("name"):
                p[c.get("name")] = _floats(c.get("value")) if "value" in c.attrib else [float(c.get(k, 0)) for k in "xyz"]
        return p

    def spectrum(self, el, name, default, within_emitter):
        """create_texture_from_spectrum (xml.cpp:1087-1170)."""
        node = next((c for c in el if c.tag in ("spectrum", "float", "rgb") and c.get("name") == name), None)
        if node is None:
            if default is None:
                return None
            return self.s.add_uniform(default)
        if node.tag == "rgb":
            raise NotImplementedError("xml: <rgb> values (sRGB upsampling is image appearance, outside the radar path)")
        text = node.get("value")
        if ":" not in text:
            value = float(text)
            if node.tag == "spectrum" and within_emitter and self.width == 4:
                # d65 * scale: zero outside 360..830 nm, i.e. at every radar / ultrasound wavelength
                return self.s.add_spectrum(N.BF_SPEC_ZERO_OUTSIDE_VISIBLE, [value])
            return self.s.add_uniform(value)
        pairs = [p.split(":") for p in re.split(r"[\s,]+", text.strip()) if p]
        lam = np.array([float(a) for a, _ in pairs], np.float32)
        val = np.array([float(b) for _, b in pairs], np.float32)
        if np.any(np.diff(lam) < 0):
            raise BeifongError("Wavelengths must be specified in increasing order!")
        if within_emitter or self.width != 4:
            val = val * np.float32(CIE_Y_NORMALIZATION)
        if self.width == 4:
            d = np.diff(lam)
            regular = len(lam) < 3 or bool(np.all(np.abs(d[1:] - d[0]) <= np.finfo(np.float32).eps / 2))
            return self.s.add_spectrum(N.BF_SPEC_REGULAR if regular else N.BF_SPEC_IRREGULAR, val, lam)
        # non-spectral: pre-integrated against the CIE curves, which vanish outside the visible range
        if lam.min() > 830.0 or lam.max() < 360.0:
            return self.s.add_uniform(0.0)
        raise NotImplementedError("xml: tabulated visible-range spectra in mono mode need the CIE tables (image appearance)")

    # ---- BSDFs ------------------------------------------------------------------------------------------
    def bsdf(self, el):
        if el.tag == "ref":
            rid = el.get("id")
            if rid not in self.named_bsdf_nodes:
                raise BeifongError(f"xml: reference to unknown id \"{rid}\"")
            if rid not in self.bsdf_ids:
                self.bsdf_ids[rid] = self.bsdf(self.named_bsdf_nodes[rid])
            return self.bsdf_ids[rid]
        kind = el.get("type")
        two_sided = False
        if kind == "twosided":
            inner = [c for c in el if c.tag in ("bsdf", "ref")]
            if len(inner) != 1:
                raise NotImplementedError("xml: twosided with two different nested BSDFs")
            el = inner[0] if inner[0].tag == "bsdf" else self.named_bsdf_nodes[inner[0].get("id")]
            kind, two_sided = el.get("type"), True
        p = self.props(el)
        if kind == "diffuse":
            return self.s.add_bsdf(N.BF_BSDF_DIFFUSE, two_sided=two_sided, reflectance=self.spectrum(el, "reflectance", 0.5, False))
        if kind in ("roughconductor", "conductor"):
            if "material" in p and p["material"] != "none":
                raise NotImplementedError("xml: named conductor materials need the visible-range IOR tables")
            eta, k = self.spectrum(el, "eta", 0.0, False), self.spectrum(el, "k", 1.0, False)
            spec = self.spectrum(el, "specular_reflectance", None, False)
            dist = {"beckmann": N.BF_MICROFACET_BECKMANN, "ggx": N.BF_MICROFACET_GGX}[p.get("distribution", "beckmann")]
            alpha = p.get("alpha", 0.1)
            return self.s.add_bsdf(N.BF_BSDF_ROUGHCONDUCTOR if kind == "roughconductor" else N.BF_BSDF_CONDUCTOR, two_sided=two_sided,
                                   eta=eta, k=k, specular_reflectance=N.BF_NONE if spec is None else spec, distribution=dist,
                                   sample_visible=p.get("sample_visible", True), alpha_u=p.get("alpha_u", alpha), alpha_v=p.get("alpha_v", alpha))
        raise NotImplementedError(f"xml: bsdf plugin \"{kind}\" is outside the radar path")

    # ---- endpoints -----------------------------------------------------------------------------------------
    def signal(self, p):
        """Signal block of wigner/phased endpoints (wignertransmitter.cpp:69-102, phasedreceiver.cpp:67-113)."""
        st = p.get("signaltype", "cw")
        g = p.get
        if st == "linfmcw":
            return FlatScene.signal(N.BF_SIG_LINFMCW, g("sig_is_delta", True), g("resample_freq", False), g("amplitude", 1.0), g("crf", 1.0),
                                    g("chirp_len", 1.0), g("freq_centre", 1.0), g("freq_sweep", 1.0), g("phase", 0.0), g("gain", 1.0))
        if st == "pulse":
            return FlatScene.signal(N.BF_SIG_PULSE, g("sig_is_delta", False), g("resample_freq", False), g("amplitude", 1.0), g("prf", 1.0),
                                    g("pulse_len", 1.0), g("freq_centre", 1.0), g("freq_ext", 1.0), g("phase", 0.0), g("gain", 1.0))
        if st == "cw":
            return FlatScene.signal(N.BF_SIG_CW, g("sig_is_delta", True), g("resample_freq", False), g("amplitude", 1.0), 0.0, 0.0,
                                    g("freq_centre", 1.0), g("freq_ext", 0.0), g("phase", 0.0), g("gain", 1.0))
        t_ext = g("pulse_len", 1.0)
        return FlatScene.signal(N.BF_SIG_OTHER, g("sig_is_delta", True), g("resample_freq", False), g("amplitude", 1.0), g("prf", 1.0),
                                t_ext, g("freq_centre", 1.0), 1.0 / t_ext, g("phase", 0.0), g("gain", 1.0))

    def rx_signal(self, p, receive_type):
        """Receivers read their signal block by receive_type (wignerreceiver.cpp:61-110 == phasedreceiver.cpp:67-113):
        raw / raw_resample only have freq_centre, freq_ext and gain; m_sig_is_delta is left uninitialised there and is
        defined here as false (a raw receiver samples the band, so `extents` carries the bandwidth)."""
        if receive_type in ("raw", "raw_resample"):
            g = p.get
            return FlatScene.signal(N.BF_SIG_UNSET, False, False, 1.0, 1.0, 1.0, g("freq_centre", 1.0), g("freq_ext", 1.0), 0.0, g("gain", 1.0))
        return self.signal(p)

    def array(self, el, p):
        """Phased-array parameters (phasedtransmitter.cpp:104-117)."""
        loc = next((c for c in el if c.tag == "transform" and c.get("name") == "array_loc"), None)
        z = [0.0, 0.0, 0.0]
        return FlatScene.array(int(p.get("n_elems", 1)), p.get("steering_vector", z), (_transform(loc) if loc is not None else T()).matrix32(),
                               p.get("elem_dims", z), p.get("elem_spacing", z), p.get("elem_axis", z))

    def antenna_texture(self, el, name, within_emitter=True):
        tex = self.spectrum(el, name, None, within_emitter)
        if tex is None:                                     # Texture::D65(1.f): d65 in spectral mode, uniform 1 otherwise
            tex = self.s.add_spectrum(N.BF_SPEC_ZERO_OUTSIDE_VISIBLE, [1.0]) if self.width == 4 else self.s.add_uniform(1.0)
        return tex

    def transmitter(self, el, shape_id):
        kind, p = el.get("type"), self.props(el)
        if kind == "spot":                                  # src/emitters/spot.cpp:72-95
            tw = next((c for c in el if c.tag == "transform" and c.get("name") == "to_world"), None)
            cutoff = p.get("cutoff_angle", 20.0)
            d = FlatScene.endpoint(N.BF_TX_SPOT, texture=self.antenna_texture(el, "intensity"), cutoff_angle_deg=cutoff,
                                   beam_width_deg=p.get("beam_width", cutoff * 3.0 / 4.0),
                                   to_world=(_transform(tw) if tw is not None else T()).matrix32())
            return self.s.add_transmitter(None, d)
        if shape_id is None:
            raise BeifongError(f"xml: emitter \"{kind}\" must be nested in a shape")
        if kind == "area":
            d = FlatScene.endpoint(N.BF_TX_AREALIGHT, texture=self.antenna_texture(el, "radiance"))
        elif kind == "areatransmitter":
            d = FlatScene.endpoint(N.BF_TX_AREA, texture=self.antenna_texture(el, "radiance"))
        elif kind == "wignertransmitter":
            name = "antenna_texture" if any(c.get("name") == "antenna_texture" for c in el) else "radiance"
            d = FlatScene.endpoint(N.BF_TX_WIGNER, texture=self.antenna_texture(el, name), signal=self.signal(p))
        elif kind == "phasedtransmitter":
            if any(c.tag == "transform" and c.get("name") == "to_world" for c in el):   # phasedtransmitter.cpp:54-57
                raise BeifongError("Found a 'to_world' transformation -- this is not allowed. The area light inherits this "
                                   "transformation from its parent shape.")
            d = FlatScene.endpoint(N.BF_TX_PHASED, texture=self.antenna_texture(el, "antenna_texture"), signal=self.signal(p),
                                   array=self.array(el, p))
        else:
            raise NotImplementedError(f"xml: emitter / transmitter plugin \"{kind}\" is not on the accelerated path")
        return self.s.add_transmitter(shape_id, d)

    def receiver(self, el, shape_id):
        if self.have_rx:
            raise NotImplementedError("xml: one sensor / receiver per load (pass receiver_index to pick another)")
        kind, p = el.get("type"), self.props(el)
        film = next((c for c in el if c.tag in ("film", "adc")), None)
        fp = self.props(film) if film is not None else {}
        filt = dict(filter_radius=0.5, filter_kind=0)
        if film is not None:
            rf = next((c for c in film if c.tag == "rfilter"), None)
            # Film / ADC default reconstruction filter is gaussian (adc.cpp:60-76).  ADCs: box, gaussian (radius = 4 stddev,
            # gaussian.cpp:37-45), tent (radius 1, tent.cpp:30-34), mitchell, catmullrom, lanczos; films (gen 1/2): box only
            rtype = rf.get("type") if rf is not None else "gaussian"
            if rtype == "gaussian" and film.tag == "adc":
                filt = dict(filter_radius=4.0 * float(self.props(rf).get("stddev", 0.5)) if rf is not None else 2.0, filter_kind=1)
            elif rtype == "tent" and film.tag == "adc":
                filt = dict(filter_radius=1.0, filter_kind=2)
            elif rtype in ("mitchell", "catmullrom") and film.tag == "adc":
                if rtype == "mitchell" and any(k in self.props(rf) for k in ("B", "C")):
                    raise NotImplementedError("xml: mitchell filter with non-default B / C")
                filt = dict(filter_radius=2.0, filter_kind=3 if rtype == "mitchell" else 4)
            elif rtype == "lanczos" and film.tag == "adc":
                filt = dict(filter_radius=float(int(self.props(rf).get("lobes", 3))), filter_kind=5)
            elif rtype != "box":
                raise NotImplementedError(f"xml: reconstruction filter \"{rtype}\" is not on the accelerated path here")
        sampler = next((c for c in el if c.tag == "sampler"), None)
        if sampler is not None:
            sp = self.props(sampler)
            self.sample_count, self.seed = int(sp.get("sample_count", 4)), int(sp.get("seed", 0))
        common = dict(shutter_open=p.get("shutter_open", 0.0), shutter_open_time=p.get("shutter_close", 0.0) - p.get("shutter_open", 0.0))
        if kind in ("fluxmeter", "perspective"):
            w, h = int(fp.get("width", 768)), int(fp.get("height", 576))
            if kind == "fluxmeter":
                d = FlatScene.endpoint(N.BF_RX_FLUXMETER, film_width=w, film_height=h, **common)
            else:
                if p.get("fov_axis", "x") != "x":
                    raise NotImplementedError("xml: perspective fov_axis other than x")
                tw = next((c for c in el if c.tag == "transform" and c.get("name") == "to_world"), None)
                near, far, fov = p.get("near_clip", 1e-2), p.get("far_clip", 1e4), p.get("fov", 0.0)
                d = FlatScene.endpoint(N.BF_RX_PERSPECTIVE, film_width=w, film_height=h, near_clip=near, far_clip=far,
                                       to_world=(_transform(tw) if tw is not None else T()).matrix32(),
                                       sample_to_camera=meshes.perspective_sample_to_camera(fov, near, far, w, h), **common)
                shape_id = None
            self.film = ("film", w, h)
        elif kind in ("omnidirectional", "wignerreceiver", "phasedreceiver"):
            start = p.get("adc_sampling_start", 0.0)
            if p.get("adc_sampling_end", 0.0) - start < 0:
                raise BeifongError("ADC sampling time must be less than or equal to the adc sampling end time!")
            rk = {"omnidirectional": N.BF_RX_OMNIDIRECTIONAL, "wignerreceiver": N.BF_RX_WIGNER, "phasedreceiver": N.BF_RX_PHASED}[kind]
            if p.get("receive_type", "raw") not in RECEIVE_TYPES:
                raise BeifongError(f"xml: unknown receive_type \"{p.get('receive_type')}\"")
            d = FlatScene.endpoint(rk, texture=self.antenna_texture(el, "antenna_texture", False),
                                   signal=self.rx_signal(p, p.get("receive_type", "raw")),
                                   receive_type=RECEIVE_TYPES[p.get("receive_type", "raw")], adc_sampling_start=start,
                                   adc_sampling_time=p.get("adc_sampling_end", 0.0) - start,
                                   array=self.array(el, p) if kind == "phasedreceiver" else None)
            t_bins, f_bins = int(fp.get("t_bins", 1024)), int(fp.get("f_bins", 1024))   # adc.cpp:18-21
            self.s.set_adc(t_bins=t_bins, f_bins=f_bins, t_bandwidth=fp.get("t_bandwidth", 3.81e-6), f_bandwidth=fp.get("f_bandwidth", 250e6),
                           window_t_bins=int(fp.get("window_t_bins", t_bins)), window_f_bins=int(fp.get("window_f_bins", f_bins)),
                           window_offset_t=int(fp.get("window_offset_t", 0)), window_offset_f=int(fp.get("window_offset_f", 0)), **filt)
            self.film = ("adc", t_bins, f_bins)
        else:
            raise NotImplementedError(f"xml: sensor / receiver plugin \"{kind}\" is not on the accelerated path")
        self.s.set_receiver(shape_id, d)
        self.have_rx = True

    # ---- shapes ------------------------------------------------------------------------------------------
    def shape(self, el):
        kind, p = el.get("type"), self.props(el)
        tw = next((c for c in el if c.tag == "transform" and c.get("name") == "to_world"), None)
        m = _transform(tw) if tw is not None else T()
        b = next((c for c in el if c.tag == "bsdf" or (c.tag == "ref" and c.get("id") in self.named_bsdf_nodes)), None)
        if b is not None:
            bsdf = self.bsdf(b)
        else:
            if self.default_bsdf is None:                   # Shape's default BSDF: diffuse (shape.cpp:52-60)
                self.default_bsdf = self.s.add_diffuse(0.5, two_sided=False)
            bsdf = self.default_bsdf
        if kind == "rectangle":
            if p.get("flip_normals", False):                # rectangle.cpp:76-81
                m = m * T.scale([1.0, 1.0, -1.0])
            sid = self.s.add_rect(m.matrix32(), bsdf)
        elif kind in ("ply", "obj", "serialized"):
            path = p.get("filename")
            if path is None:
                raise BeifongError(f"xml: <shape type=\"{kind}\"> without a filename")
            if not os.path.isabs(path):
                path = os.path.join(self.base_dir, path)
            if not os.path.exists(path):
                raise BeifongError(f"\"{path}\": file does not exist!")       # ply.cpp:120-122
            if kind == "serialized":
                try:
                    pos, faces, nrm = meshes.load_serialized(path, to_world=m.matrix32(), shape_index=int(p.get("shape_index", 0)))
                except RuntimeError as e:
                    raise BeifongError(str(e))
            else:
                pos, faces, nrm = (meshes.load_ply if kind == "ply" else meshes.load_obj)(path, to_world=m.matrix32())
            if p.get("face_normals", False):
                nrm = None
            elif nrm is None:
                nrm = meshes.vertex_normals(pos, faces)     # mesh.cpp:201-278, run by the loaders when the file has none
            sid = self.s.add_mesh(pos, faces, bsdf, normals=nrm)
        else:
            raise NotImplementedError(f"xml: shape plugin \"{kind}\" is outside the radar path")
        if el.get("id"):
            self.shape_ids[el.get("id")] = sid
        for c in el:
            if c.tag in ("emitter", "transmitter"):
                self.transmitter(c, sid)
            elif c.tag in ("sensor", "receiver"):
                self.receiver(c, sid)
        return sid

    def integrator(self, el):
        kind, p = el.get("type"), self.props(el)
        inner = next((c for c in el if c.tag == "integrator"), None)
        ip = self.props(inner) if inner is not None else {}
        depth = dict(max_depth=int(ip.get("max_depth", p.get("max_depth", -1))), rr_depth=int(ip.get("rr_depth", p.get("rr_depth", 5))))
        nested = inner.get("type") if inner is not None else None
        if kind == "time" and nested == "pathtime":
            self.s.set_integrator(N.BF_GEN_TIME, **depth)
        elif kind == "range" and nested == "pathlength":
            self.s.set_integrator(N.BF_GEN_RANGE, range_bins=int(p["bins"]), dr=float(p["dr"]), **depth)
        elif kind == "phase" and nested == "pathtimefrequency":
            self.s.set_integrator(N.BF_GEN_PHASE, phase_bins=int(p.get("bins", 1)), **depth)
        elif kind == "pathtimefrequency":
            self.s.set_integrator(N.BF_GEN_PHASE, phase_bins=0, max_depth=int(p.get("max_depth", -1)), rr_depth=int(p.get("rr_depth", 5)))
        else:
            raise NotImplementedError(f"xml: integrator \"{kind}\"" + (f"({nested})" if nested else "") + " is not a radar integrator")


XML_INCLUDE_MAX_RECURSION = 15          # include/mitsuba/core/xml.h:8


def _expand_includes(node, base_dir, depth=0):
    """<include filename="..."/> (xml.cpp:653-705): the children of an included <scene> (or the included root element
    itself) take the place of the tag; nested up to MTS_XML_INCLUDE_MAX_RECURSION levels."""
    out = []
    for child in list(node):
        if child.tag != "include":
            _expand_includes(child, base_dir, depth)
            out.append(child)
            continue
        if set(child.attrib) != {"filename"}:
            raise BeifongError("xml: <include> takes exactly the attribute \"filename\"")
        path = child.get("filename")
        if not os.path.isabs(path):
            path = os.path.join(base_dir, path)
        if not os.path.exists(path):
            raise BeifongError(f"included file \"{path}\" not found")
        if depth + 1 > XML_INCLUDE_MAX_RECURSION:
            raise BeifongError(f"Exceeded <include> recursion limit of {XML_INCLUDE_MAX_RECURSION}")
        try:
            inc = ET.parse(path).getroot()
        except ET.ParseError as e:
            raise BeifongError(f"error while loading \"{path}\": {e}")
        _expand_includes(inc, os.path.dirname(path), depth + 1)
        out.extend(list(inc) if inc.tag == "scene" else [inc])
    node[:] = out


def load_string(text, base_dir=".", spectrum_width=1, consts=None, **params):
    """Parse a scene; `params` override <default> values like the CLI's -Dname=value (mitsuba.cpp:176-177)."""
    root = ET.fromstring(text)
    if root.tag != "scene":
        raise BeifongError("xml: root element must be <scene>")
    _expand_includes(root, base_dir)
    defaults = {el.get("name"): el.get("value") for el in root if el.tag == "default"}
    defaults.update({k: v for k, v in params.items()})
    _substitute(root, defaults)
    ld = _Loader(root, base_dir, spectrum_width, dict(ULTRASOUND_DEFAULTS) if consts is None else consts)
    integ = None
    # <alias id as> (xml.cpp:594-613): a second id for a named object; BSDF aliases are resolved before the shapes that use them
    for el in root:
        if el.tag == "alias" and el.get("id") in ld.named_bsdf_nodes:
            if el.get("as") in ld.named_bsdf_nodes:
                raise BeifongError(f"\"alias\" has duplicate id \"{el.get('as')}\"")
            ld.named_bsdf_nodes[el.get("as")] = ld.named_bsdf_nodes[el.get("id")]
    for el in root:
        if el.tag == "shape":
            ld.shape(el)
        elif el.tag in ("emitter", "transmitter"):
            ld.transmitter(el, None)
        elif el.tag in ("sensor", "receiver"):
            ld.receiver(el, None)
        elif el.tag == "integrator":
            integ = el
        elif el.tag in ("default", "bsdf"):
            pass
        elif el.tag == "alias":                             # shapes: the alias follows the shape it names
            src, dst = el.get("id"), el.get("as")
            if src in ld.named_bsdf_nodes:
                continue
            if dst in ld.shape_ids:
                raise BeifongError(f"\"alias\" has duplicate id \"{dst}\"")
            if src in ld.shape_ids:
                ld.shape_ids[dst] = ld.shape_ids[src]
            else:
                raise BeifongError(f"referenced id \"{src}\" not found")
        else:
            raise NotImplementedError(f"xml: top-level element <{el.tag}>")
    if integ is None:
        raise BeifongError("xml: scene without an integrator")
    ld.integrator(integ)
    return LoadedScene(ld.s, ld.sample_count, ld.seed, ld.shape_ids)


def load_file(path, spectrum_width=1, consts=None, **params):
    with open(path, "r") as f:
        return load_string(f.read(), os.path.dirname(os.path.abspath(path)), spectrum_width, consts, **params)
