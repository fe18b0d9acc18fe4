"""FlatScene: object wrapper over the C ABI handle (include/beifong_b200/abi.h).

Scene construction is host-only; `commit()` builds the BVH and uploads to the GPU and is the first
call that needs a CUDA device.  Errors of the native layer are raised as the exception types the
reference raises for the same condition: `RuntimeError` for `Throw(...)`, `NotImplementedError` for
unsupported plugins (include/mitsuba/core/logger.h:134-138).
"""
import ctypes as C
import os

import numpy as np

from . import _native as N


class BeifongError(RuntimeError):
    pass


def _raise(handle, code, what):
    msg = N.lib().bf_last_error(handle)
    msg = msg.decode() if msg else ""
    text = f"{what}: {msg} (code {code})"
    if code == N.BF_ERR_UNSUPPORTED:
        raise NotImplementedError(text)
    raise BeifongError(text)


class PackedBatch:
    """A marshalled bf_batch_item array (FlatScene.pack_batch) and the buffers it points into."""
    def __init__(self, n, array, keep):
        self.n, self.array, self.keep = n, array, keep

    def __len__(self):
        return self.n


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _up(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint32))


# Reference defaults of the compile-time physics macros (include/mitsuba/core/spectrum.h:15-40):
# 40 kHz ultrasound in air.  RADAR_94GHZ is the commented-out alternative (spectrum.h:17,25).
ULTRASOUND_DEFAULTS = dict(c=340.0, lambda_min_nm=7555556.0, lambda_max_nm=9714286.0, cie_min_nm=360.0, cie_max_nm=830.0, P=0.0)
RADAR_94GHZ = dict(c=299792458.0, lambda_min_nm=299792458.0 / 97e9 * 1e9, lambda_max_nm=299792458.0 / 91e9 * 1e9,
                   cie_min_nm=360.0, cie_max_nm=830.0, P=0.0)


class FlatScene:
    def __init__(self, spectrum_width=1, **consts):
        c = dict(ULTRASOUND_DEFAULTS)
        c.update(consts)
        self.consts = N.bf_constants(c["c"], c["lambda_min_nm"], c["lambda_max_nm"], c["cie_min_nm"], c["cie_max_nm"], c["P"],
                                     spectrum_width, 0)
        self._h = C.c_void_p()
        rc = N.lib().bf_scene_create(C.byref(self.consts), C.byref(self._h))
        if rc:
            _raise(None, rc, "bf_scene_create")
        self.committed = False

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            N.lib().bf_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc, what):
        if rc:
            _raise(self._h, rc, what)

    # ---- construction -------------------------------------------------------------------------------
    def add_spectrum(self, kind, values, wavelengths=None):
        v = _f32(np.atleast_1d(values))
        lam = _f32(np.atleast_1d(wavelengths)) if wavelengths is not None else None
        out = C.c_uint32()
        self._ck(N.lib().bf_scene_add_spectrum(self._h, kind, len(v), _fp(lam) if lam is not None else None, _fp(v), C.byref(out)),
                 "bf_scene_add_spectrum")
        return out.value

    def add_uniform(self, value):
        return self.add_spectrum(N.BF_SPEC_UNIFORM, [value])

    def add_bsdf(self, kind, two_sided=True, reflectance=N.BF_NONE, eta=N.BF_NONE, k=N.BF_NONE, specular_reflectance=N.BF_NONE,
                 distribution=N.BF_MICROFACET_BECKMANN, sample_visible=True, alpha_u=0.1, alpha_v=0.1, back_bsdf=N.BF_NONE):
        d = N.bf_bsdf_desc(kind, int(two_sided), back_bsdf, reflectance, eta, k, specular_reflectance, distribution,
                           int(sample_visible), alpha_u, alpha_v, 0)
        out = C.c_uint32()
        self._ck(N.lib().bf_scene_add_bsdf(self._h, C.byref(d), C.byref(out)), "bf_scene_add_bsdf")
        return out.value

    def add_diffuse(self, reflectance_value, two_sided=True):
        return self.add_bsdf(N.BF_BSDF_DIFFUSE, two_sided=two_sided, reflectance=self.add_uniform(reflectance_value))

    def add_mesh(self, positions, faces, bsdf, normals=None):
        p = _f32(positions).reshape(-1, 3)
        f = np.ascontiguousarray(faces, dtype=np.uint32).reshape(-1, 3)
        n = _f32(normals).reshape(-1, 3) if normals is not None else None
        out = C.c_uint32()
        self._ck(N.lib().bf_scene_add_mesh(self._h, _fp(p), len(p), _up(f), len(f), _fp(n) if n is not None else None, bsdf, C.byref(out)),
                 "bf_scene_add_mesh")
        return out.value

    def add_rect(self, to_world, bsdf):
        m = _f32(to_world).reshape(16)
        out = C.c_uint32()
        self._ck(N.lib().bf_scene_add_rect(self._h, _fp(m), bsdf, C.byref(out)), "bf_scene_add_rect")
        return out.value

    @staticmethod
    def endpoint(kind, texture=0, signal=None, receive_type=N.BF_RECV_RAW, adc_sampling_start=0.0, adc_sampling_time=0.0,
                 to_world=None, cutoff_angle_deg=20.0, beam_width_deg=None, near_clip=1e-2, far_clip=1e4, sample_to_camera=None,
                 shutter_open=0.0, shutter_open_time=0.0, film_width=1, film_height=1, array=None):
        d = N.bf_endpoint_desc()
        d.kind = kind
        d.texture = texture
        if array is not None:
            d.array = array
        if signal is not None:
            d.signal = signal
        d.receive_type = receive_type
        d.adc_sampling_start = adc_sampling_start
        d.adc_sampling_time = adc_sampling_time
        tw = np.eye(4, dtype=np.float32) if to_world is None else _f32(to_world).reshape(4, 4)
        d.to_world = (C.c_float * 16)(*tw.reshape(16))
        d.cutoff_angle_deg = cutoff_angle_deg
        d.beam_width_deg = cutoff_angle_deg * 3.0 / 4.0 if beam_width_deg is None else beam_width_deg  # spot.cpp:78
        d.near_clip, d.far_clip = near_clip, far_clip
        s2c = np.eye(4, dtype=np.float32) if sample_to_camera is None else _f32(sample_to_camera).reshape(4, 4)
        d.sample_to_camera = (C.c_float * 16)(*s2c.reshape(16))
        d.shutter_open, d.shutter_open_time = shutter_open, shutter_open_time
        d.film_width, d.film_height = film_width, film_height
        return d

    @staticmethod
    def array(n_elems=1, steering_vector=(0, 0, 0), array_loc=None, elem_dims=(0, 0, 0), elem_spacing=(0, 0, 0), elem_axis=(0, 0, 0)):
        """Phased-array block of phasedtransmitter / phasedreceiver (phasedtransmitter.cpp:104-117): same names, same
        defaults (zero vectors, identity array_loc)."""
        a = N.bf_array_desc()
        a.n_elems = int(n_elems)
        a.steering_vector = (C.c_float * 3)(*[float(x) for x in steering_vector])
        loc = np.eye(4, dtype=np.float32) if array_loc is None else _f32(array_loc).reshape(4, 4)
        a.array_loc = (C.c_float * 16)(*loc.reshape(16))
        a.elem_dims = (C.c_float * 3)(*[float(x) for x in elem_dims])
        a.elem_spacing = (C.c_float * 3)(*[float(x) for x in elem_spacing])
        a.elem_axis = (C.c_float * 3)(*[float(x) for x in elem_axis])
        return a

    @staticmethod
    def signal(signaltype=N.BF_SIG_LINFMCW, sig_is_delta=True, resample_freq=False, amplitude=1.0, repfreq=1.0, t_ext=1.0,
               f_centre=1.0, f_ext=1.0, phi0=0.0, gain=1.0):
        return N.bf_signal_desc(signaltype, int(sig_is_delta), int(resample_freq), amplitude, repfreq, t_ext, f_centre, f_ext, phi0, gain)

    def add_transmitter(self, shape, desc):
        out = C.c_uint32()
        self._ck(N.lib().bf_scene_add_transmitter(self._h, N.BF_NONE if shape is None else shape, C.byref(desc), C.byref(out)),
                 "bf_scene_add_transmitter")
        return out.value

    def set_receiver(self, shape, desc):
        self._ck(N.lib().bf_scene_set_receiver(self._h, N.BF_NONE if shape is None else shape, C.byref(desc)), "bf_scene_set_receiver")

    def set_adc(self, t_bins=1024, f_bins=1024, t_bandwidth=3.81e-6, f_bandwidth=250e6, window_t_bins=None, window_f_bins=None,
                window_offset_t=0, window_offset_f=0, filter_radius=0.5, filter_kind=0):
        d = N.bf_adc_desc(t_bins, f_bins, t_bins if window_t_bins is None else window_t_bins,
                          f_bins if window_f_bins is None else window_f_bins, window_offset_t, window_offset_f,
                          t_bandwidth, f_bandwidth, filter_radius, filter_kind)
        self._ck(N.lib().bf_scene_set_adc(self._h, C.byref(d)), "bf_scene_set_adc")

    def set_integrator(self, gen, max_depth=-1, rr_depth=5, phase_bins=1, range_bins=0, dr=0.0, doppler=False):
        """doppler=True restores the Doppler statement the reference has commented out (pathtimefrequency.cpp:252-256)."""
        d = N.bf_integrator_desc(gen, max_depth, rr_depth, phase_bins, range_bins, dr, 0, 1 if doppler else 0)
        self._ck(N.lib().bf_scene_set_integrator(self._h, C.byref(d)), "bf_scene_set_integrator")

    def update_transform(self, shape, motion):
        """bf_scene_update_transform: re-pose a rectangle (to_world) or move a mesh (motion applied to the vertices as added);
        after commit() a mesh is refitted on the device, without a rebuild."""
        m = _f32(motion).reshape(16)
        self._ck(N.lib().bf_scene_update_transform(self._h, shape, _fp(m)), "bf_scene_update_transform")

    def set_shape_velocity(self, shape, velocity):
        """The shape's `velocity` transform (shape.cpp:39-42), a 4 x 4 matrix; read by Shape::doppler only."""
        m = _f32(velocity).reshape(16)
        self._ck(N.lib().bf_scene_set_shape_velocity(self._h, shape, _fp(m)), "bf_scene_set_shape_velocity")

    def update_rect(self, shape, to_world):
        m = _f32(to_world).reshape(16)
        self._ck(N.lib().bf_scene_update_rect(self._h, shape, _fp(m)), "bf_scene_update_rect")

    # ---- flat export (fed to the CPU oracle by the tests) ----------------------------------------------
    def export(self):
        need = C.c_size_t()
        self._ck(N.lib().bf_scene_export(self._h, None, 0, C.byref(need)), "bf_scene_export")
        buf = np.zeros(need.value, dtype=np.uint8)
        self._ck(N.lib().bf_scene_export(self._h, buf.ctypes.data_as(C.c_void_p), need.value, C.byref(need)), "bf_scene_export")
        return buf

    # ---- device ----------------------------------------------------------------------------------------
    def commit(self, device=0):
        self._ck(N.lib().bf_scene_commit(self._h, device), "bf_scene_commit")
        self.committed = True
        return self

    def reupload(self):
        self._ck(N.lib().bf_scene_reupload(self._h), "bf_scene_reupload")

    def geometry_bytes(self):
        n = C.c_size_t()
        self._ck(N.lib().bf_scene_device_bytes(self._h, C.byref(n)), "bf_scene_device_bytes")
        return n.value

    def profile_shape(self):
        w, h, c = C.c_uint32(), C.c_uint32(), C.c_uint32()
        self._ck(N.lib().bf_profile_size(self._h, C.byref(w), C.byref(h), C.byref(c)), "bf_profile_size")
        return (h.value, w.value, c.value)   # [f][t][channel]

    def receive(self, n_samples, seed=0, first_sample=0, accumulate=False):
        self._ck(N.lib().bf_receive(self._h, seed, first_sample, n_samples, int(accumulate)), "bf_receive")

    @staticmethod
    def pack_batch(items):
        """The bf_batch_item array of `items` (see receive_batch), marshalled once: a caller that submits the same batch again
        (a benchmark loop, a sweep replayed with other seeds) passes the result to receive_batch instead of the list."""
        n = len(items)
        arr = (N.bf_batch_item * max(n, 1))()
        keep = []     # ctypes buffers must outlive the calls
        for i, it in enumerate(items):
            arr[i].seed = int(it.get("seed", 0)); arr[i].first_sample = int(it.get("first_sample", 0))
            rects = it.get("rects") or []
            arr[i].n_rect_updates = len(rects)
            if rects:
                ids = (N.u32 * len(rects))(*[int(r[0]) for r in rects])
                mats = np.ascontiguousarray(np.stack([_f32(r[1]).reshape(16) for r in rects]))
                keep += [ids, mats]
                arr[i].rect_shape_ids = ids
                arr[i].rect_to_world = mats.ctypes.data_as(C.POINTER(N.f32))
            rx = it.get("receiver")
            if rx is not None:
                arr[i].receiver_shape = N.BF_NONE if rx[0] is None else int(rx[0])
                keep.append(rx[1])
                arr[i].receiver = C.pointer(rx[1])
        return PackedBatch(n, arr, keep)

    def receive_batch(self, items, n_samples, read=True, out=None):
        """bf_receive_batch: `items` is a list of dicts with keys seed, first_sample (default 0), rects = [(shape_id, to_world 4x4), ...]
        (default none) and receiver = (shape_id_or_None, bf_endpoint_desc) (default: keep), or the result of pack_batch on such a
        list.  Returns the item profiles as an array [item][f][t][channel] (read=True, synchronises; into `out` when given) or
        None (read=False: asynchronous, the profiles stay on the device)."""
        packed = items if isinstance(items, PackedBatch) else self.pack_batch(items)
        n = packed.n
        if read and out is None:
            out = np.empty((n,) + tuple(self.profile_shape()), dtype=np.float32)
        self._ck(N.lib().bf_receive_batch(self._h, n, packed.array, int(n_samples), _fp(out) if read else None, out.size if read else 0), "bf_receive_batch")
        return out if read else None

    def synchronize(self):
        self._ck(N.lib().bf_synchronize(self._h), "bf_synchronize")

    def set_stream(self, cuda_stream):
        """bf_scene_set_stream: every entry point enqueues on `cuda_stream` (a cudaStream_t as an integer; None / 0: the library's own)."""
        self._ck(N.lib().bf_scene_set_stream(self._h, C.c_void_p(int(cuda_stream) if cuda_stream else None)), "bf_scene_set_stream")

    # ---- multi-GPU through the C ABI (NCCL bound inside the library, no torch on the data path) ----
    @staticmethod
    def multi_unique_id():
        """Rank 0: a fresh NCCL unique id (bytes) to ship to the other ranks by any means."""
        buf = C.create_string_buffer(128)
        rc = N.lib().bf_multi_unique_id(buf, 128)
        if rc != 0:
            err = N.lib().bf_last_error(None)
            raise RuntimeError(f"bf_multi_unique_id failed ({rc}): {err.decode() if err else ''}")
        return buf.raw

    def multi_init(self, rank, world, unique_id):
        buf = C.create_string_buffer(bytes(unique_id), 128)
        self._ck(N.lib().bf_multi_init(self._h, int(rank), int(world), buf, 128), "bf_multi_init")

    def multi_reduce(self, root=0):
        """One ncclReduce of the per-rank profiles onto `root`, in place, on the scene's stream."""
        self._ck(N.lib().bf_multi_reduce(self._h, int(root)), "bf_multi_reduce")

    def profile(self, out=None):
        shape = self.profile_shape()
        if out is None:
            out = np.empty(shape, dtype=np.float32)
        self._ck(N.lib().bf_profile_read(self._h, _fp(out), out.size), "bf_profile_read")
        return out

    def bitmap(self, raw=False, component_format="float16"):
        """HDRADC::bitmap (hdradc.cpp:219-275): raw=True is the un-normalised fp32 storage; raw=False divides every
        channel by the weight channel on the device and drops it ([f][t][channels - 1], float16 by default as in
        the reference's `component_format`)."""
        if raw:
            return self.profile()
        fmt = {"float32": (0, np.float32), "float16": (1, np.float16)}.get(component_format)
        if fmt is None:
            raise BeifongError(f'The "component_format" parameter must either be equal to "float16" or "float32" here. Found {component_format} instead.')
        h, w, c = self.profile_shape()
        out = np.empty((h, w, max(c - 1, 0)), dtype=fmt[1])
        self._ck(N.lib().bf_profile_develop(self._h, fmt[0], out.ctypes.data_as(C.c_void_p), out.nbytes), "bf_profile_develop")
        return out

    def channel_names(self, raw=True):
        """Channel names of bitmap(raw): gen 3 ["Y", "A", "W"] + "phase.bin<k>" (integrator.cpp:359-370, phase.cpp:140-145); the developed
        bitmap drops the weight channel (hdradc.cpp:246-266)."""
        h, w, c = self.profile_shape()
        names = ["Y", "A", "W"] + [f"phase.bin{k}" for k in range(c - 3)]
        return names if raw else [n for n in names if n != "W"]

    def develop(self, filename, component_format="float16"):
        """HDRADC::develop() (hdradc.cpp:277-295): the developed bitmap written as a multichannel OpenEXR file; the extension is
        forced to ".exr" as the reference does.  Returns the path written."""
        from . import exr
        if not filename:
            raise BeifongError("Destination file not specified, cannot develop.")
        root, ext = os.path.splitext(filename)
        path = filename if ext.lower() == ".exr" else root + ".exr"
        exr.write_exr(path, self.bitmap(raw=False, component_format=component_format), self.channel_names(raw=False))
        return path

    def power_db(self, sample_count, channel=0, mode="mask"):
        """The scripts' own post-processing on the device (Receive.ipynb cells 26-27: mode "mask"; cell 31: mode "abs"):
        10 log10(v (v > 0) / sample_count + eps) resp. 10 log10(|v| / sample_count + eps), one value per bin -> [f][t] float32."""
        h, w, c = self.profile_shape()
        out = np.empty((h, w), dtype=np.float32)
        self._ck(N.lib().bf_profile_power_db(self._h, int(channel), float(sample_count), {"mask": 0, "abs": 1}[mode], _fp(out), out.size), "bf_profile_power_db")
        return out

    @staticmethod
    def range_axis(tr_min, tr_max, bins):
        """Receive.ipynb cells 26-27, 31: xvals = linspace(tr_min, tr_max, tr_bins) / 2 (two-way path length -> range)."""
        return np.linspace(tr_min, tr_max, bins) / 2

    def profile_device(self):
        """(device pointer, n_floats, cuda stream handle) of the accumulated profile."""
        p, n, st = C.c_void_p(), C.c_size_t(), C.c_void_p()
        self._ck(N.lib().bf_profile_device_ptr(self._h, C.byref(p), C.byref(n), C.byref(st)), "bf_profile_device_ptr")
        return p.value, n.value, st.value

    def trace_first_hits(self, origins, directions, mint=None, maxt=None):
        o, d = _f32(origins).reshape(-1, 3), _f32(directions).reshape(-1, 3)
        n = len(o)
        mi = _f32(mint) if mint is not None else None
        ma = _f32(maxt) if maxt is not None else None
        shape, prim = np.empty(n, np.uint32), np.empty(n, np.uint32)
        t, u, v = np.empty(n, np.float32), np.empty(n, np.float32), np.empty(n, np.float32)
        self._ck(N.lib().bf_trace_first_hits(self._h, _fp(o), _fp(d), _fp(mi) if mi is not None else None,
                                             _fp(ma) if ma is not None else None, n, _up(shape), _up(prim), _fp(t), _fp(u), _fp(v)),
                 "bf_trace_first_hits")
        return shape, prim, t, u, v

    def trace_first_hits_static(self, origins, directions, mint=None, maxt=None):
        """bf_trace_first_hits_static: the same query without any allocation inside the library (abi.h); at most
        N.BF_STATIC_TRACE_MAX_RAYS rays per call."""
        o, d = _f32(origins).reshape(-1, 3), _f32(directions).reshape(-1, 3)
        n = len(o)
        mi = _f32(mint) if mint is not None else None
        ma = _f32(maxt) if maxt is not None else None
        shape, prim = np.empty(n, np.uint32), np.empty(n, np.uint32)
        t, u, v = np.empty(n, np.float32), np.empty(n, np.float32), np.empty(n, np.float32)
        self._ck(N.lib().bf_trace_first_hits_static(self._h, _fp(o), _fp(d), _fp(mi) if mi is not None else None,
                                                    _fp(ma) if ma is not None else None, n, _up(shape), _up(prim), _fp(t), _fp(u), _fp(v)),
                 "bf_trace_first_hits_static")
        return shape, prim, t, u, v

    def trace_occluded(self, origins, directions, mint=None, maxt=None):
        o, d = _f32(origins).reshape(-1, 3), _f32(directions).reshape(-1, 3)
        n = len(o)
        mi = _f32(mint) if mint is not None else None
        ma = _f32(maxt) if maxt is not None else None
        hit = np.empty(n, np.uint8)
        self._ck(N.lib().bf_trace_occluded(self._h, _fp(o), _fp(d), _fp(mi) if mi is not None else None,
                                           _fp(ma) if ma is not None else None, n, hit.ctypes.data_as(C.POINTER(C.c_uint8))),
                 "bf_trace_occluded")
        return hit

    def trace_steps(self, origins, directions):
        o, d = _f32(origins).reshape(-1, 3), _f32(directions).reshape(-1, 3)
        steps, t = np.empty(len(o), np.uint32), np.empty(len(o), np.float32)
        self._ck(N.lib().bf_debug_trace_steps(self._h, _fp(o), _fp(d), None, None, len(o), _up(steps), _fp(t)), "bf_debug_trace_steps")
        return steps, t

    def sample_records(self, n_samples, seed=0, first_sample=0):
        recs = (N.bf_sample_record * n_samples)()
        self._ck(N.lib().bf_debug_sample_records(self._h, seed, first_sample, n_samples, recs), "bf_debug_sample_records")
        return np.frombuffer(recs, dtype=RECORD_DTYPE).copy()

    def stats(self):
        st = N.bf_stats()
        self._ck(N.lib().bf_get_stats(self._h, C.byref(st)), "bf_get_stats")
        return {k: getattr(st, k) for k, _ in N.bf_stats._fields_}

    def set_tunable(self, name, value):
        """Launch-shape knob of this scene (bf_set_tunable): never changes a result."""
        self._ck(N.lib().bf_set_tunable(self._h, name.encode(), int(value)), "bf_set_tunable")

    def set_trace_timing(self, enabled):
        self._ck(N.lib().bf_set_trace_timing(self._h, int(enabled)), "bf_set_trace_timing")

    KERNELS = ("raygen", "trace_closest", "trace_shadow", "shade", "finalize", "primary")

    def last_kernel_ms(self):
        ms, n = (C.c_float * len(self.KERNELS))(), (C.c_uint32 * len(self.KERNELS))()
        self._ck(N.lib().bf_last_kernel_ms(self._h, ms, n), "bf_last_kernel_ms")
        return {k: (ms[i], n[i]) for i, k in enumerate(self.KERNELS)}

    def last_launch_list(self, capacity=4096):
        ms, kind, n = (C.c_float * capacity)(), (C.c_uint32 * capacity)(), C.c_uint32()
        self._ck(N.lib().bf_last_launch_list(self._h, ms, kind, capacity, C.byref(n)), "bf_last_launch_list")
        return [(self.KERNELS[kind[i]], ms[i]) for i in range(n.value)]

    def last_receive_ms(self):
        a, b = C.c_float(), C.c_float()
        self._ck(N.lib().bf_last_receive_ms(self._h, C.byref(a), C.byref(b)), "bf_last_receive_ms")
        return a.value, b.value


RECORD_DTYPE = np.dtype([("first_shape", "<u4"), ("first_prim", "<u4"), ("first_t", "<f4"), ("valid", "<u4"), ("tf_t", "<f4"),
                         ("tf_f", "<f4"), ("value", "<f4"), ("phase", "<f4"), ("n_closest", "<u4"), ("n_shadow", "<u4"),
                         ("wavelength0", "<f4"), ("depth", "<u4")])
