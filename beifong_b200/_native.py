"""ctypes binding of the C ABI in include/beifong_b200/abi.h.

The shared library is built in-tree (beifong_b200/libbeifong_b200.so, see __graft_entry__.build or
beifong_b200/csrc/Makefile).  Loading fails loudly when it is missing: there is no Python or CPU
fallback for the compute path.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BF_NATIVE_LIB") or os.path.join(_HERE, "libbeifong_b200.so")   # BF_NATIVE_LIB: an alternative build of the same library (tuning experiments)

BF_NONE = 0xFFFFFFFF
BF_STATIC_TRACE_MAX_RAYS = 65536      # abi.h
BF_OK, BF_ERR_INVALID, BF_ERR_UNSUPPORTED, BF_ERR_CUDA, BF_ERR_STATE, BF_ERR_IO = 0, -1, -2, -3, -4, -5
BF_SPEC_UNIFORM, BF_SPEC_REGULAR, BF_SPEC_IRREGULAR, BF_SPEC_ZERO_OUTSIDE_VISIBLE = 0, 1, 2, 3
BF_BSDF_DIFFUSE, BF_BSDF_ROUGHCONDUCTOR, BF_BSDF_CONDUCTOR, BF_BSDF_NULL_BLACK = 0, 1, 2, 3
BF_MICROFACET_BECKMANN, BF_MICROFACET_GGX = 0, 1
BF_TX_AREA, BF_TX_WIGNER, BF_TX_PHASED, BF_TX_SPOT, BF_TX_AREALIGHT = 0, 1, 2, 3, 4
BF_RX_OMNIDIRECTIONAL, BF_RX_WIGNER, BF_RX_PHASED, BF_RX_FLUXMETER, BF_RX_PERSPECTIVE = 0, 1, 2, 3, 4
BF_SIG_CW, BF_SIG_LINFMCW, BF_SIG_PULSE, BF_SIG_OTHER, BF_SIG_UNSET = 0, 1, 2, 3, 4
BF_RECV_RAW, BF_RECV_RAW_RESAMPLE, BF_RECV_MIX_RESAMPLE, BF_RECV_MIXER = 0, 1, 2, 3
BF_GEN_TIME, BF_GEN_RANGE, BF_GEN_PHASE = 1, 2, 3

u32, i32, f32, u64 = C.c_uint32, C.c_int32, C.c_float, C.c_uint64


class bf_constants(C.Structure):
    _fields_ = [("c", f32), ("lambda_min_nm", f32), ("lambda_max_nm", f32), ("cie_min_nm", f32),
                ("cie_max_nm", f32), ("P", f32), ("spectrum_width", u32), ("reserved", u32)]


class bf_bsdf_desc(C.Structure):
    _fields_ = [("kind", u32), ("two_sided", u32), ("back_bsdf", u32), ("reflectance", u32), ("eta", u32),
                ("k", u32), ("specular_reflectance", u32), ("distribution", u32), ("sample_visible", u32),
                ("alpha_u", f32), ("alpha_v", f32), ("reserved", u32)]


class bf_signal_desc(C.Structure):
    _fields_ = [("signaltype", u32), ("sig_is_delta", u32), ("resample_freq", u32), ("amplitude", f32),
                ("repfreq", f32), ("t_ext", f32), ("f_centre", f32), ("f_ext", f32), ("phi0", f32), ("gain", f32)]


class bf_array_desc(C.Structure):
    _fields_ = [("n_elems", u32), ("steering_vector", f32 * 3), ("array_loc", f32 * 16), ("elem_dims", f32 * 3),
                ("elem_spacing", f32 * 3), ("elem_axis", f32 * 3)]


class bf_endpoint_desc(C.Structure):
    _fields_ = [("kind", u32), ("texture", u32), ("signal", bf_signal_desc), ("array", bf_array_desc),
                ("adc_sampling_start", f32), ("adc_sampling_time", f32), ("receive_type", u32),
                ("to_world", f32 * 16), ("cutoff_angle_deg", f32), ("beam_width_deg", f32),
                ("near_clip", f32), ("far_clip", f32), ("sample_to_camera", f32 * 16),
                ("shutter_open", f32), ("shutter_open_time", f32), ("film_width", u32), ("film_height", u32)]


class bf_adc_desc(C.Structure):
    _fields_ = [("t_bins", u32), ("f_bins", u32), ("window_t_bins", u32), ("window_f_bins", u32),
                ("window_offset_t", u32), ("window_offset_f", u32), ("t_bandwidth", f32), ("f_bandwidth", f32),
                ("filter_radius", f32), ("filter_kind", u32)]


class bf_integrator_desc(C.Structure):
    _fields_ = [("gen", u32), ("max_depth", i32), ("rr_depth", i32), ("phase_bins", u32), ("range_bins", u32),
                ("dr", f32), ("rng_mode", u32), ("doppler", u32)]


class bf_sample_record(C.Structure):
    _fields_ = [("first_shape", u32), ("first_prim", u32), ("first_t", f32), ("valid", u32), ("tf_t", f32),
                ("tf_f", f32), ("value", f32), ("phase", f32), ("n_closest", u32), ("n_shadow", u32),
                ("wavelength0", f32), ("depth", u32)]


class bf_batch_item(C.Structure):
    _fields_ = [("seed", u64), ("first_sample", u64), ("n_rect_updates", u32), ("receiver_shape", u32),
                ("rect_shape_ids", C.POINTER(u32)), ("rect_to_world", C.POINTER(f32)), ("receiver", C.POINTER(bf_endpoint_desc))]


class bf_stats(C.Structure):
    _fields_ = [("samples", u64), ("closest_queries", u64), ("shadow_queries", u64), ("kernel_launches", u64),
                ("bvh_nodes", u32), ("bvh_tris", u32), ("bvh_depth", u32), ("n_rects", u32),
                ("build_ms", C.c_double), ("upload_ms", C.c_double),
                ("path_closest_queries", u64), ("path_shadow_queries", u64), ("mesh_bounds_fraction", C.c_double)]


# name -> (restype, argtypes); also the list of symbols include/beifong_b200/abi.h declares
P = C.POINTER
vp = C.c_void_p
PROTOTYPES = {
    "bf_abi_version": (u32, []),
    "bf_scene_create": (C.c_int, [P(bf_constants), P(vp)]),
    "bf_scene_destroy": (None, [vp]),
    "bf_last_error": (C.c_char_p, [vp]),
    "bf_scene_add_spectrum": (C.c_int, [vp, u32, u32, P(f32), P(f32), P(u32)]),
    "bf_scene_add_bsdf": (C.c_int, [vp, P(bf_bsdf_desc), P(u32)]),
    "bf_scene_add_mesh": (C.c_int, [vp, P(f32), u32, P(u32), u32, P(f32), u32, P(u32)]),
    "bf_scene_add_rect": (C.c_int, [vp, P(f32), u32, P(u32)]),
    "bf_scene_set_shape_velocity": (C.c_int, [vp, u32, P(f32)]),
    "bf_scene_add_transmitter": (C.c_int, [vp, u32, P(bf_endpoint_desc), P(u32)]),
    "bf_scene_set_receiver": (C.c_int, [vp, u32, P(bf_endpoint_desc)]),
    "bf_scene_set_adc": (C.c_int, [vp, P(bf_adc_desc)]),
    "bf_scene_set_integrator": (C.c_int, [vp, P(bf_integrator_desc)]),
    "bf_scene_commit": (C.c_int, [vp, C.c_int]),
    "bf_scene_update_rect": (C.c_int, [vp, u32, P(f32)]),
    "bf_scene_update_transform": (C.c_int, [vp, u32, P(f32)]),
    "bf_scene_reupload": (C.c_int, [vp]),
    "bf_scene_device_bytes": (C.c_int, [vp, P(C.c_size_t)]),
    "bf_profile_channels": (C.c_int, [vp, P(u32)]),
    "bf_profile_size": (C.c_int, [vp, P(u32), P(u32), P(u32)]),
    "bf_receive": (C.c_int, [vp, u64, u64, u64, C.c_int]),
    "bf_receive_batch": (C.c_int, [vp, u32, P(bf_batch_item), u64, P(f32), C.c_size_t]),
    "bf_batch_device_ptr": (C.c_int, [vp, P(vp), P(C.c_size_t), P(vp)]),
    "bf_profile_read": (C.c_int, [vp, P(f32), C.c_size_t]),
    "bf_profile_develop": (C.c_int, [vp, u32, vp, C.c_size_t]),
    "bf_profile_power_db": (C.c_int, [vp, u32, C.c_double, C.c_int, P(f32), C.c_size_t]),
    "bf_profile_device_ptr": (C.c_int, [vp, P(vp), P(C.c_size_t), P(vp)]),
    "bf_synchronize": (C.c_int, [vp]),
    "bf_scene_set_stream": (C.c_int, [vp, vp]),
    "bf_multi_unique_id": (C.c_int, [vp, C.c_size_t]),
    "bf_multi_init": (C.c_int, [vp, C.c_int, C.c_int, vp, C.c_size_t]),
    "bf_multi_reduce": (C.c_int, [vp, C.c_int]),
    "bf_multi_reduce_comm": (C.c_int, [vp, vp, C.c_int]),
    "bf_trace_first_hits": (C.c_int, [vp, P(f32), P(f32), P(f32), P(f32), u64, P(u32), P(u32), P(f32), P(f32), P(f32)]),
    "bf_trace_first_hits_static": (C.c_int, [vp, P(f32), P(f32), P(f32), P(f32), u64, P(u32), P(u32), P(f32), P(f32), P(f32)]),
    "bf_trace_occluded": (C.c_int, [vp, P(f32), P(f32), P(f32), P(f32), u64, P(C.c_uint8)]),
    "bf_debug_trace_steps": (C.c_int, [vp, P(f32), P(f32), P(f32), P(f32), u64, P(u32), P(f32)]),
    "bf_debug_sample_records": (C.c_int, [vp, u64, u64, u64, P(bf_sample_record)]),
    "bf_get_stats": (C.c_int, [vp, P(bf_stats)]),
    "bf_last_receive_ms": (C.c_int, [vp, P(f32), P(f32)]),
    "bf_set_trace_timing": (C.c_int, [vp, C.c_int]),
    "bf_set_tunable": (C.c_int, [vp, C.c_char_p, C.c_int64]),
    "bf_last_kernel_ms": (C.c_int, [vp, P(f32), P(u32)]),
    "bf_last_launch_list": (C.c_int, [vp, P(f32), P(u32), u32, P(u32)]),
    "bf_scene_export": (C.c_int, [vp, vp, C.c_size_t, P(C.c_size_t)]),
}

_lib = None


def lib():
    """The loaded native library.  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"beifong_b200: native library not built ({LIB_PATH}); run `python -c 'import __graft_entry__ as g; "
                "g.build()'` or `make -C beifong_b200/csrc`. There is no CPU fallback.")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(handle, name)   # AttributeError if the symbol is missing
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib
