/*
 * beifong_b200 — C ABI of the B200-native radar path-integration hot path.
 *
 * This is the drop-in boundary. The reference (JacobMackay/beifong, a
 * Mitsuba 2 fork) has no C FFI for this path: the path sits behind Mitsuba's C++ plugin
 * interfaces.  Each entry point below names the reference interface it replaces
 * (file:line under /root/reference).  A Mitsuba-side plugin flattens
 * `Scene` into these calls from `SamplingIntegrator::receive()` /
 * `SamplingIntegrator::render()` (src/librender/integrator.cpp:315-768, :58-204) and
 * wraps the returned profile in a `SignalBlock` for `ADC::put`
 * (src/adcs/hdradc.cpp:172-176).
 *
 * Conventions: plain C, opaque handle, `int` status (0 = ok, <0 = error; message via
 * bf_last_error).  Caller owns every host buffer passed in; it is copied during the
 * call.  The library owns device memory until bf_scene_destroy.  A handle is not
 * thread-safe; distinct handles are.  No exceptions cross the ABI.  All arithmetic is
 * fp32 (mitsuba.conf: every enabled variant has "float": "float"), indices u32,
 * RNG state u64.  Matrices are row-major float[16].
 *
 * Scene construction is host-only; bf_scene_commit is the first call that touches
 * CUDA and fails loudly (BF_ERR_CUDA) when no device / extension is usable.  There is
 * NO CPU fallback anywhere behind this ABI.
 */
#ifndef BEIFONG_B200_ABI_H
#define BEIFONG_B200_ABI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BF_ABI_VERSION 1u
#define BF_NONE 0xffffffffu
#define BF_MAX_SPECTRUM_WIDTH 4u

/* status codes */
enum {
    BF_OK = 0,
    BF_ERR_INVALID = -1,     /* bad argument / inconsistent scene (reference: Throw(...)) */
    BF_ERR_UNSUPPORTED = -2, /* plugin or parameter outside the hot-path scope (NotImplementedError) */
    BF_ERR_CUDA = -3,        /* no device, launch failure, out of memory */
    BF_ERR_STATE = -4,       /* call order violated (e.g. receive before commit) */
    BF_ERR_IO = -5           /* file could not be read / parsed (host front end) */
};

typedef struct bf_scene bf_scene;

/* Compile-time physics macros of the reference turned into run-time scene constants.
 * include/mitsuba/core/spectrum.h:15-40 (MTS_WAVELENGTH_MIN/MAX, MTS_C, MTS_P),
 * include/mitsuba/core/math.h:40-41 (CVac, CAir).  spectrum_width: 1 = *_mono,
 * 4 = *_spectral (spectrum.h:79-80, traits.h:21,43). */
typedef struct bf_constants {
    float c;              /* propagation speed [m/s]; reference default CAir = 340 */
    float lambda_min_nm;  /* MTS_WAVELENGTH_MIN; default 7555556 */
    float lambda_max_nm;  /* MTS_WAVELENGTH_MAX; default 9714286 */
    float cie_min_nm;     /* MTS_CIE_MIN (spectrum.h:160) used by sample_uniform_spectrum */
    float cie_max_nm;     /* MTS_CIE_MAX (spectrum.h:161) */
    float P;              /* MTS_P, unused by the live integrators, carried for completeness */
    uint32_t spectrum_width; /* 1 or 4 */
    uint32_t reserved;
} bf_constants;

/* Spectrum textures (src/spectra/uniform.cpp:27-40, regular.cpp:68-85,
 * irregular.cpp:76-96; distr_1d.h:387-403,664-687). */
enum { BF_SPEC_UNIFORM = 0, BF_SPEC_REGULAR = 1, BF_SPEC_IRREGULAR = 2,
       BF_SPEC_ZERO_OUTSIDE_VISIBLE = 3 /* Texture::D65 default evaluated at radar wavelengths */ };

/* BSDFs (src/bsdfs/diffuse.cpp:78-135, twosided.cpp:94-181, roughconductor.cpp:196-394,
 * conductor.cpp). */
enum { BF_BSDF_DIFFUSE = 0, BF_BSDF_ROUGHCONDUCTOR = 1, BF_BSDF_CONDUCTOR = 2, BF_BSDF_NULL_BLACK = 3 };
enum { BF_MICROFACET_BECKMANN = 0, BF_MICROFACET_GGX = 1 };

typedef struct bf_bsdf_desc {
    uint32_t kind;                 /* BF_BSDF_* */
    uint32_t two_sided;            /* wrapped in <bsdf type="twosided"> */
    uint32_t back_bsdf;            /* second nested BSDF of twosided, BF_NONE = same as front */
    uint32_t reflectance;          /* spectrum id (diffuse) */
    uint32_t eta, k;               /* spectrum ids (conductors) */
    uint32_t specular_reflectance; /* spectrum id or BF_NONE */
    uint32_t distribution;         /* BF_MICROFACET_* */
    uint32_t sample_visible;       /* roughconductor.cpp "sample_visible", default 1 */
    float alpha_u, alpha_v;        /* default 0.1 */
    uint32_t reserved;
} bf_bsdf_desc;

/* Endpoints.  Transmitter side: src/transmitters/{areatransmitter,wignertransmitter,
 * phasedtransmitter}.cpp and the gen-1/2 emitters src/emitters/{spot,area}.cpp.
 * Receiver side: src/receivers/{omnidirectional,wignerreceiver,phasedreceiver}.cpp,
 * src/librender/receiver.cpp:16-62 and the gen-1/2 sensors
 * src/sensors/{fluxmeter,perspective}.cpp. */
enum { BF_TX_AREA = 0, BF_TX_WIGNER = 1, BF_TX_PHASED = 2, BF_TX_SPOT = 3, BF_TX_AREALIGHT = 4 };
enum { BF_RX_OMNIDIRECTIONAL = 0, BF_RX_WIGNER = 1, BF_RX_PHASED = 2, BF_RX_FLUXMETER = 3,
       BF_RX_PERSPECTIVE = 4 };
enum { BF_SIG_CW = 0, BF_SIG_LINFMCW = 1, BF_SIG_PULSE = 2, BF_SIG_OTHER = 3, BF_SIG_UNSET = 4 };
enum { BF_RECV_RAW = 0, BF_RECV_RAW_RESAMPLE = 1, BF_RECV_MIX_RESAMPLE = 2, BF_RECV_MIXER = 3 };

#define BF_MAX_ARRAY_ELEMS 64u    /* config C5's 64-element array as ONE phased endpoint: 64^2 virtual elements */

typedef struct bf_signal_desc {     /* wignertransmitter.cpp:59-102, phasedreceiver.cpp:67-113 */
    uint32_t signaltype;            /* BF_SIG_* */
    uint32_t sig_is_delta;
    uint32_t resample_freq;
    float amplitude;
    float repfreq;                  /* crf / prf */
    float t_ext;                    /* chirp_len / pulse_len */
    float f_centre;                 /* freq_centre */
    float f_ext;                    /* freq_sweep / freq_ext */
    float phi0;                     /* phase */
    float gain;
} bf_signal_desc;

typedef struct bf_array_desc {      /* phasedtransmitter.cpp:108-167 */
    uint32_t n_elems;               /* <= BF_MAX_ARRAY_ELEMS */
    float steering_vector[3];
    float array_loc[16];            /* "array_loc" transform */
    float elem_dims[3];
    float elem_spacing[3];
    float elem_axis[3];
} bf_array_desc;

typedef struct bf_endpoint_desc {
    uint32_t kind;                  /* BF_TX_* or BF_RX_* */
    uint32_t texture;               /* spectrum id: antenna_texture / radiance / intensity */
    bf_signal_desc signal;
    bf_array_desc array;
    /* receiver base (receiver.cpp:16-62) */
    float adc_sampling_start;
    float adc_sampling_time;        /* adc_sampling_end - adc_sampling_start */
    uint32_t receive_type;          /* BF_RECV_* */
    /* gen-1/2 endpoints without a parent shape */
    float to_world[16];             /* spot / perspective */
    float cutoff_angle_deg;         /* spot.cpp */
    float beam_width_deg;
    float near_clip, far_clip;      /* perspective.cpp */
    float sample_to_camera[16];     /* inverse of perspective_projection(...) (sensor.h:196-231) */
    float shutter_open, shutter_open_time;
    uint32_t film_width, film_height;
} bf_endpoint_desc;

/* ADC / film (src/librender/adc.cpp:7-76, src/adcs/hdradc.cpp:107-176;
 * gen 1/2: src/films/hdrfilm.cpp). */
typedef struct bf_adc_desc {
    uint32_t t_bins, f_bins;              /* ADC::size() */
    uint32_t window_t_bins, window_f_bins;
    uint32_t window_offset_t, window_offset_f;
    float t_bandwidth, f_bandwidth;       /* defaults 3.81e-6 s, 250e6 Hz */
    float filter_radius;                  /* box = 0.5; > 0.5+RayEpsilon selects the wide splat (signalblock.cpp:113-161) */
    uint32_t filter_kind;                 /* 0 box, 1 gaussian (radius = 4 stddev), 2 tent (radius 1), 3 mitchell (B = C = 1/3), 4 catmullrom (radius 2), 5 lanczos (radius = lobes) */
} bf_adc_desc;

/* Integrators.  gen 3: phase(pathtimefrequency) (src/integrators/phase.cpp:63-145,
 * pathtimefrequency.cpp:103-477); gen 2: range(pathlength) (range.cpp:63-221,
 * pathlength.cpp:114-342); gen 1: time(pathtime) (time.cpp:86-206, pathtime.cpp:114-302).
 * Base parameters: src/librender/integrator.cpp:1713-1728. */
enum { BF_GEN_TIME = 1, BF_GEN_RANGE = 2, BF_GEN_PHASE = 3 };

typedef struct bf_integrator_desc {
    uint32_t gen;          /* BF_GEN_* */
    int32_t max_depth;     /* -1 = infinite */
    int32_t rr_depth;      /* default 5 */
    uint32_t phase_bins;   /* gen 3: phase.cpp "bins"; 0 = bare pathtimefrequency */
    uint32_t range_bins;   /* gen 2: range.cpp "bins"; gen 1 fixed 50 */
    float dr;              /* gen 2: range.cpp "dr" */
    uint32_t rng_mode;     /* 0 = per-sample streams (sampler.cpp:89-92), the only GPU mode */
    uint32_t doppler;      /* gen 3.  0 (default) = what the reference executes today: its Doppler statements are commented out
                            * ("Took doppler out to test", pathtimefrequency.cpp:124-126, 141-144, 252-256).  1 = the statement of
                            * :254-256 restored: after the transmitter connection of every bounce the path's wavelengths are shifted
                            * by Shape::doppler of the shape hit (shape.cpp:375-389), from its `velocity` transform. */
} bf_integrator_desc;

/* Per-sample record written by bf_debug_sample_records (parity hook, no reference
 * counterpart: it exposes the locals of receive_sample, integrator.cpp:1538-1667). */
typedef struct bf_sample_record {
    uint32_t first_shape, first_prim; /* first closest hit; BF_NONE on miss */
    float first_t;
    uint32_t valid;                   /* result.second */
    float tf_t, tf_f;                 /* fractional bin coordinates after :1639 */
    float value;                      /* aovs[0] = hsum(|ray_weight| * L) or 0 */
    float phase;                      /* hsum(ray.phase) as seen by phase.cpp:117 */
    uint32_t n_closest, n_shadow;     /* ray queries issued by this sample */
    float wavelength0;                /* ray_.wavelengths[0] after the path */
    uint32_t depth;                   /* loop iterations executed */
} bf_sample_record;

typedef struct bf_stats {
    uint64_t samples;          /* path samples processed by bf_receive calls so far */
    uint64_t closest_queries;  /* closest-hit ray queries */
    uint64_t shadow_queries;   /* any-hit ray queries */
    uint64_t kernel_launches;  /* kernels launched by this library */
    uint32_t bvh_nodes, bvh_tris, bvh_depth, n_rects;
    double build_ms, upload_ms;
    uint64_t path_closest_queries; /* of closest_queries / shadow_queries: decided inside k_path (rectangles + mesh bounds, */
    uint64_t path_shadow_queries;  /* no traversal launch) */
    double mesh_bounds_fraction;   /* bf_scene_commit's estimate: share of cosine-weighted receiver rays that enter the padded mesh bounds */
} bf_stats;

uint32_t bf_abi_version(void);

/* Scene::Scene (src/librender/scene.cpp:22-120) */
int bf_scene_create(const bf_constants *consts, bf_scene **out);
void bf_scene_destroy(bf_scene *s);
const char *bf_last_error(const bf_scene *s); /* s may be NULL: error of a failed create */

/* Texture plugins `uniform` / `regular` / `irregular` */
int bf_scene_add_spectrum(bf_scene *s, uint32_t kind, uint32_t n, const float *lambda_nm,
                          const float *values, uint32_t *spectrum_id);
/* BSDF plugins */
int bf_scene_add_bsdf(bf_scene *s, const bf_bsdf_desc *desc, uint32_t *bsdf_id);

/* Mesh (src/librender/mesh.cpp, src/shapes/ply.cpp:195-322, obj.cpp): world-space
 * positions float[3V], faces u32[3F], optional vertex normals float[3V] (NULL =
 * face normals).  Primitive numbering == face order (first-hit ids are compared in it). */
int bf_scene_add_mesh(bf_scene *s, const float *positions, uint32_t n_vertices,
                      const uint32_t *faces, uint32_t n_faces, const float *normals,
                      uint32_t bsdf_id, uint32_t *shape_id);
/* Rectangle (src/shapes/rectangle.cpp:75-95): to_world already includes flip_normals. */
int bf_scene_add_rect(bf_scene *s, const float to_world[16], uint32_t bsdf_id,
                      uint32_t *shape_id);

/* The shape's `velocity` transform (Shape::Shape, src/librender/shape.cpp:39-42; default identity).  Only read by
 * Shape::doppler (shape.cpp:375-389), i.e. when the integrator's doppler flag is set.  Before or after bf_scene_commit. */
int bf_scene_set_shape_velocity(bf_scene *s, uint32_t shape_id, const float velocity[16]);

/* Endpoint attached to a shape (Shape::Shape, src/librender/shape.cpp:38-99) or free
 * standing (shape_id = BF_NONE: spot, perspective). */
int bf_scene_add_transmitter(bf_scene *s, uint32_t shape_id, const bf_endpoint_desc *desc,
                             uint32_t *transmitter_id);
int bf_scene_set_receiver(bf_scene *s, uint32_t shape_id, const bf_endpoint_desc *desc);
int bf_scene_set_adc(bf_scene *s, const bf_adc_desc *desc);
int bf_scene_set_integrator(bf_scene *s, const bf_integrator_desc *desc);

/* Scene::accel_init_* (scene_native.inl:3-10 / scene_embree.inl:24-39): SAH BVH build,
 * compressed 8-wide nodes, upload.  Re-commit after bf_scene_update_* is allowed. */
int bf_scene_commit(bf_scene *s, int device);
int bf_scene_update_rect(bf_scene *s, uint32_t shape_id, const float to_world[16]);
/* Moving geometry between frames (the reference rebuilds its whole scene per frame, python_scripts/animated_trans_rad.py:289-311).
 * Rectangle: the same as bf_scene_update_rect.  Mesh: `motion` is applied to the vertices AS GIVEN to bf_scene_add_mesh (the rest
 * pose; motions do not accumulate), with Transform::transform_affine's fp32 operations (transform.h:91-99); vertex normals get the
 * inverse transpose and are re-normalised (mesh.cpp:117-124).  Before bf_scene_commit the moved vertices go into the build; after
 * it the device triangles are re-posed and the 8-wide BVH is REFITTED (same topology, new boxes, one small kernel per tree level)
 * on the scene's stream, without a rebuild or a host synchronisation.  Hits on a refitted scene equal those of a scene built on the
 * moved vertices bit for bit (same triangles); only the traversal cost can degrade when the motion is large. */
int bf_scene_update_transform(bf_scene *s, uint32_t shape_id, const float motion[16]);
/* Copy the flattened geometry (BVH nodes, triangles, normals) again from the library's pinned host
 * image to the device without rebuilding: the host->device leg of Scene::accel_init_gpu
 * (scene_optix.inl:236-300) on its own, used to time the path end to end from host buffers. */
int bf_scene_reupload(bf_scene *s);
int bf_scene_device_bytes(const bf_scene *s, size_t *geometry_bytes);

/* Number of fp32 channels per bin: gen 3: ["Y","A","W"] + phase_bins
 * (integrator.cpp:359-370, phase.cpp:140-145); gen 1/2: ["X","Y","Z","A","W"] + bins. */
int bf_profile_channels(const bf_scene *s, uint32_t *channels);
int bf_profile_size(const bf_scene *s, uint32_t *width, uint32_t *height, uint32_t *channels);

/* SamplingIntegrator::receive / render hot loop (integrator.cpp:659-663 / :228-231):
 * sample indices [first_sample, first_sample + n_samples), per-sample PCG32 streams
 * seeded with (tea64(seed, idx), tea64(idx, seed)) (sampler.cpp:89-92).  accumulate = 0
 * clears the device profile first (SignalBlock::clear, signalblock.cpp:33-39).
 * Asynchronous on the scene's stream. */
int bf_receive(bf_scene *s, uint64_t seed, uint64_t first_sample, uint64_t n_samples,
               int accumulate);
/* Many receive() calls as one submission: the frames of a moving-target sequence (python_scripts/animated_trans_rad.py:289-311
 * re-poses the antennas and calls the integrator once per frame) or the receivers of a scene (include/mitsuba/render/scene.h:197
 * holds std::vector<ref<Receiver>>; Receive.ipynb loops over them).  Item i first re-poses the listed rectangles
 * (bf_scene_update_rect) and / or replaces the receiver (bf_scene_set_receiver), then integrates sample indices
 * [first_sample, first_sample + n_samples) with its seed into profile i of a device-side batch buffer.  No host
 * synchronisation between items.  profiles != NULL: all item profiles are copied to it ([item][f][t][channel], n_floats >=
 * n_items * profile floats) and the stream is synchronised; NULL: asynchronous, read them through bf_batch_device_ptr.
 * The scene is left in the state of the last item. */
typedef struct bf_batch_item {
    uint64_t seed, first_sample;
    uint32_t n_rect_updates;          /* rectangles re-posed before this item */
    uint32_t receiver_shape;          /* shape the new receiver is attached to (BF_NONE: free standing); used if receiver != NULL */
    const uint32_t *rect_shape_ids;   /* [n_rect_updates] */
    const float *rect_to_world;       /* [n_rect_updates][16], row major */
    const bf_endpoint_desc *receiver; /* NULL: keep the current receiver */
} bf_batch_item;
int bf_receive_batch(bf_scene *s, uint32_t n_items, const bf_batch_item *items, uint64_t n_samples,
                     float *profiles, size_t n_floats);
int bf_batch_device_ptr(bf_scene *s, void **dptr, size_t *n_floats, void **cuda_stream);
/* HDRADC::bitmap(raw=true) (hdradc.cpp:219-233): [f][t][channel] fp32, un-normalised.
 * Synchronises the scene's stream. */
int bf_profile_read(bf_scene *s, float *dst, size_t n_floats);
/* HDRADC::bitmap(raw=false) (hdradc.cpp:219-275), the step after the path: channel 2 of the storage is the
 * Struct weight field; every other channel is multiplied by 1.f / weight (struct.cpp:1761-1766, 1810-1811) and
 * stored as [f][t][channels - 1] in the ADC's component format (hdradc.cpp:134-144; the reference's default is
 * float16).  Bins that received no sample develop to NaN (0 * inf), as in the reference.  Runs on the device,
 * then copies to `dst` (dst_bytes >= bins * (channels - 1) * sizeof(component)); synchronises the scene's
 * stream.  `uint32` components and 3-channel storages (no AOV channels: the reference's converter looks up a
 * "Y" field the unnamed source does not have) return BF_ERR_UNSUPPORTED. */
enum { BF_COMPONENT_FLOAT32 = 0, BF_COMPONENT_FLOAT16 = 1 };
int bf_profile_develop(bf_scene *s, uint32_t component_format, void *dst, size_t dst_bytes);
/* The scripts' own post-processing of bitmap(raw=True) (python_scripts/Receive.ipynb cells 26-27, 31), one value per bin of one
 * channel, evaluated on the device in double as numpy does:  mode 0: 10 log10(v (v > 0) / sample_count + eps);
 * mode 1: 10 log10(|v| / sample_count + eps), eps = 2^-52.  dst: [f][t] fp32, n_floats >= bins.  Synchronises the scene's stream.
 * The range axis of those plots is linspace(tr_min, tr_max, bins) / 2 (host arithmetic, not offered here). */
int bf_profile_power_db(bf_scene *s, uint32_t channel, double sample_count, int mode, float *dst, size_t n_floats);
/* Device pointer of the accumulated profile + the stream it is produced on (for a caller
 * that reduces it with NCCL, §8e). */
int bf_profile_device_ptr(bf_scene *s, void **dptr, size_t *n_floats, void **cuda_stream);
int bf_synchronize(bf_scene *s);
/* Caller-supplied stream (the plugin runs inside a host application that owns its streams; the reference's
 * GPU variant launches on the default stream, src/librender/scene_optix.inl:284,336,348, i.e. ordered with the application's own work).  After bf_scene_commit every entry point
 * of the scene enqueues on ONE stream -- the library's own by default; this call synchronises that stream and adopts
 * `cuda_stream` (a cudaStream_t of the scene's device) instead, NULL goes back to the library's own.  bf_receive still fans
 * its waves out over internal streams, but forks from and joins back into this stream, so the call is ordered after the
 * caller's earlier work on it and the caller's later work sees the finished profile.  The caller keeps ownership: the
 * stream must stay alive until the scene is destroyed or given another one.  BF_ERR_STATE before commit. */
int bf_scene_set_stream(bf_scene *s, void *cuda_stream);

/* Multi-GPU (the reference has no equivalent): one process or thread per GPU, the scene replicated, sample
 * indices sharded by the caller (disjoint [first_sample, first_sample + n) per rank), then ONE ncclReduce of the per-rank
 * profiles onto `root`, in place, on the scene's stream behind the last wavefront (no host synchronisation).  NCCL is
 * loaded at run time (dlopen "libnccl.so.2"; BF_ERR_UNSUPPORTED with the loader's message when it is absent).
 *   bf_multi_unique_id   rank 0: ncclGetUniqueId into id (capacity >= BF_MULTI_ID_BYTES); the caller ships the bytes to the others
 *   bf_multi_init        every rank, after bf_scene_commit: ncclCommInitRank on the scene's device (blocks until all ranks joined)
 *   bf_multi_reduce      ncclReduce(profile, profile, n_floats, ncclFloat32, ncclSum, root) with the communicator of bf_multi_init
 *   bf_multi_reduce_comm the same with a caller-owned ncclComm_t (a host that already has one, e.g. torch.distributed's)
 * fp32 summation order differs from a single-GPU run: multi-GPU parity is <= 1e-3 relative L2, not bitwise. */
#define BF_MULTI_ID_BYTES 128
int bf_multi_unique_id(void *id, size_t capacity);
int bf_multi_init(bf_scene *s, int rank, int world, const void *id, size_t id_bytes);
int bf_multi_reduce(bf_scene *s, int root);
int bf_multi_reduce_comm(bf_scene *s, void *nccl_comm, int root);

/* Parity / measurement hooks */
/* Scene::ray_intersect_preliminary (scene.cpp:150-156) for caller-provided rays. */
int bf_trace_first_hits(bf_scene *s, const float *origins, const float *directions,
                        const float *mint, const float *maxt, uint64_t n_rays,
                        uint32_t *shape, uint32_t *prim, float *t, float *u, float *v);
/* Scene::ray_test (scene.cpp:170-178) */
int bf_trace_occluded(bf_scene *s, const float *origins, const float *directions,
                      const float *mint, const float *maxt, uint64_t n_rays, uint8_t *hit);
/* bf_trace_first_hits for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments, results and defaults as bf_trace_first_hits; every device-side temporary lives in one static device array of
 * the library (6.25 MiB, held by every GPU that loads the library until the process exits), so the call makes no cudaMalloc /
 * cudaFree / allocating library call (nor a host allocation).  Limits and rules:
 *   - n_rays <= BF_STATIC_TRACE_MAX_RAYS, else BF_ERR_INVALID with a message naming the limit;
 *   - device ordinal < BF_STATIC_TRACE_MAX_DEVICES;
 *   - scene edits made after bf_scene_commit must have been applied by an earlier bf_receive / bf_trace_* call (applying them may
 *     allocate): BF_ERR_STATE otherwise;
 *   - the array is shared by every call on that GPU and keeps the last call's contents (the call clears what it needs): calls
 *     that overlap on one GPU -- two scenes on two host threads -- race.  Calls on one scene are ordered on its stream.
 * origins / directions / mint / maxt and the outputs may be host or device pointers (copied with cudaMemcpyDefault on the
 * scene's stream, which is synchronised before returning).  Null outputs are skipped; null mint / maxt = the defaults. */
#define BF_STATIC_TRACE_MAX_RAYS 65536u
#define BF_STATIC_TRACE_MAX_DEVICES 64
int bf_trace_first_hits_static(bf_scene *s, const float *origins, const float *directions,
                               const float *mint, const float *maxt, uint64_t n_rays,
                               uint32_t *shape, uint32_t *prim, float *t, float *u, float *v);
/* Tuning hook: BVH traversal steps (wide nodes visited) per closest-hit query. */
int bf_debug_trace_steps(bf_scene *s, const float *origins, const float *directions, const float *mint, const float *maxt,
                         uint64_t n_rays, uint32_t *steps, float *t);
int bf_debug_sample_records(bf_scene *s, uint64_t seed, uint64_t first_sample,
                            uint64_t n_samples, bf_sample_record *records);
int bf_get_stats(const bf_scene *s, bf_stats *stats);
/* Time (ms, CUDA events on the scene's stream) of the most recent bf_receive: whole call
 * and the traverse kernels alone. */
int bf_last_receive_ms(bf_scene *s, float *total_ms, float *traverse_ms);
/* Launch-shape knobs of one scene (no reference counterpart; none changes a result).  Names: "wave_lanes" (0 = from the request
 * size), "wave_streams" (waves in flight, 1..16), "fused_primary" (0/1), "primary_defer" / "trace_defer" (-1 = the commit-time
 * estimate, 0, 1), "trace_rounds" (batches of 32 rays per active traversal warp, 0 = whole grid), "use_graph" (0/1), "mega_max"
 * (largest request run as one k_mega launch, -1 = never), "mega_batch" / "path_fused" (0/1), "l2_persist_mb" (MB of the top of the BVH
 * node array put under a persisting L2 access-policy window; 0 = off, the default: measured no gain on C3 / C5.  The carve-out
 * `cudaLimitPersistingL2CacheSize` it sets is per DEVICE, shared by every scene on it).  Defaults are read from the environment
 * variables BF_<NAME> once, in bf_scene_create. */
int bf_set_tunable(bf_scene *s, const char *name, int64_t value);
/* Record CUDA events around every traversal launch so that bf_last_receive_ms can report
 * traverse_ms (-1 when disabled, the default). */
int bf_set_trace_timing(bf_scene *s, int enabled);
/* Per-kernel device time (ms) and launch count of the most recent bf_receive, indexed by BF_K_*;
 * needs bf_set_trace_timing(s, 1).  Used for the roofline of the dominant kernel. */
enum { BF_K_RAYGEN = 0, BF_K_TRACE_CLOSEST = 1, BF_K_TRACE_SHADOW = 2, BF_K_SHADE = 3, BF_K_FINALIZE = 4,
       BF_K_PRIMARY = 5 /* fused raygen + first closest-hit query + compaction */, BF_K_COUNT = 6 };
int bf_last_kernel_ms(bf_scene *s, float *ms, uint32_t *launches);
/* The same, launch by launch in issue order (kind[i] = BF_K_*). */
int bf_last_launch_list(bf_scene *s, float *ms, uint32_t *kind, uint32_t capacity, uint32_t *n_out);

/* Flat export of everything added so far (test infrastructure feeds it to the CPU oracle;
 * also a cheap checkpoint: profile + (seed, next index) + this blob is a resumable state). */
int bf_scene_export(const bf_scene *s, void *dst, size_t capacity, size_t *required);

#ifdef __cplusplus
}
#endif
#endif /* BEIFONG_B200_ABI_H */
