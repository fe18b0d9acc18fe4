/*
 * Flat scene blob written by bf_scene_export (abi.h).  It is the recorded sequence of
 * scene-construction ABI calls, in call order, so that replaying it through
 * bf_scene_add_* rebuilds the same scene with the same shape / bsdf / spectrum ids.
 *
 * Layout: bf_flat_header, then records; each record is
 *     u32 tag, u32 reserved, u64 payload_bytes, payload (padded to 8 bytes).
 * Payloads are the POD descriptors of abi.h followed by their arrays:
 *   BF_TAG_SPECTRUM : u32 kind, u32 n, float lambda[n], float value[n]
 *   BF_TAG_BSDF     : bf_bsdf_desc
 *   BF_TAG_MESH     : u32 n_vertices, u32 n_faces, u32 bsdf, u32 has_normals,
 *                     float pos[3V], u32 faces[3F], float normals[3V] (if has_normals)
 *   BF_TAG_RECT     : float to_world[16], u32 bsdf, u32 pad
 *   BF_TAG_TX       : u32 shape, u32 pad, bf_endpoint_desc
 *   BF_TAG_RX       : u32 shape, u32 pad, bf_endpoint_desc
 *   BF_TAG_ADC      : bf_adc_desc
 *   BF_TAG_INTEGRATOR : bf_integrator_desc
 *   BF_TAG_VELOCITY : u32 shape, u32 pad, float velocity[16]
 *   BF_TAG_MESH_MOTION : u32 shape, u32 pad, float motion[16]   (bf_scene_update_transform on a mesh; the last record per shape wins)
 */
#ifndef BEIFONG_B200_FLAT_SCENE_H
#define BEIFONG_B200_FLAT_SCENE_H

#include "abi.h"

#ifdef __cplusplus
extern "C" {
#endif

#define BF_FLAT_MAGIC 0x31464642u /* "BFF1" */

typedef struct bf_flat_header {
    uint32_t magic;
    uint32_t abi_version;
    uint64_t total_bytes;
    bf_constants consts;
} bf_flat_header;

enum {
    BF_TAG_SPECTRUM = 1, BF_TAG_BSDF = 2, BF_TAG_MESH = 3, BF_TAG_RECT = 4,
    BF_TAG_TX = 5, BF_TAG_RX = 6, BF_TAG_ADC = 7, BF_TAG_INTEGRATOR = 8, BF_TAG_VELOCITY = 9, BF_TAG_MESH_MOTION = 10
};

typedef struct bf_flat_record {
    uint32_t tag;
    uint32_t reserved;
    uint64_t payload_bytes;
} bf_flat_record;

#ifdef __cplusplus
}
#endif
#endif
