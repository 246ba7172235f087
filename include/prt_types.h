/* prt_types.h -- data contract of the RtReference path (plain C, tightly packed PODs).
 *
 * Every struct here restates, byte for byte, a layout the reference shares between its
 * C++ host and its GLSL shaders (res/shader/shared/shader_structs/...). Sizes are pinned by
 * static asserts and by tests/test_contract.py.
 */
#ifndef PRT_TYPES_H
#define PRT_TYPES_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ss/scene/geometry_metadata.h:13-33 (44 B). Offsets are in u32 units, except
 * indicesOffset which is in u16 units when usesShortIndices == 1. 0xFFFFFFFF = unused. */
typedef struct prt_geometry_metadata {
    uint32_t bufferIndex;
    uint32_t indicesOffset;
    uint32_t positionsOffset;
    uint32_t normalsOffset;
    uint32_t tangentsOffset;
    uint32_t texCoord0sOffset;
    uint32_t meshletsOffset;
    uint32_t meshletBoundsOffset;
    uint32_t meshletVerticesOffset;
    uint32_t meshletTrianglesByteOffset;
    uint32_t usesShortIndices;
} prt_geometry_metadata;

/* src/scene/Mesh.hpp:17-23 (16 B) */
typedef struct prt_mesh_info {
    uint32_t vertexCount;
    uint32_t indexCount;
    uint32_t meshletCount;
    uint32_t materialIndex;
} prt_mesh_info;

/* ss/scene/material_data.h:53-68 (48 B). *TextureSampler = (sampler << 24) | texture,
 * texture 0 = "no texture" (material_data.h:19-36; materials.glsl:52-63). */
enum { PRT_ALPHA_OPAQUE = 0, PRT_ALPHA_MASK = 1, PRT_ALPHA_BLEND = 2 };
typedef struct prt_material_data {
    float baseColorFactor[4];
    float metallicFactor;
    float roughnessFactor;
    float alphaCutoff;
    uint32_t alphaMode;
    uint32_t baseColorTextureSampler;
    uint32_t metallicRoughnessTextureSampler;
    uint32_t normalTextureSampler;
    uint32_t pad;
} prt_material_data;

/* ss/scene/draw_instance.h:13-18 (12 B) */
typedef struct prt_draw_instance {
    uint32_t modelInstanceIndex;
    uint32_t meshIndex;
    uint32_t materialIndex;
} prt_draw_instance;

/* ss/scene/model_instance_transforms.h:13-17 (96 B): two glm::mat3x4 = 3 columns of vec4.
 * modelToWorld = transpose(M) so column c holds row c of the affine matrix
 * (World.cpp:405); normalToWorld = first three columns of inverse(M) (World.cpp:407). */
typedef struct prt_model_instance_transforms {
    float modelToWorld[3][4];
    float normalToWorld[3][4];
} prt_model_instance_transforms;

/* ss/scene/lights.h:18-41 */
typedef struct prt_directional_light {
    float irradiance[4];
    float direction[4];
} prt_directional_light; /* 32 B */
typedef struct prt_point_light {
    float radianceAndRadius[4];
    float position[4];
} prt_point_light; /* 32 B */
typedef struct prt_spot_light {
    float radianceAndAngleScale[4];
    float positionAndAngleOffset[4];
    float direction[4];
} prt_spot_light; /* 48 B */
#define PRT_MAX_POINT_LIGHTS 1024u /* RtRef.cpp:429-430 compile-time maxima */
#define PRT_MAX_SPOT_LIGHTS 1024u

/* ss/scene/camera.h:13-35 (532 B). mat4 are column-major: m[col][row]. */
typedef struct prt_camera_uniforms {
    float worldToCamera[4][4];
    float cameraToWorld[4][4];
    float cameraToClip[4][4];
    float clipToWorld[4][4];
    float previousWorldToCamera[4][4];
    float previousCameraToClip[4][4];
    float eye[4];
    float nearPlane[4];
    float farPlane[4];
    float leftPlane[4];
    float rightPlane[4];
    float topPlane[4];
    float bottomPlane[4];
    uint32_t resolution[2];
    float currentJitter[2];
    float previousJitter[2];
    float near_;
    float far_;
    float maxViewScale;
} prt_camera_uniforms;

/* ss/push_constants/rt_reference.h:6-16 (32 B) */
enum {
    PRT_FLAG_SKIP_HISTORY = 1u << 0, /* RtRef.cpp:68-88 pcFlags() */
    PRT_FLAG_ACCUMULATE = 1u << 1,
    PRT_FLAG_IBL = 1u << 2,
    PRT_FLAG_DEPTH_OF_FIELD = 1u << 3,
    PRT_FLAG_CLAMP_INDIRECT = 1u << 4
};
typedef struct prt_reference_pc {
    uint32_t drawType;
    uint32_t flags;
    uint32_t frameIndex;
    float apertureDiameter;
    float focusDistance;
    float focalLength;
    uint32_t rouletteStartBounce;
    uint32_t maxBounces;
} prt_reference_pc;
#define PRT_MAX_BOUNCES 6u    /* RtRef.hpp:22 sMaxBounces, MAX_BOUNCES define RtRef.cpp:410 */
#define PRT_FRAME_PERIOD 4096u /* RtRef.cpp:31 */

/* src/scene/DrawType.hpp:8-10 */
enum {
    PRT_DRAW_DEFAULT = 0, PRT_DRAW_PRIMITIVE_ID, PRT_DRAW_MESHLET_ID, PRT_DRAW_MESH_ID,
    PRT_DRAW_MATERIAL_ID, PRT_DRAW_POSITION, PRT_DRAW_SHADING_NORMAL, PRT_DRAW_TEXCOORD0,
    PRT_DRAW_ALBEDO, PRT_DRAW_ROUGHNESS, PRT_DRAW_METALLIC, PRT_DRAW_COUNT
};

/* VkAccelerationStructureInstanceKHR as filled by World.cpp:878-928 (64 B):
 * transform = the instance's modelToWorld mat3x4 reinterpreted as a row-major 3x4,
 * customIndex = index of the instance's first DrawInstance, mask 0xFF, flags 0,
 * blasReference = handle from prt_build_blas (0 = inactive instance, World.cpp:898-905). */
typedef struct prt_tlas_instance {
    float transform[3][4];
    uint32_t customIndexAndMask; /* customIndex:24 | mask:8 (mask in the high byte) */
    uint32_t sbtOffsetAndFlags;
    uint64_t blasReference;
} prt_tlas_instance;

/* One sub-model = one BLAS geometry (World.cpp:617-668; src/scene/Model.hpp SubModel) */
typedef struct prt_sub_model {
    uint32_t meshIndex;
    uint32_t materialIndex;
} prt_sub_model;

/* Sampler state the path actually uses (WorldData.cpp:686-720). Mip selection: the RT stages
 * have no derivatives, texture() reads LOD 0 (materials.glsl:44). */
enum { PRT_FILTER_NEAREST = 0, PRT_FILTER_LINEAR = 1 };
enum { PRT_ADDRESS_REPEAT = 0, PRT_ADDRESS_MIRRORED_REPEAT = 1, PRT_ADDRESS_CLAMP_TO_EDGE = 2 };
typedef struct prt_sampler_desc {
    uint32_t magFilter;
    uint32_t minFilter;
    uint32_t addressU;
    uint32_t addressV;
} prt_sampler_desc;

/* Texture payloads as the texture cache stores them (Texture.cpp:213-296, 429-432):
 * BC7 UNORM blocks (16 B per 4x4) or the RGBA8 UNORM fallback; mip levels concatenated,
 * level 0 first. */
enum { PRT_TEX_RGBA8_UNORM = 0, PRT_TEX_BC7_UNORM = 1 };
typedef struct prt_texture_desc {
    uint32_t format;
    uint32_t width;
    uint32_t height;
    uint32_t mipLevels;
    const void *data;
    uint64_t dataBytes;
} prt_texture_desc;

/* Primary-ray hit record, the payload scene.rchit:10-15 writes (16 B). */
typedef struct prt_hit {
    uint32_t drawInstanceIndex; /* 0xFFFFFFFF = miss */
    uint32_t primitiveID;
    float bary[2];
} prt_hit;

#ifdef __cplusplus
}
static_assert(sizeof(prt_geometry_metadata) == 44, "GeometryMetadata");
static_assert(sizeof(prt_material_data) == 48, "MaterialData");
static_assert(sizeof(prt_draw_instance) == 12, "DrawInstance");
static_assert(sizeof(prt_model_instance_transforms) == 96, "ModelInstanceTransforms");
static_assert(sizeof(prt_point_light) == 32, "PointLight");
static_assert(sizeof(prt_spot_light) == 48, "SpotLight");
static_assert(sizeof(prt_camera_uniforms) == 532, "CameraUniforms");
static_assert(sizeof(prt_reference_pc) == 32, "ReferencePC"); /* mirrors RtRef.cpp:485-486 */
static_assert(sizeof(prt_tlas_instance) == 64, "VkAccelerationStructureInstanceKHR");
#endif

#endif /* PRT_TYPES_H */
