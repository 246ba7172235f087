/* prt.h -- C ABI of the B200-native RtReference path tracer (libprt_b200.so).
 *
 * Drop-in boundary for prosper's render::RtReference pass (src/render/RtReference.hpp:19-107)
 * and the acceleration-structure half of scene::World it depends on (src/scene/World.hpp:49,
 * 64-65). Plain pointers and sizes only. All `const void *` inputs are HOST pointers unless the
 * name says `_device`; the library copies them to HBM on the context's stream. One context per
 * GPU; calls on a context are single-threaded and stream-ordered. Every function returns
 * PRT_OK or a negative prt_status; prt_last_error() gives the message (the C++ wrapper
 * include/prt_rt_reference.hpp re-throws it as std::runtime_error like the reference does,
 * RtRef.cpp:112-113). There is no CPU fallback: without a CUDA device prt_create fails.
 */
#ifndef PRT_H
#define PRT_H

#include "prt_types.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct prt_ctx prt_ctx;

typedef enum prt_status {
    PRT_OK = 0,
    PRT_ERR_INVALID = -1,   /* programmer error: the reference would WHEELS_ASSERT */
    PRT_ERR_CUDA = -2,      /* CUDA runtime failure */
    PRT_ERR_NO_DEVICE = -3, /* no usable sm_100 device */
    PRT_ERR_CAPACITY = -4   /* a fixed capacity from prt_config was exceeded */
} prt_status;

typedef struct prt_config {
    int32_t device;            /* CUDA ordinal */
    uint32_t maxWidth;         /* largest render area (RtRef.cpp:178-189 image extent) */
    uint32_t maxHeight;
    uint32_t samplesPerLaunch; /* path-state capacity = maxWidth*maxHeight*samplesPerLaunch */
    uint32_t maxInstances;     /* TLAS capacity (World.cpp:838-876 reserveTlasInstances) */
    uint32_t flags;            /* PRT_CFG_* */
} prt_config;
enum {
    PRT_CFG_COUNTERS = 1u << 0,     /* run the instrumented traversal kernels (untimed runs) */
    PRT_CFG_EXACT_FILTER = 1u << 1, /* material taps use the software bilinear of the any-hit stage
                                       instead of the texture units' 8-bit-weight filter (parity tests) */
    PRT_CFG_DECODE_BC7 = 1u << 2,   /* expand BC7 payloads to RGBA8 arrays at upload (a CUDA kernel)
                                       instead of sampling block-compressed CUDA arrays */
    PRT_CFG_PLOC = 1u << 3,         /* BLAS topology by PLOC (agglomerative, surface-area driven) instead of the
                                       default; same collapse, same results, different tree */
    PRT_CFG_LBVH_BLAS = 1u << 4     /* BLAS topology by the Karras Morton hierarchy (what every TLAS build uses)
                                       instead of the default top-down binned-SAH build (PRT_SAH_BINS bins per
                                       axis, one tree level per pass); same collapse, same results */
};

/* ---- lifetime: RtReference::init / ~RtReference (RtRef.cpp:90-120) ------------------------ */
prt_status prt_create(const prt_config *cfg, prt_ctx **out);
void prt_destroy(prt_ctx *ctx);
const char *prt_last_error(void);
/* The CUDA stream every call of this context is ordered on (a cudaStream_t). */
void *prt_stream(prt_ctx *ctx);
prt_status prt_synchronize(prt_ctx *ctx);

/* ---- scene data: WorldData buffers bound through WorldDSLayouts (WorldData.hpp:100-111) --- */
/* geometry buffer `index` (WorldData.cpp:1271-1320; bytes multiple of 4) */
prt_status prt_set_geometry_buffer(prt_ctx *ctx, uint32_t index, const void *data, uint64_t bytes);
/* GeometryMetadata[] + MeshInfo[] (WorldData.cpp:481-530 uploadMeshDatas) */
prt_status prt_set_meshes(prt_ctx *ctx, const prt_geometry_metadata *metadatas,
                          const prt_mesh_info *infos, uint32_t count);
/* MaterialData[] (WorldData.cpp:532-586 uploadMaterialDatas) */
prt_status prt_set_materials(prt_ctx *ctx, const prt_material_data *materials, uint32_t count);
/* sampler table (WorldData.cpp:686-720) */
prt_status prt_set_samplers(prt_ctx *ctx, const prt_sampler_desc *samplers, uint32_t count);
/* texture `index` >= 1 of the bindless table (Texture.cpp:369-489); index 0 is the reserved
 * "no texture" slot (WorldData.cpp:727-741) */
prt_status prt_set_texture(prt_ctx *ctx, uint32_t index, const prt_texture_desc *desc);
/* Texture-cache build (Texture.cpp:213-296 `compress`, where prosper calls ispc_texcomp's CompressBlocksBC7): encodes
 * one RGBA8 level (width, height multiples of 4; HOST memory) into BC7 blocks (width/4 * height/4 * 16 B; HOST) with a
 * CUDA kernel. The blocks are what prt_set_texture(PRT_TEX_BC7_UNORM) and the .dds cache take. */
prt_status prt_encode_bc7(prt_ctx *ctx, const void *rgba8, uint32_t width, uint32_t height, void *blocksOut,
                          uint64_t blocksBytes);

/* Texture-cache build, the whole chain of `compress` (Texture.cpp:213-296) on the GPU: RGBA8 level 0 (HOST) -> mip
 * levels (generateMipLevels, 165-211: stb's default 2:1 reduction = Mitchell-Netravali, 8 taps per axis, alpha-weighted,
 * colour filtered in linear light when colorSpace == 0 = sRGB, as is when 1 = linear) ->
 * BC7 blocks per level, or plain RGBA8 levels when a kept level is not whole 4x4 blocks (229-237). Levels are the
 * full chain minus the last two (217-225). dataOut (HOST) receives the levels back to back exactly as utils::Dds::data
 * holds them (Dds.cpp:95-133), so writeDds / prosper_b200.dds.write_dds can store it unchanged. */
prt_status prt_compress_texture(prt_ctx *ctx, const void *rgba8, uint32_t width, uint32_t height, uint32_t colorSpace,
                                uint32_t generateMipMaps, void *dataOut, uint64_t dataBytes, uint32_t *formatOut,
                                uint32_t *mipLevelCountOut);

/* prt_compress_texture without any device allocation. NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it). Same arguments, same kernels and bit-identical output as prt_compress_texture, but the intermediate mip
 * levels, the block staging and the filter tables live in one file-scope `static __device__` array of
 * PRT_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit; the call makes no cudaMalloc / cudaFree.
 * Limit: level 0 of at most PRT_STATIC_MAX_TEXELS texels (1024 x 1024 or any other shape of that area); more is refused
 * with PRT_ERR_INVALID and a message naming the limit. All calls (all contexts on one GPU) share the array and it keeps
 * the last call's contents: calls that overlap on the GPU race, so serialise them. rgba8 / dataOut are HOST memory. */
#define PRT_STATIC_MAX_TEXELS (1024u * 1024u)
#define PRT_STATIC_SCRATCH_BYTES (12u * 1024u * 1024u) /* tables 4352 B + chain < 2 x 4 B/texel + blocks < 2 x 1 B/texel */
prt_status prt_compress_texture_static(prt_ctx *ctx, const void *rgba8, uint32_t width, uint32_t height, uint32_t colorSpace,
                                       uint32_t generateMipMaps, void *dataOut, uint64_t dataBytes, uint32_t *formatOut,
                                       uint32_t *mipLevelCountOut);

/* Display transform (render::ToneMap, src/render/ToneMap.cpp + res/shader/tone_map.comp).
 * prt_set_tone_map_lut = ToneMap::init (ToneMap.cpp:31-42): the dim^3 3D table in DXGI R9G9B9E5_SHAREDEXP texels, x
 * fastest -- the payload of res/texture/tony_mc_mapface.dds (dim 48) as it lies in the file; HOST memory.
 * prt_tone_map = ToneMap::record (63-134): exposure, HSV contrast, Tony McMapface, 1/2.2 gamma over the current
 * Output::illumination (the fp16-rounded running mean) into R8G8B8A8 unorm, width*height*4 bytes of HOST memory;
 * prt_tone_map_device leaves the image on the GPU (valid until the next tone-map call), ordered on the stream. */
typedef struct prt_tone_map_pc { float exposure; float contrast; } prt_tone_map_pc; /* push_constants/tone_map.h */
prt_status prt_set_tone_map_lut(prt_ctx *ctx, const uint32_t *rgb9e5, uint32_t dim);
prt_status prt_tone_map(prt_ctx *ctx, const prt_tone_map_pc *pc, void *dstHost, uint64_t dstBytes);
prt_status prt_tone_map_device(prt_ctx *ctx, const prt_tone_map_pc *pc, const uint32_t **out);
/* skybox cubemap, RGBA16F, faces +X,-X,+Y,-Y,+Z,-Z (Texture.cpp:595-601; skybox.glsl:4) */
prt_status prt_set_sky_cubemap(prt_ctx *ctx, uint32_t faceSize, const void *rgba16f);
/* models: sub-models of model m are subModels[offsets[m] .. offsets[m+1]) */
prt_status prt_set_models(prt_ctx *ctx, const uint32_t *offsets, const prt_sub_model *subModels,
                          uint32_t modelCount);

/* ---- acceleration structures: World::buildAccelerationStructures (World.cpp:538-575) ------ */
/* buildNextBlas (World.cpp:585-747). *outReference goes into prt_tlas_instance.blasReference */
prt_status prt_build_blas(prt_ctx *ctx, uint32_t modelIndex, uint64_t *outReference);
/* updateTlasInstances + buildCurrentTlas (World.cpp:878-928, 749-802): full rebuild */
prt_status prt_build_tlas(prt_ctx *ctx, const prt_tlas_instance *instances, uint32_t count);

/* ---- per-frame rings: World::updateBuffers (World.cpp:468-536), Camera::updateBuffer ------- */
prt_status prt_set_draw_instances(prt_ctx *ctx, const prt_draw_instance *instances, uint32_t count);
prt_status prt_set_instance_transforms(prt_ctx *ctx, const prt_model_instance_transforms *t,
                                       uint32_t count);
prt_status prt_set_lights(prt_ctx *ctx, const prt_directional_light *directional,
                          const prt_point_light *points, uint32_t pointCount,
                          const prt_spot_light *spots, uint32_t spotCount);
prt_status prt_set_camera(prt_ctx *ctx, const prt_camera_uniforms *camera);

/* ---- the pass: RtReference::record (RtRef.cpp:161-383) -------------------------------------- */
/* Renders `sampleCount` consecutive frames (frameIndex = pc->frameIndex + k, wrapping at
 * PRT_FRAME_PERIOD like RtRef.cpp:170) into the preserved fp32 accumulation image with the
 * running mean of main.rgen:285-298; PRT_FLAG_SKIP_HISTORY applies to the first of them only.
 * width*height must fit prt_config (renderArea.offset must be 0, RtRef.cpp:327). */
prt_status prt_record(prt_ctx *ctx, const prt_reference_pc *pc, uint32_t width, uint32_t height,
                      uint32_t sampleCount);
/* m_previousIllumination: RGBA fp32, sample count in .a (DEVICE pointer, width*height*16 B) */
prt_status prt_output_f32_device(prt_ctx *ctx, const float **out);
/* Output::illumination: RGBA fp16 blit of the above (RtRef.cpp:336-377), into HOST memory */
prt_status prt_read_illumination_f16(prt_ctx *ctx, void *dstHost, uint64_t dstBytes);
prt_status prt_read_output_f32(prt_ctx *ctx, float *dstHost, uint64_t dstBytes);
/* RtReference::releasePreserved (RtRef.cpp:385-391): drops the accumulation history */
prt_status prt_release_preserved(prt_ctx *ctx);

/* ---- ReSTIR DI trace pass: render::rtdi::Trace (src/render/rtdi/Trace.hpp:19-56, Trace.cpp:176-362;
 * res/shader/rt/direct_illumination/main.rgen) and the pass that feeds it, rtdi::InitialReservoirs
 * (restir_di/initial_reservoirs.comp). Trace::Input = {G-buffer, reservoirs}: the G-buffer in the formats of
 * GBufferRenderer.cpp:31-32 -- albedo+roughness R8G8B8A8_UNORM, encoded normal+metallic A2B10G10R10_UNORM_PACK32, non-linear
 * reverse-Z depth D32 (0 = nothing drawn) -- and one R32G32 reservoir per pixel (light index bits, unbiased contribution
 * weight; restir_di/reservoir.glsl). The pass keeps its own running-mean history like RtReference does. -------------- */
typedef struct prt_rtdi_trace_pc { /* shader_structs/push_constants/restir_di/trace.h */
    uint32_t drawType;
    uint32_t frameIndex;
    uint32_t flags; /* PRT_FLAG_SKIP_HISTORY | PRT_FLAG_ACCUMULATE (Trace.cpp:255-263) */
} prt_rtdi_trace_pc;
/* the caller's (raster) G-buffer; HOST pointers, width*height texels each; fixes the pass's resolution */
prt_status prt_rtdi_set_gbuffer(prt_ctx *ctx, uint32_t width, uint32_t height, const uint32_t *albedoRoughness,
                                const uint32_t *normalMetallic, const float *nonLinearDepth);
/* or: the G-buffer a raster pass would have produced, from the tracer's own primary hits through the pixel centres
 * (gbuffer.frag:57-110; the raster G-buffer pass itself is out of scope) */
prt_status prt_rtdi_gbuffer_from_primary(prt_ctx *ctx, uint32_t width, uint32_t height, uint32_t frameIndex);
prt_status prt_rtdi_read_gbuffer(prt_ctx *ctx, uint32_t *albedoRoughness, uint32_t *normalMetallic, float *nonLinearDepth);
/* reservoirs: the caller's (2 floats per pixel; HOST), or restir_di/initial_reservoirs.comp on the device */
prt_status prt_rtdi_set_reservoirs(prt_ctx *ctx, const float *packed);
prt_status prt_rtdi_initial_reservoirs(prt_ctx *ctx, uint32_t frameIndex);
prt_status prt_rtdi_read_reservoirs(prt_ctx *ctx, float *packedOut);
/* Trace::record: one frame = one shadow ray per lit pixel through the TLAS of prt_build_tlas, running mean as main.rgen:150-164 */
prt_status prt_rtdi_trace(prt_ctx *ctx, const prt_rtdi_trace_pc *pc);
prt_status prt_rtdi_read_output_f32(prt_ctx *ctx, float *dstHost, uint64_t dstBytes);      /* RGBA fp32, count in .a */
prt_status prt_rtdi_read_illumination_f16(prt_ctx *ctx, void *dstHost, uint64_t dstBytes);  /* Trace::Output::illumination */
prt_status prt_rtdi_release_preserved(prt_ctx *ctx);                                        /* Trace::releasePreserved */

/* Primary-ray hit records (traceClosest payload, main.rgen:62-81) for frame pc->frameIndex:
 * what the reference can only show through its DrawType debug views (main.rgen:181-193). */
prt_status prt_trace_primary(prt_ctx *ctx, const prt_reference_pc *pc, uint32_t width,
                             uint32_t height, prt_hit *dstHost);
/* Traverse caller-supplied rays (origin xyz + tMin, direction xyz + tMax per ray; HOST) with the
 * closest-hit or the shadow (terminate-on-first-hit, main.rgen:49-60) kernel. */
prt_status prt_trace_rays(prt_ctx *ctx, const float *originTmin, const float *dirTmax,
                          const uint32_t *seeds, uint32_t count, int shadow, prt_hit *dstHost);

/* Test hook: what the bindless slot `textureSampler` ((sampler << 24) | texture, material_data.h:19-36)
 * returns. texelXY == NULL: texture() at `uv` (2 floats per sample) through the hardware filter
 * (exact == 0) or the software bilinear of the any-hit stage (exact != 0); otherwise the raw texels at
 * integer coordinates texelXY (2 ints per sample). outRgba = 4 floats per sample. HOST pointers. */
prt_status prt_probe_texture(prt_ctx *ctx, uint32_t textureSampler, const float *uv, const int32_t *texelXY,
                             uint32_t count, int exact, float *outRgba);

/* Test hooks: read the built structures back (80-byte wide nodes, 48-byte triangle records, 96-byte
 * instance records; layouts in prosper_b200/csrc/core_bvh.h) so the tests can validate the GPU
 * builder's output structurally on the host. Pass NULL buffers to query the counts. triCount counts triangle
 * RECORDS: a non-opaque triangle appears once per builder reference, with identical contents. */
prt_status prt_debug_read_blas(prt_ctx *ctx, uint32_t modelIndex, void *nodesOut, uint64_t nodesBytes,
                               void *trisOut, uint64_t trisBytes, uint32_t *nodeCount, uint32_t *triCount);
prt_status prt_debug_read_tlas(prt_ctx *ctx, void *nodesOut, uint64_t nodesBytes, void *instOut,
                               uint64_t instBytes, uint32_t *instanceCount);

/* ---- multi-GPU: one ctx per rank, accumulation kept as sum + count ------------- */
/* Device pointer to the fp32 RGBA *sum* image (rgb = sum of samples, a = count) suitable for
 * ncclReduce(sum); prt_resolve_sum turns a reduced sum back into the running-mean image. */
prt_status prt_export_sum_device(prt_ctx *ctx, float **outSum);
/* STREAM-ORDERED on prt_stream(ctx), no host synchronisation: `sumDevice` must be complete in that stream's order (make
 * prt_stream() the collective's stream, or have it wait on the collective's event) before this is called. */
prt_status prt_import_sum_device(prt_ctx *ctx, const float *sumDevice);
/* Same as prt_export_sum_device but writes the sum image into caller-owned DEVICE memory (width*height*16 B,
 * e.g. a buffer registered with the caller's communicator). STREAM-ORDERED on prt_stream(ctx): the copy is enqueued,
 * not awaited, so a collective enqueued behind prt_stream() overlaps nothing on the host. */
prt_status prt_export_sum_to_device(prt_ctx *ctx, float *dstDevice);

/* ---- metrics (the reference's "RtReference"/"BuildTLAS" profiler scopes, App.cpp:575) ------ */
typedef struct prt_counters {
    uint64_t closestRays;   /* traceRayEXT equivalents, main.rgen:73 */
    uint64_t shadowRays;    /* main.rgen:57 */
    uint64_t nodesVisited;  /* 80-byte wide nodes fetched (TLAS + BLAS) */
    uint64_t trisTested;    /* 48-byte triangle records fetched */
    uint64_t instancesEntered; /* 96-byte instance records fetched */
    uint64_t anyHitCalls;   /* scene.rahit invocations */
    uint64_t shadedHits;    /* evaluateSurface calls */
    uint64_t textureTaps;   /* bilinear material/alpha fetches */
    uint64_t pixelSamples;  /* rgen invocations */
    uint64_t kernelLaunches;
    uint64_t stackOverflows; /* traversal stack overflows (must stay 0; checked by the tests) */
    /* the part of nodesVisited / trisTested / instancesEntered / anyHitCalls spent by shadow rays
     * (so closest-ray traffic = total - shadow part; bench.py's per-kernel roofline needs the split) */
    uint64_t shadowNodesVisited, shadowTrisTested, shadowInstancesEntered, shadowAnyHitCalls;
} prt_counters;
prt_status prt_get_counters(prt_ctx *ctx, prt_counters *out);
prt_status prt_reset_counters(prt_ctx *ctx);
typedef struct prt_timings {
    float lastRecordMs;   /* CUDA-event time of the last prt_record ("RtReference" scope) */
    float lastTlasMs;     /* ... of the last prt_build_tlas ("BuildTLAS" scope) */
    float lastBlasMs;     /* ... of the last prt_build_blas */
    float lastTraceClosestMs; /* sum over ALL trace-closest launches of the last prt_record */
    uint32_t lastTraceClosestLaunches;
    uint32_t lastRecordLaunches;
    uint32_t lastTlasLaunches;
    float lastShadeMs;        /* likewise for the shade and trace-shadow stages */
    uint32_t lastShadeLaunches;
    float lastTraceShadowMs;
    uint32_t lastTraceShadowLaunches;
    uint32_t ringWraps;       /* laps of the pinned upload ring so far (gfx::RingBuffer wraps to 0, RingBuffer.cpp:56-133) */
    uint32_t tlasGraphReplays; /* prt_build_tlas calls so far that replayed the captured CUDA graph */
    uint32_t tlasGraphCaptures; /* ... that captured it (first repeat of an instance count, or a scratch buffer moved) */
} prt_timings;
prt_status prt_get_timings(prt_ctx *ctx, prt_timings *out);

/* BVH introspection for tests (sizes in elements) */
typedef struct prt_bvh_stats {
    uint32_t tlasNodes;
    uint32_t blasNodesTotal;
    uint64_t blasTrianglesTotal;
    uint32_t blasCount;
    uint32_t tlasInstances;
    uint64_t blasReferencesTotal; /* triangle records in the BLASes: a non-opaque triangle is referenced from several leaves */
} prt_bvh_stats;
prt_status prt_get_bvh_stats(prt_ctx *ctx, prt_bvh_stats *out);

#ifdef __cplusplus
}
#endif
#endif /* PRT_H */
