/* oracle/oracle.h -- TEST INFRASTRUCTURE. C interface of the CPU oracle (liboracle.so).
 *
 * The oracle is a CPU restatement of prosper's RtReference path: res/shader/rt/reference/
 * main.rgen, rt/scene.{rchit,rahit,rmiss}, rt/ray.glsl, brdf.glsl, common/{random,sampling,
 * math}.glsl, scene/{geometry,instances,materials,lighting,camera}.glsl. The reference has no
 * CPU implementation of this path and its BVH build/traversal live inside the Vulkan driver
 * (World.cpp:740,798; RtRef.cpp:328), and no graphics API can be brought up on these boxes, so
 * PARITY IS UNPINNED by the reference itself: the pins are the known-answer vectors of
 * SURVEY.md 8(c) (tests/golden/), analytic properties, and brute-force-vs-BVH agreement.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference) may load
 * this library. The product (libprt_b200.so) never does.
 */
#ifndef ORACLE_H
#define ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* The same byte layouts as include/prt_types.h, restated so the oracle stands alone. */
typedef struct orc_texture {
    uint32_t format; /* 0 RGBA8 UNORM, 1 BC7 UNORM */
    uint32_t width, height, mipLevels;
    const void *data; /* level 0 first */
    uint64_t dataBytes;
} orc_texture;

typedef struct orc_scene {
    uint32_t geometryBufferCount;
    const uint32_t *const *geometryBuffers;
    const uint64_t *geometryBufferBytes;
    uint32_t meshCount;
    const void *metadatas; /* GeometryMetadata[meshCount], 44 B each */
    const void *meshInfos; /* MeshInfo[meshCount], 16 B each */
    uint32_t materialCount;
    const void *materials; /* MaterialData[], 48 B each */
    uint32_t samplerCount;
    const void *samplers; /* {mag,min,addrU,addrV} u32 x4 */
    uint32_t textureCount; /* including the reserved slot 0 */
    const orc_texture *textures;
    uint32_t modelCount;
    const uint32_t *modelOffsets; /* modelCount+1 */
    const void *subModels;        /* {meshIndex, materialIndex} */
    uint32_t instanceCount;
    const void *tlasInstances; /* VkAccelerationStructureInstanceKHR-shaped, 64 B; blasReference = model+1 */
    uint32_t drawInstanceCount;
    const void *drawInstances; /* DrawInstance[], 12 B */
    uint32_t transformCount;
    const void *transforms; /* ModelInstanceTransforms[], 96 B */
    const void *directionalLight; /* 32 B */
    uint32_t pointLightCount;
    const void *pointLights; /* 32 B each */
    uint32_t spotLightCount;
    const void *spotLights; /* 48 B each */
    const void *camera;     /* CameraUniforms, 532 B */
    uint32_t skyFaceSize;   /* 0 = no cubemap */
    const void *skyRgba16f; /* 6 faces */
} orc_scene;

typedef struct orc_ctx orc_ctx;

/* Copies nothing: the scene's arrays must outlive the context. Builds the oracle's own
 * binary SAH BVH (or none when bruteForce != 0: every ray tests every triangle). */
orc_ctx *orc_create(const orc_scene *scene, int bruteForce);
void orc_destroy(orc_ctx *ctx);
double orc_bvh_build_seconds(const orc_ctx *ctx);

/* main.rgen main() for every pixel and `sampleCount` consecutive frameIndex values starting at
 * pc->frameIndex (wrapping at 4096), accumulated with the running mean of main.rgen:285-298
 * into rgba (width*height*4 floats, in/out). pc = ReferencePC (32 B). threads <= 0: all cores.
 * counters (optional, 4 x u64): closest rays, shadow rays, shaded hits, pixel samples. */
void orc_render(orc_ctx *ctx, const void *pc, uint32_t width, uint32_t height,
                uint32_t sampleCount, float *rgba, int threads, uint64_t *counters);

/* traceClosest() of the primary ray of every pixel for frame pc->frameIndex. hits = width*height
 * records {drawInstanceIndex, primitiveID, bary.x, bary.y}. alphaMargin (optional) receives, per
 * pixel, the smallest |alpha - threshold| any any-hit decision on that ray had (1e30 if none):
 * pixels below a small margin are the "ties" excluded from bit-exact ID comparison. */
void orc_trace_primary(orc_ctx *ctx, const void *pc, uint32_t width, uint32_t height,
                       void *hits, float *alphaMargin, int threads);

/* Generic rays: originTmin/dirTmax = 4 floats per ray, seeds = payload.randomSeed per ray. */
void orc_trace_rays(orc_ctx *ctx, const float *originTmin, const float *dirTmax,
                    const uint32_t *seeds, uint32_t count, int shadow, void *hits,
                    float *alphaMargin, int threads);

/* ReSTIR DI trace pass: res/shader/rt/direct_illumination/main.rgen for one frame over a G-buffer in the formats
 * of GBufferRenderer.cpp:31-32 (RGBA8 albedo+roughness, A2B10G10R10 encoded normal+metallic, non-linear depth) and one
 * R32G32 reservoir per pixel; tracePC = {drawType, frameIndex, flags}. rgba in/out (running mean, main.rgen:150-164). */
void orc_rtdi_trace(orc_ctx *ctx, const uint32_t tracePC[3], uint32_t width, uint32_t height, const uint32_t *albedoRoughness,
                    const uint32_t *normalMetallic, const float *nonLinearDepth, const float *reservoirs, float *rgba, int threads,
                    uint64_t *shadowRays);
/* restir_di/initial_reservoirs.comp: 5-candidate RIS over uniformly picked lights, no visibility */
void orc_rtdi_initial_reservoirs(orc_ctx *ctx, uint32_t frameIndex, uint32_t width, uint32_t height, const uint32_t *albedoRoughness,
                                 const uint32_t *normalMetallic, const float *nonLinearDepth, float *reservoirsOut, int threads);
/* the G-buffer texels gbuffer.frag:57-110 writes, from the primary hit through every pixel centre */
void orc_rtdi_gbuffer(orc_ctx *ctx, uint32_t frameIndex, uint32_t width, uint32_t height, uint32_t *albedoRoughness, uint32_t *normalMetallic,
                      float *nonLinearDepth, int threads);

/* ---- unit entry points for the known-answer / analytic tests ---- */
uint32_t orc_pcg(uint32_t v);                        /* random.glsl:7-12 */
void orc_pcg3d(const uint32_t in[3], uint32_t out[3]); /* random.glsl:17-28 */
float orc_rng_to_01(uint32_t u);                     /* random.glsl:42 */
void orc_cosine_sample_hemisphere(const float n[3], const float u[2], float out[3]);
void orc_sample_vndf(const float ve[3], float alpha, const float u[2], float out[3]);
float orc_vndf_pdf(const float ve[3], const float le[3], float alpha);
void orc_eval_brdf_times_nol(const float l[3], const float v[3], const float n[3],
                             const float albedo[3], float roughness, float metallic, float out[3]);
void orc_offset_ray(const float p[3], const float n[3], float out[3]); /* ray.glsl:83-103 */
void orc_pinhole_ray(const void *camera, const float uv[2], float o[3], float d[3]);
void orc_thin_lens_ray(const void *camera, const float uv[2], const float lens[2], float aperture,
                       float focusDistance, float focalLength, float o[3], float d[3]);
void orc_unpack_snorm10(uint32_t packed, float out[3]); /* geometry.glsl:95-103 */
float orc_half_to_float(uint16_t h);
uint16_t orc_float_to_half(float f);
float orc_srgb_to_linear(float x); /* materials.glsl:26-29 */
/* BC7 block decode, modes 4/5/6 (what the reference's `alpha_ultrafast` cache holds,
 * ispc_texcomp.cpp:198-229). Returns 0, or -1 for another mode. out = 16 RGBA8 texels. */
int orc_bc7_decode_block(const uint8_t block[16], uint8_t out[64]);
/* bilinear texture() at LOD 0 of scene texture `tex` with sampler `smp` */
void orc_sample_texture(orc_ctx *ctx, uint32_t tex, uint32_t smp, float u, float v, float out[4]);
void orc_sample_sky(orc_ctx *ctx, const float d[3], float out[3]);
/* ray vs one triangle with the exact contract; returns 1 on hit and writes t,u,v */
int orc_intersect_triangle(const float o[3], const float d[3], float tmin, float tmax,
                           const float v0[3], const float v1[3], const float v2[3], float tuv[3]);
/* world->object of a row-major 3x4 (exact contract); out = 12 floats row-major */
void orc_invert_transform(const float m[12], float out[12]);

/* Display transform: res/shader/tone_map.comp:17-62 with rgbToHsv / hsvToRgb of common/math.glsl:18-77 over an
 * RGBA16F illumination image; lut = dim^3 R9G9B9E5 texels (tony_mc_mapface.dds payload), sampled trilinearly with
 * clamp-to-edge; out = RGBA8 unorm (ToneMap.cpp:77). */
void orc_tone_map(const uint16_t *illuminationF16, uint32_t width, uint32_t height, float exposure, float contrast,
                  const uint32_t *lutRgb9e5, uint32_t dim, uint8_t *outRgba8);
/* one R9G9B9E5_SHAREDEXP texel -> rgb */
void orc_decode_rgb9e5(uint32_t texel, float out[3]);

#ifdef __cplusplus
}
#endif
#endif
