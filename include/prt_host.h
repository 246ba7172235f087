/* prt_host.h -- C ABI of the scene host (libprt_host.so): glTF ingestion and the per-frame scene update that feed prt.h.
 *
 * It is the stand-alone form of the part of prosper's `scene::World` the RtReference path consumes:
 *   - loadGltf / loadTextures / loadMaterials / loadModels / loadAnimations / loadScenes / gatherScene
 *     (src/scene/WorldData.cpp:103-138, 681-958, 960-1120, 1124-1543),
 *   - the mesh pipeline getMeshData -> generateTangents (MikkTSpace) -> optimizeMeshData (meshoptimizer vertex cache,
 *     overdraw, vertex fetch) -> generateMeshlets -> packMeshData -> writeCache
 *     (src/scene/DeferredLoadingContext.cpp:123-490, 609-800) -- the same calls in the same order, so the triangle order
 *     that defines gl_PrimitiveID is the reference's,
 *   - updateAnimations / updateScene / updateBuffers / updateTlasInstances
 *     (src/scene/World.cpp:349-536, 878-927; src/scene/Animations.hpp:71-125; src/scene/Accessors.cpp:25-74).
 * Built from this repo's host/gltf_host.cpp against the reference's own vendored C / C++ libraries where they lie
 * (ext/cgltf, ext/glm, ext/meshoptimizer, ext/mikktspace, ext/stb); nothing of the reference is copied.
 *
 * Plain pointers and sizes; every array getter fills caller storage sized by prth_counts. Status 0 = ok, otherwise
 * prth_last_error() explains. Not thread-safe per scene. The structs are the ones of prt_types.h. */
#ifndef PRT_HOST_H
#define PRT_HOST_H
#include <stdint.h>
#include "prt_types.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct prth_scene prth_scene;

typedef struct prth_counts {
    uint32_t meshes;          /* glTF primitives = MeshInfo entries (WorldData.cpp:829-835) */
    uint32_t materials;       /* 1 + glTF materials (index 0 = the default material, WorldData.cpp:762) */
    uint32_t samplers;        /* 1 + glTF samplers (index 0 = linear / repeat, WorldData.cpp:686-699) */
    uint32_t textures;        /* 1 + glTF images (index 0 = res/texture/empty.png, WorldData.cpp:724-738) */
    uint32_t models;          /* glTF meshes */
    uint32_t subModels;       /* sum of primitives */
    uint32_t nodes;           /* scene nodes of the current scene after gatherScene */
    uint32_t modelInstances;
    uint32_t drawInstances;
    uint32_t pointLights;
    uint32_t spotLights;
    uint32_t animatedNodes;   /* nodes with dynamicTransform after propagation */
    uint32_t hasCamera;       /* the current scene holds a perspective camera node */
    float endTimeS;           /* Scene::endTimeS */
} prth_counts;

typedef struct prth_camera {
    float eye[3], target[3], up[3];
    float fov, zN, zF;        /* CameraParameters (Camera.hpp:40-45) */
} prth_camera;

const char *prth_last_error(void);

/* Parses `gltfPath` (.gltf + .bin + external images), gathers materials / models / animations / the default scene and
 * runs the mesh pipeline: `cacheDir`/prosper_cache/cache<i>.prosper_mesh is (re)written for every mesh whose cache is
 * missing, of another version or older than the source file (cacheValid, DeferredLoadingContext.cpp:560-574).
 * `sourceWriteTime` overrides the .gltf's modification time in the cache headers when non-zero (fixtures). */
int prth_load_gltf(const char *gltfPath, const char *cacheDir, int64_t sourceWriteTime, prth_scene **out);
void prth_destroy(prth_scene *scene);
int prth_get_counts(const prth_scene *scene, prth_counts *out);

/* static scene description */
int prth_get_materials(const prth_scene *scene, prt_material_data *out);
int prth_get_samplers(const prth_scene *scene, prt_sampler_desc *out);
int prth_get_mesh_infos(const prth_scene *scene, prt_mesh_info *out);           /* after the cache pass */
int prth_get_models(const prth_scene *scene, uint32_t *modelOffsets /* models + 1 */, prt_sub_model *subModels);
/* Image `texture` - 1 of the glTF (texture 0 is the built-in empty texture: uri "" and linear 0). `linear` = the image
 * is in linearColorImages (DeferredLoadingContext.cpp:1132-1173). Returns the uri relative to the .gltf. */
int prth_get_texture_source(const prth_scene *scene, uint32_t texture, char *uriOut, uint32_t uriCapacity, uint32_t *linear);
/* stbi_load(path, 4 channels) as Texture2D::init does (Texture.cpp:380-390). `rgbaOut` may be NULL to query the size. */
int prth_decode_image(const char *path, uint32_t *width, uint32_t *height, uint8_t *rgbaOut, uint64_t capacity);
int64_t prth_source_write_time(const prth_scene *scene);

/* per frame: World::updateAnimations(timeS) + updateScene + updateBuffers + updateTlasInstances */
int prth_update(prth_scene *scene, float timeS);
int prth_get_instances(const prth_scene *scene, prt_model_instance_transforms *transforms, float *scales,
                       prt_draw_instance *drawInstances, prt_tlas_instance *tlasInstances);
int prth_get_lights(const prth_scene *scene, prt_directional_light *directional, prt_point_light *points, prt_spot_light *spots);
int prth_get_camera(const prth_scene *scene, prth_camera *out); /* the node camera when hasCamera, else the defaults */

#ifdef __cplusplus
}
#endif
#endif
