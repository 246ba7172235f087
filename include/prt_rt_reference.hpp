// prt_rt_reference.hpp -- header-only C++ host binding over prt.h with the shape of prosper's own classes, for a host that
// "stays C++" (BASELINE.json north_star): render::RtReference (src/render/RtReference.hpp:19-107), the part of scene::World
// the pass consumes (src/scene/World.hpp:49, 64-65; World.cpp:468-575) and scene::Camera's uniform block
// (src/scene/Camera.cpp:62-204). No Vulkan types: what was a vk::CommandBuffer is the context's CUDA stream, what was an
// ImageHandle is a host fp16 image or a device pointer. Errors throw std::runtime_error(prt_last_error()) the way
// RtReference::init does (RtRef.cpp:112-113); programmer errors the reference WHEELS_ASSERTs come back as the same exception.
// tests/cpp/test_abi.cpp drives a render through this header without Python.
#ifndef PRT_RT_REFERENCE_HPP
#define PRT_RT_REFERENCE_HPP

#include <cmath>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

extern "C" {
#include "prt.h"
}

namespace prt_host
{

inline void check(prt_status st)
{
    if (st != PRT_OK) throw std::runtime_error(prt_last_error());
}

enum class DrawType : uint32_t // scene::DrawType (src/scene/DrawType.hpp): the values main.rgen:181-193 switches on
{
    Default = 0,
};

struct CameraTransform // Camera.hpp:22-27
{
    float eye[3]{1.f, 0.5f, 1.f};
    float target[3]{0.f, 0.f, 0.f};
    float up[3]{0.f, 1.f, 0.f};
};

struct CameraParameters // Camera.hpp:40-48
{
    float fov{59.f * 3.14159265358979323846f / 180.f};
    float zN{0.1f};
    float zF{100.f};
    float apertureDiameter{0.00001f};
    float focusDistance{1.f};
    float focalLength{0.f};
};

// scene::Camera: lookAt + perspective -> CameraUniforms (Camera.cpp:82-204, 366-395). Column-major 4x4 like glm.
class Camera
{
  public:
    void init(uint32_t width, uint32_t height)
    {
        m_width = width;
        m_height = height;
        update();
    }
    void lookAt(const CameraTransform &t)
    {
        m_transform = t;
        update();
    }
    void setParameters(const CameraParameters &p)
    {
        m_parameters = p;
        update();
    }
    [[nodiscard]] const prt_camera_uniforms &uniforms() const { return m_uniforms; }
    [[nodiscard]] const CameraParameters &parameters() const { return m_parameters; }
    [[nodiscard]] bool changedThisFrame() const { return m_changed; }
    void endFrame() { m_changed = false; }

  private:
    static void mul(const float a[4][4], const float b[4][4], float out[4][4]) // out = a * b, column-major [col][row]
    {
        float r[4][4];
        for (int c = 0; c < 4; ++c)
            for (int row = 0; row < 4; ++row)
            {
                float s = 0.f;
                for (int k = 0; k < 4; ++k) s += a[k][row] * b[c][k];
                r[c][row] = s;
            }
        std::memcpy(out, r, sizeof(r));
    }
    static void invertAffineOrGeneral(const float m[4][4], float out[4][4])
    {
        // general 4x4 inverse by cofactors (double precision inside)
        double a[16], inv[16];
        for (int c = 0; c < 4; ++c)
            for (int r = 0; r < 4; ++r) a[c * 4 + r] = m[c][r];
        inv[0] = a[5] * a[10] * a[15] - a[5] * a[11] * a[14] - a[9] * a[6] * a[15] + a[9] * a[7] * a[14] + a[13] * a[6] * a[11] - a[13] * a[7] * a[10];
        inv[4] = -a[4] * a[10] * a[15] + a[4] * a[11] * a[14] + a[8] * a[6] * a[15] - a[8] * a[7] * a[14] - a[12] * a[6] * a[11] + a[12] * a[7] * a[10];
        inv[8] = a[4] * a[9] * a[15] - a[4] * a[11] * a[13] - a[8] * a[5] * a[15] + a[8] * a[7] * a[13] + a[12] * a[5] * a[11] - a[12] * a[7] * a[9];
        inv[12] = -a[4] * a[9] * a[14] + a[4] * a[10] * a[13] + a[8] * a[5] * a[14] - a[8] * a[6] * a[13] - a[12] * a[5] * a[10] + a[12] * a[6] * a[9];
        inv[1] = -a[1] * a[10] * a[15] + a[1] * a[11] * a[14] + a[9] * a[2] * a[15] - a[9] * a[3] * a[14] - a[13] * a[2] * a[11] + a[13] * a[3] * a[10];
        inv[5] = a[0] * a[10] * a[15] - a[0] * a[11] * a[14] - a[8] * a[2] * a[15] + a[8] * a[3] * a[14] + a[12] * a[2] * a[11] - a[12] * a[3] * a[10];
        inv[9] = -a[0] * a[9] * a[15] + a[0] * a[11] * a[13] + a[8] * a[1] * a[15] - a[8] * a[3] * a[13] - a[12] * a[1] * a[11] + a[12] * a[3] * a[9];
        inv[13] = a[0] * a[9] * a[14] - a[0] * a[10] * a[13] - a[8] * a[1] * a[14] + a[8] * a[2] * a[13] + a[12] * a[1] * a[10] - a[12] * a[2] * a[9];
        inv[2] = a[1] * a[6] * a[15] - a[1] * a[7] * a[14] - a[5] * a[2] * a[15] + a[5] * a[3] * a[14] + a[13] * a[2] * a[7] - a[13] * a[3] * a[6];
        inv[6] = -a[0] * a[6] * a[15] + a[0] * a[7] * a[14] + a[4] * a[2] * a[15] - a[4] * a[3] * a[14] - a[12] * a[2] * a[7] + a[12] * a[3] * a[6];
        inv[10] = a[0] * a[5] * a[15] - a[0] * a[7] * a[13] - a[4] * a[1] * a[15] + a[4] * a[3] * a[13] + a[12] * a[1] * a[7] - a[12] * a[3] * a[5];
        inv[14] = -a[0] * a[5] * a[14] + a[0] * a[6] * a[13] + a[4] * a[1] * a[14] - a[4] * a[2] * a[13] - a[12] * a[1] * a[6] + a[12] * a[2] * a[5];
        inv[3] = -a[1] * a[6] * a[11] + a[1] * a[7] * a[10] + a[5] * a[2] * a[11] - a[5] * a[3] * a[10] - a[9] * a[2] * a[7] + a[9] * a[3] * a[6];
        inv[7] = a[0] * a[6] * a[11] - a[0] * a[7] * a[10] - a[4] * a[2] * a[11] + a[4] * a[3] * a[10] + a[8] * a[2] * a[7] - a[8] * a[3] * a[6];
        inv[11] = -a[0] * a[5] * a[11] + a[0] * a[7] * a[9] + a[4] * a[1] * a[11] - a[4] * a[3] * a[9] - a[8] * a[1] * a[7] + a[8] * a[3] * a[5];
        inv[15] = a[0] * a[5] * a[10] - a[0] * a[6] * a[9] - a[4] * a[1] * a[10] + a[4] * a[2] * a[9] + a[8] * a[1] * a[6] - a[8] * a[2] * a[5];
        const double det = a[0] * inv[0] + a[1] * inv[4] + a[2] * inv[8] + a[3] * inv[12];
        for (int c = 0; c < 4; ++c)
            for (int r = 0; r < 4; ++r) out[c][r] = (float)(inv[c * 4 + r] / det);
    }
    static void normalize3(float v[3])
    {
        const float l = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
        v[0] /= l; v[1] /= l; v[2] /= l;
    }
    static void cross3(const float a[3], const float b[3], float o[3])
    {
        o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
    }
    void update()
    {
        if (m_width == 0 || m_height == 0) return;
        prt_camera_uniforms u{};
        // updateWorldToCamera (Camera.cpp:366-395)
        const float *eye = m_transform.eye;
        float fwd[3] = {m_transform.target[0] - eye[0], m_transform.target[1] - eye[1], m_transform.target[2] - eye[2]};
        normalize3(fwd);
        const float z[3] = {-fwd[0], -fwd[1], -fwd[2]};
        float right[3], newUp[3];
        cross3(m_transform.up, z, right);
        normalize3(right);
        cross3(z, right, newUp);
        normalize3(newUp);
        const float w2c[4][4] = {{right[0], newUp[0], z[0], 0.f}, {right[1], newUp[1], z[1], 0.f}, {right[2], newUp[2], z[2], 0.f},
                                 {-(right[0] * eye[0] + right[1] * eye[1] + right[2] * eye[2]),
                                  -(newUp[0] * eye[0] + newUp[1] * eye[1] + newUp[2] * eye[2]), -(z[0] * eye[0] + z[1] * eye[1] + z[2] * eye[2]), 1.f}};
        std::memcpy(u.worldToCamera, w2c, sizeof(w2c));
        invertAffineOrGeneral(w2c, u.cameraToWorld);
        // perspective (Camera.cpp:105-160): reverse-Z, flipped y, z in [0, 1]
        const float ar = (float)m_width / (float)m_height;
        const float zN = m_parameters.zF, zF = m_parameters.zN; // "Swap near and far for the magical properties of reverse-z"
        const float tf = 1.f / std::tan(m_parameters.fov * 0.5f);
        const float flip[4][4] = {{1.f, 0.f, 0.f, 0.f}, {0.f, -1.f, 0.f, 0.f}, {0.f, 0.f, 0.5f, 0.f}, {0.f, 0.f, 0.5f, 1.f}};
        const float proj[4][4] = {{tf / ar, 0.f, 0.f, 0.f}, {0.f, tf, 0.f, 0.f}, {0.f, 0.f, (zF + zN) / (zN - zF), -1.f},
                                  {0.f, 0.f, 2 * zF * zN / (zN - zF), 0.f}};
        mul(flip, proj, u.cameraToClip);
        float w2clip[4][4];
        mul(u.cameraToClip, w2c, w2clip);
        invertAffineOrGeneral(w2clip, u.clipToWorld);
        std::memcpy(u.previousWorldToCamera, u.worldToCamera, sizeof(u.worldToCamera));
        std::memcpy(u.previousCameraToClip, u.cameraToClip, sizeof(u.cameraToClip));
        u.eye[0] = eye[0]; u.eye[1] = eye[1]; u.eye[2] = eye[2]; u.eye[3] = 1.f;
        u.resolution[0] = m_width; u.resolution[1] = m_height;
        u.near_ = m_parameters.zN; u.far_ = m_parameters.zF;
        u.maxViewScale = 1.f;
        // focalLength = sensorHeight * tf * 0.5 (Camera.cpp:150-152), sensor width 36 mm
        m_parameters.focalLength = 0.036f / ar * tf * 0.5f;
        m_uniforms = u;
        m_changed = true;
    }

    uint32_t m_width{0}, m_height{0};
    CameraTransform m_transform;
    CameraParameters m_parameters;
    prt_camera_uniforms m_uniforms{};
    bool m_changed{true};
};

// The part of scene::World the pass reads, as arrays the caller fills (prosper's WorldData holds the same arrays).
class World
{
  public:
    explicit World(const prt_config &cfg) { check(prt_create(&cfg, &m_ctx)); }
    ~World() { prt_destroy(m_ctx); }
    World(const World &) = delete;
    World &operator=(const World &) = delete;

    [[nodiscard]] prt_ctx *ctx() const { return m_ctx; }

    // load-time data (WorldData uploads)
    void setGeometryBuffer(uint32_t index, const void *data, uint64_t bytes) { check(prt_set_geometry_buffer(m_ctx, index, data, bytes)); }
    void setMeshes(const std::vector<prt_geometry_metadata> &md, const std::vector<prt_mesh_info> &infos)
    {
        check(prt_set_meshes(m_ctx, md.data(), infos.data(), (uint32_t)md.size()));
    }
    void setMaterials(const std::vector<prt_material_data> &m) { check(prt_set_materials(m_ctx, m.data(), (uint32_t)m.size())); }
    void setSamplers(const std::vector<prt_sampler_desc> &s) { check(prt_set_samplers(m_ctx, s.data(), (uint32_t)s.size())); }
    void setTexture(uint32_t index, const prt_texture_desc &d) { check(prt_set_texture(m_ctx, index, &d)); }
    void setSkyCubemap(uint32_t faceSize, const void *rgba16f) { check(prt_set_sky_cubemap(m_ctx, faceSize, rgba16f)); }
    void setModels(const std::vector<uint32_t> &offsets, const std::vector<prt_sub_model> &subModels)
    {
        check(prt_set_models(m_ctx, offsets.data(), subModels.data(), (uint32_t)offsets.size() - 1));
        m_blasReferences.assign(offsets.size() - 1, 0);
    }

    // per-frame data, filled by the caller's updateScene / updateBuffers
    std::vector<prt_draw_instance> drawInstances;
    std::vector<prt_model_instance_transforms> transforms;
    std::vector<prt_tlas_instance> tlasInstances;
    prt_directional_light directionalLight{{2.f, 2.f, 2.f, 0.f}, {-1.f, -1.f, -1.f, 0.f}}; // lights.h:9,18-19
    std::vector<prt_point_light> pointLights;
    std::vector<prt_spot_light> spotLights;

    // World::updateBuffers (World.cpp:468-536) + Camera::updateBuffer: ring uploads
    void updateBuffers(const Camera &cam)
    {
        check(prt_set_draw_instances(m_ctx, drawInstances.data(), (uint32_t)drawInstances.size()));
        check(prt_set_instance_transforms(m_ctx, transforms.data(), (uint32_t)transforms.size()));
        check(prt_set_lights(m_ctx, &directionalLight, pointLights.data(), (uint32_t)pointLights.size(), spotLights.data(),
                             (uint32_t)spotLights.size()));
        check(prt_set_camera(m_ctx, &cam.uniforms()));
    }
    // World::buildAccelerationStructures (World.cpp:538-575): pending BLAS builds, then the TLAS. Returns "new BLAS added".
    [[nodiscard]] bool buildAccelerationStructures()
    {
        bool blasAdded = false;
        for (uint32_t m = 0; m < m_blasReferences.size(); ++m)
            if (m_blasReferences[m] == 0)
            {
                check(prt_build_blas(m_ctx, m, &m_blasReferences[m]));
                blasAdded = true;
            }
        check(prt_build_tlas(m_ctx, tlasInstances.data(), (uint32_t)tlasInstances.size()));
        return blasAdded;
    }
    [[nodiscard]] uint64_t blasReference(uint32_t model) const { return m_blasReferences.at(model); }

  private:
    prt_ctx *m_ctx{nullptr};
    std::vector<uint64_t> m_blasReferences;
};

// render::RtReference (RtReference.hpp:19-107)
class RtReference
{
  public:
    static const uint32_t sMaxBounces = 6; // RtReference.hpp:22

    RtReference() noexcept = default;
    RtReference(const RtReference &) = delete;
    RtReference &operator=(const RtReference &) = delete;

    void init(World &world)
    {
        if (m_initialized) throw std::runtime_error("RtReference::init: already initialized");
        m_world = &world;
        m_initialized = true;
    }
    // nothing to compile at run time: the kernels are in the library
    bool recompileShaders() { return true; }
    // the ImGui panel's state (RtRef.cpp:147-159) as setters
    void drawUi(bool accumulate, bool clampIndirect, uint32_t rouletteStartBounce, uint32_t maxBounces)
    {
        m_accumulationDirty |= accumulate != m_accumulate || clampIndirect != m_clampIndirect || rouletteStartBounce != m_rouletteStartBounce ||
                               maxBounces != m_maxBounces;
        m_accumulate = accumulate;
        m_clampIndirect = clampIndirect;
        m_maxBounces = maxBounces < sMaxBounces ? maxBounces : sMaxBounces;
        m_rouletteStartBounce = rouletteStartBounce < m_maxBounces ? rouletteStartBounce : m_maxBounces;
    }

    struct Options // RtReference.hpp:44-50
    {
        bool depthOfField{false};
        bool ibl{false};
        bool colorDirty{false};
        DrawType drawType{DrawType::Default};
    };
    struct Output // RtReference.hpp:51-54: the fp16 RGBA illumination image, on the host
    {
        uint32_t width{0}, height{0};
        std::vector<uint16_t> illumination; // width * height * 4 halves
    };

    // RtReference::record (RtRef.cpp:161-383); renderArea.offset must be 0 (RtRef.cpp:327)
    [[nodiscard]] Output record(const Camera &cam, uint32_t width, uint32_t height, const Options &options)
    {
        if (!m_initialized) throw std::runtime_error("RtReference::record: not initialized");
        m_frameIndex = (m_frameIndex + 1) % PRT_FRAME_PERIOD; // RtRef.cpp:170
        const CameraParameters &cp = cam.parameters();
        prt_reference_pc pc{};
        pc.drawType = (uint32_t)options.drawType;
        const bool skipHistory = cam.changedThisFrame() || options.colorDirty || m_accumulationDirty; // RtRef.cpp:282-284
        pc.flags = (skipHistory ? PRT_FLAG_SKIP_HISTORY : 0u) | (m_accumulate ? PRT_FLAG_ACCUMULATE : 0u) | (options.ibl ? PRT_FLAG_IBL : 0u) |
                   (options.depthOfField ? PRT_FLAG_DEPTH_OF_FIELD : 0u) | (m_clampIndirect ? PRT_FLAG_CLAMP_INDIRECT : 0u); // pcFlags, RtRef.cpp:68-88
        pc.frameIndex = m_frameIndex;
        pc.apertureDiameter = cp.apertureDiameter;
        pc.focusDistance = cp.focusDistance;
        pc.focalLength = cp.focalLength;
        pc.rouletteStartBounce = m_rouletteStartBounce;
        pc.maxBounces = m_maxBounces;
        check(prt_record(m_world->ctx(), &pc, width, height, 1));
        m_accumulationDirty = false;
        Output out;
        out.width = width;
        out.height = height;
        out.illumination.resize((size_t)width * height * 4);
        check(prt_read_illumination_f16(m_world->ctx(), out.illumination.data(), out.illumination.size() * 2));
        return out;
    }
    void releasePreserved() // RtRef.cpp:385-391
    {
        if (m_world) check(prt_release_preserved(m_world->ctx()));
    }

  private:
    World *m_world{nullptr};
    bool m_initialized{false};
    bool m_accumulationDirty{true};
    bool m_accumulate{true};
    bool m_clampIndirect{true};
    uint32_t m_frameIndex{0};
    uint32_t m_rouletteStartBounce{3};
    uint32_t m_maxBounces{sMaxBounces};
};

} // namespace prt_host

#endif // PRT_RT_REFERENCE_HPP
