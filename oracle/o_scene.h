// oracle/o_scene.h -- TEST INFRASTRUCTURE (CPU oracle).
// Restatement of res/shader/scene/{geometry,instances,materials,lighting}.glsl and the texture
// fetch the Vulkan sampler would do (BC7 / RGBA8 UNORM, bilinear, LOD 0).
#pragma once
#include "o_shading.h"
#include "oracle.h"
#include <vector>

namespace orc {

struct GeometryMetadata // ss/scene/geometry_metadata.h:13-33
{
    uint32_t bufferIndex, indicesOffset, positionsOffset, normalsOffset, tangentsOffset, texCoord0sOffset,
        meshletsOffset, meshletBoundsOffset, meshletVerticesOffset, meshletTrianglesByteOffset, usesShortIndices;
};
struct MeshInfo { uint32_t vertexCount, indexCount, meshletCount, materialIndex; }; // Mesh.hpp:17-23
struct MaterialData // ss/scene/material_data.h:53-68
{
    float baseColorFactor[4];
    float metallicFactor, roughnessFactor, alphaCutoff;
    uint32_t alphaMode;
    uint32_t baseColorTextureSampler, metallicRoughnessTextureSampler, normalTextureSampler, pad;
};
struct DrawInstance { uint32_t modelInstanceIndex, meshIndex, materialIndex; };
struct ModelInstanceTransforms { float modelToWorld[3][4]; float normalToWorld[3][4]; };
struct DirectionalLight { float irradiance[4]; float direction[4]; };
struct PointLight { float radianceAndRadius[4]; float position[4]; };
struct SpotLight { float radianceAndAngleScale[4]; float positionAndAngleOffset[4]; float direction[4]; };
struct TlasInstance { float transform[3][4]; uint32_t customIndexAndMask; uint32_t sbtAndFlags; uint64_t blasReference; };
struct SubModel { uint32_t meshIndex, materialIndex; };
struct SamplerDesc { uint32_t magFilter, minFilter, addressU, addressV; };
struct ReferencePC
{
    uint32_t drawType, flags, frameIndex;
    float apertureDiameter, focusDistance, focalLength;
    uint32_t rouletteStartBounce, maxBounces;
};
static_assert(sizeof(GeometryMetadata) == 44 && sizeof(MaterialData) == 48 && sizeof(DrawInstance) == 12 &&
              sizeof(ModelInstanceTransforms) == 96 && sizeof(TlasInstance) == 64 && sizeof(ReferencePC) == 32,
              "data contract");

enum : uint32_t { AlphaModeOpaque = 0, AlphaModeMask = 1, AlphaModeBlend = 2 };

// ---- BC7 (modes 4,5,6) ----------------------------------------------------------------------
struct BitReader
{
    const uint8_t *p;
    uint32_t pos = 0;
    uint32_t get(uint32_t n)
    {
        uint32_t v = 0;
        for (uint32_t i = 0; i < n; ++i, ++pos)
            v |= (uint32_t)((p[pos >> 3] >> (pos & 7)) & 1u) << i;
        return v;
    }
};

inline int bc7DecodeBlock(const uint8_t *block, uint8_t *out /*16 x RGBA8*/)
{
    static const uint8_t w2[4] = {0, 21, 43, 64};
    static const uint8_t w3[8] = {0, 9, 18, 27, 37, 46, 55, 64};
    static const uint8_t w4[16] = {0, 4, 9, 13, 17, 21, 26, 30, 34, 38, 43, 47, 51, 55, 60, 64};
    auto lerp = [](uint32_t a, uint32_t b, uint32_t w) { return (uint8_t)(((64 - w) * a + w * b + 32) >> 6); };
    BitReader br{block};
    uint32_t mode = 0;
    while (mode < 8 && br.get(1) == 0) mode++;
    if (mode == 6)
    {
        uint32_t e[4][2];
        for (int c = 0; c < 4; ++c)
            for (int k = 0; k < 2; ++k) e[c][k] = br.get(7);
        const uint32_t p0 = br.get(1), p1 = br.get(1);
        for (int c = 0; c < 4; ++c)
        {
            e[c][0] = (e[c][0] << 1) | p0;
            e[c][1] = (e[c][1] << 1) | p1;
        }
        for (int i = 0; i < 16; ++i)
        {
            const uint32_t idx = br.get(i == 0 ? 3 : 4);
            for (int c = 0; c < 4; ++c) out[i * 4 + c] = lerp(e[c][0], e[c][1], w4[idx]);
        }
        return 0;
    }
    if (mode == 5 || mode == 4)
    {
        const uint32_t rot = br.get(2);
        const uint32_t idxMode = mode == 4 ? br.get(1) : 0;
        const uint32_t cb = mode == 4 ? 5 : 7, ab = mode == 4 ? 6 : 8;
        uint32_t e[4][2];
        for (int c = 0; c < 3; ++c)
            for (int k = 0; k < 2; ++k)
            {
                const uint32_t v = br.get(cb);
                e[c][k] = (v << (8 - cb)) | (v >> (2 * cb - 8));
            }
        for (int k = 0; k < 2; ++k)
        {
            const uint32_t v = br.get(ab);
            e[3][k] = ab == 8 ? v : ((v << 2) | (v >> 4));
        }
        uint32_t i0[16], i1[16];
        const uint32_t b0 = 2, b1 = mode == 4 ? 3 : 2;
        for (int i = 0; i < 16; ++i) i0[i] = br.get(i == 0 ? b0 - 1 : b0);
        for (int i = 0; i < 16; ++i) i1[i] = br.get(i == 0 ? b1 - 1 : b1);
        for (int i = 0; i < 16; ++i)
        {
            uint32_t cw, aw;
            if (mode == 5) { cw = w2[i0[i]]; aw = w2[i1[i]]; }
            else if (idxMode == 0) { cw = w2[i0[i]]; aw = w3[i1[i]]; }
            else { cw = w3[i1[i]]; aw = w2[i0[i]]; }
            uint8_t px[4];
            for (int c = 0; c < 3; ++c) px[c] = lerp(e[c][0], e[c][1], cw);
            px[3] = lerp(e[3][0], e[3][1], aw);
            if (rot == 1) std::swap(px[3], px[0]);
            else if (rot == 2) std::swap(px[3], px[1]);
            else if (rot == 3) std::swap(px[3], px[2]);
            for (int c = 0; c < 4; ++c) out[i * 4 + c] = px[c];
        }
        return 0;
    }
    for (int i = 0; i < 64; ++i) out[i] = 0;
    return -1;
}

// ---- scene view -----------------------------------------------------------------------------
struct Scene
{
    orc_scene s;
    std::vector<std::vector<uint8_t>> decoded; // level-0 RGBA8 per texture (decoded once)

    const GeometryMetadata &metadata(uint32_t i) const { return ((const GeometryMetadata *)s.metadatas)[i]; }
    const MeshInfo &meshInfo(uint32_t i) const { return ((const MeshInfo *)s.meshInfos)[i]; }
    const MaterialData &material(uint32_t i) const { return ((const MaterialData *)s.materials)[i]; }
    const DrawInstance &drawInstance(uint32_t i) const { return ((const DrawInstance *)s.drawInstances)[i]; }
    const ModelInstanceTransforms &transforms(uint32_t i) const { return ((const ModelInstanceTransforms *)s.transforms)[i]; }
    const TlasInstance &tlasInstance(uint32_t i) const { return ((const TlasInstance *)s.tlasInstances)[i]; }
    const SubModel &subModel(uint32_t i) const { return ((const SubModel *)s.subModels)[i]; }
    const CameraUniforms &camera() const { return *(const CameraUniforms *)s.camera; }
    const uint32_t *bufU32(uint32_t b) const { return s.geometryBuffers[b]; }

    void decodeTextures()
    {
        decoded.resize(s.textureCount);
        for (uint32_t t = 0; t < s.textureCount; ++t)
        {
            const orc_texture &tx = s.textures[t];
            if (tx.data == nullptr || tx.width == 0) continue;
            std::vector<uint8_t> &img = decoded[t];
            img.resize((size_t)tx.width * tx.height * 4);
            if (tx.format == 0)
                std::memcpy(img.data(), tx.data, img.size());
            else
            {
                const uint32_t bw = (tx.width + 3) / 4, bh = (tx.height + 3) / 4;
                const uint8_t *blocks = (const uint8_t *)tx.data;
                for (uint32_t by = 0; by < bh; ++by)
                    for (uint32_t bx = 0; bx < bw; ++bx)
                    {
                        uint8_t px[64];
                        bc7DecodeBlock(blocks + ((size_t)by * bw + bx) * 16, px);
                        for (uint32_t y = 0; y < 4 && by * 4 + y < tx.height; ++y)
                            for (uint32_t x = 0; x < 4 && bx * 4 + x < tx.width; ++x)
                                std::memcpy(&img[((size_t)(by * 4 + y) * tx.width + bx * 4 + x) * 4], &px[(y * 4 + x) * 4], 4);
                    }
            }
        }
    }
};

// ---- scene/geometry.glsl --------------------------------------------------------------------
inline uint32_t loadIndex(const Scene &sc, uint32_t bufferIndex, uint32_t bufferOffset, uint32_t index, uint32_t usesShort)
{ // geometry.glsl:51-59
    if (usesShort == 1)
        return (uint32_t)((const uint16_t *)sc.bufU32(bufferIndex))[bufferOffset + index];
    return sc.bufU32(bufferIndex)[bufferOffset + index];
}

inline vec2 unpackHalf2x16(uint32_t p) { return {halfToFloat((uint16_t)(p & 0xFFFFu)), halfToFloat((uint16_t)(p >> 16))}; }

inline vec2 loadR16G16(const Scene &sc, uint32_t bufferIndex, uint32_t bufferOffset, uint32_t index)
{ // geometry.glsl:71-80
    if (bufferOffset == 0xFFFFFFFFu) return {0.f, 0.f};
    return unpackHalf2x16(sc.bufU32(bufferIndex)[bufferOffset + index]);
}

inline vec3 loadR16G16B16A16(const Scene &sc, uint32_t bufferIndex, uint32_t bufferOffset, uint32_t index)
{ // geometry.glsl:82-93
    if (bufferOffset == 0xFFFFFFFFu) return {0.f, 0.f, 0.f};
    const uint32_t px = sc.bufU32(bufferIndex)[bufferOffset + index * 2];
    const uint32_t py = sc.bufU32(bufferIndex)[bufferOffset + index * 2 + 1];
    const vec2 a = unpackHalf2x16(px), b = unpackHalf2x16(py);
    return {a.x, a.y, b.x};
}

inline vec3 unpackSnormR10G10B10(uint32_t packed) // geometry.glsl:95-103
{
    const int32_t x = (int32_t)(packed << 22) >> 22;
    const int32_t y = (int32_t)(packed << 12) >> 22;
    const int32_t z = (int32_t)(packed << 2) >> 22;
    const vec3 v = vmax(vec3{(float)x, (float)y, (float)z} / 511.f, {-1.f, -1.f, -1.f});
    return normalize(v);
}

inline vec3 loadR10G10B10Snorm(const Scene &sc, uint32_t bufferIndex, uint32_t bufferOffset, uint32_t index)
{ // geometry.glsl:105-114
    if (bufferOffset == 0xFFFFFFFFu) return {0.f, 0.f, 0.f};
    return unpackSnormR10G10B10(sc.bufU32(bufferIndex)[bufferOffset + index]);
}

inline vec4 loadTangentWithSign(const Scene &sc, uint32_t bufferIndex, uint32_t bufferOffset, uint32_t index)
{ // geometry.glsl:116-127
    if (bufferOffset == 0xFFFFFFFFu) return {0.f, 0.f, 0.f, 0.f};
    const uint32_t packed = sc.bufU32(bufferIndex)[bufferOffset + index];
    const vec3 t = unpackSnormR10G10B10(packed);
    return {t.x, t.y, t.z, (float)((int32_t)packed >> 30)};
}

struct Vertex { vec3 Position; vec3 Normal; vec4 Tangent; vec2 TexCoord0; }; // vertex.glsl:4-10

inline Vertex loadVertex(const Scene &sc, const GeometryMetadata &m, uint32_t index) // geometry.glsl:220-234
{
    Vertex r;
    r.Position = loadR16G16B16A16(sc, m.bufferIndex, m.positionsOffset, index);
    r.Normal = loadR10G10B10Snorm(sc, m.bufferIndex, m.normalsOffset, index);
    r.Tangent = loadTangentWithSign(sc, m.bufferIndex, m.tangentsOffset, index);
    r.TexCoord0 = loadR16G16(sc, m.bufferIndex, m.texCoord0sOffset, index);
    return r;
}

inline Vertex loadVertexThroughIndexBuffer(const Scene &sc, const GeometryMetadata &m, uint32_t index)
{ // geometry.glsl:236-244
    return loadVertex(sc, m, loadIndex(sc, m.bufferIndex, m.indicesOffset, index, m.usesShortIndices));
}

inline vec2 loadUV(const Scene &sc, uint32_t meshIndex, uint32_t index) // geometry.glsl:246-256
{
    const GeometryMetadata &m = sc.metadata(meshIndex);
    const uint32_t vi = loadIndex(sc, m.bufferIndex, m.indicesOffset, index, m.usesShortIndices);
    return loadR16G16(sc, m.bufferIndex, m.texCoord0sOffset, vi);
}

// position only, for the acceleration structure (World.cpp:617-640: fp16x4, stride 8)
inline vec3 loadPosition(const Scene &sc, const GeometryMetadata &m, uint32_t vertexIndex)
{
    return loadR16G16B16A16(sc, m.bufferIndex, m.positionsOffset, vertexIndex);
}

inline vec2 baryInterpolate(vec2 v0, vec2 v1, vec2 v2, float a, float b, float c) { return v0 * a + v1 * b + v2 * c; }
inline vec3 baryInterpolate(vec3 v0, vec3 v1, vec3 v2, float a, float b, float c) { return v0 * a + v1 * b + v2 * c; }
inline vec4 baryInterpolate(vec4 v0, vec4 v1, vec4 v2, float a, float b, float c) { return v0 * a + v1 * b + v2 * c; }

inline Vertex interpolate(const Vertex &v0, const Vertex &v1, const Vertex &v2, vec2 bary) // geometry.glsl:271-286
{
    const float a = 1 - bary.x - bary.y, b = bary.x, c = bary.y;
    Vertex r;
    r.Position = baryInterpolate(v0.Position, v1.Position, v2.Position, a, b, c);
    r.Normal = baryInterpolate(v0.Normal, v1.Normal, v2.Normal, a, b, c);
    r.Tangent = baryInterpolate(v0.Tangent, v1.Tangent, v2.Tangent, a, b, c);
    r.TexCoord0 = baryInterpolate(v0.TexCoord0, v1.TexCoord0, v2.TexCoord0, a, b, c);
    return r;
}
inline vec2 interpolate(vec2 v0, vec2 v1, vec2 v2, vec2 bary) // geometry.glsl:288-295
{
    const float a = 1 - bary.x - bary.y, b = bary.x, c = bary.y;
    return baryInterpolate(v0, v1, v2, a, b, c);
}

// ---- scene/instances.glsl:36-53 ---------------------------------------------------------------
// GLSL `vec * mat`: result[i] = dot(vec, column i).
inline Vertex transform(const Vertex &v, const ModelInstanceTransforms &t)
{
    Vertex r;
    const vec4 p = {v.Position.x, v.Position.y, v.Position.z, 1.0f};
    auto col4 = [](const float (*m)[4], int c) { return vec4{m[c][0], m[c][1], m[c][2], m[c][3]}; };
    auto col3 = [](const float (*m)[4], int c) { return vec3{m[c][0], m[c][1], m[c][2]}; };
    r.Position = {dot(p, col4(t.modelToWorld, 0)), dot(p, col4(t.modelToWorld, 1)), dot(p, col4(t.modelToWorld, 2))};
    r.Normal = normalize(vec3{dot(v.Normal, col3(t.normalToWorld, 0)), dot(v.Normal, col3(t.normalToWorld, 1)),
                              dot(v.Normal, col3(t.normalToWorld, 2))});
    if (v.Tangent.w != 0)
    {
        const vec3 tg = xyz(v.Tangent);
        const vec3 tw = normalize(vec3{dot(tg, col3(t.modelToWorld, 0)), dot(tg, col3(t.modelToWorld, 1)),
                                       dot(tg, col3(t.modelToWorld, 2))});
        r.Tangent = {tw.x, tw.y, tw.z, v.Tangent.w};
    }
    else
        r.Tangent = v.Tangent;
    r.TexCoord0 = v.TexCoord0;
    return r;
}

} // namespace orc
