// gltf_host.cpp -- libprt_host.so: glTF ingestion + per-frame scene update for the RtReference path (include/prt_host.h).
//
// Stand-alone restatement of the part of prosper's scene::World that feeds the path tracer, written against the
// reference's own vendored third-party libraries compiled where they lie under /root/reference/ext (cgltf, glm,
// meshoptimizer, mikktspace, stb; see host/Makefile) so that parsing, tangent generation, index optimisation and the
// vertex packing are the very code the reference runs. Each function cites the reference function it follows.
// No Vulkan, no allocator framework, no worker thread: one synchronous load, then updates on the caller's thread.
#include "../include/prt_host.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <optional>
#include <stdexcept>
#include <string>
#include <vector>

#define GLM_FORCE_XYZW_ONLY // the reference's compile definitions (CMakeLists.txt:146-147)
#define GLM_ENABLE_EXPERIMENTAL
#include <glm/glm.hpp>
#include <glm/gtc/matrix_access.hpp>
#include <glm/gtc/matrix_transform.hpp>
#include <glm/gtc/packing.hpp>
#include <glm/gtc/quaternion.hpp>
#include <glm/gtc/type_ptr.hpp>
#include <glm/gtx/matrix_decompose.hpp>

#define CGLTF_IMPLEMENTATION
#include <cgltf.h>
#define STB_IMAGE_IMPLEMENTATION
#define STBI_ONLY_PNG
#define STBI_ONLY_JPEG
#include <stb_image.h>
#include <meshoptimizer.h>
extern "C" {
#include <mikktspace.h>
}

using namespace glm;

namespace
{

thread_local std::string g_error;
int fail(const std::string &what)
{
    g_error = what;
    return 1;
}
#define HOST_ASSERT(cond)                                                                                                      \
    do                                                                                                                         \
    {                                                                                                                          \
        if (!(cond)) throw std::runtime_error(std::string("assertion failed: ") + #cond);                                      \
    } while (0)

const uint64_t sMeshCacheMagic = 0x4853'4D52'5053'5250; // PRSPRMSH (DeferredLoadingContext.cpp:24)
const uint32_t sMeshCacheVersion = 4;                    // DeferredLoadingContext.cpp:27
const uint32_t sGeometryBufferSize = 64u * 1024u * 1024u; // DeferredLoadingContext.cpp:22
const float sConeWeight = 0.5f;                          // DeferredLoadingContext.cpp:30
const size_t sMaxMsVertices = 64;                        // utils/Utils.hpp:18
const size_t sMaxMsTriangles = 124;                      // utils/Utils.hpp:19

// ---- mesh pipeline (DeferredLoadingContext.cpp) ------------------------------------------------------------------
struct MeshletBounds // DeferredLoadingContext.cpp:50-58
{
    vec3 center{};
    float radius{0.f};
    i8vec3 coneAxisS8{};
    int8_t coneCutoffS8{0};
};
static_assert(sizeof(MeshletBounds) == 5 * sizeof(uint32_t), "MeshletBounds");
static_assert(sizeof(meshopt_Meshlet) == 4 * sizeof(uint32_t), "meshopt_Meshlet");

struct MeshData // DeferredLoadingContext.cpp:69-80
{
    std::vector<uint32_t> indices;
    std::vector<vec3> positions;
    std::vector<vec3> normals;
    std::vector<vec4> tangents;
    std::vector<vec2> texCoord0s;
    std::vector<meshopt_Meshlet> meshlets;
    std::vector<MeshletBounds> meshletBounds;
    std::vector<uint32_t> meshletVertices;
    std::vector<uint8_t> meshletTriangles;
};

struct PackedMeshData // DeferredLoadingContext.cpp:82-104
{
    std::vector<uint32_t> indices;
    std::vector<uint64_t> positions;
    std::vector<uint32_t> normals;
    std::vector<uint32_t> tangents;
    std::vector<uint32_t> texCoord0s;
    std::vector<meshopt_Meshlet> meshlets;
    std::vector<MeshletBounds> meshletBounds;
    std::vector<uint32_t> meshletVertices;
    std::vector<uint8_t> meshletTriangles;
};

struct InputGeometryMetadata // DeferredLoadingContext.hpp
{
    const cgltf_accessor *indices{nullptr};
    const cgltf_accessor *positions{nullptr};
    const cgltf_accessor *normals{nullptr};
    const cgltf_accessor *tangents{nullptr};
    const cgltf_accessor *texCoord0s{nullptr};
    uint32_t sourceMeshIndex{0};
    uint32_t sourcePrimitiveIndex{0};
};

struct MeshCacheHeader // DeferredLoadingContext.hpp:60-78
{
    int64_t sourceWriteTime{0};
    uint32_t indexCount{0};
    uint32_t vertexCount{0};
    uint32_t meshletCount{0};
    uint32_t positionsOffset{0xFFFF'FFFF};
    uint32_t normalsOffset{0xFFFF'FFFF};
    uint32_t tangentsOffset{0xFFFF'FFFF};
    uint32_t texCoord0sOffset{0xFFFF'FFFF};
    uint32_t meshletsOffset{0xFFFF'FFFF};
    uint32_t meshletBoundsOffset{0xFFFF'FFFF};
    uint32_t meshletVerticesOffset{0xFFFF'FFFF};
    uint32_t meshletTrianglesByteOffset{0xFFFF'FFFF};
    uint32_t usesShortIndices{0};
    uint32_t blobByteCount{0};
};

template <typename T> void remapVertexAttribute(std::vector<T> &src, const std::vector<uint32_t> &remapIndices, size_t uniqueVertexCount)
{ // DeferredLoadingContext.cpp:36-48
    std::vector<T> remapped(uniqueVertexCount);
    meshopt_remapVertexBuffer(remapped.data(), src.data(), src.size(), sizeof(T), remapIndices.data());
    src = std::move(remapped);
}

template <int N> void unpackVector(const cgltf_accessor *accessor, std::vector<vec<N, float, defaultp>> &out)
{ // DeferredLoadingContext.cpp:110-119
    HOST_ASSERT(accessor != nullptr);
    out.resize(accessor->count);
    const cgltf_size unpackedCount = cgltf_accessor_unpack_floats(accessor, &out.data()[0][0], out.size() * N);
    HOST_ASSERT(unpackedCount == out.size() * N);
}

MeshData getMeshData(const InputGeometryMetadata &metadata, const prt_mesh_info &meshInfo)
{ // DeferredLoadingContext.cpp:123-158
    MeshData ret;
    HOST_ASSERT(metadata.indices != nullptr);
    HOST_ASSERT(meshInfo.indexCount == metadata.indices->count);
    ret.indices.resize(meshInfo.indexCount);
    const cgltf_size unpackedCount =
        cgltf_accessor_unpack_indices(metadata.indices, ret.indices.data(), sizeof(ret.indices[0]), ret.indices.size());
    HOST_ASSERT(unpackedCount == meshInfo.indexCount);
    unpackVector(metadata.positions, ret.positions);
    unpackVector(metadata.normals, ret.normals);
    if (metadata.tangents != nullptr) unpackVector(metadata.tangents, ret.tangents);
    if (metadata.texCoord0s != nullptr) unpackVector(metadata.texCoord0s, ret.texCoord0s);
    return ret;
}

// mikktspace interface (DeferredLoadingContext.cpp:163-250)
int mikkTGetNumFaces(const SMikkTSpaceContext *pContext)
{
    const MeshData *meshData = static_cast<const MeshData *>(pContext->m_pUserData);
    return (int)(meshData->positions.size() / 3);
}
int mikkTGetNumVerticesOfFace(const SMikkTSpaceContext *, const int) { return 3; }
// "Go through faces in the opposite order, this seems to fix the glTF normal map handedness problem"
int mikkTVertexIndex(const int iFace, const int iVert) { return (iFace * 3) + (2 - iVert); }
void mikkTGetPosition(const SMikkTSpaceContext *pContext, float fvPosOut[], const int iFace, const int iVert)
{
    const MeshData *meshData = static_cast<const MeshData *>(pContext->m_pUserData);
    memcpy(fvPosOut, &meshData->positions[mikkTVertexIndex(iFace, iVert)].x, sizeof(vec3));
}
void mikkTGetNormal(const SMikkTSpaceContext *pContext, float fvNormOut[], const int iFace, const int iVert)
{
    const MeshData *meshData = static_cast<const MeshData *>(pContext->m_pUserData);
    memcpy(fvNormOut, &meshData->normals[mikkTVertexIndex(iFace, iVert)].x, sizeof(vec3));
}
void mikkTGetTexCoord(const SMikkTSpaceContext *pContext, float fvTexcOut[], const int iFace, const int iVert)
{
    const MeshData *meshData = static_cast<const MeshData *>(pContext->m_pUserData);
    memcpy(fvTexcOut, &meshData->texCoord0s[mikkTVertexIndex(iFace, iVert)].x, sizeof(vec2));
}
void mikkTSetTSpaceBasic(const SMikkTSpaceContext *pContext, const float fvTangent[], const float fSign, const int iFace, const int iVert)
{
    MeshData *meshData = static_cast<MeshData *>(pContext->m_pUserData);
    meshData->tangents[mikkTVertexIndex(iFace, iVert)] = vec4{fvTangent[0], fvTangent[1], fvTangent[2], fSign};
}

template <typename T> void flattenAttribute(std::vector<T> &attribute, const std::vector<uint32_t> &indices)
{ // DeferredLoadingContext.cpp:254-262
    std::vector<T> flattened;
    flattened.reserve(indices.size());
    for (const uint32_t i : indices) flattened.push_back(attribute[i]);
    attribute = std::move(flattened);
}

void generateTangents(MeshData &meshData)
{ // DeferredLoadingContext.cpp:266-338
    HOST_ASSERT(meshData.tangents.empty());
    HOST_ASSERT(meshData.positions.size() == meshData.normals.size());
    HOST_ASSERT(meshData.positions.size() == meshData.texCoord0s.size());
    const size_t flattenedVertexCount = meshData.indices.size();
    flattenAttribute(meshData.positions, meshData.indices);
    flattenAttribute(meshData.normals, meshData.indices);
    flattenAttribute(meshData.texCoord0s, meshData.indices);
    meshData.indices.clear();
    meshData.tangents.resize(flattenedVertexCount);

    SMikkTSpaceInterface sMikkTInterface{};
    sMikkTInterface.m_getNumFaces = &mikkTGetNumFaces;
    sMikkTInterface.m_getNumVerticesOfFace = mikkTGetNumVerticesOfFace;
    sMikkTInterface.m_getPosition = mikkTGetPosition;
    sMikkTInterface.m_getNormal = mikkTGetNormal;
    sMikkTInterface.m_getTexCoord = mikkTGetTexCoord;
    sMikkTInterface.m_setTSpaceBasic = mikkTSetTSpaceBasic;
    SMikkTSpaceContext mikkCtx{};
    mikkCtx.m_pInterface = &sMikkTInterface;
    mikkCtx.m_pUserData = &meshData;
    genTangSpaceDefault(&mikkCtx);

    const meshopt_Stream vertexStreams[4] = {
        {meshData.positions.data(), sizeof(vec3), sizeof(vec3)},
        {meshData.normals.data(), sizeof(vec3), sizeof(vec3)},
        {meshData.tangents.data(), sizeof(vec4), sizeof(vec4)},
        {meshData.texCoord0s.data(), sizeof(vec2), sizeof(vec2)},
    };
    std::vector<uint32_t> remapTable(flattenedVertexCount);
    const size_t uniqueVertexCount =
        meshopt_generateVertexRemapMulti(remapTable.data(), nullptr, flattenedVertexCount, flattenedVertexCount, vertexStreams, 4);
    meshData.indices.resize(flattenedVertexCount);
    meshopt_remapIndexBuffer(meshData.indices.data(), nullptr, flattenedVertexCount, remapTable.data());
    remapVertexAttribute(meshData.positions, remapTable, uniqueVertexCount);
    remapVertexAttribute(meshData.normals, remapTable, uniqueVertexCount);
    remapVertexAttribute(meshData.tangents, remapTable, uniqueVertexCount);
    remapVertexAttribute(meshData.texCoord0s, remapTable, uniqueVertexCount);
}

void optimizeMeshData(MeshData &meshData, prt_mesh_info &meshInfo)
{ // DeferredLoadingContext.cpp:340-376
    const size_t indexCount = meshData.indices.size();
    const size_t vertexCount = meshData.positions.size();
    std::vector<uint32_t> tmpIndices(meshData.indices.size());
    meshopt_optimizeVertexCache(tmpIndices.data(), meshData.indices.data(), indexCount, vertexCount);
    const float vertexCacheDegradationThreshod = 1.00f;
    meshopt_optimizeOverdraw(meshData.indices.data(), tmpIndices.data(), tmpIndices.size(), &meshData.positions.data()[0].x, vertexCount,
                             sizeof(vec3), vertexCacheDegradationThreshod);
    std::vector<uint32_t> remapIndices(vertexCount);
    const size_t uniqueVertexCount = meshopt_optimizeVertexFetchRemap(remapIndices.data(), meshData.indices.data(), indexCount, vertexCount);
    meshopt_remapIndexBuffer(tmpIndices.data(), meshData.indices.data(), indexCount, remapIndices.data());
    meshData.indices = std::move(tmpIndices);
    remapVertexAttribute(meshData.positions, remapIndices, uniqueVertexCount);
    remapVertexAttribute(meshData.normals, remapIndices, uniqueVertexCount);
    if (!meshData.tangents.empty()) remapVertexAttribute(meshData.tangents, remapIndices, uniqueVertexCount);
    if (!meshData.texCoord0s.empty()) remapVertexAttribute(meshData.texCoord0s, remapIndices, uniqueVertexCount);
    meshInfo.vertexCount = (uint32_t)uniqueVertexCount;
}

void generateMeshlets(MeshData &meshData)
{ // DeferredLoadingContext.cpp:378-440
    const size_t maxMeshlets = meshopt_buildMeshletsBound(meshData.indices.size(), sMaxMsVertices, sMaxMsTriangles);
    HOST_ASSERT(maxMeshlets > 0);
    meshData.meshlets.resize(maxMeshlets);
    meshData.meshletVertices.resize(maxMeshlets * sMaxMsVertices);
    meshData.meshletTriangles.resize(maxMeshlets * sMaxMsTriangles * 3);
    const size_t meshletCount = meshopt_buildMeshlets(meshData.meshlets.data(), meshData.meshletVertices.data(), meshData.meshletTriangles.data(),
                                                      meshData.indices.data(), meshData.indices.size(), &meshData.positions[0].x,
                                                      meshData.positions.size(), sizeof(vec3), sMaxMsVertices, sMaxMsTriangles, sConeWeight);
    HOST_ASSERT(meshletCount > 0);
    meshData.meshlets.resize(meshletCount);
    const meshopt_Meshlet &lastMeshlet = meshData.meshlets.back();
    meshData.meshletVertices.resize(lastMeshlet.vertex_offset + lastMeshlet.vertex_count);
    const uint32_t trianglesSize = (lastMeshlet.triangle_offset + lastMeshlet.triangle_count + 3u) & ~3u; // aligned_offset(.., 4)
    meshData.meshletTriangles.resize((size_t)trianglesSize * 3);
    meshData.meshletBounds.reserve(meshData.meshlets.size());
    for (const meshopt_Meshlet &meshlet : meshData.meshlets)
    {
        const meshopt_Bounds bounds =
            meshopt_computeMeshletBounds(&meshData.meshletVertices[meshlet.vertex_offset], &meshData.meshletTriangles[meshlet.triangle_offset],
                                         meshlet.triangle_count, &meshData.positions[0].x, meshData.positions.size(), sizeof(vec3));
        MeshletBounds b;
        b.center = vec3{bounds.center[0], bounds.center[1], bounds.center[2]};
        b.radius = bounds.radius;
        b.coneAxisS8 = i8vec3{bounds.cone_axis_s8[0], bounds.cone_axis_s8[1], bounds.cone_axis_s8[2]};
        b.coneCutoffS8 = bounds.cone_cutoff_s8;
        meshData.meshletBounds.push_back(b);
    }
}

PackedMeshData packMeshData(MeshData &&meshData)
{ // DeferredLoadingContext.cpp:442-490
    PackedMeshData ret;
    ret.indices = std::move(meshData.indices);
    ret.positions.reserve(meshData.positions.size());
    for (const vec3 &p : meshData.positions) ret.positions.push_back(packHalf4x16(vec4{p, 1.f}));
    ret.normals.reserve(meshData.normals.size());
    for (const vec3 &n : meshData.normals) ret.normals.push_back(packSnorm3x10_1x2(vec4{n, 0.f}));
    ret.tangents.reserve(meshData.tangents.size());
    for (const vec4 &ts : meshData.tangents) ret.tangents.push_back(packSnorm3x10_1x2(ts));
    ret.texCoord0s.reserve(meshData.texCoord0s.size());
    for (const vec2 &uv : meshData.texCoord0s) ret.texCoord0s.push_back(packHalf2x16(uv));
    ret.meshlets = std::move(meshData.meshlets);
    ret.meshletBounds = std::move(meshData.meshletBounds);
    ret.meshletVertices = std::move(meshData.meshletVertices);
    ret.meshletTriangles = std::move(meshData.meshletTriangles);
    return ret;
}

std::filesystem::path getCachePath(const std::filesystem::path &sceneDir, uint32_t meshIndex)
{ // DeferredLoadingContext.cpp:492-499
    return sceneDir / "prosper_cache" / ("cache" + std::to_string(meshIndex) + ".prosper_mesh");
}

template <typename T> void writeRaw(std::ofstream &f, const T &v) { f.write(reinterpret_cast<const char *>(&v), sizeof(T)); }
template <typename T> void readRaw(std::ifstream &f, T &v) { f.read(reinterpret_cast<char *>(&v), sizeof(T)); }
template <typename T> void writeRawSpan(std::ofstream &f, const std::vector<T> &v)
{
    if (!v.empty()) f.write(reinterpret_cast<const char *>(v.data()), (std::streamsize)(v.size() * sizeof(T)));
}

std::optional<MeshCacheHeader> readCache(const std::filesystem::path &cachePath)
{ // DeferredLoadingContext.cpp:504-558 (header only)
    std::optional<MeshCacheHeader> ret;
    if (!std::filesystem::exists(cachePath)) return ret;
    std::ifstream cacheFile{cachePath, std::ios_base::binary};
    uint64_t magic{0};
    readRaw(cacheFile, magic);
    if (magic != sMeshCacheMagic) throw std::runtime_error("Expected a valid mesh cache in file '" + cachePath.string() + "'");
    uint32_t version{0};
    readRaw(cacheFile, version);
    if (sMeshCacheVersion != version) return ret;
    MeshCacheHeader h;
    readRaw(cacheFile, h.sourceWriteTime);
    readRaw(cacheFile, h.indexCount);
    readRaw(cacheFile, h.vertexCount);
    readRaw(cacheFile, h.meshletCount);
    readRaw(cacheFile, h.positionsOffset);
    readRaw(cacheFile, h.normalsOffset);
    readRaw(cacheFile, h.tangentsOffset);
    readRaw(cacheFile, h.texCoord0sOffset);
    readRaw(cacheFile, h.meshletsOffset);
    readRaw(cacheFile, h.meshletBoundsOffset);
    readRaw(cacheFile, h.meshletVerticesOffset);
    readRaw(cacheFile, h.meshletTrianglesByteOffset);
    readRaw(cacheFile, h.usesShortIndices);
    readRaw(cacheFile, h.blobByteCount);
    ret = h;
    return ret;
}

bool cacheValid(const std::filesystem::path &cachePath, int64_t sceneWriteTime)
{ // DeferredLoadingContext.cpp:560-574
    const std::optional<MeshCacheHeader> header = readCache(cachePath);
    return header.has_value() && header->sourceWriteTime == sceneWriteTime;
}

void writeCache(const std::filesystem::path &sceneDir, int64_t sceneWriteTime, uint32_t meshIndex, PackedMeshData &&meshData,
                const prt_mesh_info &meshInfo)
{ // DeferredLoadingContext.cpp:609-802
    HOST_ASSERT(meshData.indices.size() == meshInfo.indexCount);
    HOST_ASSERT(meshData.positions.size() == meshInfo.vertexCount);
    HOST_ASSERT(meshData.normals.size() == meshInfo.vertexCount);
    HOST_ASSERT(meshData.tangents.size() == meshInfo.vertexCount || meshData.tangents.empty());
    HOST_ASSERT(meshData.texCoord0s.size() == meshInfo.vertexCount || meshData.texCoord0s.empty());
    const bool hasTangents = !meshData.tangents.empty();
    const bool hasTexCoord0s = !meshData.texCoord0s.empty();
    const bool usesShortIndices = meshInfo.vertexCount <= 0xFFFF;
    std::vector<uint8_t> packedIndices, packedMeshletVertices;
    if (usesShortIndices)
    {
        {
            size_t byteCount = meshInfo.indexCount * sizeof(uint16_t);
            byteCount = (byteCount + 3) & ~size_t(3); // "pad to 4byte boundary to make things simpler later"
            packedIndices.resize(byteCount);
            uint16_t *dst = reinterpret_cast<uint16_t *>(packedIndices.data());
            for (size_t i = 0; i < meshInfo.indexCount; ++i) dst[i] = (uint16_t)meshData.indices[i];
        }
        {
            size_t byteCount = meshData.meshletVertices.size() * sizeof(uint16_t);
            byteCount = (byteCount + 3) & ~size_t(3);
            packedMeshletVertices.resize(byteCount);
            uint16_t *dst = reinterpret_cast<uint16_t *>(packedMeshletVertices.data());
            for (size_t i = 0; i < meshData.meshletVertices.size(); ++i) dst[i] = (uint16_t)meshData.meshletVertices[i];
        }
    }
    else
    {
        packedIndices.resize(meshInfo.indexCount * sizeof(uint32_t));
        memcpy(packedIndices.data(), meshData.indices.data(), packedIndices.size());
        // the reference sizes this copy by indexCount (DeferredLoadingContext.cpp:676-681); the meshlet vertex count is what
        // the data holds, so copy that and pad to the reference's section size
        packedMeshletVertices.assign(meshInfo.indexCount * sizeof(uint32_t), 0);
        memcpy(packedMeshletVertices.data(), meshData.meshletVertices.data(),
               std::min(packedMeshletVertices.size(), meshData.meshletVertices.size() * sizeof(uint32_t)));
    }
    uint32_t byteCount = (uint32_t)packedIndices.size();
    auto computeOffset = [&byteCount](size_t sectionBytes) -> uint32_t {
        const uint32_t offset = byteCount;
        byteCount += (uint32_t)sectionBytes;
        HOST_ASSERT(byteCount % sizeof(uint32_t) == 0 && "Mesh data is not aligned properly");
        return offset;
    };
    const uint32_t elementSize = (uint32_t)sizeof(uint32_t);
    const uint32_t positionsOffset = computeOffset(meshData.positions.size() * 8) / elementSize;
    const uint32_t normalsOffset = computeOffset(meshData.normals.size() * 4) / elementSize;
    const uint32_t tangentsOffset = computeOffset(meshData.tangents.size() * 4) / elementSize;
    const uint32_t texCoord0sOffset = computeOffset(meshData.texCoord0s.size() * 4) / elementSize;
    const uint32_t meshletsOffset = computeOffset(meshData.meshlets.size() * sizeof(meshopt_Meshlet)) / elementSize;
    const uint32_t meshletBoundsOffset = computeOffset(meshData.meshletBounds.size() * sizeof(MeshletBounds)) / elementSize;
    const uint32_t meshletVerticesOffset =
        computeOffset(packedMeshletVertices.size()) / (uint32_t)(usesShortIndices ? sizeof(uint16_t) : sizeof(uint32_t));
    const uint32_t meshletTrianglesOffset = computeOffset(meshData.meshletTriangles.size());
    HOST_ASSERT(byteCount < sGeometryBufferSize && "The default size for geometry buffers doesn't fit the mesh");

    MeshCacheHeader header;
    header.sourceWriteTime = sceneWriteTime;
    header.indexCount = meshInfo.indexCount;
    header.vertexCount = meshInfo.vertexCount;
    header.meshletCount = (uint32_t)meshData.meshlets.size();
    header.positionsOffset = positionsOffset;
    header.normalsOffset = normalsOffset;
    header.tangentsOffset = hasTangents ? tangentsOffset : 0xFFFF'FFFF;
    header.texCoord0sOffset = hasTexCoord0s ? texCoord0sOffset : 0xFFFF'FFFF;
    header.meshletsOffset = meshletsOffset;
    header.meshletBoundsOffset = meshletBoundsOffset;
    header.meshletVerticesOffset = meshletVerticesOffset;
    header.meshletTrianglesByteOffset = meshletTrianglesOffset;
    header.usesShortIndices = usesShortIndices ? 1u : 0u;
    header.blobByteCount = byteCount;

    const std::filesystem::path cachePath = getCachePath(sceneDir, meshIndex);
    std::filesystem::create_directories(cachePath.parent_path());
    std::filesystem::remove(cachePath);
    std::filesystem::path cacheTmpPath = cachePath;
    cacheTmpPath.replace_extension("prosper_mesh_TMP");
    {
        std::ofstream cacheFile{cacheTmpPath, std::ios_base::binary};
        writeRaw(cacheFile, sMeshCacheMagic);
        writeRaw(cacheFile, sMeshCacheVersion);
        writeRaw(cacheFile, header.sourceWriteTime);
        writeRaw(cacheFile, header.indexCount);
        writeRaw(cacheFile, header.vertexCount);
        writeRaw(cacheFile, header.meshletCount);
        writeRaw(cacheFile, header.positionsOffset);
        writeRaw(cacheFile, header.normalsOffset);
        writeRaw(cacheFile, header.tangentsOffset);
        writeRaw(cacheFile, header.texCoord0sOffset);
        writeRaw(cacheFile, header.meshletsOffset);
        writeRaw(cacheFile, header.meshletBoundsOffset);
        writeRaw(cacheFile, header.meshletVerticesOffset);
        writeRaw(cacheFile, header.meshletTrianglesByteOffset);
        writeRaw(cacheFile, header.usesShortIndices);
        writeRaw(cacheFile, header.blobByteCount);
        const std::streampos blobStart = cacheFile.tellp();
        writeRawSpan(cacheFile, packedIndices);
        writeRawSpan(cacheFile, meshData.positions);
        writeRawSpan(cacheFile, meshData.normals);
        writeRawSpan(cacheFile, meshData.tangents);
        writeRawSpan(cacheFile, meshData.texCoord0s);
        writeRawSpan(cacheFile, meshData.meshlets);
        writeRawSpan(cacheFile, meshData.meshletBounds);
        writeRawSpan(cacheFile, packedMeshletVertices);
        writeRawSpan(cacheFile, meshData.meshletTriangles);
        HOST_ASSERT((std::streamoff)(cacheFile.tellp() - blobStart) == (std::streamoff)header.blobByteCount);
    }
    std::filesystem::rename(cacheTmpPath, cachePath);
}

// ---- animations (Animations.hpp, Accessors.cpp) ---------------------------------------------------------------------
enum class InterpolationType : uint8_t { Step, Linear, CubicSpline };

struct KeyFrameInterpolation // Accessors.hpp:9-17
{
    float t{0.f};
    float stepDuration{0.f};
    uint32_t firstFrame{0};
};

struct TimeAccessor // Accessors.cpp:9-74
{
    std::vector<float> data;
    float startTimeS{0.f}, endTimeS{0.f};
    KeyFrameInterpolation interpolation(float timeS) const
    {
        const uint32_t count = (uint32_t)data.size();
        if (timeS <= startTimeS || timeS < data[0]) return KeyFrameInterpolation{0.f, 0.f, 0};
        if (timeS >= endTimeS) return KeyFrameInterpolation{0.f, 0.f, count - 1};
        KeyFrameInterpolation ret;
        const uint32_t lastFrame = count - 1;
        while (ret.firstFrame < lastFrame)
        {
            const float frameTimeS = data[ret.firstFrame];
            if (frameTimeS > timeS) break;
            ret.firstFrame++;
        }
        ret.firstFrame--; // "We went one over in the loop"
        if (ret.firstFrame < lastFrame)
        {
            const float firstTime = data[ret.firstFrame];
            const float secondTime = data[ret.firstFrame + 1];
            const float duration = secondTime - firstTime;
            ret.stepDuration = timeS - firstTime;
            ret.t = ret.stepDuration / duration;
        }
        return ret;
    }
};

template <typename T> struct Animation // Animations.hpp:23-125
{
    InterpolationType interpolation{InterpolationType::Step};
    TimeAccessor timeFrames;
    std::vector<T> valueFrames;
    std::vector<T *> targets;

    void update(float timeS)
    {
        const KeyFrameInterpolation interp = timeFrames.interpolation(timeS);
        T firstValue{};
        if (interp.t == 0.f && interpolation == InterpolationType::CubicSpline)
            firstValue = valueFrames[(interp.firstFrame * 3) + 1]; // three values per frame, the property value is the middle one
        else
            firstValue = valueFrames[interp.firstFrame];
        T value{};
        if (interp.t == 0.f || interpolation == InterpolationType::Step)
            value = firstValue;
        else if (interpolation == InterpolationType::Linear)
        {
            const T secondValue = valueFrames[interp.firstFrame + 1];
            if constexpr (std::is_same_v<T, quat>)
                value = glm::slerp(firstValue, secondValue, interp.t);
            else
                value = (1.f - interp.t) * firstValue + interp.t * secondValue;
        }
        else
        {
            // three values per keyframe: in-tangent, property, out-tangent
            const uint32_t firstIndex = interp.firstFrame * 3;
            const T vk = valueFrames[firstIndex + 1];
            const T bk = valueFrames[firstIndex + 2];
            const T vk1 = valueFrames[firstIndex + 3 + 1];
            const T ak1 = valueFrames[firstIndex + 3];
            const float t3 = interp.t * interp.t * interp.t;
            const float t2 = interp.t * interp.t;
            const float t2Times2 = 2.f * interp.t * interp.t;
            const float t3Times2 = 2.f * interp.t * interp.t * interp.t;
            const float t2Times3 = 3.f * interp.t * interp.t;
            const float td = interp.stepDuration;
            const float t = interp.t;
            value = (t3Times2 - t2Times3 + 1.f) * vk + td * (t3 - t2Times2 + t) * bk + (-t3Times2 + t2Times3) * vk1 + td * (t3 - t2) * ak1;
            if constexpr (std::is_same_v<T, quat>) value = glm::normalize(value);
        }
        for (T *target : targets) *target = value;
    }
};

struct NodeAnimations // WorldData.hpp
{
    std::optional<uint32_t> translation, rotation, scale; // indices into Animations::vec3 / ::quat
};

// ---- scene (Scene.hpp, WorldData.cpp) ------------------------------------------------------------------------------
struct TmpNode // WorldData.cpp TmpNode
{
    std::string gltfName;
    std::vector<uint32_t> children;
    std::optional<vec3> translation;
    std::optional<quat> rotation;
    std::optional<vec3> scale;
    std::optional<uint32_t> modelIndex;
    std::optional<uint32_t> camera;
    std::optional<uint32_t> light;
};

struct SceneNode // Scene::Node
{
    uint32_t gltfSourceNode{0xFFFF'FFFF};
    std::optional<uint32_t> parent;
    uint32_t firstChild{0xFFFF'FFFF};
    uint32_t lastChild{0};
    std::optional<vec3> translation;
    std::optional<quat> rotation;
    std::optional<vec3> scale;
    std::optional<uint32_t> modelIndex;
    std::optional<uint32_t> modelInstance;
    std::optional<uint32_t> camera;
    bool directionalLight{false};
    std::optional<uint32_t> pointLight;
    std::optional<uint32_t> spotLight;
    bool dynamicTransform{false};
};

struct ModelInstance
{
    uint32_t modelIndex{0};
    prt_model_instance_transforms transforms{};
};

struct CameraParameters // Camera.hpp:40-45
{
    float fov{glm::radians(59.f)};
    float zN{0.1f};
    float zF{100.f};
};

} // namespace

struct prth_scene
{
    cgltf_data *gltf{nullptr};
    std::filesystem::path sceneDir;
    int64_t sceneWriteTime{0};

    std::vector<prt_sampler_desc> samplers;
    std::vector<uint32_t> texture2DSamplers; // packed (sampler << 24) | texture, index = glTF texture + 1
    std::vector<std::string> imageUris;      // glTF image order
    std::vector<uint8_t> imageLinear;
    std::vector<prt_material_data> materials;
    std::vector<std::pair<InputGeometryMetadata, prt_mesh_info>> meshes;
    std::vector<uint32_t> modelOffsets;
    std::vector<prt_sub_model> subModels;
    std::vector<Animation<vec3>> animVec3;
    std::vector<Animation<quat>> animQuat;
    std::vector<CameraParameters> cameras;

    // the current scene
    std::vector<SceneNode> nodes;
    std::vector<uint32_t> rootNodes;
    std::vector<ModelInstance> modelInstances;
    uint32_t drawInstanceCount{0};
    float endTimeS{0.f};
    prt_directional_light directional{{2.f, 2.f, 2.f, 0.f}, {-1.f, -1.f, -1.f, 0.f}}; // lights.h:9,18-19 defaults
    std::vector<prt_point_light> pointLights;
    std::vector<prt_spot_light> spotLights;
    bool hasCamera{false};
    prth_camera camera{{1.f, 0.5f, 1.f}, {0.f, 0.f, 0.f}, {0.f, 1.f, 0.f}, glm::radians(59.f), 0.1f, 100.f}; // Camera.hpp:24-26, 42-44

    // updateBuffers output
    std::vector<prt_model_instance_transforms> transforms;
    std::vector<float> scales;
    std::vector<prt_draw_instance> drawInstances;
    std::vector<prt_tlas_instance> tlasInstances;

    ~prth_scene()
    {
        if (gltf) cgltf_free(gltf);
    }
};

namespace
{

// GL enums of glTF samplers (WorldData.cpp:170-217)
uint32_t filterMode(cgltf_int glEnum)
{
    switch (glEnum)
    {
    case 9728: case 9984: case 9986: return PRT_FILTER_NEAREST;
    default: return PRT_FILTER_LINEAR; // 9729, 9985, 9987 and the "invalid" fall-through
    }
}
uint32_t addressMode(cgltf_int glEnum)
{
    switch (glEnum)
    {
    case 33648: return PRT_ADDRESS_MIRRORED_REPEAT;
    case 10497: return PRT_ADDRESS_REPEAT;
    default: return PRT_ADDRESS_CLAMP_TO_EDGE; // 33071 and the "invalid" fall-through
    }
}

cgltf_data *loadGltf(const std::filesystem::path &path)
{ // WorldData.cpp:103-138
    cgltf_file_type gltfType = cgltf_file_type_invalid;
    if (path.extension() == ".gltf") gltfType = cgltf_file_type_gltf;
    else if (path.extension() == ".glb") gltfType = cgltf_file_type_glb;
    else throw std::runtime_error("Unknown extension '" + path.extension().string() + "'");
    cgltf_options options{};
    options.type = gltfType;
    cgltf_data *data = nullptr;
    cgltf_result result = cgltf_parse_file(&options, path.string().c_str(), &data);
    if (result != cgltf_result_success) throw std::runtime_error("Failed to parse gltf '" + path.string() + "': " + std::to_string((int)result));
    result = cgltf_load_buffers(&options, data, path.string().c_str());
    if (result != cgltf_result_success)
    {
        cgltf_free(data);
        throw std::runtime_error("Failed to load glTF buffers: " + std::to_string((int)result));
    }
    return data;
}

void loadTextures(prth_scene &s)
{ // WorldData.cpp:681-754
    const cgltf_data &gltfData = *s.gltf;
    s.samplers.push_back(prt_sampler_desc{PRT_FILTER_LINEAR, PRT_FILTER_LINEAR, PRT_ADDRESS_REPEAT, PRT_ADDRESS_REPEAT});
    HOST_ASSERT(gltfData.samplers_count < 0xFE && "Too many samplers to pack in u32 texture index");
    for (cgltf_size i = 0; i < gltfData.samplers_count; ++i)
    {
        const cgltf_sampler &sampler = gltfData.samplers[i];
        s.samplers.push_back(prt_sampler_desc{filterMode(sampler.mag_filter), filterMode(sampler.min_filter), addressMode(sampler.wrap_s),
                                              addressMode(sampler.wrap_t)});
    }
    for (cgltf_size i = 0; i < gltfData.images_count; ++i)
    {
        const cgltf_image &image = gltfData.images[i];
        if (image.uri == nullptr)
            throw std::runtime_error("Embedded glTF textures aren't supported. Scene should be glTF + bin + textures.");
        s.imageUris.emplace_back(image.uri);
    }
    s.imageLinear.assign(gltfData.images_count, 0);
    s.texture2DSamplers.push_back(0u);
    for (cgltf_size i = 0; i < gltfData.textures_count; ++i)
    {
        const cgltf_texture &texture = gltfData.textures[i];
        HOST_ASSERT(texture.image != nullptr);
        const uint32_t imageIndex = (uint32_t)(cgltf_image_index(&gltfData, texture.image) + 1);
        const uint32_t samplerIndex = texture.sampler == nullptr ? 0 : (uint32_t)(cgltf_sampler_index(&gltfData, texture.sampler) + 1);
        s.texture2DSamplers.push_back((samplerIndex << 24) | imageIndex);
    }
}

std::optional<uint32_t> getImageIndex(const cgltf_data *gltfData, const cgltf_texture *texture)
{ // DeferredLoadingContext.cpp:576-583
    if (texture == nullptr || texture->image == nullptr) return {};
    return (uint32_t)cgltf_image_index(gltfData, texture->image);
}

void assignColorSpaces(prth_scene &s)
{ // DeferredLoadingContext::launch, DeferredLoadingContext.cpp:1132-1173: base colour = sRGB (wins), the rest linear
    const cgltf_data *gltfData = s.gltf;
    std::vector<uint8_t> srgb(gltfData->images_count, 0), linear(gltfData->images_count, 0);
    for (cgltf_size i = 0; i < gltfData->materials_count; ++i)
    {
        const cgltf_material &material = gltfData->materials[i];
        if (material.has_pbr_metallic_roughness == 1)
        {
            const cgltf_pbr_metallic_roughness &pbrParams = material.pbr_metallic_roughness;
            if (const auto baseColorIndex = getImageIndex(gltfData, pbrParams.base_color_texture.texture))
            {
                linear[*baseColorIndex] = 0;
                srgb[*baseColorIndex] = 1;
            }
            if (const auto mrIndex = getImageIndex(gltfData, pbrParams.metallic_roughness_texture.texture))
                if (!srgb[*mrIndex]) linear[*mrIndex] = 1;
        }
        if (const auto normalIndex = getImageIndex(gltfData, material.normal_texture.texture))
            if (!srgb[*normalIndex]) linear[*normalIndex] = 1;
    }
    s.imageLinear = linear;
}

void loadMaterials(prth_scene &s)
{ // WorldData.cpp:756-834 (the deferred context's material list is what the scene ends up with)
    const cgltf_data &gltfData = *s.gltf;
    const prt_material_data defaults{{1.f, 1.f, 1.f, 1.f}, 1.f, 1.f, 0.5f, PRT_ALPHA_OPAQUE, 0u, 0u, 0u, 0u}; // material_data.h:53-68
    s.materials.push_back(defaults);
    for (cgltf_size i = 0; i < gltfData.materials_count; ++i)
    {
        const cgltf_material &material = gltfData.materials[i];
        prt_material_data mat = defaults;
        if (material.has_pbr_metallic_roughness == 0) continue; // as the reference does (later material indices shift)
        auto getTexture2dSampler = [&](cgltf_texture_view const &tv) -> uint32_t {
            if (tv.texture != nullptr) return s.texture2DSamplers[cgltf_texture_index(&gltfData, tv.texture) + 1];
            return 0u;
        };
        const cgltf_pbr_metallic_roughness &pbrParams = material.pbr_metallic_roughness;
        mat.baseColorTextureSampler = getTexture2dSampler(pbrParams.base_color_texture);
        memcpy(mat.baseColorFactor, pbrParams.base_color_factor, sizeof(mat.baseColorFactor));
        mat.metallicRoughnessTextureSampler = getTexture2dSampler(pbrParams.metallic_roughness_texture);
        mat.metallicFactor = pbrParams.metallic_factor;
        mat.roughnessFactor = pbrParams.roughness_factor;
        mat.normalTextureSampler = getTexture2dSampler(material.normal_texture);
        if (material.alpha_mode == cgltf_alpha_mode_mask) mat.alphaMode = PRT_ALPHA_MASK;
        else if (material.alpha_mode == cgltf_alpha_mode_blend) mat.alphaMode = PRT_ALPHA_BLEND;
        mat.alphaCutoff = material.alpha_cutoff;
        s.materials.push_back(mat);
    }
}

void loadModels(prth_scene &s)
{ // WorldData.cpp:836-925
    const cgltf_data &gltfData = *s.gltf;
    uint32_t meshIndex = 0;
    s.modelOffsets.push_back(0);
    for (cgltf_size mi = 0; mi < gltfData.meshes_count; ++mi)
    {
        const cgltf_mesh &mesh = gltfData.meshes[mi];
        for (cgltf_size pi = 0; pi < mesh.primitives_count; ++pi)
        {
            const cgltf_primitive &primitive = mesh.primitives[pi];
            HOST_ASSERT(primitive.indices != nullptr);
            InputGeometryMetadata inputMetadata;
            inputMetadata.indices = primitive.indices;
            inputMetadata.sourceMeshIndex = (uint32_t)mi;
            inputMetadata.sourcePrimitiveIndex = (uint32_t)pi;
            for (cgltf_size ai = 0; ai < primitive.attributes_count; ++ai)
            {
                const cgltf_attribute &attr = primitive.attributes[ai];
                HOST_ASSERT(attr.data != nullptr);
                if (strcmp("POSITION", attr.name) == 0) inputMetadata.positions = attr.data;
                else if (strcmp("NORMAL", attr.name) == 0) inputMetadata.normals = attr.data;
                else if (strcmp("TANGENT", attr.name) == 0) inputMetadata.tangents = attr.data;
                else if (strcmp("TEXCOORD_0", attr.name) == 0) inputMetadata.texCoord0s = attr.data;
            }
            HOST_ASSERT(inputMetadata.positions != nullptr);
            HOST_ASSERT(inputMetadata.normals != nullptr);
            HOST_ASSERT(inputMetadata.positions->count == inputMetadata.normals->count);
            HOST_ASSERT(inputMetadata.tangents == nullptr || inputMetadata.tangents->count == inputMetadata.positions->count);
            HOST_ASSERT(inputMetadata.texCoord0s == nullptr || inputMetadata.texCoord0s->count == inputMetadata.positions->count);
            const uint32_t material = primitive.material != nullptr ? (uint32_t)(cgltf_material_index(&gltfData, primitive.material) + 1) : 0;
            prt_mesh_info meshInfo{};
            meshInfo.vertexCount = (uint32_t)inputMetadata.positions->count;
            meshInfo.indexCount = (uint32_t)inputMetadata.indices->count;
            meshInfo.materialIndex = material;
            s.meshes.emplace_back(inputMetadata, meshInfo);
            s.subModels.push_back(prt_sub_model{meshIndex++, material});
        }
        s.modelOffsets.push_back((uint32_t)s.subModels.size());
    }
}

std::vector<float> accessorFloats(const cgltf_accessor &accessor, size_t components)
{ // appendAccessorData (WorldData.cpp:220-236): the raw floats of a tightly packed f32 accessor
    std::vector<float> out(accessor.count * components);
    const cgltf_size n = cgltf_accessor_unpack_floats(&accessor, out.data(), out.size());
    HOST_ASSERT(n == out.size());
    return out;
}

std::vector<NodeAnimations> loadAnimations(prth_scene &s)
{ // WorldData.cpp:960-1120
    const cgltf_data &gltfData = *s.gltf;
    std::vector<NodeAnimations> ret(gltfData.nodes_count);
    for (cgltf_size ai = 0; ai < gltfData.animations_count; ++ai)
    {
        const cgltf_animation &animation = gltfData.animations[ai];
        std::vector<std::pair<bool, uint32_t>> concreteAnimations; // (isQuat, index), by glTF sampler index
        for (cgltf_size si = 0; si < animation.samplers_count; ++si)
        {
            const cgltf_animation_sampler &sampler = animation.samplers[si];
            InterpolationType interpolation{InterpolationType::Step};
            if (sampler.interpolation == cgltf_interpolation_type_step) interpolation = InterpolationType::Step;
            else if (sampler.interpolation == cgltf_interpolation_type_linear) interpolation = InterpolationType::Linear;
            else if (sampler.interpolation == cgltf_interpolation_type_cubic_spline) interpolation = InterpolationType::CubicSpline;
            else HOST_ASSERT(!"Unsupported interpolation type");
            HOST_ASSERT(sampler.input != nullptr);
            const cgltf_accessor &inputAccessor = *sampler.input;
            HOST_ASSERT(!inputAccessor.is_sparse);
            HOST_ASSERT(inputAccessor.component_type == cgltf_component_type_r_32f);
            HOST_ASSERT(inputAccessor.type == cgltf_type_scalar);
            HOST_ASSERT(inputAccessor.has_min);
            HOST_ASSERT(inputAccessor.has_max);
            TimeAccessor timeFrames;
            timeFrames.data = accessorFloats(inputAccessor, 1);
            timeFrames.startTimeS = (float)inputAccessor.min[0];
            timeFrames.endTimeS = (float)inputAccessor.max[0];
            HOST_ASSERT(sampler.output != nullptr);
            const cgltf_accessor &outputAccessor = *sampler.output;
            HOST_ASSERT(!outputAccessor.is_sparse);
            HOST_ASSERT(outputAccessor.component_type == cgltf_component_type_r_32f);
            if (outputAccessor.type == cgltf_type_vec3)
            {
                const std::vector<float> raw = accessorFloats(outputAccessor, 3);
                Animation<vec3> a;
                a.interpolation = interpolation;
                a.timeFrames = std::move(timeFrames);
                a.valueFrames.resize(outputAccessor.count);
                memcpy(a.valueFrames.data(), raw.data(), raw.size() * sizeof(float));
                s.animVec3.push_back(std::move(a));
                concreteAnimations.emplace_back(false, (uint32_t)s.animVec3.size() - 1);
            }
            else if (outputAccessor.type == cgltf_type_vec4)
            {
                const std::vector<float> raw = accessorFloats(outputAccessor, 4);
                Animation<quat> a;
                a.interpolation = interpolation;
                a.timeFrames = std::move(timeFrames);
                a.valueFrames.resize(outputAccessor.count);
                // ValueAccessor<quat>::read reinterprets the glTF bytes as glm::quat (Accessors.cpp:104-109); make_quat is the
                // same byte copy
                for (cgltf_size k = 0; k < outputAccessor.count; ++k) a.valueFrames[k] = make_quat(&raw[k * 4]);
                s.animQuat.push_back(std::move(a));
                concreteAnimations.emplace_back(true, (uint32_t)s.animQuat.size() - 1);
            }
            else
                HOST_ASSERT(!"Unsupported animation output type");
        }
        for (cgltf_size ci = 0; ci < animation.channels_count; ++ci)
        {
            const cgltf_animation_channel &channel = animation.channels[ci];
            HOST_ASSERT(channel.target_node != nullptr);
            const uint32_t nodeIndex = (uint32_t)cgltf_node_index(&gltfData, channel.target_node);
            HOST_ASSERT(channel.sampler != nullptr);
            const uint32_t samplerIndex = (uint32_t)cgltf_animation_sampler_index(&animation, channel.sampler);
            NodeAnimations &target = ret[nodeIndex];
            if (channel.target_path == cgltf_animation_path_type_translation) target.translation = concreteAnimations[samplerIndex].second;
            else if (channel.target_path == cgltf_animation_path_type_rotation) target.rotation = concreteAnimations[samplerIndex].second;
            else if (channel.target_path == cgltf_animation_path_type_scale) target.scale = concreteAnimations[samplerIndex].second;
        }
    }
    return ret;
}

void gatherScene(prth_scene &s, const cgltf_scene &gltfScene, const std::vector<TmpNode> &nodes)
{ // WorldData.cpp:1384-1543
    const cgltf_data &gltfData = *s.gltf;
    struct NodePair { uint32_t tmpNode, sceneNode; };
    std::vector<NodePair> nodeStack;
    bool directionalLightFound = false;
    for (cgltf_size ri = 0; ri < gltfScene.nodes_count; ++ri)
    {
        const cgltf_node *node = gltfScene.nodes[ri];
        HOST_ASSERT(node != nullptr);
        const cgltf_size nodeIndex = cgltf_node_index(&gltfData, node);
        s.rootNodes.push_back((uint32_t)s.nodes.size());
        s.nodes.emplace_back();
        nodeStack.clear();
        nodeStack.push_back(NodePair{(uint32_t)nodeIndex, (uint32_t)s.nodes.size() - 1});
        while (!nodeStack.empty())
        {
            const NodePair indices = nodeStack.back();
            nodeStack.pop_back();
            const TmpNode &tmpNode = nodes[indices.tmpNode];
            const uint32_t childCount = (uint32_t)tmpNode.children.size();
            const uint32_t firstChild = (uint32_t)s.nodes.size();
            const uint32_t lastChild = firstChild + childCount - 1; // "If no children, firstChild <= lastChild false as intended"
            s.nodes.resize(s.nodes.size() + childCount);
            SceneNode &sceneNode = s.nodes[indices.sceneNode];
            sceneNode.gltfSourceNode = indices.tmpNode;
            sceneNode.firstChild = firstChild;
            sceneNode.lastChild = lastChild;
            for (uint32_t i = 0; i < childCount; ++i)
            {
                const uint32_t childIndex = sceneNode.firstChild + i;
                s.nodes[childIndex].parent = indices.sceneNode;
                nodeStack.push_back(NodePair{tmpNode.children[i], childIndex});
            }
            sceneNode.translation = tmpNode.translation;
            sceneNode.rotation = tmpNode.rotation;
            sceneNode.scale = tmpNode.scale;
            sceneNode.modelIndex = tmpNode.modelIndex;
            sceneNode.camera = tmpNode.camera;
            if (sceneNode.modelIndex.has_value())
            {
                sceneNode.modelInstance = (uint32_t)s.modelInstances.size();
                ModelInstance mi;
                mi.modelIndex = *sceneNode.modelIndex;
                s.modelInstances.push_back(mi);
                s.drawInstanceCount += s.modelOffsets[*sceneNode.modelIndex + 1] - s.modelOffsets[*sceneNode.modelIndex];
            }
            if (tmpNode.light.has_value())
            {
                const cgltf_light &light = gltfData.lights[*tmpNode.light];
                if (light.type == cgltf_light_type_directional)
                {
                    // gltf blender exporter puts W/m^2 into intensity
                    const vec4 irradiance = vec4{(float)light.color[0], (float)light.color[1], (float)light.color[2], 0.f} * (float)light.intensity;
                    memcpy(s.directional.irradiance, &irradiance.x, 16);
                    sceneNode.directionalLight = true;
                    directionalLightFound = true;
                }
                else if (light.type == cgltf_light_type_point)
                {
                    auto radiance = vec3{(float)light.color[0], (float)light.color[1], (float)light.color[2]} * (float)light.intensity
                                    // gltf blender exporter puts W into intensity
                                    / (4.f * glm::pi<float>());
                    const auto luminance = dot(radiance, vec3{0.2126, 0.7152, 0.0722});
                    const auto minLuminance = 0.01f;
                    const auto radius = light.range > 0.f ? light.range : sqrt(luminance / minLuminance);
                    sceneNode.pointLight = (uint32_t)s.pointLights.size();
                    prt_point_light pl{};
                    pl.radianceAndRadius[0] = radiance.x; pl.radianceAndRadius[1] = radiance.y; pl.radianceAndRadius[2] = radiance.z;
                    pl.radianceAndRadius[3] = radius;
                    s.pointLights.push_back(pl);
                }
                else if (light.type == cgltf_light_type_spot)
                {
                    sceneNode.spotLight = (uint32_t)s.spotLights.size();
                    // angular attenuation from the gltf spec
                    const auto angleScale = 1.f / max(0.001f, cos(light.spot_inner_cone_angle) - cos(light.spot_outer_cone_angle));
                    const auto angleOffset = -cos(light.spot_outer_cone_angle) * angleScale;
                    vec4 r = vec4{(float)light.color[0], (float)light.color[1], (float)light.color[2], 0.f} * (float)light.intensity;
                    r /= 4.f * glm::pi<float>();
                    r.w = angleScale;
                    prt_spot_light sl{};
                    memcpy(sl.radianceAndAngleScale, &r.x, 16);
                    sl.positionAndAngleOffset[3] = angleOffset;
                    s.spotLights.push_back(sl);
                }
            }
        }
    }
    // "Honor scene lighting"
    if (!directionalLightFound && (!s.pointLights.empty() || !s.spotLights.empty()))
        memset(s.directional.irradiance, 0, sizeof(s.directional.irradiance));
}

void loadScenes(prth_scene &s, const std::vector<NodeAnimations> &nodeAnimations)
{ // WorldData.cpp:1124-1382 (the default scene only: it is the one the renderer shows)
    const cgltf_data &gltfData = *s.gltf;
    std::vector<TmpNode> nodes;
    nodes.reserve(gltfData.nodes_count);
    for (cgltf_size ni = 0; ni < gltfData.nodes_count; ++ni)
    {
        const cgltf_node &gltfNode = gltfData.nodes[ni];
        nodes.emplace_back();
        TmpNode &node = nodes.back();
        node.gltfName = gltfNode.name != nullptr ? gltfNode.name : "";
        for (cgltf_size ci = 0; ci < gltfNode.children_count; ++ci)
            node.children.push_back((uint32_t)cgltf_node_index(&gltfData, gltfNode.children[ci]));
        if (gltfNode.mesh != nullptr) node.modelIndex = (uint32_t)cgltf_mesh_index(&gltfData, gltfNode.mesh);
        if (gltfNode.camera != nullptr)
        {
            const uint32_t cameraIndex = (uint32_t)cgltf_camera_index(&gltfData, gltfNode.camera);
            const cgltf_camera &cam = *gltfNode.camera;
            if (cam.type == cgltf_camera_type_perspective)
            {
                if (s.cameras.size() <= cameraIndex) s.cameras.resize(cameraIndex + 1);
                s.cameras[cameraIndex] = CameraParameters{(float)cam.data.perspective.yfov, (float)cam.data.perspective.znear,
                                                          (float)cam.data.perspective.zfar};
                node.camera = cameraIndex;
            }
        }
        if (gltfNode.light != nullptr) node.light = (uint32_t)cgltf_light_index(&gltfData, gltfNode.light);
        vec3 translation{0.f};
        vec3 scale{1.f};
        quat rotation{1.f, 0.f, 0.f, 0.f};
        if (gltfNode.has_matrix == 1)
        {
            // "Spec defines the matrix to be decomposeable to T * R * S"
            const mat4 matrix = make_mat4(&gltfNode.matrix[0]);
            vec3 skew;
            vec4 perspective;
            decompose(matrix, scale, rotation, translation, skew, perspective);
        }
        if (gltfNode.has_translation == 1) translation = make_vec3(&gltfNode.translation[0]);
        if (gltfNode.has_rotation == 1) rotation = make_quat(&gltfNode.rotation[0]);
        if (gltfNode.has_scale == 1) scale = make_vec3(&gltfNode.scale[0]);
        const float srtThreshold = 0.001f; // "Skip transform components that are close to identity"
        if (any(lessThan(translation, vec3{-srtThreshold})) || any(greaterThan(translation, vec3{srtThreshold}))) node.translation = translation;
        const vec3 eulers = eulerAngles(rotation);
        if (any(lessThan(eulers, vec3{-srtThreshold})) || any(greaterThan(eulers, vec3{srtThreshold}))) node.rotation = rotation;
        if (any(lessThan(scale, vec3{1.f - srtThreshold})) || any(greaterThan(scale, vec3{1.f + srtThreshold}))) node.scale = scale;
    }
    HOST_ASSERT(gltfData.scenes_count > 0);
    const cgltf_scene &gltfScene = gltfData.scene != nullptr ? *gltfData.scene : gltfData.scenes[0];
    gatherScene(s, gltfScene, nodes);
    // register the animation targets (nodes won't move in memory anymore)
    for (SceneNode &node : s.nodes)
    {
        const NodeAnimations &animations = nodeAnimations[node.gltfSourceNode];
        if (animations.translation.has_value())
        {
            node.dynamicTransform = true;
            Animation<vec3> &animation = s.animVec3[*animations.translation];
            s.endTimeS = std::max(s.endTimeS, animation.timeFrames.endTimeS);
            if (!node.translation.has_value()) node.translation = vec3{0.f};
            animation.targets.push_back(&*node.translation);
        }
        if (animations.rotation.has_value())
        {
            node.dynamicTransform = true;
            Animation<quat> &animation = s.animQuat[*animations.rotation];
            s.endTimeS = std::max(s.endTimeS, animation.timeFrames.endTimeS);
            if (!node.rotation.has_value()) node.rotation = quat{1.f, 0.f, 0.f, 0.f};
            animation.targets.push_back(&*node.rotation);
        }
        if (animations.scale.has_value())
        {
            node.dynamicTransform = true;
            Animation<vec3> &animation = s.animVec3[*animations.scale];
            s.endTimeS = std::max(s.endTimeS, animation.timeFrames.endTimeS);
            if (!node.scale.has_value()) node.scale = vec3{1.f};
            animation.targets.push_back(&*node.scale);
        }
    }
    // propagate dynamic flags (a child of a dynamic node is dynamic); parents precede their children in `nodes`
    for (SceneNode &node : s.nodes)
        if (node.parent.has_value() && s.nodes[*node.parent].dynamicTransform) node.dynamicTransform = true;
    if (s.cameras.empty()) s.cameras.emplace_back(); // "Make sure we always have a camera"
}

void processMeshes(prth_scene &s)
{ // loadNextMesh for every mesh, DeferredLoadingContext.cpp:804-865 (cache write; reading it back is the caller's upload)
    for (uint32_t meshIndex = 0; meshIndex < s.meshes.size(); ++meshIndex)
    {
        const InputGeometryMetadata &metadata = s.meshes[meshIndex].first;
        prt_mesh_info info = s.meshes[meshIndex].second;
        const std::filesystem::path cachePath = getCachePath(s.sceneDir, meshIndex);
        if (!cacheValid(cachePath, s.sceneWriteTime))
        {
            MeshData meshData = getMeshData(metadata, info);
            if (meshData.tangents.empty() && !meshData.texCoord0s.empty())
            {
                generateTangents(meshData);
                info.vertexCount = (uint32_t)meshData.positions.size();
            }
            optimizeMeshData(meshData, info);
            generateMeshlets(meshData);
            PackedMeshData packedMeshData = packMeshData(std::move(meshData));
            writeCache(s.sceneDir, s.sceneWriteTime, meshIndex, std::move(packedMeshData), info);
        }
        const std::optional<MeshCacheHeader> cacheHeader = readCache(cachePath); // "Always read from the cache"
        HOST_ASSERT(cacheHeader.has_value());
        HOST_ASSERT(cacheHeader->indexCount == info.indexCount);
        info.vertexCount = cacheHeader->vertexCount; // tangent generation can change the vertex count
        info.meshletCount = cacheHeader->meshletCount;
        s.meshes[meshIndex].second = info;
    }
}

bool relativeEq(float a, float b, float maxRelativeDiff)
{ // World.cpp:34-40
    const float diff = std::abs(a - b);
    const float maxMagnitude = std::max(std::abs(a), std::abs(b));
    return diff < maxRelativeDiff * maxMagnitude;
}

void updateScene(prth_scene &s)
{ // World::Impl::updateScene, World.cpp:359-466. `nodes` holds parents before their children, so one forward pass with the
  // parent's matrix kept per node visits every node with the same matrices as the reference's explicit stack.
    std::vector<mat4> world(s.nodes.size(), mat4{1.f});
    for (size_t nodeIndex = 0; nodeIndex < s.nodes.size(); ++nodeIndex)
    {
        SceneNode &node = s.nodes[nodeIndex];
        mat4 modelToWorld4x4 = node.parent.has_value() ? world[*node.parent] : mat4{1.f};
        if (node.translation.has_value()) modelToWorld4x4 = translate(modelToWorld4x4, *node.translation);
        if (node.rotation.has_value()) modelToWorld4x4 *= mat4_cast(*node.rotation);
        if (node.scale.has_value()) modelToWorld4x4 = scale(modelToWorld4x4, *node.scale);
        world[nodeIndex] = modelToWorld4x4;
        const mat3x4 modelToWorld = transpose(modelToWorld4x4);
        const mat3x4 normalToWorld = inverse(modelToWorld4x4); // "No transpose as mat4->mat3x4 effectively does it"
        if (node.modelInstance.has_value())
        {
            prt_model_instance_transforms &t = s.modelInstances[*node.modelInstance].transforms;
            static_assert(sizeof(mat3x4) == sizeof(t.modelToWorld), "mat3x4");
            memcpy(t.modelToWorld, &modelToWorld[0][0], sizeof(t.modelToWorld));
            memcpy(t.normalToWorld, &normalToWorld[0][0], sizeof(t.normalToWorld));
        }
        if (node.camera.has_value() && *node.camera == 0 && !s.hasCamera) s.hasCamera = true;
        if (node.camera.has_value() && *node.camera == 0)
        {
            const vec3 eye = vec3{modelToWorld4x4 * vec4{0.f, 0.f, 0.f, 1.f}};
            const vec3 target = vec3{modelToWorld4x4 * vec4{0.f, 0.f, -1.f, 1.f}};
            const vec3 up = mat3{modelToWorld4x4} * vec3{0.f, 1.f, 0.f};
            memcpy(s.camera.eye, &eye.x, 12); memcpy(s.camera.target, &target.x, 12); memcpy(s.camera.up, &up.x, 12);
            s.camera.fov = s.cameras[0].fov; s.camera.zN = s.cameras[0].zN; s.camera.zF = s.cameras[0].zF;
        }
        if (node.directionalLight)
        {
            const vec4 d = vec4{mat3{modelToWorld4x4} * vec3{0.f, 0.f, -1.f}, 0.f};
            memcpy(s.directional.direction, &d.x, 16);
        }
        if (node.pointLight.has_value())
        {
            const vec4 p = modelToWorld4x4 * vec4{0.f, 0.f, 0.f, 1.f};
            memcpy(s.pointLights[*node.pointLight].position, &p.x, 16);
        }
        if (node.spotLight.has_value())
        {
            prt_spot_light &sl = s.spotLights[*node.spotLight];
            const vec3 position = vec3{modelToWorld4x4 * vec4{0.f, 0.f, 0.f, 1.f}};
            sl.positionAndAngleOffset[0] = position.x; sl.positionAndAngleOffset[1] = position.y; sl.positionAndAngleOffset[2] = position.z;
            const vec4 d = vec4{mat3{modelToWorld4x4} * vec3{0.f, 0.f, -1.f}, 0.f};
            memcpy(sl.direction, &d.x, 16);
        }
    }
}

void updateBuffers(prth_scene &s)
{ // World::Impl::updateBuffers + updateTlasInstances, World.cpp:468-536, 878-927
    s.drawInstances.clear();
    s.transforms.clear();
    s.scales.clear();
    s.tlasInstances.clear();
    uint32_t rti = 0;
    for (uint32_t mi = 0; mi < s.modelInstances.size(); ++mi)
    {
        const ModelInstance &instance = s.modelInstances[mi];
        s.transforms.push_back(instance.transforms);
        mat3x4 modelToWorld;
        memcpy(&modelToWorld[0][0], instance.transforms.modelToWorld, sizeof(mat3x4));
        // "lengths of rows instead of columns because of the transposed 3x4"
        const vec3 scale{length(row(modelToWorld, 0)), length(row(modelToWorld, 1)), length(row(modelToWorld, 2))};
        float uniformScale = 0.f; // zero = non-uniform
        const float tolerance = 0.0001f;
        if (relativeEq(scale.x, scale.y, tolerance) && relativeEq(scale.x, scale.z, tolerance)) uniformScale = scale.x;
        s.scales.push_back(uniformScale);
        const uint32_t first = s.modelOffsets[instance.modelIndex], last = s.modelOffsets[instance.modelIndex + 1];
        for (uint32_t k = first; k < last; ++k) s.drawInstances.push_back(prt_draw_instance{mi, s.subModels[k].meshIndex, s.subModels[k].materialIndex});
        prt_tlas_instance ti{};
        memcpy(ti.transform, instance.transforms.modelToWorld, sizeof(ti.transform));
        ti.customIndexAndMask = (rti & 0xFFFFFFu) | (0xFFu << 24);
        ti.sbtOffsetAndFlags = 0;
        ti.blasReference = (uint64_t)instance.modelIndex + 1; // the C ABI's BLAS handle: model index + 1, 0 = inactive (prt.h)
        s.tlasInstances.push_back(ti);
        rti += last - first;
    }
}

} // namespace

extern "C" const char *prth_last_error(void) { return g_error.c_str(); }

extern "C" int prth_load_gltf(const char *gltfPath, const char *cacheDir, int64_t sourceWriteTime, prth_scene **out)
{
    if (!gltfPath || !out) return fail("prth_load_gltf: null argument");
    *out = nullptr;
    prth_scene *s = new prth_scene;
    try
    {
        const std::filesystem::path path{gltfPath};
        s->gltf = loadGltf(path);
        s->sceneDir = cacheDir && cacheDir[0] ? std::filesystem::path{cacheDir} : path.parent_path();
        s->sceneWriteTime = sourceWriteTime != 0 ? sourceWriteTime
                                                 : (int64_t)std::filesystem::last_write_time(path).time_since_epoch().count();
        loadTextures(*s);
        loadMaterials(*s);
        assignColorSpaces(*s);
        loadModels(*s);
        const std::vector<NodeAnimations> nodeAnimations = loadAnimations(*s);
        loadScenes(*s, nodeAnimations);
        processMeshes(*s);
        for (Animation<vec3> &a : s->animVec3) a.update(0.f);
        for (Animation<quat> &a : s->animQuat) a.update(0.f);
        updateScene(*s);
        updateBuffers(*s);
    }
    catch (const std::exception &e)
    {
        delete s;
        return fail(std::string("prth_load_gltf: ") + e.what());
    }
    *out = s;
    return 0;
}

extern "C" void prth_destroy(prth_scene *scene) { delete scene; }

extern "C" int prth_get_counts(const prth_scene *s, prth_counts *out)
{
    if (!s || !out) return fail("prth_get_counts: null argument");
    out->meshes = (uint32_t)s->meshes.size();
    out->materials = (uint32_t)s->materials.size();
    out->samplers = (uint32_t)s->samplers.size();
    out->textures = (uint32_t)s->imageUris.size() + 1;
    out->models = (uint32_t)s->modelOffsets.size() - 1;
    out->subModels = (uint32_t)s->subModels.size();
    out->nodes = (uint32_t)s->nodes.size();
    out->modelInstances = (uint32_t)s->modelInstances.size();
    out->drawInstances = s->drawInstanceCount;
    out->pointLights = (uint32_t)s->pointLights.size();
    out->spotLights = (uint32_t)s->spotLights.size();
    out->animatedNodes = (uint32_t)std::count_if(s->nodes.begin(), s->nodes.end(), [](const SceneNode &n) { return n.dynamicTransform; });
    out->hasCamera = s->hasCamera ? 1u : 0u;
    out->endTimeS = s->endTimeS;
    return 0;
}

extern "C" int prth_get_materials(const prth_scene *s, prt_material_data *out)
{
    if (!s || !out) return fail("prth_get_materials: null argument");
    memcpy(out, s->materials.data(), s->materials.size() * sizeof(prt_material_data));
    return 0;
}
extern "C" int prth_get_samplers(const prth_scene *s, prt_sampler_desc *out)
{
    if (!s || !out) return fail("prth_get_samplers: null argument");
    memcpy(out, s->samplers.data(), s->samplers.size() * sizeof(prt_sampler_desc));
    return 0;
}
extern "C" int prth_get_mesh_infos(const prth_scene *s, prt_mesh_info *out)
{
    if (!s || !out) return fail("prth_get_mesh_infos: null argument");
    for (size_t i = 0; i < s->meshes.size(); ++i) out[i] = s->meshes[i].second;
    return 0;
}
extern "C" int prth_get_models(const prth_scene *s, uint32_t *modelOffsets, prt_sub_model *subModels)
{
    if (!s || !modelOffsets || !subModels) return fail("prth_get_models: null argument");
    memcpy(modelOffsets, s->modelOffsets.data(), s->modelOffsets.size() * sizeof(uint32_t));
    memcpy(subModels, s->subModels.data(), s->subModels.size() * sizeof(prt_sub_model));
    return 0;
}
extern "C" int prth_get_texture_source(const prth_scene *s, uint32_t texture, char *uriOut, uint32_t uriCapacity, uint32_t *linear)
{
    if (!s || !uriOut || uriCapacity == 0) return fail("prth_get_texture_source: null argument");
    if (texture > s->imageUris.size()) return fail("prth_get_texture_source: texture index out of range");
    const std::string uri = texture == 0 ? std::string{} : s->imageUris[texture - 1];
    if (uri.size() + 1 > uriCapacity) return fail("prth_get_texture_source: uri buffer too small");
    memcpy(uriOut, uri.c_str(), uri.size() + 1);
    if (linear) *linear = texture == 0 ? 0u : s->imageLinear[texture - 1];
    return 0;
}
extern "C" int prth_decode_image(const char *path, uint32_t *width, uint32_t *height, uint8_t *rgbaOut, uint64_t capacity)
{
    if (!path || !width || !height) return fail("prth_decode_image: null argument");
    int w = 0, h = 0, channels = 0;
    if (rgbaOut == nullptr)
    {
        if (!stbi_info(path, &w, &h, &channels)) return fail(std::string("Failed to load texture '") + path + "'");
        *width = (uint32_t)w; *height = (uint32_t)h;
        return 0;
    }
    stbi_uc *px = stbi_load(path, &w, &h, &channels, 4); // desiredChannels = 4 (Texture.cpp:384)
    if (px == nullptr) return fail(std::string("Failed to load texture '") + path + "'");
    const uint64_t bytes = (uint64_t)w * (uint64_t)h * 4u;
    if (bytes > capacity)
    {
        stbi_image_free(px);
        return fail("prth_decode_image: destination too small");
    }
    memcpy(rgbaOut, px, bytes);
    stbi_image_free(px);
    *width = (uint32_t)w; *height = (uint32_t)h;
    return 0;
}
extern "C" int64_t prth_source_write_time(const prth_scene *s) { return s ? s->sceneWriteTime : 0; }

extern "C" int prth_update(prth_scene *s, float timeS)
{
    if (!s) return fail("prth_update: null argument");
    try
    {
        for (Animation<vec3> &a : s->animVec3) a.update(timeS); // World::Impl::updateAnimations, World.cpp:349-357
        for (Animation<quat> &a : s->animQuat) a.update(timeS);
        updateScene(*s);
        updateBuffers(*s);
    }
    catch (const std::exception &e)
    {
        return fail(std::string("prth_update: ") + e.what());
    }
    return 0;
}
extern "C" int prth_get_instances(const prth_scene *s, prt_model_instance_transforms *transforms, float *scales, prt_draw_instance *drawInstances,
                                  prt_tlas_instance *tlasInstances)
{
    if (!s) return fail("prth_get_instances: null argument");
    if (transforms) memcpy(transforms, s->transforms.data(), s->transforms.size() * sizeof(prt_model_instance_transforms));
    if (scales) memcpy(scales, s->scales.data(), s->scales.size() * sizeof(float));
    if (drawInstances) memcpy(drawInstances, s->drawInstances.data(), s->drawInstances.size() * sizeof(prt_draw_instance));
    if (tlasInstances) memcpy(tlasInstances, s->tlasInstances.data(), s->tlasInstances.size() * sizeof(prt_tlas_instance));
    return 0;
}
extern "C" int prth_get_lights(const prth_scene *s, prt_directional_light *directional, prt_point_light *points, prt_spot_light *spots)
{
    if (!s) return fail("prth_get_lights: null argument");
    if (directional) *directional = s->directional;
    if (points) memcpy(points, s->pointLights.data(), s->pointLights.size() * sizeof(prt_point_light));
    if (spots) memcpy(spots, s->spotLights.data(), s->spotLights.size() * sizeof(prt_spot_light));
    return 0;
}
extern "C" int prth_get_camera(const prth_scene *s, prth_camera *out)
{
    if (!s || !out) return fail("prth_get_camera: null argument");
    *out = s->camera;
    return 0;
}
