// core_build.h -- per-primitive bodies of the builder's input stages (shared by the CUDA kernels in
// kernels_build.cuh and the CPU kernel-logic simulator under tests/hostsim).
#pragma once
#include "core_bvh.h"
#include "core_scene.h"
#include "core_trace.h"

namespace prt {

// Host-compacted active instances (inactive ones, blasReference == 0, never reach the device:
// World.cpp:898-905).
struct TlasInput
{
    float transform[12]; // row-major 3x4 object -> world
    uint32_t customIndexAndMask;
    uint32_t index; // position in the caller's array
    const WideNode *nodes;
    const TriRec *tris;
    const AlphaUv *uvs;
};

// One BLAS triangle (World.cpp:617-668): fp16x4 positions, stride 8 (Mesh.hpp:11-12), through the
// u16/u32 index buffer. Writes the triangle record and its unpadded box.
HD void gatherTriangle(const SceneView &sv, uint32_t meshIndex, uint32_t p, uint32_t geometryIndex, uint32_t opaque, TriRec &t, V3 &lo,
                       V3 &hi, V3 *corners = nullptr)
{
    const prt_geometry_metadata md = sv.metadatas[meshIndex];
    const uint32_t i0 = loadIndex(sv, md.bufferIndex, md.indicesOffset, p * 3 + 0, md.usesShortIndices);
    const uint32_t i1 = loadIndex(sv, md.bufferIndex, md.indicesOffset, p * 3 + 1, md.usesShortIndices);
    const uint32_t i2 = loadIndex(sv, md.bufferIndex, md.indicesOffset, p * 3 + 2, md.usesShortIndices);
    const V3 v0 = loadR16G16B16A16(sv, md.bufferIndex, md.positionsOffset, i0);
    const V3 v1 = loadR16G16B16A16(sv, md.bufferIndex, md.positionsOffset, i1);
    const V3 v2 = loadR16G16B16A16(sv, md.bufferIndex, md.positionsOffset, i2);
    const V3 e1 = xsub3(v1, v0), e2 = xsub3(v2, v0);
    t.v0[0] = v0.x; t.v0[1] = v0.y; t.v0[2] = v0.z; t.primitive = p;
    t.e1[0] = e1.x; t.e1[1] = e1.y; t.e1[2] = e1.z; t.geometry = geometryIndex | (opaque << 31);
    t.e2[0] = e2.x; t.e2[1] = e2.y; t.e2[2] = e2.z; t.pad = 0;
    lo = vmin(v0, vmin(v1, v2));
    hi = vmax(v0, vmax(v1, v2));
    if (corners) { corners[0] = v0; corners[1] = v1; corners[2] = v2; }
}

// ---- reference splitting of non-opaque triangles (early split clipping, Ernst & Greiner 2007) ------------------------------
// A ray is not stopped by alpha-tested geometry: it passes every card of a plant and meets the box of every triangle on the
// way, and the box of a tilted triangle is mostly empty (on the bench atrium 90 % of the closest rays' triangle tests are on
// foliage cards and four in five of those miss; tools/cpu_item_probe.py). So every non-opaque triangle enters the BLAS builder
// as 4^PRT_BLAS_ALPHA_SPLIT references: the sub-triangles of `levels` rounds of midpoint subdivision, each with its own box and
// all with the SAME triangle record, so the exact ray/triangle test, its barycentrics and the primitive ID are untouched. The
// union of the sub-boxes covers the triangle (midpoints are rounded, the builder's pad of max|coord| / 16384 is five orders of
// magnitude above that rounding), a hit found through two references is the same (t, IDs) twice -- the second is an exact tie
// with itself and never "better" -- and an any-hit decision is a pure function of (ray, triangle), which the trace kernel
// remembers for the triangle it rejected last (core_trace.h) so that the alpha test is not repeated for a sibling reference.
// Measured: on the CPU probe 23.5 -> 17.2 triangle tests per closest ray and 10.7 -> 6.8 per shadow ray with node items unchanged
// (13.7 / 5.0 with 16 references); on the GPU (one box, `profiles/r02_ab_reference_split.txt`) the shadow stage gains 8 % but the
// closest stage does not move (the tree grows from 39 k to 64 k nodes and 258 k to 430 k records) and the rejected-triangle
// lookup alone costs the closest kernel 6 %: 1070 Mrays/s without, 1023 with the lookup and no split, 1029 / 1010 with 4 / 16
// references. Kept as an experiment switch, OFF by default.
// PRT_BLAS_ALPHA_SPLIT (core_trace.h) = rounds of midpoint subdivision: 0 = off (the default), 1 = 4 references, 2 = 16.
HD uint32_t blasRefsPerTriangle(uint32_t opaque) { return opaque ? 1u : (1u << (2 * PRT_BLAS_ALPHA_SPLIT)); }

// Box of reference `k` (base-4 digits, most significant round first: 0 = corner v0, 1 = corner v1, 2 = corner v2, 3 = the
// middle triangle) of the triangle (v0, v1, v2).
HD void triangleRefBox(V3 a, V3 b, V3 c, uint32_t k, V3 &lo, V3 &hi)
{
    for (int round = PRT_BLAS_ALPHA_SPLIT - 1; round >= 0; --round)
    {
        const V3 mab = (a + b) * 0.5f, mbc = (b + c) * 0.5f, mca = (c + a) * 0.5f;
        switch ((k >> (2 * round)) & 3u)
        {
        case 0: b = mab; c = mca; break;
        case 1: a = mab; c = mbc; break;
        case 2: a = mca; b = mbc; break;
        default: a = mab; b = mbc; c = mca; break;
        }
    }
    lo = vmin(a, vmin(b, c));
    hi = vmax(a, vmax(b, c));
}

// The any-hit record of triangle p (scene.rahit:24-30: loadUV of its three corners).
HD void gatherTriangleUvs(const SceneView &sv, uint32_t meshIndex, uint32_t p, AlphaUv &r)
{
    const V2 uv0 = loadUV(sv, meshIndex, p * 3 + 0), uv1 = loadUV(sv, meshIndex, p * 3 + 1), uv2 = loadUV(sv, meshIndex, p * 3 + 2);
    r.uv[0] = uv0.x; r.uv[1] = uv0.y; r.uv[2] = uv1.x; r.uv[3] = uv1.y; r.uv[4] = uv2.x; r.uv[5] = uv2.y;
    r.pad[0] = 0; r.pad[1] = 0;
}

// Pads a primitive box (padMode 1: the oracle's rule for triangles, pad = max(max|coord| / 16384,
// 1e-6), so a box never culls a triangle the exact test accepts) and returns its Morton key.
HD uint64_t padBoxAndKey(Box &b, V3 sceneLo, V3 sceneHi, int padMode, uint64_t keyMask)
{
    if (padMode == 1)
    {
        const float maxAbs = fmaxf_(fmaxf_(fmaxf_(fabsf(sceneLo.x), fabsf(sceneHi.x)), fmaxf_(fabsf(sceneLo.y), fabsf(sceneHi.y))),
                                    fmaxf_(fabsf(sceneLo.z), fabsf(sceneHi.z)));
        const float pad = fmaxf_(maxAbs * (1.0f / 16384.0f), 1e-6f);
        for (int a = 0; a < 3; ++a) { b.lo[a] -= pad; b.hi[a] += pad; }
    }
    const V3 c = (boxLo(b) + boxHi(b)) * 0.5f;
    return mortonKey(c, sceneLo, sceneHi) & keyMask;
}

// One TLAS instance (World.cpp:878-928): exact world->object matrix + padded world box of the BLAS root.
HD void tlasInstanceRecord(const TlasInput &ti, InstRec &r, V3 &lo, V3 &hi)
{
    invertTransformExact(ti.transform, r.w2o);
    r.nodes = ti.nodes;
    r.tris = ti.tris;
    r.customIndex = ti.customIndexAndMask & 0xFFFFFFu;
    r.mask = ti.customIndexAndMask >> 24;
    r.index = ti.index;
    r.pad = 0;
    r.uvs = ti.uvs;
    r.pad2 = 0;
    WideNode root;
    loadNode(ti.nodes, root);
    const Box ob = wideNodeBounds(root);
    lo = {3.4e38f, 3.4e38f, 3.4e38f};
    hi = {-3.4e38f, -3.4e38f, -3.4e38f};
    const float *m = ti.transform;
    for (int c = 0; c < 8; ++c)
    {
        const V3 p = {(c & 1) ? ob.hi[0] : ob.lo[0], (c & 2) ? ob.hi[1] : ob.lo[1], (c & 4) ? ob.hi[2] : ob.lo[2]};
        const V3 w = {m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3], m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7],
                      m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]};
        lo = vmin(lo, w);
        hi = vmax(hi, w);
    }
    const V3 e = hi - lo;
    const float pad = fmaxf_(fmaxf_(e.x, fmaxf_(e.y, e.z)) * (1.0f / 16384.0f), 1e-6f);
    lo = lo - V3{pad, pad, pad};
    hi = hi + V3{pad, pad, pad};
}

} // namespace prt
