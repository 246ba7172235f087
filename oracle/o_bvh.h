// oracle/o_bvh.h -- TEST INFRASTRUCTURE (CPU oracle).
// The oracle's own acceleration structure: a plain binary BVH (binned SAH, 16 bins) per model
// and one over the instances. It stands in for the Vulkan driver's opaque BLAS/TLAS
// (World.cpp:740, 798). It shares no code with the product's GPU builder; the two only have
// to agree on the *results* of traversal, which the exact contract below makes independent of
// the tree shape:
//   * ray/triangle test = intersectTriangle() (Moeller-Trumbore, every operation one IEEE
//     binary32 rounding, fixed association),
//   * closest hit = smallest t, ties broken by the smallest (instance, geometry, primitive),
//   * boxes are padded and the slab test is widened so a box never culls a triangle the
//     triangle test would accept.
#pragma once
#include "o_math.h"
#include <vector>
#include <cfloat>

namespace orc {

struct Aabb
{
    vec3 lo{FLT_MAX, FLT_MAX, FLT_MAX};
    vec3 hi{-FLT_MAX, -FLT_MAX, -FLT_MAX};
    void grow(vec3 p) { lo = vmin(lo, p); hi = vmax(hi, p); }
    void grow(const Aabb &b) { lo = vmin(lo, b.lo); hi = vmax(hi, b.hi); }
    float halfArea() const
    {
        const vec3 e = hi - lo;
        return e.x * e.y + e.y * e.z + e.z * e.x;
    }
    vec3 centre() const { return (lo + hi) * 0.5f; }
    bool valid() const { return lo.x <= hi.x; }
};

struct BvhNode
{
    Aabb box;
    uint32_t left = 0;  // internal: index of left child (right = left + 1); leaf: first item
    uint32_t count = 0; // 0 = internal, else number of items
};

struct Bvh
{
    std::vector<BvhNode> nodes;
    std::vector<uint32_t> items; // permutation of the input boxes
    Aabb bounds() const { return nodes.empty() ? Aabb{} : nodes[0].box; }
};

inline float axisOf(vec3 v, int a) { return a == 0 ? v.x : (a == 1 ? v.y : v.z); }

inline void buildRecursive(Bvh &bvh, const std::vector<Aabb> &boxes, uint32_t nodeIndex, uint32_t first,
                           uint32_t count, uint32_t maxLeaf)
{
    BvhNode &node0 = bvh.nodes[nodeIndex];
    Aabb box, cbox;
    for (uint32_t i = first; i < first + count; ++i)
    {
        box.grow(boxes[bvh.items[i]]);
        cbox.grow(boxes[bvh.items[i]].centre());
    }
    node0.box = box;
    if (count <= maxLeaf)
    {
        node0.left = first;
        node0.count = count;
        return;
    }
    constexpr int BINS = 16;
    int bestAxis = -1, bestBin = 0;
    float bestCost = FLT_MAX;
    for (int a = 0; a < 3; ++a)
    {
        const float lo = axisOf(cbox.lo, a), hi = axisOf(cbox.hi, a);
        if (!(hi > lo)) continue;
        Aabb bb[BINS];
        uint32_t bc[BINS] = {};
        const float scale = BINS / (hi - lo);
        for (uint32_t i = first; i < first + count; ++i)
        {
            const Aabb &b = boxes[bvh.items[i]];
            int bin = (int)((axisOf(b.centre(), a) - lo) * scale);
            bin = std::min(std::max(bin, 0), BINS - 1);
            bb[bin].grow(b);
            bc[bin]++;
        }
        float rightArea[BINS];
        uint32_t rightCount[BINS];
        Aabb acc;
        uint32_t n = 0;
        for (int i = BINS - 1; i > 0; --i)
        {
            if (bc[i]) acc.grow(bb[i]);
            n += bc[i];
            rightArea[i] = acc.valid() ? acc.halfArea() : 0.f;
            rightCount[i] = n;
        }
        acc = Aabb{};
        n = 0;
        for (int i = 0; i < BINS - 1; ++i)
        {
            if (bc[i]) acc.grow(bb[i]);
            n += bc[i];
            if (n == 0 || rightCount[i + 1] == 0) continue;
            const float cost = acc.halfArea() * n + rightArea[i + 1] * rightCount[i + 1];
            if (cost < bestCost) { bestCost = cost; bestAxis = a; bestBin = i; }
        }
    }
    uint32_t mid;
    if (bestAxis < 0)
        mid = first + count / 2; // all centres coincide: split the list in half
    else
    {
        const float lo = axisOf(cbox.lo, bestAxis), hi = axisOf(cbox.hi, bestAxis);
        const float scale = BINS / (hi - lo);
        auto it = std::partition(bvh.items.begin() + first, bvh.items.begin() + first + count, [&](uint32_t id) {
            int bin = (int)((axisOf(boxes[id].centre(), bestAxis) - lo) * scale);
            bin = std::min(std::max(bin, 0), BINS - 1);
            return bin <= bestBin;
        });
        mid = (uint32_t)(it - bvh.items.begin());
        if (mid == first || mid == first + count) mid = first + count / 2;
    }
    const uint32_t left = (uint32_t)bvh.nodes.size();
    bvh.nodes.emplace_back();
    bvh.nodes.emplace_back();
    bvh.nodes[nodeIndex].left = left;
    bvh.nodes[nodeIndex].count = 0;
    buildRecursive(bvh, boxes, left, first, mid - first, maxLeaf);
    buildRecursive(bvh, boxes, left + 1, mid, first + count - mid, maxLeaf);
}

inline Bvh buildBvh(const std::vector<Aabb> &boxes, uint32_t maxLeaf)
{
    Bvh bvh;
    if (boxes.empty()) return bvh;
    bvh.items.resize(boxes.size());
    for (uint32_t i = 0; i < boxes.size(); ++i) bvh.items[i] = i;
    bvh.nodes.reserve(boxes.size() * 2);
    bvh.nodes.emplace_back();
    buildRecursive(bvh, boxes, 0, 0, (uint32_t)boxes.size(), maxLeaf);
    return bvh;
}

// Conservative slab test. Returns true when [tNear, tFar] (widened by a few ulp) overlaps
// [tMin, tMax]. NaNs from 0 * inf drop out of fminf/fmaxf, which keeps the test conservative.
inline bool slab(const Aabb &b, vec3 o, vec3 invD, float tMin, float tMax, float &tNearOut)
{
    const float x0 = (b.lo.x - o.x) * invD.x, x1 = (b.hi.x - o.x) * invD.x;
    const float y0 = (b.lo.y - o.y) * invD.y, y1 = (b.hi.y - o.y) * invD.y;
    const float z0 = (b.lo.z - o.z) * invD.z, z1 = (b.hi.z - o.z) * invD.z;
    float tn = std::fmax(std::fmax(std::fmin(x0, x1), std::fmin(y0, y1)), std::fmax(std::fmin(z0, z1), tMin));
    float tf = std::fmin(std::fmin(std::fmax(x0, x1), std::fmax(y0, y1)), std::fmin(std::fmax(z0, z1), tMax));
    tf = tf + std::fabs(tf) * 4e-7f;
    tn = tn - std::fabs(tn) * 4e-7f;
    tNearOut = tn;
    return tn <= tf;
}

// EXACT CONTRACT: Moeller-Trumbore, two-sided, no epsilon; the hit
// interval is open (tMin < t < tMax) like VK_KHR_ray_tracing triangle intersection.
// Returns barycentrics of v1 and v2, matching the rchit hit attribute (scene.rchit:8,14).
inline bool intersectTriangle(vec3 o, vec3 d, float tMin, float tMax, vec3 v0, vec3 v1, vec3 v2, float &t, float &u,
                              float &v)
{
    const vec3 e1 = v1 - v0;
    const vec3 e2 = v2 - v0;
    const vec3 P = cross(d, e2);
    const float det = dot(e1, P);
    if (det == 0.0f) return false;
    const float inv = 1.0f / det;
    const vec3 T = o - v0;
    const float uu = dot(T, P) * inv;
    if (!(uu >= 0.0f) || uu > 1.0f) return false;
    const vec3 Q = cross(T, e1);
    const float vv = dot(d, Q) * inv;
    if (!(vv >= 0.0f) || uu + vv > 1.0f) return false;
    const float tt = dot(e2, Q) * inv;
    if (!(tt > tMin) || !(tt < tMax)) return false;
    t = tt;
    u = uu;
    v = vv;
    return true;
}

// EXACT CONTRACT: inverse of the affine 3x4 (row-major rows r0,r1,r2 = [a b c | t]) through
// cofactors, each entry one division by det; translation = -(inv3x3 * t) with dot() association.
inline void invertTransform(const float m[12], float out[12])
{
    const float a = m[0], b = m[1], c = m[2];
    const float d = m[4], e = m[5], f = m[6];
    const float g = m[8], h = m[9], i = m[10];
    const float c00 = (e * i) - (f * h);
    const float c01 = (f * g) - (d * i);
    const float c02 = (d * h) - (e * g);
    const float det = ((a * c00) + (b * c01)) + (c * c02);
    const float i00 = c00 / det, i01 = ((c * h) - (b * i)) / det, i02 = ((b * f) - (c * e)) / det;
    const float i10 = c01 / det, i11 = ((a * i) - (c * g)) / det, i12 = ((c * d) - (a * f)) / det;
    const float i20 = c02 / det, i21 = ((b * g) - (a * h)) / det, i22 = ((a * e) - (b * d)) / det;
    const float tx = m[3], ty = m[7], tz = m[11];
    out[0] = i00; out[1] = i01; out[2] = i02;
    out[3] = -(((i00 * tx) + (i01 * ty)) + (i02 * tz));
    out[4] = i10; out[5] = i11; out[6] = i12;
    out[7] = -(((i10 * tx) + (i11 * ty)) + (i12 * tz));
    out[8] = i20; out[9] = i21; out[10] = i22;
    out[11] = -(((i20 * tx) + (i21 * ty)) + (i22 * tz));
}

// EXACT CONTRACT: world -> object ray. The direction is NOT renormalised, so t is the same
// parameter in both spaces (Vulkan semantics).
inline vec3 transformPoint(const float w2o[12], vec3 p)
{
    return {(((w2o[0] * p.x) + (w2o[1] * p.y)) + (w2o[2] * p.z)) + w2o[3],
            (((w2o[4] * p.x) + (w2o[5] * p.y)) + (w2o[6] * p.z)) + w2o[7],
            (((w2o[8] * p.x) + (w2o[9] * p.y)) + (w2o[10] * p.z)) + w2o[11]};
}
inline vec3 transformVector(const float w2o[12], vec3 d)
{
    return {((w2o[0] * d.x) + (w2o[1] * d.y)) + (w2o[2] * d.z), ((w2o[4] * d.x) + (w2o[5] * d.y)) + (w2o[6] * d.z),
            ((w2o[8] * d.x) + (w2o[9] * d.y)) + (w2o[10] * d.z)};
}

} // namespace orc
