// core_bvh.h -- acceleration-structure data layout and the per-thread bodies of the builder kernels.
//
// Replaces what the Vulkan driver does behind vkCmdBuildAccelerationStructuresKHR for prosper
// (BLAS: src/scene/World.cpp:585-747; TLAS, rebuilt every frame: World.cpp:749-802, 878-961).
// Pipeline (kernels_build.cuh wires these bodies to launches): primitive boxes -> a binary hierarchy (BLAS: top-down
// binned SAH; TLAS: 63-bit Morton keys -> radix sort -> Karras hierarchy -> bottom-up refit) -> surface-area-driven
// collapse into 8-wide nodes -> quantised 80-byte nodes + primitive records in node order.
#pragma once
#include <cstddef>
#include "hd.h"

namespace prt {

#if defined(__CUDA_ARCH__)
HD uint32_t atomAdd(uint32_t *p, uint32_t v) { return atomicAdd(p, v); }
HD void atomMin(uint32_t *p, uint32_t v) { atomicMin(p, v); }
HD void atomMax(uint32_t *p, uint32_t v) { atomicMax(p, v); }
HD void memFence() { __threadfence(); }
#else
HD uint32_t atomAdd(uint32_t *p, uint32_t v) { const uint32_t o = *p; *p = o + v; return o; }
HD void atomMin(uint32_t *p, uint32_t v) { if (v < *p) *p = v; }
HD void atomMax(uint32_t *p, uint32_t v) { if (v > *p) *p = v; }
HD void memFence() {}
#endif

struct Box
{
    float lo[3];
    float hi[3];
};
HD V3 boxLo(const Box &b) { return {b.lo[0], b.lo[1], b.lo[2]}; }
HD V3 boxHi(const Box &b) { return {b.hi[0], b.hi[1], b.hi[2]}; }
HD Box makeBox(V3 lo, V3 hi) { return Box{{lo.x, lo.y, lo.z}, {hi.x, hi.y, hi.z}}; }
HD Box boxUnion(const Box &a, const Box &b) { return makeBox(vmin(boxLo(a), boxLo(b)), vmax(boxHi(a), boxHi(b))); }
HD Box loadBoxL2(const Box *p)
{
#if defined(__CUDA_ARCH__)
    Box b;
    for (int k = 0; k < 3; ++k) { b.lo[k] = __ldcg(&p->lo[k]); b.hi[k] = __ldcg(&p->hi[k]); }
    return b;
#else
    return *p;
#endif
}
HD float boxHalfArea(const Box &b)
{
    const float ex = b.hi[0] - b.lo[0], ey = b.hi[1] - b.lo[1], ez = b.hi[2] - b.lo[2];
    return ex * ey + ey * ez + ez * ex;
}

// Compressed 8-wide node, 80 bytes (Ylitie/Karras/Laine 2017 layout) stored at a 128-byte stride: a node never straddles
// an L1 line and is fetched with two 256-bit loads and one 128-bit load (three L1 wavefronts per lane instead of the
// five to seven of five 128-bit loads at a 16-byte alignment -- the L1 wavefront queue is a co-bottleneck of the trace
// kernels).
// Child boxes are quantised to 8 bits per plane on a per-node grid: plane = p + q * 2^(e-127).
// meta[i]: 0 = empty slot; internal child: (1 << 5) | (24 + i); leaf: (unary count << 5) | first
// primitive offset relative to primBase (offset + count <= 24). imask bit i = slot i is internal;
// internal children are stored consecutively from childBase in slot order.
constexpr int WIDE_NODE_BYTES = 80; // payload; sizeof(WideNode) is the 128-byte stride
struct alignas(128) WideNode
{
    float px, py, pz;
    uint8_t ex, ey, ez, imask;
    uint32_t childBase;
    uint32_t primBase;
    uint8_t meta[8];
    uint8_t qlo[3][8];
    uint8_t qhi[3][8];
};
static_assert(sizeof(WideNode) == 128 && offsetof(WideNode, qhi) + 24 == WIDE_NODE_BYTES, "WideNode: 80 bytes of payload at a 128-byte stride");

// Triangle record in node order, 48 bytes at a 64-byte stride = one 256-bit and one 128-bit load inside one L1 line. e1 = v1 - v0 and e2 = v2 - v0 are
// the same single-rounding subtractions the exact ray/triangle test would do, hoisted to build time.
constexpr int TRI_REC_BYTES = 48; // payload; sizeof(TriRec) is the 64-byte stride
struct alignas(64) TriRec
{
    float v0[3];
    uint32_t primitive; // gl_PrimitiveID
    float e1[3];
    uint32_t geometry; // gl_GeometryIndexEXT | opaque << 31
    float e2[3];
    uint32_t pad;
};
static_assert(sizeof(TriRec) == 64, "TriRec: 48 bytes of payload at a 64-byte stride");

// Any-hit record of one triangle, parallel to the TriRec array (same index): the three texture coordinates the
// any-hit stage would otherwise fetch through the draw instance, the mesh metadata and the index buffer
// (scene.rahit:24-30 loadUV x 3). Filled for non-opaque geometry only. 32 bytes = two 128-bit loads.
struct alignas(16) AlphaUv
{
    float uv[6]; // uv0.xy, uv1.xy, uv2.xy: the fp16x2 attributes widened to float (exact)
    uint32_t pad[2];
};
static_assert(sizeof(AlphaUv) == 32, "AlphaUv");

// Instance record in TLAS node order, 96 bytes = six 128-bit loads.
struct alignas(16) InstRec
{
    float w2o[12];          // world -> object, row-major 3x4 (exact contract inverse)
    const WideNode *nodes;  // BLAS root
    const TriRec *tris;
    uint32_t customIndex;   // first DrawInstance of the model instance (World.cpp:906-913)
    uint32_t mask;
    uint32_t index;         // position in the caller's instance array (tie-break key)
    uint32_t pad;
    const AlphaUv *uvs;     // any-hit records of the BLAS (nullptr: fetch uvs through the scene buffers)
    uint64_t pad2;
};
static_assert(sizeof(InstRec) == 96, "InstRec");

// ---- Morton keys ---------------------------------------------------------------------------------
HD uint64_t expandBits21(uint32_t v)
{
    uint64_t x = v & 0x1FFFFFu;
    x = (x | (x << 32)) & 0x001F00000000FFFFull;
    x = (x | (x << 16)) & 0x001F0000FF0000FFull;
    x = (x | (x << 8)) & 0x100F00F00F00F00Full;
    x = (x | (x << 4)) & 0x10C30C30C30C30C3ull;
    x = (x | (x << 2)) & 0x1249249249249249ull;
    return x;
}

HD uint64_t mortonKey(V3 c, V3 lo, V3 hi)
{
    uint32_t q[3];
    const float cc[3] = {c.x, c.y, c.z}, l[3] = {lo.x, lo.y, lo.z}, h[3] = {hi.x, hi.y, hi.z};
    for (int a = 0; a < 3; ++a)
    {
        const float ext = h[a] - l[a];
        float t = ext > 0.f ? (cc[a] - l[a]) / ext : 0.f;
        t = fminf_(fmaxf_(t, 0.f), 1.f) * 2097152.0f;
        const uint32_t v = (uint32_t)t;
        q[a] = v > 2097151u ? 2097151u : v;
    }
    return (expandBits21(q[0]) << 2) | (expandBits21(q[1]) << 1) | expandBits21(q[2]);
}

// ---- binary hierarchy over the sorted keys (Karras 2012) ------------------------------------------
// Node references: ref >= 0 is internal node `ref`; ref < 0 is the leaf at sorted position ~ref.
struct Lbvh
{
    uint32_t count;       // primitives
    const uint64_t *keys; // sorted
    const uint32_t *order; // sorted position -> primitive index
    const Box *primBox;   // by primitive index
    int32_t *left;        // [count-1]
    int32_t *right;
    uint32_t *first;      // sorted range covered by the internal node
    uint32_t *last;
    int32_t *parentOfInternal; // [count-1], -1 for the root
    int32_t *parentOfLeaf;     // [count]
    Box *nodeBox;              // [count-1]
    uint32_t *visit;           // [count-1], zeroed
};

HD int lbvhDelta(const Lbvh &t, int i, int j)
{
    if (j < 0 || j >= (int)t.count) return -1;
    const uint64_t a = t.keys[i], b = t.keys[j];
    if (a == b) return 64 + (31 - bfind((uint32_t)i ^ (uint32_t)j)); // tie-break on position
    return clz64(a ^ b);
}

HD void lbvhBuildNode(const Lbvh &t, int i)
{
    const int d = (lbvhDelta(t, i, i + 1) - lbvhDelta(t, i, i - 1)) < 0 ? -1 : 1;
    const int dmin = lbvhDelta(t, i, i - d);
    int lmax = 2;
    while (lbvhDelta(t, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int s = lmax / 2; s >= 1; s /= 2)
        if (lbvhDelta(t, i, i + (l + s) * d) > dmin) l += s;
    const int j = i + l * d;
    const int dnode = lbvhDelta(t, i, j);
    int s = 0;
    int step = l;
    do
    {
        step = (step + 1) >> 1;
        if (lbvhDelta(t, i, i + (s + step) * d) > dnode) s += step;
    } while (step > 1);
    const int gamma = i + s * d + (d < 0 ? d : 0);
    const int lo = i < j ? i : j, hi = i < j ? j : i;
    t.first[i] = (uint32_t)lo;
    t.last[i] = (uint32_t)hi;
    if (lo == gamma) { t.left[i] = ~gamma; t.parentOfLeaf[gamma] = i; }
    else { t.left[i] = gamma; t.parentOfInternal[gamma] = i; }
    if (hi == gamma + 1) { t.right[i] = ~(gamma + 1); t.parentOfLeaf[gamma + 1] = i; }
    else { t.right[i] = gamma + 1; t.parentOfInternal[gamma + 1] = i; }
    if (i == 0) t.parentOfInternal[0] = -1;
}

HD Box lbvhRefBox(const Lbvh &t, int32_t ref) { return ref >= 0 ? t.nodeBox[ref] : t.primBox[t.order[~ref]]; }
HD uint32_t lbvhRefCount(const Lbvh &t, int32_t ref) { return ref >= 0 ? t.last[ref] - t.first[ref] + 1u : 1u; }
HD uint32_t lbvhRefFirst(const Lbvh &t, int32_t ref) { return ref >= 0 ? t.first[ref] : (uint32_t)~ref; }

// One thread per leaf walks to the root; the second arrival at a node owns it (atomic visit flag).
HD void lbvhRefitFromLeaf(const Lbvh &t, uint32_t leaf)
{
    int32_t node = t.parentOfLeaf[leaf];
    while (node >= 0)
    {
        memFence();
        if (atomAdd(&t.visit[node], 1u) == 0u) return;
        memFence();
        // the sibling's box was written by another thread: read it around L1
        const Box a = t.left[node] >= 0 ? loadBoxL2(&t.nodeBox[t.left[node]]) : t.primBox[t.order[~t.left[node]]];
        const Box b = t.right[node] >= 0 ? loadBoxL2(&t.nodeBox[t.right[node]]) : t.primBox[t.order[~t.right[node]]];
        t.nodeBox[node] = boxUnion(a, b);
        node = t.parentOfInternal[node];
    }
}

// ---- PLOC: parallel locally-ordered clustering (Meister & Bittner 2018) --------------------------------
// Bottom-up agglomerative build over the Morton-sorted primitives: every cluster looks `radius` positions to
// both sides for the neighbour that minimises the surface area of the merged box, mutual nearest neighbours
// merge, the cluster array is compacted (order preserved) and the search repeats until one cluster is left.
// Produces the same binary-tree arrays as the Karras builder (left/right/parents/nodeBox, root = node 0) with a
// surface-area-driven topology; plocFinish* then renumber the leaves in depth-first order so that every subtree
// covers a contiguous range [first, last] again, which is what the collapse wants.
struct Ploc
{
    Lbvh t;              // t.visit holds the subtree sizes while building; t.first/last are written at the end
    const int32_t *refIn; // clusters of this iteration, in spatial order
    const Box *boxIn;
    int32_t *refOut;      // compacted clusters of the next iteration
    Box *boxOut;
    uint32_t *nearest;    // [count]
    int32_t *mergedRef;   // [count] result of the merge step before compaction (PLOC_DROPPED = slot vanishes)
    Box *mergedBox;
    uint32_t *nodeAlloc;  // internal nodes created so far
    uint32_t radius;
};
constexpr int32_t PLOC_DROPPED = 0x7FFFFFFF;

HD uint32_t plocRefSize(const Lbvh &t, int32_t ref) { return ref >= 0 ? t.visit[ref] : 1u; }

HD void plocNearest(const Ploc &p, uint32_t i, uint32_t m)
{
    const Box b = p.boxIn[i];
    const uint32_t lo = i > p.radius ? i - p.radius : 0u, hi = i + p.radius < m ? i + p.radius : m - 1u;
    float best = 3.4e38f;
    uint32_t bj = i;
    for (uint32_t j = lo; j <= hi; ++j)
    {
        if (j == i) continue;
        const float a = boxHalfArea(boxUnion(b, p.boxIn[j]));
        if (a < best) { best = a; bj = j; } // ties keep the lower j: the globally best pair is always mutual
    }
    p.nearest[i] = bj;
}

HD void plocMerge(const Ploc &p, uint32_t i, uint32_t m)
{
    const uint32_t j = p.nearest[i];
    if (m > 1u && p.nearest[j] == i)
    {
        if (i > j) { p.mergedRef[i] = PLOC_DROPPED; return; }
        const Lbvh &t = p.t;
        // nodes are numbered downwards so that the last merge (the root) is node 0
        const int32_t node = (int32_t)(t.count - 2u - atomAdd(p.nodeAlloc, 1u));
        const int32_t l = p.refIn[i], r = p.refIn[j];
        t.left[node] = l; t.right[node] = r;
        if (l >= 0) t.parentOfInternal[l] = node; else t.parentOfLeaf[~l] = node;
        if (r >= 0) t.parentOfInternal[r] = node; else t.parentOfLeaf[~r] = node;
        t.parentOfInternal[node] = -1;
        const Box u = boxUnion(p.boxIn[i], p.boxIn[j]);
        t.nodeBox[node] = u;
        t.visit[node] = plocRefSize(t, l) + plocRefSize(t, r);
        p.mergedRef[i] = node;
        p.mergedBox[i] = u;
        return;
    }
    p.mergedRef[i] = p.refIn[i];
    p.mergedBox[i] = p.boxIn[i];
}

// Start of the depth-first leaf range of `ref`: the sizes of all left siblings on the way to the root.
HD uint32_t plocRangeStart(const Lbvh &t, int32_t ref)
{
    uint32_t start = 0;
    int32_t node = ref >= 0 ? t.parentOfInternal[ref] : t.parentOfLeaf[~ref];
    while (node >= 0)
    {
        if (t.right[node] == ref) start += plocRefSize(t, t.left[node]);
        ref = node;
        node = t.parentOfInternal[node];
    }
    return start;
}
// one call per old sorted position: its depth-first position and the re-ordered primitive index
HD void plocFinishLeaf(const Lbvh &t, uint32_t oldPos, uint32_t *newPos, uint32_t *orderOut)
{
    const uint32_t pos = plocRangeStart(t, ~(int32_t)oldPos);
    newPos[oldPos] = pos;
    orderOut[pos] = t.order[oldPos];
}
// one call per internal node, after every plocFinishLeaf: range + leaf references in the new numbering
HD void plocFinishNode(const Lbvh &t, uint32_t node, const uint32_t *newPos)
{
    const uint32_t start = plocRangeStart(t, (int32_t)node);
    t.first[node] = start;
    t.last[node] = start + t.visit[node] - 1u;
}
HD void plocRelabelNode(const Lbvh &t, uint32_t node, const uint32_t *newPos)
{
    if (t.left[node] < 0) t.left[node] = ~(int32_t)newPos[~t.left[node]];
    if (t.right[node] < 0) t.right[node] = ~(int32_t)newPos[~t.right[node]];
}

// ---- binned SAH: top-down, level-synchronous (Wald 2007 binning, one level of the whole tree per pass) --------
// The default BLAS builder (PRT_CFG_LBVH_BLAS and PRT_CFG_PLOC select the others). Every level runs four per-slot passes over the primitive index array:
// centroid bounds per active range -> SAH_BINS bins along each axis (box + count per bin) -> the
// range's first slot sweeps the bins for the cheapest split (area_L * n_L + area_R * n_R) and creates the two child
// ranges -> partition into the children's sub-ranges (ping-pong arrays). A range is named by key = first >> 1
// (ranges of a level are disjoint and hold >= 2 primitives, so keys are unique); an internal node is named by the
// gap it splits (first + leftCount - 1, a bijection between internal nodes and the n - 1 gaps between leaves),
// with the root's gap and 0 swapped so that the root is node 0 as the collapse expects. Neither needs an allocator,
// so tree and numbering do not depend on atomic order. Fills the same arrays as the Karras builder, nodeBox
// included (the union of a range's bin boxes), so no refit pass follows. Ranges whose centroids coincide, and every
// range from level SAH_MEDIAN_LEVEL on, are binned by primitive index and split at the most balanced border.
#ifndef PRT_SAH_BINS
#define PRT_SAH_BINS 64 // measured on the atrium: 8 / 16 / 32 / 64 bins trace at 884 / 867 / 890 / 896 Mrays/s
#endif
#ifndef PRT_SAH_AXES
#define PRT_SAH_AXES 1 // 1: bin along the longest centroid axis; 3: along every axis, cheapest split of all (measured: no faster to trace)
#endif
#ifndef PRT_SAH_GRANULE
#define PRT_SAH_GRANULE 1 // primitives per unit of leaf cost in the SAH sum (experiment: 3 = one leaf slot of the wide node)
#endif
constexpr uint32_t SAH_GRANULE = PRT_SAH_GRANULE;
constexpr uint32_t SAH_BINS = PRT_SAH_BINS;
constexpr uint32_t SAH_AXES = PRT_SAH_AXES;
constexpr uint32_t SAH_RANGE_BINS = SAH_BINS * SAH_AXES; // most bins one range uses: [axis slot][sahBinsFor(count)]
constexpr uint32_t SAH_SLOT_BINS = 2u * SAH_AXES;        // bins reserved per slot: a range owns [first, first + count) * SAH_SLOT_BINS
constexpr uint32_t SAH_FINAL = 0xFFFFFFFFu;   // slot key: the slot is a finished leaf
constexpr uint32_t SAH_MEDIAN = 0xFFFFFFFFu;  // SahSplit::axisBin: halve by position (safety net, see sahSplitPass)
constexpr uint32_t SAH_INDEX_AXIS = 3u;       // axis code: the primitive index (coincident centroids, depth bound)
constexpr uint32_t SAH_CENT = 8u;             // words of centroid bounds per range: lo xyz, hi xyz, min / max primitive index
constexpr uint32_t SAH_MEDIAN_LEVEL = 48;
constexpr uint32_t SAH_MAX_LEVELS = 96;

// monotone float <-> uint map, so that bounds are plain atomicMin / atomicMax
HD uint32_t ordF(float f) { const uint32_t u = f2u(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
HD float unordF(uint32_t o) { return u2f((o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o); }

struct SahBin // 32 B; planes in ordF form; empty: lo = 0xFFFFFFFF, hi = 0, count = 0
{
    uint32_t lo[3], hi[3], count, pad;
};
struct SahRange
{
    uint32_t first, count;
    int32_t parent;   // internal node above, -1 for the root range
    uint32_t isRight;
};
struct SahSplit // what the partition pass needs of a range, written by its first slot
{
    uint32_t first, count, leftCount, axisBin; // axisBin = axis << 8 | first bin of the right side, or SAH_MEDIAN
    float cmin, scale;
    uint32_t cursorL, cursorR;
};
struct Sah
{
    Lbvh t; // outputs; t.order is set by the caller to the final index array
    const uint32_t *idxIn;
    uint32_t *idxOut;
    const uint32_t *keyIn; // slot -> key of the active range holding it, or SAH_FINAL
    uint32_t *keyOut;
    const SahRange *rangeIn;
    SahRange *rangeOut;
    SahSplit *split;  // [keys]
    SahBin *bins;     // [count * SAH_SLOT_BINS]; a range's bins start at first * SAH_SLOT_BINS (sahBinsFor(count) <= 2 * count)
    uint32_t *cent;   // [keys][SAH_CENT] centroid bounds in ordF form + primitive index bounds
    uint32_t *active; // [SAH_MAX_LEVELS + 1] ranges created by each level
    uint32_t *rootGap;
    uint32_t level;
};

// the binned axes of a range: axis slot k bins along axis[k] with (c - cmin[k]) * scale[k]; scale 0 = no extent
// byIndex: the range is binned by primitive index instead (its centroids coincide, or the level is past
// SAH_MEDIAN_LEVEL) and split at the most balanced bin border -- an order-free rule, so the tree stays deterministic
struct SahAxes
{
    uint32_t axis[SAH_AXES];
    float cmin[SAH_AXES], scale[SAH_AXES];
    bool byIndex;
    uint32_t idxMin, idxSpan;
};
// Bins a range of `count` primitives uses: small ranges (most ranges of the deep levels) sweep and clear only a few.
HD uint32_t sahBinsFor(uint32_t count) { return count >= SAH_BINS / 2u ? SAH_BINS : (count < 2u ? 4u : count * 2u); }
HD float sahCentroid(const Box &b, uint32_t axis) { return (b.lo[axis] + b.hi[axis]) * 0.5f; }
HD SahAxes sahAxes(const uint32_t *cent, uint32_t level, uint32_t nb)
{
    float lo[3], ext[3];
    uint32_t longest = 0;
    for (uint32_t a = 0; a < 3; ++a)
    {
        lo[a] = unordF(cent[a]);
        ext[a] = unordF(cent[3 + a]) - lo[a];
        if (ext[a] > ext[longest]) longest = a;
    }
    SahAxes r;
    for (uint32_t k = 0; k < SAH_AXES; ++k)
    {
        const uint32_t a = SAH_AXES == 3 ? k : longest;
        float scale = ext[a] > 0.0f ? (float)nb / ext[a] : 0.0f;
        if (!(scale < 3.0e38f)) scale = 0.0f; // denormal extent: treat as coincident
        r.axis[k] = a; r.cmin[k] = lo[a]; r.scale[k] = scale;
    }
    r.byIndex = level >= SAH_MEDIAN_LEVEL || !(ext[longest] > 0.0f) || !((float)nb / ext[longest] < 3.0e38f);
    r.idxMin = cent[6];
    r.idxSpan = cent[7] - cent[6] + 1u;
    return r;
}
HD uint32_t sahIndexBinOf(uint32_t prim, uint32_t idxMin, uint32_t idxSpan, uint32_t nb)
{
    return (uint32_t)(((uint64_t)(prim - idxMin) * nb) / idxSpan);
}
HD uint32_t sahBinOf(float c, float cmin, float scale, uint32_t nb)
{
    const float f = (c - cmin) * scale;
    const int b = f > 0.0f ? (int)fminf_(f, (float)(nb - 1u)) : 0;
    return (uint32_t)b;
}
HD void sahBinAdd(SahBin *dst, const Box &b)
{
    for (int a = 0; a < 3; ++a) { atomMin(&dst->lo[a], ordF(b.lo[a])); atomMax(&dst->hi[a], ordF(b.hi[a])); }
    atomAdd(&dst->count, 1u);
}
HD void sahBinMerge(SahBin *dst, const SahBin &src)
{
    for (int a = 0; a < 3; ++a) { atomMin(&dst->lo[a], src.lo[a]); atomMax(&dst->hi[a], src.hi[a]); }
    atomAdd(&dst->count, src.count);
}
HD void sahBinClear(SahBin *b)
{
    for (int a = 0; a < 3; ++a) { b->lo[a] = 0xFFFFFFFFu; b->hi[a] = 0u; }
    b->count = 0; b->pad = 0;
}
HD void sahCentClear(uint32_t *cent) { for (int a = 0; a < 3; ++a) { cent[a] = 0xFFFFFFFFu; cent[3 + a] = 0u; } cent[6] = 0xFFFFFFFFu; cent[7] = 0u; }

// pass 1, per slot
HD void sahCentroidPass(const Sah &s, uint32_t slot)
{
    const uint32_t key = s.keyIn[slot];
    if (key == SAH_FINAL) return;
    const uint32_t prim = s.idxIn[slot];
    const Box b = s.t.primBox[prim];
    for (uint32_t a = 0; a < 3; ++a)
    {
        const uint32_t o = ordF(sahCentroid(b, a));
        atomMin(&s.cent[(size_t)key * SAH_CENT + a], o);
        atomMax(&s.cent[(size_t)key * SAH_CENT + 3u + a], o);
    }
    atomMin(&s.cent[(size_t)key * SAH_CENT + 6u], prim);
    atomMax(&s.cent[(size_t)key * SAH_CENT + 7u], prim);
}
// pass 2, per slot: adds the slot's primitive to one bin per axis slot of `dst` (the range's bins or a block's copy)
HD void sahBinInto(const Sah &s, uint32_t slot, uint32_t key, uint32_t count, SahBin *dst)
{
    const uint32_t nb = sahBinsFor(count);
    const SahAxes ax = sahAxes(&s.cent[(size_t)key * SAH_CENT], s.level, nb);
    const uint32_t prim = s.idxIn[slot];
    const Box b = s.t.primBox[prim];
    if (ax.byIndex) { sahBinAdd(&dst[sahIndexBinOf(prim, ax.idxMin, ax.idxSpan, nb)], b); return; }
    for (uint32_t k = 0; k < SAH_AXES; ++k)
        sahBinAdd(&dst[k * nb + sahBinOf(sahCentroid(b, ax.axis[k]), ax.cmin[k], ax.scale[k], nb)], b);
}
HD void sahBinPass(const Sah &s, uint32_t slot)
{
    const uint32_t key = s.keyIn[slot];
    if (key == SAH_FINAL) return;
    const SahRange r = s.rangeIn[key];
    sahBinInto(s, slot, key, r.count, &s.bins[(size_t)r.first * SAH_SLOT_BINS]);
}
HD uint32_t sahNodeOfGap(uint32_t gap, uint32_t rootGap) { return gap == rootGap ? 0u : (gap == 0u ? rootGap : gap); }
HD float sahOrdHalfArea(const uint32_t *lo, const uint32_t *hi)
{
    const float ex = unordF(hi[0]) - unordF(lo[0]), ey = unordF(hi[1]) - unordF(lo[1]), ez = unordF(hi[2]) - unordF(lo[2]);
    return ex * ey + ey * ez + ez * ex;
}
// The result of a range's split: its SahSplit record, its internal node (named by the gap it splits), its two children
// (leaves or ranges of the next level). Returns the position of the first of the `made` child ranges in the next level's
// count; `childKeys` receives their keys (the cooperative builder queues them, k_sah_all_levels).
HD uint32_t sahSplitEmit(const Sah &s, uint32_t key, const SahRange &r, const SahAxes &ax, uint32_t bestBin, uint32_t bestLeft, uint32_t bestSlot,
                         const Box &nodeBox, uint32_t childKeys[2], uint32_t &made)
{
    SahSplit sp;
    sp.first = r.first; sp.count = r.count; sp.cursorL = 0; sp.cursorR = 0;
    sp.cmin = ax.byIndex ? u2f(ax.idxMin) : ax.cmin[bestSlot];
    sp.scale = ax.byIndex ? u2f(ax.idxSpan) : ax.scale[bestSlot];
    // distinct primitive indices always leave two non-empty bins, so a range of >= 2 primitives always finds a split;
    // halving by position is kept as the safety net only (it would depend on the order the partition left behind)
    if (bestLeft == 0) { sp.leftCount = r.count / 2u; sp.axisBin = SAH_MEDIAN; }
    else { sp.leftCount = bestLeft; sp.axisBin = ((ax.byIndex ? SAH_INDEX_AXIS : ax.axis[bestSlot]) << 8) | bestBin; }
    s.split[key] = sp;

    const uint32_t gap = r.first + sp.leftCount - 1u;
    if (r.parent < 0) *s.rootGap = gap;
    const uint32_t node = r.parent < 0 ? 0u : sahNodeOfGap(gap, *s.rootGap);
    const Lbvh &t = s.t;
    t.first[node] = r.first;
    t.last[node] = r.first + r.count - 1u;
    t.nodeBox[node] = nodeBox;
    t.parentOfInternal[node] = r.parent;
    if (r.parent >= 0) { if (r.isRight) t.right[r.parent] = (int32_t)node; else t.left[r.parent] = (int32_t)node; }
    made = 0;
    if (sp.leftCount == 1u) { t.left[node] = ~(int32_t)r.first; t.parentOfLeaf[r.first] = (int32_t)node; }
    else { s.rangeOut[r.first >> 1] = SahRange{r.first, sp.leftCount, (int32_t)node, 0u}; childKeys[made++] = r.first >> 1; }
    const uint32_t rf = r.first + sp.leftCount, rc = r.count - sp.leftCount;
    if (rc == 1u) { t.right[node] = ~(int32_t)rf; t.parentOfLeaf[rf] = (int32_t)node; }
    else { s.rangeOut[rf >> 1] = SahRange{rf, rc, (int32_t)node, 1u}; childKeys[made++] = rf >> 1; }
    return made ? atomAdd(&s.active[s.level + 1u], made) : 0u;
}
// pass 3, per slot; only the first slot of an active range works
HD void sahSplitPass(const Sah &s, uint32_t slot)
{
    const uint32_t key = s.keyIn[slot];
    if (key == SAH_FINAL) return;
    const SahRange r = s.rangeIn[key];
    if (slot != r.first) return;
    const uint32_t nb = sahBinsFor(r.count);
    const SahAxes ax = sahAxes(&s.cent[(size_t)key * SAH_CENT], s.level, nb);
    sahCentClear(&s.cent[(size_t)key * SAH_CENT]);
    float bestCost = 3.4e38f;
    uint32_t bestBin = 0, bestLeft = 0, bestSlot = 0;
    Box nodeBox{};
    for (uint32_t k = 0; k < (ax.byIndex ? 1u : SAH_AXES); ++k)
    {
        SahBin *bins = &s.bins[(size_t)r.first * SAH_SLOT_BINS + k * nb];
        // suffix sweep: area and count of bins [i, nb)
        float areaR[SAH_BINS];
        uint32_t countR[SAH_BINS];
        uint32_t lo[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, hi[3] = {0u, 0u, 0u}, cnt = 0;
        for (int i = (int)nb - 1; i >= 0; --i)
        {
            const SahBin b = bins[i];
            if (b.count)
            {
                for (int a = 0; a < 3; ++a) { lo[a] = b.lo[a] < lo[a] ? b.lo[a] : lo[a]; hi[a] = b.hi[a] > hi[a] ? b.hi[a] : hi[a]; }
                cnt += b.count;
            }
            areaR[i] = cnt ? sahOrdHalfArea(lo, hi) : 0.0f;
            countR[i] = cnt;
        }
        if (k == 0) // every axis slot holds all primitives of the range
            for (int a = 0; a < 3; ++a) { nodeBox.lo[a] = unordF(lo[a]); nodeBox.hi[a] = unordF(hi[a]); }
        // prefix sweep + choice (ties keep the earlier axis and bin)
        uint32_t llo[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, lhi[3] = {0u, 0u, 0u}, lcnt = 0;
        for (uint32_t i = 1; i < nb; ++i)
        {
            const SahBin b = bins[i - 1];
            if (b.count)
            {
                for (int a = 0; a < 3; ++a) { llo[a] = b.lo[a] < llo[a] ? b.lo[a] : llo[a]; lhi[a] = b.hi[a] > lhi[a] ? b.hi[a] : lhi[a]; }
                lcnt += b.count;
            }
            if (lcnt == 0 || countR[i] == 0) continue;
            const float cost = ax.byIndex ? (float)(lcnt > countR[i] ? lcnt - countR[i] : countR[i] - lcnt)
                                          : sahOrdHalfArea(llo, lhi) * (float)((lcnt + SAH_GRANULE - 1u) / SAH_GRANULE) +
                                                areaR[i] * (float)((countR[i] + SAH_GRANULE - 1u) / SAH_GRANULE);
            if (cost < bestCost) { bestCost = cost; bestBin = i; bestLeft = lcnt; bestSlot = k; }
        }
        for (uint32_t i = 0; i < nb; ++i) sahBinClear(&bins[i]);
    }
    uint32_t childKeys[2], made;
    sahSplitEmit(s, key, r, ax, bestBin, bestLeft, bestSlot, nodeBox, childKeys, made);
}
// pass 4, per slot: side of the slot's primitive (0 left, 1 right); false: the slot is final / halved by position
HD bool sahSlotSide(const Sah &s, uint32_t slot, uint32_t prim, const SahSplit &sp, uint32_t &side)
{
    if (sp.axisBin == SAH_MEDIAN) { side = (slot - sp.first) >= sp.leftCount ? 1u : 0u; return false; }
    if ((sp.axisBin >> 8) == SAH_INDEX_AXIS)
    {
        side = sahIndexBinOf(prim, f2u(sp.cmin), f2u(sp.scale), sahBinsFor(sp.count)) >= (sp.axisBin & 0xFFu) ? 1u : 0u;
        return true;
    }
    const Box b = s.t.primBox[prim];
    side = sahBinOf(sahCentroid(b, sp.axisBin >> 8), sp.cmin, sp.scale, sahBinsFor(sp.count)) >= (sp.axisBin & 0xFFu) ? 1u : 0u;
    return true;
}
HD void sahPlace(const Sah &s, uint32_t pos, uint32_t prim, const SahSplit &sp, uint32_t side)
{
    const uint32_t childFirst = side ? sp.first + sp.leftCount : sp.first;
    const uint32_t childCount = side ? sp.count - sp.leftCount : sp.leftCount;
    s.idxOut[pos] = prim;
    s.keyOut[pos] = childCount >= 2u ? (childFirst >> 1) : SAH_FINAL;
}
HD void sahPartitionPass(const Sah &s, uint32_t slot)
{
    const uint32_t key = s.keyIn[slot];
    const uint32_t prim = s.idxIn[slot];
    if (key == SAH_FINAL) { s.idxOut[slot] = prim; s.keyOut[slot] = SAH_FINAL; return; }
    SahSplit *spp = &s.split[key];
    const SahSplit sp = *spp;
    uint32_t side, pos = slot;
    if (sahSlotSide(s, slot, prim, sp, side))
        pos = side ? sp.first + sp.leftCount + atomAdd(&spp->cursorR, 1u) : sp.first + atomAdd(&spp->cursorL, 1u);
    sahPlace(s, pos, prim, sp, side);
}

// ---- collapse to 8-wide + quantise -----------------------------------------------------------------
// Work item of the level-synchronous collapse: binary subtree `ref` becomes wide node `wide`.
struct CollapseItem
{
    int32_t ref;
    uint32_t wide;
};

struct Collapse
{
    Lbvh t;
    uint32_t maxLeaf;      // primitives per leaf slot: 3 for a BLAS, 1 for the TLAS
    WideNode *nodes;       // out
    uint32_t *primOut;     // out: node-order position -> primitive index
    uint32_t *nodeCount;   // atomic allocators (nodeCount starts at 1: the root)
    uint32_t *primCount;
    const CollapseItem *queueIn;
    CollapseItem *queueOut;
    uint32_t *queueOutCount;
};

// Smallest biased exponent byte e with 255 * 2^(e-127) >= ext.
HD uint32_t gridExponent(float ext)
{
    if (!(ext > 0.f)) return 1u;
    const float s = ext / 255.0f;
    uint32_t bits = f2u(s);
    uint32_t e = bits >> 23;
    if (bits & 0x7FFFFFu) e += 1u;
    if (e < 1u) e = 1u;
    if (e > 254u) e = 254u;
    return e;
}

// Conservative 8-bit quantisation of [lo, hi] on the grid p + q * scale. Returns false when hi does
// not fit in 255 steps (caller bumps the exponent).
HD bool quantisePlanes(float p, float scale, float lo, float hi, uint32_t &qlo, uint32_t &qhi)
{
    float fl = floorf((lo - p) / scale);
    if (fl < 0.f) fl = 0.f;
    if (fl > 255.f) fl = 255.f;
    while (fl > 0.f && fl * scale + p > lo) fl -= 1.f;
    float fh = ceilf((hi - p) / scale);
    if (fh < 0.f) fh = 0.f;
    while (fh <= 255.f && fh * scale + p < hi) fh += 1.f;
    if (fh > 255.f) return false;
    qlo = (uint32_t)fl;
    qhi = (uint32_t)fh;
    return true;
}

#ifndef PRT_BOTTOM_CAP
#define PRT_BOTTOM_CAP 0
#endif

HD void collapseItem(const Collapse &c, uint32_t itemIndex)
{
#if defined(__CUDA_ARCH__)
    // written by other SMs one level earlier: read around L1
    const uint2 raw = __ldcg(reinterpret_cast<const uint2 *>(c.queueIn + itemIndex));
    const CollapseItem item = {(int32_t)raw.x, raw.y};
#else
    const CollapseItem item = c.queueIn[itemIndex];
#endif
    const Lbvh &t = c.t;
    // children of the wide node with their box, area and primitive count fetched ONCE when they enter the set
    // (the open-largest loop below used to re-read every child's box and range on every round)
    int32_t ch[8];
    Box cb[8];
    float carea[8];
    uint32_t ccnt[8];
    int n = 0;
    auto add = [&](int32_t ref) {
        ch[n] = ref;
        cb[n] = lbvhRefBox(t, ref);
        carea[n] = boxHalfArea(cb[n]);
        ccnt[n] = lbvhRefCount(t, ref);
        n++;
    };
    auto replace = [&](int i, int32_t ref) {
        ch[i] = ref;
        cb[i] = lbvhRefBox(t, ref);
        carea[i] = boxHalfArea(cb[i]);
        ccnt[i] = lbvhRefCount(t, ref);
    };
    if (item.ref >= 0 && lbvhRefCount(t, item.ref) > c.maxLeaf)
    {
        add(t.left[item.ref]);
        add(t.right[item.ref]);
    }
    else
        add(item.ref); // a tree that is a single leaf (count <= maxLeaf at the root)
    // phase 1: open the largest-area child that is too big for a leaf slot; phase 2: use spare slots
    // to split multi-primitive leaves (tighter boxes, same memory)
#if PRT_BOTTOM_CAP
    // a subtree that fits one wide node's leaf storage is kept whole (it becomes ONE bottom node) instead of
    // being opened from above into several sparse ones
    const uint32_t total = item.ref >= 0 ? lbvhRefCount(t, item.ref) : 1u;
    const uint32_t cap = (uint32_t)PRT_BOTTOM_CAP * c.maxLeaf;
    const uint32_t openAbove = total > cap ? cap : c.maxLeaf;
#else
    const uint32_t openAbove = c.maxLeaf;
#endif
    for (int phase = 0; phase < 2; ++phase)
    {
        while (n < 8)
        {
            int best = -1;
            float bestArea = -1.f;
            for (int i = 0; i < n; ++i)
            {
                if (ch[i] < 0) continue;
                if (phase == 0 && !(ccnt[i] > openAbove)) continue;
                if (phase == 1 && ccnt[i] > c.maxLeaf) continue;
                if (carea[i] > bestArea) { bestArea = carea[i]; best = i; }
            }
            if (best < 0) break;
            const int32_t r = ch[best];
            const int32_t l2 = t.left[r], r2 = t.right[r];
            replace(best, l2);
            add(r2);
        }
    }
    // node bounds
    Box nb = cb[0];
    for (int i = 1; i < n; ++i) nb = boxUnion(nb, cb[i]);
    // slot assignment (greedy on the Ylitie cost: child offset projected on the slot's octant diagonal)
    const V3 nc = (boxLo(nb) + boxHi(nb)) * 0.5f;
    float off[8][3];
    for (int i = 0; i < n; ++i)
    {
        const V3 cc = (boxLo(cb[i]) + boxHi(cb[i])) * 0.5f - nc;
        off[i][0] = cc.x; off[i][1] = cc.y; off[i][2] = cc.z;
    }
    int slotOf[8];
    int childAt[8];
    for (int i = 0; i < 8; ++i) { slotOf[i] = -1; childAt[i] = -1; }
    for (int k = 0; k < n; ++k)
    {
        int bi = -1, bs = -1;
        float bc = -3.4e38f;
        for (int i = 0; i < n; ++i)
        {
            if (slotOf[i] >= 0) continue;
            for (int s = 0; s < 8; ++s)
            {
                if (childAt[s] >= 0) continue;
                const float cost = ((s & 1) ? off[i][0] : -off[i][0]) + ((s & 2) ? off[i][1] : -off[i][1]) +
                                   ((s & 4) ? off[i][2] : -off[i][2]);
                if (cost > bc) { bc = cost; bi = i; bs = s; }
            }
        }
        slotOf[bi] = bs;
        childAt[bs] = bi;
    }
    // allocate internal children and leaf primitives
    uint32_t nInternal = 0, nPrims = 0;
    for (int i = 0; i < n; ++i)
    {
        const uint32_t cnt = ccnt[i];
        if (cnt > c.maxLeaf) nInternal++;
        else nPrims += cnt;
    }
    const uint32_t childBase = nInternal ? atomAdd(c.nodeCount, nInternal) : 0u;
    const uint32_t primBase = nPrims ? atomAdd(c.primCount, nPrims) : 0u;
    const uint32_t queueBase = nInternal ? atomAdd(c.queueOutCount, nInternal) : 0u;
    WideNode node;
    node.px = nb.lo[0]; node.py = nb.lo[1]; node.pz = nb.lo[2];
    node.childBase = childBase;
    node.primBase = primBase;
    uint32_t imask = 0, internalSeen = 0, primSeen = 0;
    for (int s = 0; s < 8; ++s)
    {
        node.meta[s] = 0;
        const int i = childAt[s];
        if (i < 0) continue;
        const uint32_t cnt = ccnt[i];
        if (cnt > c.maxLeaf)
        {
            imask |= 1u << s;
            node.meta[s] = (uint8_t)((1u << 5) | (24u + (uint32_t)s));
            c.queueOut[queueBase + internalSeen] = CollapseItem{ch[i], childBase + internalSeen};
            internalSeen++;
        }
        else
        {
            const uint32_t first = lbvhRefFirst(t, ch[i]);
            for (uint32_t k = 0; k < cnt; ++k) c.primOut[primBase + primSeen + k] = t.order[first + k];
            node.meta[s] = (uint8_t)((((1u << cnt) - 1u) << 5) | primSeen);
            primSeen += cnt;
        }
    }
    node.imask = (uint8_t)imask;
    // quantisation grid
    uint32_t e[3];
    const float plo[3] = {nb.lo[0], nb.lo[1], nb.lo[2]};
    for (int a = 0; a < 3; ++a)
    {
        e[a] = gridExponent(nb.hi[a] - nb.lo[a]);
        for (;;)
        {
            const float scale = u2f(e[a] << 23);
            bool ok = true;
            for (int s = 0; s < 8 && ok; ++s)
            {
                const int i = childAt[s];
                uint32_t ql = 255, qh = 0;
                if (i >= 0) ok = quantisePlanes(plo[a], scale, cb[i].lo[a], cb[i].hi[a], ql, qh);
                node.qlo[a][s] = (uint8_t)ql;
                node.qhi[a][s] = (uint8_t)qh;
            }
            if (ok || e[a] >= 254u) break;
            e[a]++;
        }
    }
    node.ex = (uint8_t)e[0]; node.ey = (uint8_t)e[1]; node.ez = (uint8_t)e[2];
    c.nodes[item.wide] = node;
}

// Dequantised bounds of a wide node (root box of a BLAS for the TLAS builder; tests).
HD Box wideNodeBounds(const WideNode &n)
{
    const float sx = u2f((uint32_t)n.ex << 23), sy = u2f((uint32_t)n.ey << 23), sz = u2f((uint32_t)n.ez << 23);
    V3 lo = {3.4e38f, 3.4e38f, 3.4e38f}, hi = {-3.4e38f, -3.4e38f, -3.4e38f};
    for (int s = 0; s < 8; ++s)
    {
        if (n.meta[s] == 0) continue;
        lo = vmin(lo, V3{n.px + n.qlo[0][s] * sx, n.py + n.qlo[1][s] * sy, n.pz + n.qlo[2][s] * sz});
        hi = vmax(hi, V3{n.px + n.qhi[0][s] * sx, n.py + n.qhi[1][s] * sy, n.pz + n.qhi[2][s] * sz});
    }
    return makeBox(lo, hi);
}

} // namespace prt
