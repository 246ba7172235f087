// core_trace.h -- two-level traversal of the compressed 8-wide BVH: the CUDA stand-in for
// traceRayEXT (main.rgen:57, 73) with scene.rchit / scene.rahit / scene.rmiss inlined.
//
// Results are independent of tree shape and traversal order by construction:
//   * ray/triangle test and the world->object ray transform follow the exact contract (single
//     roundings, fixed association) shared with the oracle,
//   * closest hit = smallest t, ties broken by the smallest (instance, geometry, primitive),
//   * box tests are conservative (quantised boxes only grow; slab interval widened by a few ulp),
//   * any-hit acceptance uses one random number per RAY (scene.rahit:35), so it is order-free.
#pragma once
#include <cstring>
#include "core_bvh.h"
#include "core_scene.h"

#ifndef PRT_TRACE_PREFETCH
#define PRT_TRACE_PREFETCH 0 // measured: 3 % slower on the atrium
#endif

namespace prt {

constexpr uint32_t MISS_INDEX = 0xFFFFFFFFu; // main.rgen:47
constexpr int TRACE_STACK = 64;

struct HitRec
{
    uint32_t drawInstanceIndex; // scene.rchit:12: gl_InstanceCustomIndexEXT + gl_GeometryIndexEXT
    uint32_t primitiveID;
    float u, v; // barycentrics of v1, v2 (scene.rchit:8,14)
    float t;
    uint32_t instance; // tie-break keys, not part of the payload
    uint32_t geometry;
};

struct TraceCounters
{
    uint32_t nodes, tris, instances, anyHits, overflow;
};

struct Accel
{
    const WideNode *tlasNodes; // nullptr = empty scene
    const InstRec *instances;  // TLAS node order
    uint32_t instanceCount;    // 1: rays start inside the only instance (its BLAS root culls what the TLAS root would)
};

struct NodeRay
{
    V3 o, d;   // exact ray in the current space
    float ix, iy, iz; // reciprocal direction for the box test (zero components nudged)
    uint32_t octinv;  // bit k set: direction non-negative on axis k
};

HD void setupNodeRay(NodeRay &r, V3 o, V3 d)
{
    r.o = o;
    r.d = d;
    const float eps = 1e-18f;
    const float dx = fabsf(d.x) > eps ? d.x : (f2u(d.x) >> 31 ? -eps : eps);
    const float dy = fabsf(d.y) > eps ? d.y : (f2u(d.y) >> 31 ? -eps : eps);
    const float dz = fabsf(d.z) > eps ? d.z : (f2u(d.z) >> 31 ? -eps : eps);
    r.ix = 1.0f / dx;
    r.iy = 1.0f / dy;
    r.iz = 1.0f / dz;
    r.octinv = (dx >= 0.f ? 1u : 0u) | (dy >= 0.f ? 2u : 0u) | (dz >= 0.f ? 4u : 0u);
}

// EXACT CONTRACT: world -> object (oracle o_bvh.h transformPoint / transformVector)
HD V3 xformPoint(const float *m, V3 p)
{
    return {xadd(xadd(xadd(xmul(m[0], p.x), xmul(m[1], p.y)), xmul(m[2], p.z)), m[3]),
            xadd(xadd(xadd(xmul(m[4], p.x), xmul(m[5], p.y)), xmul(m[6], p.z)), m[7]),
            xadd(xadd(xadd(xmul(m[8], p.x), xmul(m[9], p.y)), xmul(m[10], p.z)), m[11])};
}
HD V3 xformVector(const float *m, V3 d)
{
    return {xadd(xadd(xmul(m[0], d.x), xmul(m[1], d.y)), xmul(m[2], d.z)),
            xadd(xadd(xmul(m[4], d.x), xmul(m[5], d.y)), xmul(m[6], d.z)),
            xadd(xadd(xmul(m[8], d.x), xmul(m[9], d.y)), xmul(m[10], d.z))};
}

// EXACT CONTRACT: inverse of a row-major affine 3x4 (oracle o_bvh.h invertTransform)
HD void invertTransformExact(const float *m, float *out)
{
    const float a = m[0], b = m[1], c = m[2];
    const float d = m[4], e = m[5], f = m[6];
    const float g = m[8], h = m[9], i = m[10];
    const float c00 = xsub(xmul(e, i), xmul(f, h));
    const float c01 = xsub(xmul(f, g), xmul(d, i));
    const float c02 = xsub(xmul(d, h), xmul(e, g));
    const float det = xadd(xadd(xmul(a, c00), xmul(b, c01)), xmul(c, c02));
    const float i00 = xdiv(c00, det), i01 = xdiv(xsub(xmul(c, h), xmul(b, i)), det), i02 = xdiv(xsub(xmul(b, f), xmul(c, e)), det);
    const float i10 = xdiv(c01, det), i11 = xdiv(xsub(xmul(a, i), xmul(c, g)), det), i12 = xdiv(xsub(xmul(c, d), xmul(a, f)), det);
    const float i20 = xdiv(c02, det), i21 = xdiv(xsub(xmul(b, g), xmul(a, h)), det), i22 = xdiv(xsub(xmul(a, e), xmul(b, d)), det);
    const float tx = m[3], ty = m[7], tz = m[11];
    out[0] = i00; out[1] = i01; out[2] = i02;
    out[3] = -xadd(xadd(xmul(i00, tx), xmul(i01, ty)), xmul(i02, tz));
    out[4] = i10; out[5] = i11; out[6] = i12;
    out[7] = -xadd(xadd(xmul(i10, tx), xmul(i11, ty)), xmul(i12, tz));
    out[8] = i20; out[9] = i21; out[10] = i22;
    out[11] = -xadd(xadd(xmul(i20, tx), xmul(i21, ty)), xmul(i22, tz));
}

// EXACT CONTRACT: Moeller-Trumbore, two-sided, no epsilon, open interval (oracle intersectTriangle)
HD bool intersectTriangleExact(V3 o, V3 d, float tMin, float tMax, V3 v0, V3 e1, V3 e2, float &t, float &u, float &v)
{
    // Straight-line form: the lanes of a warp test different triangles, so the early exits of the textbook version
    // never save the warp anything; every value is computed with the same operations in the same order and the
    // decision is one conjunction (a zero determinant makes inv infinite, which the comparisons below reject, and is
    // excluded explicitly as before).
    const V3 P = xcross(d, e2);
    const float det = xdot(e1, P);
    const float inv = xdiv(1.0f, det);
    const V3 T = xsub3(o, v0);
    const float uu = xmul(xdot(T, P), inv);
    const V3 Q = xcross(T, e1);
    const float vv = xmul(xdot(d, Q), inv);
    const float tt = xmul(xdot(e2, Q), inv);
    const bool hit = (det != 0.0f) & (uu >= 0.0f) & !(uu > 1.0f) & (vv >= 0.0f) & !(xadd(uu, vv) > 1.0f) & (tt > tMin) & (tt < tMax);
    t = tt; u = uu; v = vv;
    return hit;
}

HD bool betterHit(float t, uint32_t inst, uint32_t geom, uint32_t prim, const HitRec &h)
{
    if (h.drawInstanceIndex == MISS_INDEX) return true;
    if (t != h.t) return t < h.t;
    if (inst != h.instance) return inst < h.instance;
    if (geom != h.geometry) return geom < h.geometry;
    return prim < h.primitiveID;
}

HD void loadNode(const WideNode *p, WideNode &n)
{
#if defined(__CUDA_ARCH__)
    const uint4 *s = reinterpret_cast<const uint4 *>(p);
    uint4 *d = reinterpret_cast<uint4 *>(&n);
    d[0] = __ldg(s + 0); d[1] = __ldg(s + 1); d[2] = __ldg(s + 2); d[3] = __ldg(s + 3); d[4] = __ldg(s + 4);
#else
    n = *p;
#endif
}

// Hint the line holding `p` into L1 (no register cost): issued as soon as the next item's address is known,
// one traversal step before the item is loaded.
HD void prefetchL1(const void *p)
{
#if defined(__CUDA_ARCH__)
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#else
    (void)p;
#endif
}

HD AlphaMat loadAlphaMat(const AlphaMat *p)
{
#if defined(__CUDA_ARCH__)
    AlphaMat m;
    const uint4 *s = reinterpret_cast<const uint4 *>(p);
    uint4 *d = reinterpret_cast<uint4 *>(&m);
    d[0] = __ldg(s + 0); d[1] = __ldg(s + 1);
    return m;
#else
    return *p;
#endif
}
HD AlphaUv loadAlphaUv(const AlphaUv *p)
{
#if defined(__CUDA_ARCH__)
    AlphaUv m;
    const uint4 *s = reinterpret_cast<const uint4 *>(p);
    uint4 *d = reinterpret_cast<uint4 *>(&m);
    d[0] = __ldg(s + 0); d[1] = __ldg(s + 1);
    return m;
#else
    return *p;
#endif
}

// Intersects the ray with the eight quantised child boxes. Returns the hit mask: internal children
// in bits 24..31 at position 24 + (slot ^ octinv) (nearer octant = higher bit), leaf primitives in
// bits 0..23.
HD uint32_t intersectNode(const WideNode &n, const NodeRay &r, float tMin, float tMax)
{
    const float sx = u2f((uint32_t)n.ex << 23) * r.ix, sy = u2f((uint32_t)n.ey << 23) * r.iy, sz = u2f((uint32_t)n.ez << 23) * r.iz;
    const float ox = (n.px - r.o.x) * r.ix, oy = (n.py - r.o.y) * r.iy, oz = (n.pz - r.o.z) * r.iz;
    // Entry/exit planes by ray octant instead of min/max per axis: with a non-negative direction the scaled
    // step is non-negative and lo * s + o <= hi * s + o, so the near plane is `lo` (and vice versa) -- the same
    // values min(x0, x1) / max(x0, x1) would pick. Branch-free so that the eight slots stay converged; an empty
    // slot (meta == 0) contributes no bits whatever its planes say.
    // (selected a 32-bit word at a time so that the node stays in registers)
    uint32_t lo[6], hi[6];
    memcpy(lo, n.qlo, 24);
    memcpy(hi, n.qhi, 24);
    const bool px = (r.octinv & 1u) != 0, py = (r.octinv & 2u) != 0, pz = (r.octinv & 4u) != 0;
    const uint32_t nw[6] = {px ? lo[0] : hi[0], px ? lo[1] : hi[1], py ? lo[2] : hi[2], py ? lo[3] : hi[3], pz ? lo[4] : hi[4], pz ? lo[5] : hi[5]};
    const uint32_t fw[6] = {px ? hi[0] : lo[0], px ? hi[1] : lo[1], py ? hi[2] : lo[2], py ? hi[3] : lo[3], pz ? hi[4] : lo[4], pz ? hi[5] : lo[5]};
    uint32_t mask = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int s = 0; s < 8; ++s)
    {
        const uint32_t meta = n.meta[s];
        const int w = s >> 2, sh = (s & 3) * 8;
        const float x0 = (float)((nw[0 + w] >> sh) & 255u) * sx + ox, x1 = (float)((fw[0 + w] >> sh) & 255u) * sx + ox;
        const float y0 = (float)((nw[2 + w] >> sh) & 255u) * sy + oy, y1 = (float)((fw[2 + w] >> sh) & 255u) * sy + oy;
        const float z0 = (float)((nw[4 + w] >> sh) & 255u) * sz + oz, z1 = (float)((fw[4 + w] >> sh) & 255u) * sz + oz;
        float tn = fmaxf_(fmaxf_(x0, y0), fmaxf_(z0, tMin));
        float tf = fminf_(fminf_(x1, y1), fminf_(z1, tMax));
        tn = tn - fabsf(tn) * 4e-7f;
        tf = tf + fabsf(tf) * 4e-7f;
        const uint32_t low5 = meta & 31u;
        const uint32_t bits = ((meta & 0x18u) == 0x18u) ? (1u << (low5 ^ r.octinv)) : ((meta >> 5) << low5);
        mask |= (tn <= tf) ? bits : 0u;
    }
    return mask;
}

// Traversal state of one ray, resumable one step at a time so that the persistent trace kernels can
// retire and refill lanes between steps (kernels_render.cuh); traceRay below runs it to completion.
struct TraceState
{
    float tMin, tMax;
    NodeRay ray;     // ray in the current space (world in the TLAS, object in a BLAS)
    const WideNode *nodes;
    const TriRec *tris;
    uint32_t instIndex;
    uint32_t gx, gy; // node group: base index, hit bits | imask
    uint32_t px, py; // primitive group: base index, hit bits
    int sp;
    int blasSp;      // stack depth at BLAS entry; -1 while in the TLAS
    HitRec best;
};
// Rarely read part of the traversal state (instance entry/exit, any-hit): lives next to the stack in local
// memory so that it does not hold registers across the hot node / triangle loop.
struct TraceCold
{
    V3 o, d;            // world-space ray
    uint32_t seed;      // payload.randomSeed (scene.rahit:35)
    uint32_t instCustom;
    const AlphaUv *uvs; // any-hit records of the current BLAS (may be nullptr)
    float pendT, pendU, pendV; // candidate waiting for its any-hit test (TRACE_PENDING set in TraceState::px)
    uint32_t pendTri;          // its index in the BLAS triangle array
};
constexpr uint32_t TRACE_PENDING = 0x80000000u;

// kept apart from TraceState so that the scalar state stays in registers (a dynamically indexed member
// would drag the whole struct into local memory)
struct TraceStack
{
    uint32_t x[TRACE_STACK], y[TRACE_STACK];
    TraceCold cold;
    HD void put(int i, uint32_t a, uint32_t b) { x[i] = a; y[i] = b; }
    HD void get(int i, uint32_t &a, uint32_t &b) const { a = x[i]; b = y[i]; }
};

#if defined(__CUDACC__)
// The first TRACE_SMEM_STACK entries of every thread's stack live in shared memory ([entry][thread], 8-byte
// columns: conflict-free), deeper ones in a per-thread local array that is almost never touched.
constexpr int TRACE_SMEM_STACK = 12;
struct SharedTraceStack
{
    uint2 *column; // &smem[0][threadIdx.x]
    uint32_t stride; // threads per block
    uint32_t ox[TRACE_STACK - TRACE_SMEM_STACK], oy[TRACE_STACK - TRACE_SMEM_STACK];
    TraceCold cold;
    __device__ __forceinline__ void put(int i, uint32_t a, uint32_t b)
    {
        if (i < TRACE_SMEM_STACK) column[(uint32_t)i * stride] = make_uint2(a, b);
        else { ox[i - TRACE_SMEM_STACK] = a; oy[i - TRACE_SMEM_STACK] = b; }
    }
    __device__ __forceinline__ void get(int i, uint32_t &a, uint32_t &b) const
    {
        if (i < TRACE_SMEM_STACK) { const uint2 v = column[(uint32_t)i * stride]; a = v.x; b = v.y; }
        else { a = ox[i - TRACE_SMEM_STACK]; b = oy[i - TRACE_SMEM_STACK]; }
    }
};
#endif

// Returns false when there is nothing to traverse (the ray is a miss).
template <class Stack>
HD bool traceBegin(TraceState &s, Stack &st, const Accel &as, V3 o, V3 d, float tMin, float tMax, uint32_t seed,
                   TraceCounters *cnt = nullptr)
{
    s.best.drawInstanceIndex = MISS_INDEX;
    s.best.primitiveID = MISS_INDEX;
    s.best.u = 0.f; s.best.v = 0.f; s.best.t = tMax; s.best.instance = MISS_INDEX; s.best.geometry = 0;
    st.cold.o = o; st.cold.d = d; st.cold.seed = seed; st.cold.instCustom = 0; st.cold.uvs = nullptr;
    s.tMin = tMin; s.tMax = tMax;
    s.sp = 0;
    s.blasSp = -1;
    setupNodeRay(s.ray, o, d);
    s.nodes = as.tlasNodes;
    s.tris = nullptr;
    s.instIndex = 0;
    s.gx = 0; s.gy = 0x80000000u;
    s.px = 0; s.py = 0;
    if (as.tlasNodes == nullptr) return false;
    if (as.instanceCount == 1u)
    {
        // single-instance scene (a baked glTF such as Sponza): skip the one-node TLAS
        const InstRec &inst = as.instances[0];
        if (cnt) cnt->instances++;
        if ((inst.mask & 0xFFu) == 0) return false;
        s.blasSp = 0;
        setupNodeRay(s.ray, xformPoint(inst.w2o, o), xformVector(inst.w2o, d));
        s.nodes = inst.nodes;
        s.tris = inst.tris;
        st.cold.uvs = inst.uvs;
        st.cold.instCustom = inst.customIndex;
        s.instIndex = inst.index;
    }
    return true;
}

// One step of traceRayEXT = ONE item of work, so that the lanes of a warp stay in lock step on items of
// similar cost: the any-hit test of a parked candidate, or one pending primitive of the current primitive group
// (a triangle test, or entering an instance), or one node (take the nearest pending child of the current node
// group, intersect its eight boxes). Pops when both groups are exhausted. traceItemKind tells which it will be.
// SHADOW = gl_RayFlagsTerminateOnFirstHitEXT (main.rgen:52). Returns true when the ray is finished.
HD int traceItemKind(const TraceState &s) { return (s.px & TRACE_PENDING) ? 2 : (s.py != 0 ? 1 : 0); }
// A lane with pending primitives AND pending nodes may serve either kind: postponing its primitive group (it goes
// on the stack) is legal because the result does not depend on traversal order.
HD bool traceCanPostpone(const TraceState &s) { return !(s.px & TRACE_PENDING) && s.py != 0 && (s.gy & 0xFF000000u) != 0; }

// preferNode: take a node item even though primitives are pending (only meaningful when traceCanPostpone(s)).
// kind: the caller's knowledge of traceItemKind(s) (-1: look it up). The trace kernels pass the warp-uniform kind
// their vote selected, which turns the three-way choice below into a uniform branch.
template <bool SHADOW, bool COUNT, class Stack>
HD bool traceStep(TraceState &s, Stack &st, const Accel &as, const SceneView &sv, TraceCounters *cnt, bool preferNode = false,
                  int kind = -1)
{
    if (preferNode && traceCanPostpone(s))
    {
        if (s.sp < TRACE_STACK) { st.put(s.sp, s.px, s.py); s.sp++; s.py = 0; }
        // (a full stack keeps the primitives first: nothing is lost, only the order stays as it was)
        kind = 0;
    }
    if (kind < 0) kind = traceItemKind(s);
    if (kind == 2)
    {
        // scene.rahit for the candidate parked by the triangle item below (its own item kind, so that the lanes of
        // a warp run the alpha test together instead of one or two at a time inside the triangle item)
        s.px &= ~TRACE_PENDING;
        const uint32_t tri = st.cold.pendTri;
        const float t = st.cold.pendT, u = st.cold.pendU, v = st.cold.pendV;
        const TriRec &tr = s.tris[tri];
        const uint32_t geom = tr.geometry & 0x7FFFFFFFu;
        const uint32_t drawInstanceIndex = st.cold.instCustom + geom;
        if (COUNT) cnt->anyHits++;
        const AlphaUv *uvs = st.cold.uvs;
        const uint32_t seed = st.cold.seed;
        bool accept = true;
        bool fast = uvs != nullptr && sv.alphaMats != nullptr;
        if (fast)
        {
            const AlphaMat am = loadAlphaMat(sv.alphaMats + drawInstanceIndex);
            fast = (am.bits & ALPHAMAT_USABLE) != 0;
            if (fast) accept = anyHitAcceptFast(am, loadAlphaUv(uvs + tri), u, v, seed);
        }
        if (!fast) accept = anyHitAccept(sv, drawInstanceIndex, tr.primitive, u, v, seed);
        if (accept)
        {
            s.best.drawInstanceIndex = drawInstanceIndex;
            s.best.primitiveID = tr.primitive;
            s.best.u = u; s.best.v = v; s.best.t = t;
            s.best.instance = s.instIndex;
            s.best.geometry = geom;
            if (SHADOW) return true;
        }
    }
    else if (kind == 1)
    {
        const int pbit = bfind(s.py);
        s.py &= ~(1u << pbit);
        if (s.blasSp < 0)
        {
            // TLAS leaf: enter the instance (traversal state of the TLAS goes on the stack)
            const InstRec &inst = as.instances[s.px + (uint32_t)pbit];
            if (COUNT) cnt->instances++;
            if ((inst.mask & 0xFFu) != 0)
            {
                if (s.gy & 0xFF000000u)
                {
                    if (s.sp < TRACE_STACK) { st.put(s.sp, s.gx, s.gy); s.sp++; }
                    else if (COUNT) cnt->overflow++;
                }
                if (s.py != 0)
                {
                    if (s.sp < TRACE_STACK) { st.put(s.sp, s.px, s.py); s.sp++; }
                    else if (COUNT) cnt->overflow++;
                }
                s.blasSp = s.sp;
                setupNodeRay(s.ray, xformPoint(inst.w2o, st.cold.o), xformVector(inst.w2o, st.cold.d));
                s.nodes = inst.nodes;
                s.tris = inst.tris;
                st.cold.uvs = inst.uvs;
                st.cold.instCustom = inst.customIndex;
                s.instIndex = inst.index;
                s.gx = 0; s.gy = 0x80000000u;
                s.py = 0;
                return false;
            }
        }
        else
        {
            const TriRec &tr = s.tris[s.px + (uint32_t)pbit];
            if (COUNT) cnt->tris++;
            float t, u, v;
            if (intersectTriangleExact(s.ray.o, s.ray.d, s.tMin, s.tMax, V3{tr.v0[0], tr.v0[1], tr.v0[2]}, V3{tr.e1[0], tr.e1[1], tr.e1[2]},
                                       V3{tr.e2[0], tr.e2[1], tr.e2[2]}, t, u, v))
            {
                const uint32_t geom = tr.geometry & 0x7FFFFFFFu;
                if (betterHit(t, s.instIndex, geom, tr.primitive, s.best))
                {
                    if (!(tr.geometry >> 31))
                    {
                        // non-opaque geometry: park the candidate, its any-hit test is the lane's next item
                        st.cold.pendT = t; st.cold.pendU = u; st.cold.pendV = v;
                        st.cold.pendTri = s.px + (uint32_t)pbit;
                        s.px |= TRACE_PENDING;
                        return false;
                    }
                    s.best.drawInstanceIndex = st.cold.instCustom + geom;
                    s.best.primitiveID = tr.primitive;
                    s.best.u = u; s.best.v = v; s.best.t = t;
                    s.best.instance = s.instIndex;
                    s.best.geometry = geom;
                    if (SHADOW) return true;
                }
            }
        }
    }
    else if (s.gy & 0xFF000000u)
    {
        const int bit = bfind(s.gy);
        const uint32_t imask = s.gy;
        s.gy &= ~(1u << bit);
        if (s.gy & 0xFF000000u)
        {
            if (s.sp < TRACE_STACK) { st.put(s.sp, s.gx, s.gy); s.sp++; }
            else if (COUNT) cnt->overflow++;
        }
        const uint32_t slot = ((uint32_t)bit - 24u) ^ s.ray.octinv;
        const uint32_t rel = (uint32_t)popc(imask & 0xFFu & ((1u << slot) - 1u));
        WideNode n;
        loadNode(s.nodes + s.gx + rel, n);
        if (COUNT) cnt->nodes++;
        const float tFar = s.best.drawInstanceIndex == MISS_INDEX ? s.tMax : fminf_(s.tMax, s.best.t);
        const uint32_t hits = intersectNode(n, s.ray, s.tMin, tFar);
        s.gx = n.childBase;
        s.gy = (hits & 0xFF000000u) | n.imask;
        s.px = n.primBase;
        s.py = hits & 0x00FFFFFFu;
#if PRT_TRACE_PREFETCH
        if (s.py != 0)
        {
            if (s.blasSp >= 0) prefetchL1(s.tris + s.px + (uint32_t)bfind(s.py));
        }
        else if (s.gy & 0xFF000000u)
        {
            const uint32_t nslot = ((uint32_t)bfind(s.gy) - 24u) ^ s.ray.octinv;
            const WideNode *next = s.nodes + s.gx + (uint32_t)popc(s.gy & 0xFFu & ((1u << nslot) - 1u));
            prefetchL1(next);
            prefetchL1(reinterpret_cast<const char *>(next) + 64);
        }
#endif
    }

    if (s.py == 0 && (s.gy & 0xFF000000u) == 0)
    {
        if (s.sp == 0) return true;
        if (s.sp == s.blasSp)
        {
            // BLAS exhausted: back to world space
            s.blasSp = -1;
            setupNodeRay(s.ray, st.cold.o, st.cold.d);
            s.nodes = as.tlasNodes;
        }
        s.sp--;
        st.get(s.sp, s.gx, s.gy);
        if ((s.gy & 0xFF000000u) == 0)
        {
            // a postponed primitive group (pushed when an instance was entered)
            s.px = s.gx; s.py = s.gy;
            s.gx = 0; s.gy = 0;
        }
    }
    return false;
}

// traceRayEXT to completion. Returns true on a hit.
template <bool SHADOW, bool COUNT>
HDNI bool traceRay(const Accel &as, const SceneView &sv, V3 o, V3 d, float tMin, float tMax, uint32_t seed, HitRec &best,
                   TraceCounters *cnt)
{
    TraceState s;
    TraceStack st;
    if (traceBegin(s, st, as, o, d, tMin, tMax, seed, COUNT ? cnt : nullptr))
        while (!traceStep<SHADOW, COUNT>(s, st, as, sv, cnt)) {}
    best = s.best;
    return best.drawInstanceIndex != MISS_INDEX;
}

} // namespace prt
