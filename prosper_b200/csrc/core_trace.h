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
#define PRT_TRACE_PREFETCH 0 // measured twice (r01: -3 %, r02 pooled kernel: -14 %): prefetch.global.L1 of the next item costs more L1 than it hides
#endif

// Reference splitting of non-opaque triangles in the BLAS builder (core_build.h): rounds of midpoint subdivision, 0 = off.
#ifndef PRT_BLAS_ALPHA_SPLIT
#define PRT_BLAS_ALPHA_SPLIT 0
#endif

namespace prt {

constexpr int PEND_ROWS = PRT_BLAS_ALPHA_SPLIT ? 2 : 1; // 128-bit words of RayCtx::pend per ray
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
    uint32_t planeBias;        // 0x47000000 (the float 32768): handed in as data so that the node test's PRMTs keep it in a
                               // register and take their byte selectors as immediates (intersectNode)
    uint64_t arena;            // 128-byte aligned address at or below every node / triangle array: a ray names the arrays of
                               // its level by their distance from it in 128-byte units (ROW_M), 32 bits each = 512 GB of reach
};
HD uint32_t arenaUnits(const Accel &as, const void *p) { return (uint32_t)(((uint64_t)(uintptr_t)p - as.arena) >> 7); }
HD const WideNode *arenaNodes(const Accel &as, uint32_t units) { return reinterpret_cast<const WideNode *>((uintptr_t)(as.arena + ((uint64_t)units << 7))); }
HD const TriRec *arenaTris(const Accel &as, uint32_t units) { return reinterpret_cast<const TriRec *>((uintptr_t)(as.arena + ((uint64_t)units << 7))); }
constexpr uint32_t PLANE_BIAS = 0x47000000u;

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

// __byte_perm (PRMT): result byte k = byte (sel >> 4k) & 7 of the eight bytes b:a.
HD uint32_t bytePerm(uint32_t a, uint32_t b, uint32_t sel)
{
#if defined(__CUDA_ARCH__)
    return __byte_perm(a, b, sel);
#else
    const uint64_t v = ((uint64_t)b << 32) | a;
    uint32_t r = 0;
    for (int k = 0; k < 4; ++k) r |= (uint32_t)((v >> (8 * ((sel >> (4 * k)) & 7u))) & 255u) << (8 * k);
    return r;
#endif
}
HD float fmaf_(float a, float b, float c) { return fmaf(a, b, c); }

// The 80-byte node as the five 128-bit words it is fetched in.
struct NodeWords
{
    uint4 w0; // px, py, pz, ex | ey << 8 | ez << 16 | imask << 24
    uint4 w1; // childBase, primBase, meta[0..3], meta[4..7]
    uint4 w2; // qlo[x][0..3], qlo[x][4..7], qlo[y][0..3], qlo[y][4..7]
    uint4 w3; // qlo[z][0..3], qlo[z][4..7], qhi[x][0..3], qhi[x][4..7]
    uint4 w4; // qhi[y][0..3], qhi[y][4..7], qhi[z][0..3], qhi[z][4..7]
};
static_assert(sizeof(NodeWords) == WIDE_NODE_BYTES, "NodeWords mirrors the payload of WideNode");

// L1 policy of the node and triangle gathers (experiment switches): 0 = default, 1 = L1::no_allocate, 2 = L1::evict_first,
// 3 = L1::evict_last. The idea: upper-level nodes are reused by every ray of an SM, a triangle record almost never.
#ifndef PRT_NODE_L1
#define PRT_NODE_L1 0
#endif
#ifndef PRT_TRI_L1
#define PRT_TRI_L1 0
#endif
#if defined(__CUDA_ARCH__)
#define PRT_LDG_HINT_0 ""
#define PRT_LDG_HINT_1 ".L1::no_allocate"
#define PRT_LDG_HINT_2 ".L1::evict_first"
#define PRT_LDG_HINT_3 ".L1::evict_last"
#define PRT_LDG_HINT_(h) PRT_LDG_HINT_##h
#define PRT_LDG_HINT(h) PRT_LDG_HINT_(h)
// One 256-bit read-only load (LDG.E.256, sm_100): `p` must be 32-byte aligned.
template <int HINT>
__device__ __forceinline__ void ldg256(const void *p, uint4 &a, uint4 &b)
{
#define PRT_LDG256(h)                                                                                                       \
    asm("ld.global.nc" h ".v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"                                                        \
        : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w)                            \
        : "l"(p))
    if (HINT == 1) PRT_LDG256(PRT_LDG_HINT_1);
    else if (HINT == 2) PRT_LDG256(PRT_LDG_HINT_2);
    else if (HINT == 3) PRT_LDG256(PRT_LDG_HINT_3);
    else PRT_LDG256(PRT_LDG_HINT_0);
#undef PRT_LDG256
}
template <int HINT>
__device__ __forceinline__ uint4 ldg128(const void *p)
{
    uint4 a;
#define PRT_LDG128(h) asm("ld.global.nc" h ".v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w) : "l"(p))
    if (HINT == 1) PRT_LDG128(PRT_LDG_HINT_1);
    else if (HINT == 2) PRT_LDG128(PRT_LDG_HINT_2);
    else if (HINT == 3) PRT_LDG128(PRT_LDG_HINT_3);
    else PRT_LDG128(PRT_LDG_HINT_0);
#undef PRT_LDG128
    return a;
}
#endif

HD void loadNodeWords(const WideNode *p, NodeWords &n)
{
#if defined(__CUDA_ARCH__)
    const char *s = reinterpret_cast<const char *>(p);
    ldg256<PRT_NODE_L1>(s, n.w0, n.w1);
    ldg256<PRT_NODE_L1>(s + 32, n.w2, n.w3);
    n.w4 = ldg128<PRT_NODE_L1>(s + 64);
#else
    memcpy(&n, p, sizeof(NodeWords));
#endif
}

// Intersects the ray with the eight quantised child boxes. Returns the hit mask: internal children
// in bits 24..31 at position 24 + (slot ^ octinv) (nearer octant = higher bit), leaf primitives in
// bits 0..23.
//
// A plane sits at p + q * 2^(e-127), q in 0..255, so its ray parameter is q * s + op with s = 2^(e-127) / d and
// op = (p - o) / d, per axis. The byte q goes straight from the packed word into the mantissa of the float
// 32768 + q (one PRMT; no integer-to-float conversion, which is a quarter-rate XU instruction and was half of that
// pipe's load), and the bias moves into the constant: t = (32768 + q) * s + (op - 32768 * s), one FMA per plane.
//
// CONSERVATIVE by construction: every rounding on the way is bounded by operand magnitude and
// the bound is folded into the constants -- near planes use b - err, far planes b + err, with
//   err = |1/d| * 2^-22 * (|p| + |o| + 256 * 2^(e-127))      plane position (builder rounding) and p - o
//       + 2^-21 * (|op| + 33023 * |s|)                        the products, the bias fold and the final FMA,
// so a point inside a child's real box always passes its slab test whatever the tree shape, including on-grid flat
// boxes and origins next to a plane of a large node. 32768 * |s| is 128 node extents, so the widening is about
// 1e-4 of the node's extent along the ray: no measurable loss of culling.
HD uint32_t intersectNode(const NodeWords &n, V3 o, V3 idir, uint32_t octinv, float tMin, float tMax, uint32_t bias)
{
    const float kBias = 32768.f, kRel = 4.76837158e-7f /* 2^-21 */, kPos = 2.38418579e-7f /* 2^-22 */;
    const float cx = u2f((n.w0.w & 0xFFu) << 23), cy = u2f((n.w0.w & 0xFF00u) << 15), cz = u2f((n.w0.w & 0xFF0000u) << 7);
    const float px = u2f(n.w0.x), py = u2f(n.w0.y), pz = u2f(n.w0.z);
    const float sx = cx * idir.x, sy = cy * idir.y, sz = cz * idir.z;
    const float opx = (px - o.x) * idir.x, opy = (py - o.y) * idir.y, opz = (pz - o.z) * idir.z;
    const float bx = fmaf_(-kBias, sx, opx), by = fmaf_(-kBias, sy, opy), bz = fmaf_(-kBias, sz, opz);
    const float ex = fmaf_(fabsf(idir.x) * kPos, fmaf_(256.f, cx, fabsf(px) + fabsf(o.x)), fmaf_(fabsf(sx), 33023.f * kRel, fabsf(opx) * kRel));
    const float ey = fmaf_(fabsf(idir.y) * kPos, fmaf_(256.f, cy, fabsf(py) + fabsf(o.y)), fmaf_(fabsf(sy), 33023.f * kRel, fabsf(opy) * kRel));
    const float ez = fmaf_(fabsf(idir.z) * kPos, fmaf_(256.f, cz, fabsf(pz) + fabsf(o.z)), fmaf_(fabsf(sz), 33023.f * kRel, fabsf(opz) * kRel));
    const float bnx = bx - ex, bfx = bx + ex, bny = by - ey, bfy = by + ey, bnz = bz - ez, bfz = bz + ez;
    // Entry/exit planes by ray octant instead of min/max per axis: with a non-negative direction the scaled
    // step is non-negative, so the near plane is `lo` (and vice versa). Selected a packed word at a time.
    const bool dx = (octinv & 1u) != 0, dy = (octinv & 2u) != 0, dz = (octinv & 4u) != 0;
    const uint32_t nx[2] = {dx ? n.w2.x : n.w3.z, dx ? n.w2.y : n.w3.w}, fx[2] = {dx ? n.w3.z : n.w2.x, dx ? n.w3.w : n.w2.y};
    const uint32_t ny[2] = {dy ? n.w2.z : n.w4.x, dy ? n.w2.w : n.w4.y}, fy[2] = {dy ? n.w4.x : n.w2.z, dy ? n.w4.y : n.w2.w};
    const uint32_t nz[2] = {dz ? n.w3.x : n.w4.z, dz ? n.w3.y : n.w4.w}, fz[2] = {dz ? n.w4.z : n.w3.x, dz ? n.w4.w : n.w3.y};
    // meta bytes four at a time: a slot's bits are childBits << bitIndex with bitIndex = low5 (^ octinv for an internal
    // child, whose low5 is 24 + slot) and childBits = meta >> 5 (1 for an internal child, the unary primitive count
    // of a leaf, 0 for an empty slot, which therefore never contributes)
    const uint32_t oct4 = octinv * 0x01010101u;
    uint32_t mask = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int h = 0; h < 2; ++h)
    {
        const uint32_t meta4 = h ? n.w1.w : n.w1.z;
        const uint32_t inner4 = ((meta4 & (meta4 << 1)) & 0x10101010u) >> 4; // 1 per internal byte
        const uint32_t index4 = (meta4 ^ (oct4 & (inner4 * 7u))) & 0x1F1F1F1Fu;
        const uint32_t bits4 = (meta4 >> 5) & 0x07070707u;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int k = 0; k < 4; ++k)
        {
            const uint32_t sel = 0x7404u | ((uint32_t)k << 4); // bytes: 0x47, 0x00, q, 0x00 = the float 32768 + q
            const float x0 = fmaf_(u2f(bytePerm(nx[h], bias, sel)), sx, bnx), x1 = fmaf_(u2f(bytePerm(fx[h], bias, sel)), sx, bfx);
            const float y0 = fmaf_(u2f(bytePerm(ny[h], bias, sel)), sy, bny), y1 = fmaf_(u2f(bytePerm(fy[h], bias, sel)), sy, bfy);
            const float z0 = fmaf_(u2f(bytePerm(nz[h], bias, sel)), sz, bnz), z1 = fmaf_(u2f(bytePerm(fz[h], bias, sel)), sz, bfz);
            const float tn = fmaxf_(fmaxf_(x0, y0), fmaxf_(z0, tMin));
            const float tf = fminf_(fminf_(x1, y1), fminf_(z1, tMax));
            const uint32_t bits = ((bits4 >> (8 * k)) & 0xFFu) << ((index4 >> (8 * k)) & 0xFFu);
            mask |= (tn <= tf) ? bits : 0u;
        }
    }
    return mask;
}

// ---- pooled traversal state --------------------------------------------------------------------------------------
// A ray in flight is six 128-bit rows plus a stack. The rows live wherever the caller puts them (`rows[i * stride]`):
// shared memory, [row][ray][thread], in the persistent trace kernels -- every thread owns PRT_POOL_RAYS rays and advances, each round, whichever of them wants
// the item kind the warp voted for (kernels_render.cuh) -- and plain arrays in traceRay below (the CPU simulator,
// prt_trace_rays). Every item reads the rows it needs and writes back only what it changed, so nothing but a kind per
// ray stays in registers between items. All row accesses are whole 128-bit rows: [thread]-contiguous uint4 columns are
// conflict-free, while a 32-bit access at a 16-byte lane stride is a 4-way bank conflict (measured: 30 % of the
// shared-memory wavefronts of the first pooled kernel were conflicts, with the L1 data pipe 80 % busy).
//
// The committed hit itself is not traversal state: a commit writes the 16-byte record straight to the caller's hit
// queue (plus the two tie-break keys beside it), the rows keep only its t (`curT`) and a flag.
enum : int { ROW_O = 0, ROW_D = 1, ROW_I = 2, ROW_G = 3, ROW_M = 4, POOL_ROWS = 5 };
//  ROW_O  origin.xyz (current space: world in the TLAS, object in a BLAS), curT = committed hit ? its t : tMax
//  ROW_D  direction.xyz (current space), tMin
//  ROW_I  reciprocal direction.xyz (zero components nudged), octinv | TRACE_HAS_HIT | a lower bound of tMin in the top 16 bits
//         (the node test takes its tMin from here: the same row it reads anyway, and a lower bound only ever visits more)
//  ROW_G  node group (base index, hit bits << 24 | state | imask), primitive group (base index | TRACE_PENDING, hit bits);
//         state = sp << 8 | (stack depth at BLAS entry + 1) << 16 (0 while in the TLAS): the one row every item writes
//  ROW_M  node array, triangle array of the current level (arenaUnits), queue slot, instance position: written on entering
//         and leaving an instance only
// A node item reads G M I O and writes G; a triangle item reads G M O D and writes G (and O I on a commit). Round 1's layout
// had a sixth row for two 64-bit level pointers and kept sp in ROW_M: five reads and two writes per item, and 4 KB more
// shared memory per block (= less L1) -- the L1 data pipe is what bounds this kernel.
enum : int { KIND_NODE = 0, KIND_TRI = 1, KIND_ANYHIT = 2, KIND_INSTANCE = 3, KIND_COUNT = 4, KIND_IDLE = 7 };
constexpr uint32_t TRACE_PENDING = 0x80000000u; // ROW_G.z: a candidate waits for its any-hit test (parked in RayCtx::pend)
constexpr uint32_t TRACE_HAS_HIT = 0x100u;      // ROW_I.w
constexpr uint32_t TRACE_STATE = 0x00FFFF00u;   // ROW_G.y: sp << 8 | blasP1 << 16
HD uint32_t stateSp(uint32_t gy) { return (gy >> 8) & 0xFFu; }
HD uint32_t stateBlas(uint32_t gy) { return (gy >> 16) & 0xFFu; }
static_assert(TRACE_STACK + 1 < 256, "sp and blasP1 are bytes of ROW_G.y");
// The stack is per-thread local memory (L1). Measured alternative: the first 8 entries in shared memory next to the rows
// and the level pointers fetched through the instance record -- 942 instead of 990 Mrays/s, because the larger carve-out
// leaves 28 KB of L1 for nodes and triangles (hit rate 16 % instead of 31-45 %) and the two-tier entry access costs 10 %
// more instructions.
struct RayCtx
{
    uint4 *rows;      // row i at rows[i * stride]
    uint32_t stride;
    uint2 *stack;     // TRACE_STACK entries
    uint4 *pend;      // [0] parked any-hit candidate: t, u, v bits and the triangle index; [1] (PRT_BLAS_ALPHA_SPLIT builds only) the
                      // triangle this ray's any-hit test rejected last: instance position, geometry, primitive, valid (traceTriItem)
    HD uint4 &row(int i) const { return rows[(uint32_t)i * stride]; }
    HD uint2 &entry(uint32_t i) const { return stack[i]; }
};

// Where a ray comes from and where its result goes, by queue slot.
struct TraceIO
{
    const float4 *rayO; // world-space origin.xyz, tMin
    const float4 *rayD; // world-space direction.xyz, tMax
    const uint2 *ids;   // .y = payload.randomSeed (scene.rahit:35)
    uint4 *hits;        // closest: drawInstanceIndex, primitiveID, barycentric bits (written at every commit)
    uint2 *tie;         // closest: (instance index, geometry index) of the committed hit -- the tie-break keys
    unsigned long long *overflow; // sticky count of dropped stack entries (never silent, in every build)
};

HD void stackPush(const RayCtx &c, uint32_t &sp, uint32_t a, uint32_t b, const TraceIO &io)
{
    if (sp < (uint32_t)TRACE_STACK) { c.entry(sp) = make_uint2(a, b); sp++; }
    else if (io.overflow)
    {
#if defined(__CUDA_ARCH__)
        atomicAdd(io.overflow, 1ull); // global memory
#else
        (*io.overflow)++;
#endif
    }
}

// `flags` = the TRACE_HAS_HIT bit to keep. The bound of tMin kept in ROW_I.w is tMin with its low 16 bits cleared (towards
// minus infinity), so that the node test never starts later than the exact interval.
HD void storeRaySpace(const RayCtx &c, V3 o, V3 d, float curT, uint32_t flags, float tMin)
{
    NodeRay r;
    setupNodeRay(r, o, d);
    uint32_t lower = f2u(tMin) & 0xFFFF0000u;
    if ((f2u(tMin) >> 31) && (f2u(tMin) & 0xFFFFu)) lower += 0x10000u; // negative: one step further from zero
    c.row(ROW_O) = make_uint4(f2u(o.x), f2u(o.y), f2u(o.z), f2u(curT));
    c.row(ROW_D) = make_uint4(f2u(d.x), f2u(d.y), f2u(d.z), f2u(tMin));
    c.row(ROW_I) = make_uint4(f2u(r.ix), f2u(r.iy), f2u(r.iz), r.octinv | flags | lower);
}
// The stack's top entry, loaded at the START of an item so that a pop at its end does not wait for local memory (the
// pop runs at ~6 of 32 lanes and its load was 19 % of the kernel's stall samples).
HD uint2 stackTop(const RayCtx &c, uint32_t sp) { return c.entry(sp ? sp - 1u : 0u); }

HD int kindOf(const uint4 &g, uint32_t blasP1)
{
    return (g.z & TRACE_PENDING) ? KIND_ANYHIT : (g.w != 0 ? (blasP1 ? KIND_TRI : KIND_INSTANCE) : KIND_NODE);
}

// Starts a ray in `c`. Returns its first item kind, KIND_IDLE when there is nothing to traverse (a miss).
HD int traceBegin(const RayCtx &c, const Accel &as, const TraceIO &io, uint32_t qslot, TraceCounters *cnt)
{
    if (as.tlasNodes == nullptr) return KIND_IDLE;
#if PRT_BLAS_ALPHA_SPLIT
    c.pend[1] = make_uint4(0u, 0u, 0u, 0u); // the slot's previous ray may have left a rejected triangle behind
#endif
    const float4 wo = io.rayO[qslot], wd = io.rayD[qslot];
    const V3 o = {wo.x, wo.y, wo.z}, d = {wd.x, wd.y, wd.z};
    if (as.instanceCount == 1u)
    {
        // single-instance scene (a baked glTF such as Sponza): skip the one-node TLAS
        const InstRec &inst = as.instances[0];
        if (cnt) cnt->instances++;
        if ((inst.mask & 0xFFu) == 0) return KIND_IDLE;
        storeRaySpace(c, xformPoint(inst.w2o, o), xformVector(inst.w2o, d), wd.w, 0u, wo.w);
        c.row(ROW_G) = make_uint4(0u, 0x80000000u | (1u << 16), 0u, 0u);
        c.row(ROW_M) = make_uint4(arenaUnits(as, inst.nodes), arenaUnits(as, inst.tris), qslot, 0u);
        return KIND_NODE;
    }
    storeRaySpace(c, o, d, wd.w, 0u, wo.w);
    c.row(ROW_G) = make_uint4(0u, 0x80000000u, 0u, 0u);
    c.row(ROW_M) = make_uint4(arenaUnits(as, as.tlasNodes), 0u, qslot, 0u);
    return KIND_NODE;
}

// End of every item: when both groups are exhausted pop the next one (leaving the BLAS when its entries are used up);
// an empty stack finishes the ray. Stores the group row and the state row, returns the next item kind.
HD int traceFinishItem(const RayCtx &c, const Accel &as, const TraceIO &io, uint4 g, uint32_t sp, uint32_t blasP1, uint2 top)
{
    if (g.w == 0 && (g.y & 0xFF000000u) == 0 && !(g.z & TRACE_PENDING))
    {
        if (sp == 0) return KIND_IDLE;
        if (sp + 1u == blasP1)
        {
            // BLAS exhausted: back to world space
            blasP1 = 0;
            const uint4 m = c.row(ROW_M);
            const float4 wo = io.rayO[m.z], wd = io.rayD[m.z];
            storeRaySpace(c, V3{wo.x, wo.y, wo.z}, V3{wd.x, wd.y, wd.z}, u2f(c.row(ROW_O).w), c.row(ROW_I).w & TRACE_HAS_HIT, wo.w);
            c.row(ROW_M) = make_uint4(arenaUnits(as, as.tlasNodes), 0u, m.z, m.w);
        }
        sp--;
        const uint2 e = top;
        if (e.y & 0xFF000000u) { g.x = e.x; g.y = e.y; g.z = 0; g.w = 0; }
        else { g.x = 0; g.y = 0; g.z = e.x; g.w = e.y; } // a postponed primitive group (pushed when an instance was entered)
    }
    g.y = (g.y & ~TRACE_STATE) | (sp << 8) | (blasP1 << 16);
    c.row(ROW_G) = g;
    return kindOf(g, blasP1);
}

// gl_RayFlagsTerminateOnFirstHitEXT (SHADOW, main.rgen:52): the first accepted hit ends the ray. Otherwise the record
// goes to the hit queue and curT shrinks. `O` is the caller's copy of ROW_O. Returns true when the ray is finished.
template <bool SHADOW>
HD bool commitHit(const RayCtx &c, const Accel &as, const TraceIO &io, uint4 O, uint32_t qslot, uint32_t instPos, uint32_t geom, uint32_t prim,
                  float t, float u, float v)
{
    uint4 I = c.row(ROW_I);
    I.w |= TRACE_HAS_HIT;
    c.row(ROW_I) = I;
    if (SHADOW) return true;
    const InstRec &inst = as.instances[instPos];
    io.hits[qslot] = make_uint4(inst.customIndex + geom, prim, f2u(u), f2u(v));
    io.tie[qslot] = make_uint2(inst.index, geom);
    O.w = f2u(t);
    c.row(ROW_O) = O;
    return false;
}

#ifndef PRT_NODE_PER_ITEM
#define PRT_NODE_PER_ITEM 1 // nodes one node item may descend through (see traceNodeItem)
#endif
#ifndef PRT_TRI_PER_ITEM
#define PRT_TRI_PER_ITEM 2 // primitives one triangle item tests (1: round 1's one-per-item)
#endif

// KIND_NODE: take the nearest pending child of the node group, intersect its eight boxes.
template <bool COUNT>
HD int traceNodeItem(const RayCtx &c, const Accel &as, const TraceIO &io, TraceCounters *cnt)
{
    uint4 g = c.row(ROW_G);
    uint32_t sp = stateSp(g.y);
    const uint32_t blasP1 = stateBlas(g.y);
    g.y &= ~TRACE_STATE;
    uint2 top = stackTop(c, sp);
    const WideNode *nodes = arenaNodes(as, c.row(ROW_M).x);
    const uint4 I = c.row(ROW_I);
    const uint32_t octinv = I.w & 7u;
    const uint4 O = c.row(ROW_O);
    // Up to PRT_NODE_PER_ITEM nodes per item: a node whose test leaves node children only (no primitives, so no other kind
    // and no pop comes first) is followed at once by its nearest child -- the rows are read and written once for both.
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int k = 0; k < PRT_NODE_PER_ITEM; ++k)
    {
        if (k > 0 && !(g.w == 0u && (g.y & 0xFF000000u))) break;
        const int bit = bfind(g.y);
        const uint32_t imask = g.y;
        g.y &= ~(1u << bit);
        if (g.y & 0xFF000000u)
        {
            if (sp < (uint32_t)TRACE_STACK) top = make_uint2(g.x, g.y);
            stackPush(c, sp, g.x, g.y, io);
        }
        const uint32_t slot = ((uint32_t)bit - 24u) ^ octinv;
        const uint32_t rel = (uint32_t)popc(imask & 0xFFu & ((1u << slot) - 1u));
        NodeWords n;
        loadNodeWords(nodes + g.x + rel, n);
        if (COUNT) cnt->nodes++;
        const uint32_t hits = intersectNode(n, V3{u2f(O.x), u2f(O.y), u2f(O.z)}, V3{u2f(I.x), u2f(I.y), u2f(I.z)}, octinv, u2f(I.w & 0xFFFF0000u), u2f(O.w), as.planeBias);
        g.x = n.w1.x;
        g.y = (hits & 0xFF000000u) | (n.w0.w >> 24);
        g.z = n.w1.y;
        g.w = hits & 0x00FFFFFFu;
    }
    return traceFinishItem(c, as, io, g, sp, blasP1, top);
}

// KIND_TRI: one pending primitive of the primitive group inside a BLAS = one exact ray/triangle test.
template <bool SHADOW, bool COUNT>
HD int traceTriItem(const RayCtx &c, const Accel &as, const TraceIO &io, TraceCounters *cnt)
{
    uint4 g = c.row(ROW_G);
    const uint32_t sp = stateSp(g.y), blasP1 = stateBlas(g.y);
    g.y &= ~TRACE_STATE;
    const uint4 m = c.row(ROW_M);
    const TriRec *tris = arenaTris(as, m.y);
    // One item tests up to PRT_TRI_PER_ITEM primitives of the group: the ray rows are read and written once per item, not
    // once per triangle. It ends in a pop exactly when it tests the group's last primitives with no node child pending.
    const bool willPop = popc(g.w) <= PRT_TRI_PER_ITEM && (g.y & 0xFF000000u) == 0;
    const uint2 top = willPop ? stackTop(c, sp) : make_uint2(0u, 0u);
    uint4 O = c.row(ROW_O);
    const uint4 D = c.row(ROW_D);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int k = 0; k < PRT_TRI_PER_ITEM; ++k)
    {
        if (k > 0 && g.w == 0u) break;
        const int pbit = bfind(g.w);
        g.w &= ~(1u << pbit);
        const uint32_t triIndex = g.z + (uint32_t)pbit;
#if defined(__CUDA_ARCH__)
        uint4 a, b;
        ldg256<PRT_TRI_L1>(tris + triIndex, a, b);
        const uint4 e = ldg128<PRT_TRI_L1>(reinterpret_cast<const uint4 *>(tris + triIndex) + 2);
#else
        uint4 a, b, e;
        memcpy(&a, &tris[triIndex], 16); memcpy(&b, (const char *)&tris[triIndex] + 16, 16); memcpy(&e, (const char *)&tris[triIndex] + 32, 16);
#endif
        if (COUNT) cnt->tris++;
        const float curT = u2f(O.w);
        float t, u, v;
        // the open interval (tMin, tMax) and "closer than the committed hit" in one comparison against curT (the committed
        // t is below tMax); an exact tie goes to the smallest (instance, geometry, primitive) -- betterHit's order
        const bool inside = intersectTriangleExact(V3{u2f(O.x), u2f(O.y), u2f(O.z)}, V3{u2f(D.x), u2f(D.y), u2f(D.z)}, u2f(D.w), 3.402823466e38f,
                                                   V3{u2f(a.x), u2f(a.y), u2f(a.z)}, V3{u2f(b.x), u2f(b.y), u2f(b.z)}, V3{u2f(e.x), u2f(e.y), u2f(e.z)},
                                                   t, u, v);
        bool candidate = inside & (t < curT);
        const uint32_t geom = b.w & 0x7FFFFFFFu;
        if (inside && t == curT && !SHADOW)
        {
            const uint32_t flags = c.row(ROW_I).w;
            if (flags & TRACE_HAS_HIT)
            {
                const uint2 key = io.tie[m.z];
                const uint32_t bestPrim = io.hits[m.z].y;
                const uint32_t inst = as.instances[m.w].index;
                candidate = inst != key.x ? inst < key.x : (geom != key.y ? geom < key.y : a.w < bestPrim);
            }
        }
        if (candidate)
        {
            if (!(b.w >> 31))
            {
#if PRT_BLAS_ALPHA_SPLIT
                // non-opaque geometry. The any-hit decision is a pure function of (ray, triangle) -- one random number per
                // ray, scene.rahit:35 -- so the triangle this ray rejected last is skipped when it comes up again through
                // another builder reference (core_build.h: with reference splitting a non-opaque triangle sits in several leaves).
                const uint4 rej = c.pend[1];
                if (rej.w != 0u && rej.x == m.w && rej.y == geom && rej.z == a.w) continue;
#endif
                // non-opaque geometry: park the candidate, its any-hit test is the ray's next item
                c.pend[0] = make_uint4(f2u(t), f2u(u), f2u(v), triIndex);
                g.z |= TRACE_PENDING;
                break;
            }
            if (commitHit<SHADOW>(c, as, io, O, m.z, m.w, geom, a.w, t, u, v)) return KIND_IDLE;
            O.w = f2u(t); // the next primitive of this item is tested against the new hit
        }
    }
    return traceFinishItem(c, as, io, g, sp, blasP1, top);
}

// KIND_ANYHIT: scene.rahit for the candidate parked by the triangle item (its own kind, so that the rays of a warp
// run the alpha test together instead of one or two at a time inside the triangle item).
template <bool SHADOW, bool COUNT>
HD int traceAnyHitItem(const RayCtx &c, const Accel &as, const SceneView &sv, const TraceIO &io, TraceCounters *cnt)
{
    uint4 g = c.row(ROW_G);
    const uint32_t sp = stateSp(g.y), blasP1 = stateBlas(g.y);
    g.y &= ~TRACE_STATE;
    const uint4 m = c.row(ROW_M);
    g.z &= ~TRACE_PENDING;
    const uint4 pd = c.pend[0];
    const float t = u2f(pd.x), u = u2f(pd.y), v = u2f(pd.z);
    const uint32_t tri = pd.w;
    const InstRec &inst = as.instances[m.w];
    const TriRec *tris = inst.tris;
    const uint32_t geom = tris[tri].geometry & 0x7FFFFFFFu, prim = tris[tri].primitive;
    const uint32_t drawInstanceIndex = inst.customIndex + geom;
    const uint32_t seed = io.ids[m.z].y;
    if (COUNT) cnt->anyHits++;
    const AlphaUv *uvs = inst.uvs;
    bool accept = true;
    bool fast = uvs != nullptr && sv.alphaMats != nullptr;
    if (fast)
    {
        const AlphaMat am = loadAlphaMat(sv.alphaMats + drawInstanceIndex);
        fast = (am.bits & ALPHAMAT_USABLE) != 0;
        if (fast) accept = anyHitAcceptFast(am, loadAlphaUv(uvs + tri), u, v, seed);
    }
    if (!fast) accept = anyHitAccept(sv, drawInstanceIndex, prim, u, v, seed);
#if PRT_BLAS_ALPHA_SPLIT
    if (!accept) c.pend[1] = make_uint4(m.w, geom, prim, 1u);
#endif
    if (accept && commitHit<SHADOW>(c, as, io, c.row(ROW_O), m.z, m.w, geom, prim, t, u, v)) return KIND_IDLE;
    return traceFinishItem(c, as, io, g, sp, blasP1, stackTop(c, sp));
}

// KIND_INSTANCE: one pending primitive of a TLAS leaf = enter that instance (the TLAS groups go on the stack).
template <bool COUNT>
HD int traceInstanceItem(const RayCtx &c, const Accel &as, const TraceIO &io, TraceCounters *cnt)
{
    uint4 g = c.row(ROW_G);
    uint32_t sp = stateSp(g.y);
    g.y &= ~TRACE_STATE;
    const uint4 m = c.row(ROW_M);
    const int pbit = bfind(g.w);
    g.w &= ~(1u << pbit);
    const uint32_t instPos = g.z + (uint32_t)pbit;
    const InstRec &inst = as.instances[instPos];
    if (COUNT) cnt->instances++;
    if ((inst.mask & 0xFFu) == 0) return traceFinishItem(c, as, io, g, sp, 0u, stackTop(c, sp));
    if (g.y & 0xFF000000u) stackPush(c, sp, g.x, g.y, io);
    if (g.w != 0) stackPush(c, sp, g.z, g.w, io);
    const float4 wo = io.rayO[m.z], wd = io.rayD[m.z];
    storeRaySpace(c, xformPoint(inst.w2o, V3{wo.x, wo.y, wo.z}), xformVector(inst.w2o, V3{wd.x, wd.y, wd.z}), u2f(c.row(ROW_O).w),
                  c.row(ROW_I).w & TRACE_HAS_HIT, wo.w);
    c.row(ROW_G) = make_uint4(0u, 0x80000000u | (sp << 8) | ((sp + 1u) << 16), 0u, 0u);
    c.row(ROW_M) = make_uint4(arenaUnits(as, inst.nodes), arenaUnits(as, inst.tris), m.z, instPos);
    return KIND_NODE;
}

template <bool SHADOW, bool COUNT>
HD int traceItem(int kind, const RayCtx &c, const Accel &as, const SceneView &sv, const TraceIO &io, TraceCounters *cnt)
{
    switch (kind)
    {
    case KIND_NODE: return traceNodeItem<COUNT>(c, as, io, cnt);
    case KIND_TRI: return traceTriItem<SHADOW, COUNT>(c, as, io, cnt);
    case KIND_ANYHIT: return traceAnyHitItem<SHADOW, COUNT>(c, as, sv, io, cnt);
    default: return traceInstanceItem<COUNT>(c, as, io, cnt);
    }
}

// traceRayEXT to completion for one ray (CPU simulator, prt_trace_rays). Returns true on a hit.
template <bool SHADOW, bool COUNT>
HDNI bool traceRay(const Accel &as, const SceneView &sv, V3 o, V3 d, float tMin, float tMax, uint32_t seed, HitRec &best,
                   TraceCounters *cnt, unsigned long long *overflow)
{
    uint4 rows[POOL_ROWS];
    uint2 stack[TRACE_STACK];
    uint4 pend[PEND_ROWS];
    const RayCtx c = {rows, 1u, stack, pend};
    const float4 wo = make_float4(o.x, o.y, o.z, tMin), wd = make_float4(d.x, d.y, d.z, tMax);
    const uint2 id = make_uint2(0u, seed);
    uint4 hit = make_uint4(MISS_INDEX, MISS_INDEX, 0u, 0u);
    uint2 tie = make_uint2(MISS_INDEX, 0u);
    const TraceIO io = {&wo, &wd, &id, &hit, &tie, overflow};
    rows[ROW_I] = make_uint4(0u, 0u, 0u, 0u);
    rows[ROW_O] = make_uint4(0u, 0u, 0u, f2u(tMax));
    int kind = traceBegin(c, as, io, 0u, COUNT ? cnt : nullptr);
    while (kind != KIND_IDLE) kind = traceItem<SHADOW, COUNT>(kind, c, as, sv, io, cnt);
    const bool found = (rows[ROW_I].w & TRACE_HAS_HIT) != 0;
    best.drawInstanceIndex = SHADOW ? (found ? 0u : MISS_INDEX) : hit.x;
    best.primitiveID = hit.y;
    best.u = u2f(hit.z); best.v = u2f(hit.w);
    best.t = u2f(rows[ROW_O].w);
    best.instance = tie.x; best.geometry = tie.y;
    return found;
}

} // namespace prt
