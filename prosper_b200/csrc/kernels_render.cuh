// kernels_render.cuh -- the wavefront form of res/shader/rt/reference/main.rgen.
//
// One rgen invocation (pixel, frameIndex) = one path with a fixed slot in the path-state arrays.
// The bounce loop main.rgen:241-283 is cut at its two traceRayEXT calls into stages that communicate
// through compacted queues in HBM:
//   raygen -> [ trace_closest -> shade -> trace_shadow ] x maxBounces -> accumulate
// Live counts stay in device memory (one counter per bounce), so the host enqueues the whole frame
// without a single synchronisation. Compaction = warp ballot + one atomic per warp per queue.
#pragma once
#include "core_shade.h"
#include "core_trace.h"

namespace prt {

constexpr int RENDER_THREADS = 128;
constexpr uint32_t COUNT_SHADOW_BASE = 8; // counts[b] = closest rays of bounce b, counts[8 + b] = shadow rays
constexpr uint32_t CURSOR_BASE = 16;      // counts[16 + i] = fetch cursor of the persistent trace kernel over queue i
constexpr uint32_t COUNT_WORDS = 32;
// Persistent trace kernels (k_trace below): every thread owns PRT_POOL_RAYS rays.
#ifndef PRT_POOL_RAYS
#define PRT_POOL_RAYS 2
#endif
#ifndef PRT_POOL_REFILL
#define PRT_POOL_REFILL (8 * PRT_POOL_RAYS) // idle ray slots per warp (of 32 * PRT_POOL_RAYS) that trigger a refill
#endif
#ifndef PRT_TRACE_SINGLE_KIND
#define PRT_TRACE_SINGLE_KIND 0 // 1: only the most wanted kind runs per census
#endif
#ifndef PRT_TRACE_MIN_BLOCKS
#define PRT_TRACE_MIN_BLOCKS 6
#endif

struct PathBuffers
{
    uint4 *rng;          // per path: pcg_state.xyz
    float4 *throughput;  // per path
    float4 *color;       // per path: the rgen `color` accumulator
    float4 *rayO[2];     // ray queues (ping-pong): origin.xyz, tMin
    float4 *rayD[2];     // direction.xyz, tMax
    uint2 *rayId[2];     // path index, payload.randomSeed
    uint4 *hits;         // per closest-ray slot: drawInstanceIndex, primitiveID, bary bits
    uint2 *tie;          // per closest-ray slot: (instance index, geometry index) of the committed hit (tie-break keys)
    float4 *shadowO;     // shadow queue
    float4 *shadowD;
    uint2 *shadowId;
    float4 *shadowC;     // contribution if unoccluded
    uint32_t *counts;    // [16]
};

struct GlobalCounters
{
    unsigned long long closestRays, shadowRays, nodesVisited, trisTested, instancesEntered, anyHitCalls, shadedHits, textureTaps,
        pixelSamples, stackOverflows, shadowNodes, shadowTris, shadowInstances, shadowAnyHits;
};

__device__ __forceinline__ uint32_t appendSlot(uint32_t *counter, bool pred)
{
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, pred);
    if (m == 0) return 0;
    const int leader = __ffs(m) - 1;
    uint32_t base = 0;
    if ((int)lane == leader) base = atomicAdd(counter, (uint32_t)__popc(m));
    base = __shfl_sync(0xFFFFFFFFu, base, leader);
    return base + (uint32_t)__popc(m & ((1u << lane) - 1u));
}

__device__ __forceinline__ void flushTraceCounters(GlobalCounters *g, const TraceCounters &c, bool shadow = false)
{
    const uint32_t n = __reduce_add_sync(0xFFFFFFFFu, c.nodes), t = __reduce_add_sync(0xFFFFFFFFu, c.tris);
    const uint32_t i = __reduce_add_sync(0xFFFFFFFFu, c.instances), a = __reduce_add_sync(0xFFFFFFFFu, c.anyHits);
    const uint32_t o = __reduce_add_sync(0xFFFFFFFFu, c.overflow);
    if ((threadIdx.x & 31u) == 0)
    {
        if (n) atomicAdd(&g->nodesVisited, (unsigned long long)n);
        if (t) atomicAdd(&g->trisTested, (unsigned long long)t);
        if (i) atomicAdd(&g->instancesEntered, (unsigned long long)i);
        if (a) atomicAdd(&g->anyHitCalls, (unsigned long long)a);
        if (o) atomicAdd(&g->stackOverflows, (unsigned long long)o);
        if (shadow)
        {
            if (n) atomicAdd(&g->shadowNodes, (unsigned long long)n);
            if (t) atomicAdd(&g->shadowTris, (unsigned long long)t);
            if (i) atomicAdd(&g->shadowInstances, (unsigned long long)i);
            if (a) atomicAdd(&g->shadowAnyHits, (unsigned long long)a);
        }
    }
}

// main.rgen:225-240
__global__ void __launch_bounds__(RENDER_THREADS)
    k_raygen(SceneView sv, prt_reference_pc pc, uint32_t width, uint32_t height, uint32_t samples, PathBuffers pb)
{
    const uint32_t pixels = width * height;
    const uint32_t total = pixels * samples;
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p == 0)
    {
        pb.counts[0] = total;
        for (uint32_t i = 1; i < COUNT_WORDS; ++i) pb.counts[i] = 0;
    }
    if (p >= total) return;
    const uint32_t k = p / pixels, pix = p - k * pixels;
    const uint32_t px = pix % width, py = pix / width;
    const uint32_t frameIndex = (pc.frameIndex + k) % PRT_FRAME_PERIOD; // RtRef.cpp:170
    Rng rng;
    const Ray ray = generatePrimaryRay(sv, pc, frameIndex, px, py, width, height, rng);
    pb.rng[p] = make_uint4(rng.state.x, rng.state.y, rng.state.z, 0u);
    pb.throughput[p] = make_float4(1.f, 1.f, 1.f, 0.f);
    pb.color[p] = make_float4(0.f, 0.f, 0.f, 0.f);
    pb.rayO[0][p] = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.tMin);
    pb.rayD[0][p] = make_float4(ray.d.x, ray.d.y, ray.d.z, ray.tMax);
    pb.rayId[0][p] = make_uint2(p, pcg(rng.state.x ^ rng.state.z)); // main.rgen:71
}

// ---- ray reordering -------------------------------------------------------------------------------------------------
// Secondary rays leave k_shade in path order, i.e. spatially scattered. A counting sort of the queue INDICES by
// (direction octant, Morton cell of the origin on a 2^bits grid over the TLAS root's quantisation frame) hands each warp
// of the persistent trace kernels 32 rays that start close together and walk the tree in the same child order, which is
// what the per-step vote and L1 want. Results do not depend on the order (every ray writes its own slot), so hit IDs and
// radiance are unchanged bit for bit. Three small launches per queue: bin (key + histogram), scan, scatter.
// a kind runs in a round when at least PRT_BAR_*_Q quarters as many lanes want it as want the most wanted kind
#ifndef PRT_BAR_NODE_Q
#define PRT_BAR_NODE_Q 2
#endif
#ifndef PRT_BAR_TRI_Q
#define PRT_BAR_TRI_Q 2
#endif
#ifndef PRT_BAR_ANYHIT
#define PRT_BAR_ANYHIT 12 // lanes (of 32) holding a parked any-hit candidate that make the any-hit test run although another kind is wanted more
                          // (8 / 12 / 16: 1068 / 1082 / 1070 Mrays/s; 12 won by 1.3 - 2.1 % in four A/B pairs on three boxes)
#endif
#ifndef PRT_RAY_SORT_CELL_BITS
#define PRT_RAY_SORT_CELL_BITS 5
#endif
#ifndef PRT_RAY_SORT_CELL_MAJOR
#define PRT_RAY_SORT_CELL_MAJOR 0
#endif
constexpr uint32_t RAY_SORT_BINS = 8u << (3 * PRT_RAY_SORT_CELL_BITS);
constexpr uint32_t RAY_SCAN_THREADS = 1024;

__device__ __forceinline__ uint32_t spreadBits3(uint32_t v)
{
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}

__global__ void k_ray_bin(Accel as, const float4 *__restrict__ rayO, const float4 *__restrict__ rayD,
                          const uint32_t *__restrict__ countPtr, uint32_t *__restrict__ keys, uint32_t *__restrict__ hist)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= *countPtr) return;
    const WideNode *root = as.tlasNodes;
    const float4 o = rayO[i], d = rayD[i];
    const float cells = (float)(1u << PRT_RAY_SORT_CELL_BITS);
    // the root's quantisation frame spans origin + [0, 256) * 2^(e - 127) per axis
    const float ix = cells / (256.f * __uint_as_float((uint32_t)root->ex << 23));
    const float iy = cells / (256.f * __uint_as_float((uint32_t)root->ey << 23));
    const float iz = cells / (256.f * __uint_as_float((uint32_t)root->ez << 23));
    const float top = cells - 1.f;
    const uint32_t cx = (uint32_t)fminf(fmaxf((o.x - root->px) * ix, 0.f), top);
    const uint32_t cy = (uint32_t)fminf(fmaxf((o.y - root->py) * iy, 0.f), top);
    const uint32_t cz = (uint32_t)fminf(fmaxf((o.z - root->pz) * iz, 0.f), top);
    const uint32_t octant = (d.x < 0.f ? 1u : 0u) | (d.y < 0.f ? 2u : 0u) | (d.z < 0.f ? 4u : 0u);
#if PRT_RAY_SORT_CELL_MAJOR
    const uint32_t key = (((spreadBits3(cx) << 2) | (spreadBits3(cy) << 1) | spreadBits3(cz)) << 3) | octant;
#else
    const uint32_t key = (octant << (3 * PRT_RAY_SORT_CELL_BITS)) | (spreadBits3(cx) << 2) | (spreadBits3(cy) << 1) | spreadBits3(cz);
#endif
    keys[i] = key;
    atomicAdd(&hist[key], 1u);
}

// exclusive scan of the histogram in place, one block
__global__ void __launch_bounds__(RAY_SCAN_THREADS) k_ray_scan(uint32_t *__restrict__ hist)
{
    constexpr uint32_t PER = RAY_SORT_BINS / RAY_SCAN_THREADS;
    __shared__ uint32_t warpSums[RAY_SCAN_THREADS / 32];
    uint32_t *mine = hist + threadIdx.x * PER;
    uint32_t sum = 0;
    for (uint32_t k = 0; k < PER; k += 4)
    {
        const uint4 v = *reinterpret_cast<const uint4 *>(mine + k);
        sum += v.x + v.y + v.z + v.w;
    }
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint32_t incl = sum;
    for (int off = 1; off < 32; off <<= 1)
    {
        const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, off);
        if ((int)lane >= off) incl += t;
    }
    if (lane == 31) warpSums[warp] = incl;
    __syncthreads();
    if (warp == 0)
    {
        uint32_t w = warpSums[lane];
        for (int off = 1; off < 32; off <<= 1)
        {
            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, w, off);
            if ((int)lane >= off) w += t;
        }
        warpSums[lane] = w;
    }
    __syncthreads();
    uint32_t run = incl - sum + (warp ? warpSums[warp - 1] : 0u);
    for (uint32_t k = 0; k < PER; ++k)
    {
        const uint32_t v = mine[k];
        mine[k] = run;
        run += v;
    }
}

__global__ void k_ray_scatter(const uint32_t *__restrict__ countPtr, const uint32_t *__restrict__ keys, uint32_t *__restrict__ hist,
                              uint32_t *__restrict__ perm)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= *countPtr) return;
    perm[atomicAdd(&hist[keys[i]], 1u)] = i;
}

// traceClosest (main.rgen:62-81) and shadow() (main.rgen:49-60, fused with the addBounce of evaluateDirectLighting's
// result): persistent warps over the closest-ray queue of `bounce`, or over its shadow queue (SHADOW).
//
// Every thread owns R rays whose state sits in shared memory ([row][ray][thread]: conflict-free 128-bit rows, core_trace.h)
// and whose stacks sit in its local memory. One round = ONE item (a node, a triangle, an any-hit test or an instance
// entry) for at most one ray per thread: a census of the kinds the warp's rays are waiting for (one REDUX), the most
// wanted kind runs, and a thread advances whichever of its rays wants it. With one ray per thread 14.5 of 32 lanes
// took part in an item (the r01 kernel: the lanes whose kind lost the vote idled); with R rays a lane idles only when
// none of its R rays wants the winning kind. Finished rays retire in place; when `refillAt` slots of the warp are
// idle they are refilled from the queue with one atomic (Aila & Laine's dynamic fetch).
template <int R, bool SHADOW, bool COUNT>
__global__ void __launch_bounds__(RENDER_THREADS, PRT_TRACE_MIN_BLOCKS)
    k_trace(Accel as, SceneView sv, PathBuffers pb, int q, uint32_t bounce, GlobalCounters *gc, const uint32_t *__restrict__ perm,
            uint32_t refillAt)
{
    __shared__ uint4 rowMem[POOL_ROWS * R * RENDER_THREADS];
    uint2 stack[R][TRACE_STACK]; // per-thread local memory, L1-resident
    uint4 pend[R][PEND_ROWS]; // parked any-hit candidate (+ the last rejected triangle in reference-splitting builds) per ray
    const uint32_t lane = threadIdx.x & 31u, ltMask = (1u << lane) - 1u;
    const uint32_t count = pb.counts[(SHADOW ? COUNT_SHADOW_BASE : 0u) + bounce];
    uint32_t *cursor = &pb.counts[CURSOR_BASE + (SHADOW ? COUNT_SHADOW_BASE : 0u) + bounce];
    const TraceIO io = {SHADOW ? pb.shadowO : pb.rayO[q], SHADOW ? pb.shadowD : pb.rayD[q], SHADOW ? pb.shadowId : pb.rayId[q],
                        SHADOW ? nullptr : pb.hits, SHADOW ? nullptr : pb.tie, &gc->stackOverflows};
    TraceCounters tc = {0, 0, 0, 0, 0};
    auto ctx = [&](int r) { return RayCtx{rowMem + r * RENDER_THREADS + threadIdx.x, (uint32_t)(R * RENDER_THREADS), stack[r], pend[r]}; };
    auto retire = [&](uint32_t slot, bool hit) {
        if (SHADOW)
        {
            if (!hit) // unoccluded
            {
                const float4 c = pb.shadowC[slot];
                const uint32_t path = pb.shadowId[slot].x;
                float4 col = pb.color[path];
                col.x += c.x; col.y += c.y; col.z += c.z;
                pb.color[path] = col;
            }
        }
        else if (!hit) pb.hits[slot] = make_uint4(MISS_INDEX, MISS_INDEX, 0u, 0u);
    };
    int kind[R]; // the item each owned ray waits for (registers: every index below is a compile-time constant)
#pragma unroll
    for (int r = 0; r < R; ++r) kind[r] = KIND_IDLE;
    bool exhausted = false;
    for (;;)
    {
        // census: lanes that hold a ray of each kind (6 bits per kind), idle slots (from bit 24)
        uint32_t mine = 0;
#pragma unroll
        for (int r = 0; r < R; ++r) mine = kind[r] == KIND_IDLE ? mine + (1u << 24) : mine | (1u << (6 * kind[r]));
        const uint32_t census = __reduce_add_sync(0xFFFFFFFFu, mine);
        const uint32_t idle = census >> 24;
        if (!exhausted && idle >= refillAt)
        {
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(cursor, idle);
            base = __shfl_sync(0xFFFFFFFFu, base, 0);
            exhausted = base + idle >= count;
#pragma unroll
            for (int r = 0; r < R; ++r)
            {
                const bool want = kind[r] == KIND_IDLE;
                const uint32_t m = __ballot_sync(0xFFFFFFFFu, want);
                const uint32_t next = base + (uint32_t)__popc(m & ltMask);
                base += (uint32_t)__popc(m);
                if (want && next < count)
                {
                    const uint32_t slot = perm ? perm[next] : next;
                    kind[r] = traceBegin(ctx(r), as, io, slot, COUNT ? &tc : nullptr);
                    if (kind[r] == KIND_IDLE) retire(slot, false);
                }
            }
            continue;
        }
        if ((census & 0xFFFFFFu) == 0) break; // nothing in flight and nothing left to fetch
        // One census serves a whole round: every kind that at least half as many lanes want as the most wanted one runs,
        // nodes first, and a lane may advance one ray per kind (after its node item a ray often wants a triangle item,
        // which it then gets in the same round). The any-hit test is long and rare: it waits for a quarter of the warp.
        const uint32_t n0 = census & 63u, n1 = (census >> 6) & 63u, n2 = (census >> 12) & 63u, n3 = (census >> 18) & 63u;
        const uint32_t most = max(max(n0, n1), max(n2, n3));
        const uint32_t barNode = (most * PRT_BAR_NODE_Q + 3u) >> 2, barTri = (most * PRT_BAR_TRI_Q + 3u) >> 2, bar = (most + 1u) >> 1;
        auto runKind = [&](const int run) {
            int r = -1;
#pragma unroll
            for (int i = R - 1; i >= 0; --i)
                if (kind[i] == run) r = i;
            if (r >= 0)
            {
                const RayCtx c = ctx(r);
                const int k = traceItem<SHADOW, COUNT>(run, c, as, sv, io, &tc);
                if (k == KIND_IDLE) retire(c.row(ROW_M).z, (c.row(ROW_I).w & TRACE_HAS_HIT) != 0);
#pragma unroll
                for (int i = 0; i < R; ++i)
                    if (i == r) kind[i] = k;
            }
        };
#if PRT_TRACE_SINGLE_KIND
        if (n2 == most) runKind(KIND_ANYHIT);
        else if (n1 == most) runKind(KIND_TRI);
        else if (n3 == most) runKind(KIND_INSTANCE);
        else runKind(KIND_NODE);
#else
        if (n0 >= barNode) runKind(KIND_NODE);
        if (n3 >= bar) runKind(KIND_INSTANCE);
        if (n1 >= barTri) runKind(KIND_TRI);
        if (n2 == most || n2 >= (uint32_t)PRT_BAR_ANYHIT) runKind(KIND_ANYHIT);
#endif
    }
    if (COUNT) flushTraceCounters(gc, tc, SHADOW);
}

// texture() through the texture units for one tap of every lane of the warp. All 32 lanes must call
// this together. The bindless handle differs from lane to lane; the loop hands the texture unit ONE
// handle per iteration (the lowest pending lane's), so the fetch itself always sees a warp-uniform
// handle. Besides being how the hardware wants it, this side-steps a ptxas 12.9 code-generation bug
// we hit on sm_100a: in its own divergent-handle ("waterfall") loop the coordinate moves were sunk
// into the loop over registers aliased with the result, corrupting .xy of lanes served earlier
// (DESIGN.md section 6).
__device__ __forceinline__ V4 warpTextureTap(const SceneView &sv, uint32_t textureSampler, bool wanted, V2 uv)
{
    V4 out = {1.f, 1.f, 1.f, 1.f};
    const TexSlot *slot = wanted ? texSlot(sv, textureSampler) : nullptr;
    unsigned long long handle = slot ? (unsigned long long)slot->linear : 0ull;
    bool pending = handle != 0ull;
    for (;;)
    {
        const uint32_t m = __ballot_sync(0xFFFFFFFFu, pending);
        if (m == 0) break;
        const int leader = __ffs(m) - 1;
        const unsigned long long h = __shfl_sync(0xFFFFFFFFu, handle, leader);
        if (pending && handle == h)
        {
            const float4 t = tex2D<float4>((cudaTextureObject_t)h, uv.x, uv.y);
            out = V4{t.x, t.y, t.z, t.w};
            pending = false;
        }
    }
    return out;
}

// One loop iteration of main.rgen:241-283 minus the traces.
#ifndef PRT_SHADE_MIN_BLOCKS
#define PRT_SHADE_MIN_BLOCKS 8 // 64 registers (48 B of spills): 88 ms per step against 98 (6 blocks) and 96 (10 blocks); the kernel waits on
                               // its gather chain -- prefetching the path state to L2 at the top changed nothing (88.7 vs 88.4)
#endif
__global__ void __launch_bounds__(RENDER_THREADS, PRT_SHADE_MIN_BLOCKS)
    k_shade(SceneView sv, prt_reference_pc pc, PathBuffers pb, int q, uint32_t bounce, GlobalCounters *gc, int countShaded)
{
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t count = pb.counts[bounce];
    ShadeResult r;
    r.shadowRay = false;
    r.alive = false;
    uint32_t path = 0;
    Rng rng;
    rng.state = {0u, 0u, 0u};
    const bool active = slot < count;
    uint4 h = make_uint4(MISS_INDEX, MISS_INDEX, 0u, 0u);
    float4 d = make_float4(0.f, 0.f, 0.f, 0.f);
    if (active)
    {
        h = pb.hits[slot];
        d = pb.rayD[q][slot];
        path = pb.rayId[q][slot].x;
    }
    const bool shaded = active && h.x != MISS_INDEX;
    SurfaceGeom geom;
    geom.materialIndex = 0;
    MaterialTaps taps;
    taps.base = taps.metallicRoughness = taps.normal = V4{1.f, 1.f, 1.f, 1.f};
    if (shaded) geom = surfaceGeometry(sv, h.x, h.y, __uint_as_float(h.z), __uint_as_float(h.w));
    uint32_t tapCount = 0;
    if (shaded)
    {
        const prt_material_data &md = sv.materials[geom.materialIndex];
        tapCount = ((md.baseColorTextureSampler & 0xFFFFFFu) > 0) + ((md.metallicRoughnessTextureSampler & 0xFFFFFFu) > 0) +
                   ((md.normalTextureSampler & 0xFFFFFFu) > 0);
    }
    if (sv.exactMaterialFilter)
    {
        if (shaded) taps = softwareMaterialTaps(sv, geom.materialIndex, geom.v.TexCoord0);
    }
    else
    {
        // warp-convergent: every lane takes part in the three taps
        uint32_t tb = 0, tm = 0, tn = 0;
        if (shaded)
        {
            const prt_material_data &md = sv.materials[geom.materialIndex];
            tb = md.baseColorTextureSampler; tm = md.metallicRoughnessTextureSampler; tn = md.normalTextureSampler;
        }
        const V2 uv = shaded ? geom.v.TexCoord0 : V2{0.f, 0.f};
        taps.base = warpTextureTap(sv, tb, (tb & 0xFFFFFFu) > 0, uv);
        taps.metallicRoughness = warpTextureTap(sv, tm, (tm & 0xFFFFFFu) > 0, uv);
        taps.normal = warpTextureTap(sv, tn, (tn & 0xFFFFFFu) > 0, uv);
    }
    if (active)
    {
        const uint4 rs = pb.rng[path];
        const float4 tp = pb.throughput[path];
        const float4 cl = pb.color[path];
        rng.state = {rs.x, rs.y, rs.z};
        V3 throughput = {tp.x, tp.y, tp.z}, color = {cl.x, cl.y, cl.z};
        shadePath(sv, pc, bounce, V3{d.x, d.y, d.z}, h.x, h.y, geom, taps, rng, throughput, color, r);
        pb.color[path] = make_float4(color.x, color.y, color.z, 0.f);
        if (r.alive)
        {
            pb.rng[path] = make_uint4(rng.state.x, rng.state.y, rng.state.z, 0u);
            pb.throughput[path] = make_float4(throughput.x, throughput.y, throughput.z, 0.f);
        }
    }
    const uint32_t ss = appendSlot(&pb.counts[COUNT_SHADOW_BASE + bounce], r.shadowRay);
    if (r.shadowRay)
    {
        pb.shadowO[ss] = make_float4(r.shadowO.x, r.shadowO.y, r.shadowO.z, 0.1f); // main.rgen:217 tMin
        pb.shadowD[ss] = make_float4(r.shadowD.x, r.shadowD.y, r.shadowD.z, r.shadowTMax);
        pb.shadowId[ss] = make_uint2(path, r.shadowSeed);
        pb.shadowC[ss] = make_float4(r.shadowContribution.x, r.shadowContribution.y, r.shadowContribution.z, 0.f);
    }
    const uint32_t ns = appendSlot(&pb.counts[bounce + 1], r.alive);
    if (r.alive)
    {
        pb.rayO[q ^ 1][ns] = make_float4(r.next.o.x, r.next.o.y, r.next.o.z, r.next.tMin);
        pb.rayD[q ^ 1][ns] = make_float4(r.next.d.x, r.next.d.y, r.next.d.z, r.next.tMax);
        pb.rayId[q ^ 1][ns] = make_uint2(path, pcg(rng.state.x ^ rng.state.z));
    }
    if (countShaded)
    {
        const uint32_t n = __popc(__ballot_sync(0xFFFFFFFFu, shaded));
        const uint32_t tp = __reduce_add_sync(0xFFFFFFFFu, tapCount);
        if ((threadIdx.x & 31u) == 0 && n) atomicAdd(&gc->shadedHits, (unsigned long long)n);
        if ((threadIdx.x & 31u) == 0 && tp) atomicAdd(&gc->textureTaps, (unsigned long long)tp);
    }
}

// Adds the per-bounce live counts of the finished batch to the 64-bit totals.
__global__ void k_tally(PathBuffers pb, GlobalCounters *gc, uint32_t pixelSamples)
{
    if (threadIdx.x == 0)
    {
        unsigned long long c = 0, s = 0;
        for (uint32_t b = 0; b < MAX_BOUNCES; ++b)
        {
            c += pb.counts[b];
            s += pb.counts[COUNT_SHADOW_BASE + b];
        }
        gc->closestRays += c;
        gc->shadowRays += s;
        gc->pixelSamples += pixelSamples;
    }
}

// main.rgen:285-298, frame after frame in frameIndex order (the same order of operations as the
// reference's one-frame-per-dispatch running mean).
__global__ void k_accumulate(PathBuffers pb, float4 *__restrict__ accum, uint32_t pixels, uint32_t samples, uint32_t flags)
{
    const uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= pixels) return;
    float4 h = accum[pix];
    const bool accumulate = (flags >> 1) & 1u;
    for (uint32_t k = 0; k < samples; ++k)
    {
        const float4 c = pb.color[k * pixels + pix];
        const bool skipHistory = k == 0 && (flags & 1u);
        if (skipHistory || !accumulate) h = make_float4(c.x, c.y, c.z, 1.f);
        else
        {
            const float n = h.w + 1.f;
            h.x = h.x + (c.x - h.x) / n;
            h.y = h.y + (c.y - h.y) / n;
            h.z = h.z + (c.z - h.z) / n;
            h.w = n;
        }
    }
    accum[pix] = h;
}

// The fp32 -> fp16 blit that produces Output::illumination (RtRef.cpp:336-377)
__global__ void k_blit_f16(const float4 *__restrict__ accum, uint2 *__restrict__ out, uint32_t pixels)
{
    const uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= pixels) return;
    const float4 c = accum[pix];
    const uint32_t lo = (uint32_t)floatToHalf(c.x) | ((uint32_t)floatToHalf(c.y) << 16);
    const uint32_t hi = (uint32_t)floatToHalf(c.z) | ((uint32_t)floatToHalf(c.w) << 16);
    out[pix] = make_uint2(lo, hi);
}

// running mean (rgb, n) <-> sum (rgb * n, n) for the multi-GPU reduce
__global__ void k_mean_to_sum(const float4 *__restrict__ accum, float4 *__restrict__ sum, uint32_t pixels)
{
    const uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= pixels) return;
    const float4 c = accum[pix];
    sum[pix] = make_float4(c.x * c.w, c.y * c.w, c.z * c.w, c.w);
}
__global__ void k_sum_to_mean(const float4 *__restrict__ sum, float4 *__restrict__ accum, uint32_t pixels)
{
    const uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= pixels) return;
    const float4 s = sum[pix];
    const float inv = s.w > 0.f ? 1.f / s.w : 0.f;
    accum[pix] = make_float4(s.x * inv, s.y * inv, s.z * inv, s.w);
}

// prt_trace_rays: caller-supplied rays
template <bool SHADOW>
__global__ void __launch_bounds__(RENDER_THREADS)
    k_trace_user(Accel as, SceneView sv, const float4 *__restrict__ o, const float4 *__restrict__ d, const uint32_t *__restrict__ seeds,
                 uint32_t n, prt_hit *__restrict__ out, GlobalCounters *gc)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    TraceCounters tc = {0, 0, 0, 0, 0};
    if (i < n)
    {
        const float4 oo = o[i], dd = d[i];
        HitRec hit;
        traceRay<SHADOW, true>(as, sv, V3{oo.x, oo.y, oo.z}, V3{dd.x, dd.y, dd.z}, oo.w, dd.w, seeds ? seeds[i] : 0u, hit, &tc, &gc->stackOverflows);
        prt_hit h;
        h.drawInstanceIndex = hit.drawInstanceIndex;
        h.primitiveID = hit.primitiveID;
        h.bary[0] = hit.u;
        h.bary[1] = hit.v;
        out[i] = h;
    }
    flushTraceCounters(gc, tc, SHADOW);
}

__global__ void k_hits_to_records(const uint4 *__restrict__ hits, prt_hit *__restrict__ out, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 h = hits[i];
    prt_hit r;
    r.drawInstanceIndex = h.x;
    r.primitiveID = h.y;
    r.bary[0] = __uint_as_float(h.z);
    r.bary[1] = __uint_as_float(h.w);
    out[i] = r;
}

} // namespace prt
