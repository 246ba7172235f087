// kernels_build.cuh -- launches of the acceleration-structure builder (bodies in core_bvh.h).
// BLAS: World::Impl::buildNextBlas (World.cpp:585-747). TLAS: updateTlasInstances + buildCurrentTlas
// (World.cpp:878-928, 749-802), rebuilt from scratch every frame like the reference does.
#pragma once
#include <cooperative_groups.h>
#include "../../include/prt_types.h"
#include "core_build.h"
#include "kernels_sort.cuh"

namespace prt {

// order-preserving float <-> uint mapping for atomicMin/atomicMax on bounds
__device__ __forceinline__ uint32_t floatOrdered(float f)
{
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ __forceinline__ float orderedFloat(uint32_t k)
{
    const uint32_t b = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k;
#ifdef __CUDA_ARCH__
    return __uint_as_float(b);
#else
    float f;
    memcpy(&f, &b, 4);
    return f;
#endif
}

// bounds[0..2] = min xyz, bounds[3..5] = max xyz (ordered-uint encoding), bounds[6] = max |coord|
struct BuildBounds
{
    uint32_t v[8];
};

__global__ void k_init_bounds(BuildBounds *b, uint32_t *counters, uint32_t nCounters)
{
    if (threadIdx.x < 3) b->v[threadIdx.x] = 0xFFFFFFFFu;
    else if (threadIdx.x < 8) b->v[threadIdx.x] = 0u;
    if (threadIdx.x < nCounters) counters[threadIdx.x] = 0u;
}

// Every lane of the warp must call this (invalid lanes pass valid = false).
__device__ __forceinline__ void atomicBounds(BuildBounds *b, V3 lo, V3 hi, bool valid)
{
    const float l[3] = {lo.x, lo.y, lo.z}, h[3] = {hi.x, hi.y, hi.z};
#pragma unroll
    for (int a = 0; a < 3; ++a)
    {
        const uint32_t mn = __reduce_min_sync(0xFFFFFFFFu, valid ? floatOrdered(l[a]) : 0xFFFFFFFFu);
        const uint32_t mx = __reduce_max_sync(0xFFFFFFFFu, valid ? floatOrdered(h[a]) : 0u);
        if ((threadIdx.x & 31u) == 0u)
        {
            atomicMin(&b->v[a], mn);
            atomicMax(&b->v[3 + a], mx);
        }
    }
}

// One launch per sub-model (= BLAS geometry, World.cpp:617-668): triangle records + unpadded boxes.
__global__ void k_blas_gather(SceneView sv, uint32_t meshIndex, uint32_t primCount, uint32_t geometryIndex, uint32_t opaque,
                              uint32_t outOffset, TriRec *__restrict__ tris, AlphaUv *__restrict__ uvs, Box *__restrict__ boxes,
                              BuildBounds *bounds)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= primCount)
    {
        atomicBounds(bounds, V3{0.f, 0.f, 0.f}, V3{0.f, 0.f, 0.f}, false);
        return;
    }
    TriRec t;
    V3 lo, hi, corner[3];
    gatherTriangle(sv, meshIndex, p, geometryIndex, opaque, t, lo, hi, corner);
    AlphaUv uv = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, {0u, 0u}};
    if (!opaque) gatherTriangleUvs(sv, meshIndex, p, uv);
    // a non-opaque triangle enters the builder as several references with the same record (core_build.h)
    const uint32_t refs = blasRefsPerTriangle(opaque);
    for (uint32_t k = 0; k < refs; ++k)
    {
        const uint32_t slot = outOffset + p * refs + k;
        V3 rl = lo, rh = hi;
        if (refs > 1u) triangleRefBox(corner[0], corner[1], corner[2], k, rl, rh);
        tris[slot] = t;
        uvs[slot] = uv;
        boxes[slot] = makeBox(rl, rh);
    }
    atomicBounds(bounds, lo, hi, true);
}

// Pads the primitive boxes (so a box never culls a triangle the exact test accepts: the oracle's
// rule, pad = max(max|coord| / 16384, 1e-6)) and computes the Morton keys of the box centres.
__global__ void k_pad_and_keys(Box *__restrict__ boxes, uint32_t n, const BuildBounds *__restrict__ bounds, int padMode,
                               uint64_t keyMask, uint64_t *__restrict__ keys, uint32_t *__restrict__ order)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    V3 lo = {orderedFloat(bounds->v[0]), orderedFloat(bounds->v[1]), orderedFloat(bounds->v[2])};
    V3 hi = {orderedFloat(bounds->v[3]), orderedFloat(bounds->v[4]), orderedFloat(bounds->v[5])};
    Box b = boxes[i];
    keys[i] = padBoxAndKey(b, lo, hi, padMode, keyMask);
    if (padMode == 1) boxes[i] = b;
    order[i] = i;
}

__global__ void k_lbvh_nodes(Lbvh t)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i + 1 >= t.count) return;
    lbvhBuildNode(t, (int)i);
    t.visit[i] = 0;
}

__global__ void k_lbvh_refit(Lbvh t)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= t.count) return;
    lbvhRefitFromLeaf(t, i);
}

// ---- PLOC (core_bvh.h): per iteration nearest -> merge (+ per-block survivor counts) -> block offsets -> scatter.
// The live cluster count stays in device memory (counts[it & 1]), so iterations are enqueued without host waits.
constexpr int PLOC_THREADS = 256;

__global__ void k_ploc_init(Lbvh t, int32_t *ref, Box *box, uint32_t *counts, uint32_t *nodeAlloc)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { counts[0] = t.count; counts[1] = 0; *nodeAlloc = 0; }
    if (i >= t.count) return;
    ref[i] = ~(int32_t)i;
    box[i] = t.primBox[t.order[i]];
    t.parentOfLeaf[i] = -1;
}

__global__ void __launch_bounds__(PLOC_THREADS) k_ploc_nearest(Ploc p, const uint32_t *__restrict__ count)
{
    const uint32_t m = *count, i = blockIdx.x * blockDim.x + threadIdx.x;
    if (m <= 1u || i >= m) return;
    plocNearest(p, i, m);
}

// merge + count of surviving clusters per block
__global__ void __launch_bounds__(PLOC_THREADS) k_ploc_merge(Ploc p, const uint32_t *__restrict__ count, uint32_t *__restrict__ blockCounts)
{
    const uint32_t m = *count, i = blockIdx.x * blockDim.x + threadIdx.x;
    if (m <= 1u) return;
    bool keep = false;
    if (i < m)
    {
        plocMerge(p, i, m);
        keep = p.mergedRef[i] != PLOC_DROPPED;
    }
    const int n = __syncthreads_count(keep);
    if (threadIdx.x == 0) blockCounts[blockIdx.x] = (uint32_t)n;
}

// exclusive scan of the per-block counts (one block; the array is count / 256 long) + next live count
__global__ void __launch_bounds__(1024) k_ploc_offsets(const uint32_t *__restrict__ count, uint32_t *__restrict__ blockCounts, uint32_t *__restrict__ nextCount)
{
    __shared__ uint32_t warpSums[32];
    __shared__ uint32_t carry;
    const uint32_t m = *count;
    if (m <= 1u) { if (threadIdx.x == 0) *nextCount = m; return; }
    const uint32_t blocks = (m + PLOC_THREADS - 1) / PLOC_THREADS;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < blocks; base += blockDim.x)
    {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < blocks ? blockCounts[i] : 0u;
        uint32_t x = v;
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o); if ((threadIdx.x & 31) >= o) x += y; }
        if ((threadIdx.x & 31) == 31) warpSums[threadIdx.x >> 5] = x;
        __syncthreads();
        if (threadIdx.x < 32)
        {
            uint32_t w = warpSums[threadIdx.x], z = w;
            for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, z, o); if ((int)threadIdx.x >= o) z += y; }
            warpSums[threadIdx.x] = z - w; // exclusive
        }
        __syncthreads();
        const uint32_t excl = carry + warpSums[threadIdx.x >> 5] + x - v;
        if (i < blocks) blockCounts[i] = excl;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) *nextCount = carry;
}

__global__ void __launch_bounds__(PLOC_THREADS) k_ploc_scatter(Ploc p, const uint32_t *__restrict__ count, const uint32_t *__restrict__ blockOffsets)
{
    __shared__ uint32_t warpBase[PLOC_THREADS / 32];
    const uint32_t m = *count, i = blockIdx.x * blockDim.x + threadIdx.x;
    if (m <= 1u) return;
    const bool keep = i < m && p.mergedRef[i] != PLOC_DROPPED;
    const uint32_t ballot = __ballot_sync(0xFFFFFFFFu, keep);
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    if (lane == 0) warpBase[warp] = (uint32_t)__popc(ballot);
    __syncthreads();
    if (threadIdx.x == 0)
    {
        uint32_t run = 0;
        for (int w = 0; w < PLOC_THREADS / 32; ++w) { const uint32_t c = warpBase[w]; warpBase[w] = run; run += c; }
    }
    __syncthreads();
    if (keep)
    {
        const uint32_t dst = blockOffsets[blockIdx.x] + warpBase[warp] + (uint32_t)__popc(ballot & ((1u << lane) - 1u));
        p.refOut[dst] = p.mergedRef[i];
        p.boxOut[dst] = p.mergedBox[i];
    }
}

__global__ void k_ploc_finish_leaves(Lbvh t, uint32_t *newPos, uint32_t *orderOut)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= t.count) return;
    plocFinishLeaf(t, i, newPos, orderOut);
}
__global__ void k_ploc_finish_nodes(Lbvh t, const uint32_t *newPos)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i + 1 >= t.count) return;
    plocFinishNode(t, i, newPos);
    plocRelabelNode(t, i, newPos);
}

// ---- binned SAH (core_bvh.h): ONE cooperative launch builds every level --------------------------------------------
// Per level: bins (per slot) -> split (one warp per active range, from the level's queue) -> partition (per slot; it
// also gathers the centroid bounds of the child ranges, so only level 0 has a centroid pass of its own), with a grid
// barrier between the passes; the host neither launches per level nor reads a counter back. Near the root most tiles lie
// inside one range: those reduce in the warp / bin in shared memory first, so the range's global counters see one update
// per warp or block instead of one per primitive. The tree is the one the per-slot passes of core_bvh.h give (hostsim).
constexpr int SAH_THREADS = 256;
static_assert(SAH_RANGE_BINS <= SAH_THREADS, "one thread per bin clears and flushes the block's bins");
static_assert(SAH_AXES == 1 && SAH_BINS <= 64, "the warp split takes two bins of one axis per lane");

struct SahGrid
{
    uint32_t *idx[2], *key[2], *queue[2]; // ping-pong: slot -> primitive, slot -> range key, active range keys of a level
    SahRange *range[2];
    uint32_t keys;                        // count / 2 + 1
    uint32_t *levelsOut;                  // levels run; SAH_MAX_LEVELS + 1 when the budget ran out
};

// true when every slot of tile [base, base + SAH_THREADS) belongs to the same active range (ranges are contiguous)
__device__ __forceinline__ bool sahTileUniform(const Sah &s, uint32_t base, uint32_t &key)
{
    const uint32_t lastSlot = min(base + (uint32_t)SAH_THREADS, s.t.count) - 1u;
    key = s.keyIn[base];
    return key != SAH_FINAL && key == s.keyIn[lastSlot];
}

__device__ __forceinline__ void sahCentroidTile(const Sah &s, uint32_t base)
{
    const uint32_t slot = base + threadIdx.x;
    uint32_t key;
    if (!sahTileUniform(s, base, key))
    {
        if (slot < s.t.count) sahCentroidPass(s, slot);
        return;
    }
    uint32_t lo[4] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, hi[4] = {0u, 0u, 0u, 0u}; // xyz + primitive index
    if (slot < s.t.count)
    {
        const uint32_t prim = s.idxIn[slot];
        const Box b = s.t.primBox[prim];
        for (uint32_t a = 0; a < 3; ++a) lo[a] = hi[a] = ordF(sahCentroid(b, a));
        lo[3] = hi[3] = prim;
    }
    for (uint32_t a = 0; a < 4; ++a)
    {
        lo[a] = __reduce_min_sync(0xFFFFFFFFu, lo[a]);
        hi[a] = __reduce_max_sync(0xFFFFFFFFu, hi[a]);
    }
    if ((threadIdx.x & 31u) == 0u)
    {
        uint32_t *cent = &s.cent[(size_t)key * SAH_CENT];
        for (uint32_t a = 0; a < 3; ++a) { atomicMin(&cent[a], lo[a]); atomicMax(&cent[3u + a], hi[a]); }
        atomicMin(&cent[6], lo[3]);
        atomicMax(&cent[7], hi[3]);
    }
}

__device__ __forceinline__ void sahBinTile(const Sah &s, uint32_t base, SahBin *sb)
{
    const uint32_t slot = base + threadIdx.x;
    uint32_t key;
    if (!sahTileUniform(s, base, key)) // block-uniform branch
    {
        if (slot < s.t.count) sahBinPass(s, slot);
        return;
    }
    const SahRange r = s.rangeIn[key];
    const uint32_t used = sahBinsFor(r.count) * SAH_AXES; // same [axis slot][bin] layout as the range's own bins
    if (threadIdx.x < used) sahBinClear(&sb[threadIdx.x]);
    __syncthreads();
    if (slot < s.t.count) sahBinInto(s, slot, key, r.count, sb);
    __syncthreads();
    if (threadIdx.x < used && sb[threadIdx.x].count) sahBinMerge(&s.bins[(size_t)r.first * SAH_SLOT_BINS + threadIdx.x], sb[threadIdx.x]);
    __syncthreads(); // the next tile of this block clears sb again
}

// running union of bins: planes in ordF form + primitive count (the identity is an empty bin)
struct SahAcc
{
    uint32_t lo[3], hi[3], cnt;
};
__device__ __forceinline__ SahAcc sahAccEmpty() { return SahAcc{{0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, {0u, 0u, 0u}, 0u}; }
__device__ __forceinline__ SahAcc sahAccUnion(const SahAcc &x, const SahAcc &y)
{
    SahAcc r;
    for (int a = 0; a < 3; ++a) { r.lo[a] = min(x.lo[a], y.lo[a]); r.hi[a] = max(x.hi[a], y.hi[a]); }
    r.cnt = x.cnt + y.cnt;
    return r;
}
template <bool UP> __device__ __forceinline__ SahAcc sahAccShift(const SahAcc &x, uint32_t d)
{
    SahAcc r;
    for (int a = 0; a < 3; ++a)
    {
        r.lo[a] = UP ? __shfl_up_sync(0xFFFFFFFFu, x.lo[a], d) : __shfl_down_sync(0xFFFFFFFFu, x.lo[a], d);
        r.hi[a] = UP ? __shfl_up_sync(0xFFFFFFFFu, x.hi[a], d) : __shfl_down_sync(0xFFFFFFFFu, x.hi[a], d);
    }
    r.cnt = UP ? __shfl_up_sync(0xFFFFFFFFu, x.cnt, d) : __shfl_down_sync(0xFFFFFFFFu, x.cnt, d);
    return r;
}
// cost of the border with `l` on its left and `r` on its right, as sahSplitPass computes it; >= 3.4e38 = no candidate
__device__ __forceinline__ float sahBorderCost(const SahAcc &l, const SahAcc &r, bool byIndex)
{
    if (l.cnt == 0u || r.cnt == 0u) return 3.4e38f;
    if (byIndex) return (float)(l.cnt > r.cnt ? l.cnt - r.cnt : r.cnt - l.cnt);
    const float areaR = sahOrdHalfArea(r.lo, r.hi);
    return sahOrdHalfArea(l.lo, l.hi) * (float)((l.cnt + SAH_GRANULE - 1u) / SAH_GRANULE) + areaR * (float)((r.cnt + SAH_GRANULE - 1u) / SAH_GRANULE);
}

// sahSplitPass by one warp: lane L holds bins 2L and 2L + 1, prefix / suffix unions come from shuffle scans, every lane
// prices its two borders, the cheapest (ties: the lowest border, as the serial sweep keeps the first) wins.
__device__ __forceinline__ void sahSplitWarp(const Sah &s, uint32_t key, uint32_t *queueOut)
{
    const uint32_t lane = threadIdx.x & 31u;
    const SahRange r = s.rangeIn[key];
    const uint32_t nb = sahBinsFor(r.count);
    uint32_t *cent = &s.cent[(size_t)key * SAH_CENT];
    const SahAxes ax = sahAxes(cent, s.level, nb);
    __syncwarp();
    if (lane < SAH_CENT) cent[lane] = (lane < 3u || lane == 6u) ? 0xFFFFFFFFu : 0u; // sahCentClear: the left child reuses this key
    SahAcc a = sahAccEmpty(), b = sahAccEmpty();
    if (2u * lane < nb) // nb is even
    {
        uint4 *q = reinterpret_cast<uint4 *>(&s.bins[(size_t)r.first * SAH_SLOT_BINS + 2u * lane]);
        const uint4 a0 = q[0], a1 = q[1], b0 = q[2], b1 = q[3];
        a = SahAcc{{a0.x, a0.y, a0.z}, {a0.w, a1.x, a1.y}, a1.z};
        b = SahAcc{{b0.x, b0.y, b0.z}, {b0.w, b1.x, b1.y}, b1.z};
        const uint4 c0 = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u), c1 = make_uint4(0u, 0u, 0u, 0u); // sahBinClear
        q[0] = c0; q[1] = c1; q[2] = c0; q[3] = c1;
    }
    const SahAcc u = sahAccUnion(a, b);
    SahAcc pre = u, suf = u; // inclusive scans over the lanes
    for (uint32_t d = 1; d < 32u; d <<= 1)
    {
        const SahAcc p = sahAccShift<true>(pre, d), q = sahAccShift<false>(suf, d);
        if (lane >= d) pre = sahAccUnion(pre, p);
        if (lane + d < 32u) suf = sahAccUnion(suf, q);
    }
    SahAcc pex = sahAccShift<true>(pre, 1), sex = sahAccShift<false>(suf, 1);
    if (lane == 0u) pex = sahAccEmpty();
    if (lane == 31u) sex = sahAccEmpty();
    float cost = 3.4e38f;
    uint32_t bin = 0u, left = 0u;
    if (lane >= 1u && 2u * lane < nb) // border 2L: bins [0, 2L) | [2L, nb)
    {
        const float c = sahBorderCost(pex, suf, ax.byIndex);
        if (c < cost) { cost = c; bin = 2u * lane; left = pex.cnt; }
    }
    if (2u * lane + 1u < nb) // border 2L + 1
    {
        const SahAcc l = sahAccUnion(pex, a);
        const float c = sahBorderCost(l, sahAccUnion(b, sex), ax.byIndex);
        if (c < cost) { cost = c; bin = 2u * lane + 1u; left = l.cnt; }
    }
    for (uint32_t d = 16u; d; d >>= 1)
    {
        const float oc = __shfl_xor_sync(0xFFFFFFFFu, cost, d);
        const uint32_t ob = __shfl_xor_sync(0xFFFFFFFFu, bin, d), ol = __shfl_xor_sync(0xFFFFFFFFu, left, d);
        if (oc < cost || (oc == cost && ob < bin)) { cost = oc; bin = ob; left = ol; }
    }
    if (lane == 0u)
    {
        Box nodeBox;
        for (int k = 0; k < 3; ++k) { nodeBox.lo[k] = unordF(suf.lo[k]); nodeBox.hi[k] = unordF(suf.hi[k]); }
        uint32_t childKeys[2], made;
        const uint32_t at = sahSplitEmit(s, key, r, ax, bin, left, 0u, nodeBox, childKeys, made);
        for (uint32_t k = 0; k < made; ++k) queueOut[at + k] = childKeys[k];
    }
}

__device__ __forceinline__ void sahCentroidAdd(uint32_t *cent, const uint32_t lo[4], const uint32_t hi[4])
{
    for (uint32_t a = 0; a < 3; ++a) { atomicMin(&cent[a], lo[a]); atomicMax(&cent[3u + a], hi[a]); }
    atomicMin(&cent[6], lo[3]);
    atomicMax(&cent[7], hi[3]);
}

// sahPartitionPass + the centroid pass of the NEXT level: the slot's primitive moves into its child's sub-range and adds
// its centroid (and index) to that child's bounds; a warp that lies inside one range does both once per side.
__device__ __forceinline__ void sahPartitionTile(const Sah &s, uint32_t base)
{
    const uint32_t slot = base + threadIdx.x, lane = threadIdx.x & 31u;
    const bool valid = slot < s.t.count;
    const uint32_t key = valid ? s.keyIn[slot] : SAH_FINAL;
    const uint32_t key0 = __shfl_sync(0xFFFFFFFFu, key, 0);
    const bool warpUniform = __all_sync(0xFFFFFFFFu, valid && key == key0 && key != SAH_FINAL);
    if (!valid) return;
    const uint32_t prim = s.idxIn[slot];
    if (key == SAH_FINAL) { s.idxOut[slot] = prim; s.keyOut[slot] = SAH_FINAL; return; }
    SahSplit *spp = &s.split[key];
    const SahSplit sp = *spp;
    uint32_t side, pos = slot;
    const bool binned = sahSlotSide(s, slot, prim, sp, side);
    const uint32_t childFirst = side ? sp.first + sp.leftCount : sp.first;
    const uint32_t childCount = side ? sp.count - sp.leftCount : sp.leftCount;
    uint32_t lo[4], hi[4];
    if (childCount >= 2u)
    {
        const Box b = s.t.primBox[prim];
        for (uint32_t a = 0; a < 3; ++a) lo[a] = hi[a] = ordF(sahCentroid(b, a));
        lo[3] = hi[3] = prim;
    }
    if (warpUniform)
    {
        const uint32_t right = __ballot_sync(0xFFFFFFFFu, side != 0u), left = ~right, mine = side ? right : left;
        if (binned) // uniform across the warp: axisBin belongs to the range
        {
            uint32_t baseL = 0, baseR = 0;
            if (lane == 0)
            {
                if (left) baseL = atomicAdd(&spp->cursorL, (uint32_t)__popc(left));
                if (right) baseR = atomicAdd(&spp->cursorR, (uint32_t)__popc(right));
            }
            baseL = __shfl_sync(0xFFFFFFFFu, baseL, 0);
            baseR = __shfl_sync(0xFFFFFFFFu, baseR, 0);
            const uint32_t below = (1u << lane) - 1u;
            pos = side ? sp.first + sp.leftCount + baseR + __popc(right & below) : sp.first + baseL + __popc(left & below);
        }
        if (childCount >= 2u) // uniform across each side
        {
            for (uint32_t a = 0; a < 4; ++a) { lo[a] = __reduce_min_sync(mine, lo[a]); hi[a] = __reduce_max_sync(mine, hi[a]); }
            if (lane == (uint32_t)__ffs((int)mine) - 1u) sahCentroidAdd(&s.cent[(size_t)(childFirst >> 1) * SAH_CENT], lo, hi);
        }
    }
    else
    {
        if (binned) pos = side ? sp.first + sp.leftCount + atomicAdd(&spp->cursorR, 1u) : sp.first + atomicAdd(&spp->cursorL, 1u);
        if (childCount >= 2u) sahCentroidAdd(&s.cent[(size_t)(childFirst >> 1) * SAH_CENT], lo, hi);
    }
    sahPlace(s, pos, prim, sp, side);
}

__global__ void __launch_bounds__(SAH_THREADS) k_sah_all_levels(Sah s, SahGrid g)
{
    __shared__ SahBin sb[SAH_RANGE_BINS];
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    const uint32_t n = s.t.count, stride = gridDim.x * blockDim.x, tid = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t tileStride = gridDim.x * SAH_THREADS, warps = stride >> 5, warp = tid >> 5;
    // start state: one range holding every primitive in index order
    for (uint32_t i = tid; i < n; i += stride)
    {
        g.idx[0][i] = i;
        g.key[0][i] = 0u;
        for (uint32_t b = 0; b < SAH_SLOT_BINS; ++b) sahBinClear(&s.bins[(size_t)i * SAH_SLOT_BINS + b]);
    }
    for (uint32_t i = tid; i < g.keys; i += stride) sahCentClear(&s.cent[i * SAH_CENT]);
    for (uint32_t i = tid; i <= SAH_MAX_LEVELS; i += stride) s.active[i] = i == 0 ? 1u : 0u;
    if (tid == 0) { g.range[0][0] = SahRange{0u, n, -1, 0u}; g.queue[0][0] = 0u; *s.rootGap = 0u; }
    __threadfence();
    grid.sync();
    s.idxIn = g.idx[0]; s.keyIn = g.key[0];
    for (uint32_t base = blockIdx.x * SAH_THREADS; base < n; base += tileStride) sahCentroidTile(s, base);
    __threadfence();
    grid.sync();
    uint32_t level = 0;
    for (; level < SAH_MAX_LEVELS; ++level)
    {
        const uint32_t ranges = __ldcg(&s.active[level]);
        if (ranges == 0u) break; // grid-uniform
        const uint32_t a = level & 1u, b = a ^ 1u;
        s.level = level;
        s.idxIn = g.idx[a]; s.idxOut = g.idx[b];
        s.keyIn = g.key[a]; s.keyOut = g.key[b];
        s.rangeIn = g.range[a]; s.rangeOut = g.range[b];
        for (uint32_t base = blockIdx.x * SAH_THREADS; base < n; base += tileStride) sahBinTile(s, base, sb);
        __threadfence();
        grid.sync();
        for (uint32_t q = warp; q < ranges; q += warps) sahSplitWarp(s, __ldcg(&g.queue[a][q]), g.queue[b]);
        __threadfence();
        grid.sync();
        for (uint32_t base = blockIdx.x * SAH_THREADS; base < n; base += tileStride) sahPartitionTile(s, base);
        __threadfence();
        grid.sync();
    }
    const bool finished = level < SAH_MAX_LEVELS || __ldcg(&s.active[SAH_MAX_LEVELS]) == 0u;
    if (level & 1u) // the collapse reads the final order from idx[0]
        for (uint32_t i = tid; i < n; i += stride) g.idx[0][i] = g.idx[1][i];
    if (tid == 0) *g.levelsOut = finished ? level : SAH_MAX_LEVELS + 1u;
}

// ---- collapse to 8-wide: collapseItem (core_bvh.h) by EIGHT lanes, one per child -------------------------------------
// The serial form walks 8-entry local arrays a few thousand instructions long per wide node; with a level-synchronous
// collapse that latency, times the depth of the tree, is the whole TLAS rebuild. Here lane j of an 8-lane group holds
// child j (reference, box, area, primitive count): the open-largest loop, the greedy slot assignment and the plane
// quantisation become 8-lane arg-max / ballot steps. Same choices in the same order as the serial form (ties: lowest
// child, then lowest slot), so the node is the same; only the allocator order differs, as it does between any two runs.
__device__ __forceinline__ void collapseItemGroup(const Collapse &c, uint32_t itemIndex, uint32_t gm)
{
    const uint32_t lane = threadIdx.x & 31u, j = lane & 7u, shift = lane & 24u; // gm = 0xFF << shift
    const uint2 raw = __ldcg(reinterpret_cast<const uint2 *>(c.queueIn + itemIndex)); // written by other SMs one level earlier
    const CollapseItem item = {(int32_t)raw.x, raw.y};
    const Lbvh &t = c.t;
    int32_t ref = 0;
    Box box = Box{{0.f, 0.f, 0.f}, {0.f, 0.f, 0.f}};
    float area = 0.f;
    uint32_t cnt = 0u, n;
    int32_t kidL = 0, kidR = 0; // the child's own children, fetched with it: opening it later costs no dependent load
    auto fetch = [&](int32_t r) {
        ref = r;
        if (r >= 0) { kidL = t.left[r]; kidR = t.right[r]; }
        box = lbvhRefBox(t, r);
        area = boxHalfArea(box);
        cnt = lbvhRefCount(t, r);
    };
    if (item.ref >= 0 && lbvhRefCount(t, item.ref) > c.maxLeaf)
    {
        if (j == 0u) fetch(t.left[item.ref]);
        if (j == 1u) fetch(t.right[item.ref]);
        n = 2u;
    }
    else
    {
        if (j == 0u) fetch(item.ref); // a tree that is a single leaf (count <= maxLeaf at the root)
        n = 1u;
    }
    // phase 1: open the largest-area child that is too big for a leaf slot; phase 2: use spare slots to split
    // multi-primitive leaves (tighter boxes, same memory)
    for (int phase = 0; phase < 2; ++phase)
    {
        while (n < 8u)
        {
            const bool eligible = j < n && ref >= 0 && (phase == 0 ? cnt > c.maxLeaf : cnt <= c.maxLeaf);
            float a = eligible ? area : -1.f;
            uint32_t who = j;
            for (uint32_t d = 4u; d; d >>= 1)
            {
                const float oa = __shfl_xor_sync(gm, a, d, 8);
                const uint32_t ow = __shfl_xor_sync(gm, who, d, 8);
                if (oa > a || (oa == a && ow < who)) { a = oa; who = ow; }
            }
            if (!(a > -1.f)) break; // group-uniform
            const int32_t l2 = __shfl_sync(gm, kidL, who, 8), r2 = __shfl_sync(gm, kidR, who, 8);
            if (j == who) fetch(l2);
            if (j == n) fetch(r2);
            n++;
        }
    }
    const bool live = j < n;
    // node bounds
    float blo[3], bhi[3];
    for (int k = 0; k < 3; ++k)
    {
        blo[k] = live ? box.lo[k] : 3.4e38f;
        bhi[k] = live ? box.hi[k] : -3.4e38f;
        for (uint32_t d = 4u; d; d >>= 1)
        {
            blo[k] = fminf(blo[k], __shfl_xor_sync(gm, blo[k], d, 8));
            bhi[k] = fmaxf(bhi[k], __shfl_xor_sync(gm, bhi[k], d, 8));
        }
    }
    // slot assignment (greedy on the Ylitie cost: child offset projected on the slot's octant diagonal)
    const V3 nc = (V3{blo[0], blo[1], blo[2]} + V3{bhi[0], bhi[1], bhi[2]}) * 0.5f;
    const V3 off = (boxLo(box) + boxHi(box)) * 0.5f - nc;
    uint32_t freeSlots = 0xFFu, slot = 8u; // slot of this lane's child
    for (uint32_t k = 0; k < n; ++k)
    {
        float bc = -3.4e38f;
        uint32_t bs = 8u, who = j;
        if (live && slot == 8u)
            for (uint32_t sl = 0; sl < 8u; ++sl)
            {
                if (!((freeSlots >> sl) & 1u)) continue;
                const float cost = ((sl & 1u) ? off.x : -off.x) + ((sl & 2u) ? off.y : -off.y) + ((sl & 4u) ? off.z : -off.z);
                if (cost > bc) { bc = cost; bs = sl; }
            }
        for (uint32_t d = 4u; d; d >>= 1)
        {
            const float oc = __shfl_xor_sync(gm, bc, d, 8);
            const uint32_t os = __shfl_xor_sync(gm, bs, d, 8), ow = __shfl_xor_sync(gm, who, d, 8);
            // a lane without a candidate carries bs == 8 and loses every comparison
            if (os != 8u && (bs == 8u || oc > bc || (oc == bc && ow < who))) { bc = oc; bs = os; who = ow; }
        }
        if (j == who) slot = bs;
        freeSlots &= ~(1u << bs);
    }
    // from here lane j is SLOT j: pull the child assigned to it
    uint32_t src = 8u;
    for (uint32_t i = 0; i < 8u; ++i)
        if (__shfl_sync(gm, slot, i, 8) == j) src = i;
    const bool used = src < 8u;
    const uint32_t from = used ? src : 0u;
    const int32_t sref = __shfl_sync(gm, ref, from, 8);
    const uint32_t pulled = __shfl_sync(gm, cnt, from, 8);
    const uint32_t scnt = used ? pulled : 0u;
    Box sbox;
    for (int k = 0; k < 3; ++k) { sbox.lo[k] = __shfl_sync(gm, box.lo[k], from, 8); sbox.hi[k] = __shfl_sync(gm, box.hi[k], from, 8); }
    // allocate internal children and leaf primitives (children in slot order)
    const bool internal = used && scnt > c.maxLeaf;
    const uint32_t imask = (__ballot_sync(gm, internal) >> shift) & 0xFFu;
    const uint32_t leafCnt = used && !internal ? scnt : 0u;
    uint32_t primSeen = leafCnt; // inclusive prefix over the slots
    for (uint32_t d = 1u; d < 8u; d <<= 1)
    {
        const uint32_t o = __shfl_up_sync(gm, primSeen, d, 8);
        if (j >= d) primSeen += o;
    }
    const uint32_t nPrims = __shfl_sync(gm, primSeen, 7, 8), nInternal = (uint32_t)__popc(imask);
    primSeen -= leafCnt;
    const uint32_t internalSeen = (uint32_t)__popc(imask & ((1u << j) - 1u));
    uint32_t childBase = 0u, primBase = 0u, queueBase = 0u;
    if (j == 0u)
    {
        if (nInternal) { childBase = atomicAdd(c.nodeCount, nInternal); queueBase = atomicAdd(c.queueOutCount, nInternal); }
        if (nPrims) primBase = atomicAdd(c.primCount, nPrims);
    }
    childBase = __shfl_sync(gm, childBase, 0, 8);
    primBase = __shfl_sync(gm, primBase, 0, 8);
    queueBase = __shfl_sync(gm, queueBase, 0, 8);
    uint32_t meta = 0u;
    if (internal)
    {
        meta = (1u << 5) | (24u + j);
        c.queueOut[queueBase + internalSeen] = CollapseItem{sref, childBase + internalSeen};
    }
    else if (used)
    {
        const uint32_t first = lbvhRefFirst(t, sref);
        for (uint32_t k = 0; k < scnt; ++k) c.primOut[primBase + primSeen + k] = t.order[first + k];
        meta = ((((1u << scnt) - 1u) << 5) | primSeen) & 0xFFu;
    }
    // quantisation grid: every slot quantises its own planes, the exponent grows until all eight fit
    uint32_t e[3], ql[3], qh[3];
    for (int a = 0; a < 3; ++a)
    {
        e[a] = gridExponent(bhi[a] - blo[a]);
        for (;;)
        {
            const float scale = u2f(e[a] << 23);
            bool ok = true;
            ql[a] = 255u; qh[a] = 0u;
            if (used) ok = quantisePlanes(blo[a], scale, sbox.lo[a], sbox.hi[a], ql[a], qh[a]);
            if (__all_sync(gm, ok) || e[a] >= 254u) break;
            e[a]++;
        }
    }
    // the 80-byte record: lanes 0 and 4 of the group assemble the two words of each byte row
    const uint32_t quad = 0xFu << (lane & 28u), bs = 8u * (j & 3u);
    uint32_t rows[7];
    rows[0] = __reduce_or_sync(quad, meta << bs);
    for (int a = 0; a < 3; ++a)
    {
        rows[1 + a] = __reduce_or_sync(quad, (ql[a] & 0xFFu) << bs);
        rows[4 + a] = __reduce_or_sync(quad, (qh[a] & 0xFFu) << bs);
    }
    uint32_t *out = reinterpret_cast<uint32_t *>(&c.nodes[item.wide]);
    if ((j & 3u) == 0u)
    {
        const uint32_t w = j >> 2; // 0: slots 0-3, 1: slots 4-7
        for (int r = 0; r < 7; ++r) out[6 + 2 * r + w] = rows[r];
    }
    if (j == 0u)
    {
        out[0] = f2u(blo[0]); out[1] = f2u(blo[1]); out[2] = f2u(blo[2]);
        out[3] = e[0] | (e[1] << 8) | (e[2] << 16) | (imask << 24);
        out[4] = childBase;
        out[5] = primBase;
    }
}

// All levels in one block (TLAS: a few thousand instances, launch latency matters more than width).
// queues = two arrays of capacity `queueCapacity` used alternately; queueCounts[l] = items of level l.
__global__ void __launch_bounds__(1024) k_collapse_all_levels(Collapse c, CollapseItem *queueA, CollapseItem *queueB,
                                                             uint32_t *queueCounts, uint32_t maxLevels)
{
    for (uint32_t level = 0; level + 1 < maxLevels; ++level)
    {
        const uint32_t count = __ldcg(&queueCounts[level]);
        if (count == 0) break;
        c.queueIn = (level & 1u) ? queueB : queueA;
        c.queueOut = (level & 1u) ? queueA : queueB;
        c.queueOutCount = &queueCounts[level + 1];
        for (uint32_t i = threadIdx.x >> 3; i < count; i += blockDim.x >> 3) collapseItemGroup(c, i, 0xFFu << (threadIdx.x & 24u));
        __threadfence_block();
        __syncthreads();
    }
}

// All levels in one cooperative launch (TLAS with thousands of instances): the blocks of the grid share every
// level's queue and meet at a grid barrier between levels, so the host enqueues ONE kernel whatever the depth.
__global__ void __launch_bounds__(128) k_collapse_all_levels_grid(Collapse c, CollapseItem *queueA, CollapseItem *queueB,
                                                                  uint32_t *queueCounts, uint32_t maxLevels)
{
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    const uint32_t stride = gridDim.x * blockDim.x, tid = blockIdx.x * blockDim.x + threadIdx.x;
    for (uint32_t level = 0; level + 1 < maxLevels; ++level)
    {
        const uint32_t count = __ldcg(&queueCounts[level]);
        if (count == 0) break; // grid-uniform: every thread reads the same finished counter
        c.queueIn = (level & 1u) ? queueB : queueA;
        c.queueOut = (level & 1u) ? queueA : queueB;
        c.queueOutCount = &queueCounts[level + 1];
        for (uint32_t i = tid >> 3; i < count; i += stride >> 3) collapseItemGroup(c, i, 0xFFu << (threadIdx.x & 24u));
        __threadfence();
        grid.sync();
    }
}

__global__ void k_collapse_seed(CollapseItem *queue, uint32_t *queueCounts, uint32_t levels, uint32_t *nodeCount,
                                uint32_t *primCount, uint32_t primTotal)
{
    if (threadIdx.x == 0)
    {
        queue[0] = CollapseItem{primTotal > 1 ? 0 : ~0, 0u};
        queueCounts[0] = 1;
        *nodeCount = 1;
        *primCount = 0;
    }
    for (uint32_t l = 1 + threadIdx.x; l < levels; l += blockDim.x) queueCounts[l] = 0;
}

__global__ void k_emit_tris(const uint32_t *__restrict__ primOut, uint32_t n, const TriRec *__restrict__ in, TriRec *__restrict__ out,
                            const AlphaUv *__restrict__ uvIn, AlphaUv *__restrict__ uvOut)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t src = primOut[i];
    const uint4 *s = reinterpret_cast<const uint4 *>(in + src);
    uint4 *d = reinterpret_cast<uint4 *>(out + i);
    d[0] = s[0]; d[1] = s[1]; d[2] = s[2];
    const uint4 *us = reinterpret_cast<const uint4 *>(uvIn + src);
    uint4 *ud = reinterpret_cast<uint4 *>(uvOut + i);
    ud[0] = us[0]; ud[1] = us[1];
}

// AlphaMat table: one record per draw instance (any-hit stage)
__global__ void k_alpha_mats(SceneView sv, uint32_t n, AlphaMat *__restrict__ out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = makeAlphaMat(sv, i);
}

// ---- TLAS ------------------------------------------------------------------------------------------------
__global__ void k_tlas_instances(const TlasInput *__restrict__ in, uint32_t n, InstRec *__restrict__ recs, Box *__restrict__ boxes,
                                 BuildBounds *bounds)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
    {
        atomicBounds(bounds, V3{0.f, 0.f, 0.f}, V3{0.f, 0.f, 0.f}, false);
        return;
    }
    InstRec r;
    V3 lo, hi;
    tlasInstanceRecord(in[i], r, lo, hi);
    recs[i] = r;
    boxes[i] = makeBox(lo, hi);
    atomicBounds(bounds, lo, hi, true);
}

// The same from the caller's own 64-byte records (every instance active): the BLAS pointers come from a device table indexed
// by blasReference - 1, so the host neither rewrites nor compacts the array -- it uploads it as it is.
struct BlasTableEntry
{
    const WideNode *nodes;
    const TriRec *tris;
    const AlphaUv *uvs;
};
__global__ void k_tlas_instances_direct(const prt_tlas_instance *__restrict__ in, const BlasTableEntry *__restrict__ table, uint32_t n,
                                        InstRec *__restrict__ recs, Box *__restrict__ boxes, BuildBounds *bounds)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
    {
        atomicBounds(bounds, V3{0.f, 0.f, 0.f}, V3{0.f, 0.f, 0.f}, false);
        return;
    }
    const uint4 *q = reinterpret_cast<const uint4 *>(in + i);
    const uint4 a = q[0], b = q[1], c = q[2], d = q[3];
    TlasInput ti;
    const uint32_t m[12] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, c.x, c.y, c.z, c.w};
    for (int k = 0; k < 12; ++k) ti.transform[k] = __uint_as_float(m[k]);
    ti.customIndexAndMask = d.x;
    ti.index = i;
    const BlasTableEntry e = table[(((uint64_t)d.w << 32) | d.z) - 1ull];
    ti.nodes = e.nodes; ti.tris = e.tris; ti.uvs = e.uvs;
    InstRec r;
    V3 lo, hi;
    tlasInstanceRecord(ti, r, lo, hi);
    recs[i] = r;
    boxes[i] = makeBox(lo, hi);
    atomicBounds(bounds, lo, hi, true);
}

__global__ void k_emit_instances(const uint32_t *__restrict__ primOut, uint32_t n, const InstRec *__restrict__ in, InstRec *__restrict__ out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 *s = reinterpret_cast<const uint4 *>(in + primOut[i]);
    uint4 *d = reinterpret_cast<uint4 *>(out + i);
    static_assert(sizeof(InstRec) == 6 * sizeof(uint4), "InstRec copy");
    d[0] = s[0]; d[1] = s[1]; d[2] = s[2]; d[3] = s[3]; d[4] = s[4]; d[5] = s[5];
}

} // namespace prt
