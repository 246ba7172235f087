// kernels_build.cuh -- launches of the acceleration-structure builder (bodies in core_bvh.h).
// BLAS: World::Impl::buildNextBlas (World.cpp:585-747). TLAS: updateTlasInstances + buildCurrentTlas
// (World.cpp:878-928, 749-802), rebuilt from scratch every frame like the reference does.
#pragma once
#include <cooperative_groups.h>
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
    V3 lo, hi;
    gatherTriangle(sv, meshIndex, p, geometryIndex, opaque, t, lo, hi);
    tris[outOffset + p] = t;
    AlphaUv uv = {{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, {0u, 0u}};
    if (!opaque) gatherTriangleUvs(sv, meshIndex, p, uv);
    uvs[outOffset + p] = uv;
    boxes[outOffset + p] = makeBox(lo, hi);
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

// One collapse level: items [0, *countIn) of queueIn. Launched with enough threads for the worst
// case; the live count is read from device memory so the host never waits between levels.
__global__ void k_collapse_level(Collapse c, const uint32_t *__restrict__ countIn)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= *countIn) return;
    collapseItem(c, i);
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
        for (uint32_t i = threadIdx.x; i < count; i += blockDim.x) collapseItem(c, i);
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
        for (uint32_t i = tid; i < count; i += stride) collapseItem(c, i);
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
