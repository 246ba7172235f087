// kernels_sort.cuh -- LSD radix sort of (64-bit Morton key, 32-bit primitive index) pairs, 8 bits per
// pass, stable. Three launches per pass: per-block digit histograms, one exclusive scan over the
// [digit][block] table, and a scatter that ranks keys inside a block with warp match + per-warp
// digit counters in shared memory.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace prt {

constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = 8;                        // keys per thread and block: large inputs (BLAS Morton builders)
constexpr int SORT_ITEMS_SMALL = 2;                  // TLAS-sized inputs: four times the blocks, a quarter of the serial chunks
constexpr uint32_t SORT_SMALL_N = 1u << 15;         // at most 64 blocks: each scatter block scans the digit table itself
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS_SMALL; // sizes the histogram table: the smaller tile gives the larger table

template <int ITEMS>
__global__ void __launch_bounds__(SORT_THREADS) k_sort_hist(const uint64_t *__restrict__ keys, uint32_t n, int shift,
                                                            uint32_t *__restrict__ hist, uint32_t numBlocks)
{
    constexpr int SORT_ITEMS = ITEMS, SORT_TILE = SORT_THREADS * ITEMS;
    __shared__ uint32_t sh[256];
    sh[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * SORT_TILE;
#pragma unroll
    for (int c = 0; c < SORT_ITEMS; ++c)
    {
        const uint32_t i = base + c * SORT_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&sh[(uint32_t)(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[threadIdx.x * numBlocks + blockIdx.x] = sh[threadIdx.x];
}

// Exclusive scan of `total` counters in place, one block of 1024 threads walking the array.
__global__ void __launch_bounds__(1024) k_scan_exclusive(uint32_t *__restrict__ data, uint32_t total)
{
    __shared__ uint32_t warpSums[32];
    __shared__ uint32_t carry;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < total; base += 1024u)
    {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < total ? data[i] : 0u;
        uint32_t x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
        {
            const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
            if (lane >= (uint32_t)o) x += y;
        }
        if (lane == 31u) warpSums[warp] = x;
        __syncthreads();
        if (warp == 0)
        {
            uint32_t s = warpSums[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1)
            {
                const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, s, o);
                if (lane >= (uint32_t)o) s += y;
            }
            warpSums[lane] = s; // inclusive
        }
        __syncthreads();
        const uint32_t warpOffset = warp ? warpSums[warp - 1] : 0u;
        const uint32_t c = carry;
        if (i < total) data[i] = c + warpOffset + x - v;
        __syncthreads();
        if (threadIdx.x == 1023u) carry = c + warpOffset + x;
        __syncthreads();
    }
}

// SCAN: `hist` holds the raw per-block digit counts and every block derives its own bases from them (a few dozen blocks:
// 256 x numBlocks reads per block cost less than the scan launch they replace); otherwise `hist` was scanned by
// k_scan_exclusive.
template <int ITEMS, bool SCAN>
__global__ void __launch_bounds__(SORT_THREADS)
    k_sort_scatter(const uint64_t *__restrict__ keysIn, const uint32_t *__restrict__ valsIn, uint64_t *__restrict__ keysOut,
                   uint32_t *__restrict__ valsOut, uint32_t n, int shift, const uint32_t *__restrict__ hist, uint32_t numBlocks)
{
    constexpr int SORT_ITEMS = ITEMS, SORT_TILE = SORT_THREADS * ITEMS;
    __shared__ uint32_t binBase[256];
    __shared__ uint32_t warpCnt[SORT_THREADS / 32][256];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    if (SCAN)
    {
        // base of digit d in this block = keys of smaller digits anywhere + keys of digit d in earlier blocks
        static_assert(SORT_THREADS == 256, "one thread per digit");
        uint32_t total = 0, before = 0;
        for (uint32_t b = 0; b < numBlocks; ++b)
        {
            const uint32_t v = hist[tid * numBlocks + b];
            total += v;
            if (b < blockIdx.x) before += v;
        }
        uint32_t x = total; // inclusive scan over the 256 digits
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
        {
            const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
            if (lane >= (uint32_t)o) x += y;
        }
        if (lane == 31u) warpCnt[0][warp] = x;
        __syncthreads();
        uint32_t offset = 0;
        for (uint32_t w = 0; w < warp; ++w) offset += warpCnt[0][w];
        __syncthreads(); // warpCnt is cleared below
        binBase[tid] = offset + x - total + before;
    }
    else
        binBase[tid] = hist[tid * numBlocks + blockIdx.x];
    const uint32_t tile = blockIdx.x * SORT_TILE;
    for (int c = 0; c < SORT_ITEMS; ++c)
    {
#pragma unroll
        for (int w = 0; w < SORT_THREADS / 32; ++w) warpCnt[w][tid] = 0;
        __syncthreads();
        const uint32_t i = tile + c * SORT_THREADS + tid;
        const bool valid = i < n;
        const uint64_t key = valid ? keysIn[i] : 0ull;
        const uint32_t d = valid ? ((uint32_t)(key >> shift) & 255u) : (0x100u + lane); // invalid lanes never match
        const uint32_t peers = __match_any_sync(0xFFFFFFFFu, d);
        const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
        if (valid && rank == 0) warpCnt[warp][d] = __popc(peers);
        __syncthreads();
        uint32_t run = binBase[tid];
#pragma unroll
        for (int w = 0; w < SORT_THREADS / 32; ++w)
        {
            const uint32_t cnt = warpCnt[w][tid];
            warpCnt[w][tid] = run;
            run += cnt;
        }
        binBase[tid] = run;
        __syncthreads();
        if (valid)
        {
            const uint32_t pos = warpCnt[warp][d] + rank;
            keysOut[pos] = key;
            valsOut[pos] = valsIn[i];
        }
        __syncthreads();
    }
}

// Sorts n pairs on key bits [firstBit, 64). keys/vals are ping-ponged with the alt buffers; returns
// true when the result ended up in the alt buffers. hist must hold 256 * ceil(n / SORT_TILE) words.
inline bool radixSortPairs(uint64_t *keys, uint32_t *vals, uint64_t *keysAlt, uint32_t *valsAlt, uint32_t n, int firstBit,
                           uint32_t *hist, cudaStream_t stream, uint32_t *launches)
{
    if (n == 0) return false;
    const bool small = n <= SORT_SMALL_N;
    const uint32_t tile = (uint32_t)SORT_THREADS * (small ? SORT_ITEMS_SMALL : SORT_ITEMS);
    const uint32_t numBlocks = (n + tile - 1) / tile;
    bool flipped = false;
    for (int shift = firstBit; shift < 64; shift += 8)
    {
        uint64_t *kin = flipped ? keysAlt : keys, *kout = flipped ? keys : keysAlt;
        uint32_t *vin = flipped ? valsAlt : vals, *vout = flipped ? vals : valsAlt;
        if (small)
        {
            k_sort_hist<SORT_ITEMS_SMALL><<<numBlocks, SORT_THREADS, 0, stream>>>(kin, n, shift, hist, numBlocks);
            k_sort_scatter<SORT_ITEMS_SMALL, true><<<numBlocks, SORT_THREADS, 0, stream>>>(kin, vin, kout, vout, n, shift, hist, numBlocks);
            if (launches) *launches += 2;
        }
        else
        {
            k_sort_hist<SORT_ITEMS><<<numBlocks, SORT_THREADS, 0, stream>>>(kin, n, shift, hist, numBlocks);
            k_scan_exclusive<<<1, 1024, 0, stream>>>(hist, 256u * numBlocks);
            k_sort_scatter<SORT_ITEMS, false><<<numBlocks, SORT_THREADS, 0, stream>>>(kin, vin, kout, vout, n, shift, hist, numBlocks);
            if (launches) *launches += 3;
        }
        flipped = !flipped;
    }
    return flipped;
}

} // namespace prt
