// kernels_tex.cuh -- texture-side kernels: BC7 -> RGBA8 expansion (PRT_CFG_DECODE_BC7 upload path) and
// a texel probe used by the tests to pin what the texture units return against the software decode.
#pragma once
#include <cuda_runtime.h>
#include "core_bc7.h"
#include "core_scene.h"

namespace prt {

// One thread per 4x4 block of a BC7 level -> RGBA8 texels in a pitched linear image.
__global__ void k_bc7_decode(const uint2 *__restrict__ blocks64, uint32_t blocksX, uint32_t blocksY, uint32_t width, uint32_t height,
                             uint32_t *__restrict__ rgba)
{
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= blocksX * blocksY) return;
    const uint64_t *src = reinterpret_cast<const uint64_t *>(blocks64) + 2ull * b;
    uint32_t px[16];
    bc7DecodeBlock(src[0], src[1], px);
    const uint32_t bx = b % blocksX, by = b / blocksX;
    for (uint32_t y = 0; y < 4; ++y)
        for (uint32_t x = 0; x < 4; ++x)
        {
            const uint32_t tx = bx * 4 + x, ty = by * 4 + y;
            if (tx < width && ty < height) rgba[(size_t)ty * width + tx] = px[y * 4 + x];
        }
}

// Level-0 alpha plane for scene.rahit's sampleAlpha: one byte per texel, exactly the alpha the payload decodes to.
// One thread per texel (BC7: 16 threads share a block fetch through L1). minAlpha collects the smallest alpha so
// a texture without any transparency needs no plane at all.
__global__ void k_alpha_plane(TexSlot slot, uint8_t *__restrict__ alpha, uint32_t *__restrict__ minAlpha)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t a = 255u;
    if (i < slot.width * slot.height)
    {
        const uint32_t x = i % slot.width, y = i / slot.width;
        if (slot.format == PRT_TEX_BC7_UNORM)
        {
            const uint32_t bw = (slot.width + 3u) >> 2;
            const uint4 b = __ldg(reinterpret_cast<const uint4 *>(slot.payload + ((size_t)(y >> 2) * bw + (x >> 2)) * 16));
            const uint64_t lo = (uint64_t)b.x | ((uint64_t)b.y << 32), hi = (uint64_t)b.z | ((uint64_t)b.w << 32);
            a = bc7DecodeTexel(lo, hi, ((y & 3u) << 2) | (x & 3u)) >> 24;
        }
        else
            a = __ldg(reinterpret_cast<const uint32_t *>(slot.payload) + i) >> 24;
        alpha[i] = (uint8_t)a;
    }
    a = __reduce_min_sync(0xFFFFFFFFu, a);
    if ((threadIdx.x & 31u) == 0 && a < 255u) atomicMin(minAlpha, a);
}

// out[i] = texel xy[i] as the texture units decode it (point view of the CUDA array; uniform handle)
__global__ void k_probe_texels(TexSlot slot, const int2 *__restrict__ xy, uint32_t n, float4 *__restrict__ out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = tex2D<float4>(slot.texels, (float)xy[i].x + 0.5f, (float)xy[i].y + 0.5f);
}

// out[i] = texture() at uv[i]: software filter (exact != 0) or texture units (uniform handle)
__global__ void k_probe_sample(SceneView sv, uint32_t textureSampler, const float2 *__restrict__ uv, uint32_t n, int exact,
                               float4 *__restrict__ out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const TexSlot *s = texSlot(sv, textureSampler);
    if (s == nullptr) { out[i] = make_float4(1.f, 1.f, 1.f, 1.f); return; }
    if (exact)
    {
        const V4 t = sampleTextureExact(*s, V2{uv[i].x, uv[i].y});
        out[i] = make_float4(t.x, t.y, t.z, t.w);
    }
    else
        out[i] = tex2D<float4>(s->linear, uv[i].x, uv[i].y);
}

} // namespace prt
