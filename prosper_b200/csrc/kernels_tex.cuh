// kernels_tex.cuh -- texture-side kernels: BC7 -> RGBA8 expansion (PRT_CFG_DECODE_BC7 upload path) and
// a texel probe used by the tests to pin what the texture units return against the software decode.
#pragma once
#include <cuda_runtime.h>
#include "core_bc7.h"
#include "core_bc7enc.h"
#include "core_scene.h"
#include "core_tonemap.h"

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

// ---- BC7 encode (texture-cache build, Texture.cpp:213-296: the step prosper runs through ispc_texcomp) ------------
// One thread per 4x4 block through bc7EncodeBlock (core_bc7enc.h): the reference's `alpha_ultrafast` profile -- modes 4
// (both index layouts), 5 and 6, principal-axis endpoints, least-squares refinement, cheapest candidate. The host build
// of the same header (tests/hostsim) emits the same bits; quality is gated against the decoders, which are pinned on
// hand-assembled blocks and on the texture units (tests/test_bc7.py).
__global__ void __launch_bounds__(64) k_bc7_encode(const uint32_t *__restrict__ rgba, uint32_t width, uint32_t height, uint2 *__restrict__ blocks64)
{
    const uint32_t bx = width >> 2, by = height >> 2;
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= bx * by) return;
    const uint32_t x0 = (b % bx) * 4, y0 = (b / bx) * 4;
    uint32_t texels[16];
    for (int row = 0; row < 4; ++row)
    {
        const uint4 t = __ldg(reinterpret_cast<const uint4 *>(rgba + (size_t)(y0 + row) * width + x0)); // width % 4 == 0: aligned
        texels[4 * row + 0] = t.x; texels[4 * row + 1] = t.y; texels[4 * row + 2] = t.z; texels[4 * row + 3] = t.w;
    }
    uint64_t lo, hi;
    bc7EncodeBlock(texels, lo, hi);
    blocks64[2 * (size_t)b + 0] = make_uint2((uint32_t)lo, (uint32_t)(lo >> 32));
    blocks64[2 * (size_t)b + 1] = make_uint2((uint32_t)hi, (uint32_t)(hi >> 32));
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


// ---- display transform: tone_map.comp over Output::illumination -> RGBA8 unorm ----
// One thread per pixel: 16 B in, 4 B out, eight LUT taps that stay in L1/L2 (the 48^3 table is 442 KB packed).
// The pass reads the fp16 illumination image in the reference (RtRef.cpp:336-377 blits to R16G16B16A16 first), so
// the fp32 running mean goes through the same rounding here instead of through a separate blit launch.
__global__ void k_tone_map(const float4 *__restrict__ accum, ToneLut lut, float exposure, float contrast,
                           uint32_t *__restrict__ out, uint32_t pixels)
{
    const uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= pixels) return;
    const float4 c = accum[pix];
    const V3 color{halfToFloat(floatToHalf(c.x)), halfToFloat(floatToHalf(c.y)), halfToFloat(floatToHalf(c.z))};
    out[pix] = toneMapPixel(lut, color, exposure, contrast);
}

// ---- texture-cache build: the mip chain of generateMipLevels (Texture.cpp:165-211) ----
// What stbir_resize_uint8_srgb / _linear (ext/stb/stb_image_resize2.h, the reference's call at Texture.cpp:193-208) do for
// an exact 2:1 reduction of an RGBA image with default settings: decode (sRGB table for colour, /255 for alpha or for
// linear textures), "fancy" alpha weighting (filter R G B A and R*A G*A B*A, divide the weighted colour by the filtered
// alpha unless it is ~0), the default downsampling filter = Mitchell-Netravali B=C=1/3 stretched by 2, i.e. eight fixed
// taps per axis at +-0.25, 0.75, 1.25, 1.75 (normalised), clamped edges, re-encode to 8 bits. One thread per destination
// texel; every operation is a single rounding in a fixed order (tables, __fmul_rn/__fadd_rn/__fdiv_rn), so the result
// is bit-identical to the numpy restatement in prosper_b200/texture_cache.py, which tests pin against the real stb code.
// tables: decode colour [256], decode linear [256], colour midpoints [255], linear midpoints [255], taps [8].
struct MipTables
{
    const float *colour, *linear, *colourMid, *linearMid, *taps;
};
__device__ __forceinline__ uint32_t mipEncode(float v, const float *__restrict__ midpoints)
{
    uint32_t lo = 0, hi = 255; // number of midpoints <= v
    while (lo < hi)
    {
        const uint32_t mid = (lo + hi) >> 1;
        if (v < midpoints[mid]) hi = mid;
        else lo = mid + 1;
    }
    return lo;
}
__global__ void k_downsample_mitchell(const uint32_t *__restrict__ src, int sw, int sh, uint32_t *__restrict__ dst, int dw,
                                      int dh, MipTables tb)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= dw * dh) return;
    const int dx = idx % dw, dy = idx / dw;
    float acc[7];
#pragma unroll 1
    for (int j = 0; j < 8; ++j)
    {
        const int y = min(max(2 * dy - 3 + j, 0), sh - 1);
        float h[7];
#pragma unroll
        for (int i = 0; i < 8; ++i)
        {
            const int x = min(max(2 * dx - 3 + i, 0), sw - 1);
            const uint32_t t = src[(size_t)y * sw + x];
            float v[7];
            v[0] = tb.colour[t & 0xFFu];
            v[1] = tb.colour[(t >> 8) & 0xFFu];
            v[2] = tb.colour[(t >> 16) & 0xFFu];
            v[3] = tb.linear[t >> 24];
            v[4] = __fmul_rn(v[0], v[3]);
            v[5] = __fmul_rn(v[1], v[3]);
            v[6] = __fmul_rn(v[2], v[3]);
            const float w = tb.taps[i];
#pragma unroll
            for (int c = 0; c < 7; ++c) h[c] = i == 0 ? __fmul_rn(w, v[c]) : __fadd_rn(h[c], __fmul_rn(w, v[c]));
        }
        const float w = tb.taps[j];
#pragma unroll
        for (int c = 0; c < 7; ++c) acc[c] = j == 0 ? __fmul_rn(w, h[c]) : __fadd_rn(acc[c], __fmul_rn(w, h[c]));
    }
    const float alpha = acc[3];
    float rgb[3] = {acc[0], acc[1], acc[2]};
    if (!(alpha < 7.52316385e-37f)) // stbir__small_float = 2^-120
    {
        const float ia = __fdiv_rn(1.0f, alpha);
        rgb[0] = __fmul_rn(acc[4], ia);
        rgb[1] = __fmul_rn(acc[5], ia);
        rgb[2] = __fmul_rn(acc[6], ia);
    }
    dst[idx] = mipEncode(rgb[0], tb.colourMid) | (mipEncode(rgb[1], tb.colourMid) << 8) | (mipEncode(rgb[2], tb.colourMid) << 16) |
               (mipEncode(alpha, tb.linearMid) << 24);
}

} // namespace prt
