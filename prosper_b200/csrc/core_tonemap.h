// Display transform of the reference: res/shader/tone_map.comp + the HSV helpers of
// res/shader/common/math.glsl:18-77, driven by render::ToneMap (src/render/ToneMap.cpp:31-42, 63-134).
// Per-pixel bodies, compiled for the device (kernels_tex.cuh::k_tone_map) and, with g++, for the CPU tests.
#pragma once
#include "hd.h"

namespace prt {

// GLSL mod(x, y) = x - y * floor(x / y)
HD float glslMod(float x, float y) { return x - y * floorf(x / y); }

// math.glsl:18-42 (hue in [0, 6), not degrees)
HD V3 rgbToHsv(V3 rgb)
{
    const float value = fmaxf(fmaxf(rgb.x, rgb.y), rgb.z);
    const float valueMinusChroma = fminf(fminf(rgb.x, rgb.y), rgb.z);
    const float chroma = value - valueMinusChroma;
    float hue;
    if (chroma == 0.f)
        hue = 0.f;
    else if (value == rgb.x)
        hue = glslMod((rgb.y - rgb.z) / chroma, 6.f);
    else if (value == rgb.y)
        hue = (rgb.z - rgb.x) / chroma + 2.f;
    else
        hue = (rgb.x - rgb.y) / chroma + 4.f;
    const float saturation = value == 0.f ? 0.f : chroma / value;
    return V3{hue, saturation, value};
}

// math.glsl:45-77
HD V3 hsvToRgb(V3 hsv)
{
    const float hue = hsv.x, value = hsv.z;
    const float chroma = value * hsv.y;
    const float x = chroma * (1.f - fabsf(glslMod(hue, 2.f) - 1.f));
    V3 rgb;
    if (hue < 1.f) rgb = V3{chroma, x, 0.f};
    else if (hue < 2.f) rgb = V3{x, chroma, 0.f};
    else if (hue < 3.f) rgb = V3{0.f, chroma, x};
    else if (hue < 4.f) rgb = V3{0.f, x, chroma};
    else if (hue < 5.f) rgb = V3{x, 0.f, chroma};
    else rgb = V3{chroma, 0.f, x};
    const float m = value - chroma;
    return V3{rgb.x + m, rgb.y + m, rgb.z + m};
}

// One texel of DXGI_FORMAT_R9G9B9E5_SHAREDEXP (the format of res/texture/tony_mc_mapface.dds): three 9-bit
// mantissas without a hidden bit, one 5-bit exponent with bias 15 -> m * 2^(e - 15 - 9). Exact in fp32.
HD V3 decodeRgb9e5(uint32_t v)
{
    const int e = (int)(v >> 27) - 24;
    const float scale = u2f((uint32_t)(e + 127) << 23); // 2^e, e in [-24, 7]
    return V3{(float)(v & 0x1FFu) * scale, (float)((v >> 9) & 0x1FFu) * scale, (float)((v >> 18) & 0x1FFu) * scale};
}

struct ToneLut
{
    const uint32_t *texels; // dim^3 R9G9B9E5 words, x fastest: kept packed (442 KB for 48^3) so the eight taps of a pixel
    uint32_t dim;           // mostly hit L1; the decode is three conversions and exact
};

// textureLod(sampler3D, uv, 0) with a linear clamp-to-edge sampler (ToneMap.cpp:95-98): unnormalised
// coordinate u * dim - 0.5, the two nearest texels per axis, weights from the fraction.
HD V3 sampleLutTrilinear(const ToneLut &lut, V3 uv)
{
    const int dim = (int)lut.dim;
    const float c[3] = {uv.x * (float)dim - 0.5f, uv.y * (float)dim - 0.5f, uv.z * (float)dim - 0.5f};
    int i0[3], i1[3];
    float w[3];
    for (int a = 0; a < 3; ++a)
    {
        const float f = floorf(c[a]);
        w[a] = c[a] - f;
        // NaN / out-of-range coordinates clamp like the sampler does
        const int i = (f >= -1.f) ? ((f < (float)dim) ? (int)f : dim) : -1;
        i0[a] = i < 0 ? 0 : (i > dim - 1 ? dim - 1 : i);
        i1[a] = i + 1 < 0 ? 0 : (i + 1 > dim - 1 ? dim - 1 : i + 1);
        if (!(w[a] >= 0.f)) w[a] = 0.f;
    }
    V3 acc{0.f, 0.f, 0.f};
    for (int k = 0; k < 8; ++k)
    {
        const int x = (k & 1) ? i1[0] : i0[0], y = (k & 2) ? i1[1] : i0[1], z = (k & 4) ? i1[2] : i0[2];
        const float wt = ((k & 1) ? w[0] : 1.f - w[0]) * ((k & 2) ? w[1] : 1.f - w[1]) * ((k & 4) ? w[2] : 1.f - w[2]);
        const V3 t = decodeRgb9e5(lut.texels[((size_t)z * dim + y) * dim + x]);
        acc.x += wt * t.x;
        acc.y += wt * t.y;
        acc.z += wt * t.z;
    }
    return acc;
}

// tone_map.comp:17-27
HD V3 tonyMcMapface(const ToneLut &lut, V3 stimulus)
{
    const V3 encoded{stimulus.x / (stimulus.x + 1.f), stimulus.y / (stimulus.y + 1.f), stimulus.z / (stimulus.z + 1.f)};
    const float dims = (float)lut.dim; // LUT_DIMS = 48 for the shipped table
    const float s = (dims - 1.f) / dims, o = 0.5f / dims;
    return sampleLutTrilinear(lut, V3{encoded.x * s + o, encoded.y * s + o, encoded.z * s + o});
}

// store to an R8G8B8A8_UNORM image (ToneMap.cpp:77): clamp, scale, round to nearest even; NaN -> 0
HD uint32_t toUnorm8(float x)
{
    if (!(x > 0.f)) return 0u;
    if (x > 1.f) x = 1.f;
    return (uint32_t)rintf(x * 255.f);
}

// tone_map.comp:30-62 for one pixel; `color` is the fp16 illumination texel the pass reads
HD uint32_t toneMapPixel(const ToneLut &lut, V3 color, float exposure, float contrast)
{
    color = V3{color.x * exposure, color.y * exposure, color.z * exposure};
    V3 hsv = rgbToHsv(color);
    if (contrast != 1.f) hsv.z = powf(hsv.z, contrast); // pow(x, 1) == x exactly
    color = tonyMcMapface(lut, hsvToRgb(hsv));
    const float g = 1.f / 2.2f;
    return toUnorm8(powf(color.x, g)) | (toUnorm8(powf(color.y, g)) << 8) | (toUnorm8(powf(color.z, g)) << 16) | 0xFF000000u;
}

} // namespace prt
