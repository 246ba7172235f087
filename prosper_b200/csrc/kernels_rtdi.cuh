// kernels_rtdi.cuh -- the ReSTIR DI trace pass as stages around the shared shadow-ray kernel:
//   [k_rtdi_gbuffer]  primary hits -> G-buffer texels (stands in for the raster G-buffer pass, which is out of scope)
//   [k_rtdi_initial_reservoirs]  restir_di/initial_reservoirs.comp
//   k_rtdi_shade  main.rgen:112-148 up to the traceRayEXT: one shadow ray per lit pixel into the shadow queue
//   k_trace<.., SHADOW>  the same persistent traversal kernel the reference pass uses (same TLAS, any-hit, miss)
//   k_rtdi_accumulate  main.rgen:150-164
#pragma once
#include "core_rtdi.h"
#include "kernels_render.cuh"

namespace prt {

// G-buffer from the primary hits of k_trace (queue 0 holds one ray per pixel through the pixel CENTRE, the raster sample point)
__global__ void __launch_bounds__(RENDER_THREADS)
    k_rtdi_primary_rays(SceneView sv, uint32_t width, uint32_t height, uint32_t frameIndex, PathBuffers pb)
{
    const uint32_t pixels = width * height;
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p == 0)
    {
        pb.counts[0] = pixels;
        for (uint32_t i = 1; i < COUNT_WORDS; ++i) pb.counts[i] = 0;
    }
    if (p >= pixels) return;
    const uint32_t px = p % width, py = p / width;
    const Ray ray = pinholeCameraRay(*sv.camera, V2{((float)px + 0.5f) / (float)width, ((float)py + 0.5f) / (float)height});
    pb.rayO[0][p] = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.tMin);
    pb.rayD[0][p] = make_float4(ray.d.x, ray.d.y, ray.d.z, ray.tMax);
    pb.rayId[0][p] = make_uint2(p, pcg(px ^ frameIndex)); // payload.randomSeed for stochastic (BLEND) any-hit decisions
}

__global__ void __launch_bounds__(RENDER_THREADS)
    k_rtdi_gbuffer(SceneView sv, PathBuffers pb, uint32_t width, uint32_t height, uint32_t *albedoRoughness, uint32_t *normalMetallic, float *depth)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = p < width * height;
    uint4 h = make_uint4(MISS_INDEX, MISS_INDEX, 0u, 0u);
    float4 d = make_float4(0.f, 0.f, 0.f, 0.f);
    if (active)
    {
        h = pb.hits[p];
        d = pb.rayD[0][p];
    }
    const bool shaded = active && h.x != MISS_INDEX;
    SurfaceGeom geom;
    geom.materialIndex = 0;
    MaterialTaps taps;
    taps.base = taps.metallicRoughness = taps.normal = V4{1.f, 1.f, 1.f, 1.f};
    if (shaded) geom = surfaceGeometry(sv, h.x, h.y, __uint_as_float(h.z), __uint_as_float(h.w));
    if (sv.exactMaterialFilter)
    {
        if (shaded) taps = softwareMaterialTaps(sv, geom.materialIndex, geom.v.TexCoord0);
    }
    else
    {
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
    if (!active) return;
    uint32_t ar = 0, nm = 0;
    float z = 0.f; // cleared G-buffer: nothing drawn (reverse-Z far plane)
    if (shaded)
    {
        const VisibleSurface surface = evaluateSurface(sv, V3{d.x, d.y, d.z}, geom, taps);
        if (surface.material.alpha != 0) gbufferTexel(sv, surface, ar, nm, z); // "Early out if alpha test failed / zero alpha", gbuffer.frag:62-63
    }
    albedoRoughness[p] = ar;
    normalMetallic[p] = nm;
    depth[p] = z;
}

__global__ void __launch_bounds__(RENDER_THREADS)
    k_rtdi_initial_reservoirs(SceneView sv, GBufferView gb, uint32_t frameIndex, float2 *reservoirs)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= gb.width * gb.height) return;
    const uint32_t px = p % gb.width, py = p / gb.width;
    Rng rng;
    rng.state = {px, py, frameIndex};
    const VisibleSurface surface = gbufferSurface(sv, gb, px, py);
    const LightReservoir r = initialLightCandidate(sv, surface, rng);
    reservoirs[p] = make_float2(__int_as_float(r.lightIndex), r.unbiasedContributionWeight);
}

// rt/direct_illumination/main.rgen:112-148 up to the shadow ray; debug draw types write their colour directly
__global__ void __launch_bounds__(RENDER_THREADS)
    k_rtdi_shade(SceneView sv, GBufferView gb, RtdiTracePC pc, const float2 *__restrict__ reservoirs, PathBuffers pb)
{
    const uint32_t pixels = gb.width * gb.height;
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    bool wantRay = false; // (the queue counters were zeroed by the caller before this launch)
    V3 l = {0.f, 1.f, 0.f}, contribution = {0.f, 0.f, 0.f}, origin = {0.f, 0.f, 0.f};
    float dist = 0.f;
    uint32_t seed = 0;
    if (p < pixels)
    {
        const uint32_t px = p % gb.width, py = p / gb.width;
        const VisibleSurface surface = gbufferSurface(sv, gb, px, py);
        V3 color = {0.f, 0.f, 0.f};
        if (pc.drawType != PRT_DRAW_DEFAULT)
        {
            // DrawType_Position shows the reconstructed position, everything else the albedo the G-buffer pass stored
            color = pc.drawType == PRT_DRAW_POSITION ? surface.positionWS : surface.material.albedo;
        }
        else
        {
            const float2 packed = reservoirs[p];
            const LightReservoir reservoir = {__float_as_int(packed.x), packed.y};
            wantRay = restirShade(sv, surface, reservoir, l, dist, contribution);
            origin = surface.positionWS;
            seed = pcg(px ^ py); // payload.randomSeed = pcg(pcg_state.x ^ pcg_state.y) with pcg_state = (px, frameIndex), main.rgen:50
        }
        pb.color[p] = make_float4(color.x, color.y, color.z, 0.f);
    }
    const uint32_t ss = appendSlot(&pb.counts[COUNT_SHADOW_BASE], wantRay);
    if (wantRay)
    {
        pb.shadowO[ss] = make_float4(origin.x, origin.y, origin.z, 0.1f); // main.rgen:104 tMin
        pb.shadowD[ss] = make_float4(l.x, l.y, l.z, dist);
        pb.shadowId[ss] = make_uint2(p, seed);
        pb.shadowC[ss] = make_float4(contribution.x, contribution.y, contribution.z, 0.f);
    }
}

// main.rgen:150-164
__global__ void k_rtdi_accumulate(PathBuffers pb, float4 *__restrict__ accum, uint32_t pixels, uint32_t flags, uint32_t debugView)
{
    const uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= pixels) return;
    const float4 c = pb.color[pix];
    float4 h = accum[pix];
    const bool skipHistory = (flags & 1u) != 0, accumulate = ((flags >> 1) & 1u) != 0;
    if (debugView || skipHistory || !accumulate) h = make_float4(c.x, c.y, c.z, 1.f);
    else
    {
        const float n = h.w + 1.f;
        h.x = h.x + (c.x - h.x) / n;
        h.y = h.y + (c.y - h.y) / n;
        h.z = h.z + (c.z - h.z) / n;
        h.w = n;
    }
    accum[pix] = h;
}

} // namespace prt
