// tests/cpp/test_abi.cpp -- drives libprt_b200.so through include/prt_rt_reference.hpp WITHOUT Python: the C++ host side
// north_star asks for ("the host stays C++"). A quad above a larger floor quad, one material, sky off, the default
// directional light; checks primary hit IDs against what the geometry dictates, the accumulation counter, the
// frame-index / skipHistory logic of RtReference::record, and the error path. Exit code 0 = pass.
// Built by __graft_entry__.build() (g++), run on the GPU box by tests/test_gpu_cpp_host.py.
#include <cstdio>
#include <cstdlib>
#include "../../include/prt_rt_reference.hpp"

using namespace prt_host;

static uint16_t half(float f) // exact for the values used here (small integers and halves)
{
    uint32_t x;
    std::memcpy(&x, &f, 4);
    const uint32_t s = (x >> 16) & 0x8000u, e = (x >> 23) & 0xFF, m = x & 0x7FFFFFu;
    if (e == 0) return (uint16_t)s;
    const int he = (int)e - 127 + 15;
    if (he <= 0) return (uint16_t)s;
    return (uint16_t)(s | ((uint32_t)he << 10) | (m >> 13));
}
static uint32_t snorm3x10(float x, float y, float z, float w)
{
    auto q = [](float v, float scale) { return (int)std::lround(std::fmax(-1.f, std::fmin(1.f, v)) * scale); };
    return ((uint32_t)q(x, 511.f) & 0x3FFu) | (((uint32_t)q(y, 511.f) & 0x3FFu) << 10) | (((uint32_t)q(z, 511.f) & 0x3FFu) << 20) |
           (((uint32_t)q(w, 1.f) & 0x3u) << 30);
}
#define EXPECT(cond)                                                                                                           \
    do                                                                                                                         \
    {                                                                                                                          \
        if (!(cond))                                                                                                           \
        {                                                                                                                      \
            std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond);                                             \
            return 1;                                                                                                          \
        }                                                                                                                      \
    } while (0)

// one quad (two triangles, 4 vertices) in the y = `y` plane, half-extent `h`, packed the way packMeshData does
static void pushQuad(std::vector<uint32_t> &buf, float y, float h, prt_geometry_metadata &md, prt_mesh_info &info, uint32_t material)
{
    const float p[4][3] = {{-h, y, -h}, {h, y, -h}, {h, y, h}, {-h, y, h}};
    const uint16_t idx[6] = {0, 2, 1, 0, 3, 2};
    md = prt_geometry_metadata{};
    md.bufferIndex = 0;
    md.indicesOffset = (uint32_t)buf.size() * 2; // in u16 units
    for (int i = 0; i < 6; i += 2) buf.push_back((uint32_t)idx[i] | ((uint32_t)idx[i + 1] << 16));
    md.positionsOffset = (uint32_t)buf.size();
    for (auto &v : p)
    {
        buf.push_back((uint32_t)half(v[0]) | ((uint32_t)half(v[1]) << 16));
        buf.push_back((uint32_t)half(v[2]) | ((uint32_t)half(1.f) << 16));
    }
    md.normalsOffset = (uint32_t)buf.size();
    for (int i = 0; i < 4; ++i) buf.push_back(snorm3x10(0.f, 1.f, 0.f, 0.f));
    md.tangentsOffset = (uint32_t)buf.size();
    for (int i = 0; i < 4; ++i) buf.push_back(snorm3x10(1.f, 0.f, 0.f, 1.f));
    md.texCoord0sOffset = (uint32_t)buf.size();
    const float uv[4][2] = {{0, 0}, {1, 0}, {1, 1}, {0, 1}};
    for (auto &t : uv) buf.push_back((uint32_t)half(t[0]) | ((uint32_t)half(t[1]) << 16));
    md.meshletsOffset = md.meshletBoundsOffset = md.meshletVerticesOffset = md.meshletTrianglesByteOffset = 0xFFFFFFFFu;
    md.usesShortIndices = 1;
    info = prt_mesh_info{4, 6, 0, material};
}

int main()
{
    const uint32_t W = 64, H = 48;
    try
    {
        prt_config cfg{0, W, H, 1, 0, PRT_CFG_COUNTERS};
        World world(cfg);
        std::vector<uint32_t> buf;
        std::vector<prt_geometry_metadata> mds(2);
        std::vector<prt_mesh_info> infos(2);
        pushQuad(buf, 0.f, 4.f, mds[0], infos[0], 1);  // floor, model 0
        pushQuad(buf, 1.f, 0.5f, mds[1], infos[1], 2); // small quad above it, model 1
        world.setGeometryBuffer(0, buf.data(), buf.size() * 4);
        world.setMeshes(mds, infos);
        const prt_material_data dflt{{1, 1, 1, 1}, 1.f, 1.f, 0.5f, PRT_ALPHA_OPAQUE, 0, 0, 0, 0};
        prt_material_data grey = dflt, red = dflt;
        grey.baseColorFactor[0] = grey.baseColorFactor[1] = grey.baseColorFactor[2] = 0.5f;
        grey.metallicFactor = 0.f;
        red.baseColorFactor[1] = red.baseColorFactor[2] = 0.1f;
        red.metallicFactor = 0.f;
        world.setMaterials({dflt, grey, red});
        world.setSamplers({prt_sampler_desc{PRT_FILTER_LINEAR, PRT_FILTER_LINEAR, PRT_ADDRESS_REPEAT, PRT_ADDRESS_REPEAT}});
        world.setModels({0, 1, 2}, {prt_sub_model{0, 1}, prt_sub_model{1, 2}});
        // two instances: draw instance 0 = floor, 1 = quad (identity transforms)
        prt_model_instance_transforms ident{};
        for (int i = 0; i < 3; ++i) ident.modelToWorld[i][i] = ident.normalToWorld[i][i] = 1.f;
        world.transforms = {ident, ident};
        world.drawInstances = {prt_draw_instance{0, 0, 1}, prt_draw_instance{1, 1, 2}};
        Camera cam;
        cam.init(W, H);
        CameraTransform t;
        t.eye[0] = 0.f; t.eye[1] = 5.f; t.eye[2] = 0.01f;
        cam.lookAt(t); // looking straight down at the origin
        world.updateBuffers(cam);
        world.tlasInstances.resize(2);
        for (uint32_t i = 0; i < 2; ++i)
        {
            std::memcpy(world.tlasInstances[i].transform, ident.modelToWorld, sizeof(ident.modelToWorld));
            world.tlasInstances[i].customIndexAndMask = i | (0xFFu << 24);
            world.tlasInstances[i].blasReference = i + 1; // placeholder until the BLAS exists: 0 would mean inactive
        }
        EXPECT(world.buildAccelerationStructures());
        for (uint32_t i = 0; i < 2; ++i) world.tlasInstances[i].blasReference = world.blasReference(i);
        EXPECT(!world.buildAccelerationStructures()); // nothing new to build, TLAS rebuilt with the real references

        // primary hits: the centre pixel sees the small quad (draw instance 1), a corner sees the floor or the sky
        std::vector<prt_hit> hits((size_t)W * H);
        prt_reference_pc pc{};
        pc.flags = PRT_FLAG_ACCUMULATE | PRT_FLAG_CLAMP_INDIRECT | PRT_FLAG_SKIP_HISTORY;
        pc.frameIndex = 1; pc.rouletteStartBounce = 3; pc.maxBounces = 4; pc.focusDistance = 1.f; pc.apertureDiameter = 1e-5f;
        check(prt_trace_primary(world.ctx(), &pc, W, H, hits.data()));
        const prt_hit c = hits[(size_t)(H / 2) * W + W / 2];
        EXPECT(c.drawInstanceIndex == 1 && c.primitiveID < 2);
        EXPECT(c.bary[0] >= 0.f && c.bary[1] >= 0.f && c.bary[0] + c.bary[1] <= 1.f);
        uint32_t quadPx = 0, floorPx = 0, missPx = 0;
        for (const prt_hit &h : hits)
        {
            if (h.drawInstanceIndex == 1) quadPx++;
            else if (h.drawInstanceIndex == 0) floorPx++;
            else if (h.drawInstanceIndex == 0xFFFFFFFFu) missPx++;
            else EXPECT(!"unexpected draw instance");
        }
        EXPECT(quadPx > 20 && floorPx > quadPx && quadPx + floorPx + missPx == W * H);

        // the pass: three records accumulate, a camera move resets the history (RtRef.cpp:282-284)
        RtReference pass;
        pass.init(world);
        EXPECT(pass.recompileShaders());
        pass.drawUi(true, true, 3, 4);
        RtReference::Options opts;
        std::vector<float> img((size_t)W * H * 4);
        for (int k = 0; k < 3; ++k)
        {
            const RtReference::Output out = pass.record(cam, W, H, opts);
            cam.endFrame();
            EXPECT(out.illumination.size() == (size_t)W * H * 4);
        }
        check(prt_read_output_f32(world.ctx(), img.data(), img.size() * 4));
        EXPECT(img[3] == 3.f);
        float sum = 0.f;
        for (size_t i = 0; i < img.size(); i += 4)
        {
            EXPECT(std::isfinite(img[i]) && img[i] >= 0.f);
            sum += img[i] + img[i + 1] + img[i + 2];
        }
        EXPECT(sum > 0.f); // the directional light reaches the quads
        // the red quad is redder than the grey floor
        const size_t centre = ((size_t)(H / 2) * W + W / 2) * 4;
        EXPECT(img[centre] > 2.f * img[centre + 1]);
        t.eye[1] = 5.5f;
        cam.lookAt(t);
        world.updateBuffers(cam);
        (void)pass.record(cam, W, H, opts);
        check(prt_read_output_f32(world.ctx(), img.data(), img.size() * 4));
        EXPECT(img[3] == 1.f);
        pass.releasePreserved();

        prt_counters counters{};
        check(prt_get_counters(world.ctx(), &counters));
        EXPECT(counters.pixelSamples == 4ull * W * H + (uint64_t)W * H /* trace_primary */ || counters.pixelSamples == 4ull * W * H);
        EXPECT(counters.stackOverflows == 0);

        // errors come back as exceptions carrying prt_last_error()
        bool threw = false;
        try
        {
            (void)pass.record(cam, 4096, 4096, opts); // larger than prt_config
        }
        catch (const std::runtime_error &e)
        {
            threw = std::string(e.what()).size() > 0;
        }
        EXPECT(threw);
    }
    catch (const std::exception &e)
    {
        std::fprintf(stderr, "exception: %s\n", e.what());
        return 2;
    }
    std::printf("cpp host ok\n");
    return 0;
}
