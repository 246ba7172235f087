// tests/hostsim/hostsim.cpp -- TEST INFRASTRUCTURE: kernel-logic simulator.
//
// Compiles the per-thread bodies of the CUDA kernels (prosper_b200/csrc/core_*.h, the very same
// source the sm_100a kernels are built from) with g++ and drives them serially, so the CPU-only CI
// (`pytest -m "not gpu"`) exercises the GPU builder, the compressed-node traversal and the shading
// stages against the oracle without a GPU. It is NOT a CPU fallback: nothing in prosper_b200/
// loads it and libprt_b200.so does not contain it. What it cannot cover (radix sort, scans, warp
// compaction, texture units) is covered by the `-m gpu` tests.
#include "../../prosper_b200/csrc/core_bc7.h"
#include "../../prosper_b200/csrc/core_bc7enc.h"
#include "../../prosper_b200/csrc/core_build.h"
#include "../../prosper_b200/csrc/core_shade.h"
#include "../../oracle/oracle.h" // orc_scene: the scene description struct shared with the oracle binding

#include <algorithm>
#include <atomic>
#include <mutex>
#include <thread>
#include <vector>

#include <sys/mman.h>

using namespace prt;

namespace {

// Node and triangle arrays come from ONE address range, because a ray names the arrays of its level by their distance from
// Accel::arena in 128-byte units (32 bits = 512 GB; the GPU's allocations lie that close together, a process heap does
// not). A 64 GB reservation without commit, bump-allocated; a freed array gives its pages back, its addresses are not reused.
struct ArenaRange
{
    char *base = nullptr;
    size_t used = 0;
    std::mutex lock;
    static constexpr size_t BYTES = 64ull << 30;
    static ArenaRange &get()
    {
        static ArenaRange a;
        return a;
    }
    void *take(size_t bytes)
    {
        std::lock_guard<std::mutex> g(lock);
        if (!base)
        {
            void *p = mmap(nullptr, BYTES, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
            if (p == MAP_FAILED) throw std::bad_alloc();
            base = (char *)p;
        }
        const size_t at = (used + 4095) & ~(size_t)4095;
        if (at + bytes > BYTES) throw std::bad_alloc();
        used = at + bytes;
        return base + at;
    }
};
template <class T> struct ArenaAlloc
{
    using value_type = T;
    ArenaAlloc() = default;
    template <class U> ArenaAlloc(const ArenaAlloc<U> &) {}
    T *allocate(size_t n) { return (T *)ArenaRange::get().take(n * sizeof(T)); }
    void deallocate(T *p, size_t n)
    {
        const uintptr_t lo = ((uintptr_t)p + 4095) & ~(uintptr_t)4095, hi = ((uintptr_t)p + n * sizeof(T)) & ~(uintptr_t)4095;
        if (hi > lo) madvise((void *)lo, hi - lo, MADV_DONTNEED);
    }
    template <class U> bool operator==(const ArenaAlloc<U> &) const { return true; }
    template <class U> bool operator!=(const ArenaAlloc<U> &) const { return false; }
};
using NodeVec = std::vector<WideNode, ArenaAlloc<WideNode>>;
using TriVec = std::vector<TriRec, ArenaAlloc<TriRec>>;

struct HostBvh
{
    NodeVec nodes;
    std::vector<uint32_t> primOut;
    uint32_t levels = 0;
};

HostBvh buildHierarchyHost(std::vector<Box> &boxes, V3 lo, V3 hi, bool tlas, bool ploc = false, bool sah = false)
{
    const uint32_t n = (uint32_t)boxes.size();
    HostBvh out;
    std::vector<uint64_t> keys(n);
    std::vector<uint32_t> order(n);
    const uint64_t keyMask = tlas ? 0xFFFFFFFF00000000ull : ~0ull;
    for (uint32_t i = 0; i < n; ++i)
    {
        keys[i] = padBoxAndKey(boxes[i], lo, hi, tlas ? 0 : 1, keyMask);
        order[i] = i;
    }
    // stand-in for the radix sort (stable, like the LSD passes)
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return keys[a] < keys[b]; });
    std::vector<uint64_t> sorted(n);
    for (uint32_t i = 0; i < n; ++i) sorted[i] = keys[order[i]];
    std::vector<int32_t> left(n), right(n), parentI(n), parentL(n);
    std::vector<uint32_t> first(n), last(n), visit(n, 0);
    std::vector<Box> nodeBox(n);
    Lbvh t{};
    t.count = n; t.keys = sorted.data(); t.order = order.data(); t.primBox = boxes.data();
    t.left = left.data(); t.right = right.data(); t.first = first.data(); t.last = last.data();
    t.parentOfInternal = parentI.data(); t.parentOfLeaf = parentL.data(); t.nodeBox = nodeBox.data(); t.visit = visit.data();
    std::vector<uint32_t> orderOut(n), newPos(n);
    if (n > 1 && !tlas && sah)
    {
        // binned SAH, the k_sah_* sequence of buildHierarchy (prt_api.cu): four per-slot passes per level
        const uint32_t keys = n / 2u + 1u;
        std::vector<uint32_t> idx[2] = {std::vector<uint32_t>(n), std::vector<uint32_t>(n)};
        std::vector<uint32_t> key[2] = {std::vector<uint32_t>(n, 0u), std::vector<uint32_t>(n)};
        std::vector<SahRange> range[2] = {std::vector<SahRange>(keys), std::vector<SahRange>(keys)};
        std::vector<SahSplit> split(keys);
        std::vector<SahBin> bins((size_t)SAH_SLOT_BINS * n);
        std::vector<uint32_t> cent((size_t)SAH_CENT * keys), active(SAH_MAX_LEVELS + 2u, 0u);
        uint32_t rootGap = 0;
        for (uint32_t i = 0; i < n; ++i) idx[0][i] = i;
        for (uint32_t k = 0; k < keys; ++k) sahCentClear(&cent[k * SAH_CENT]);
        for (SahBin &b : bins) sahBinClear(&b);
        range[0][0] = SahRange{0u, n, -1, 0u};
        active[0] = 1;
        Sah s{};
        s.t = t; s.split = split.data(); s.bins = bins.data(); s.cent = cent.data(); s.active = active.data(); s.rootGap = &rootGap;
        uint32_t level = 0;
        for (; level < SAH_MAX_LEVELS && active[level] > 0; ++level)
        {
            const int a = (int)(level & 1u), b = a ^ 1;
            s.level = level;
            s.idxIn = idx[a].data(); s.idxOut = idx[b].data(); s.keyIn = key[a].data(); s.keyOut = key[b].data();
            s.rangeIn = range[a].data(); s.rangeOut = range[b].data();
            for (uint32_t i = 0; i < n; ++i) sahCentroidPass(s, i);
            for (uint32_t i = 0; i < n; ++i) sahBinPass(s, i);
            for (uint32_t i = 0; i < n; ++i) sahSplitPass(s, i);
            for (uint32_t i = 0; i < n; ++i) sahPartitionPass(s, i);
        }
        orderOut = idx[level & 1u];
        t.order = orderOut.data();
    }
    else if (n > 1 && !tlas && ploc)
    {
        // PLOC, the k_ploc_* sequence of buildHierarchy (prt_api.cu) with the compaction done serially
        std::vector<int32_t> ref[2] = {std::vector<int32_t>(n), std::vector<int32_t>(n)}, mergedRef(n);
        std::vector<Box> box[2] = {std::vector<Box>(n), std::vector<Box>(n)}, mergedBox(n);
        std::vector<uint32_t> nearest(n);
        uint32_t nodeAlloc = 0, m = n;
        for (uint32_t i = 0; i < n; ++i) { ref[0][i] = ~(int32_t)i; box[0][i] = boxes[order[i]]; parentL[i] = -1; }
        Ploc p{};
        p.t = t; p.nearest = nearest.data(); p.mergedRef = mergedRef.data(); p.mergedBox = mergedBox.data();
        p.nodeAlloc = &nodeAlloc; p.radius = 16;
        for (uint32_t it = 0; m > 1; ++it)
        {
            const int a = (int)(it & 1u), b = a ^ 1;
            p.refIn = ref[a].data(); p.boxIn = box[a].data(); p.refOut = ref[b].data(); p.boxOut = box[b].data();
            for (uint32_t i = 0; i < m; ++i) plocNearest(p, i, m);
            for (uint32_t i = 0; i < m; ++i) plocMerge(p, i, m);
            uint32_t next = 0;
            for (uint32_t i = 0; i < m; ++i)
                if (mergedRef[i] != PLOC_DROPPED) { ref[b][next] = mergedRef[i]; box[b][next] = mergedBox[i]; next++; }
            if (next >= m) break; // no progress: a builder bug, caught by the prim-count check below
            m = next;
        }
        for (uint32_t i = 0; i < n; ++i) plocFinishLeaf(t, i, newPos.data(), orderOut.data());
        for (uint32_t i = 0; i + 1 < n; ++i) { plocFinishNode(t, i, newPos.data()); plocRelabelNode(t, i, newPos.data()); }
        t.order = orderOut.data();
    }
    else if (n > 1)
    {
        for (uint32_t i = 0; i + 1 < n; ++i) lbvhBuildNode(t, (int)i);
        for (uint32_t i = 0; i < n; ++i) lbvhRefitFromLeaf(t, i);
    }
    out.nodes.resize(std::max(n, 1u));
    out.primOut.resize(n);
    uint32_t nodeCount = 1, primCount = 0;
    std::vector<CollapseItem> qa(n + 1), qb(n + 1);
    qa[0] = CollapseItem{n > 1 ? 0 : ~0, 0u};
    uint32_t countIn = 1;
    Collapse c{};
    c.t = t; c.maxLeaf = tlas ? 1u : 3u; c.nodes = out.nodes.data(); c.primOut = out.primOut.data();
    c.nodeCount = &nodeCount; c.primCount = &primCount;
    for (uint32_t level = 0; countIn > 0; ++level)
    {
        uint32_t countOut = 0;
        c.queueIn = (level & 1) ? qb.data() : qa.data();
        c.queueOut = (level & 1) ? qa.data() : qb.data();
        c.queueOutCount = &countOut;
        for (uint32_t i = 0; i < countIn; ++i) collapseItem(c, i);
        countIn = countOut;
        out.levels = level + 1;
    }
    out.nodes.resize(nodeCount);
    if (primCount != n) out.primOut.clear(); // signals a builder bug to the caller
    return out;
}

struct HostBlas
{
    NodeVec nodes;
    TriVec tris;
    std::vector<AlphaUv> uvs; // any-hit records, parallel to tris
};

} // namespace

struct hs_ctx
{
    orc_scene s;
    std::vector<TexSlot> slots;
    std::vector<std::vector<uint8_t>> alphaPlanes; // per texture (k_alpha_plane)
    std::vector<AlphaMat> alphaMats;               // per draw instance (k_alpha_mats)
    std::vector<HostBlas> blas;
    NodeVec tlasNodes;
    std::vector<InstRec> tlasInst;
    SceneView sv{};
    Accel as{};
    uint32_t blasLevels = 0, tlasLevels = 0;
    int buildError = 0;
    std::atomic<uint64_t> nodes{0}, tris{0}, instances{0}, anyHits{0}, overflow{0};
    uint64_t uniqueTris = 0; // triangles gathered (a BLAS holds more records: non-opaque triangles are referenced several times)
};

static void flush(hs_ctx *c, const TraceCounters &tc)
{
    c->nodes += tc.nodes; c->tris += tc.tris; c->instances += tc.instances; c->anyHits += tc.anyHits; c->overflow += tc.overflow;
}

template <typename F> static void parallelFor(uint32_t count, int threads, F &&fn)
{
    int n = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
    n = std::max(1, std::min(n, (int)count));
    std::atomic<uint32_t> next{0};
    auto worker = [&]() {
        for (;;)
        {
            const uint32_t i = next.fetch_add(1);
            if (i >= count) break;
            fn(i);
        }
    };
    if (n == 1) { worker(); return; }
    std::vector<std::thread> pool;
    for (int i = 0; i < n; ++i) pool.emplace_back(worker);
    for (auto &t : pool) t.join();
}

// the any-hit records of a BLAS from its triangle records (what k_blas_gather + k_emit_tris leave behind)
static void buildUvsHost(hs_ctx *c, uint32_t model)
{
    const orc_scene &s = c->s;
    const prt_sub_model *subs = (const prt_sub_model *)s.subModels;
    HostBlas &b = c->blas[model];
    b.uvs.assign(b.tris.size(), AlphaUv{{0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, {0u, 0u}});
    for (size_t i = 0; i < b.tris.size(); ++i)
    {
        const TriRec &t = b.tris[i];
        if (t.geometry >> 31) continue;
        gatherTriangleUvs(c->sv, subs[s.modelOffsets[model] + (t.geometry & 0x7FFFFFFFu)].meshIndex, t.primitive, b.uvs[i]);
    }
}

static void buildTlasHost(hs_ctx *c)
{
    const orc_scene &s = c->s;
    c->tlasNodes.clear();
    c->tlasInst.clear();
    c->as = Accel{};
    // TLAS (prt_build_tlas)
    const prt_tlas_instance *inst = (const prt_tlas_instance *)s.tlasInstances;
    std::vector<InstRec> recs;
    std::vector<Box> boxes;
    V3 lo = {3.4e38f, 3.4e38f, 3.4e38f}, hi = {-3.4e38f, -3.4e38f, -3.4e38f};
    for (uint32_t i = 0; i < s.instanceCount; ++i)
    {
        if (inst[i].blasReference == 0) continue;
        const uint64_t model = inst[i].blasReference - 1;
        if (model >= s.modelCount || c->blas[model].tris.empty()) continue;
        TlasInput ti;
        memcpy(ti.transform, inst[i].transform, sizeof(ti.transform));
        ti.customIndexAndMask = inst[i].customIndexAndMask;
        ti.index = i;
        ti.nodes = c->blas[model].nodes.data();
        ti.tris = c->blas[model].tris.data();
        ti.uvs = c->blas[model].uvs.data();
        InstRec r;
        V3 l, h;
        tlasInstanceRecord(ti, r, l, h);
        recs.push_back(r);
        boxes.push_back(makeBox(l, h));
        lo = vmin(lo, l);
        hi = vmax(hi, h);
    }
    if (!recs.empty())
    {
        HostBvh h = buildHierarchyHost(boxes, lo, hi, true);
        if (h.primOut.size() != recs.size()) c->buildError = 2;
        else
        {
            c->tlasLevels = h.levels;
            c->tlasNodes = std::move(h.nodes);
            c->tlasInst.resize(recs.size());
            for (size_t i = 0; i < recs.size(); ++i) c->tlasInst[i] = recs[h.primOut[i]];
            c->as.tlasNodes = c->tlasNodes.data();
            c->as.instances = c->tlasInst.data();
            c->as.instanceCount = (uint32_t)c->tlasInst.size();
            c->as.planeBias = PLANE_BIAS;
            c->as.arena = (uint64_t)(uintptr_t)ArenaRange::get().base;
        }
    }
}

extern "C" {

// flags: bit 0 = software filter for material taps (PRT_CFG_EXACT_FILTER), bit 1 = PLOC builder (PRT_CFG_PLOC),
// bit 2 = Karras BLAS builder (PRT_CFG_LBVH_BLAS) instead of the default binned SAH
hs_ctx *hs_create(const orc_scene *scene, int flags)
{
    const int exactMaterialFilter = flags & 1;
    const bool ploc = (flags & 2) != 0;
    const bool sah = (flags & (2 | 4)) == 0;
    hs_ctx *c = new hs_ctx;
    c->s = *scene;
    const orc_scene &s = c->s;
    // texture slots point at the cached payloads: the software filter decodes texels on the fly
    const uint32_t ns = std::max(s.samplerCount, 1u);
    c->slots.assign((size_t)s.textureCount * ns, TexSlot{});
    c->alphaPlanes.resize(s.textureCount);
    for (uint32_t t = 0; t < s.textureCount; ++t)
    {
        const orc_texture &tx = s.textures[t];
        if (!tx.data || !tx.width) continue;
        // level-0 alpha plane (k_alpha_plane): one byte per texel as the payload decodes, dropped when all 255
        std::vector<uint8_t> &plane = c->alphaPlanes[t];
        plane.resize((size_t)tx.width * tx.height);
        uint32_t lowest = 255;
        {
            TexSlot view{};
            view.payload = (const uint8_t *)tx.data; view.format = tx.format; view.width = tx.width; view.height = tx.height;
            for (uint32_t y = 0; y < tx.height; ++y)
                for (uint32_t x = 0; x < tx.width; ++x)
                {
                    const uint32_t a = (uint32_t)(fetchTexel(view, (int)x, (int)y).w * 255.f + 0.5f);
                    plane[(size_t)y * tx.width + x] = (uint8_t)a;
                    lowest = std::min(lowest, a);
                }
        }
        for (uint32_t sm = 0; sm < ns; ++sm)
        {
            TexSlot &slot = c->slots[(size_t)t * ns + sm];
            slot.payload = (const uint8_t *)tx.data;
            slot.format = tx.format;
            slot.width = tx.width; slot.height = tx.height;
            const prt_sampler_desc sd = s.samplerCount ? ((const prt_sampler_desc *)s.samplers)[sm] : prt_sampler_desc{1, 1, 0, 0};
            slot.magFilter = sd.magFilter; slot.addressU = sd.addressU; slot.addressV = sd.addressV;
            slot.valid = 1;
            slot.alphaOpaque = lowest == 255 ? 1u : 0u;
            slot.alpha = lowest == 255 ? nullptr : plane.data();
        }
    }
    SceneView &sv = c->sv;
    sv.geometryBuffers = (const uint32_t *const *)s.geometryBuffers;
    sv.metadatas = (const prt_geometry_metadata *)s.metadatas;
    sv.materials = (const prt_material_data *)s.materials;
    sv.drawInstances = (const prt_draw_instance *)s.drawInstances;
    sv.transforms = (const prt_model_instance_transforms *)s.transforms;
    sv.texSlots = c->slots.data();
    sv.textureCount = s.textureCount;
    sv.samplerCount = ns;
    sv.directional = (const prt_directional_light *)s.directionalLight;
    sv.pointLights = (const prt_point_light *)s.pointLights;
    sv.spotLights = (const prt_spot_light *)s.spotLights;
    sv.pointLightCount = s.pointLightCount;
    sv.spotLightCount = s.spotLightCount;
    sv.camera = (const prt_camera_uniforms *)s.camera;
    sv.sky = (const uint16_t *)s.skyRgba16f;
    sv.skyFaceSize = s.skyRgba16f ? s.skyFaceSize : 0;
    sv.exactMaterialFilter = exactMaterialFilter ? 1u : 0u;
    sv.alphaMats = nullptr;
    c->alphaMats.resize(s.drawInstanceCount);
    for (uint32_t i = 0; i < s.drawInstanceCount; ++i) c->alphaMats[i] = makeAlphaMat(sv, i);
    sv.alphaMats = c->alphaMats.empty() ? nullptr : c->alphaMats.data();

    // BLAS per model (the k_blas_gather -> buildHierarchy -> k_emit_tris sequence of prt_build_blas)
    const prt_mesh_info *infos = (const prt_mesh_info *)s.meshInfos;
    const prt_sub_model *subs = (const prt_sub_model *)s.subModels;
    c->blas.resize(s.modelCount);
    for (uint32_t m = 0; m < s.modelCount; ++m)
    {
        std::vector<TriRec> tris;
        std::vector<Box> boxes;
        V3 lo = {3.4e38f, 3.4e38f, 3.4e38f}, hi = {-3.4e38f, -3.4e38f, -3.4e38f};
        for (uint32_t g = s.modelOffsets[m]; g < s.modelOffsets[m + 1]; ++g)
        {
            const prt_mesh_info &info = infos[subs[g].meshIndex];
            const uint32_t opaque = sv.materials[info.materialIndex].alphaMode == PRT_ALPHA_OPAQUE ? 1u : 0u;
            for (uint32_t p = 0; p < info.indexCount / 3; ++p)
            {
                TriRec t;
                V3 l, h, corner[3];
                gatherTriangle(sv, subs[g].meshIndex, p, g - s.modelOffsets[m], opaque, t, l, h, corner);
                // k_blas_gather: a non-opaque triangle enters the builder as several references (core_build.h)
                const uint32_t refs = blasRefsPerTriangle(opaque);
                for (uint32_t k = 0; k < refs; ++k)
                {
                    V3 rl = l, rh = h;
                    if (refs > 1u) triangleRefBox(corner[0], corner[1], corner[2], k, rl, rh);
                    tris.push_back(t);
                    boxes.push_back(makeBox(rl, rh));
                }
                c->uniqueTris++;
                lo = vmin(lo, l);
                hi = vmax(hi, h);
            }
        }
        if (tris.empty()) continue;
        HostBvh h = buildHierarchyHost(boxes, lo, hi, false, ploc, sah);
        if (h.primOut.size() != tris.size()) { c->buildError = 1; continue; }
        c->blasLevels = std::max(c->blasLevels, h.levels);
        c->blas[m].nodes = std::move(h.nodes);
        c->blas[m].tris.resize(tris.size());
        for (size_t i = 0; i < tris.size(); ++i) c->blas[m].tris[i] = tris[h.primOut[i]];
        buildUvsHost(c, m);
    }
    buildTlasHost(c);
    return c;
}

void hs_destroy(hs_ctx *c) { delete c; }

// Swap in a BLAS built elsewhere (read back from the GPU with prt_debug_read_blas) and rebuild the TLAS:
// lets the tests traverse the GPU builder's output with the CPU copy of the traversal code.
void hs_replace_blas(hs_ctx *c, uint32_t model, const void *nodes, uint32_t nodeCount, const void *tris, uint32_t triCount)
{
    HostBlas &b = c->blas[model];
    b.nodes.assign((const WideNode *)nodes, (const WideNode *)nodes + nodeCount);
    b.tris.assign((const TriRec *)tris, (const TriRec *)tris + triCount);
    buildUvsHost(c, model);
    buildTlasHost(c);
}

// out[0..7]: buildError, blas wide nodes, blas triangles, tlas wide nodes, tlas instances, blas levels, tlas levels,
// blas triangle records (>= triangles: builder references of non-opaque triangles)
void hs_stats(hs_ctx *c, uint64_t *out)
{
    uint64_t bn = 0, bt = 0;
    for (const HostBlas &b : c->blas) { bn += b.nodes.size(); bt += b.tris.size(); }
    out[0] = (uint64_t)c->buildError; out[1] = bn; out[2] = c->uniqueTris; out[3] = c->tlasNodes.size(); out[4] = c->tlasInst.size();
    out[5] = c->blasLevels; out[6] = c->tlasLevels; out[7] = bt;
}

// out[0..4]: nodes visited, triangles tested, instances entered, any-hit calls, stack overflows
void hs_trace_counters(hs_ctx *c, uint64_t *out)
{
    out[0] = c->nodes; out[1] = c->tris; out[2] = c->instances; out[3] = c->anyHits; out[4] = c->overflow;
}

void hs_trace_primary(hs_ctx *c, const void *pcIn, uint32_t width, uint32_t height, prt_hit *hits, int threads)
{
    const prt_reference_pc pc = *(const prt_reference_pc *)pcIn;
    parallelFor(height, threads, [&](uint32_t y) {
        TraceCounters tc = {0, 0, 0, 0, 0};
        unsigned long long ovf = 0;
        for (uint32_t x = 0; x < width; ++x)
        {
            Rng rng;
            const Ray ray = generatePrimaryRay(c->sv, pc, pc.frameIndex, x, y, width, height, rng);
            HitRec h;
            traceRay<false, true>(c->as, c->sv, ray.o, ray.d, ray.tMin, ray.tMax, pcg(rng.state.x ^ rng.state.z), h, &tc, &ovf);
            hits[(size_t)y * width + x] = prt_hit{h.drawInstanceIndex, h.primitiveID, {h.u, h.v}};
        }
        tc.overflow += (uint32_t)ovf;
        flush(c, tc);
    });
}

void hs_trace_rays(hs_ctx *c, const float *originTmin, const float *dirTmax, const uint32_t *seeds, uint32_t count, int shadow,
                   prt_hit *hits, int threads)
{
    const uint32_t chunk = 1024;
    parallelFor((count + chunk - 1) / chunk, threads, [&](uint32_t r) {
        TraceCounters tc = {0, 0, 0, 0, 0};
        unsigned long long ovf = 0;
        for (uint32_t i = r * chunk; i < std::min(count, (r + 1) * chunk); ++i)
        {
            const V3 o = {originTmin[i * 4], originTmin[i * 4 + 1], originTmin[i * 4 + 2]};
            const V3 d = {dirTmax[i * 4], dirTmax[i * 4 + 1], dirTmax[i * 4 + 2]};
            HitRec h;
            if (shadow) traceRay<true, true>(c->as, c->sv, o, d, originTmin[i * 4 + 3], dirTmax[i * 4 + 3], seeds ? seeds[i] : 0u, h, &tc, &ovf);
            else traceRay<false, true>(c->as, c->sv, o, d, originTmin[i * 4 + 3], dirTmax[i * 4 + 3], seeds ? seeds[i] : 0u, h, &tc, &ovf);
            hits[i] = prt_hit{h.drawInstanceIndex, h.primitiveID, {h.u, h.v}};
        }
        tc.overflow += (uint32_t)ovf;
        flush(c, tc);
    });
}

// The wavefront stages executed path by path: raygen -> [trace_closest -> shade -> trace_shadow] -> accumulate.
// counters (optional, 4 x u64): closest rays, shadow rays, shaded hits, pixel samples.
void hs_render(hs_ctx *c, const void *pcIn, uint32_t width, uint32_t height, uint32_t sampleCount, float *rgba, int threads,
               uint64_t *counters)
{
    const prt_reference_pc pc0 = *(const prt_reference_pc *)pcIn;
    std::atomic<uint64_t> closest{0}, shadowRays{0}, shaded{0};
    parallelFor(height, threads, [&](uint32_t y) {
        TraceCounters tc = {0, 0, 0, 0, 0};
        unsigned long long ovf = 0;
        uint64_t nc = 0, nsr = 0, nsh = 0;
        for (uint32_t x = 0; x < width; ++x)
        {
            float *out = rgba + ((size_t)y * width + x) * 4;
            for (uint32_t k = 0; k < sampleCount; ++k)
            {
                prt_reference_pc pc = pc0;
                pc.frameIndex = (pc0.frameIndex + k) % PRT_FRAME_PERIOD;
                if (k > 0) pc.flags &= ~1u;
                Rng rng;
                Ray ray = generatePrimaryRay(c->sv, pc, pc.frameIndex, x, y, width, height, rng);
                uint32_t seed = pcg(rng.state.x ^ rng.state.z);
                V3 throughput = {1.f, 1.f, 1.f}, color = {0.f, 0.f, 0.f};
                const uint32_t bounces = std::min<uint32_t>(pc.maxBounces, MAX_BOUNCES);
                for (uint32_t bounce = 0; bounce < bounces; ++bounce)
                {
                    HitRec h;
                    traceRay<false, true>(c->as, c->sv, ray.o, ray.d, ray.tMin, ray.tMax, seed, h, &tc, &ovf);
                    nc++;
                    ShadeResult r;
                    SurfaceGeom geom;
                    MaterialTaps taps;
                    if (h.drawInstanceIndex != MISS_INDEX)
                    {
                        geom = surfaceGeometry(c->sv, h.drawInstanceIndex, h.primitiveID, h.u, h.v);
                        taps = softwareMaterialTaps(c->sv, geom.materialIndex, geom.v.TexCoord0);
                    }
                    shadePath(c->sv, pc, bounce, ray.d, h.drawInstanceIndex, h.primitiveID, geom, taps, rng, throughput, color, r);
                    if (h.drawInstanceIndex != MISS_INDEX) nsh++;
                    if (r.shadowRay)
                    {
                        HitRec sh;
                        nsr++;
                        if (!traceRay<true, true>(c->as, c->sv, r.shadowO, r.shadowD, 0.1f, r.shadowTMax, r.shadowSeed, sh, &tc, &ovf))
                            color = color + r.shadowContribution;
                    }
                    if (!r.alive) break;
                    ray = r.next;
                    seed = pcg(rng.state.x ^ rng.state.z);
                }
                if (flagSkipHistory(pc) || !flagAccumulate(pc))
                {
                    out[0] = color.x; out[1] = color.y; out[2] = color.z; out[3] = 1.f;
                }
                else
                {
                    const float n = out[3] + 1.f;
                    out[0] = out[0] + (color.x - out[0]) / n;
                    out[1] = out[1] + (color.y - out[1]) / n;
                    out[2] = out[2] + (color.z - out[2]) / n;
                    out[3] = n;
                }
            }
        }
        tc.overflow += (uint32_t)ovf;
        flush(c, tc);
        closest += nc; shadowRays += nsr; shaded += nsh;
    });
    if (counters)
    {
        counters[0] += closest; counters[1] += shadowRays; counters[2] += shaded;
        counters[3] += (uint64_t)width * height * sampleCount;
    }
}

// unit entry points
int hs_bc7_decode_block(const uint8_t block[16], uint8_t out[64])
{
    uint64_t lo, hi;
    memcpy(&lo, block, 8);
    memcpy(&hi, block + 8, 8);
    return bc7DecodeBlock(lo, hi, (uint32_t *)out) ? 0 : -1;
}

// The product's block encoder (core_bc7enc.h) on the CPU: rgba8 [height][width][4] -> width/4 * height/4 blocks of 16 bytes.
// modes (optional): one byte per block, the mode written.
void hs_bc7_encode(const uint8_t *rgba, uint32_t width, uint32_t height, uint8_t *blocks, uint8_t *modes)
{
    const uint32_t bx = width / 4, by = height / 4;
    for (uint32_t b = 0; b < bx * by; ++b)
    {
        uint32_t texels[16];
        for (uint32_t i = 0; i < 16; ++i)
            memcpy(&texels[i], rgba + ((size_t)((b / bx) * 4 + i / 4) * width + (b % bx) * 4 + i % 4) * 4, 4);
        uint64_t lo, hi;
        const uint32_t mode = bc7EncodeBlock(texels, lo, hi);
        memcpy(blocks + (size_t)b * 16, &lo, 8);
        memcpy(blocks + (size_t)b * 16 + 8, &hi, 8);
        if (modes) modes[b] = (uint8_t)mode;
    }
}

// Builder check on bare boxes (n x {lo xyz, hi xyz}): builds the BLAS hierarchy with builder 0 = binned SAH, 1 = Karras,
// 2 = PLOC and walks the wide tree. out: [0] wide nodes, [1] collapse levels, [2] depth of the wide tree, [3] errors
// (a primitive missing or repeated in the node-order permutation, or a primitive box sticking out of a slot box on its
// path). Returns [3].
struct CheckWalk
{
    const HostBvh &h;
    const std::vector<Box> &boxes;
    std::vector<uint32_t> seen;
    uint32_t errors = 0, depth = 0;
    void walk(uint32_t nodeIndex, uint32_t level, const Box *bounds, uint32_t nBounds)
    {
        depth = std::max(depth, level);
        const WideNode &n = h.nodes[nodeIndex];
        const float sx = u2f((uint32_t)n.ex << 23), sy = u2f((uint32_t)n.ey << 23), sz = u2f((uint32_t)n.ez << 23);
        std::vector<Box> path(bounds, bounds + nBounds);
        path.push_back(Box{});
        uint32_t child = n.childBase;
        for (int s = 0; s < 8; ++s)
        {
            if (n.meta[s] == 0) continue;
            path.back() = Box{{n.px + n.qlo[0][s] * sx, n.py + n.qlo[1][s] * sy, n.pz + n.qlo[2][s] * sz},
                              {n.px + n.qhi[0][s] * sx, n.py + n.qhi[1][s] * sy, n.pz + n.qhi[2][s] * sz}};
            if (n.imask & (1u << s)) { walk(child++, level + 1, path.data(), (uint32_t)path.size()); continue; }
            const uint32_t count = (uint32_t)popc((uint32_t)n.meta[s] >> 5), offset = n.meta[s] & 31u;
            for (uint32_t j = 0; j < count; ++j)
            {
                const uint32_t slot = n.primBase + offset + j;
                if (slot >= h.primOut.size()) { errors++; continue; }
                const uint32_t prim = h.primOut[slot];
                if (prim >= boxes.size()) { errors++; continue; }
                seen[prim]++;
                for (const Box &b : path)
                    for (int a = 0; a < 3; ++a)
                        if (boxes[prim].lo[a] < b.lo[a] || boxes[prim].hi[a] > b.hi[a]) { errors++; break; }
            }
        }
    }
};

uint32_t hs_build_check(const float *boxesIn, uint32_t n, int builder, uint32_t out[4])
{
    std::vector<Box> boxes(n);
    V3 lo = {3.4e38f, 3.4e38f, 3.4e38f}, hi = {-3.4e38f, -3.4e38f, -3.4e38f};
    for (uint32_t i = 0; i < n; ++i)
    {
        for (int a = 0; a < 3; ++a) { boxes[i].lo[a] = boxesIn[i * 6 + a]; boxes[i].hi[a] = boxesIn[i * 6 + 3 + a]; }
        lo = vmin(lo, boxLo(boxes[i]));
        hi = vmax(hi, boxHi(boxes[i]));
    }
    const HostBvh h = buildHierarchyHost(boxes, lo, hi, false, builder == 2, builder == 0); // pads the boxes in place
    CheckWalk w{h, boxes, std::vector<uint32_t>(n, 0u)};
    if (h.primOut.size() != n) w.errors++;
    else if (n > 0) w.walk(0, 1, nullptr, 0);
    for (uint32_t i = 0; i < n; ++i) if (w.seen[i] != 1u) w.errors++;
    out[0] = (uint32_t)h.nodes.size(); out[1] = h.levels; out[2] = w.depth; out[3] = w.errors;
    return w.errors;
}

} // extern "C"
