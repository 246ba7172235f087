// prt_api.cu -- libprt_b200.so: the C ABI of include/prt.h over the sm_100a kernels.
//
// Host side of the drop-in boundary for render::RtReference (src/render/RtReference.{hpp,cpp}) and
// the acceleration-structure half of scene::World (src/scene/World.cpp:538-961). No Vulkan, no CPU
// fallback: every entry point fails with PRT_ERR_NO_DEVICE / PRT_ERR_CUDA when the GPU is missing.
#include "../../include/prt.h"

#include <cuda.h> // types of the virtual-memory-management calls only; the entry points come from cudaGetDriverEntryPoint
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "kernels_build.cuh"
#include "kernels_render.cuh"
#include "kernels_tex.cuh"
#include "kernels_rtdi.cuh"

using namespace prt;

static thread_local std::string g_lastError;

static prt_status fail(prt_status st, const std::string &msg)
{
    g_lastError = msg;
    return st;
}

#define CK(call)                                                                                                        \
    do                                                                                                                  \
    {                                                                                                                   \
        const cudaError_t e__ = (call);                                                                                 \
        if (e__ != cudaSuccess)                                                                                         \
            return fail(PRT_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__) + " (" __FILE__ ":" +         \
                                          std::to_string(__LINE__) + ")");                                              \
    } while (0)
#define CKS(expr)                                                                                                       \
    do                                                                                                                  \
    {                                                                                                                   \
        const prt_status s__ = (expr);                                                                                  \
        if (s__ != PRT_OK) return s__;                                                                                  \
    } while (0)
#define REQUIRE(cond, msg)                                                                                              \
    do                                                                                                                  \
    {                                                                                                                   \
        if (!(cond)) return fail(PRT_ERR_INVALID, std::string(msg) + " [" #cond "]");                                   \
    } while (0)

static inline uint32_t blocksFor(uint64_t n, uint32_t threads) { return (uint32_t)std::max<uint64_t>(1, (n + threads - 1) / threads); }

// Bumped whenever a DBuf moves: a captured TLAS graph holds raw pointers and is re-captured when this changed.
static unsigned long long g_allocGeneration = 0;

// Grow-only device array
template <class T> struct DBuf
{
    T *p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t n)
    {
        if (n <= cap) return cudaSuccess;
        g_allocGeneration++;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        const cudaError_t e = cudaMalloc((void **)&p, std::max<size_t>(n, 1) * sizeof(T));
        if (e == cudaSuccess) cap = n;
        return e;
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

// Pinned bump ring for the per-frame uploads: the CUDA form of gfx::RingBuffer
// (src/gfx/RingBuffer.hpp:14-76; RingBuffer.cpp:56-133): 256-byte aligned writes, wraps to 0, and
// instead of asserting on a stomp it waits for the copies of the previous lap to retire.
struct UploadRing
{
    uint8_t *base = nullptr;
    size_t capacity = 0, head = 0;
    uint32_t wraps = 0; // laps completed (prt_timings::ringWraps; the wrap path has its own test)
    cudaEvent_t lap = nullptr;
    static constexpr size_t kAlignment = 256; // RingBuffer.hpp:18 sAlignment

    cudaError_t init(size_t bytes)
    {
        capacity = bytes;
        cudaError_t e = cudaHostAlloc((void **)&base, bytes, cudaHostAllocDefault);
        if (e != cudaSuccess) return e;
        return cudaEventCreateWithFlags(&lap, cudaEventDisableTiming);
    }
    void destroy()
    {
        if (base) cudaFreeHost(base);
        if (lap) cudaEventDestroy(lap);
        base = nullptr;
    }
    // Copies `bytes` from src into the ring and enqueues ring -> dst on `stream`.
    cudaError_t upload(void *dst, const void *src, size_t bytes, cudaStream_t stream)
    {
        if (bytes == 0) return cudaSuccess;
        if (bytes > capacity / 2) return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, stream); // staged by the driver
        size_t start = (head + kAlignment - 1) / kAlignment * kAlignment;
        if (start + bytes > capacity)
        {
            // wrap: everything enqueued from the previous lap must retire before its bytes are reused
            cudaError_t e = cudaEventRecord(lap, stream);
            if (e != cudaSuccess) return e;
            e = cudaEventSynchronize(lap);
            if (e != cudaSuccess) return e;
            start = 0;
            wraps++;
        }
        memcpy(base + start, src, bytes);
        head = start + bytes;
        return cudaMemcpyAsync(dst, base + start, bytes, cudaMemcpyHostToDevice, stream);
    }
};

struct TexRes
{
    cudaArray_t array = nullptr; // level 0 (texture() in the RT stages reads LOD 0, materials.glsl:44)
    uint8_t *payload = nullptr;  // level 0 as cached (BC7 blocks / RGBA8 texels) for the software filter
    uint8_t *alpha = nullptr;    // level-0 alpha plane (k_alpha_plane); nullptr + alphaOpaque when all 255
    bool alphaOpaque = false;
    uint32_t width = 0, height = 0, format = 0, mipLevels = 0;
};

// Every node / triangle array the trace kernels walk is allocated inside ONE reserved address range, so that a ray can name
// the arrays of its level by 32-bit distances from Accel::arena (core_trace.h). Plain cudaMalloc does not give that: under
// torch + NCCL two allocations of one process were seen more than 512 GB apart. The range is a 256 GB reservation
// (cuMemAddressReserve: address space only), physical memory is created and mapped block by block (cuMemCreate / cuMemMap);
// a freed block returns its memory, its addresses are not handed out again.
struct AccelHeap
{
    static constexpr size_t RESERVE = 256ull << 30;
    CUdeviceptr base = 0;
    size_t used = 0, gran = 0;
    int device = 0;
    struct Block { CUdeviceptr p; size_t bytes; CUmemGenericAllocationHandle h; };
    std::vector<Block> blocks;
    CUresult (*memAddressReserve)(CUdeviceptr *, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*memAddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*memCreate)(CUmemGenericAllocationHandle *, size_t, const CUmemAllocationProp *, unsigned long long) = nullptr;
    CUresult (*memRelease)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*memMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*memUnmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*memSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc *, size_t) = nullptr;
    CUresult (*memGetAllocationGranularity)(size_t *, const CUmemAllocationProp *, CUmemAllocationGranularity_flags) = nullptr;

    CUmemAllocationProp prop() const
    {
        CUmemAllocationProp p{};
        p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
        p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        p.location.id = device;
        return p;
    }
    prt_status init(int dev)
    {
        device = dev;
        struct { const char *name; void **fn; } wanted[] = {
            {"cuMemAddressReserve", (void **)&memAddressReserve}, {"cuMemAddressFree", (void **)&memAddressFree},
            {"cuMemCreate", (void **)&memCreate}, {"cuMemRelease", (void **)&memRelease}, {"cuMemMap", (void **)&memMap},
            {"cuMemUnmap", (void **)&memUnmap}, {"cuMemSetAccess", (void **)&memSetAccess},
            {"cuMemGetAllocationGranularity", (void **)&memGetAllocationGranularity}};
        for (auto &w : wanted)
        {
            cudaDriverEntryPointQueryResult found;
            CK(cudaGetDriverEntryPoint(w.name, w.fn, cudaEnableDefault, &found));
            if (found != cudaDriverEntryPointSuccess || !*w.fn) return fail(PRT_ERR_CUDA, std::string("driver entry point missing: ") + w.name);
        }
        const CUmemAllocationProp p = prop();
        if (memGetAllocationGranularity(&gran, &p, CU_MEM_ALLOC_GRANULARITY_MINIMUM) != CUDA_SUCCESS || gran == 0)
            return fail(PRT_ERR_CUDA, "cuMemGetAllocationGranularity failed");
        if (memAddressReserve(&base, RESERVE, gran, 0, 0) != CUDA_SUCCESS) return fail(PRT_ERR_CUDA, "cuMemAddressReserve(256 GB) failed");
        return PRT_OK;
    }
    prt_status alloc(void **out, size_t bytes)
    {
        const size_t size = (std::max<size_t>(bytes, 1) + gran - 1) / gran * gran;
        if (used + size > RESERVE) return fail(PRT_ERR_CUDA, "acceleration-structure address range (256 GB) exhausted");
        Block b{base + used, size, 0};
        const CUmemAllocationProp p = prop();
        if (memCreate(&b.h, size, &p, 0) != CUDA_SUCCESS) return fail(PRT_ERR_CUDA, "cuMemCreate failed (out of device memory?)");
        if (memMap(b.p, size, 0, b.h, 0) != CUDA_SUCCESS) { memRelease(b.h); return fail(PRT_ERR_CUDA, "cuMemMap failed"); }
        CUmemAccessDesc access{};
        access.location = p.location;
        access.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
        if (memSetAccess(b.p, size, &access, 1) != CUDA_SUCCESS) { memUnmap(b.p, size); memRelease(b.h); return fail(PRT_ERR_CUDA, "cuMemSetAccess failed"); }
        used += size;
        blocks.push_back(b);
        *out = (void *)(uintptr_t)b.p;
        return PRT_OK;
    }
    void release(void *ptr) // the stream must be idle
    {
        for (size_t i = 0; i < blocks.size(); ++i)
            if (blocks[i].p == (CUdeviceptr)(uintptr_t)ptr)
            {
                memUnmap(blocks[i].p, blocks[i].bytes);
                memRelease(blocks[i].h);
                blocks.erase(blocks.begin() + (long)i);
                return;
            }
    }
    void destroy()
    {
        for (const Block &b : blocks) { memUnmap(b.p, b.bytes); memRelease(b.h); }
        blocks.clear();
        if (base) memAddressFree(base, RESERVE);
        base = 0;
    }
};

struct BlasRes
{
    WideNode *nodes = nullptr;
    TriRec *tris = nullptr;
    AlphaUv *uvs = nullptr; // any-hit records, parallel to tris
    uint32_t nodeCount = 0, triCount = 0; // triCount = triangle RECORDS (builder references, core_build.h)
    uint32_t primCount = 0;               // triangles of the model
    uint32_t nodeCap = 0, triCap = 0; // a rebuild reuses the buffers while they are large enough
    bool built = false;
};

struct prt_ctx
{
    prt_config cfg{};
    cudaStream_t stream = nullptr;
    UploadRing ring;

    // scene data (WorldData buffers)
    std::vector<uint32_t *> geomBufs;
    DBuf<const uint32_t *> geomTable;
    std::vector<prt_geometry_metadata> hMeta;
    std::vector<prt_mesh_info> hInfo;
    DBuf<prt_geometry_metadata> dMeta;
    std::vector<prt_material_data> hMat;
    DBuf<prt_material_data> dMat;
    std::vector<prt_sampler_desc> hSamplers;
    std::vector<TexRes> textures;
    std::vector<TexSlot> hSlots;
    DBuf<TexSlot> dSlots;
    bool slotsDirty = true;
    DBuf<uint16_t> dSky;
    uint32_t skyFace = 0;
    std::vector<uint32_t> modelOffsets;
    std::vector<prt_sub_model> subModels;
    std::vector<BlasRes> blas;

    // per-frame rings
    DBuf<prt_draw_instance> dDraw;
    DBuf<prt_model_instance_transforms> dXf;
    DBuf<prt_directional_light> dDir;
    DBuf<prt_point_light> dPoints;
    DBuf<prt_spot_light> dSpots;
    uint32_t pointCount = 0, spotCount = 0;
    DBuf<prt_camera_uniforms> dCam;
    bool haveCamera = false, haveLights = false;

    // TLAS
    AccelHeap heap;
    WideNode *tlasNodesP = nullptr; // in `heap`, like every BLAS's nodes and triangles
    uint32_t tlasNodesCap = 0;
    DBuf<InstRec> tlasInst;
    uint32_t tlasInstances = 0;
    std::vector<TlasInput> hTlasInput;

    // builder scratch
    DBuf<TriRec> bTris;
    DBuf<AlphaUv> bUvs;
    DBuf<AlphaMat> dAlphaMats; // per draw instance, rebuilt when draw instances / materials / textures change
    uint32_t drawCount = 0;
    bool alphaMatsDirty = true;
    DBuf<InstRec> bInst;
    DBuf<TlasInput> bTlasIn;
    DBuf<prt_tlas_instance> bTlasRaw; // the caller's records as they are (all instances active)
    DBuf<BlasTableEntry> dBlasTable;  // model -> BLAS buffers, for k_tlas_instances_direct
    DBuf<Box> bBoxes, bNodeBox;
    DBuf<uint64_t> bKeys[2];
    DBuf<uint32_t> bOrder[2], bHist, bFirst, bLast, bVisit, bPrimOut, bScalars;
    DBuf<int32_t> bLeft, bRight, bParentI, bParentL;
    DBuf<CollapseItem> bQueue[2];
    // PLOC scratch (BLAS builds)
    DBuf<int32_t> bPlocRef[2], bPlocMergedRef;
    DBuf<Box> bPlocBox[2], bPlocMergedBox;
    DBuf<uint32_t> bPlocNearest, bPlocBlocks, bNewPos;
    // binned-SAH scratch (BLAS builds)
    DBuf<uint32_t> bSahKey[2], bSahQueue[2], bSahCent;
    DBuf<SahRange> bSahRange[2];
    DBuf<SahSplit> bSahSplit;
    DBuf<SahBin> bSahBins;
    DBuf<WideNode> bNodes;
    DBuf<BuildBounds> bBounds;

    // path state
    PathBuffers pb{};
    uint64_t pathCapacity = 0;
    DBuf<float4> accum, sum;
    DBuf<uint2> f16;
    DBuf<uint32_t> sortKeys, sortPerm, sortHist; // ray reordering (kernels_render.cuh)
    uint32_t raySort = 0;                        // off: measured no gain. bit 0: closest queues of bounce >= 1, bit 1: shadow queues of bounce >= 1, bit 2: shadow queue of bounce 0
    // ReSTIR DI trace pass (8(f)2): G-buffer, reservoirs and its own accumulation image
    DBuf<uint32_t> gbAlbedoRoughness, gbNormalMetallic;
    DBuf<float> gbDepth;
    DBuf<float2> rtdiReservoirs;
    DBuf<float4> rtdiAccum;
    uint32_t rtdiWidth = 0, rtdiHeight = 0;
    bool rtdiGbufferValid = false, rtdiReservoirsValid = false, rtdiHistoryValid = false;
    DBuf<uint32_t> toneLut; // R9G9B9E5 3D LUT of the display transform (8(f)3)
    uint32_t toneLutDim = 0;
    DBuf<uint32_t> ldr;
    DBuf<float> mipTables; // decode tables, code midpoints and filter taps of the mip chain (8(f)1)
    DBuf<prt_hit> userHits;
    DBuf<float4> userO, userD;
    DBuf<uint32_t> userSeeds;
    uint32_t accumWidth = 0, accumHeight = 0;
    bool historyValid = false;
    GlobalCounters *dCounters = nullptr;
    uint64_t kernelLaunches = 0;
    uint32_t traceBlocks = 148; // grid of the persistent trace kernels
    uint32_t poolRays = PRT_POOL_RAYS;   // rays per thread of the persistent trace kernels (kernels_render.cuh)
    uint32_t refillAt = PRT_POOL_REFILL; // idle ray slots per warp that trigger a refill
    // TLAS rebuild as a CUDA graph (prt_build_tlas); PRT_TLAS_GRAPH=0 keeps the eager launches (experiment switch)
    bool tlasGraphs = true, tlasGraphDirect = false, tlasEagerDirect = false;
    bool blasTableDirty = true; // dBlasTable lags behind `blas`
    cudaGraphExec_t tlasGraphExec = nullptr;
    uint32_t tlasGraphN = 0, tlasGraphLaunches = 0, tlasEagerN = 0, tlasGraphReplays = 0, tlasGraphCaptures = 0;
    unsigned long long tlasGraphGeneration = 0, tlasEagerGeneration = 0;
    uint32_t sahBlocks = 0;     // grid of k_sah_all_levels (co-resident by construction)
    uint32_t coopBlocks = 0;    // co-resident blocks of k_collapse_all_levels_grid (0: cooperative launch unsupported)

    // timing
    cudaEvent_t evRecord[2]{}, evTlas[2]{}, evBlas[2]{};
    bool recordTimed = false, tlasTimed = false, blasTimed = false;
    uint32_t lastRecordLaunches = 0, lastTlasLaunches = 0;
    // per-stage timing of the last prt_record: an event between every two stage launches (pool grows on demand);
    // stageOf[i] = stage of the launch between evStage[i] and evStage[i + 1] (0 closest, 1 shade, 2 shadow, 3 other)
    std::vector<cudaEvent_t> evStage;
    std::vector<uint8_t> stageOf;
    uint32_t evUsed = 0;

    cudaError_t mark(int stageOfNextLaunch)
    {
        if (evUsed == evStage.size())
        {
            cudaEvent_t e;
            cudaError_t err = cudaEventCreate(&e);
            if (err != cudaSuccess) return err;
            evStage.push_back(e);
            stageOf.push_back(3);
        }
        stageOf[evUsed] = (uint8_t)stageOfNextLaunch;
        return cudaEventRecord(evStage[evUsed++], stream);
    }

    SceneView view() const
    {
        SceneView sv{};
        sv.geometryBuffers = geomTable.p;
        sv.metadatas = dMeta.p;
        sv.materials = dMat.p;
        sv.drawInstances = dDraw.p;
        sv.transforms = dXf.p;
        sv.texSlots = dSlots.p;
        sv.textureCount = (uint32_t)textures.size();
        sv.samplerCount = (uint32_t)std::max<size_t>(hSamplers.size(), 1);
        sv.directional = dDir.p;
        sv.pointLights = dPoints.p;
        sv.spotLights = dSpots.p;
        sv.pointLightCount = pointCount;
        sv.spotLightCount = spotCount;
        sv.camera = dCam.p;
        sv.sky = skyFace ? dSky.p : nullptr;
        sv.skyFaceSize = skyFace;
        sv.exactMaterialFilter = (cfg.flags & PRT_CFG_EXACT_FILTER) ? 1u : 0u;
        sv.alphaMats = (!alphaMatsDirty && drawCount) ? dAlphaMats.p : nullptr;
        return sv;
    }
    Accel accel() const
    {
        Accel a{};
        a.tlasNodes = tlasInstances ? tlasNodesP : nullptr;
        a.instances = tlasInst.p;
        a.instanceCount = tlasInstances;
        a.planeBias = PLANE_BIAS;
        a.arena = accelArena;
        return a;
    }
    uint64_t accelArena = 0; // see Accel::arena; refreshed by prt_build_tlas (every BLAS an instance may name exists by then)
};

extern "C" const char *prt_last_error(void) { return g_lastError.c_str(); }

// The persistent trace kernels are instantiated for 1 and 2 rays per thread (3 measured slower: its rows squeeze L1); PRT_POOL_RAYS (build default) or the
// environment (experiment switch) picks one per context.
template <int R, bool SHADOW>
static void launchTraceR(prt_ctx *c, bool count, uint32_t grid, const Accel &as, const SceneView &sv, int q, uint32_t b, const uint32_t *perm)
{
    if (count) k_trace<R, SHADOW, true><<<grid, RENDER_THREADS, 0, c->stream>>>(as, sv, c->pb, q, b, c->dCounters, perm, c->refillAt);
    else k_trace<R, SHADOW, false><<<grid, RENDER_THREADS, 0, c->stream>>>(as, sv, c->pb, q, b, c->dCounters, perm, c->refillAt);
}
template <bool SHADOW>
static void launchTrace(prt_ctx *c, bool count, uint32_t grid, const Accel &as, const SceneView &sv, int q, uint32_t b, const uint32_t *perm)
{
    switch (c->poolRays)
    {
    case 1: launchTraceR<1, SHADOW>(c, count, grid, as, sv, q, b, perm); break;
    case 3: launchTraceR<3, SHADOW>(c, count, grid, as, sv, q, b, perm); break;
    default: launchTraceR<2, SHADOW>(c, count, grid, as, sv, q, b, perm); break;
    }
}
template <int R>
static cudaError_t configureTrace(prt_ctx *c, int sms)
{
    // Blocks per SM from the launch bounds (registers) and the ray rows (shared memory), then a shared-memory carve-out
    // that is just large enough for them: what is left of the 256 KB is L1, which serves the nodes, the triangles and
    // the per-thread stacks (with the maximum carve-out the L1 hit rate fell from 70 % to 32 %).
    int perSm = 0, dev = 0;
    cudaError_t err = cudaSuccess;
    cudaFuncAttributes fa;
    if ((err = cudaFuncGetAttributes(&fa, k_trace<R, false, false>)) != cudaSuccess) return err;
    if ((err = cudaGetDevice(&dev)) != cudaSuccess) return err;
    cudaDeviceProp prop;
    if ((err = cudaGetDeviceProperties(&prop, dev)) != cudaSuccess) return err;
    const size_t perBlock = fa.sharedSizeBytes + prop.reservedSharedMemPerBlock;
    const int byRegs = (int)(prop.regsPerMultiprocessor / std::max(1, fa.numRegs * RENDER_THREADS));
    const int bySmem = (int)(prop.sharedMemPerMultiprocessor / perBlock);
    perSm = std::max(1, std::min(std::min(byRegs, bySmem), PRT_TRACE_MIN_BLOCKS));
    if (const char *e = std::getenv("PRT_TRACE_BLOCKS_PER_SM")) perSm = std::min(perSm, std::max(1, std::atoi(e))); // experiment switch
    int carve = (int)std::min<size_t>(100, (perSm * perBlock * 100 + prop.sharedMemPerMultiprocessor - 1) / prop.sharedMemPerMultiprocessor);
    if (const char *e = std::getenv("PRT_TRACE_CARVEOUT")) carve = std::atoi(e); // experiment switch
    if ((err = cudaFuncSetAttribute(k_trace<R, false, false>, cudaFuncAttributePreferredSharedMemoryCarveout, carve)) != cudaSuccess) return err;
    if ((err = cudaFuncSetAttribute(k_trace<R, true, false>, cudaFuncAttributePreferredSharedMemoryCarveout, carve)) != cudaSuccess) return err;
    if ((err = cudaFuncSetAttribute(k_trace<R, false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, carve)) != cudaSuccess) return err;
    if ((err = cudaFuncSetAttribute(k_trace<R, true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, carve)) != cudaSuccess) return err;
    int occ = 0;
    if ((err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_trace<R, false, false>, RENDER_THREADS, 0)) != cudaSuccess) return err;
    perSm = std::max(1, std::min(perSm, occ));
    c->traceBlocks = (uint32_t)std::max(1, sms * std::max(1, perSm));
    return cudaSuccess;
}

// ---- lifetime: RtReference::init / ~RtReference (RtRef.cpp:90-120) ------------------------------------
static prt_status allocPathBuffers(prt_ctx *c, uint64_t paths)
{
    if (paths <= c->pathCapacity) return PRT_OK;
    PathBuffers &pb = c->pb;
    void **ptrs[] = {(void **)&pb.rng,      (void **)&pb.throughput, (void **)&pb.color,   (void **)&pb.rayO[0], (void **)&pb.rayO[1],
                     (void **)&pb.rayD[0],  (void **)&pb.rayD[1],    (void **)&pb.rayId[0], (void **)&pb.rayId[1], (void **)&pb.hits,
                     (void **)&pb.shadowO,  (void **)&pb.shadowD,    (void **)&pb.shadowId, (void **)&pb.shadowC, (void **)&pb.tie};
    const size_t elem[] = {16, 16, 16, 16, 16, 16, 16, 8, 8, 16, 16, 16, 8, 16, 8};
    for (size_t i = 0; i < sizeof(elem) / sizeof(elem[0]); ++i)
    {
        if (*ptrs[i]) CK(cudaFree(*ptrs[i]));
        *ptrs[i] = nullptr;
        CK(cudaMalloc(ptrs[i], paths * elem[i]));
    }
    if (!pb.counts)
    {
        CK(cudaMalloc((void **)&pb.counts, COUNT_WORDS * sizeof(uint32_t)));
        CK(cudaMemset(pb.counts, 0, COUNT_WORDS * sizeof(uint32_t)));
    }
    CK(c->sortKeys.ensure(paths));
    CK(c->sortPerm.ensure(paths));
    CK(c->sortHist.ensure(RAY_SORT_BINS));
    c->pathCapacity = paths;
    return PRT_OK;
}

static prt_status createContext(const prt_config *cfg, prt_ctx **out);
extern "C" prt_status prt_create(const prt_config *cfg, prt_ctx **out)
{
    if (!cfg || !out) return fail(PRT_ERR_INVALID, "prt_create: null argument");
    *out = nullptr;
    const prt_status st = createContext(cfg, out);
    if (st != PRT_OK && *out)
    {
        // a half-initialised context never reaches the caller
        const std::string why = g_lastError;
        prt_destroy(*out);
        *out = nullptr;
        g_lastError = why;
    }
    return st;
}

static prt_status createContext(const prt_config *cfg, prt_ctx **out)
{
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0)
        return fail(PRT_ERR_NO_DEVICE, "prt_create: no CUDA device (this library has no CPU fallback)");
    if (cfg->device < 0 || cfg->device >= count) return fail(PRT_ERR_NO_DEVICE, "prt_create: device ordinal out of range");
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, cfg->device));
    if (prop.major < 10) return fail(PRT_ERR_NO_DEVICE, std::string("prt_create: ") + prop.name + " is not an sm_100 device");
    CK(cudaSetDevice(cfg->device));
    prt_ctx *c = new prt_ctx;
    c->cfg = *cfg;
    if (c->cfg.samplesPerLaunch == 0) c->cfg.samplesPerLaunch = 1;
    *out = c;
    CK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    CKS(c->heap.init(cfg->device));
    {
        size_t ringBytes = 32u << 20;
        if (const char *e = std::getenv("PRT_RING_BYTES")) ringBytes = std::max<size_t>(4096, (size_t)std::atoll(e)); // test switch: forces wraps
        CK(c->ring.init(ringBytes));
    }
    CK(cudaMalloc((void **)&c->dCounters, sizeof(GlobalCounters)));
    CK(cudaMemset(c->dCounters, 0, sizeof(GlobalCounters)));
    for (int i = 0; i < 2; ++i)
    {
        CK(cudaEventCreate(&c->evRecord[i]));
        CK(cudaEventCreate(&c->evTlas[i]));
        CK(cudaEventCreate(&c->evBlas[i]));
    }
    CK(c->dDir.ensure(1));
    CK(c->dCam.ensure(1));
    CK(c->dPoints.ensure(PRT_MAX_POINT_LIGHTS));
    CK(c->dSpots.ensure(PRT_MAX_SPOT_LIGHTS));
    CK(c->bBounds.ensure(1));
    CK(c->bScalars.ensure(512)); // [256 ..] the binned-SAH per-level counts
    const uint64_t pixels = (uint64_t)cfg->maxWidth * cfg->maxHeight;
    if (pixels)
    {
        CKS(allocPathBuffers(c, pixels * c->cfg.samplesPerLaunch));
        CK(c->accum.ensure(pixels));
        CK(cudaMemsetAsync(c->accum.p, 0, pixels * sizeof(float4), c->stream));
    }
    {
        // persistent trace kernels: one resident wave of blocks (SM count x blocks per SM)
        int sms = 0;
        CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->cfg.device));
        if (const char *e = std::getenv("PRT_POOL_RAYS")) c->poolRays = (uint32_t)std::min(3, std::max(1, std::atoi(e))); // experiment switch
        c->refillAt = 8u * c->poolRays;
        CK(c->poolRays == 1 ? configureTrace<1>(c, sms) : c->poolRays == 3 ? configureTrace<3>(c, sms) : configureTrace<2>(c, sms));
        if (const char *e = std::getenv("PRT_POOL_REFILL")) c->refillAt = (uint32_t)std::max(1, std::atoi(e)); // experiment switch
        if (const char *e = std::getenv("PRT_RAY_SORT")) c->raySort = (uint32_t)std::atoi(e); // experiment switch
        int coop = 0, perSmCollapse = 0;
        CK(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, c->cfg.device));
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSmCollapse, k_collapse_all_levels_grid, 128, 0));
        c->coopBlocks = coop ? (uint32_t)(sms * std::max(0, perSmCollapse)) : 0u;
        if (const char *e = std::getenv("PRT_TLAS_GRAPH")) c->tlasGraphs = std::atoi(e) != 0;
        int perSmSah = 0;
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSmSah, k_sah_all_levels, SAH_THREADS, 0));
        int sahPerSm = std::min(perSmSah, 4); // more blocks only lengthen the grid barriers (3 per level)
        if (const char *e = std::getenv("PRT_SAH_BLOCKS_PER_SM")) sahPerSm = std::min(perSmSah, std::max(1, std::atoi(e))); // experiment switch
        c->sahBlocks = coop ? (uint32_t)(sms * std::max(0, sahPerSm)) : 0u;
    }
    CK(cudaStreamSynchronize(c->stream));
    return PRT_OK;
}

extern "C" void prt_destroy(prt_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->cfg.device);
    cudaDeviceSynchronize();
    if (c->tlasGraphExec) cudaGraphExecDestroy(c->tlasGraphExec);
    for (uint32_t *b : c->geomBufs)
        if (b) cudaFree(b);
    for (TexSlot &s : c->hSlots)
    {
        if (s.texels) cudaDestroyTextureObject(s.texels);
        if (s.linear) cudaDestroyTextureObject(s.linear);
    }
    for (TexRes &t : c->textures)
    {
        if (t.array) cudaFreeArray(t.array);
        if (t.payload) cudaFree(t.payload);
        if (t.alpha) cudaFree(t.alpha);
    }
    for (BlasRes &b : c->blas)
    {
        if (b.nodes) c->heap.release(b.nodes);
        if (b.tris) c->heap.release(b.tris);
        if (b.uvs) cudaFree(b.uvs);
    }
    c->heap.destroy(); // the TLAS nodes with it
    PathBuffers &pb = c->pb;
    void *ptrs[] = {pb.rng,      pb.throughput, pb.color,   pb.rayO[0], pb.rayO[1], pb.rayD[0],  pb.rayD[1], pb.rayId[0],
                    pb.rayId[1], pb.hits,       pb.shadowO, pb.shadowD, pb.shadowId, pb.shadowC, pb.counts,  c->dCounters, pb.tie};
    for (void *p : ptrs)
        if (p) cudaFree(p);
    c->geomTable.release(); c->dMeta.release(); c->dMat.release(); c->dSlots.release(); c->dSky.release();
    c->dDraw.release(); c->dXf.release(); c->dDir.release(); c->dPoints.release(); c->dSpots.release(); c->dCam.release();
    c->tlasInst.release();
    c->bTris.release(); c->bUvs.release(); c->dAlphaMats.release(); c->bInst.release(); c->bTlasIn.release(); c->bTlasRaw.release(); c->dBlasTable.release(); c->bBoxes.release(); c->bNodeBox.release();
    for (int i = 0; i < 2; ++i) { c->bKeys[i].release(); c->bOrder[i].release(); c->bQueue[i].release(); }
    for (int i = 0; i < 2; ++i) { c->bPlocRef[i].release(); c->bPlocBox[i].release(); }
    c->bPlocMergedRef.release(); c->bPlocMergedBox.release(); c->bPlocNearest.release(); c->bPlocBlocks.release(); c->bNewPos.release();
    for (int i = 0; i < 2; ++i) { c->bSahKey[i].release(); c->bSahRange[i].release(); c->bSahQueue[i].release(); }
    c->bSahCent.release(); c->bSahSplit.release(); c->bSahBins.release();
    c->bHist.release(); c->bFirst.release(); c->bLast.release(); c->bVisit.release(); c->bPrimOut.release(); c->bScalars.release();
    c->bLeft.release(); c->bRight.release(); c->bParentI.release(); c->bParentL.release(); c->bNodes.release(); c->bBounds.release();
    c->accum.release(); c->sum.release(); c->f16.release(); c->userHits.release(); c->userO.release(); c->userD.release();
    c->userSeeds.release();
    c->gbAlbedoRoughness.release(); c->gbNormalMetallic.release(); c->gbDepth.release(); c->rtdiReservoirs.release(); c->rtdiAccum.release();
    c->toneLut.release(); c->ldr.release(); c->mipTables.release(); c->sortKeys.release(); c->sortPerm.release(); c->sortHist.release();
    for (int i = 0; i < 2; ++i)
    {
        if (c->evRecord[i]) cudaEventDestroy(c->evRecord[i]);
        if (c->evTlas[i]) cudaEventDestroy(c->evTlas[i]);
        if (c->evBlas[i]) cudaEventDestroy(c->evBlas[i]);
    }
    for (cudaEvent_t e : c->evStage) cudaEventDestroy(e);
    c->ring.destroy();
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

extern "C" void *prt_stream(prt_ctx *c) { return c ? (void *)c->stream : nullptr; }

extern "C" prt_status prt_synchronize(prt_ctx *c)
{
    REQUIRE(c, "prt_synchronize");
    CK(cudaStreamSynchronize(c->stream));
    return PRT_OK;
}

// ---- scene data ------------------------------------------------------------------------------------------
extern "C" prt_status prt_set_geometry_buffer(prt_ctx *c, uint32_t index, const void *data, uint64_t bytes)
{
    REQUIRE(c && data, "prt_set_geometry_buffer");
    REQUIRE(bytes % 4 == 0, "geometry buffers are u32 arrays (WorldData.cpp:1271-1320)");
    REQUIRE(index < 100, "at most 100 geometry buffers (WorldData.cpp:31)");
    CK(cudaSetDevice(c->cfg.device));
    if (c->geomBufs.size() <= index) c->geomBufs.resize(index + 1, nullptr);
    if (c->geomBufs[index]) CK(cudaFree(c->geomBufs[index]));
    c->geomBufs[index] = nullptr;
    // 16 bytes of slack: loadIndex() on a short-index mesh may read the padding half-word
    CK(cudaMalloc((void **)&c->geomBufs[index], bytes + 16));
    CK(cudaMemcpyAsync(c->geomBufs[index], data, bytes, cudaMemcpyHostToDevice, c->stream));
    CK(c->geomTable.ensure(100));
    std::vector<const uint32_t *> table(100, nullptr);
    for (size_t i = 0; i < c->geomBufs.size(); ++i) table[i] = c->geomBufs[i];
    CK(cudaMemcpyAsync(c->geomTable.p, table.data(), 100 * sizeof(uint32_t *), cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream)); // `table` and `data` may go away after return
    return PRT_OK;
}

extern "C" prt_status prt_set_meshes(prt_ctx *c, const prt_geometry_metadata *metadatas, const prt_mesh_info *infos, uint32_t count)
{
    REQUIRE(c && (count == 0 || (metadatas && infos)), "prt_set_meshes");
    CK(cudaSetDevice(c->cfg.device));
    c->hMeta.assign(metadatas, metadatas + count);
    c->hInfo.assign(infos, infos + count);
    CK(c->dMeta.ensure(count));
    CK(c->ring.upload(c->dMeta.p, metadatas, count * sizeof(prt_geometry_metadata), c->stream));
    return PRT_OK;
}

extern "C" prt_status prt_set_materials(prt_ctx *c, const prt_material_data *materials, uint32_t count)
{
    REQUIRE(c && (count == 0 || materials), "prt_set_materials");
    CK(cudaSetDevice(c->cfg.device));
    c->hMat.assign(materials, materials + count);
    CK(c->dMat.ensure(count));
    CK(c->ring.upload(c->dMat.p, materials, count * sizeof(prt_material_data), c->stream));
    c->alphaMatsDirty = true;
    return PRT_OK;
}

extern "C" prt_status prt_set_samplers(prt_ctx *c, const prt_sampler_desc *samplers, uint32_t count)
{
    REQUIRE(c && (count == 0 || samplers), "prt_set_samplers");
    REQUIRE(count <= 256, "sampler index is 8 bits (material_data.h:19-36)");
    c->hSamplers.assign(samplers, samplers + count);
    c->slotsDirty = true;
    return PRT_OK;
}

extern "C" prt_status prt_set_texture(prt_ctx *c, uint32_t index, const prt_texture_desc *desc)
{
    REQUIRE(c && desc && desc->data, "prt_set_texture");
    REQUIRE(index >= 1 && index < (1u << 24), "texture index 0 is the reserved 'no texture' slot (WorldData.cpp:727-741)");
    REQUIRE(desc->width > 0 && desc->height > 0, "empty texture");
    REQUIRE(desc->format == PRT_TEX_RGBA8_UNORM || desc->format == PRT_TEX_BC7_UNORM, "unknown texture format");
    {
        // validate the payload before anything is released or allocated
        const uint64_t need = desc->format == PRT_TEX_BC7_UNORM ? (uint64_t)((desc->width + 3) / 4) * ((desc->height + 3) / 4) * 16
                                                                 : (uint64_t)desc->width * desc->height * 4;
        REQUIRE(desc->dataBytes >= need, desc->format == PRT_TEX_BC7_UNORM ? "BC7 payload shorter than level 0" : "RGBA8 payload shorter than level 0");
    }
    CK(cudaSetDevice(c->cfg.device));
    if (c->textures.size() <= index) c->textures.resize(index + 1);
    TexRes &t = c->textures[index];
    if (t.array)
    {
        CK(cudaStreamSynchronize(c->stream));
        CK(cudaFreeArray(t.array));
        t.array = nullptr;
        if (t.payload) CK(cudaFree(t.payload));
        t.payload = nullptr;
        if (t.alpha) CK(cudaFree(t.alpha));
        t.alpha = nullptr;
        t.alphaOpaque = false;
    }
    t.width = desc->width; t.height = desc->height; t.format = desc->format; t.mipLevels = desc->mipLevels;
    const uint32_t bx = (desc->width + 3) / 4, by = (desc->height + 3) / 4;
    if (desc->format == PRT_TEX_BC7_UNORM && !(c->cfg.flags & PRT_CFG_DECODE_BC7))
    {
        REQUIRE(desc->dataBytes >= (uint64_t)bx * by * 16, "BC7 payload shorter than level 0");
        // stored UNORM, linearised in the shader after filtering (Texture.cpp:429-432)
        const cudaChannelFormatDesc cd = cudaCreateChannelDesc<cudaChannelFormatKindUnsignedBlockCompressed7>();
        CK(cudaMallocArray(&t.array, &cd, desc->width, desc->height));
        CK(cudaMemcpy2DToArrayAsync(t.array, 0, 0, desc->data, (size_t)bx * 16, (size_t)bx * 16, by, cudaMemcpyHostToDevice, c->stream));
    }
    else
    {
        const cudaChannelFormatDesc cd = cudaCreateChannelDesc<uchar4>();
        CK(cudaMallocArray(&t.array, &cd, desc->width, desc->height));
        if (desc->format == PRT_TEX_BC7_UNORM)
        {
            REQUIRE(desc->dataBytes >= (uint64_t)bx * by * 16, "BC7 payload shorter than level 0");
            DBuf<uint2> blocks;
            DBuf<uint32_t> rgba;
            CK(blocks.ensure((size_t)bx * by * 2));
            CK(rgba.ensure((size_t)desc->width * desc->height));
            CK(cudaMemcpyAsync(blocks.p, desc->data, (size_t)bx * by * 16, cudaMemcpyHostToDevice, c->stream));
            k_bc7_decode<<<blocksFor((uint64_t)bx * by, 128), 128, 0, c->stream>>>(blocks.p, bx, by, desc->width, desc->height, rgba.p);
            CK(cudaGetLastError());
            CK(cudaMemcpy2DToArrayAsync(t.array, 0, 0, rgba.p, (size_t)desc->width * 4, (size_t)desc->width * 4, desc->height,
                                        cudaMemcpyDeviceToDevice, c->stream));
            CK(cudaStreamSynchronize(c->stream));
            blocks.release();
            rgba.release();
        }
        else
        {
            REQUIRE(desc->format == PRT_TEX_RGBA8_UNORM, "unknown texture format");
            REQUIRE(desc->dataBytes >= (uint64_t)desc->width * desc->height * 4, "RGBA8 payload shorter than level 0");
            CK(cudaMemcpy2DToArrayAsync(t.array, 0, 0, desc->data, (size_t)desc->width * 4, (size_t)desc->width * 4, desc->height,
                                        cudaMemcpyHostToDevice, c->stream));
        }
    }
    {
        const size_t bytes = desc->format == PRT_TEX_BC7_UNORM ? (size_t)bx * by * 16 : (size_t)desc->width * desc->height * 4;
        CK(cudaMalloc((void **)&t.payload, bytes));
        CK(cudaMemcpyAsync(t.payload, desc->data, bytes, cudaMemcpyHostToDevice, c->stream));
    }
    {
        // alpha plane for the any-hit stage (derived data like the BLAS; the payload stays the source of truth)
        const size_t texels = (size_t)desc->width * desc->height;
        REQUIRE(texels < (1ull << 32), "texture too large");
        CK(cudaMalloc((void **)&t.alpha, texels));
        uint32_t *minAlpha = c->bScalars.p + 250; // past the builder's per-level counts (bScalars holds 256)
        const uint32_t init = 255u;
        CK(cudaMemcpyAsync(minAlpha, &init, 4, cudaMemcpyHostToDevice, c->stream));
        TexSlot view{};
        view.payload = t.payload; view.format = t.format; view.width = t.width; view.height = t.height;
        k_alpha_plane<<<blocksFor(texels, 256), 256, 0, c->stream>>>(view, t.alpha, minAlpha);
        CK(cudaGetLastError());
        uint32_t lowest = 0;
        CK(cudaMemcpyAsync(&lowest, minAlpha, 4, cudaMemcpyDeviceToHost, c->stream));
        CK(cudaStreamSynchronize(c->stream));
        if (lowest == 255u)
        {
            CK(cudaFree(t.alpha));
            t.alpha = nullptr;
            t.alphaOpaque = true;
        }
    }
    CK(cudaStreamSynchronize(c->stream)); // desc->data may go away after return
    c->slotsDirty = true;
    return PRT_OK;
}

extern "C" prt_status prt_encode_bc7(prt_ctx *c, const void *rgba8, uint32_t width, uint32_t height, void *blocksOut, uint64_t blocksBytes)
{
    REQUIRE(c && rgba8 && blocksOut, "prt_encode_bc7");
    REQUIRE(width > 0 && height > 0 && (width & 3u) == 0 && (height & 3u) == 0,
            "BC7 levels are whole 4x4 blocks (Texture.cpp:229-237 falls back to RGBA8 otherwise)");
    const size_t nBlocks = (size_t)(width / 4) * (height / 4);
    REQUIRE(blocksBytes >= nBlocks * 16, "block buffer too small");
    CK(cudaSetDevice(c->cfg.device));
    DBuf<uint32_t> src;
    DBuf<uint2> dst;
    CK(src.ensure((size_t)width * height));
    CK(dst.ensure(nBlocks * 2));
    CK(cudaMemcpyAsync(src.p, rgba8, (size_t)width * height * 4, cudaMemcpyHostToDevice, c->stream));
    k_bc7_encode<<<blocksFor(nBlocks, 64), 64, 0, c->stream>>>(src.p, width, height, dst.p);
    c->kernelLaunches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(blocksOut, dst.p, nBlocks * 16, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    src.release();
    dst.release();
    return PRT_OK;
}

// getMipCount (Utils.hpp:144-148)
static uint32_t mipCountFor(uint32_t maxDimension) { uint32_t n = 0; while (maxDimension) { n++; maxDimension >>= 1; } return n; }

extern "C" prt_status prt_compress_texture(prt_ctx *c, const void *rgba8, uint32_t width, uint32_t height, uint32_t colorSpace,
                                           uint32_t generateMipMaps, void *dataOut, uint64_t dataBytes, uint32_t *formatOut,
                                           uint32_t *mipLevelCountOut)
{
    REQUIRE(c && rgba8 && dataOut && formatOut && mipLevelCountOut, "prt_compress_texture");
    REQUIRE(width > 0 && height > 0 && width <= 16384 && height <= 16384, "texture extent out of range");
    REQUIRE(colorSpace <= 1, "colorSpace is 0 (sRGB) or 1 (linear), TextureColorSpace in Texture.hpp");
    // Texture.cpp:217-237: full chain minus the 2x2 and 1x1 levels; BC7 only when every kept level is whole 4x4 blocks
    const uint32_t full = generateMipMaps ? mipCountFor(std::max(width, height)) : 1u;
    const uint32_t levels = full > 2 ? full - 2 : 1u;
    bool bc7 = true;
    for (uint32_t i = 0; i < levels; ++i)
        if ((std::max(width >> i, 1u) % 4) != 0 || (std::max(height >> i, 1u) % 4) != 0) bc7 = false;
    std::vector<uint64_t> rawOff(levels), outOff(levels);
    uint64_t rawTexels = 0, outBytes = 0;
    for (uint32_t i = 0; i < levels; ++i)
    {
        const uint64_t w = std::max(width >> i, 1u), h = std::max(height >> i, 1u);
        rawOff[i] = rawTexels;
        outOff[i] = outBytes;
        rawTexels += w * h;
        outBytes += bc7 ? (w / 4) * (h / 4) * 16 : w * h * 4; // Dds::Dds, Dds.cpp:95-133
    }
    REQUIRE(dataBytes >= outBytes, "destination too small for the level chain");
    CK(cudaSetDevice(c->cfg.device));
    if (c->mipTables.cap == 0)
    {
        // decode tables, the midpoints between neighbouring codes and the eight filter taps, each computed in double and
        // rounded once: [0,256) sRGB->linear, [256,512) b/255, [512,767) sRGB midpoints, [768,1023) linear midpoints,
        // [1024,1032) Mitchell(B=C=1/3) at 1.75 1.25 0.75 0.25 | 0.25 ... 1.75, normalised
        std::vector<float> t(1032, 0.f);
        auto lin = [](double s) { return s <= 0.04045 ? s / 12.92 : std::pow((s + 0.055) / 1.055, 2.4); };
        auto mitchell = [](double x) { return x < 1.0 ? (16.0 + x * x * (21.0 * x - 36.0)) / 18.0 : (32.0 + x * (-60.0 + x * (36.0 - 7.0 * x))) / 18.0; };
        for (int b = 0; b < 256; ++b) t[b] = (float)lin(b / 255.0), t[256 + b] = (float)(b / 255.0);
        for (int b = 0; b < 255; ++b) t[512 + b] = (float)lin((b + 0.5) / 255.0), t[768 + b] = (float)((b + 0.5) / 255.0);
        const double d[8] = {1.75, 1.25, 0.75, 0.25, 0.25, 0.75, 1.25, 1.75};
        double sum = 0;
        for (double x : d) sum += mitchell(x);
        for (int k = 0; k < 8; ++k) t[1024 + k] = (float)(mitchell(d[k]) / sum);
        CK(c->mipTables.ensure(t.size()));
        CK(cudaMemcpyAsync(c->mipTables.p, t.data(), t.size() * 4, cudaMemcpyHostToDevice, c->stream));
        CK(cudaStreamSynchronize(c->stream));
    }
    DBuf<uint32_t> raw;
    DBuf<uint2> blocks;
    CK(raw.ensure(rawTexels));
    if (bc7)
    {
        CK(blocks.ensure(outBytes / 8));
    }
    CK(cudaMemcpyAsync(raw.p, rgba8, (size_t)width * height * 4, cudaMemcpyHostToDevice, c->stream));
    for (uint32_t i = 0; i < levels; ++i)
    {
        const uint32_t w = std::max(width >> i, 1u), h = std::max(height >> i, 1u);
        if (i > 0)
        {
            const uint32_t pw = std::max(width >> (i - 1), 1u), ph = std::max(height >> (i - 1), 1u);
            const float *tp = c->mipTables.p;
            const bool srgb = colorSpace == 0;
            const MipTables tb{srgb ? tp : tp + 256, tp + 256, srgb ? tp + 512 : tp + 768, tp + 768, tp + 1024};
            k_downsample_mitchell<<<blocksFor((size_t)w * h, 128), 128, 0, c->stream>>>(raw.p + rawOff[i - 1], (int)pw, (int)ph,
                                                                                       raw.p + rawOff[i], (int)w, (int)h, tb);
            c->kernelLaunches++;
        }
        if (bc7)
        {
            const size_t nBlocks = (size_t)(w / 4) * (h / 4);
            k_bc7_encode<<<blocksFor(nBlocks, 64), 64, 0, c->stream>>>(raw.p + rawOff[i], w, h, blocks.p + outOff[i] / 8);
            c->kernelLaunches++;
        }
    }
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(dataOut, bc7 ? (const void *)blocks.p : (const void *)raw.p, outBytes, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    raw.release();
    blocks.release();
    *formatOut = bc7 ? PRT_TEX_BC7_UNORM : PRT_TEX_RGBA8_UNORM;
    *mipLevelCountOut = levels;
    return PRT_OK;
}

// ---- prt_compress_texture_static: the same level chain with NO device allocation ------------------------------------
// Added (no counterpart in the reference): callers that cannot reach cudaMalloc
// (graph capture, serving loops, harnesses that forbid it) get the mip + BC7 chain with its intermediate levels, its block
// staging and the filter tables in ONE file-scope device array. Every GPU that loads the library holds
// PRT_STATIC_SCRATCH_BYTES of it until exit. All calls share it (it keeps the last call's contents): calls that overlap on
// one GPU race. Nothing in it has to start at zero -- every byte a kernel reads was written earlier in the same call (level
// 0 by the upload, level i by the downsample of level i-1, the blocks by the encoder), and the tables are written once.
static __device__ __align__(256) unsigned char g_staticScratch[PRT_STATIC_SCRATCH_BYTES];
static const size_t STATIC_TABLE_BYTES = 4352; // 1032 floats, rounded up to 256 B
static const int STATIC_MAX_DEVICES = 64;
static unsigned char *g_staticScratchPtr[STATIC_MAX_DEVICES]; // DEVICE addresses of g_staticScratch, one per GPU set up
static std::mutex g_staticScratchMutex;

// Once per device, with that device current: the device address of the array (its name is only a host-side stand-in
// here) and the filter tables of prt_compress_texture, written into its first STATIC_TABLE_BYTES.
static prt_status staticScratchFor(int device, unsigned char **out)
{
    if (device < 0 || device >= STATIC_MAX_DEVICES) return fail(PRT_ERR_INVALID, "prt_compress_texture_static: device index beyond 63");
    std::lock_guard<std::mutex> lock(g_staticScratchMutex);
    if (!g_staticScratchPtr[device])
    {
        CK(cudaSetDevice(device));
        unsigned char *ptr = nullptr;
        CK(cudaGetSymbolAddress((void **)&ptr, g_staticScratch));
        if (!ptr) return fail(PRT_ERR_CUDA, "cudaGetSymbolAddress(g_staticScratch) returned a null address");
        float t[1032] = {}; // same tables, same arithmetic as prt_compress_texture
        auto lin = [](double s) { return s <= 0.04045 ? s / 12.92 : std::pow((s + 0.055) / 1.055, 2.4); };
        auto mitchell = [](double x) { return x < 1.0 ? (16.0 + x * x * (21.0 * x - 36.0)) / 18.0 : (32.0 + x * (-60.0 + x * (36.0 - 7.0 * x))) / 18.0; };
        for (int b = 0; b < 256; ++b) t[b] = (float)lin(b / 255.0), t[256 + b] = (float)(b / 255.0);
        for (int b = 0; b < 255; ++b) t[512 + b] = (float)lin((b + 0.5) / 255.0), t[768 + b] = (float)((b + 0.5) / 255.0);
        const double d[8] = {1.75, 1.25, 0.75, 0.25, 0.25, 0.75, 1.25, 1.75};
        double sum = 0;
        for (double x : d) sum += mitchell(x);
        for (int k = 0; k < 8; ++k) t[1024 + k] = (float)(mitchell(d[k]) / sum);
        CK(cudaMemcpy(ptr, t, sizeof(t), cudaMemcpyHostToDevice));
        g_staticScratchPtr[device] = ptr;
    }
    *out = g_staticScratchPtr[device];
    return PRT_OK;
}

extern "C" prt_status prt_compress_texture_static(prt_ctx *c, const void *rgba8, uint32_t width, uint32_t height, uint32_t colorSpace,
                                                  uint32_t generateMipMaps, void *dataOut, uint64_t dataBytes, uint32_t *formatOut,
                                                  uint32_t *mipLevelCountOut)
{
    REQUIRE(c && rgba8 && dataOut && formatOut && mipLevelCountOut, "prt_compress_texture_static");
    REQUIRE(width > 0 && height > 0 && width <= 16384 && height <= 16384, "texture extent out of range");
    REQUIRE((uint64_t)width * height <= PRT_STATIC_MAX_TEXELS,
            "prt_compress_texture_static takes at most PRT_STATIC_MAX_TEXELS (1024 x 1024) texels in level 0; larger images go "
            "through prt_compress_texture");
    REQUIRE(colorSpace <= 1, "colorSpace is 0 (sRGB) or 1 (linear), TextureColorSpace in Texture.hpp");
    const uint32_t full = generateMipMaps ? mipCountFor(std::max(width, height)) : 1u;
    const uint32_t levels = full > 2 ? full - 2 : 1u;
    bool bc7 = true;
    for (uint32_t i = 0; i < levels; ++i)
        if ((std::max(width >> i, 1u) % 4) != 0 || (std::max(height >> i, 1u) % 4) != 0) bc7 = false;
    uint64_t rawOff[16], outOff[16]; // 16384 has 15 levels, 13 kept
    uint64_t rawTexels = 0, outBytes = 0;
    for (uint32_t i = 0; i < levels; ++i)
    {
        const uint64_t w = std::max(width >> i, 1u), h = std::max(height >> i, 1u);
        rawOff[i] = rawTexels;
        outOff[i] = outBytes;
        rawTexels += w * h;
        outBytes += bc7 ? (w / 4) * (h / 4) * 16 : w * h * 4;
    }
    REQUIRE(dataBytes >= outBytes, "destination too small for the level chain");
    const uint64_t rawBytes = (rawTexels * 4 + 255) & ~(uint64_t)255;
    REQUIRE(STATIC_TABLE_BYTES + rawBytes + (bc7 ? outBytes : 0) <= PRT_STATIC_SCRATCH_BYTES,
            "level chain larger than the static scratch array (PRT_STATIC_SCRATCH_BYTES)");
    CK(cudaSetDevice(c->cfg.device));
    unsigned char *scratch = nullptr;
    CKS(staticScratchFor(c->cfg.device, &scratch));
    const float *tp = (const float *)scratch;
    uint32_t *raw = (uint32_t *)(scratch + STATIC_TABLE_BYTES);
    uint2 *blocks = (uint2 *)(scratch + STATIC_TABLE_BYTES + rawBytes);
    CK(cudaMemcpyAsync(raw, rgba8, (size_t)width * height * 4, cudaMemcpyHostToDevice, c->stream));
    for (uint32_t i = 0; i < levels; ++i)
    {
        const uint32_t w = std::max(width >> i, 1u), h = std::max(height >> i, 1u);
        if (i > 0)
        {
            const uint32_t pw = std::max(width >> (i - 1), 1u), ph = std::max(height >> (i - 1), 1u);
            const bool srgb = colorSpace == 0;
            const MipTables tb{srgb ? tp : tp + 256, tp + 256, srgb ? tp + 512 : tp + 768, tp + 768, tp + 1024};
            k_downsample_mitchell<<<blocksFor((size_t)w * h, 128), 128, 0, c->stream>>>(raw + rawOff[i - 1], (int)pw, (int)ph, raw + rawOff[i],
                                                                                       (int)w, (int)h, tb);
            c->kernelLaunches++;
        }
        if (bc7)
        {
            const size_t nBlocks = (size_t)(w / 4) * (h / 4);
            k_bc7_encode<<<blocksFor(nBlocks, 64), 64, 0, c->stream>>>(raw + rawOff[i], w, h, blocks + outOff[i] / 8);
            c->kernelLaunches++;
        }
    }
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(dataOut, bc7 ? (const void *)blocks : (const void *)raw, outBytes, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    *formatOut = bc7 ? PRT_TEX_BC7_UNORM : PRT_TEX_RGBA8_UNORM;
    *mipLevelCountOut = levels;
    return PRT_OK;
}

extern "C" prt_status prt_set_sky_cubemap(prt_ctx *c, uint32_t faceSize, const void *rgba16f)
{
    REQUIRE(c, "prt_set_sky_cubemap");
    CK(cudaSetDevice(c->cfg.device));
    if (faceSize == 0 || rgba16f == nullptr)
    {
        c->skyFace = 0;
        return PRT_OK;
    }
    const size_t n = (size_t)6 * faceSize * faceSize * 4;
    CK(c->dSky.ensure(n));
    CK(cudaMemcpyAsync(c->dSky.p, rgba16f, n * 2, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    c->skyFace = faceSize;
    return PRT_OK;
}

extern "C" prt_status prt_set_models(prt_ctx *c, const uint32_t *offsets, const prt_sub_model *subModels, uint32_t modelCount)
{
    REQUIRE(c && offsets, "prt_set_models");
    c->modelOffsets.assign(offsets, offsets + modelCount + 1);
    const uint32_t n = offsets[modelCount];
    REQUIRE(n == 0 || subModels, "prt_set_models: sub-models missing");
    c->subModels.assign(subModels, subModels + n);
    if (c->blas.size() < modelCount) c->blas.resize(modelCount);
    return PRT_OK;
}

// (texture, sampler) -> CUDA texture objects; rebuilt lazily when textures or samplers changed
static prt_status refreshTextureSlots(prt_ctx *c)
{
    if (!c->slotsDirty) return PRT_OK;
    CK(cudaStreamSynchronize(c->stream));
    for (TexSlot &s : c->hSlots)
    {
        if (s.texels) cudaDestroyTextureObject(s.texels);
        if (s.linear) cudaDestroyTextureObject(s.linear);
    }
    std::vector<prt_sampler_desc> samplers = c->hSamplers;
    if (samplers.empty()) samplers.push_back(prt_sampler_desc{PRT_FILTER_LINEAR, PRT_FILTER_LINEAR, PRT_ADDRESS_REPEAT, PRT_ADDRESS_REPEAT});
    const size_t ns = samplers.size(), nt = c->textures.size();
    c->hSlots.assign(nt * ns, TexSlot{});
    auto addr = [](uint32_t m) {
        return m == PRT_ADDRESS_REPEAT ? cudaAddressModeWrap : (m == PRT_ADDRESS_MIRRORED_REPEAT ? cudaAddressModeMirror : cudaAddressModeClamp);
    };
    for (size_t t = 0; t < nt; ++t)
    {
        const TexRes &tr = c->textures[t];
        if (!tr.array) continue;
        for (size_t s = 0; s < ns; ++s)
        {
            TexSlot &slot = c->hSlots[t * ns + s];
            cudaResourceDesc rd{};
            rd.resType = cudaResourceTypeArray;
            rd.res.array.array = tr.array;
            cudaTextureDesc td{};
            td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
            td.filterMode = cudaFilterModePoint;
            td.readMode = cudaReadModeNormalizedFloat;
            td.normalizedCoords = 0;
            CK(cudaCreateTextureObject(&slot.texels, &rd, &td, nullptr));
            td.addressMode[0] = addr(samplers[s].addressU);
            td.addressMode[1] = addr(samplers[s].addressV);
            td.filterMode = samplers[s].magFilter == PRT_FILTER_NEAREST ? cudaFilterModePoint : cudaFilterModeLinear;
            td.normalizedCoords = 1;
            CK(cudaCreateTextureObject(&slot.linear, &rd, &td, nullptr));
            slot.payload = tr.payload;
            slot.alpha = tr.alpha;
            slot.alphaOpaque = tr.alphaOpaque ? 1u : 0u;
            slot.format = tr.format;
            slot.width = tr.width; slot.height = tr.height;
            slot.magFilter = samplers[s].magFilter;
            slot.addressU = samplers[s].addressU; slot.addressV = samplers[s].addressV;
            slot.valid = 1;
        }
    }
    CK(c->dSlots.ensure(std::max<size_t>(c->hSlots.size(), 1)));
    if (!c->hSlots.empty())
        CK(cudaMemcpyAsync(c->dSlots.p, c->hSlots.data(), c->hSlots.size() * sizeof(TexSlot), cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    c->slotsDirty = false;
    c->alphaMatsDirty = true;
    return PRT_OK;
}

// The any-hit stage's per-draw-instance records (AlphaMat), stream-ordered after the uploads they derive from.
static prt_status refreshAlphaMats(prt_ctx *c)
{
    if (!c->alphaMatsDirty) return PRT_OK;
    if (c->drawCount && !c->hMat.empty())
    {
        CK(c->dAlphaMats.ensure(c->drawCount));
        SceneView sv = c->view();
        k_alpha_mats<<<blocksFor(c->drawCount, 128), 128, 0, c->stream>>>(sv, c->drawCount, c->dAlphaMats.p);
        c->kernelLaunches++;
        CK(cudaGetLastError());
        c->alphaMatsDirty = false;
    }
    return PRT_OK;
}

// ---- acceleration structures -------------------------------------------------------------------------------
// Shared back half of the BLAS and TLAS builds: boxes + bounds are in bBoxes/bBounds; produces the
// wide nodes in bNodes and the node-order primitive permutation in bPrimOut.
// bScalars layout: [0] nodeCount, [1] primCount, [8 .. 8 + MAX_LEVELS] per-level queue counts.
// Collapse level budget. It cannot run out: a binary tree from the SAH builder is at most SAH_MAX_LEVELS = 96 deep, one from the
// Karras builder at most 63 key bits + 32 position bits, and the wide tree is never deeper than the binary one. (PLOC gives no
// such bound; prt_build_blas checks every build's primitive count after the collapse.)
static constexpr uint32_t kMaxLevels = 128;
#ifndef PRT_PLOC_RADIUS
#define PRT_PLOC_RADIUS 16
#endif

static prt_status buildHierarchy(prt_ctx *c, uint32_t n, bool tlas, WideNode *nodesOut, uint32_t *launches)
{
    cudaStream_t st = c->stream;
    for (int i = 0; i < 2; ++i)
    {
        CK(c->bKeys[i].ensure(n));
        CK(c->bOrder[i].ensure(n));
        CK(c->bQueue[i].ensure(n + 1));
    }
    const uint32_t sortBlocks = (n + SORT_TILE - 1) / SORT_TILE;
    CK(c->bHist.ensure(256u * sortBlocks));
    CK(c->bLeft.ensure(n)); CK(c->bRight.ensure(n)); CK(c->bFirst.ensure(n)); CK(c->bLast.ensure(n));
    CK(c->bParentI.ensure(n)); CK(c->bParentL.ensure(n)); CK(c->bVisit.ensure(n)); CK(c->bNodeBox.ensure(n));
    CK(c->bPrimOut.ensure(n)); CK(c->bNodes.ensure(n));

    // TLAS keys keep the top 31 Morton bits (4 sort passes); BLAS keys all 63 (8 passes)
    const uint64_t keyMask = tlas ? 0xFFFFFFFF00000000ull : ~0ull;
    k_pad_and_keys<<<blocksFor(n, 256), 256, 0, st>>>(c->bBoxes.p, n, c->bBounds.p, tlas ? 0 : 1, keyMask, c->bKeys[0].p, c->bOrder[0].p);
    *launches += 1;
    const bool sah = n > 1 && !tlas && !(c->cfg.flags & (PRT_CFG_PLOC | PRT_CFG_LBVH_BLAS));
    const bool flipped = sah ? false : radixSortPairs(c->bKeys[0].p, c->bOrder[0].p, c->bKeys[1].p, c->bOrder[1].p, n, tlas ? 32 : 0, c->bHist.p, st, launches);
    CK(cudaGetLastError());

    Lbvh t{};
    t.count = n;
    t.keys = flipped ? c->bKeys[1].p : c->bKeys[0].p;
    t.order = flipped ? c->bOrder[1].p : c->bOrder[0].p;
    t.primBox = c->bBoxes.p;
    t.left = c->bLeft.p; t.right = c->bRight.p; t.first = c->bFirst.p; t.last = c->bLast.p;
    t.parentOfInternal = c->bParentI.p; t.parentOfLeaf = c->bParentL.p;
    t.nodeBox = c->bNodeBox.p; t.visit = c->bVisit.p;
    uint32_t *scal = c->bScalars.p;
    if (sah)
    {
        // BLAS (default): top-down binned SAH, every level inside ONE cooperative launch (k_sah_all_levels). Scalars: [243] the
        // root's gap, [244] levels run (SAH_MAX_LEVELS + 1: budget exhausted, checked by prt_build_blas with its own
        // read-back), [256 .. 256 + SAH_MAX_LEVELS] ranges per level.
        REQUIRE(c->sahBlocks > 0, "this device cannot launch the cooperative SAH builder");
        const uint32_t keys = n / 2u + 1u;
        for (int i = 0; i < 2; ++i) { CK(c->bSahKey[i].ensure(n)); CK(c->bSahRange[i].ensure(keys)); CK(c->bSahQueue[i].ensure(keys)); }
        CK(c->bSahCent.ensure((size_t)SAH_CENT * keys)); CK(c->bSahSplit.ensure(keys)); CK(c->bSahBins.ensure((size_t)SAH_SLOT_BINS * n));
        Sah s{};
        s.t = t;
        s.split = c->bSahSplit.p; s.bins = c->bSahBins.p; s.cent = c->bSahCent.p;
        s.active = scal + 256; s.rootGap = scal + 243;
        SahGrid g{};
        for (int i = 0; i < 2; ++i) { g.idx[i] = c->bOrder[i].p; g.key[i] = c->bSahKey[i].p; g.queue[i] = c->bSahQueue[i].p; g.range[i] = c->bSahRange[i].p; }
        g.keys = keys;
        g.levelsOut = scal + 244;
        void *args[] = {&s, &g};
        const uint32_t blocks = std::min<uint32_t>(c->sahBlocks, blocksFor(n, SAH_THREADS));
        CK(cudaLaunchCooperativeKernel((const void *)k_sah_all_levels, dim3(blocks), dim3(SAH_THREADS), args, 0, st));
        *launches += 1;
        t.order = c->bOrder[0].p;
    }
    else if (n > 1 && !tlas && (c->cfg.flags & PRT_CFG_PLOC))
    {
        // BLAS: PLOC (surface-area-driven topology). Scalars: [240], [241] live cluster counts (ping-pong), [242] nodes made.
        for (int i = 0; i < 2; ++i) { CK(c->bPlocRef[i].ensure(n)); CK(c->bPlocBox[i].ensure(n)); }
        CK(c->bPlocMergedRef.ensure(n)); CK(c->bPlocMergedBox.ensure(n)); CK(c->bPlocNearest.ensure(n)); CK(c->bNewPos.ensure(n));
        CK(c->bPlocBlocks.ensure(blocksFor(n, PLOC_THREADS)));
        uint32_t *counts = scal + 240, *nodeAlloc = scal + 242;
        k_ploc_init<<<blocksFor(n, 256), 256, 0, st>>>(t, c->bPlocRef[0].p, c->bPlocBox[0].p, counts, nodeAlloc);
        *launches += 1;
        Ploc p{};
        p.t = t;
        p.nearest = c->bPlocNearest.p; p.mergedRef = c->bPlocMergedRef.p; p.mergedBox = c->bPlocMergedBox.p;
        p.nodeAlloc = nodeAlloc;
        p.radius = PRT_PLOC_RADIUS;
        uint32_t bound = n; // host-side upper bound of the live cluster count
        bool done = false;
        for (uint32_t it = 0; it < 4096u && !done; ++it)
        {
            const int a = (int)(it & 1u), b = a ^ 1;
            p.refIn = c->bPlocRef[a].p; p.boxIn = c->bPlocBox[a].p; p.refOut = c->bPlocRef[b].p; p.boxOut = c->bPlocBox[b].p;
            const uint32_t grid = blocksFor(bound, PLOC_THREADS);
            k_ploc_nearest<<<grid, PLOC_THREADS, 0, st>>>(p, counts + a);
            k_ploc_merge<<<grid, PLOC_THREADS, 0, st>>>(p, counts + a, c->bPlocBlocks.p);
            k_ploc_offsets<<<1, 1024, 0, st>>>(counts + a, c->bPlocBlocks.p, counts + b);
            k_ploc_scatter<<<grid, PLOC_THREADS, 0, st>>>(p, counts + a, c->bPlocBlocks.p);
            *launches += 4;
            if ((it & 3u) == 3u)
            {
                uint32_t live = 0;
                CK(cudaMemcpyAsync(&live, counts + b, 4, cudaMemcpyDeviceToHost, st));
                CK(cudaStreamSynchronize(st));
                REQUIRE(live >= 1 && live <= bound, "PLOC cluster count out of range");
                bound = live;
                done = live == 1;
            }
        }
        REQUIRE(done, "PLOC did not converge");
        uint32_t *orderOut = flipped ? c->bOrder[0].p : c->bOrder[1].p;
        k_ploc_finish_leaves<<<blocksFor(n, 256), 256, 0, st>>>(t, c->bNewPos.p, orderOut);
        k_ploc_finish_nodes<<<blocksFor(n - 1, 256), 256, 0, st>>>(t, c->bNewPos.p);
        *launches += 2;
        t.order = orderOut;
    }
    else if (n > 1)
    {
        k_lbvh_nodes<<<blocksFor(n - 1, 256), 256, 0, st>>>(t);
        k_lbvh_refit<<<blocksFor(n, 256), 256, 0, st>>>(t);
        *launches += 2;
    }
    k_collapse_seed<<<1, 128, 0, st>>>(c->bQueue[0].p, scal + 8, kMaxLevels + 1, scal + 0, scal + 1, n);
    *launches += 1;
    Collapse col{};
    col.t = t;
#ifndef PRT_BLAS_MAX_LEAF
#define PRT_BLAS_MAX_LEAF 3
#endif
    col.maxLeaf = tlas ? 1u : (uint32_t)PRT_BLAS_MAX_LEAF;
    col.nodes = nodesOut;
    col.primOut = c->bPrimOut.p;
    col.nodeCount = scal + 0;
    col.primCount = scal + 1;
    if (tlas && n <= 512u)
    {
        k_collapse_all_levels<<<1, 1024, 0, st>>>(col, c->bQueue[0].p, c->bQueue[1].p, scal + 8, kMaxLevels);
        *launches += 1;
    }
    else
    {
        // one cooperative launch for all levels, eight lanes per wide node: enough blocks for the widest level, never more
        // than are co-resident
        REQUIRE(c->coopBlocks > 0, "this device cannot launch the cooperative collapse");
        CollapseItem *qa = c->bQueue[0].p, *qb = c->bQueue[1].p;
        uint32_t *counts = scal + 8;
        uint32_t maxLevels = kMaxLevels;
        void *args[] = {&col, &qa, &qb, &counts, &maxLevels};
        const uint32_t blocks = std::min<uint32_t>(c->coopBlocks, blocksFor((uint64_t)n * 2u, 128));
        CK(cudaLaunchCooperativeKernel((const void *)k_collapse_all_levels_grid, dim3(blocks), dim3(128), args, 0, st));
        *launches += 1;
    }
    CK(cudaGetLastError());
    return PRT_OK;
}

// World::Impl::buildNextBlas (World.cpp:585-747): one BLAS per model, one geometry per sub-model,
// OPAQUE flag from the mesh's material alpha mode (World.cpp:646-651).
extern "C" prt_status prt_build_blas(prt_ctx *c, uint32_t modelIndex, uint64_t *outReference)
{
    REQUIRE(c, "prt_build_blas");
    REQUIRE(modelIndex + 1 < c->modelOffsets.size(), "prt_build_blas: model index out of range (call prt_set_models first)");
    CK(cudaSetDevice(c->cfg.device));
    cudaStream_t st = c->stream;
    BlasRes &b = c->blas[modelIndex];
    b.built = false; // until this build has succeeded; the buffers of an earlier build are kept for reuse
    uint64_t total = 0, unique = 0;
    for (uint32_t g = c->modelOffsets[modelIndex]; g < c->modelOffsets[modelIndex + 1]; ++g)
    {
        const prt_sub_model &sm = c->subModels[g];
        REQUIRE(sm.meshIndex < c->hInfo.size(), "sub-model mesh index out of range");
        const prt_mesh_info &info = c->hInfo[sm.meshIndex];
        REQUIRE(info.materialIndex < c->hMat.size(), "mesh material index out of range");
        // builder references: a non-opaque triangle counts blasRefsPerTriangle times (core_build.h)
        total += (uint64_t)(info.indexCount / 3) * blasRefsPerTriangle(c->hMat[info.materialIndex].alphaMode == PRT_ALPHA_OPAQUE ? 1u : 0u);
        unique += info.indexCount / 3;
    }
    REQUIRE(total < (1ull << 31), "model too large");
    if (outReference) *outReference = (uint64_t)modelIndex + 1;
    if (total == 0)
    {
        b.nodeCount = b.triCount = b.primCount = 0;
        b.built = true;
        return PRT_OK;
    }
    const uint32_t n = (uint32_t)total;
    CK(cudaEventRecord(c->evBlas[0], st));
    uint32_t launches = 0;
    CK(c->bTris.ensure(n));
    CK(c->bUvs.ensure(n));
    CK(c->bBoxes.ensure(n));
    k_init_bounds<<<1, 32, 0, st>>>(c->bBounds.p, c->bScalars.p, 8);
    const SceneView sv = c->view();
    uint32_t offset = 0;
    for (uint32_t g = c->modelOffsets[modelIndex]; g < c->modelOffsets[modelIndex + 1]; ++g)
    {
        const prt_sub_model &sm = c->subModels[g];
        const prt_mesh_info &info = c->hInfo[sm.meshIndex];
        const uint32_t prims = info.indexCount / 3;
        if (prims == 0) continue;
        REQUIRE(info.materialIndex < c->hMat.size(), "mesh material index out of range");
        const uint32_t opaque = c->hMat[info.materialIndex].alphaMode == PRT_ALPHA_OPAQUE ? 1u : 0u;
        k_blas_gather<<<blocksFor(prims, 128), 128, 0, st>>>(sv, sm.meshIndex, prims, g - c->modelOffsets[modelIndex], opaque, offset,
                                                             c->bTris.p, c->bUvs.p, c->bBoxes.p, c->bBounds.p);
        launches++;
        offset += prims * blasRefsPerTriangle(opaque);
    }
    CK(cudaGetLastError());
    CK(c->bNodes.ensure(n));
    CKS(buildHierarchy(c, n, false, c->bNodes.p, &launches));
    uint32_t counts[2] = {0, 0}, sahLevels = 0;
    CK(cudaMemcpyAsync(counts, c->bScalars.p, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&sahLevels, c->bScalars.p + 244, 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (n > 1 && !(c->cfg.flags & (PRT_CFG_PLOC | PRT_CFG_LBVH_BLAS)))
        REQUIRE(sahLevels <= SAH_MAX_LEVELS, "binned SAH did not finish within its level budget");
    REQUIRE(counts[1] == n, "collapse lost primitives");
    b.nodeCount = counts[0];
    b.triCount = n;
    b.primCount = (uint32_t)unique;
    if (b.nodeCap < b.nodeCount) // first build of the model, or it grew: the stream is idle after the read-back above
    {
        if (b.nodes) c->heap.release(b.nodes);
        b.nodes = nullptr;
        b.nodeCap = 0;
        const uint32_t cap = b.nodeCount + b.nodeCount / 8u;
        CKS(c->heap.alloc((void **)&b.nodes, (size_t)cap * sizeof(WideNode)));
        b.nodeCap = cap;
    }
    if (b.triCap < n)
    {
        if (b.tris) c->heap.release(b.tris);
        if (b.uvs) CK(cudaFree(b.uvs));
        b.tris = nullptr; b.uvs = nullptr;
        b.triCap = 0;
        CKS(c->heap.alloc((void **)&b.tris, (size_t)n * sizeof(TriRec)));
        CK(cudaMalloc((void **)&b.uvs, (size_t)n * sizeof(AlphaUv)));
        b.triCap = n;
    }
    CK(cudaMemcpyAsync(b.nodes, c->bNodes.p, (size_t)b.nodeCount * sizeof(WideNode), cudaMemcpyDeviceToDevice, st));
    k_emit_tris<<<blocksFor(n, 256), 256, 0, st>>>(c->bPrimOut.p, n, c->bTris.p, b.tris, c->bUvs.p, b.uvs);
    launches++;
    CK(cudaGetLastError());
    CK(cudaEventRecord(c->evBlas[1], st));
    c->blasTimed = true;
    c->kernelLaunches += launches;
    b.built = true; // only a build that got this far is referable from prt_build_tlas
    c->blasTableDirty = true;
    return PRT_OK;
}

// The device side of a TLAS rebuild for the `n` records already uploaded (bTlasRaw: the caller's own records, all active;
// bTlasIn: the compacted ones): instance records + boxes, hierarchy, emit.
static prt_status enqueueTlasBuild(prt_ctx *c, uint32_t n, bool direct, uint32_t *launches)
{
    cudaStream_t st = c->stream;
    CK(c->bInst.ensure(n));
    CK(c->bBoxes.ensure(n));
    CK(c->tlasInst.ensure(n));
    k_init_bounds<<<1, 32, 0, st>>>(c->bBounds.p, c->bScalars.p, 8);
    if (direct) k_tlas_instances_direct<<<blocksFor(n, 128), 128, 0, st>>>(c->bTlasRaw.p, c->dBlasTable.p, n, c->bInst.p, c->bBoxes.p, c->bBounds.p);
    else k_tlas_instances<<<blocksFor(n, 128), 128, 0, st>>>(c->bTlasIn.p, n, c->bInst.p, c->bBoxes.p, c->bBounds.p);
    *launches += 2;
    CKS(buildHierarchy(c, n, true, c->tlasNodesP, launches)); // wide nodes <= n: the collapse writes them in place
    k_emit_instances<<<blocksFor(n, 128), 128, 0, st>>>(c->bPrimOut.p, n, c->bInst.p, c->tlasInst.p);
    *launches += 1;
    CK(cudaGetLastError());
    return PRT_OK;
}

// updateTlasInstances + buildCurrentTlas (World.cpp:878-928, 749-802): full rebuild, every frame.
extern "C" prt_status prt_build_tlas(prt_ctx *c, const prt_tlas_instance *instances, uint32_t count)
{
    REQUIRE(c && (count == 0 || instances), "prt_build_tlas");
    if (c->cfg.maxInstances) REQUIRE(count <= c->cfg.maxInstances, "more instances than prt_config.maxInstances");
    CK(cudaSetDevice(c->cfg.device));
    cudaStream_t st = c->stream;
    // Every instance active (the usual frame): the records go up as they are and k_tlas_instances_direct resolves the BLAS
    // pointers from a device table. Otherwise (inactive instances, World.cpp:898-905, or empty models) the host compacts.
    bool direct = count > 0;
    for (uint32_t i = 0; i < count; ++i)
    {
        const uint64_t ref = instances[i].blasReference;
        if (ref == 0) { direct = false; continue; }
        REQUIRE(ref - 1 < c->blas.size() && c->blas[ref - 1].built, "instance references a BLAS that was not built");
        if (c->blas[ref - 1].triCount == 0) direct = false;
    }
    {
        // the arena of the level references a ray carries (Accel::arena): the lowest address of any node / triangle array
        uint64_t lo = ~0ull, hi = 0;
        auto span = [&](const void *p, size_t bytes) { if (p) { lo = std::min<uint64_t>(lo, (uint64_t)(uintptr_t)p); hi = std::max<uint64_t>(hi, (uint64_t)(uintptr_t)p + bytes); } };
        if (c->tlasNodesCap < std::max<uint32_t>(count, 1u)) // grows rarely: the instance count of a scene is stable
        {
            CK(cudaStreamSynchronize(st));
            if (c->tlasNodesP) c->heap.release(c->tlasNodesP);
            c->tlasNodesP = nullptr;
            c->tlasNodesCap = 0;
            const uint32_t cap = std::max<uint32_t>(count, 1u) + count / 8u;
            CKS(c->heap.alloc((void **)&c->tlasNodesP, (size_t)cap * sizeof(WideNode)));
            c->tlasNodesCap = cap;
            g_allocGeneration++; // a captured TLAS graph holds the old address
        }
        span(c->tlasNodesP, (size_t)c->tlasNodesCap * sizeof(WideNode));
        for (const BlasRes &b : c->blas) { span(b.nodes, (size_t)b.nodeCap * sizeof(WideNode)); span(b.tris, (size_t)b.triCap * sizeof(TriRec)); }
        lo &= ~127ull;
        REQUIRE(hi - lo < (1ull << 39), "acceleration structures spread over more than 512 GB of address space");
        c->accelArena = lo;
    }
    std::vector<TlasInput> &in = c->hTlasInput;
    in.clear();
    for (uint32_t i = 0; i < count && !direct; ++i)
    {
        const prt_tlas_instance &ti = instances[i];
        if (ti.blasReference == 0) continue; // inactive instance, World.cpp:898-905
        const BlasRes &b = c->blas[ti.blasReference - 1];
        if (b.triCount == 0) continue;
        TlasInput t;
        memcpy(t.transform, ti.transform, sizeof(t.transform));
        t.customIndexAndMask = ti.customIndexAndMask;
        t.index = i;
        t.nodes = b.nodes;
        t.tris = b.tris;
        t.uvs = b.uvs;
        in.push_back(t);
    }
    const uint32_t n = direct ? count : (uint32_t)in.size();
    CK(cudaEventRecord(c->evTlas[0], st));
    uint32_t launches = 0;
    c->tlasInstances = n;
    if (n)
    {
        if (direct)
        {
            if (c->blasTableDirty)
            {
                std::vector<BlasTableEntry> table(c->blas.size());
                for (size_t m = 0; m < table.size(); ++m) table[m] = BlasTableEntry{c->blas[m].nodes, c->blas[m].tris, c->blas[m].uvs};
                CK(c->dBlasTable.ensure(table.size()));
                CK(c->ring.upload(c->dBlasTable.p, table.data(), table.size() * sizeof(BlasTableEntry), st));
                c->blasTableDirty = false;
            }
            CK(c->bTlasRaw.ensure(n));
            CK(c->ring.upload(c->bTlasRaw.p, instances, (size_t)n * sizeof(prt_tlas_instance), st));
        }
        else
        {
            CK(c->bTlasIn.ensure(n));
            CK(c->ring.upload(c->bTlasIn.p, in.data(), (size_t)n * sizeof(TlasInput), st));
        }
        // The rebuild is ~16 short dependent launches: from the second frame with the same instance count on, they are
        // replayed as one CUDA graph (captured on the first repeat, re-captured when the count or any scratch buffer moved).
        if (c->tlasGraphExec && c->tlasGraphN == n && c->tlasGraphDirect == direct && c->tlasGraphGeneration == g_allocGeneration)
        {
            CK(cudaGraphLaunch(c->tlasGraphExec, st));
            launches = c->tlasGraphLaunches;
            c->tlasGraphReplays++;
        }
        else if (c->tlasGraphs && c->tlasEagerN == n && c->tlasEagerDirect == direct && c->tlasEagerGeneration == g_allocGeneration)
        {
            if (c->tlasGraphExec) { CK(cudaGraphExecDestroy(c->tlasGraphExec)); c->tlasGraphExec = nullptr; }
            CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
            const prt_status s = enqueueTlasBuild(c, n, direct, &launches); // no allocation: the eager build of this size came first
            cudaGraph_t graph = nullptr;
            const cudaError_t e = cudaStreamEndCapture(st, &graph);
            cudaError_t ei = e;
            if (s == PRT_OK && e == cudaSuccess) ei = cudaGraphInstantiate(&c->tlasGraphExec, graph, 0);
            if (graph) cudaGraphDestroy(graph);
            if (s != PRT_OK || ei != cudaSuccess)
            {
                // a driver that cannot capture this sequence: say so once, keep the eager path from here on
                cudaGetLastError();
                c->tlasGraphExec = nullptr;
                c->tlasGraphs = false;
                fprintf(stderr, "prt: TLAS graph capture unavailable (%s); rebuilding with plain launches\n",
                        s != PRT_OK ? g_lastError.c_str() : cudaGetErrorString(ei));
                launches = 0;
                CKS(enqueueTlasBuild(c, n, direct, &launches));
            }
            else
            {
                c->tlasGraphN = n;
                c->tlasGraphDirect = direct;
                c->tlasGraphGeneration = g_allocGeneration;
                c->tlasGraphLaunches = launches;
                c->tlasGraphCaptures++;
                CK(cudaGraphLaunch(c->tlasGraphExec, st));
            }
        }
        else
        {
            CKS(enqueueTlasBuild(c, n, direct, &launches));
            c->tlasEagerN = n;
            c->tlasEagerDirect = direct;
            c->tlasEagerGeneration = g_allocGeneration; // read after the build: its own allocations are included
        }
    }
    CK(cudaEventRecord(c->evTlas[1], st));
    c->tlasTimed = true;
    c->lastTlasLaunches = launches;
    c->kernelLaunches += launches;
    return PRT_OK;
}

// ---- per-frame rings: World::updateBuffers (World.cpp:468-536), Camera::updateBuffer --------------------------
extern "C" prt_status prt_set_draw_instances(prt_ctx *c, const prt_draw_instance *instances, uint32_t count)
{
    REQUIRE(c && (count == 0 || instances), "prt_set_draw_instances");
    CK(cudaSetDevice(c->cfg.device));
    CK(c->dDraw.ensure(count));
    CK(c->ring.upload(c->dDraw.p, instances, (size_t)count * sizeof(prt_draw_instance), c->stream));
    c->drawCount = count;
    c->alphaMatsDirty = true;
    return PRT_OK;
}

extern "C" prt_status prt_set_instance_transforms(prt_ctx *c, const prt_model_instance_transforms *t, uint32_t count)
{
    REQUIRE(c && (count == 0 || t), "prt_set_instance_transforms");
    CK(cudaSetDevice(c->cfg.device));
    CK(c->dXf.ensure(count));
    CK(c->ring.upload(c->dXf.p, t, (size_t)count * sizeof(prt_model_instance_transforms), c->stream));
    return PRT_OK;
}

extern "C" prt_status prt_set_lights(prt_ctx *c, const prt_directional_light *directional, const prt_point_light *points,
                                     uint32_t pointCount, const prt_spot_light *spots, uint32_t spotCount)
{
    REQUIRE(c && directional, "prt_set_lights");
    REQUIRE(pointCount <= PRT_MAX_POINT_LIGHTS && spotCount <= PRT_MAX_SPOT_LIGHTS, "light maxima (RtRef.cpp:429-430)");
    REQUIRE((pointCount == 0 || points) && (spotCount == 0 || spots), "prt_set_lights: null array");
    CK(cudaSetDevice(c->cfg.device));
    CK(c->ring.upload(c->dDir.p, directional, sizeof(prt_directional_light), c->stream));
    CK(c->ring.upload(c->dPoints.p, points, (size_t)pointCount * sizeof(prt_point_light), c->stream));
    CK(c->ring.upload(c->dSpots.p, spots, (size_t)spotCount * sizeof(prt_spot_light), c->stream));
    c->pointCount = pointCount;
    c->spotCount = spotCount;
    c->haveLights = true;
    return PRT_OK;
}

extern "C" prt_status prt_set_camera(prt_ctx *c, const prt_camera_uniforms *camera)
{
    REQUIRE(c && camera, "prt_set_camera");
    CK(cudaSetDevice(c->cfg.device));
    CK(c->ring.upload(c->dCam.p, camera, sizeof(prt_camera_uniforms), c->stream));
    c->haveCamera = true;
    return PRT_OK;
}

// ---- the pass: RtReference::record (RtRef.cpp:161-383) ------------------------------------------------------------
static prt_status checkRenderState(prt_ctx *c, const prt_reference_pc *pc, uint32_t width, uint32_t height)
{
    REQUIRE(pc, "null push constants");
    REQUIRE(width > 0 && height > 0, "empty render area");
    REQUIRE(c->haveCamera, "prt_set_camera has not been called");
    REQUIRE(c->haveLights, "prt_set_lights has not been called");
    REQUIRE(pc->maxBounces <= PRT_MAX_BOUNCES, "maxBounces above sMaxBounces (RtRef.hpp:22)");
    REQUIRE(pc->drawType < PRT_DRAW_COUNT, "unknown draw type");
    CKS(refreshTextureSlots(c));
    CKS(refreshAlphaMats(c));
    return PRT_OK;
}

// Counting sort of a ray queue's indices by (octant, origin cell): fills c->sortPerm. The queue length lives on the device.
static void sortRays(prt_ctx *c, const Accel &as, const float4 *rayO, const float4 *rayD, const uint32_t *countPtr, uint32_t grid,
                     uint32_t &launches)
{
    cudaStream_t st = c->stream;
    cudaMemsetAsync(c->sortHist.p, 0, RAY_SORT_BINS * sizeof(uint32_t), st);
    k_ray_bin<<<grid, RENDER_THREADS, 0, st>>>(as, rayO, rayD, countPtr, c->sortKeys.p, c->sortHist.p);
    k_ray_scan<<<1, RAY_SCAN_THREADS, 0, st>>>(c->sortHist.p);
    k_ray_scatter<<<grid, RENDER_THREADS, 0, st>>>(countPtr, c->sortKeys.p, c->sortHist.p, c->sortPerm.p);
    launches += 3;
}

// Enqueues `samples` frames starting at pc.frameIndex as one wavefront batch.
static prt_status renderBatch(prt_ctx *c, prt_reference_pc pc, uint32_t width, uint32_t height, uint32_t samples, bool accumulate,
                              bool timeTrace)
{
    cudaStream_t st = c->stream;
    const uint64_t total = (uint64_t)width * height * samples;
    REQUIRE(total < (1ull << 32), "width x height x samplesPerLaunch must stay below 2^32 paths (32-bit path indices)");
    CKS(allocPathBuffers(c, total));
    const SceneView sv = c->view();
    const Accel as = c->accel();
    const bool count = (c->cfg.flags & PRT_CFG_COUNTERS) != 0;
    const uint32_t grid = blocksFor(total, RENDER_THREADS);
    const uint32_t tgrid = std::min(grid, c->traceBlocks);
    uint32_t launches = 0;
    k_raygen<<<grid, RENDER_THREADS, 0, st>>>(sv, pc, width, height, samples, c->pb);
    launches++;
    const uint32_t bounces = std::min<uint32_t>(pc.maxBounces, PRT_MAX_BOUNCES);
    const bool debugView = pc.drawType != PRT_DRAW_DEFAULT && pc.drawType != PRT_DRAW_MESHLET_ID;
    for (uint32_t b = 0; b < bounces; ++b)
    {
        const int q = (int)(b & 1u);
        CK(c->mark(0));
        const uint32_t *perm = nullptr;
        if (b >= 1 && (c->raySort & 1u))
        {
            sortRays(c, as, c->pb.rayO[q], c->pb.rayD[q], c->pb.counts + b, grid, launches);
            perm = c->sortPerm.p;
        }
        launchTrace<false>(c, count, tgrid, as, sv, q, b, perm);
        CK(c->mark(1));
        k_shade<<<grid, RENDER_THREADS, 0, st>>>(sv, pc, c->pb, q, b, c->dCounters, count ? 1 : 0);
        launches += 2;
        if (debugView) break; // main.rgen:259-264: debug views stop at the first hit
        CK(c->mark(2));
        perm = nullptr;
        if ((b >= 1 && (c->raySort & 2u)) || (b == 0 && (c->raySort & 4u)))
        {
            sortRays(c, as, c->pb.shadowO, c->pb.shadowD, c->pb.counts + COUNT_SHADOW_BASE + b, grid, launches);
            perm = c->sortPerm.p;
        }
        launchTrace<true>(c, count, tgrid, as, sv, q, b, perm);
        launches++;
    }
    CK(c->mark(3));
    k_tally<<<1, 32, 0, st>>>(c->pb, c->dCounters, (uint32_t)total);
    launches++;
    if (accumulate)
    {
        k_accumulate<<<blocksFor((uint64_t)width * height, 256), 256, 0, st>>>(c->pb, c->accum.p, width * height, samples, pc.flags);
        launches++;
    }
    CK(cudaGetLastError());
    c->lastRecordLaunches += launches;
    c->kernelLaunches += launches;
    return PRT_OK;
}

extern "C" prt_status prt_record(prt_ctx *c, const prt_reference_pc *pcIn, uint32_t width, uint32_t height, uint32_t sampleCount)
{
    REQUIRE(c, "prt_record");
    CK(cudaSetDevice(c->cfg.device));
    CKS(checkRenderState(c, pcIn, width, height));
    REQUIRE(sampleCount > 0, "sampleCount must be positive");
    const uint64_t pixels = (uint64_t)width * height;
    if (c->cfg.maxWidth && c->cfg.maxHeight) REQUIRE(pixels <= (uint64_t)c->cfg.maxWidth * c->cfg.maxHeight, "render area larger than prt_config");
    prt_reference_pc pc = *pcIn;
    // history is recreated on resize (RtRef.cpp:178-219): the first frame then ignores it
    if (c->accumWidth != width || c->accumHeight != height || !c->historyValid)
    {
        CK(c->accum.ensure(pixels));
        pc.flags |= PRT_FLAG_SKIP_HISTORY;
        c->accumWidth = width;
        c->accumHeight = height;
    }
    CK(cudaEventRecord(c->evRecord[0], c->stream));
    c->lastRecordLaunches = 0;
    c->evUsed = 0;
    const uint32_t perLaunch = std::max(1u, c->cfg.samplesPerLaunch);
    uint32_t done = 0;
    while (done < sampleCount)
    {
        const uint32_t k = std::min(perLaunch, sampleCount - done);
        prt_reference_pc bpc = pc;
        bpc.frameIndex = (pc.frameIndex + done) % PRT_FRAME_PERIOD; // RtRef.cpp:170
        if (done > 0) bpc.flags &= ~(uint32_t)PRT_FLAG_SKIP_HISTORY;
        CKS(renderBatch(c, bpc, width, height, k, true, done == 0));
        done += k;
    }
    CK(cudaEventRecord(c->evRecord[1], c->stream));
    c->recordTimed = true;
    c->historyValid = true;
    return PRT_OK;
}

extern "C" prt_status prt_output_f32_device(prt_ctx *c, const float **out)
{
    REQUIRE(c && out, "prt_output_f32_device");
    REQUIRE(c->historyValid, "nothing recorded yet");
    *out = reinterpret_cast<const float *>(c->accum.p);
    return PRT_OK;
}

extern "C" prt_status prt_read_output_f32(prt_ctx *c, float *dstHost, uint64_t dstBytes)
{
    REQUIRE(c && dstHost, "prt_read_output_f32");
    REQUIRE(c->historyValid, "nothing recorded yet");
    const uint64_t bytes = (uint64_t)c->accumWidth * c->accumHeight * 16;
    REQUIRE(dstBytes >= bytes, "destination too small");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaMemcpyAsync(dstHost, c->accum.p, bytes, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return PRT_OK;
}

extern "C" prt_status prt_read_illumination_f16(prt_ctx *c, void *dstHost, uint64_t dstBytes)
{
    REQUIRE(c && dstHost, "prt_read_illumination_f16");
    REQUIRE(c->historyValid, "nothing recorded yet");
    const uint32_t pixels = c->accumWidth * c->accumHeight;
    REQUIRE(dstBytes >= (uint64_t)pixels * 8, "destination too small");
    CK(cudaSetDevice(c->cfg.device));
    CK(c->f16.ensure(pixels));
    k_blit_f16<<<blocksFor(pixels, 256), 256, 0, c->stream>>>(c->accum.p, c->f16.p, pixels);
    c->kernelLaunches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(dstHost, c->f16.p, (size_t)pixels * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return PRT_OK;
}

extern "C" prt_status prt_set_tone_map_lut(prt_ctx *c, const uint32_t *rgb9e5, uint32_t dim)
{
    REQUIRE(c && rgb9e5, "prt_set_tone_map_lut");
    REQUIRE(dim >= 2 && dim <= 256, "LUT edge must be in [2, 256] (the shipped table is 48)");
    CK(cudaSetDevice(c->cfg.device));
    const size_t n = (size_t)dim * dim * dim;
    CK(c->toneLut.ensure(n));
    CK(cudaMemcpyAsync(c->toneLut.p, rgb9e5, n * 4, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    c->toneLutDim = dim;
    return PRT_OK;
}

static prt_status toneMapLaunch(prt_ctx *c, const prt_tone_map_pc *pc, uint32_t &pixels)
{
    REQUIRE(c->historyValid, "nothing recorded yet");
    REQUIRE(c->toneLutDim != 0, "prt_set_tone_map_lut first (ToneMap::init loads it, ToneMap.cpp:37-39)");
    pixels = c->accumWidth * c->accumHeight;
    CK(cudaSetDevice(c->cfg.device));
    CK(c->ldr.ensure(pixels));
    const ToneLut lut{c->toneLut.p, c->toneLutDim};
    k_tone_map<<<blocksFor(pixels, 256), 256, 0, c->stream>>>(c->accum.p, lut, pc->exposure, pc->contrast, c->ldr.p, pixels);
    c->kernelLaunches++;
    CK(cudaGetLastError());
    return PRT_OK;
}

extern "C" prt_status prt_tone_map(prt_ctx *c, const prt_tone_map_pc *pc, void *dstHost, uint64_t dstBytes)
{
    REQUIRE(c && pc && dstHost, "prt_tone_map");
    uint32_t pixels = 0;
    CKS(toneMapLaunch(c, pc, pixels));
    REQUIRE(dstBytes >= (uint64_t)pixels * 4, "destination too small");
    CK(cudaMemcpyAsync(dstHost, c->ldr.p, (size_t)pixels * 4, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return PRT_OK;
}

extern "C" prt_status prt_tone_map_device(prt_ctx *c, const prt_tone_map_pc *pc, const uint32_t **out)
{
    REQUIRE(c && pc && out, "prt_tone_map_device");
    uint32_t pixels = 0;
    CKS(toneMapLaunch(c, pc, pixels));
    *out = c->ldr.p;
    return PRT_OK;
}

// ---- ReSTIR DI trace pass: render::rtdi::Trace (src/render/rtdi/Trace.cpp:176-362) -------------------------------------------
static prt_status rtdiSize(prt_ctx *c, uint32_t width, uint32_t height)
{
    REQUIRE(width > 0 && height > 0, "empty render area");
    const uint64_t pixels = (uint64_t)width * height;
    if (c->cfg.maxWidth && c->cfg.maxHeight) REQUIRE(pixels <= (uint64_t)c->cfg.maxWidth * c->cfg.maxHeight, "render area larger than prt_config");
    if (c->rtdiWidth != width || c->rtdiHeight != height)
    {
        c->rtdiGbufferValid = c->rtdiReservoirsValid = c->rtdiHistoryValid = false;
        c->rtdiWidth = width;
        c->rtdiHeight = height;
    }
    CK(c->gbAlbedoRoughness.ensure(pixels)); CK(c->gbNormalMetallic.ensure(pixels)); CK(c->gbDepth.ensure(pixels));
    CK(c->rtdiReservoirs.ensure(pixels)); CK(c->rtdiAccum.ensure(pixels));
    CKS(allocPathBuffers(c, pixels));
    return PRT_OK;
}
static GBufferView gbufferView(const prt_ctx *c)
{
    return GBufferView{c->gbAlbedoRoughness.p, c->gbNormalMetallic.p, c->gbDepth.p, c->rtdiWidth, c->rtdiHeight};
}

extern "C" prt_status prt_rtdi_set_gbuffer(prt_ctx *c, uint32_t width, uint32_t height, const uint32_t *albedoRoughness,
                                           const uint32_t *normalMetallic, const float *nonLinearDepth)
{
    REQUIRE(c && albedoRoughness && normalMetallic && nonLinearDepth, "prt_rtdi_set_gbuffer");
    CK(cudaSetDevice(c->cfg.device));
    CKS(rtdiSize(c, width, height));
    const size_t pixels = (size_t)width * height;
    CK(cudaMemcpyAsync(c->gbAlbedoRoughness.p, albedoRoughness, pixels * 4, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->gbNormalMetallic.p, normalMetallic, pixels * 4, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->gbDepth.p, nonLinearDepth, pixels * 4, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    c->rtdiGbufferValid = true;
    return PRT_OK;
}

extern "C" prt_status prt_rtdi_gbuffer_from_primary(prt_ctx *c, uint32_t width, uint32_t height, uint32_t frameIndex)
{
    REQUIRE(c, "prt_rtdi_gbuffer_from_primary");
    CK(cudaSetDevice(c->cfg.device));
    prt_reference_pc pc{};
    pc.maxBounces = 1;
    CKS(checkRenderState(c, &pc, width, height));
    CKS(rtdiSize(c, width, height));
    const uint32_t pixels = width * height;
    const SceneView sv = c->view();
    const Accel as = c->accel();
    const uint32_t grid = blocksFor(pixels, RENDER_THREADS);
    k_rtdi_primary_rays<<<grid, RENDER_THREADS, 0, c->stream>>>(sv, width, height, frameIndex, c->pb);
    launchTrace<false>(c, (c->cfg.flags & PRT_CFG_COUNTERS) != 0, std::min(grid, c->traceBlocks), as, sv, 0, 0, nullptr);
    k_rtdi_gbuffer<<<grid, RENDER_THREADS, 0, c->stream>>>(sv, c->pb, width, height, c->gbAlbedoRoughness.p, c->gbNormalMetallic.p, c->gbDepth.p);
    c->kernelLaunches += 3;
    CK(cudaGetLastError());
    c->rtdiGbufferValid = true;
    return PRT_OK;
}

extern "C" prt_status prt_rtdi_read_gbuffer(prt_ctx *c, uint32_t *albedoRoughness, uint32_t *normalMetallic, float *nonLinearDepth)
{
    REQUIRE(c && albedoRoughness && normalMetallic && nonLinearDepth, "prt_rtdi_read_gbuffer");
    REQUIRE(c->rtdiGbufferValid, "no G-buffer yet (prt_rtdi_set_gbuffer / prt_rtdi_gbuffer_from_primary)");
    CK(cudaSetDevice(c->cfg.device));
    const size_t pixels = (size_t)c->rtdiWidth * c->rtdiHeight;
    CK(cudaMemcpyAsync(albedoRoughness, c->gbAlbedoRoughness.p, pixels * 4, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(normalMetallic, c->gbNormalMetallic.p, pixels * 4, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(nonLinearDepth, c->gbDepth.p, pixels * 4, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return PRT_OK;
}

extern "C" prt_status prt_rtdi_set_reservoirs(prt_ctx *c, const float *packed)
{
    REQUIRE(c && packed, "prt_rtdi_set_reservoirs");
    REQUIRE(c->rtdiGbufferValid, "set the G-buffer first: it fixes the resolution");
    CK(cudaSetDevice(c->cfg.device));
    const size_t pixels = (size_t)c->rtdiWidth * c->rtdiHeight;
    CK(cudaMemcpyAsync(c->rtdiReservoirs.p, packed, pixels * 8, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    c->rtdiReservoirsValid = true;
    return PRT_OK;
}

// rtdi::InitialReservoirs::record (src/render/rtdi/InitialReservoirs.cpp; restir_di/initial_reservoirs.comp)
extern "C" prt_status prt_rtdi_initial_reservoirs(prt_ctx *c, uint32_t frameIndex)
{
    REQUIRE(c, "prt_rtdi_initial_reservoirs");
    REQUIRE(c->rtdiGbufferValid, "no G-buffer yet (prt_rtdi_set_gbuffer / prt_rtdi_gbuffer_from_primary)");
    REQUIRE(c->haveCamera && c->haveLights, "prt_set_camera / prt_set_lights have not been called");
    CK(cudaSetDevice(c->cfg.device));
    const uint32_t pixels = c->rtdiWidth * c->rtdiHeight;
    k_rtdi_initial_reservoirs<<<blocksFor(pixels, RENDER_THREADS), RENDER_THREADS, 0, c->stream>>>(c->view(), gbufferView(c), frameIndex,
                                                                                                     c->rtdiReservoirs.p);
    c->kernelLaunches++;
    CK(cudaGetLastError());
    c->rtdiReservoirsValid = true;
    return PRT_OK;
}

extern "C" prt_status prt_rtdi_read_reservoirs(prt_ctx *c, float *packedOut)
{
    REQUIRE(c && packedOut, "prt_rtdi_read_reservoirs");
    REQUIRE(c->rtdiReservoirsValid, "no reservoirs yet");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaMemcpyAsync(packedOut, c->rtdiReservoirs.p, (size_t)c->rtdiWidth * c->rtdiHeight * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return PRT_OK;
}

// rtdi::Trace::record (Trace.cpp:176-362): one frame. The history is recreated on resize, which resets it (Trace.cpp:185-219).
extern "C" prt_status prt_rtdi_trace(prt_ctx *c, const prt_rtdi_trace_pc *pcIn)
{
    REQUIRE(c && pcIn, "prt_rtdi_trace");
    REQUIRE(c->rtdiGbufferValid, "no G-buffer yet (Trace::Input::gbuffer)");
    REQUIRE(pcIn->drawType < PRT_DRAW_COUNT, "unknown draw type");
    REQUIRE(c->rtdiReservoirsValid || pcIn->drawType != PRT_DRAW_DEFAULT, "no reservoirs yet (Trace::Input::reservoirs)");
    CK(cudaSetDevice(c->cfg.device));
    prt_reference_pc rp{};
    rp.maxBounces = 1;
    CKS(checkRenderState(c, &rp, c->rtdiWidth, c->rtdiHeight));
    RtdiTracePC pc{pcIn->drawType, pcIn->frameIndex, pcIn->flags};
    if (!c->rtdiHistoryValid) pc.flags |= 1u;
    const uint32_t pixels = c->rtdiWidth * c->rtdiHeight;
    const SceneView sv = c->view();
    const Accel as = c->accel();
    const uint32_t grid = blocksFor(pixels, RENDER_THREADS);
    const bool count = (c->cfg.flags & PRT_CFG_COUNTERS) != 0;
    CK(cudaMemsetAsync(c->pb.counts, 0, COUNT_WORDS * sizeof(uint32_t), c->stream));
    k_rtdi_shade<<<grid, RENDER_THREADS, 0, c->stream>>>(sv, gbufferView(c), pc, c->rtdiReservoirs.p, c->pb);
    uint32_t launches = 2;
    if (pc.drawType == PRT_DRAW_DEFAULT)
    {
        launchTrace<true>(c, count, std::min(grid, c->traceBlocks), as, sv, 0, 0, nullptr);
        k_tally<<<1, 32, 0, c->stream>>>(c->pb, c->dCounters, pixels);
        launches += 2;
    }
    k_rtdi_accumulate<<<blocksFor(pixels, 256), 256, 0, c->stream>>>(c->pb, c->rtdiAccum.p, pixels, pc.flags, pc.drawType != PRT_DRAW_DEFAULT ? 1u : 0u);
    c->kernelLaunches += launches;
    CK(cudaGetLastError());
    c->rtdiHistoryValid = true;
    return PRT_OK;
}

extern "C" prt_status prt_rtdi_read_output_f32(prt_ctx *c, float *dstHost, uint64_t dstBytes)
{
    REQUIRE(c && dstHost, "prt_rtdi_read_output_f32");
    REQUIRE(c->rtdiHistoryValid, "nothing traced yet");
    const uint64_t bytes = (uint64_t)c->rtdiWidth * c->rtdiHeight * 16;
    REQUIRE(dstBytes >= bytes, "destination too small");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaMemcpyAsync(dstHost, c->rtdiAccum.p, bytes, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return PRT_OK;
}

// The fp16 image later passes expect (Trace.cpp:300-345)
extern "C" prt_status prt_rtdi_read_illumination_f16(prt_ctx *c, void *dstHost, uint64_t dstBytes)
{
    REQUIRE(c && dstHost, "prt_rtdi_read_illumination_f16");
    REQUIRE(c->rtdiHistoryValid, "nothing traced yet");
    const uint32_t pixels = c->rtdiWidth * c->rtdiHeight;
    REQUIRE(dstBytes >= (uint64_t)pixels * 8, "destination too small");
    CK(cudaSetDevice(c->cfg.device));
    CK(c->f16.ensure(pixels));
    k_blit_f16<<<blocksFor(pixels, 256), 256, 0, c->stream>>>(c->rtdiAccum.p, c->f16.p, pixels);
    c->kernelLaunches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(dstHost, c->f16.p, (size_t)pixels * 8, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return PRT_OK;
}

extern "C" prt_status prt_rtdi_release_preserved(prt_ctx *c)
{
    REQUIRE(c, "prt_rtdi_release_preserved");
    c->rtdiHistoryValid = false; // Trace::releasePreserved
    return PRT_OK;
}

extern "C" prt_status prt_release_preserved(prt_ctx *c)
{
    REQUIRE(c, "prt_release_preserved");
    c->historyValid = false; // RtRef.cpp:385-391
    return PRT_OK;
}

extern "C" prt_status prt_trace_primary(prt_ctx *c, const prt_reference_pc *pcIn, uint32_t width, uint32_t height, prt_hit *dstHost)
{
    REQUIRE(c && dstHost, "prt_trace_primary");
    CK(cudaSetDevice(c->cfg.device));
    CKS(checkRenderState(c, pcIn, width, height));
    const uint64_t pixels = (uint64_t)width * height;
    CKS(allocPathBuffers(c, pixels));
    CK(c->userHits.ensure(pixels));
    const SceneView sv = c->view();
    const Accel as = c->accel();
    cudaStream_t st = c->stream;
    const uint32_t grid = blocksFor(pixels, RENDER_THREADS);
    k_raygen<<<grid, RENDER_THREADS, 0, st>>>(sv, *pcIn, width, height, 1, c->pb);
    launchTrace<false>(c, true, std::min(grid, c->traceBlocks), as, sv, 0, 0, nullptr);
    k_hits_to_records<<<blocksFor(pixels, 256), 256, 0, st>>>(c->pb.hits, c->userHits.p, (uint32_t)pixels);
    c->kernelLaunches += 3;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(dstHost, c->userHits.p, pixels * sizeof(prt_hit), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return PRT_OK;
}

extern "C" prt_status prt_trace_rays(prt_ctx *c, const float *originTmin, const float *dirTmax, const uint32_t *seeds, uint32_t count,
                                     int shadow, prt_hit *dstHost)
{
    REQUIRE(c && originTmin && dirTmax && dstHost, "prt_trace_rays");
    CK(cudaSetDevice(c->cfg.device));
    CKS(refreshTextureSlots(c));
    CKS(refreshAlphaMats(c));
    if (count == 0) return PRT_OK;
    cudaStream_t st = c->stream;
    CK(c->userO.ensure(count)); CK(c->userD.ensure(count)); CK(c->userSeeds.ensure(count)); CK(c->userHits.ensure(count));
    CK(cudaMemcpyAsync(c->userO.p, originTmin, (size_t)count * 16, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(c->userD.p, dirTmax, (size_t)count * 16, cudaMemcpyHostToDevice, st));
    if (seeds) CK(cudaMemcpyAsync(c->userSeeds.p, seeds, (size_t)count * 4, cudaMemcpyHostToDevice, st));
    const SceneView sv = c->view();
    const Accel as = c->accel();
    const uint32_t grid = blocksFor(count, RENDER_THREADS);
    if (shadow) k_trace_user<true><<<grid, RENDER_THREADS, 0, st>>>(as, sv, c->userO.p, c->userD.p, seeds ? c->userSeeds.p : nullptr, count, c->userHits.p, c->dCounters);
    else k_trace_user<false><<<grid, RENDER_THREADS, 0, st>>>(as, sv, c->userO.p, c->userD.p, seeds ? c->userSeeds.p : nullptr, count, c->userHits.p, c->dCounters);
    c->kernelLaunches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(dstHost, c->userHits.p, (size_t)count * sizeof(prt_hit), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return PRT_OK;
}

// ---- multi-GPU ------------------------------------------------------------------------------------------------------
extern "C" prt_status prt_export_sum_device(prt_ctx *c, float **outSum)
{
    REQUIRE(c && outSum, "prt_export_sum_device");
    REQUIRE(c->historyValid, "nothing recorded yet");
    CK(cudaSetDevice(c->cfg.device));
    const uint32_t pixels = c->accumWidth * c->accumHeight;
    CK(c->sum.ensure(pixels));
    k_mean_to_sum<<<blocksFor(pixels, 256), 256, 0, c->stream>>>(c->accum.p, c->sum.p, pixels);
    c->kernelLaunches++;
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(c->stream)); // the caller reduces it on its own (NCCL) stream
    *outSum = reinterpret_cast<float *>(c->sum.p);
    return PRT_OK;
}

extern "C" prt_status prt_export_sum_to_device(prt_ctx *c, float *dstDevice)
{
    REQUIRE(c && dstDevice, "prt_export_sum_to_device");
    REQUIRE(c->historyValid, "nothing recorded yet");
    CK(cudaSetDevice(c->cfg.device));
    const uint32_t pixels = c->accumWidth * c->accumHeight;
    // stream-ordered on prt_stream(): the caller's collective is enqueued behind it (no host synchronisation)
    k_mean_to_sum<<<blocksFor(pixels, 256), 256, 0, c->stream>>>(c->accum.p, reinterpret_cast<float4 *>(dstDevice), pixels);
    c->kernelLaunches++;
    CK(cudaGetLastError());
    return PRT_OK;
}

extern "C" prt_status prt_import_sum_device(prt_ctx *c, const float *sumDevice)
{
    REQUIRE(c && sumDevice, "prt_import_sum_device");
    REQUIRE(c->accumWidth && c->accumHeight, "no render area yet");
    CK(cudaSetDevice(c->cfg.device));
    const uint32_t pixels = c->accumWidth * c->accumHeight;
    // stream-ordered on prt_stream(): `sumDevice` must be complete in that stream's order (e.g. the collective's work.wait()
    // with prt_stream() current); no host synchronisation
    k_sum_to_mean<<<blocksFor(pixels, 256), 256, 0, c->stream>>>(reinterpret_cast<const float4 *>(sumDevice), c->accum.p, pixels);
    c->kernelLaunches++;
    CK(cudaGetLastError());
    c->historyValid = true;
    return PRT_OK;
}

// ---- metrics -----------------------------------------------------------------------------------------------------------
extern "C" prt_status prt_get_counters(prt_ctx *c, prt_counters *out)
{
    REQUIRE(c && out, "prt_get_counters");
    CK(cudaSetDevice(c->cfg.device));
    GlobalCounters g;
    CK(cudaMemcpyAsync(&g, c->dCounters, sizeof(g), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    out->closestRays = g.closestRays; out->shadowRays = g.shadowRays; out->nodesVisited = g.nodesVisited;
    out->trisTested = g.trisTested; out->instancesEntered = g.instancesEntered; out->anyHitCalls = g.anyHitCalls;
    out->shadedHits = g.shadedHits; out->textureTaps = g.textureTaps; out->pixelSamples = g.pixelSamples;
    out->kernelLaunches = c->kernelLaunches;
    out->stackOverflows = g.stackOverflows;
    out->shadowNodesVisited = g.shadowNodes; out->shadowTrisTested = g.shadowTris;
    out->shadowInstancesEntered = g.shadowInstances; out->shadowAnyHitCalls = g.shadowAnyHits;
    return PRT_OK;
}

extern "C" prt_status prt_reset_counters(prt_ctx *c)
{
    REQUIRE(c, "prt_reset_counters");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaMemsetAsync(c->dCounters, 0, sizeof(GlobalCounters), c->stream));
    c->kernelLaunches = 0;
    return PRT_OK;
}

extern "C" prt_status prt_get_timings(prt_ctx *c, prt_timings *out)
{
    REQUIRE(c && out, "prt_get_timings");
    CK(cudaSetDevice(c->cfg.device));
    CK(cudaStreamSynchronize(c->stream));
    memset(out, 0, sizeof(*out));
    if (c->recordTimed) CK(cudaEventElapsedTime(&out->lastRecordMs, c->evRecord[0], c->evRecord[1]));
    if (c->tlasTimed) CK(cudaEventElapsedTime(&out->lastTlasMs, c->evTlas[0], c->evTlas[1]));
    if (c->blasTimed) CK(cudaEventElapsedTime(&out->lastBlasMs, c->evBlas[0], c->evBlas[1]));
    if (c->recordTimed)
    {
        float sums[3] = {0.f, 0.f, 0.f};
        uint32_t n[3] = {0, 0, 0};
        for (uint32_t i = 0; i + 1 < c->evUsed; ++i)
        {
            const uint32_t stage = c->stageOf[i];
            if (stage > 2) continue;
            float ms = 0.f;
            CK(cudaEventElapsedTime(&ms, c->evStage[i], c->evStage[i + 1]));
            sums[stage] += ms;
            n[stage]++;
        }
        out->lastTraceClosestMs = sums[0]; out->lastTraceClosestLaunches = n[0];
        out->lastShadeMs = sums[1]; out->lastShadeLaunches = n[1];
        out->lastTraceShadowMs = sums[2]; out->lastTraceShadowLaunches = n[2];
    }
    out->lastRecordLaunches = c->lastRecordLaunches;
    out->lastTlasLaunches = c->lastTlasLaunches;
    out->ringWraps = c->ring.wraps;
    out->tlasGraphReplays = c->tlasGraphReplays;
    out->tlasGraphCaptures = c->tlasGraphCaptures;
    return PRT_OK;
}

extern "C" prt_status prt_get_bvh_stats(prt_ctx *c, prt_bvh_stats *out)
{
    REQUIRE(c && out, "prt_get_bvh_stats");
    memset(out, 0, sizeof(*out));
    out->tlasInstances = c->tlasInstances;
    for (const BlasRes &b : c->blas)
    {
        if (!b.built) continue;
        out->blasCount++;
        out->blasNodesTotal += b.nodeCount;
        out->blasTrianglesTotal += b.primCount;
        out->blasReferencesTotal += b.triCount;
    }
    if (c->tlasInstances)
    {
        uint32_t nodes = 0;
        CK(cudaSetDevice(c->cfg.device));
        CK(cudaMemcpyAsync(&nodes, c->bScalars.p, 4, cudaMemcpyDeviceToHost, c->stream));
        CK(cudaStreamSynchronize(c->stream));
        out->tlasNodes = nodes; // valid when the TLAS was the last structure built
    }
    return PRT_OK;
}

extern "C" prt_status prt_probe_texture(prt_ctx *c, uint32_t textureSampler, const float *uv, const int32_t *texelXY, uint32_t count,
                                        int exact, float *outRgba)
{
    REQUIRE(c && outRgba && (uv || texelXY), "prt_probe_texture");
    if (count == 0) return PRT_OK;
    CK(cudaSetDevice(c->cfg.device));
    CKS(refreshTextureSlots(c));
    cudaStream_t st = c->stream;
    DBuf<float4> out;
    DBuf<float2> in;
    CK(out.ensure(count));
    CK(in.ensure(count));
    const SceneView sv = c->view();
    if (texelXY)
    {
        const uint32_t tex = textureSampler & 0xFFFFFFu, smp = textureSampler >> 24;
        REQUIRE(tex < c->textures.size() && c->textures[tex].array, "prt_probe_texture: no such texture");
        const TexSlot slot = c->hSlots[(size_t)tex * sv.samplerCount + (smp < sv.samplerCount ? smp : 0)];
        CK(cudaMemcpyAsync(in.p, texelXY, (size_t)count * 8, cudaMemcpyHostToDevice, st));
        k_probe_texels<<<blocksFor(count, 128), 128, 0, st>>>(slot, reinterpret_cast<const int2 *>(in.p), count, out.p);
    }
    else
    {
        CK(cudaMemcpyAsync(in.p, uv, (size_t)count * 8, cudaMemcpyHostToDevice, st));
        k_probe_sample<<<blocksFor(count, 128), 128, 0, st>>>(sv, textureSampler, in.p, count, exact, out.p);
    }
    c->kernelLaunches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(outRgba, out.p, (size_t)count * 16, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    out.release();
    in.release();
    return PRT_OK;
}

// ---- debug read-back of the built structures (tests validate the GPU builder's output on the host) --------
extern "C" prt_status prt_debug_read_blas(prt_ctx *c, uint32_t modelIndex, void *nodesOut, uint64_t nodesBytes, void *trisOut,
                                          uint64_t trisBytes, uint32_t *nodeCount, uint32_t *triCount)
{
    REQUIRE(c && modelIndex < c->blas.size(), "prt_debug_read_blas");
    const BlasRes &b = c->blas[modelIndex];
    if (nodeCount) *nodeCount = b.nodeCount;
    if (triCount) *triCount = b.triCount;
    CK(cudaSetDevice(c->cfg.device));
    if (nodesOut)
    {
        REQUIRE(nodesBytes >= (uint64_t)b.nodeCount * WIDE_NODE_BYTES, "node buffer too small");
        CK(cudaMemcpy2DAsync(nodesOut, WIDE_NODE_BYTES, b.nodes, sizeof(WideNode), WIDE_NODE_BYTES, b.nodeCount, cudaMemcpyDeviceToHost, c->stream));
    }
    if (trisOut)
    {
        REQUIRE(trisBytes >= (uint64_t)b.triCount * TRI_REC_BYTES, "triangle buffer too small");
        CK(cudaMemcpy2DAsync(trisOut, TRI_REC_BYTES, b.tris, sizeof(TriRec), TRI_REC_BYTES, b.triCount, cudaMemcpyDeviceToHost, c->stream));
    }
    CK(cudaStreamSynchronize(c->stream));
    return PRT_OK;
}

extern "C" prt_status prt_debug_read_tlas(prt_ctx *c, void *nodesOut, uint64_t nodesBytes, void *instOut, uint64_t instBytes,
                                          uint32_t *instanceCount)
{
    REQUIRE(c, "prt_debug_read_tlas");
    const uint32_t n = c->tlasInstances;
    if (instanceCount) *instanceCount = n;
    CK(cudaSetDevice(c->cfg.device));
    if (nodesOut && n)
    {
        REQUIRE(nodesBytes >= (uint64_t)n * WIDE_NODE_BYTES, "node buffer too small (instanceCount nodes)");
        CK(cudaMemcpy2DAsync(nodesOut, WIDE_NODE_BYTES, c->tlasNodesP, sizeof(WideNode), WIDE_NODE_BYTES, n, cudaMemcpyDeviceToHost, c->stream));
    }
    if (instOut && n)
    {
        REQUIRE(instBytes >= (uint64_t)n * sizeof(InstRec), "instance buffer too small");
        CK(cudaMemcpyAsync(instOut, c->tlasInst.p, (size_t)n * sizeof(InstRec), cudaMemcpyDeviceToHost, c->stream));
    }
    CK(cudaStreamSynchronize(c->stream));
    return PRT_OK;
}
