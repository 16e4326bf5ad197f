// pt_api.cu -- context + C ABI of libptb200.so (include/ptb200.h).  Host orchestration of the
// wavefront stages == what reference src/pathtracer.h:73-310 (Pathtracer::Init/Render/Finish) does,
// minus the GL interop (headless float4 device images instead of GL textures).
// Compiled for sm_100a only; every compute entry point fails loudly without a CUDA device.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <mutex>
#include <thread>
#include <atomic>
#include <vector>
#include <utility>
#include <chrono>
#include "pt_kernels.cuh"
#include "pt_post.cuh"
#include "wide_bvh.h"

using namespace ptb;

static thread_local std::string g_createError;

#define PT_MAX_SETS 8
#define PT_SORT_MAX_BITS 5          /* ray sorting: up to 32 cells per axis x 8 octants = 262144 bins */
#define PT_SORT_MAX_BOUNCES 2       /* queues produced by the first PT_SORT_MAX_BOUNCES shade launches of a sample can be sorted */
#define PT_GATHER_RING 4            /* image slots of the N-rank framebuffer gather on rank 0 */
#define PT_GATHER_ARRIVE 256        /* flag words (uint[16]) in the radiance-cache block header: rank r pushed its tiles of this epoch */
#define PT_GATHER_CONSUMED 320      /* rank 0 has shipped epoch e to the host: the slot of e may be overwritten by e + PT_GATHER_RING */
#define PT_GATHER_COUNTER 321
#define PT_GATHER_BANDDONE 384      /* uint[16]: rank r has copied its band of epoch e from rank 0's slot to the host framebuffer */
struct pt_ctx {
    pt_config cfg;
    std::string err;
    cudaStream_t stream = nullptr;
    int numSMs = 0;
    TileMap map;
    // Per-sample working state.  PT_MAX_SETS sets: converge samples rotate through them on private streams, so the
    // short-queue end of sample k (deep bounces, k_tail) overlaps the wide launches of samples k+1.. (measured on
    // B200, sibenik 1080p: 2.46 -> 2.03 ms/sample; one eighth of the image: 0.86 -> 0.48 ms/sample).
    struct SampleSet {
        Queues q;
        float4 *masks = nullptr, *accucolors = nullptr, *lights = nullptr, *albedoSample = nullptr;
        pt_sample_cache* sampleCache = nullptr;
        uint *sortKeys[2] = {nullptr, nullptr}, *sortPerm[2] = {nullptr, nullptr}, *sortHist = nullptr;   // ray sorting: [0] extension, [1] shadow queue; hist = [bounce][queue][bins] behind the counters
        cudaStream_t stream = nullptr;
        cudaEvent_t done = nullptr;       // recorded after the sample's accumulate
        cudaEvent_t tailStart = nullptr;  // recorded right before the sample's k_tail launch
        cudaEvent_t partDone = nullptr;   // split frames: this tile range's cache additions are in
        bool doneValid = false, tailValid = false;
    } set[PT_MAX_SETS];
    int nSets = 4, cur = 0;
    bool forked = false, prevTraining = false;
    cudaEvent_t forkEv = nullptr;
    // Asynchronous read-back (pt_read_image_async): copies run on their own stream after the sample they follow; the
    // next accumulate waits for the copy, so the snapshot is exactly "after that sample".  Tickets index a ring of events.
    cudaStream_t copyStream = nullptr;
    cudaEvent_t copyReady = nullptr, copyEv[8] = {}, imgFree = nullptr;   // imgFree: the pending read-back no longer reads the image
    uint64_t copyTickets = 0;
    bool copyPending = false;
    int lastSet = -1;
    float4 *image = nullptr, *albedo = nullptr, *postBuf = nullptr, *packBuf = nullptr;
    unsigned long long* stats = nullptr;
    // scene
    DScene sc;
    bool haveScene = false;
    std::vector<void*> sceneAllocs;
    uint32_t numInstancesCap = 0, numTlasCap = 0, maxWideDepth = 0, numWideNodes = 0, numWideTris = 0;
    uint32_t numModels = 0;
    float4* modelBox = nullptr; unsigned char* tlasAlive = nullptr; uint32_t* tlasDepth = nullptr;   // device top-level BVH build (k_build_tlas)
    // textures
    FrameTex tex;
    std::vector<cudaArray_t> arrays;
    std::vector<cudaTextureObject_t> textures;
    // state mirrored from Pathtracer
    bool HNEE = true, HCACHE = true;      // pathtracer.h:213-214
    bool DNEE = false, DCACHE = false;    // __constant__ copies, synced on shouldClear (pathtracer.h:242-243)
    uint32_t sampleIdx = 0, randIdx = 0;
    unsigned long long launches = 0;      // kernels launched by pt_render since the last pt_reset_stats
    pt_allreduce_hook hook = nullptr;
    void* hookUser = nullptr;
    // radiance-cache block: [flags 4 KB][rcAdditions][rcRadiance] in ONE allocation, so that a single cudaIpc handle
    // gives the peers everything the fused reduce+propagate needs (pt_peer_export / pt_peer_connect)
    unsigned char* rcBlock = nullptr;
    size_t rcAddOffset = 0, rcRcOffset = 0;
    struct Peers { bool on = false; PeerSet ps; void* opened[PT_PEER_MAX] = {}; uint32_t epoch = 0; unsigned long long timeoutNs = 30000000000ull; } peer;
    int gridTraceExt = 0, gridTraceAny = 0, gridShade = 0, gridTail = 0;
    uint32_t tailBounce = 3;              // bounces >= this run in k_tail (one per-path launch) when a sample runs alone; PT_MAX_BOUNCES: never
    bool tailForced = false;              // PT_FLAG_TAIL_BOUNCE / PTB200_TAIL_BOUNCE given: use k_tail for overlapped samples too
    uint32_t knobMinWidth = 0; bool knobNoTailWait = false, knobNoTailPair = false;   // developer knobs (environment, read in pt_create)
    bool knobPdlAlways = false, knobNoMix = false;
    uint32_t anyFarFirst = 0;
    uint32_t split = 1, splitWidth = 32; bool splitTail = true, splitHold = false;   // cache-training frames as `split` concurrent tile ranges (PTB200_SPLIT=parts[,width[,tail[,hold]]])
    uint32_t headAt = 3;                  // bounce at which a lone sample releases the next sample's head (PTB200_HEAD_AT)
    // Feedback for the tail bounce of samples that run alone: the per-bounce queue sizes of finished samples come back through a
    // small pinned ring (asynchronous copies; the host never waits for one) and the next sample starts its per-path tail at the
    // first bounce whose queue has shrunk to about one and a half paths per tail warp.
    uint32_t* fbHost = nullptr; uint32_t* fbDev = nullptr; uint32_t fbSeen = 0, fbTail = 0; bool fbOn = true, tailFixed = false;   // mapped pinned memory: [0] sequence number, [1 + b] extension rays of bounce b
    uint32_t tailPairMax = 16;            // k_tail gives a path two lanes while the paths number <= this many per warp (PTB200_TAIL_PAIR_MAX)
    uint32_t coop = 4, coopTail = 2;      // cooperative (8 lanes per ray) traversal for queues up to coop x 4 rays (paths) per warp, latency mode only (PTB200_COOP=trace[,tail])
    bool pdl = true;                      // stage kernels are launched with programmatic stream serialization (PT_FLAG_NO_PDL / PTB200_NO_PDL: off)
    bool pdlNow = true;                   // ... for the sample being enqueued: samples that overlap others launch plainly (see pt_render)
    bool sortOn = false; uint32_t sortBits = 4, sortBounces = PT_SORT_MAX_BOUNCES;          // pt_set_ray_sorting
    bool counting = false; unsigned long long* tally = nullptr;                            // pt_set_counting: [0..2] extend, [3..5] connect
    float *skyPdf = nullptr, *skyCdf = nullptr;                                           // pt_set_skydome builds them (pathtracer.h:85-122)
    // N-rank framebuffer gather (pt_gather_image_async): image ring on rank 0, flags in the radiance-cache block header
    struct HostRing { unsigned char* host = nullptr; unsigned char* dev = nullptr; uint64_t bytes = 0; uint32_t slots = 0; uint64_t epoch = 0; } ring;   // pt_host_ring_attach
    unsigned char* gatherBlock = nullptr; size_t gatherBytes = 0; void* gatherOpened = nullptr; float4* gatherRoot = nullptr; uint32_t gatherEpoch = 0;
    // per-stage timing (pt_set_profiling): CUDA event pairs around every stage launch, resolved in pt_finish
    bool prof = false;
    std::vector<cudaEvent_t> evPool;
    struct Pending { int stage; cudaEvent_t a, b; };
    std::vector<Pending> pending;
    double profMs[PT_NUM_STAGES] = {0, 0, 0, 0, 0, 0};
    unsigned long long profLaunches[PT_NUM_STAGES] = {0, 0, 0, 0, 0, 0};
    // debug buffers
    pt_ray* dbgRays = nullptr; pt_hit* dbgHits = nullptr; uint32_t dbgCap = 0; uint* dbgCursor = nullptr;
};

#define CK(call)                                                                                          \
    do {                                                                                                  \
        cudaError_t e_ = (call);                                                                          \
        if (e_ != cudaSuccess) {                                                                          \
            char buf_[512];                                                                               \
            snprintf(buf_, sizeof(buf_), "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            ctx->err = buf_;                                                                              \
            return e_ == cudaErrorMemoryAllocation ? PT_ERR_OOM : PT_ERR_CUDA;                            \
        }                                                                                                 \
    } while (0)

static cudaEvent_t getEvent(pt_ctx* ctx) {
    if (!ctx->evPool.empty()) { cudaEvent_t e = ctx->evPool.back(); ctx->evPool.pop_back(); return e; }
    cudaEvent_t e = nullptr; cudaEventCreate(&e); return e;
}
struct StageTimer {      // records an event pair on the context's stream around one stage launch
    pt_ctx* ctx; int stage; cudaEvent_t a = nullptr, b = nullptr;
    cudaStream_t s;
    StageTimer(pt_ctx* c, int st, cudaStream_t stream) : ctx(c), stage(st), s(stream) { if (ctx->prof) { a = getEvent(ctx); b = getEvent(ctx); cudaEventRecord(a, s); } }
    ~StageTimer() { if (a) { cudaEventRecord(b, s); ctx->pending.push_back({stage, a, b}); } }
};
static void resolveProfile(pt_ctx* ctx) {   // call after the stream has been synchronised
    for (auto& p : ctx->pending) {
        float ms = 0; if (cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) { ctx->profMs[p.stage] += ms; ctx->profLaunches[p.stage]++; }
        ctx->evPool.push_back(p.a); ctx->evPool.push_back(p.b);
    }
    ctx->pending.clear();
}

static int fail(pt_ctx* ctx, int code, const char* msg) { ctx->err = msg; return code; }

// Every stage launch of pt_render goes through here: counted (pt_stats.kernel_launches) and, unless switched off, launched
// with programmatic stream serialization, so the next kernel's blocks are scheduled while this one drains (each kernel
// starts with pdlWait(), pt_device.cuh).  Errors surface through the cudaGetLastError at the end of pt_render.
template <typename... KA, typename... A>
static inline void launchK(pt_ctx* ctx, void (*kernel)(KA...), dim3 grid, dim3 block, cudaStream_t s, A&&... args) {
    ctx->launches++;
    cudaLaunchConfig_t c; memset(&c, 0, sizeof(c));
    c.gridDim = grid; c.blockDim = block; c.dynamicSmemBytes = 0; c.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    c.attrs = at; c.numAttrs = ctx->pdlNow ? 1u : 0u;
    cudaLaunchKernelEx(&c, kernel, std::forward<A>(args)...);
}

template <typename T>
static int devAlloc(pt_ctx* ctx, T** p, size_t n, bool sceneOwned) {
    *p = nullptr;
    CK(cudaMalloc((void**)p, std::max<size_t>(n, 1) * sizeof(T)));
    if (sceneOwned) ctx->sceneAllocs.push_back((void*)*p);
    return PT_OK;
}
template <typename T>
static int devUpload(pt_ctx* ctx, T** p, const T* host, size_t n, bool sceneOwned) {
    int r = devAlloc(ctx, p, n, sceneOwned);
    if (r) return r;
    if (n) CK(cudaMemcpyAsync(*p, host, n * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    return PT_OK;
}

// Work enqueued by pt_render lives on the private streams of the sample sets.  joinStreams() orders everything
// enqueued so far before whatever comes next on the context's public stream (pt_cuda_stream()).
// Bounds-checked build: the extents the kernels compare their indices with (pt_device.cuh PT_IDX).
static int setCheckLimits(pt_ctx* ctx, uint32_t numTlas, uint32_t numModels, uint32_t numMaterials) {
#if defined(PT_BOUNDS_CHECK)
    CheckLimits& l = ctx->sc.chk;
    l.numNodes = ctx->numWideNodes; l.numWtris = ctx->numWideTris; l.numTlas = numTlas; l.numModels = numModels; l.numMaterials = numMaterials;
    l.queueCap = std::max<uint32_t>(ctx->map.nSlots, 1024u); l.nSlots = std::max<uint32_t>(ctx->map.nSlots, 1024u); l.stack = PT_STACK;
    if (getenv("PTB200_CHK_SHRINK")) l.numWtris = 1;          // positive control of the checks themselves (tests/test_zy_bounds_checked.py); a shrunk NODE extent would turn the substituted index 0 into an endless traversal
#else
    (void)ctx; (void)numTlas; (void)numModels; (void)numMaterials;
#endif
    return PT_OK;
}
static int joinStreams(pt_ctx* ctx) {
    if (ctx->copyPending) { CK(cudaStreamWaitEvent(ctx->stream, ctx->copyEv[(ctx->copyTickets - 1) & 7], 0)); ctx->copyPending = false; }
    if (!ctx->forked) return PT_OK;
    for (auto& S : ctx->set) {
        if (S.doneValid) { CK(cudaStreamWaitEvent(ctx->stream, S.done, 0)); S.doneValid = false; }
        S.tailValid = false;
    }
    ctx->forked = false;
    return PT_OK;
}

extern "C" {

#if defined(PT_BOUNDS_CHECK)
PT_API const char* pt_version(void) { return "ptb200 0.1 (sm_100a, bounds-checked build)"; }
#else
PT_API const char* pt_version(void) { return "ptb200 0.1 (sm_100a)"; }
#endif

PT_API void pt_default_config(pt_config* c) {
    memset(c, 0, sizeof(*c));
    c->width = 640; c->height = 480;          // constants.h:4-5
    c->max_ray_depth = 32;                    // constants.h:13
    c->interactive_depth = 1;                 // scene.h:137
    c->guiding_frames = 100;                  // pathtracer.h:292
    c->blue_noise_frames = 1;                 // kernels.h:23
    c->device = 0; c->rank = 0; c->world_size = 1; c->flags = PT_FLAG_DEFAULT;
}

PT_API const char* pt_last_error(const pt_ctx* ctx) { return ctx ? ctx->err.c_str() : g_createError.c_str(); }

static void peerClose(pt_ctx* ctx) {
    for (auto& p : ctx->peer.opened) if (p) { cudaIpcCloseMemHandle(p); p = nullptr; }
    if (ctx->gatherOpened) { cudaIpcCloseMemHandle(ctx->gatherOpened); ctx->gatherOpened = nullptr; }
    ctx->gatherRoot = nullptr;
    ctx->peer.on = false;
}

PT_API void pt_destroy(pt_ctx* ctx) {
    if (ctx && ctx->ring.host) { cudaSetDevice(ctx->cfg.device); cudaDeviceSynchronize(); cudaHostUnregister(ctx->ring.host); ctx->ring.host = nullptr; }
    if (!ctx) return;
    cudaSetDevice(ctx->cfg.device);
    for (auto& S : ctx->set) if (S.stream) cudaStreamSynchronize(S.stream);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    resolveProfile(ctx);
    for (auto e : ctx->evPool) cudaEventDestroy(e);
    for (auto t : ctx->textures) cudaDestroyTextureObject(t);
    for (auto a : ctx->arrays) cudaFreeArray(a);
    peerClose(ctx);
    for (void* p : ctx->sceneAllocs) cudaFree(p);
    for (auto& S : ctx->set) {
        if (S.stream) cudaStreamSynchronize(S.stream);
        void* sb[] = {S.q.rayO[0], S.q.rayO[1], S.q.rayD[0], S.q.rayD[1], S.q.shO, S.q.shD, S.q.hits, S.q.cnt,
                      S.masks, S.accucolors, S.lights, S.albedoSample, &S == &ctx->set[0] ? S.sampleCache : nullptr,
                      S.sortKeys[0], S.sortKeys[1], S.sortPerm[0], S.sortPerm[1]};
        for (void* p : sb) if (p) cudaFree(p);
        if (S.done) cudaEventDestroy(S.done);
        if (S.tailStart) cudaEventDestroy(S.tailStart);
        if (S.partDone) cudaEventDestroy(S.partDone);
        if (S.stream) cudaStreamDestroy(S.stream);
    }
    if (ctx->forkEv) cudaEventDestroy(ctx->forkEv);
    if (ctx->copyReady) cudaEventDestroy(ctx->copyReady);
    if (ctx->imgFree) cudaEventDestroy(ctx->imgFree);
    if (ctx->fbHost) cudaFreeHost(ctx->fbHost);
    if (ctx->gatherBlock) cudaFree(ctx->gatherBlock);
    if (ctx->skyPdf) cudaFree(ctx->skyPdf);
    if (ctx->skyCdf) cudaFree(ctx->skyCdf);
    if (ctx->tally) cudaFree(ctx->tally);
    for (auto e : ctx->copyEv) if (e) cudaEventDestroy(e);
    if (ctx->copyStream) cudaStreamDestroy(ctx->copyStream);
    void* bufs[] = {ctx->image, ctx->albedo, ctx->postBuf, ctx->packBuf, ctx->stats, ctx->dbgRays, ctx->dbgHits, ctx->dbgCursor};
    for (void* p : bufs) if (p) cudaFree(p);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

static TileMap makeTileMap(const pt_config* cfg) {
    TileMap m;
    m.W = cfg->width; m.H = cfg->height;
    m.tilesX = (m.W + 31) / 32; m.tilesY = (m.H + 31) / 32;
    m.rank = cfg->rank; m.world = cfg->world_size;
    const uint32_t T = m.tilesX * m.tilesY;
    m.nLocalTiles = T > m.rank ? (T - m.rank + m.world - 1) / m.world : 0;
    m.nSlots = m.nLocalTiles * 1024u;
    m.slotBase = 0;
    return m;
}
// Host view of the tile partition (no GPU needed): the same mapping the kernels use.
PT_API uint32_t pt_tile_num_slots(const pt_config* cfg) {
    if (!cfg || cfg->world_size == 0 || cfg->rank >= cfg->world_size) return 0;
    return makeTileMap(cfg).nSlots;
}
PT_API int pt_tile_slot_to_pixel(const pt_config* cfg, uint32_t slot, uint32_t* x, uint32_t* y) {
    if (!cfg || !x || !y || cfg->world_size == 0 || cfg->rank >= cfg->world_size) return PT_ERR_INVALID;
    const TileMap m = makeTileMap(cfg);
    if (slot >= m.nSlots) return PT_ERR_INVALID;
    return slotToPixel(m, slot, *x, *y) ? 1 : 0;
}

PT_API int pt_create(pt_ctx** out, const pt_config* cfg) {
    if (!out || !cfg) { g_createError = "pt_create: null argument"; return PT_ERR_INVALID; }
    *out = nullptr;
    if (cfg->width == 0 || cfg->height == 0 || cfg->world_size == 0 || cfg->rank >= cfg->world_size ||
        cfg->max_ray_depth == 0 || cfg->max_ray_depth > PT_MAX_BOUNCES || cfg->interactive_depth + 1 > PT_MAX_BOUNCES) {
        g_createError = "pt_create: invalid config (size, rank/world_size or depth out of range; max depth 64)";
        return PT_ERR_INVALID;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0 || cfg->device >= ndev) {
        cudaGetLastError();
        g_createError = "pt_create: no CUDA device available -- libptb200 has no CPU fallback";
        return PT_ERR_NO_DEVICE;
    }
    pt_ctx* ctx = new pt_ctx();
    ctx->cfg = *cfg;
    for (auto& S : ctx->set) memset(&S.q, 0, sizeof(S.q));
    memset(&ctx->sc, 0, sizeof(ctx->sc)); memset(&ctx->tex, 0, sizeof(ctx->tex));
    ctx->tex.skyConst = make_float3(1.0f, 1.0f, 1.0f);
    auto bail = [&](int code) { g_createError = ctx->err; pt_destroy(ctx); return code; };
    auto body = [&]() -> int {
        CK(cudaSetDevice(cfg->device));
        CK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
        CK(cudaDeviceGetAttribute(&ctx->numSMs, cudaDevAttrMultiProcessorCount, cfg->device));
        TileMap& m = ctx->map;
        m = makeTileMap(cfg);
        const size_t cap = std::max<uint32_t>(m.nSlots, 1024u);
        int r;
        ctx->nSets = m.nSlots <= 1200000u ? 8 : 4;       // small shards (multi-GPU) are latency bound per launch: more samples in flight
                                                         // (measured, sibenik 1080p: 1/2 of the image 0.813 -> 0.768 ms/sample with 8; full image +2% for 2x the memory)
        if ((cfg->flags & PT_FLAG_NO_PIPELINE) || getenv("PTB200_NO_PIPELINE")) ctx->nSets = 1;
        else if (const char* e = getenv("PTB200_SETS")) ctx->nSets = std::min(PT_MAX_SETS, std::max(1, atoi(e)));    // developer knob
        CK(cudaEventCreateWithFlags(&ctx->forkEv, cudaEventDisableTiming));
        for (int k = 0; k < ctx->nSets; k++) {
            pt_ctx::SampleSet& S = ctx->set[k];
            CK(cudaStreamCreateWithFlags(&S.stream, cudaStreamNonBlocking));
            CK(cudaEventCreateWithFlags(&S.done, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&S.tailStart, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&S.partDone, cudaEventDisableTiming));
            for (int i = 0; i < 2; i++) {
                if ((r = devAlloc(ctx, &S.q.rayO[i], cap, false))) return r;
                if ((r = devAlloc(ctx, &S.q.rayD[i], cap, false))) return r;
            }
            if ((r = devAlloc(ctx, &S.q.shO, cap, false))) return r;
            if ((r = devAlloc(ctx, &S.q.shD, cap, false))) return r;
            if ((r = devAlloc(ctx, &S.q.hits, cap, false))) return r;
            // counters, then the ray-sorting histograms [bounce][queue][bins]: one allocation, zeroed by ONE memset per sample
            if ((r = devAlloc(ctx, &S.q.cnt, (size_t)CNT_TOTAL + (size_t)PT_SORT_MAX_BOUNCES * 2 * (8u << (3 * PT_SORT_MAX_BITS)), false))) return r;
            S.sortHist = S.q.cnt + CNT_TOTAL;
            CK(cudaMemsetAsync(S.sortHist, 0, (size_t)PT_SORT_MAX_BOUNCES * 2 * (8u << (3 * PT_SORT_MAX_BITS)) * sizeof(uint), ctx->stream));
            if ((r = devAlloc(ctx, &S.masks, cap, false))) return r;
            if ((r = devAlloc(ctx, &S.accucolors, cap, false))) return r;
            if ((r = devAlloc(ctx, &S.lights, cap, false))) return r;
            if ((r = devAlloc(ctx, &S.albedoSample, cap, false))) return r;
            // ONE sample cache for both sets: entries a path does not rewrite (emissive hits) must stay those of the
            // previous frame, as in the reference (kernels.h:563-576, 863-882); training frames never overlap.
            if (k == 0) { if ((r = devAlloc(ctx, &S.sampleCache, cap * PT_MAX_CACHE_DEPTH, false))) return r; }
            else S.sampleCache = ctx->set[0].sampleCache;
            CK(cudaMemsetAsync(S.q.cnt, 0, CNT_TOTAL * sizeof(uint), ctx->stream));
            CK(cudaMemsetAsync(S.lights, 0, cap * sizeof(float4), ctx->stream));
            CK(cudaMemsetAsync(S.masks, 0, cap * sizeof(float4), ctx->stream));
            CK(cudaMemsetAsync(S.accucolors, 0, cap * sizeof(float4), ctx->stream));
            CK(cudaMemsetAsync(S.albedoSample, 0, cap * sizeof(float4), ctx->stream));
            if (k == 0) CK(cudaMemsetAsync(S.sampleCache, 0, cap * PT_MAX_CACHE_DEPTH * sizeof(pt_sample_cache), ctx->stream));
        }
        if ((r = devAlloc(ctx, &ctx->image, (size_t)m.W * m.H, false))) return r;
        if ((r = devAlloc(ctx, &ctx->albedo, (size_t)m.W * m.H, false))) return r;
        if ((r = devAlloc(ctx, &ctx->stats, (size_t)8, false))) return r;
        if ((r = devAlloc(ctx, &ctx->tally, (size_t)8, false))) return r;
        CK(cudaMemsetAsync(ctx->tally, 0, 8 * sizeof(unsigned long long), ctx->stream));
        CK(cudaEventCreateWithFlags(&ctx->imgFree, cudaEventDisableTiming));
        if ((r = devAlloc(ctx, &ctx->dbgCursor, (size_t)1, false))) return r;
        CK(cudaMemsetAsync(ctx->image, 0, (size_t)m.W * m.H * sizeof(float4), ctx->stream));
        CK(cudaMemsetAsync(ctx->albedo, 0, (size_t)m.W * m.H * sizeof(float4), ctx->stream));
        CK(cudaMemsetAsync(ctx->stats, 0, 8 * sizeof(unsigned long long), ctx->stream));
        int occ = 0;
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_trace<false>, 128, 0));
        ctx->gridTraceExt = ctx->numSMs * std::max(occ, 1);
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_trace<true>, 128, 0));
        ctx->gridTraceAny = ctx->numSMs * std::max(occ, 1);
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_shade, 128, 0));
        ctx->gridShade = ctx->numSMs * std::max(occ, 1);
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_tail, 128, 0));
        ctx->gridTail = ctx->numSMs * std::max(occ, 1);
        if (const char* e = getenv("PTB200_TRACE_GRID")) ctx->gridTraceExt = ctx->gridTraceAny = ctx->numSMs * std::max(1, atoi(e));   // developer knob
        if (const char* e = getenv("PTB200_SHADE_GRID")) ctx->gridShade = ctx->numSMs * std::max(1, atoi(e));   // developer knob
        if (const char* e = getenv("PTB200_TAIL_GRID")) ctx->gridTail = ctx->numSMs * std::max(1, atoi(e));   // developer knob
        if (cfg->flags & PT_FLAG_WAVEFRONT_ONLY) ctx->tailBounce = PT_MAX_BOUNCES;
        else if ((cfg->flags >> 8) & 0xffu) { ctx->tailBounce = std::max(1u, ((cfg->flags >> 8) & 0xffu) - 1u); ctx->tailForced = true; }
        if (const char* e = getenv("PTB200_TAIL_BOUNCE")) { ctx->tailBounce = std::max(1, atoi(e)); ctx->tailForced = true; }      // developer knob (tools/tune.sh)
        if (const char* e = getenv("PTB200_TAIL_ALONE")) { ctx->tailBounce = std::max(1, atoi(e)); ctx->tailFixed = true; }   // developer knob: tail bounce of samples that run alone only
        ctx->fbOn = !getenv("PTB200_NO_FEEDBACK") && !ctx->tailForced && !ctx->tailFixed && !(cfg->flags & PT_FLAG_WAVEFRONT_ONLY);
        if (ctx->fbOn) {
            // written by k_accumulate with plain stores (a copy on the sample's stream would queue behind the image read-backs on the
            // device-to-host copy engine and stall the next sample, which waits for this one's `done`)
            CK(cudaHostAlloc((void**)&ctx->fbHost, (PT_MAX_BOUNCES + 2) * sizeof(uint32_t), cudaHostAllocMapped));
            memset(ctx->fbHost, 0, (PT_MAX_BOUNCES + 2) * sizeof(uint32_t));
            CK(cudaHostGetDevicePointer((void**)&ctx->fbDev, ctx->fbHost, 0));
        }
        if (const char* e = getenv("PTB200_MIN_WIDTH")) ctx->knobMinWidth = (uint32_t)std::max(1, std::min(32, atoi(e)));          // developer knobs, read once
        ctx->knobNoTailWait = getenv("PTB200_NO_TAIL_WAIT") != nullptr;
        ctx->knobNoTailPair = getenv("PTB200_NO_TAIL_PAIR") != nullptr;
        ctx->pdl = !(cfg->flags & PT_FLAG_NO_PDL) && !getenv("PTB200_NO_PDL");
        ctx->knobPdlAlways = getenv("PTB200_PDL_ALWAYS") != nullptr;
        ctx->knobNoMix = getenv("PTB200_NO_MIX") != nullptr;
        if (const char* e = getenv("PTB200_SPLIT")) {                                     // developer knob
            unsigned a = 1, b = 32, c = 1, d = 0;
            sscanf(e, "%u,%u,%u,%u", &a, &b, &c, &d);
            ctx->split = std::max(1u, std::min((unsigned)PT_MAX_SETS, a)); ctx->splitWidth = std::max(1u, std::min(32u, b)); ctx->splitTail = c != 0; ctx->splitHold = d != 0;
        }
        ctx->anyFarFirst = getenv("PTB200_ANY_FAR_FIRST") ? (uint32_t)atoi(getenv("PTB200_ANY_FAR_FIRST")) : 0u;                   // developer knob
        // a whole 1080p image: the head (primary rays + first extend, ~0.3 ms) needs the launches of bounce >= 2 to hide under; on a
        // shard it is short and only disturbs the latency-bound launches, so it waits for the tail (profiles/r2/ab_r2i_ab.log)
        ctx->headAt = m.nSlots > 1500000u ? 2u : (uint32_t)PT_MAX_BOUNCES;
        if (const char* e = getenv("PTB200_HEAD_AT")) ctx->headAt = (uint32_t)std::max(1, atoi(e));                               // developer knob
        if (const char* e = getenv("PTB200_TAIL_PAIR_MAX")) ctx->tailPairMax = (uint32_t)std::max(1, atoi(e));                     // developer knob
        if (const char* e = getenv("PTB200_COOP")) {                                                                               // developer knob
            ctx->coop = ctx->coopTail = (uint32_t)std::max(0, atoi(e));
            if (const char* c = strchr(e, ',')) ctx->coopTail = (uint32_t)std::max(0, atoi(c + 1));
        }
        if (cfg->flags & PT_FLAG_SORT_RAYS) ctx->sortOn = true;
        if (const char* e = getenv("PTB200_SORT")) {                                       // developer knob: "bits[,bounces]", 0 = off
            ctx->sortOn = atoi(e) > 0;
            if (ctx->sortOn) ctx->sortBits = (uint32_t)std::min(PT_SORT_MAX_BITS, std::max(1, atoi(e)));
            if (const char* c = strchr(e, ',')) ctx->sortBounces = (uint32_t)std::min(PT_SORT_MAX_BOUNCES, std::max(1, atoi(c + 1)));
        }
        CK(cudaStreamSynchronize(ctx->stream));
        return PT_OK;
    };
    int r = body();
    if (r) return bail(r);
    *out = ctx;
    return PT_OK;
}

static int makeTexture(pt_ctx* ctx, const float* data, uint32_t w, uint32_t h, int channels, bool linear, cudaTextureObject_t* out) {
    // loadTexture / loadTextureL / loadTextureHDR (use_cuda.h:108-263): normalised coords, wrap, element type
    cudaChannelFormatDesc desc = channels == 4 ? cudaCreateChannelDesc(32, 32, 32, 32, cudaChannelFormatKindFloat)
                                               : cudaCreateChannelDesc(32, 0, 0, 0, cudaChannelFormatKindFloat);
    cudaArray_t arr = nullptr;
    CK(cudaMallocArray(&arr, &desc, w, h));
    ctx->arrays.push_back(arr);
    const size_t pitch = (size_t)w * channels * sizeof(float);
    CK(cudaMemcpy2DToArray(arr, 0, 0, data, pitch, pitch, h, cudaMemcpyHostToDevice));
    cudaResourceDesc res; memset(&res, 0, sizeof(res));
    res.resType = cudaResourceTypeArray; res.res.array.array = arr;
    cudaTextureDesc td; memset(&td, 0, sizeof(td));
    td.normalizedCoords = 1;
    td.addressMode[0] = cudaAddressModeWrap; td.addressMode[1] = cudaAddressModeWrap;
    td.filterMode = linear ? cudaFilterModeLinear : cudaFilterModePoint;
    td.readMode = cudaReadModeElementType;
    CK(cudaCreateTextureObject(out, &res, &td, nullptr));
    ctx->textures.push_back(*out);
    return PT_OK;
}

PT_API int pt_set_skydome(pt_ctx* ctx, const float* rgba, uint32_t w, uint32_t h) {
    if (!ctx) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    if (!rgba) { ctx->tex.sky = 0; ctx->tex.skySample = 0; return PT_OK; }
    { int jr = joinStreams(ctx); if (jr) return jr; }
    cudaTextureObject_t t; int r = makeTexture(ctx, rgba, w, h, 4, true, &t);
    if (r) return r;
    ctx->tex.sky = t;
    // the CDF over the dome's texels, pathtracer.h:85-122: energy = max(r,g,b), running float sum in texel order, both
    // arrays divided by the total afterwards (same operation order, so the same floats as the reference's host loop)
    const size_t n = (size_t)w * h;
    std::vector<float> values(n), cum(n);
    float totalEnergy = 0.0f;
    for (size_t i = 0; i < n; i++) {
        const float energy = fmaxf(rgba[4 * i], fmaxf(rgba[4 * i + 1], rgba[4 * i + 2]));
        values[i] = energy; totalEnergy += energy; cum[i] = totalEnergy;
    }
    for (size_t i = 0; i < n; i++) { values[i] /= totalEnergy; cum[i] /= totalEnergy; }
    CK(cudaStreamSynchronize(ctx->stream));
    if (ctx->skyPdf) { cudaFree(ctx->skyPdf); ctx->skyPdf = nullptr; }
    if (ctx->skyCdf) { cudaFree(ctx->skyCdf); ctx->skyCdf = nullptr; }
    CK(cudaMalloc((void**)&ctx->skyPdf, n * sizeof(float))); CK(cudaMalloc((void**)&ctx->skyCdf, n * sizeof(float)));
    CK(cudaMemcpy(ctx->skyPdf, values.data(), n * sizeof(float), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(ctx->skyCdf, cum.data(), n * sizeof(float), cudaMemcpyHostToDevice));
    ctx->tex.skyPdf = ctx->skyPdf; ctx->tex.skyCdf = ctx->skyCdf; ctx->tex.skyW = w; ctx->tex.skyH = h;
    if (!(totalEnergy > 0.0f)) ctx->tex.skySample = 0;
    return PT_OK;
}
// Opt-in: importance-sample the skydome through the CDF above (the reference builds the CDF and ships
// binarySearch, kernels.h:65-79, but no kernel samples it -- README.md:55).  Off by default, so parity is untouched.
PT_API int pt_set_sky_sampling(pt_ctx* ctx, int on) {
    if (!ctx) return PT_ERR_INVALID;
    if (on && (!ctx->tex.sky || !ctx->skyCdf)) return fail(ctx, PT_ERR_INVALID, "pt_set_sky_sampling: no skydome texture set (pt_set_skydome)");
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    ctx->tex.skySample = on ? 1u : 0u;
    return PT_OK;
}
// Opt-in: sort the extension / shadow queues the first `bounces` shade launches of a sample produce (counting sort on
// origin cell x direction octant, 2^bits cells per axis) before they are traced.  Results are identical; scheduling only.
PT_API int pt_set_ray_sorting(pt_ctx* ctx, int on, uint32_t bits, uint32_t bounces) {
    if (!ctx) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    ctx->sortOn = on != 0;
    if (bits) ctx->sortBits = std::min<uint32_t>(PT_SORT_MAX_BITS, bits);
    if (bounces) ctx->sortBounces = std::min<uint32_t>(PT_SORT_MAX_BOUNCES, bounces);
    return PT_OK;
}
// Measurement mode: every bounce runs as separate wavefront launches of the counting traversal kernels; pt_get_tally returns
// node steps / triangle tests / top-level steps of the extension rays [0..2] and of the shadow rays [3..5] since it was
// switched on.  Images are identical to the normal schedule.
PT_API int pt_set_counting(pt_ctx* ctx, int on) {
    if (!ctx) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    ctx->counting = on != 0;
    if (on) CK(cudaMemsetAsync(ctx->tally, 0, 8 * sizeof(unsigned long long), ctx->stream));
    return PT_OK;
}
PT_API int pt_get_tally(pt_ctx* ctx, uint64_t* out6 /* 8 values */) {
    if (!ctx || !out6) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    unsigned long long h[8];
    CK(cudaMemcpyAsync(h, ctx->tally, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    for (int i = 0; i < 8; i++) out6[i] = h[i];
    return PT_OK;
}
// Test / measurement hook: the per-bounce queue counters of the most recent sample that ran on working set 0 (clearing,
// profiled and counting frames always do): out[b] = extension rays of bounce b, out[max_bounces_cap + b] = shadow rays.
PT_API int pt_debug_read_counters(pt_ctx* ctx, uint32_t* ext_out, uint32_t* shadow_out, uint32_t n) {
    if (!ctx || !ext_out || !shadow_out || n > PT_MAX_BOUNCES + 1) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    uint32_t h[CNT_TOTAL];
    CK(cudaMemcpyAsync(h, ctx->set[0].q.cnt, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    for (uint32_t b = 0; b < n; b++) { ext_out[b] = h[CNT_EXT(b)]; shadow_out[b] = h[CNT_SHADOW(b)]; }
    return PT_OK;
}
PT_API int pt_set_sky_constant(pt_ctx* ctx, float r, float g, float b) {
    if (!ctx) return PT_ERR_INVALID;
    ctx->tex.sky = 0; ctx->tex.skyConst = make_float3(r, g, b);
    return PT_OK;
}
PT_API int pt_set_blue_noise(pt_ctx* ctx, const float* lum, uint32_t w, uint32_t h) {
    if (!ctx) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    if (!lum) { ctx->tex.blue = 0; return PT_OK; }
    cudaTextureObject_t t; int r = makeTexture(ctx, lum, w, h, 1, false, &t);
    if (r) return r;
    ctx->tex.blue = t;
    ctx->tex.invBlueSize = make_float2(1.0f / (float)w, 1.0f / (float)h);      // pathtracer.h:79
    return PT_OK;
}
PT_API int pt_texture_create(pt_ctx* ctx, const float* rgba, uint32_t w, uint32_t h, uint64_t* out_handle) {
    if (!ctx || !rgba || !out_handle) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    cudaTextureObject_t t; int r = makeTexture(ctx, rgba, w, h, 4, true, &t);
    if (r) return r;
    *out_handle = (uint64_t)t;
    return PT_OK;
}
PT_API int pt_set_allreduce_hook(pt_ctx* ctx, pt_allreduce_hook hook, void* user) {
    if (!ctx) return PT_ERR_INVALID;
    ctx->hook = hook; ctx->hookUser = user;
    return PT_OK;
}
PT_API int pt_set_guiding_frames(pt_ctx* ctx, uint32_t frames) {
    if (!ctx) return PT_ERR_INVALID;
    // samples still in flight were enqueued under the old `training` predicate (pt_render): order them before anything
    // enqueued under the new one, so a training frame never reads sample-cache entries of an overlapped sample
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    ctx->cfg.guiding_frames = frames;
    return PT_OK;
}
PT_API int pt_set_toggles(pt_ctx* ctx, int nee, int guiding) {
    if (!ctx) return PT_ERR_INVALID;
    if ((nee != 0) != ctx->HNEE || (guiding != 0) != ctx->HCACHE) {        // HCACHE feeds the `training` predicate: see pt_set_guiding_frames
        CK(cudaSetDevice(ctx->cfg.device));
        int jr = joinStreams(ctx); if (jr) return jr;
    }
    ctx->HNEE = nee != 0; ctx->HCACHE = guiding != 0;
    return PT_OK;
}
// out4: failures since the last report, code of the first one (PT_CHK_*), its index, its limit; a build without PT_BOUNDS_CHECK
// answers PT_ERR_INVALID.  Synchronises the device.  The counters restart at zero.
PT_API int pt_debug_check_report(pt_ctx* ctx, uint32_t* out4) {        // ctx may be NULL: the current device
    if (!out4) return PT_ERR_INVALID;
#if defined(PT_BOUNDS_CHECK)
    if (ctx && cudaSetDevice(ctx->cfg.device) != cudaSuccess) return PT_ERR_CUDA;
    const uint32_t zero[4] = {0, 0, 0, 0};
    if (cudaDeviceSynchronize() != cudaSuccess || cudaMemcpyFromSymbol(out4, g_chkFail, 4 * sizeof(uint32_t)) != cudaSuccess ||
        cudaMemcpyToSymbol(g_chkFail, zero, sizeof(zero)) != cudaSuccess) return PT_ERR_CUDA;
    return PT_OK;
#else
    if (ctx) ctx->err = "pt_debug_check_report: not a bounds-checked build (make variant DEFS=-DPT_BOUNDS_CHECK)";
    return PT_ERR_INVALID;
#endif
}
PT_API void* pt_cuda_stream(pt_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
PT_API int pt_join(pt_ctx* ctx) {
    if (!ctx) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    return joinStreams(ctx);
}

// Host-side validation of what the traversal will index with: instance model ids, top-level leaf ids and child indices
// (one bad record would be an out-of-bounds device read in tlasStep, i.e. a sticky CUDA fault).  Walks the tree from the
// root; returns its depth in edges, or -1 with *why set.
static int checkTopLevel(const pt_instance* inst, uint32_t n, const pt_tlas_node* tl, uint32_t n_nodes, uint32_t numModels, const char** why) {
    for (uint32_t i = 0; i < n; i++) if (inst[i].model_id >= numModels) { *why = "instance model_id out of range"; return -1; }
    if (n == 0 || n_nodes == 0) return 0;
    if (n_nodes > 65536u) { *why = "top-level BVH has more than 65536 nodes (children are 16-bit)"; return -1; }
    std::vector<std::pair<uint32_t, uint32_t>> stack{{0u, 0u}};
    uint32_t visited = 0, deepest = 0;
    while (!stack.empty()) {
        const uint32_t idx = stack.back().first, depth = stack.back().second; stack.pop_back();
        if (++visited > n_nodes) { *why = "top-level BVH is not a tree (a node is reachable twice)"; return -1; }
        deepest = std::max(deepest, depth);
        const pt_tlas_node& nd = tl[idx];
        if (nd.children == 0) { if (nd.leaf >= n) { *why = "top-level BVH leaf names an instance out of range"; return -1; } continue; }
        const uint32_t a = nd.children & 0xffffu, b = (nd.children >> 16) & 0xffffu;
        if (a >= n_nodes || b >= n_nodes) { *why = "top-level BVH child index out of range"; return -1; }
        stack.push_back({a, depth + 1}); stack.push_back({b, depth + 1});
    }
    return (int)deepest;
}

// Test hook: the wide BVH as it sits on the device (80-byte nodes, 48-byte triangle records).  NULL pointers: sizes only.
// Host-only (no GPU, no context): the BVH2 -> 8-wide conversion of ONE model as pt_upload_scene's host converter does it (wide_bvh.cpp).
// nodes / wtris may be NULL (counts only); otherwise *n_nodes / *n_tris say how many entries the buffers hold.  optimal = 0: round 1's
// largest-area collapse.  sah_cost (optional): sum over all children of area / root area x (1 per inner child, triangle count per leaf child).
PT_API int pt_host_wide_bvh(const pt_bvh_node* bvh2, uint32_t n_bvh2, const pt_triangle_v* triangles, int optimal,
                            void* nodes, uint32_t* n_nodes, void* wtris, uint32_t* n_tris, double* sah_cost, uint32_t* max_depth) {
    if (!bvh2 || !n_bvh2 || !triangles || !n_nodes || !n_tris) return PT_ERR_INVALID;
    std::vector<WideNode> vn; std::vector<WideTri> vt; WideBuildStats st;
    buildWideBVH(bvh2, n_bvh2, triangles, vn, vt, &st, optimal != 0);
    if (nodes) { if (*n_nodes < vn.size()) return PT_ERR_INVALID; memcpy(nodes, vn.data(), vn.size() * sizeof(WideNode)); }
    if (wtris) { if (*n_tris < vt.size()) return PT_ERR_INVALID; memcpy(wtris, vt.data(), vt.size() * sizeof(WideTri)); }
    *n_nodes = (uint32_t)vn.size(); *n_tris = (uint32_t)vt.size();
    if (sah_cost) *sah_cost = st.sah_cost;
    if (max_depth) *max_depth = st.max_depth;
    return PT_OK;
}
PT_API int pt_debug_read_wide_bvh(pt_ctx* ctx, void* nodes, uint32_t* n_nodes, void* tris, uint32_t* n_tris) {
    if (!ctx || !ctx->haveScene) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    if (n_nodes) *n_nodes = ctx->numWideNodes;
    if (n_tris) *n_tris = ctx->numWideTris;
    if (nodes && ctx->numWideNodes) CK(cudaMemcpy(nodes, ctx->sc.nodes, (size_t)ctx->numWideNodes * sizeof(WideNode), cudaMemcpyDeviceToHost));
    if (tris && ctx->numWideTris) CK(cudaMemcpy(tris, ctx->sc.wtris, (size_t)ctx->numWideTris * sizeof(WideTri), cudaMemcpyDeviceToHost));
    return PT_OK;
}

static double wallMs() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

PT_API int pt_upload_scene(pt_ctx* ctx, const pt_scene_desc* d) {
    if (!ctx || !d) return PT_ERR_INVALID;
    const bool timing = getenv("PTB200_TIMING") != nullptr;          // developer knob: phase wall times on stderr
    double tPhase = wallMs();
    auto phase = [&](const char* what) { if (timing) { cudaStreamSynchronize(ctx->stream); const double t = wallMs(); fprintf(stderr, "[pt_upload_scene] %-28s %8.2f ms\n", what, t - tPhase); tPhase = t; } };
    if (d->num_models == 0 && d->num_instances > 0) return fail(ctx, PT_ERR_INVALID, "pt_upload_scene: instances without models");
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    CK(cudaStreamSynchronize(ctx->stream));
    // the peers map this rank's radiance-cache block (cudaIpc): freeing it under them would turn their next training frame
    // into stores to freed memory.  Replacing the scene of a connected context needs the collective disconnect first.
    if (ctx->peer.on) return fail(ctx, PT_ERR_INVALID, "pt_upload_scene: peers are connected -- call pt_peer_disconnect on every rank first");
    for (void* p : ctx->sceneAllocs) cudaFree(p);
    ctx->sceneAllocs.clear();
    ctx->rcBlock = nullptr;
    ctx->haveScene = false;
    DScene& sc = ctx->sc;
    memset(&sc, 0, sizeof(sc));
    // Wide BVH per model, one node array / one triangle-record array for the whole scene.  Built on the device from
    // the uploaded BVH2 (wide_bvh_device.cu); PT_FLAG_HOST_WIDE_BVH / PTB200_HOST_WIDE_BVH selects the host converter
    // (wide_bvh.cpp), which produces the same bytes and is what the device converter is tested against.
    std::vector<WideNode> nodes; std::vector<WideTri> wtris; std::vector<DModel> models(d->num_models);
    uint32_t maxWide = 0;
    const bool hostWide = (ctx->cfg.flags & PT_FLAG_HOST_WIDE_BVH) || getenv("PTB200_HOST_WIDE_BVH");
    const bool wideGreedy = getenv("PTB200_WIDE_GREEDY") != nullptr;          // developer knob: round 1's largest-area collapse instead of the SAH-optimal one (A/B)
    float4 *dn = nullptr, *dt = nullptr; pt_triangle_v* dv = nullptr;
    int r;
    for (uint32_t m = 0; m < d->num_models; m++)
        if (!d->models[m].bvh || d->models[m].nrBvhNodes == 0) return fail(ctx, PT_ERR_INVALID, "pt_upload_scene: model without BVH");
    if ((r = devUpload(ctx, &dv, d->vertices, d->num_triangles, true))) return r;
    phase("vertices H2D");
    ctx->numWideNodes = ctx->numWideTris = 0;
    if (hostWide) {
        for (uint32_t m = 0; m < d->num_models; m++) {
            const pt_model& pm = d->models[m];
            WideBuildStats st;
            models[m].rootNode = buildWideBVH(pm.bvh, pm.nrBvhNodes, d->vertices, nodes, wtris, &st, !wideGreedy);
            models[m].triangleStart = pm.triangleStart; models[m].nrTriangles = pm.nrTriangles; models[m].pad = 0;
            maxWide = std::max(maxWide, st.max_depth);
            if (st.tris != pm.nrTriangles) return fail(ctx, PT_ERR_INVALID, "pt_upload_scene: BVH leaves do not cover the model's triangles");
        }
        if ((r = devUpload(ctx, (WideNode**)&dn, nodes.data(), nodes.size(), true))) return r;
        if ((r = devUpload(ctx, (WideTri**)&dt, wtris.data(), wtris.size(), true))) return r;
        ctx->numWideNodes = (uint32_t)nodes.size(); ctx->numWideTris = (uint32_t)wtris.size();
    } else {
        size_t nodeCap = 0, maxBound = 0, maxBvh2 = 0, triCap = 0;
        for (uint32_t m = 0; m < d->num_models; m++) {
            const size_t b = wideNodeBound(d->models[m].nrBvhNodes, d->models[m].nrTriangles);
            nodeCap += b; maxBound = std::max(maxBound, b); maxBvh2 = std::max<size_t>(maxBvh2, d->models[m].nrBvhNodes); triCap += d->models[m].nrTriangles;
        }
        const size_t tmpBytes = wideScanTempBytes(maxBound + 1);
        struct Scratch { std::vector<void*> p; ~Scratch() { for (void* q : p) cudaFree(q); } } scratch;
        auto scratchAlloc = [&](void** out, size_t bytes) { cudaError_t e = cudaMalloc(out, std::max<size_t>(bytes, 256)); if (e == cudaSuccess) scratch.p.push_back(*out); return e; };
        void *tmpNodes, *refs, *counts, *scan, *tmp, *dBvh2, *dpScratch = nullptr;
        if (!wideGreedy) CK(scratchAlloc(&dpScratch, wideDpScratchBytes((uint32_t)maxBvh2)));
        CK(scratchAlloc(&tmpNodes, nodeCap * sizeof(WideNode))); CK(scratchAlloc(&refs, maxBound * 8)); CK(scratchAlloc(&counts, (maxBound + 1) * 8));
        CK(scratchAlloc(&scan, (maxBound + 1) * 8)); CK(scratchAlloc(&tmp, tmpBytes)); CK(scratchAlloc(&dBvh2, maxBvh2 * sizeof(pt_bvh_node)));
        if ((r = devAlloc(ctx, (WideTri**)&dt, triCap, true))) return r;
        uint32_t nodeCursor = 0, triCursor = 0;
        for (uint32_t m = 0; m < d->num_models; m++) {
            const pt_model& pm = d->models[m];
            CK(cudaMemcpyAsync(dBvh2, pm.bvh, (size_t)pm.nrBvhNodes * sizeof(pt_bvh_node), cudaMemcpyHostToDevice, ctx->stream));
            uint32_t nUsed = 0, tUsed = 0, depth = 0; const char* err = "";
            r = buildWideBVHDevice(ctx->stream, (const pt_bvh_node*)dBvh2, pm.nrBvhNodes, dv, (WideNode*)tmpNodes, nodeCursor, (WideTri*)dt, triCursor,
                                   refs, counts, scan, tmp, tmpBytes, &nUsed, &tUsed, &depth, &err, dpScratch);
            if (r) return fail(ctx, r, err);
            if (tUsed != pm.nrTriangles) return fail(ctx, PT_ERR_INVALID, "pt_upload_scene: BVH leaves do not cover the model's triangles");
            models[m].rootNode = nodeCursor; models[m].triangleStart = pm.triangleStart; models[m].nrTriangles = pm.nrTriangles; models[m].pad = 0;
            nodeCursor += nUsed; triCursor += tUsed; maxWide = std::max(maxWide, depth);
        }
        if ((r = devAlloc(ctx, (WideNode**)&dn, nodeCursor, true))) return r;
        CK(cudaMemcpyAsync(dn, tmpNodes, (size_t)nodeCursor * sizeof(WideNode), cudaMemcpyDeviceToDevice, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        ctx->numWideNodes = nodeCursor; ctx->numWideTris = triCursor;
    }
    phase(hostWide ? "wide BVH (host) + H2D" : "wide BVH (device)");
    for (uint32_t m = 0; m < d->num_models; m++)
        if ((uint64_t)d->models[m].triangleStart + d->models[m].nrTriangles > d->num_triangles) return fail(ctx, PT_ERR_INVALID, "pt_upload_scene: model triangle range exceeds num_triangles");
    // unified traversal stack: TLAS entries below the node groups of the instance being traversed
    ctx->maxWideDepth = maxWide;
    {
        const char* why = "";
        const int depth = checkTopLevel(d->instances, d->num_instances, d->topBvh, d->num_tlas_nodes, d->num_models, &why);
        if (depth < 0) { ctx->err = std::string("pt_upload_scene: ") + why; return PT_ERR_INVALID; }
        if (d->num_instances && d->num_tlas_nodes && (uint32_t)depth + maxWide + 2 > PT_STACK)
            return fail(ctx, PT_ERR_INVALID, "pt_upload_scene: top-level + bottom-level BVH deeper than the traversal stack");
    }
    // emissive triangles (pathtracer.h:155-167)
    std::vector<pt_triangle_light> lights;
    for (uint32_t i = 0; i < d->num_instances; i++) {
        const pt_instance& inst = d->instances[i];
        const pt_model& model = d->models[inst.model_id];
        for (uint32_t t = model.triangleStart; t < model.triangleStart + model.nrTriangles; t++) {
            const uint32_t mat_id = inst.material_id == 0xffffffffu ? d->vertexData[t].material : inst.material_id;
            if (mat_id >= d->num_materials) return fail(ctx, PT_ERR_INVALID, "pt_upload_scene: material id out of range");
            const pt_float3 e = d->materials[mat_id].emission;
            if (fmaxf(e.x, fmaxf(e.y, e.z)) < PT_EPS) continue;
            lights.push_back(pt_triangle_light{t, i});
        }
    }
    phase("light scan (host)");
    pt_triangle_d* dd; pt_material* dm; pt_instance* di; pt_tlas_node* dl; DModel* dmo;
    pt_sphere* ds; pt_plane* dp; pt_triangle_light* dli;
    if ((r = devUpload(ctx, &dd, d->vertexData, d->num_triangles, true))) return r;
    if ((r = devUpload(ctx, &dm, d->materials, d->num_materials, true))) return r;
    if ((r = devUpload(ctx, &di, d->instances, d->num_instances, true))) return r;
    if ((r = devUpload(ctx, &dl, d->topBvh, d->num_tlas_nodes, true))) return r;
    if ((r = devUpload(ctx, &dmo, models.data(), models.size(), true))) return r;
    {   // what the device top-level BVH build needs: every model's BVH2 root box (scene.h:88-90)
        std::vector<float4> boxes(2 * (size_t)d->num_models);
        for (uint32_t m = 0; m < d->num_models; m++) {
            const pt_bvh_node& rn = d->models[m].bvh[0];
            boxes[2 * m] = make_float4(rn.vmin.x, rn.vmin.y, rn.vmin.z, 0.f); boxes[2 * m + 1] = make_float4(rn.vmax.x, rn.vmax.y, rn.vmax.z, 0.f);
        }
        if ((r = devUpload(ctx, &ctx->modelBox, boxes.data(), boxes.size(), true))) return r;
        CK(cudaStreamSynchronize(ctx->stream));
        if ((r = devAlloc(ctx, &ctx->tlasAlive, (size_t)d->num_tlas_nodes + 4, true))) return r;
        if ((r = devAlloc(ctx, &ctx->tlasDepth, (size_t)d->num_tlas_nodes + 4, true))) return r;
    }
    if ((r = devUpload(ctx, &ds, d->spheres, d->num_spheres, true))) return r;
    if ((r = devUpload(ctx, &dp, d->planes, d->num_planes, true))) return r;
    if ((r = devUpload(ctx, &dli, lights.data(), lights.size(), true))) return r;
    if ((r = devAlloc(ctx, &sc.lightTab, lights.size() * PT_LIGHT_STRIDE, true))) return r;
    {
        const size_t addBytes = ((size_t)d->num_triangles * 16 * sizeof(float) + 255) & ~(size_t)255;
        ctx->rcAddOffset = 4096; ctx->rcRcOffset = 4096 + addBytes;
        const size_t total = std::max<size_t>(ctx->rcRcOffset + (size_t)d->num_triangles * 3 * sizeof(float4), (size_t)4 << 20);   // >= 4 MB: never sub-allocated, the IPC handle maps exactly this range
        if ((r = devAlloc(ctx, &ctx->rcBlock, total, true))) return r;
        CK(cudaMemsetAsync(ctx->rcBlock, 0, 4096, ctx->stream));
        sc.rcAdditions = reinterpret_cast<float*>(ctx->rcBlock + ctx->rcAddOffset);
        sc.rcRadiance = reinterpret_cast<float4*>(ctx->rcBlock + ctx->rcRcOffset);
    }
    phase("other H2D + allocations");
    sc.nodes = dn; sc.wtris = dt; sc.vertices = dv; sc.vertexData = dd; sc.materials = dm; sc.instances = di; sc.tlas = dl;
    sc.models = dmo; sc.spheres = ds; sc.planes = dp; sc.lights = dli;
    sc.num_triangles = d->num_triangles; sc.num_instances = d->num_instances; sc.num_spheres = d->num_spheres;
    sc.num_planes = d->num_planes; sc.num_lights = (uint)lights.size();
    ctx->numInstancesCap = d->num_instances; ctx->numTlasCap = d->num_tlas_nodes; ctx->numModels = d->num_models;
    { int cr = setCheckLimits(ctx, d->num_tlas_nodes, d->num_models, d->num_materials); if (cr) return cr; }
    {   // ray-sorting cells span the top-level root box and the spheres (planes are unbounded: origins outside are clamped)
        float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
        if (d->num_instances && d->num_tlas_nodes) {
            const pt_tlas_node& rt = d->topBvh[0];
            const float a[3] = {rt.vmin.x, rt.vmin.y, rt.vmin.z}, b[3] = {rt.vmax.x, rt.vmax.y, rt.vmax.z};
            for (int k = 0; k < 3; k++) { lo[k] = fminf(lo[k], a[k]); hi[k] = fmaxf(hi[k], b[k]); }
        }
        for (uint32_t i = 0; i < d->num_spheres; i++) {
            const float c[3] = {d->spheres[i].pos.x, d->spheres[i].pos.y, d->spheres[i].pos.z};
            for (int k = 0; k < 3; k++) { lo[k] = fminf(lo[k], c[k] - d->spheres[i].radius); hi[k] = fmaxf(hi[k], c[k] + d->spheres[i].radius); }
        }
        float sc3[3];
        for (int k = 0; k < 3; k++) { if (!(hi[k] > lo[k])) { lo[k] = 0.0f; hi[k] = 1.0f; } sc3[k] = 1.0f / (hi[k] - lo[k]); }
        sc.sortLo = make_float3(lo[0], lo[1], lo[2]);
        sc.sortScale = make_float3(sc3[0], sc3[1], sc3[2]);
    }
    if (d->num_triangles) {
        k_rc_init<<<(d->num_triangles + 255) / 256, 256, 0, ctx->stream>>>(sc.rcRadiance, sc.rcAdditions, d->num_triangles);   // pathtracer.h:220
        CK(cudaGetLastError());
    }
    if (sc.num_lights) { k_prepare_lights<<<(sc.num_lights + 127) / 128, 128, 0, ctx->stream>>>(sc); CK(cudaGetLastError()); }
    CK(cudaStreamSynchronize(ctx->stream));      // host vectors above go out of scope
    phase("cache init + light table");
    ctx->haveScene = true;
    return (int)lights.size();
}

PT_API int pt_update_instances(pt_ctx* ctx, const pt_instance* inst, uint32_t n, const pt_tlas_node* tlas, uint32_t n_nodes) {
    if (!ctx || !ctx->haveScene) return PT_ERR_INVALID;
    if (n != ctx->numInstancesCap || n_nodes != ctx->numTlasCap) return fail(ctx, PT_ERR_INVALID, "pt_update_instances: counts differ from the uploaded scene");
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    if (!inst || !tlas) return PT_ERR_INVALID;
    {
        const char* why = "";
        const int depth = checkTopLevel(inst, n, tlas, n_nodes, ctx->numModels, &why);
        if (depth < 0) { ctx->err = std::string("pt_update_instances: ") + why; return PT_ERR_INVALID; }
        if ((uint32_t)depth + ctx->maxWideDepth + 2 > PT_STACK) return fail(ctx, PT_ERR_INVALID, "pt_update_instances: top-level BVH too deep");
    }
    // pathtracer.h:245-246; pageable sources are staged before the call returns
    CK(cudaMemcpyAsync((void*)ctx->sc.instances, inst, n * sizeof(pt_instance), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync((void*)ctx->sc.tlas, tlas, n_nodes * sizeof(pt_tlas_node), cudaMemcpyHostToDevice, ctx->stream));
    if (ctx->sc.num_lights) { k_prepare_lights<<<(ctx->sc.num_lights + 127) / 128, 128, 0, ctx->stream>>>(ctx->sc); CK(cudaGetLastError()); }   // per-light NEE terms follow the instance transforms
    return PT_OK;
}

// Scene::update's instance half on the host, its BuildTopLevelBVH half (scene.h:81-118) on the device: the caller hands
// over the instance matrices only; the top-level BVH is rebuilt in stream order by k_build_tlas (no host work, no
// synchronisation unless the tree could outgrow the traversal stack).
PT_API int pt_update_instances_device_tlas(pt_ctx* ctx, const pt_instance* inst, uint32_t n) {
    if (!ctx || !ctx->haveScene || !inst) return PT_ERR_INVALID;
    if (n == 0 || n != ctx->numInstancesCap || 2 * n - 1 != ctx->numTlasCap) return fail(ctx, PT_ERR_INVALID, "pt_update_instances_device_tlas: counts differ from the uploaded scene");
    if (n > 32768u) return fail(ctx, PT_ERR_INVALID, "pt_update_instances_device_tlas: more than 32768 instances (top-level children are 16-bit node indices)");
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    for (uint32_t i = 0; i < n; i++) if (inst[i].model_id * 2 + 1 >= ctx->numModels * 2) return fail(ctx, PT_ERR_INVALID, "pt_update_instances_device_tlas: instance model_id out of range");
    CK(cudaMemcpyAsync((void*)ctx->sc.instances, inst, n * sizeof(pt_instance), cudaMemcpyHostToDevice, ctx->stream));
    k_build_tlas<<<1, 256, 0, ctx->stream>>>(ctx->sc.instances, ctx->sc.models, ctx->modelBox, n, (pt_tlas_node*)ctx->sc.tlas, ctx->tlasAlive, ctx->tlasDepth, ctx->tlasDepth + ctx->numTlasCap);
    CK(cudaGetLastError());
    if (n - 1 + ctx->maxWideDepth + 2 > PT_STACK) {          // a chain of n leaves is n-1 deep: only then the depth has to be looked at
        uint32_t depth = 0;
        CK(cudaMemcpyAsync(&depth, ctx->tlasDepth + ctx->numTlasCap, sizeof depth, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        if (depth + ctx->maxWideDepth + 2 > PT_STACK) return fail(ctx, PT_ERR_INVALID, "pt_update_instances_device_tlas: top-level BVH too deep");
    }
    if (ctx->sc.num_lights) { k_prepare_lights<<<(ctx->sc.num_lights + 127) / 128, 128, 0, ctx->stream>>>(ctx->sc); CK(cudaGetLastError()); }
    return PT_OK;
}

// Test hook: the top-level BVH as it sits on the device.
PT_API int pt_debug_read_tlas(pt_ctx* ctx, pt_tlas_node* out, uint32_t n_nodes) {
    if (!ctx || !ctx->haveScene || !out || n_nodes != ctx->numTlasCap) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    CK(cudaMemcpyAsync(out, ctx->sc.tlas, (size_t)n_nodes * sizeof(pt_tlas_node), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return PT_OK;
}

static DCamera toDCamera(const pt_camera* c) {
    DCamera d;
    d.eye = make_float3(c->eye.x, c->eye.y, c->eye.z); d.viewDir = make_float3(c->viewDir.x, c->viewDir.y, c->viewDir.z);
    d.lt = make_float3(c->lt.x, c->lt.y, c->lt.z); d.u = make_float3(c->u.x, c->u.y, c->u.z); d.v = make_float3(c->v.x, c->v.y, c->v.z);
    d.d = c->d; d.focalLength = c->focalLength; d.aperture = c->aperture;
    return d;
}

// Ray sorting buffers (bin keys + fetch permutation per queue entry, both queues) are allocated when sorting is first enabled.
static int ensureSortBuffers(pt_ctx* ctx) {
    const size_t cap = std::max<uint32_t>(ctx->map.nSlots, 1024u);
    for (int k = 0; k < ctx->nSets; k++) {
        pt_ctx::SampleSet& S = ctx->set[k];
        for (int qd = 0; qd < 2; qd++) {
            int r;
            if (!S.sortKeys[qd] && (r = devAlloc(ctx, &S.sortKeys[qd], cap, false))) return r;
            if (!S.sortPerm[qd] && (r = devAlloc(ctx, &S.sortPerm[qd], cap, false))) return r;
        }
    }
    return PT_OK;
}

// Pathtracer::Render, pathtracer.h:224-302.  Asynchronous: only enqueues work.
PT_API int pt_render(pt_ctx* ctx, const pt_camera* cam, int should_clear) {
    if (!ctx || !cam) return PT_ERR_INVALID;
    if (!ctx->haveScene) return fail(ctx, PT_ERR_INVALID, "pt_render: no scene uploaded");
    CK(cudaSetDevice(ctx->cfg.device));
    const TileMap& map = ctx->map;
    const DCamera dcam = toDCamera(cam);
    const uint32_t npix = map.W * map.H;
    // Clearing frames, profiled frames and counting frames run alone on the public stream; converge samples are pipelined.
    const bool pipelined = !should_clear && !ctx->prof && !ctx->counting && ctx->nSets >= 2;
    if (!pipelined) { int jr = joinStreams(ctx); if (jr) return jr; }
    else if (!ctx->forked) {
        CK(cudaEventRecord(ctx->forkEv, ctx->stream));
        for (int k = 0; k < ctx->nSets; k++) CK(cudaStreamWaitEvent(ctx->set[k].stream, ctx->forkEv, 0));
        ctx->forked = true;
    }
    if (ctx->sortOn) { int sr = ensureSortBuffers(ctx); if (sr) return sr; }
    ctx->pdlNow = ctx->pdl;
    if (should_clear) {
        ctx->DNEE = ctx->HNEE; ctx->DCACHE = ctx->HCACHE;                                  // :242-243
        launchK(ctx, k_clear_image, dim3((npix + 255) / 256), dim3(256), ctx->stream, ctx->image, ctx->albedo, npix);    // :248-249
        ctx->sampleIdx = 0; ctx->randIdx = 0;                                              // :250-251
    }
    const uint32_t nsamples = should_clear ? ctx->cfg.interactive_depth : 1;               // :255
    const bool deviceNEE = ctx->DNEE && ctx->sc.num_lights > 0;                            // globals.h:23
    const bool skyNEE = ctx->tex.skySample && ctx->tex.sky;                                // opt-in skydome sampling: shadow rays towards the sky
    const uint32_t nbins = 8u << (3 * ctx->sortBits);
    ctx->sc.sortBits = ctx->sortBits;
    for (uint32_t sample = 0; sample < nsamples; sample++) {
        const bool training = !should_clear && ctx->HCACHE && ctx->sampleIdx < ctx->cfg.guiding_frames && ctx->sc.num_triangles;   // :292 (a rank without pixels still takes part in the exchange)
        // A cache-training frame cannot overlap its neighbours (each reads the cache the one before it wrote), so it overlaps
        // ITSELF: its tiles run as `parts` independent wavefront chains on the working sets' streams, each chain's short launches
        // and drain phases filled by the others; the chains meet at the cache update.  Paths of one frame never read each other's
        // results (additions are atomics applied by the propagate step), so per-pixel results are the unsplit frame's.
        const uint32_t parts = (pipelined && training && ctx->split > 1u) ? std::min(std::min(ctx->split, (uint32_t)ctx->nSets), std::max(1u, (uint32_t)ctx->map.nLocalTiles)) : 1u;
        const bool splitNow = parts > 1u;
        const int firstSet = pipelined ? ctx->cur : 0;
        pt_ctx::SampleSet& other = ctx->set[pipelined ? (ctx->cur + ctx->nSets - 1) % ctx->nSets : 0];     // the previous sample's (last) set
        if (pipelined) ctx->cur = (ctx->cur + (int)parts) % ctx->nSets;
        ctx->lastSet = pipelined ? (firstSet + (int)parts - 1) % ctx->nSets : -1;
        const uint32_t randIdx0 = ctx->randIdx;
      for (uint32_t part = 0; part < parts; part++) {
        pt_ctx::SampleSet& S = ctx->set[pipelined ? (firstSet + (int)part) % ctx->nSets : 0];
        cudaStream_t s = pipelined ? S.stream : ctx->stream;
        const bool lastPart = part + 1 == parts;
        TileMap map = ctx->map;                                                            // this part's tile range
        const uint32_t tile0 = (uint32_t)((uint64_t)map.nLocalTiles * part / parts), tile1 = (uint32_t)((uint64_t)map.nLocalTiles * (part + 1) / parts);
        map.slotBase = tile0 * 1024u;
        const uint32_t partSlots = (tile1 - tile0) * 1024u;
        ctx->randIdx = randIdx0;                                                           // every part draws the frame's numbers
        // Short queues: a sample that runs alone (clearing / profiled / cache-training frames) spreads its rays over all
        // warps (latency); samples that overlap others keep their warps full (throughput: 0.41 -> 0.32 ms/sample on a
        // 1/8-image shard), the other samples in flight fill the machine.
        ctx->sc.minWidth = (pipelined && !training && !ctx->prevTraining) ? 32u : 1u;
        if (splitNow) ctx->sc.minWidth = ctx->splitWidth;
        if (ctx->knobMinWidth) ctx->sc.minWidth = ctx->knobMinWidth;
        ctx->sc.tailPair = (ctx->sc.minWidth < 32u && !ctx->knobNoTailPair) ? ctx->tailPairMax : 0u;     // latency mode only: a sample that runs alone
        ctx->sc.coop = ctx->sc.minWidth < 32u ? ctx->coop : 0u;                            // likewise: overlapped samples fill short launches with each other's rays
        ctx->sc.coopTail = ctx->sc.minWidth < 32u ? ctx->coopTail : 0u;
        ctx->sc.mixPair = (ctx->sc.minWidth < 32u && !ctx->knobNoMix) ? 1u : 0u;
        ctx->sc.anyFarFirst = ctx->anyFarFirst;
        // A sample that cannot overlap the previous one's shading (cache training) holds its cache-independent head (primary
        // rays + first extend) back until the previous sample reaches its tail, whose low occupancy it then fills, instead
        // of competing with the previous sample's wide launches.
        if (pipelined && (training || ctx->prevTraining) && other.tailValid && !ctx->knobNoTailWait && (!splitNow || ctx->splitHold)) CK(cudaStreamWaitEvent(s, other.tailStart, 0));
        const bool overlapped = pipelined && !training && !ctx->prevTraining;
        const bool crowded = overlapped || splitNow;                                       // other launches of this or other samples are in flight: schedule for throughput
        // Programmatic dependent launch pays where a sample runs alone (its launch chain is the critical path: measured 1-2 % per
        // training frame).  Among overlapped samples an early-scheduled successor holds SM slots the other samples' kernels
        // could use (measured: steady state 4 % slower at 1080p, 18 % on a 1/8 shard), so those launch in plain stream order.
        ctx->pdlNow = ctx->pdl && (!crowded || ctx->knobPdlAlways);
        const bool sorting = ctx->sortOn && !ctx->counting;
        CK(cudaMemsetAsync(S.q.cnt, 0, (CNT_TOTAL + (sorting ? (size_t)ctx->sortBounces * 2 * nbins : 0)) * sizeof(uint), s));   // :265 (+ cursors, + sort histograms)
        if (partSlots) { StageTimer tm(ctx, PT_STAGE_OTHER, s); launchK(ctx, k_primary, dim3(partSlots / 256), dim3(256), s, dcam, map, ctx->randIdx, S.q, S.masks, S.accucolors); }   // :258,:266
        ctx->randIdx++;                                                                    // :267
        uint32_t max_bounces;                                                              // :270-274 (host _NEE == HNEE)
        if (ctx->HNEE) max_bounces = should_clear ? ctx->cfg.interactive_depth : ctx->cfg.max_ray_depth;
        else max_bounces = should_clear ? ctx->cfg.interactive_depth + 1 : ctx->cfg.max_ray_depth;
        // k_tail (one per-path launch for every bounce >= tailBounce) shortens the dependent-launch chain of a sample that
        // runs alone (clearing / profiled / cache-training frames: 0.95 -> 0.72 ms on a 1/8 shard).  Samples that overlap
        // others stay wavefront to the last bounce: their short launches are filled in by the neighbours (measured
        // 1.77 -> 1.58 ms/sample at 1080p).  Either way the per-path operation order is the same: identical images.
        uint32_t tailFrom = ctx->fbTail ? ctx->fbTail : ctx->tailBounce;                   // the last decision stands until newer queue sizes have landed (the host runs many samples ahead)
        if (ctx->fbOn && !overlapped && !should_clear) {
            // newest finished sample whose queue sizes have landed: the tail starts at the first bounce with few enough paths for the
            // per-path kernel (measured optimum on 1/1 .. 1/8 of 1080p: where the queue has fallen to 1.2-2.5 k paths, profiles/r2/ab_r2g_ab.log)
            const uint32_t few = (uint32_t)ctx->gridTail * 4u * 3u / 2u;
            const volatile uint32_t* c = ctx->fbHost;
            const uint32_t seq = c[0];
            if (seq != ctx->fbSeen) {                         // a newer sample has finished (a torn read only costs one sample a worse schedule)
                tailFrom = max_bounces;
                for (uint32_t b = 1; b < max_bounces; b++) if (c[1 + b] <= few) { tailFrom = b; break; }
                ctx->fbTail = tailFrom; ctx->fbSeen = seq;
            }
        }
        const uint32_t waveBounces = ctx->counting ? max_bounces : std::min(max_bounces, (((overlapped || (splitNow && !ctx->splitTail))) && !ctx->tailForced) ? (uint32_t)PT_MAX_BOUNCES : tailFrom);
        const bool connects = (ctx->HNEE && deviceNEE) || skyNEE;                          // :286 (no shadow rays exist when the device _NEE is false)
        const bool fuse = connects && !ctx->prof && !ctx->counting;                        // connect(b) rides in the launch of extend(b+1)
        // ray sorting: shade(b) bins the rays it pushes, the two small kernels after it turn the bins into fetch permutations
        auto sortsExt = [&](uint32_t bounce) { return sorting && bounce < ctx->sortBounces && bounce + 1 < waveBounces; };
        auto sortsSh = [&](uint32_t bounce) { return sorting && bounce < ctx->sortBounces && connects; };
        auto histOf = [&](uint32_t bounce, int queue) { return S.sortHist + ((size_t)bounce * 2 + queue) * nbins; };
        auto shadeParams = [&](uint32_t bounce) {
            ShadeParams P;
            P.sc = ctx->sc; P.map = map; P.tex = ctx->tex;
            P.masks = S.masks; P.accucolors = S.accucolors; P.lights = S.lights;
            P.sampleCache = S.sampleCache; P.albedoSample = S.albedoSample;
            P.sampleIdx = ctx->sampleIdx; P.randIdx = ctx->randIdx; P.blueFrames = ctx->cfg.blue_noise_frames;
            P.bounce = (int)bounce; P.nee = deviceNEE; P.cache = ctx->DCACHE;
            // Overlapped (non-training) samples do not write the sample cache: nothing reads it until the cache trains again
            // after a clearing frame, and then the entries a path does not rewrite (emissive hits, kernels.h:563-576) must be
            // those of ONE earlier frame -- here the last clearing / training frame -- not of whichever of the samples in
            // flight wrote last.  (The reference, one sample at a time, leaves the last converge sample's entries there.)
            P.cacheWrite = !overlapped;
            P.sortKeys[0] = sortsExt(bounce) ? S.sortKeys[0] : nullptr; P.sortHist[0] = sortsExt(bounce) ? histOf(bounce, 0) : nullptr;
            P.sortKeys[1] = sortsSh(bounce) ? S.sortKeys[1] : nullptr; P.sortHist[1] = sortsSh(bounce) ? histOf(bounce, 1) : nullptr;
            return P;
        };
        auto launchConnect = [&](uint32_t bounce) {
            StageTimer tm(ctx, PT_STAGE_CONNECT, s);
            if (ctx->counting)
                launchK(ctx, k_trace_count<true>, dim3(ctx->gridTraceAny), dim3(128), s, ctx->sc, S.q.shO, S.q.shD, S.q.cnt + CNT_SHADOW(bounce),
                        S.q.cnt + CUR_CONN(bounce), nullptr, S.lights, S.accucolors, ctx->tally + 3);
            else
                launchK(ctx, k_trace<true>, dim3(ctx->gridTraceAny), dim3(128), s, ctx->sc, S.q.shO, S.q.shD, S.q.cnt + CNT_SHADOW(bounce),
                        S.q.cnt + CUR_CONN(bounce), nullptr, S.lights, S.accucolors, sortsSh(bounce) ? S.sortPerm[1] : nullptr);
        };
        // From here on this sample's launches are short (few rays, latency bound): the next sample's cache-independent head (primary
        // rays + first extend) is released to fill the machine under them -- at bounce headAt, or at the tail if that comes first.
        S.tailValid = false;
        auto releaseHead = [&]() -> int { if (pipelined && !S.tailValid) { CK(cudaEventRecord(S.tailStart, s)); S.tailValid = true; } return PT_OK; };
        for (uint32_t bounce = 0; bounce < waveBounces; bounce++) {                        // :276-290
            const int buf = bounce & 1;
            if (bounce == ctx->headAt && !overlapped && lastPart) { int hr = releaseHead(); if (hr) return hr; }
            const uint* extPerm = (bounce > 0 && sortsExt(bounce - 1)) ? S.sortPerm[0] : nullptr;
            if (fuse && bounce > 0 && ctx->sc.mixPair) {
                StageTimer tm(ctx, PT_STAGE_EXTEND, s);
                launchK(ctx, k_trace_pair_mixed, dim3(ctx->gridTraceExt), dim3(128), s, ctx->sc, S.q.rayO[buf], S.q.rayD[buf], S.q.cnt + CNT_EXT(bounce), S.q.cnt + CUR_EXT(bounce), S.q.hits,
                        S.q.shO, S.q.shD, S.q.cnt + CNT_SHADOW(bounce - 1), S.lights, S.accucolors,
                        extPerm, sortsSh(bounce - 1) ? S.sortPerm[1] : nullptr);
            } else if (fuse && bounce > 0) {
                StageTimer tm(ctx, PT_STAGE_EXTEND, s);
                launchK(ctx, k_trace_pair, dim3(ctx->gridTraceExt), dim3(128), s, ctx->sc, S.q.rayO[buf], S.q.rayD[buf], S.q.cnt + CNT_EXT(bounce), S.q.cnt + CUR_EXT(bounce), S.q.hits,
                        S.q.shO, S.q.shD, S.q.cnt + CNT_SHADOW(bounce - 1), S.q.cnt + CUR_CONN(bounce - 1), S.lights, S.accucolors,
                        extPerm, sortsSh(bounce - 1) ? S.sortPerm[1] : nullptr);
            } else if (ctx->counting) {
                launchK(ctx, k_trace_count<false>, dim3(ctx->gridTraceExt), dim3(128), s, ctx->sc, S.q.rayO[buf], S.q.rayD[buf], S.q.cnt + CNT_EXT(bounce),
                        S.q.cnt + CUR_EXT(bounce), S.q.hits, nullptr, nullptr, ctx->tally);
            } else {
                StageTimer tm(ctx, PT_STAGE_EXTEND, s);
                launchK(ctx, k_trace<false>, dim3(ctx->gridTraceExt), dim3(128), s, ctx->sc, S.q.rayO[buf], S.q.rayD[buf], S.q.cnt + CNT_EXT(bounce),
                        S.q.cnt + CUR_EXT(bounce), S.q.hits, nullptr, nullptr, extPerm);
            }
            // the previous sample's propagate step rewrites the radiance cache (and shares the sample cache) shading uses:
            // while the cache trains, only primary rays + the first extend overlap the previous sample
            if (bounce == 0 && pipelined && (ctx->prevTraining || training) && other.doneValid) CK(cudaStreamWaitEvent(s, other.done, 0));
            { StageTimer tm(ctx, PT_STAGE_SHADE, s); launchK(ctx, k_shade, dim3(ctx->gridShade), dim3(128), s, shadeParams(bounce), S.q, buf); }
            if (sortsExt(bounce) || sortsSh(bounce)) {
                StageTimer tm(ctx, PT_STAGE_OTHER, s);
                const bool e = sortsExt(bounce), c = sortsSh(bounce);
                launchK(ctx, k_sort_scan, dim3(2), dim3(1024), s, e ? histOf(bounce, 0) : nullptr, c ? histOf(bounce, 1) : nullptr, nbins);
                launchK(ctx, k_sort_scatter, dim3(ctx->numSMs * 4), dim3(256), s, e ? S.sortKeys[0] : nullptr, e ? histOf(bounce, 0) : nullptr, S.sortPerm[0], S.q.cnt + CNT_EXT(bounce + 1),
                        c ? S.sortKeys[1] : nullptr, c ? histOf(bounce, 1) : nullptr, S.sortPerm[1], S.q.cnt + CNT_SHADOW(bounce));
            }
            ctx->randIdx++;                                                                // :283
            if (connects && (!fuse || bounce + 1 == waveBounces)) launchConnect(bounce);
        }
        if (max_bounces > waveBounces) {                                                   // the same bounces, one launch (k_tail)
            { int hr = releaseHead(); if (hr) return hr; }
            StageTimer tm(ctx, PT_STAGE_TAIL, s);
            launchK(ctx, k_tail, dim3(ctx->gridTail), dim3(128), s, shadeParams(waveBounces), S.q, (int)(waveBounces & 1), (int)waveBounces, (int)max_bounces,
                    S.q.cnt + CUR_EXT(waveBounces));
            ctx->randIdx += max_bounces - waveBounces;                                     // :283, once per bounce
        }
        // image / albedo / statistics / radiance-cache updates happen in sample order
        if (pipelined && other.doneValid) CK(cudaStreamWaitEvent(s, other.done, 0));
        auto accumulate = [&]() -> int {
            if (ctx->copyPending && pipelined) CK(cudaStreamWaitEvent(s, ctx->imgFree, 0));    // a read-back in flight still reads the image
            if (partSlots) { StageTimer tm(ctx, PT_STAGE_OTHER, s);
                              launchK(ctx, k_accumulate, dim3(partSlots / 256), dim3(256), s, map, S.accucolors, S.albedoSample, ctx->image, ctx->albedo, S.q.cnt, ctx->stats, max_bounces, waveBounces,
                                      (ctx->fbOn && !overlapped && !should_clear && part == 0 && max_bounces == ctx->cfg.max_ray_depth) ? ctx->fbDev : nullptr); }   // :299
            return PT_OK;
        };
        bool accumulated = false;
        if (training) {                                                                    // :292-296
            StageTimer tm(ctx, PT_STAGE_GUIDING, s);
            if (partSlots) launchK(ctx, k_rc_update, dim3(partSlots / 256), dim3(256), s, map, S.accucolors, S.sampleCache, ctx->sc.rcAdditions, ctx->sc.num_triangles);
            if (splitNow && !lastPart) {                                                   // the last part's stream collects the others and runs the exchange
                CK(cudaEventRecord(S.partDone, s));
                int ar = accumulate(); if (ar) return ar;
                CK(cudaEventRecord(S.done, s)); S.doneValid = true;
                continue;
            }
            if (splitNow) for (uint32_t p = 0; p + 1 < parts; p++) CK(cudaStreamWaitEvent(s, ctx->set[(firstSet + (int)p) % ctx->nSets].partDone, 0));
            if (ctx->peer.on) {
                // reduce + propagate + broadcast over peer memory (pt_kernels.cuh): the owner of a triangle range sums every
                // rank's additions out of their HBM, propagates once and stores the result into every rank's cache
                const PeerSet& ps = ctx->peer.ps;
                const uint32_t T = ctx->sc.num_triangles, epoch = ++ctx->peer.epoch;
                const uint32_t lo = (uint32_t)((uint64_t)T * ps.rank / ps.world), hi = (uint32_t)((uint64_t)T * (ps.rank + 1) / ps.world);
                launchK(ctx, k_peer_arrive, dim3(1), dim3(32), s, ps, epoch);
                if (!ctx->prof) { int ar = accumulate(); if (ar) return ar; accumulated = true; }   // the image update fills the wait for the slowest rank
                launchK(ctx, k_peer_reduce_propagate, dim3(std::max(1u, (hi - lo + 255u) / 256u)), dim3(256), s, ps, epoch, T, lo, hi, ctx->peer.timeoutNs);
                launchK(ctx, k_peer_finish, dim3((unsigned)std::min<size_t>(296, ((size_t)16 * T + 255) / 256)), dim3(256), s, ps, epoch, T, ctx->peer.timeoutNs);
            } else {
                if (ctx->hook) {
                    int hr = ctx->hook(ctx->hookUser, ctx->sc.rcAdditions, (size_t)16 * ctx->sc.num_triangles, (void*)s);
                    if (hr) return fail(ctx, PT_ERR_INVALID, "pt_render: all-reduce hook failed");
                }
                launchK(ctx, k_rc_propagate, dim3((ctx->sc.num_triangles + 255) / 256), dim3(256), s, ctx->sc.rcRadiance, ctx->sc.rcAdditions, ctx->sc.num_triangles);
            }
        }
        if (!accumulated) { int ar = accumulate(); if (ar) return ar; }
        if (splitNow) for (uint32_t p = 0; p + 1 < parts; p++) CK(cudaStreamWaitEvent(s, ctx->set[(firstSet + (int)p) % ctx->nSets].done, 0));   // their image updates
        if (pipelined) { CK(cudaEventRecord(S.done, s)); S.doneValid = true; }
      }   // parts
        ctx->prevTraining = training;
        ctx->sampleIdx++;                                                                  // :300
    }
    CK(cudaGetLastError());
    return PT_OK;
}

PT_API int pt_finish(pt_ctx* ctx) {                                                        // pathtracer.h:305-310
    if (!ctx) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    CK(cudaStreamSynchronize(ctx->stream));
    resolveProfile(ctx);
    if (ctx->peer.on) {
        uint32_t e = 0;
        CK(cudaMemcpy(&e, ctx->rcBlock + PT_PEER_ERROR * sizeof(uint32_t), sizeof(e), cudaMemcpyDeviceToHost));
        if (e) {                                                                           // reported once: the word is cleared so that a later exchange starts clean
            CK(cudaMemset(ctx->rcBlock + PT_PEER_ERROR * sizeof(uint32_t), 0, sizeof(uint32_t)));
            return fail(ctx, PT_ERR_CUDA, "pt_finish: a peer did not reach the radiance-cache exchange in time (ranks must render the same frames); that exchange was skipped on this rank");
        }
    }
    return PT_OK;
}

// ---- multi-GPU exchange over NVLink peer memory --------------------------------------------------------------------------
// pt_peer_export: the cudaIpc handle of this rank's radiance-cache block (flags | additions | cache).  The caller gathers the
// handles of all ranks (any transport: torch.distributed, MPI, a file) and passes them, in rank order, to pt_peer_connect,
// which maps every peer's block and switches the training frames from "hook + propagate" to the fused peer kernels.
PT_API int pt_peer_export(pt_ctx* ctx, void* handle_out) {
    if (!ctx || !handle_out) return PT_ERR_INVALID;
    if (!ctx->haveScene || !ctx->rcBlock) return fail(ctx, PT_ERR_INVALID, "pt_peer_export: no scene uploaded");
    static_assert(2 * sizeof(cudaIpcMemHandle_t) == PT_PEER_HANDLE_BYTES, "handle size");
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    CK(cudaStreamSynchronize(ctx->stream));
    peerClose(ctx);
    uint32_t flags[1024] = {};
    flags[PT_PEER_MAGIC] = 0x50544230u + ctx->cfg.rank;
    CK(cudaMemcpy(ctx->rcBlock, flags, sizeof(flags), cudaMemcpyHostToDevice));
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, ctx->rcBlock));
    memcpy(handle_out, &h, sizeof(h));
    // second handle: the framebuffer-gather block.  Rank 0 owns the image ring every rank pushes its tiles into
    // (pt_gather_image_async); the other ranks export a minimal block so that the handle layout is uniform.
    if (!ctx->gatherBlock) {
        ctx->gatherBytes = std::max<size_t>(ctx->cfg.rank == 0 ? (size_t)PT_GATHER_RING * ctx->map.W * ctx->map.H * sizeof(float4) : 0, (size_t)4 << 20);
        CK(cudaMalloc((void**)&ctx->gatherBlock, ctx->gatherBytes));
        CK(cudaMemset(ctx->gatherBlock, 0, ctx->gatherBytes));
    }
    ctx->gatherEpoch = 0;
    CK(cudaIpcGetMemHandle(&h, ctx->gatherBlock));
    memcpy((unsigned char*)handle_out + sizeof(h), &h, sizeof(h));
    return PT_OK;
}
PT_API int pt_peer_connect(pt_ctx* ctx, const void* handles, uint32_t world) {
    if (!ctx || !handles) return PT_ERR_INVALID;
    if (!ctx->haveScene || !ctx->rcBlock) return fail(ctx, PT_ERR_INVALID, "pt_peer_connect: no scene uploaded");
    if (world != ctx->cfg.world_size || world > PT_PEER_MAX) return fail(ctx, PT_ERR_INVALID, "pt_peer_connect: world size differs from pt_config.world_size (or exceeds 16)");
    CK(cudaSetDevice(ctx->cfg.device));
    peerClose(ctx);
    PeerSet ps; memset(&ps, 0, sizeof(ps));
    ps.rank = ctx->cfg.rank; ps.world = world; ps.addOffset = ctx->rcAddOffset; ps.rcOffset = ctx->rcRcOffset;
    for (uint32_t r = 0; r < world; r++) {
        if (r == ctx->cfg.rank) { ps.block[r] = ctx->rcBlock; continue; }
        cudaIpcMemHandle_t h; memcpy(&h, (const unsigned char*)handles + (size_t)r * PT_PEER_HANDLE_BYTES, sizeof(h));
        void* p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        uint32_t magic = 0;
        if (e == cudaSuccess) { ctx->peer.opened[r] = p; e = cudaMemcpy(&magic, (unsigned char*)p + PT_PEER_MAGIC * sizeof(uint32_t), sizeof(magic), cudaMemcpyDeviceToHost); }
        if (e != cudaSuccess || magic != 0x50544230u + r) {
            cudaGetLastError(); peerClose(ctx);
            char buf[256]; snprintf(buf, sizeof(buf), "pt_peer_connect: cannot map the block of rank %u (%s)", r, e != cudaSuccess ? cudaGetErrorString(e) : "wrong contents");
            return fail(ctx, PT_ERR_CUDA, buf);
        }
        ps.block[r] = (unsigned char*)p;
    }
    if (const char* t = getenv("PTB200_PEER_TIMEOUT_MS")) ctx->peer.timeoutNs = (unsigned long long)(atof(t) * 1e6);
    if (ctx->cfg.rank == 0) ctx->gatherRoot = reinterpret_cast<float4*>(ctx->gatherBlock);
    else {                                           // rank 0's image ring
        cudaIpcMemHandle_t h; memcpy(&h, (const unsigned char*)handles + sizeof(h), sizeof(h));
        void* p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) { cudaGetLastError(); peerClose(ctx); return fail(ctx, PT_ERR_CUDA, "pt_peer_connect: cannot map the framebuffer ring of rank 0"); }
        ctx->gatherOpened = p; ctx->gatherRoot = reinterpret_cast<float4*>(p);
    }
    ctx->peer.ps = ps; ctx->peer.epoch = 0; ctx->gatherEpoch = 0; ctx->peer.on = true;
    return PT_OK;
}
PT_API int pt_peer_disconnect(pt_ctx* ctx) {
    if (!ctx) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    CK(cudaStreamSynchronize(ctx->stream));
    peerClose(ctx);
    return PT_OK;
}

PT_API int pt_set_profiling(pt_ctx* ctx, int on) {
    if (!ctx) return PT_ERR_INVALID;
    { int jr = joinStreams(ctx); if (jr) return jr; }
    ctx->prof = on != 0;
    for (int i = 0; i < PT_NUM_STAGES; i++) { ctx->profMs[i] = 0; ctx->profLaunches[i] = 0; }
    return PT_OK;
}
PT_API int pt_get_profile(pt_ctx* ctx, double* ms, uint64_t* launches) {
    if (!ctx) return PT_ERR_INVALID;
    for (int i = 0; i < PT_NUM_STAGES; i++) { if (ms) ms[i] = ctx->profMs[i]; if (launches) launches[i] = ctx->profLaunches[i]; }
    return PT_OK;
}

PT_API int pt_read_image(pt_ctx* ctx, float* host_rgbw, int which) {
    if (!ctx || !host_rgbw) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    CK(cudaMemcpyAsync(host_rgbw, which == 0 ? ctx->image : ctx->albedo, (size_t)ctx->map.W * ctx->map.H * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return PT_OK;
}

// Asynchronous read-back: the image as it is after everything enqueued so far, copied on a private stream while later
// samples render (only their accumulate waits for the copy).  host should be pinned memory.  *ticket (optional) can be
// handed to pt_wait_read; pt_finish / pt_join also cover it.  tiles != 0: this rank's tiles only, slot order (pt_read_tiles).
static int readAsync(pt_ctx* ctx, float* host, int which, bool tiles, uint64_t* ticket) {
    if (!ctx || !host) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    if (!ctx->copyStream) {
        CK(cudaStreamCreateWithFlags(&ctx->copyStream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&ctx->copyReady, cudaEventDisableTiming));
        for (auto& e : ctx->copyEv) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    if (ctx->forked && ctx->lastSet >= 0 && ctx->set[ctx->lastSet].doneValid) CK(cudaStreamWaitEvent(ctx->copyStream, ctx->set[ctx->lastSet].done, 0));
    else { CK(cudaEventRecord(ctx->copyReady, ctx->stream)); CK(cudaStreamWaitEvent(ctx->copyStream, ctx->copyReady, 0)); }
    const float4* src = which == 0 ? ctx->image : ctx->albedo;
    if (tiles) {
        const uint32_t n = ctx->map.nSlots;
        if (n) {
            if (!ctx->packBuf) CK(cudaMalloc((void**)&ctx->packBuf, (size_t)n * sizeof(float4)));
            k_pack_tiles<<<n / 256, 256, 0, ctx->copyStream>>>(ctx->map, src, ctx->packBuf);
            CK(cudaGetLastError());
            CK(cudaMemcpyAsync(host, ctx->packBuf, (size_t)n * sizeof(float4), cudaMemcpyDeviceToHost, ctx->copyStream));
        }
    } else {
        CK(cudaMemcpyAsync(host, src, (size_t)ctx->map.W * ctx->map.H * sizeof(float4), cudaMemcpyDeviceToHost, ctx->copyStream));
    }
    CK(cudaEventRecord(ctx->imgFree, ctx->copyStream));
    CK(cudaEventRecord(ctx->copyEv[ctx->copyTickets & 7], ctx->copyStream));
    if (ticket) *ticket = ctx->copyTickets;
    ctx->copyTickets++;
    ctx->copyPending = true;
    return PT_OK;
}
// N ranks, ONE framebuffer: every rank calls this after the same Render calls.  Each rank's tiles of the accumulated image go
// over NVLink into rank 0's image ring (k_gather_push), rank 0 ships the assembled image to `host_full` (pinned, W*H float4;
// ignored on the other ranks) with one copy.  Asynchronous like pt_read_image_async: later samples keep rendering, only their
// accumulate step waits until this rank's tiles have been read.  Needs pt_peer_connect; with world_size 1 it is
// pt_read_image_async.
PT_API int pt_gather_image_async(pt_ctx* ctx, float* host_full, int which, uint64_t* ticket) {
    if (!ctx) return PT_ERR_INVALID;
    if (ctx->cfg.world_size == 1) return readAsync(ctx, host_full, which, false, ticket);
    if (!ctx->peer.on || !ctx->gatherRoot) return fail(ctx, PT_ERR_INVALID, "pt_gather_image_async: peers are not connected (pt_peer_connect)");
    if (ctx->cfg.rank == 0 && !host_full) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    if (!ctx->copyStream) {
        CK(cudaStreamCreateWithFlags(&ctx->copyStream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&ctx->copyReady, cudaEventDisableTiming));
        for (auto& e : ctx->copyEv) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    if (ctx->forked && ctx->lastSet >= 0 && ctx->set[ctx->lastSet].doneValid) CK(cudaStreamWaitEvent(ctx->copyStream, ctx->set[ctx->lastSet].done, 0));
    else { CK(cudaEventRecord(ctx->copyReady, ctx->stream)); CK(cudaStreamWaitEvent(ctx->copyStream, ctx->copyReady, 0)); }
    const uint32_t epoch = ++ctx->gatherEpoch;
    const size_t npix = (size_t)ctx->map.W * ctx->map.H;
    float4* slot = ctx->gatherRoot + (size_t)(epoch % PT_GATHER_RING) * npix;
    const PeerSet& ps = ctx->peer.ps;
    k_gather_push<<<std::max(1u, ctx->map.nSlots / 256), 256, 0, ctx->copyStream>>>(ps, ctx->map, which == 0 ? ctx->image : ctx->albedo, slot, epoch, PT_GATHER_RING,
                                                                                  PT_GATHER_ARRIVE, PT_GATHER_CONSUMED, PT_GATHER_COUNTER, ctx->peer.timeoutNs);
    CK(cudaEventRecord(ctx->imgFree, ctx->copyStream));
    if (ctx->cfg.rank == 0) {
        k_gather_wait<<<1, 32, 0, ctx->copyStream>>>(ps, epoch, PT_GATHER_ARRIVE, ctx->peer.timeoutNs);
        CK(cudaMemcpyAsync(host_full, slot, npix * sizeof(float4), cudaMemcpyDeviceToHost, ctx->copyStream));
        k_gather_release<<<1, 32, 0, ctx->copyStream>>>(ps, epoch, PT_GATHER_CONSUMED);
    }
    CK(cudaGetLastError());
    CK(cudaEventRecord(ctx->copyEv[ctx->copyTickets & 7], ctx->copyStream));
    if (ticket) *ticket = ctx->copyTickets;
    ctx->copyTickets++;
    ctx->copyPending = true;
    return PT_OK;
}
// ---- N ranks, one framebuffer in host memory, every rank through its own PCIe link ----------------------------------------------
// pt_gather_image_async assembles the image in rank 0's HBM and ships 16 B/pixel through ONE link (33 MB per 1080p image: ~0.7 ms,
// more than an 8-rank frame takes to render).  Here the framebuffer lives in a block of process-shared host memory (a file in
// /dev/shm mapped by every rank's process): header + `slots` images of W*H float4.  Every rank maps it into its GPU's address space
// (cudaHostRegister) and stores its own tiles into slot epoch % slots with a kernel, then publishes the epoch in the header; whoever
// consumes the frames (normally rank 0's process) waits until all ranks have published an epoch, reads the image in place and
// releases it.  Header: uint32 arrived[rank] at byte 4*rank (rank < 64), uint32 consumed at byte 256; images from byte 4096.
static volatile uint32_t* ringWord(pt_ctx* ctx, size_t byteOffset) { return reinterpret_cast<volatile uint32_t*>(ctx->ring.host + byteOffset); }
PT_API int pt_host_ring_attach(pt_ctx* ctx, void* base, uint64_t bytes, uint32_t slots) {
    if (!ctx || !base || slots == 0) return PT_ERR_INVALID;
    if (ctx->ring.host) return fail(ctx, PT_ERR_INVALID, "pt_host_ring_attach: already attached");
    if (ctx->cfg.world_size > 64) return fail(ctx, PT_ERR_INVALID, "pt_host_ring_attach: at most 64 ranks");
    const uint64_t need = PT_HOST_RING_HEADER_BYTES + (uint64_t)slots * ctx->map.W * ctx->map.H * sizeof(float4);
    if (bytes < need) return fail(ctx, PT_ERR_INVALID, "pt_host_ring_attach: block smaller than header + slots * width * height * 16");
    CK(cudaSetDevice(ctx->cfg.device));
    CK(cudaHostRegister(base, need, cudaHostRegisterPortable | cudaHostRegisterMapped));
    void* dev = nullptr;
    cudaError_t e = cudaHostGetDevicePointer(&dev, base, 0);
    if (e != cudaSuccess) { cudaHostUnregister(base); CK(e); }
    ctx->ring.host = (unsigned char*)base; ctx->ring.dev = (unsigned char*)dev; ctx->ring.bytes = need; ctx->ring.slots = slots;
    ctx->ring.epoch = *ringWord(ctx, 4 * ctx->cfg.rank);                 // a block that was used before: carry on from its epoch
    return PT_OK;
}
PT_API int pt_host_ring_detach(pt_ctx* ctx) {
    if (!ctx) return PT_ERR_INVALID;
    if (!ctx->ring.host) return PT_OK;
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    CK(cudaStreamSynchronize(ctx->stream));
    if (ctx->copyStream) CK(cudaStreamSynchronize(ctx->copyStream));
    CK(cudaHostUnregister(ctx->ring.host));
    ctx->ring.host = nullptr; ctx->ring.dev = nullptr; ctx->ring.slots = 0;
    return PT_OK;
}
// Every rank, after the same Render calls: this rank's tiles of the accumulated image go into the ring slot of the next epoch.
// Asynchronous like pt_read_image_async.  The host waits (bounded) until the consumer has released the image that occupied the slot.
PT_API int pt_deliver_image_async(pt_ctx* ctx, int which, uint64_t* epoch_out) {
    if (!ctx) return PT_ERR_INVALID;
    if (!ctx->ring.host) return fail(ctx, PT_ERR_INVALID, "pt_deliver_image_async: no host ring attached (pt_host_ring_attach)");
    CK(cudaSetDevice(ctx->cfg.device));
    const uint64_t epoch = ctx->ring.epoch + 1;
    if (epoch > ctx->ring.slots) {                                         // flow control: slot reuse waits for the consumer
        const double t0 = wallMs(), limit = (double)ctx->peer.timeoutNs * 1e-6;
        while ((uint64_t)*ringWord(ctx, 256) + ctx->ring.slots < epoch) {
            if (wallMs() - t0 > limit) return fail(ctx, PT_ERR_INVALID, "pt_deliver_image_async: the consumer did not release a ring slot in time");
            std::this_thread::yield();
        }
    }
    if (!ctx->copyStream) {
        CK(cudaStreamCreateWithFlags(&ctx->copyStream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&ctx->copyReady, cudaEventDisableTiming));
        for (auto& e : ctx->copyEv) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    if (ctx->forked && ctx->lastSet >= 0 && ctx->set[ctx->lastSet].doneValid) CK(cudaStreamWaitEvent(ctx->copyStream, ctx->set[ctx->lastSet].done, 0));
    else { CK(cudaEventRecord(ctx->copyReady, ctx->stream)); CK(cudaStreamWaitEvent(ctx->copyStream, ctx->copyReady, 0)); }
    const size_t npix = (size_t)ctx->map.W * ctx->map.H;
    const float4* src = which == 0 ? ctx->image : ctx->albedo;
    float4* hostSlot = reinterpret_cast<float4*>(ctx->ring.host + PT_HOST_RING_HEADER_BYTES) + (size_t)(epoch % ctx->ring.slots) * npix;
    volatile uint* devArrived = reinterpret_cast<volatile uint*>(ctx->ring.dev);
    if (ctx->cfg.world_size == 1) {
        CK(cudaMemcpyAsync(hostSlot, src, npix * sizeof(float4), cudaMemcpyDeviceToHost, ctx->copyStream));
        CK(cudaEventRecord(ctx->imgFree, ctx->copyStream));
        k_publish_host<<<1, 1, 0, ctx->copyStream>>>(devArrived, 0u, (uint)epoch);
    } else if (ctx->peer.on && ctx->gatherRoot) {
        // assemble in rank 0's HBM ring (as pt_gather_image_async), then every rank ships one band of the assembled frame
        const uint32_t ge = ++ctx->gatherEpoch;
        float4* rootSlot = ctx->gatherRoot + (size_t)(ge % PT_GATHER_RING) * npix;
        const PeerSet& ps = ctx->peer.ps;
        k_gather_push<<<std::max(1u, ctx->map.nSlots / 256), 256, 0, ctx->copyStream>>>(ps, ctx->map, src, rootSlot, ge, PT_GATHER_RING,
                                                                                      PT_GATHER_ARRIVE, PT_GATHER_CONSUMED, PT_GATHER_COUNTER, ctx->peer.timeoutNs);
        CK(cudaEventRecord(ctx->imgFree, ctx->copyStream));
        k_flags_wait<<<1, 32, 0, ctx->copyStream>>>(ps, 0u, PT_GATHER_ARRIVE, ge, ctx->peer.timeoutNs);
        const uint32_t H = ctx->map.H, world = ctx->cfg.world_size, rank = ctx->cfg.rank;
        const size_t row0 = (size_t)H * rank / world, row1 = (size_t)H * (rank + 1) / world;
        if (row1 > row0) CK(cudaMemcpyAsync(hostSlot + row0 * ctx->map.W, rootSlot + row0 * ctx->map.W, (row1 - row0) * ctx->map.W * sizeof(float4), cudaMemcpyDefault, ctx->copyStream));
        k_band_done<<<1, 1, 0, ctx->copyStream>>>(ps, PT_GATHER_BANDDONE, ge, devArrived, (uint)epoch);
        if (rank == 0) {
            k_flags_wait<<<1, 32, 0, ctx->copyStream>>>(ps, 0u, PT_GATHER_BANDDONE, ge, ctx->peer.timeoutNs);
            k_gather_release<<<1, 32, 0, ctx->copyStream>>>(ps, ge, PT_GATHER_CONSUMED);
        }
    } else {
        // no peer mapping: every rank stores its own tiles into the host image from a kernel (zero-copy stores, slower than a copy engine)
        float4* slot = reinterpret_cast<float4*>(ctx->ring.dev + PT_HOST_RING_HEADER_BYTES) + (size_t)(epoch % ctx->ring.slots) * npix;
        const uint32_t n = ctx->map.nSlots;
        if (n) {
            if (!ctx->packBuf) CK(cudaMalloc((void**)&ctx->packBuf, (size_t)n * sizeof(float4)));
            k_pack_tiles<<<n / 256, 256, 0, ctx->copyStream>>>(ctx->map, src, ctx->packBuf);
        }
        CK(cudaEventRecord(ctx->imgFree, ctx->copyStream));                  // later samples may accumulate again; the PCIe stores read the copy
        if (n) k_push_host<<<ctx->map.nLocalTiles, 256, 0, ctx->copyStream>>>(ctx->map, ctx->packBuf, slot);
        k_publish_host<<<1, 1, 0, ctx->copyStream>>>(devArrived, ctx->cfg.rank, (uint)epoch);
    }
    CK(cudaGetLastError());
    CK(cudaEventRecord(ctx->copyEv[ctx->copyTickets & 7], ctx->copyStream));
    ctx->copyTickets++;
    ctx->copyPending = true;
    ctx->ring.epoch = epoch;
    if (epoch_out) *epoch_out = epoch;
    return PT_OK;
}
// Consumer: blocks (bounded) until every rank has published `epoch`; *image = the W*H float4 framebuffer in the shared block.
PT_API int pt_host_ring_wait(pt_ctx* ctx, uint64_t epoch, const float** image) {
    if (!ctx || !ctx->ring.host || epoch == 0) return PT_ERR_INVALID;
    const double t0 = wallMs(), limit = (double)ctx->peer.timeoutNs * 1e-6;
    for (uint32_t r = 0; r < ctx->cfg.world_size; r++)
        while ((uint64_t)*ringWord(ctx, 4 * r) < epoch) {
            if (wallMs() - t0 > limit) return fail(ctx, PT_ERR_INVALID, "pt_host_ring_wait: a rank did not deliver its tiles in time");
            std::this_thread::yield();
        }
    std::atomic_thread_fence(std::memory_order_acquire);
    if (image) *image = reinterpret_cast<const float*>(ctx->ring.host + PT_HOST_RING_HEADER_BYTES) + (size_t)(epoch % ctx->ring.slots) * ctx->map.W * ctx->map.H * 4;
    return PT_OK;
}
PT_API int pt_host_ring_release(pt_ctx* ctx, uint64_t epoch) {
    if (!ctx || !ctx->ring.host) return PT_ERR_INVALID;
    std::atomic_thread_fence(std::memory_order_release);
    *ringWord(ctx, 256) = (uint32_t)epoch;
    return PT_OK;
}
PT_API int pt_read_image_async(pt_ctx* ctx, float* host_rgbw, int which, uint64_t* ticket) { return readAsync(ctx, host_rgbw, which, false, ticket); }
PT_API int pt_read_tiles_async(pt_ctx* ctx, float* host_rgbw_slots, int which, uint64_t* ticket) { return readAsync(ctx, host_rgbw_slots, which, true, ticket); }
// Blocks the host until the read-back with this ticket has landed (valid for the 8 most recent tickets; older ones have landed).
PT_API int pt_wait_read(pt_ctx* ctx, uint64_t ticket) {
    if (!ctx || ticket >= ctx->copyTickets) return PT_ERR_INVALID;
    if (ctx->copyTickets - ticket > 8) return PT_OK;
    CK(cudaEventSynchronize(ctx->copyEv[ticket & 7]));
    return PT_OK;
}

// This rank's share of the image only: nSlots float4 in slot order (pt_tile_slot_to_pixel maps a slot to its pixel).
// With N ranks every rank moves 1/N of the framebuffer instead of a full-size image that is zero outside its tiles.
PT_API int pt_read_tiles(pt_ctx* ctx, float* host_rgbw_slots, int which) {
    if (!ctx || !host_rgbw_slots) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    const uint32_t n = ctx->map.nSlots;
    if (n == 0) return PT_OK;
    if (!ctx->packBuf) CK(cudaMalloc((void**)&ctx->packBuf, (size_t)n * sizeof(float4)));
    k_pack_tiles<<<n / 256, 256, 0, ctx->stream>>>(ctx->map, which == 0 ? ctx->image : ctx->albedo, ctx->packBuf);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(host_rgbw_slots, ctx->packBuf, (size_t)n * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return PT_OK;
}

// Headless stand-in for the GLFW window (north_star): the display value rgb / w (main.cpp:58-62) as a colour PFM
// ("PF", little-endian floats, bottom row first).
PT_API int pt_write_pfm(pt_ctx* ctx, const char* path, int which) {
    if (!ctx || !path) return PT_ERR_INVALID;
    const uint32_t W = ctx->map.W, H = ctx->map.H;
    std::vector<float> img((size_t)W * H * 4);
    int r = pt_read_image(ctx, img.data(), which);
    if (r) return r;
    FILE* f = fopen(path, "wb");
    if (!f) return fail(ctx, PT_ERR_INVALID, "pt_write_pfm: cannot open the output file");
    fprintf(f, "PF\n%u %u\n-1.0\n", W, H);
    std::vector<float> row((size_t)W * 3);
    for (uint32_t y = 0; y < H; y++) {
        const float* src = img.data() + (size_t)(H - 1 - y) * W * 4;
        for (uint32_t x = 0; x < W; x++) {
            const float w = src[x * 4 + 3] > 0.0f ? src[x * 4 + 3] : 1.0f;
            row[x * 3] = src[x * 4] / w; row[x * 3 + 1] = src[x * 4 + 1] / w; row[x * 3 + 2] = src[x * 4 + 2] / w;
        }
        fwrite(row.data(), sizeof(float), row.size(), f);
    }
    fclose(f);
    return PT_OK;
}

// The display post-process of src/main.cpp:316-376 (gauss_horz -> gauss_vert -> quad_fs_blurred, or quad_fs alone),
// headless: result = the RGBA the fragment shader would write, one float4 per pixel.  lum / albedo: host float4
// images to process (e.g. the gathered multi-GPU framebuffer) or NULL for the context's own accumulators.
PT_API int pt_postprocess(pt_ctx* ctx, const float* lum_host, const float* albedo_host, float nr_samples, int blur, float* out_host_rgba) {
    if (!ctx || !out_host_rgba || (lum_host == nullptr) != (albedo_host == nullptr)) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    const int W = (int)ctx->map.W, H = (int)ctx->map.H;
    const size_t n = (size_t)W * H;
    cudaStream_t s = ctx->stream;
    if (!ctx->postBuf) CK(cudaMalloc((void**)&ctx->postBuf, 4 * n * sizeof(float4)));
    float4 *tmpA = ctx->postBuf, *tmpB = ctx->postBuf + n, *inL = ctx->postBuf + 2 * n, *inA = ctx->postBuf + 3 * n;
    const float4 *lum = ctx->image, *alb = ctx->albedo;
    if (lum_host) {
        CK(cudaMemcpyAsync(inL, lum_host, n * sizeof(float4), cudaMemcpyHostToDevice, s));
        CK(cudaMemcpyAsync(inA, albedo_host, n * sizeof(float4), cudaMemcpyHostToDevice, s));
        lum = inL; alb = inA;
    }
    const dim3 blk(32, 8), grd((W + 31) / 32, (H + 7) / 8);
    if (blur) {
        k_post_horz<<<grd, blk, 0, s>>>(lum, alb, tmpA, W, H, nr_samples);              // main.cpp:322-329
        k_post_vert<<<grd, blk, 0, s>>>(tmpA, tmpB, W, H, nr_samples);                  // main.cpp:331-336
        k_post_resolve<<<grd, blk, 0, s>>>(tmpB, alb, tmpA, W, H, 1);                   // main.cpp:370-376, quad_fs_blurred
    } else k_post_resolve<<<grd, blk, 0, s>>>(lum, alb, tmpA, W, H, 0);                 // quad_fs
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out_host_rgba, tmpA, n * sizeof(float4), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return PT_OK;
}

// Measured ceiling for the BVH gathers: GB/s of random `access_bytes` (16 or 32) reads over `array_bytes` of device
// memory (<= L2 size: the L2 gather rate; larger: the HBM random-access rate).
PT_API int pt_measure_gather_bandwidth(pt_ctx* ctx, uint64_t array_bytes, uint32_t access_bytes, uint32_t repeats, double* out_gbs) {
    if (!ctx || !out_gbs || (access_bytes != 16 && access_bytes != 32) || array_bytes < 4096) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    uint32_t nElems = 1; while ((uint64_t)nElems * 2 * access_bytes <= array_bytes) nElems *= 2;
    float4* data = nullptr; float* sink = nullptr;
    CK(cudaMalloc((void**)&data, (size_t)nElems * access_bytes));
    cudaError_t e = cudaMalloc((void**)&sink, sizeof(float));
    if (e != cudaSuccess) { cudaFree(data); CK(e); }
    cudaStream_t s = ctx->stream;
    cudaMemsetAsync(data, 0, (size_t)nElems * access_bytes, s);
    const uint32_t iters = 256, blocks = ctx->numSMs * 8, threads = 256;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (uint32_t r = 0; r < std::max(repeats, 2u) + 1; r++) {
        cudaEventRecord(e0, s);
        if (access_bytes == 16) k_gather_probe<1><<<blocks, threads, 0, s>>>(data, nElems, iters, sink);
        else k_gather_probe<2><<<blocks, threads, 0, s>>>(data, nElems, iters, sink);
        cudaEventRecord(e1, s);
        cudaStreamSynchronize(s);
        float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
        if (r > 0) best = std::min(best, ms);            // first pass warms the cache
    }
    e = cudaGetLastError();
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(data); cudaFree(sink);
    CK(e);
    *out_gbs = (double)blocks * threads * iters * access_bytes / (best * 1e-3) / 1e9;
    return PT_OK;
}

PT_API int pt_read_radiance_cache(pt_ctx* ctx, pt_radiance_cache* out) {
    if (!ctx || !out || !ctx->haveScene) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    const uint32_t T = ctx->sc.num_triangles;
    std::vector<float> rc((size_t)T * 12), add((size_t)T * 16);
    CK(cudaMemcpyAsync(rc.data(), ctx->sc.rcRadiance, rc.size() * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(add.data(), ctx->sc.rcAdditions, add.size() * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    for (uint32_t i = 0; i < T; i++) {
        for (int b = 0; b < 8; b++) {
            out[i].radianceCache[b] = rc[(size_t)i * 12 + b];
            out[i].additionCache[b] = add[(size_t)b * T + i];
            out[i].additionCacheCount[b] = add[(size_t)(8 + b) * T + i];
        }
        out[i].radianceTotal = rc[(size_t)i * 12 + 8];
    }
    return PT_OK;
}

PT_API int pt_get_stats(pt_ctx* ctx, pt_stats* out) {
    if (!ctx || !out) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    unsigned long long h[8];
    CK(cudaMemcpyAsync(h, ctx->stats, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    out->primary_rays = h[0]; out->extend_rays = h[1]; out->shadow_rays = h[2]; out->shade_calls = h[3]; out->samples = h[4];
    out->wave_extend_rays = h[5]; out->wave_shadow_rays = h[6]; out->kernel_launches = ctx->launches;
    return PT_OK;
}
PT_API int pt_reset_stats(pt_ctx* ctx) {
    if (!ctx) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    CK(cudaMemsetAsync(ctx->stats, 0, 8 * sizeof(unsigned long long), ctx->stream));
    ctx->launches = 0;
    return PT_OK;
}
PT_API int pt_get_rand_state(pt_ctx* ctx, uint32_t* sampleIdx, uint32_t* randIdx) {
    if (!ctx) return PT_ERR_INVALID;
    if (sampleIdx) *sampleIdx = ctx->sampleIdx;
    if (randIdx) *randIdx = ctx->randIdx;
    return PT_OK;
}

PT_API int pt_debug_primary_rays(pt_ctx* ctx, const pt_camera* cam, uint32_t randIdx, pt_ray* host_out) {
    if (!ctx || !cam || !host_out) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    const uint32_t W = ctx->map.W, H = ctx->map.H;
    pt_ray* d = nullptr;
    CK(cudaMalloc((void**)&d, (size_t)W * H * sizeof(pt_ray)));
    k_primary_debug<<<dim3((W + 15) / 16, (H + 15) / 16), dim3(16, 16), 0, ctx->stream>>>(toDCamera(cam), W, H, randIdx, d);
    cudaError_t e = cudaMemcpyAsync(host_out, d, (size_t)W * H * sizeof(pt_ray), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d);
    CK(e);
    return PT_OK;
}

static int debugTrace(pt_ctx* ctx, const pt_ray* host_rays, uint32_t n, int any_hit, uint32_t repeats, float* out_ms, pt_hit* host_out) {
    if (!ctx || !host_rays || !ctx->haveScene) return PT_ERR_INVALID;
    CK(cudaSetDevice(ctx->cfg.device));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    if (n > ctx->dbgCap) {
        if (ctx->dbgRays) cudaFree(ctx->dbgRays);
        if (ctx->dbgHits) cudaFree(ctx->dbgHits);
        ctx->dbgRays = nullptr; ctx->dbgHits = nullptr; ctx->dbgCap = 0;
        CK(cudaMalloc((void**)&ctx->dbgRays, (size_t)n * sizeof(pt_ray)));
        CK(cudaMalloc((void**)&ctx->dbgHits, (size_t)n * sizeof(pt_hit)));
        ctx->dbgCap = n;
    }
    cudaStream_t s = ctx->stream;
    ctx->sc.minWidth = 1u; ctx->sc.coop = ctx->coop; ctx->sc.coopTail = ctx->coopTail; ctx->sc.anyFarFirst = ctx->anyFarFirst;          // the stage hook runs alone: latency-mode scheduling, cooperative form where the queue is short
    CK(cudaMemcpyAsync(ctx->dbgRays, host_rays, (size_t)n * sizeof(pt_ray), cudaMemcpyHostToDevice, s));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float total = 0.0f;
    for (uint32_t r = 0; r < std::max(repeats, 1u); r++) {
        CK(cudaMemsetAsync(ctx->dbgCursor, 0, sizeof(uint), s));
        CK(cudaEventRecord(e0, s));
        if (any_hit) k_trace_debug<true><<<ctx->gridTraceAny, 128, 0, s>>>(ctx->sc, ctx->dbgRays, n, ctx->dbgCursor, ctx->dbgHits);
        else k_trace_debug<false><<<ctx->gridTraceExt, 128, 0, s>>>(ctx->sc, ctx->dbgRays, n, ctx->dbgCursor, ctx->dbgHits);
        CK(cudaEventRecord(e1, s));
        CK(cudaStreamSynchronize(s));
        float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (r > 0 || repeats <= 1) total += ms;     // first repeat is the warm-up when repeats > 1
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (out_ms) *out_ms = repeats > 1 ? total / (float)(repeats - 1) : total;
    if (host_out) {
        CK(cudaMemcpyAsync(host_out, ctx->dbgHits, (size_t)n * sizeof(pt_hit), cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
    }
    return PT_OK;
}
PT_API int pt_debug_trace(pt_ctx* ctx, const pt_ray* host_rays, uint32_t n, int any_hit, pt_hit* host_out) {
    return debugTrace(ctx, host_rays, n, any_hit, 1, nullptr, host_out);
}
PT_API int pt_debug_trace_resident(pt_ctx* ctx, const pt_ray* host_rays, uint32_t n, int any_hit, uint32_t repeats, float* out_ms, pt_hit* host_out) {
    return debugTrace(ctx, host_rays, n, any_hit, repeats, out_ms, host_out);
}
}  // extern "C"

// ---- pt_trace_static: the traversal hook for callers that may not allocate ------------------------------------------
// No counterpart in the reference.  Same work and same kernel as
// pt_debug_trace (k_trace_debug: persistent-warp traversal fed by a work cursor), but the scratch it needs -- the work
// cursor, the staged rays and the hit records -- lives in ONE file-scope device array instead of cudaMalloc'd buffers.
// Every GPU that loads the library holds the array (PT_STATIC_SCRATCH_BYTES, about 14 MiB) until the process exits.
// All calls share it and it keeps the last call's contents: the cursor is cleared by every call; calls that overlap on
// the GPU (two host threads, two contexts on one device) race on it -- serialise them.
struct StaticTraceScratch {
    uint cursor; uint pad[63];                                   // 256 B: the cursor has a line pair to itself
    pt_ray rays[PT_STATIC_TRACE_MAX_RAYS];
    pt_hit hits[PT_STATIC_TRACE_MAX_RAYS];
};
static_assert(sizeof(StaticTraceScratch) == PT_STATIC_SCRATCH_BYTES, "PT_STATIC_SCRATCH_BYTES in ptb200.h");
static __device__ __align__(256) unsigned char g_staticScratch[sizeof(StaticTraceScratch)];

// In host code g_staticScratch is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress,
// once per device, with that device current.
static constexpr int kMaxStaticDevices = 64;
static StaticTraceScratch* g_staticScratchPtr[kMaxStaticDevices];
static std::mutex g_staticScratchMutex;

static cudaError_t staticScratchFor(int device, StaticTraceScratch** out) {
    if (device < 0 || device >= kMaxStaticDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_staticScratchMutex);
    if (!g_staticScratchPtr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_staticScratch);       // the caller made `device` current
        if (e != cudaSuccess) return e;
        g_staticScratchPtr[device] = (StaticTraceScratch*)p;
    }
    *out = g_staticScratchPtr[device];
    return cudaSuccess;
}

extern "C" PT_API int pt_trace_static(pt_ctx* ctx, const pt_ray* host_rays, uint32_t n, int any_hit, pt_hit* host_out) {
    if (!ctx) return PT_ERR_INVALID;
    if (!host_rays || !host_out || !ctx->haveScene) { ctx->err = "pt_trace_static: needs a scene, host_rays and host_out"; return PT_ERR_INVALID; }
    if (n > PT_STATIC_TRACE_MAX_RAYS) {
        char buf[160];
        snprintf(buf, sizeof(buf), "pt_trace_static: %u rays exceed the static scratch (PT_STATIC_TRACE_MAX_RAYS = %u); use pt_debug_trace", n, (unsigned)PT_STATIC_TRACE_MAX_RAYS);
        ctx->err = buf;
        return PT_ERR_INVALID;
    }
    if (n == 0) return PT_OK;
    CK(cudaSetDevice(ctx->cfg.device));
    StaticTraceScratch* scr = nullptr;
    CK(staticScratchFor(ctx->cfg.device, &scr));
    { int jr = joinStreams(ctx); if (jr) return jr; }
    cudaStream_t s = ctx->stream;
    DScene sc = ctx->sc;                                         // by value: the context's own scheduling fields stay as they are
    sc.minWidth = 1u; sc.coop = ctx->coop; sc.coopTail = ctx->coopTail; sc.anyFarFirst = ctx->anyFarFirst;
    CK(cudaMemsetAsync(&scr->cursor, 0, sizeof(uint), s));       // the one word that must start at zero
    CK(cudaMemcpyAsync(scr->rays, host_rays, (size_t)n * sizeof(pt_ray), cudaMemcpyHostToDevice, s));
    if (any_hit) k_trace_debug<true><<<ctx->gridTraceAny, 128, 0, s>>>(sc, scr->rays, n, &scr->cursor, scr->hits);
    else k_trace_debug<false><<<ctx->gridTraceExt, 128, 0, s>>>(sc, scr->rays, n, &scr->cursor, scr->hits);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(host_out, scr->hits, (size_t)n * sizeof(pt_hit), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return PT_OK;
}
