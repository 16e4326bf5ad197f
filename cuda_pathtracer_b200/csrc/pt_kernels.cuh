// pt_kernels.cuh -- the wavefront stages as CUDA kernels for sm_100a.
//   k_primary      clear_state + generate_primary_rays (reference kernels.h:485-501), compacted append
//   k_trace<ANY>   extend / connect (kernels.h:503-510, 799-810): persistent warps, dynamic ray fetch
//   k_shade        kernel_shade (kernels.h:513-796) + warp-aggregated compaction of survivors
//   k_accumulate   add_to_screen (kernels.h:812-824) + per-sample ray statistics
//   k_rc_*         init / update / propagate radiance cache (kernels.h:848-905)
// Queues: SoA float4 streams (origin|tmax, direction|slot), ping-pong by bounce parity; per-bounce
// counters and work cursors live in one small device array that is zeroed once per sample, which
// replaces the reference's kernel_clear_rays / kernel_swap_and_clear <<<1,1>>> launches.
#pragma once
#include "pt_shade.cuh"

namespace ptb {

// counter array layout (uint32 each)
#define CNT_EXT(b) (b)                              /* rays queued for extend at bounce b          */
#define CNT_SHADOW(b) (PT_MAX_BOUNCES + 1 + (b))    /* shadow rays queued at bounce b              */
#define CUR_EXT(b) (2 * (PT_MAX_BOUNCES + 1) + (b)) /* dynamic-fetch cursors of the persistent kernels */
#define CUR_CONN(b) (3 * (PT_MAX_BOUNCES + 1) + (b))
#define CNT_TOTAL (4 * (PT_MAX_BOUNCES + 1))

struct Queues {
    float4* rayO[2]; float4* rayD[2];   // extension rays, ping-pong
    float4* shO; float4* shD;           // shadow rays
    float4* hits;                       // HitInfoPacked per extension ray (types.h:310-334)
    uint* cnt;
};

// One atomic per warp: ballot the survivors, the first one reserves popc slots, everyone writes at
// base + rank.  Must be called by all 32 lanes.
__device__ __forceinline__ uint warpAppend(bool pred, uint* counter) {
    const uint m = __ballot_sync(0xffffffffu, pred);
    if (m == 0) return 0;
    const uint leader = __ffs(m) - 1;
    uint base = 0;
    if (laneId() == leader) base = atomicAdd(counter, __popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + __popc(m & ((1u << laneId()) - 1u));
}

// One atomic per BLOCK (<= 8 warps): warps take consecutive ranges in warp order.  Must be called by every thread of the
// block; `sm` = 9 uints of shared memory, reusable after the call returns.
__device__ __forceinline__ uint blockAppend(bool pred, uint* counter, uint* sm) {
    const uint m = __ballot_sync(0xffffffffu, pred);
    const uint w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    if (laneId() == 0) sm[w] = __popc(m);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint total = 0;
        for (uint i = 0; i < nw; i++) { const uint c = sm[i]; sm[i] = total; total += c; }
        sm[8] = total ? atomicAdd(counter, total) : 0u;
    }
    __syncthreads();
    const uint idx = sm[8] + sm[w] + __popc(m & ((1u << laneId()) - 1u));
    __syncthreads();
    return idx;
}

// ---- primary rays: types.h:641-658 (Camera::getRay), 669-676 (distort) ----------------------------------
__device__ __forceinline__ float3 camDistort(const DCamera& c, float3 p) {
    float3 center = c.eye + c.d * c.viewDir;
    float3 fromCenter = p - center;
    float r = length(p - center);
    float rd = r + 0.2f * r * r * r;
    return center + fromCenter * (rd / fmaxf(0.0001f, r));
}
__device__ __forceinline__ void camGetRay(const DCamera& c, uint x, uint y, uint& seed, uint W, uint H, float3& o, float3& d) {
    float xf = ((float)x + rnd(seed)) / (float)W;
    float yf = ((float)y + rnd(seed)) / (float)H;
    float3 origin = camDistort(c, c.lt + xf * c.u + yf * c.v);
    float3 direction = origin - c.eye;
    const float correction = length(direction);
    direction = direction / correction;
    const float3 focalPoint = origin + (c.focalLength - c.d) * direction;
    const float offsetR = sqrtf(rnd(seed));
    const float offsetA = rnd(seed) * 2.0f * PT_PI;
    const float fx = offsetR * sinf(offsetA), fy = offsetR * cosf(offsetA);
    origin += c.aperture * (fx * c.u + fy * c.v);
    direction = normalize(focalPoint - origin);
    if (dot(direction, c.viewDir) < 0) direction = -direction;
    o = origin - correction * direction; d = direction;
}

__global__ void __launch_bounds__(256) k_primary(DCamera cam, TileMap map, uint randIdx, Queues q, float4* masks, float4* accucolors) {
    const uint slot = map.slotBase + blockIdx.x * blockDim.x + threadIdx.x;      // grid covers whole tiles (multiples of 1024 slots)
    uint x, y;
    const bool valid = slotToPixel(map, slot, x, y);
    float3 o = f3(0), d = f3(0);
    pdlWait(); pdlLaunch();
    if (valid) {
        // (the per-sample state reset of kernels.h:485-491 happens in the first shade: every valid slot gets a primary ray)
        uint seed = getSeed(x, y, randIdx, map.W);
        camGetRay(cam, x, y, seed, map.W, map.H, o, d);
    }
    // one atomic per 256-thread block: 65k same-address atomics per 1080p frame were this kernel's critical path
    __shared__ uint sm[9];
    const uint idx = blockAppend(valid, q.cnt + CNT_EXT(0), sm);
    if (valid) {
        q.rayO[0][idx] = make_float4(o.x, o.y, o.z, 9999999.0f);
        q.rayD[0][idx] = make_float4(d.x, d.y, d.z, __uint_as_float(slot));
    }
}

// debug variant: rays written at their PIXEL index in reference RayPacked form (pt_debug_primary_rays)
__global__ void k_primary_debug(DCamera cam, uint W, uint H, uint randIdx, pt_ray* out) {
    const uint x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= W || y >= H) return;
    uint seed = getSeed(x, y, randIdx, W);
    float3 o, d; camGetRay(cam, x, y, seed, W, H, o, d);
    pt_ray r; r.origin.x = o.x; r.origin.y = o.y; r.origin.z = o.z; r.origin.w = 9999999.0f;
    r.direction.x = d.x; r.direction.y = d.y; r.direction.z = d.z; r.direction.w = __uint_as_float(x + y * W);
    out[x + y * W] = r;
}

// ---- extend / connect ---------------------------------------------------------------------------------------
// Persistent warps running the flat two-level traversal with dynamic ray fetch (pt_trace.cuh traceLoop).
// count/cursor are device-side: the host never reads queue sizes (reference README.md:201).
#ifndef PT_TRACE_MINBLOCKS
#define PT_TRACE_MINBLOCKS 8
#endif
// Short queue in latency mode: 8 lanes per ray (traceLoopCoop) instead of one.  Block-uniform.
__device__ __forceinline__ bool useCoop(const DScene& sc, uint n) {
    return sc.coop != 0u && n <= ((gridDim.x * blockDim.x) >> 5) * 4u * sc.coop;
}
template <bool ANY>
__global__ void __launch_bounds__(128, PT_TRACE_MINBLOCKS) k_trace(DScene sc, const float4* __restrict__ rayO, const float4* __restrict__ rayD,
                                               const uint* __restrict__ count, uint* cursor, float4* __restrict__ hits,
                                               const float4* __restrict__ lights, float4* accucolors, const uint* __restrict__ perm) {
    pdlWait(); pdlLaunch();
    QueueIO io; io.rayO = rayO; io.rayD = rayD; io.hits = hits; io.lights = lights; io.accucolors = accucolors; io.perm = perm; io.limits(sc);
    NoTally tally;
    const uint n = *count;
    if (useCoop(sc, n)) traceLoopCoop<ANY>(sc, n, cursor, io, tally);
    else traceLoop<ANY>(sc, n, cursor, io, tally);
}
// counting variant (pt_set_counting): the same traversal, plus the node steps / triangle tests / top-level steps it
// performs, summed into tallyOut[0..2] (+ the longest single ray) -- the bytes the wide layout actually requests (80 B per node step, 48 B per
// triangle test, 32 B per top-level node) for the roofline line of bench.py
template <bool ANY>
__global__ void __launch_bounds__(128, PT_TRACE_MINBLOCKS) k_trace_count(DScene sc, const float4* __restrict__ rayO, const float4* __restrict__ rayD,
                                               const uint* __restrict__ count, uint* cursor, float4* __restrict__ hits,
                                               const float4* __restrict__ lights, float4* accucolors, unsigned long long* tallyOut) {
    pdlWait(); pdlLaunch();
    QueueIO io; io.rayO = rayO; io.rayD = rayD; io.hits = hits; io.lights = lights; io.accucolors = accucolors; io.perm = nullptr; io.limits(sc);
    Tally tally;
    traceLoop<ANY>(sc, *count, cursor, io, tally);
    tally.flush(tallyOut, tallyOut + (ANY ? 4 : 6));        // tally buffer: ext[0..2], conn[3..5], longest ext ray [6], longest shadow ray [7]
}

// extend(b+1) and connect(b) in ONE launch: both queues exist once shade(b) has run and neither kernel touches what the
// other writes (hits vs accucolors), so the dependent-launch chain per bounce is shade -> trace instead of
// shade -> connect -> extend.  Warps drain the extension queue first, then the shadow queue fills the gaps its
// stragglers leave.  (The per-stage profiling pass launches them separately to time them apart.)
__global__ void __launch_bounds__(128, PT_TRACE_MINBLOCKS) k_trace_pair(DScene sc, const float4* __restrict__ rayO, const float4* __restrict__ rayD,
                                                    const uint* __restrict__ extCount, uint* extCursor, float4* __restrict__ hits,
                                                    const float4* __restrict__ shO, const float4* __restrict__ shD,
                                                    const uint* __restrict__ shCount, uint* shCursor,
                                                    const float4* __restrict__ lights, float4* accucolors,
                                                    const uint* __restrict__ extPerm, const uint* __restrict__ shPerm) {
    pdlWait(); pdlLaunch();
    NoTally tally;
    QueueIO e; e.rayO = rayO; e.rayD = rayD; e.hits = hits; e.lights = nullptr; e.accucolors = nullptr; e.perm = extPerm; e.limits(sc);
    const uint ne = *extCount, ns = *shCount;
    QueueIO c; c.rayO = shO; c.rayD = shD; c.hits = nullptr; c.lights = lights; c.accucolors = accucolors; c.perm = shPerm; c.limits(sc);
    if (useCoop(sc, ne + ns)) {
        traceLoopCoop<false>(sc, ne, extCursor, e, tally);
        traceLoopCoop<true>(sc, ns, shCursor, c, tally);
        return;
    }
    traceLoop<false>(sc, ne, extCursor, e, tally);
    traceLoop<true>(sc, ns, shCursor, c, tally);
}

// The same launch for a sample that runs alone (latency mode): both queues as ONE work list (traceLoopMixed), cooperative when
// short.  A kernel of its own so that k_trace_pair's register allocation stays what it is for the throughput launches.
__global__ void __launch_bounds__(128, PT_TRACE_MINBLOCKS) k_trace_pair_mixed(DScene sc, const float4* __restrict__ rayO, const float4* __restrict__ rayD,
                                                    const uint* __restrict__ extCount, uint* extCursor, float4* __restrict__ hits,
                                                    const float4* __restrict__ shO, const float4* __restrict__ shD,
                                                    const uint* __restrict__ shCount,
                                                    const float4* __restrict__ lights, float4* accucolors,
                                                    const uint* __restrict__ extPerm, const uint* __restrict__ shPerm) {
    pdlWait(); pdlLaunch();
    NoTally tally;
    QueueIO e; e.rayO = rayO; e.rayD = rayD; e.hits = hits; e.lights = nullptr; e.accucolors = nullptr; e.perm = extPerm; e.limits(sc);
    QueueIO c; c.rayO = shO; c.rayD = shD; c.hits = nullptr; c.lights = lights; c.accucolors = accucolors; c.perm = shPerm; c.limits(sc);
    const uint ne = *extCount, ns = *shCount;
    if (useCoop(sc, ne + ns)) traceLoopCoopMixed(sc, ne, ns, extCursor, e, c, tally);
    else traceLoopMixed(sc, ne, ns, extCursor, e, c, tally);
}

// stage-parity hook: reference-format rays in, reference-format hits out (pt_debug_trace)
template <bool ANY>
__global__ void __launch_bounds__(128) k_trace_debug(DScene sc, const pt_ray* __restrict__ rays, uint n, uint* cursor, pt_hit* __restrict__ out) {
    DebugIO io; io.rays = rays; io.out = out;
    NoTally tally;
    if (useCoop(sc, n)) traceLoopCoop<ANY>(sc, n, cursor, io, tally);
    else traceLoop<ANY>(sc, n, cursor, io, tally);
}

__global__ void k_prepare_lights(DScene sc) {
    const uint i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < sc.num_lights) prepareLight(sc, i);
}

// ---- shade ----------------------------------------------------------------------------------------------------
#ifndef PT_SHADE_MINBLOCKS
#define PT_SHADE_MINBLOCKS 8
#endif
// Ray sorting (the reference's experiment sorts by pixel with a CUB radix sort and permutes the rays,
// src/raySorting.h:1-68).  Here: a counting sort on (Morton cell of the origin, direction octant) whose histogram is
// built by the shade kernel while it appends the rays; k_sort_scan turns it into bin offsets, k_sort_scatter into a
// permutation, and the traversal kernels fetch rays through that permutation (hits land at the ray's own queue index, so
// shading order -- and every result -- is untouched).  Queue sizes stay on the device.
__device__ __forceinline__ uint sortBin(const DScene& sc, float3 o, float3 d) {
    const float cells = (float)(1u << sc.sortBits), top = cells - 1.0f;          // sortScale = 1 / extent of the scene box
    const uint cx = (uint)fminf(fmaxf((o.x - sc.sortLo.x) * sc.sortScale.x * cells, 0.0f), top);
    const uint cy = (uint)fminf(fmaxf((o.y - sc.sortLo.y) * sc.sortScale.y * cells, 0.0f), top);
    const uint cz = (uint)fminf(fmaxf((o.z - sc.sortLo.z) * sc.sortScale.z * cells, 0.0f), top);
    uint m = 0;
    for (uint b = 0; b < sc.sortBits; b++) m |= (((cx >> b) & 1u) << (3u * b)) | (((cy >> b) & 1u) << (3u * b + 1u)) | (((cz >> b) & 1u) << (3u * b + 2u));
    return (m << 3) | (d.x < 0.0f ? 1u : 0u) | (d.y < 0.0f ? 2u : 0u) | (d.z < 0.0f ? 4u : 0u);
}
__global__ void __launch_bounds__(1024) k_sort_scan(uint* histExt, uint* histSh, uint nbins) {      // block 0: extension queue, block 1: shadow queue
    pdlWait(); pdlLaunch();
    uint* h = blockIdx.x ? histSh : histExt;
    if (!h) return;
    __shared__ uint warpSums[32];
    const uint per = (nbins + 1023u) / 1024u, b0 = threadIdx.x * per, lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint sum = 0;
    for (uint k = 0; k < per && b0 + k < nbins; k++) sum += h[b0 + k];
    uint incl = sum;
    for (uint off = 1; off < 32; off <<= 1) { const uint v = __shfl_up_sync(0xffffffffu, incl, off); if (lane >= off) incl += v; }
    if (lane == 31) warpSums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        const uint w = warpSums[lane]; uint wi = w;
        for (uint off = 1; off < 32; off <<= 1) { const uint v = __shfl_up_sync(0xffffffffu, wi, off); if (lane >= off) wi += v; }
        warpSums[lane] = wi - w;
    }
    __syncthreads();
    uint run = warpSums[warp] + incl - sum;
    for (uint k = 0; k < per && b0 + k < nbins; k++) { const uint c = h[b0 + k]; h[b0 + k] = run; run += c; }
}
__global__ void __launch_bounds__(256) k_sort_scatter(const uint* __restrict__ keysExt, uint* histExt, uint* permExt, const uint* __restrict__ nExt,
                                                      const uint* __restrict__ keysSh, uint* histSh, uint* permSh, const uint* __restrict__ nSh) {
    pdlWait(); pdlLaunch();
    const uint stride = gridDim.x * blockDim.x, t0 = blockIdx.x * blockDim.x + threadIdx.x;
    if (keysExt) { const uint n = *nExt; for (uint i = t0; i < n; i += stride) permExt[atomicAdd(histExt + keysExt[i], 1u)] = i; }
    if (keysSh) { const uint n = *nSh; for (uint i = t0; i < n; i += stride) permSh[atomicAdd(histSh + keysSh[i], 1u)] = i; }
}

__global__ void __launch_bounds__(128, PT_SHADE_MINBLOCKS) k_shade(ShadeParams P, Queues q, int inBuf) {
    pdlWait(); pdlLaunch();
    const uint n = q.cnt[CNT_EXT(P.bounce)];
    const float4* rayO = q.rayO[inBuf]; const float4* rayD = q.rayD[inBuf];
    float4* outO = q.rayO[inBuf ^ 1]; float4* outD = q.rayD[inBuf ^ 1];
    __shared__ uint sm[9];
    const uint nRound = (n + blockDim.x - 1u) / blockDim.x * blockDim.x;          // block-uniform trip count (blockAppend synchronises)
    for (uint base = blockIdx.x * blockDim.x; base < nRound; base += gridDim.x * blockDim.x) {
        const uint i = base + threadIdx.x;
        ShadeOut so; so.pushExt = so.pushShadow = false;
        uint slot = 0;
        if (i < n) {
            const float4 o4 = rayO[PT_IDX(i, PT_LIMS(P.sc, queueCap), PT_CHK_QUEUE)], d4 = rayD[PT_IDX(i, PT_LIMS(P.sc, queueCap), PT_CHK_QUEUE)], h4 = q.hits[PT_IDX(i, PT_LIMS(P.sc, queueCap), PT_CHK_QUEUE)];
            slot = __float_as_uint(d4.w);
            Hit h; h.type = __float_as_uint(h4.x); h.id = __float_as_uint(h4.y); h.inst = __float_as_uint(h4.z); h.t = h4.w;
            shadeOne(P, f3(o4), f3(d4), slot, h, so);
        }
        uint si = blockAppend(so.pushShadow, q.cnt + CNT_SHADOW(P.bounce), sm);
        if (so.pushShadow) {
            si = PT_IDX(si, PT_LIMS(P.sc, queueCap), PT_CHK_QUEUE);
            q.shO[si] = make_float4(so.shadow.o.x, so.shadow.o.y, so.shadow.o.z, so.shadow.len);
            q.shD[si] = make_float4(so.shadow.d.x, so.shadow.d.y, so.shadow.d.z, __uint_as_float(slot));
            if (P.sortKeys[1]) { const uint bin = sortBin(P.sc, so.shadow.o, so.shadow.d); P.sortKeys[1][si] = bin; atomicAdd(P.sortHist[1] + bin, 1u); }
        }
        uint ei = blockAppend(so.pushExt, q.cnt + CNT_EXT(P.bounce + 1), sm);
        if (so.pushExt) {
            ei = PT_IDX(ei, PT_LIMS(P.sc, queueCap), PT_CHK_QUEUE);
            outO[ei] = make_float4(so.ext.o.x, so.ext.o.y, so.ext.o.z, so.ext.len);
            outD[ei] = make_float4(so.ext.d.x, so.ext.d.y, so.ext.d.z, __uint_as_float(slot));
            if (P.sortKeys[0]) { const uint bin = sortBin(P.sc, so.ext.o, so.ext.d); P.sortKeys[0][ei] = bin; atomicAdd(P.sortHist[0] + bin, 1u); }
        }
    }
}

// ---- tail bounces: one launch, one lane per PATH ----------------------------------------------------------------
// From bounce b0 on the queues hold a few thousand rays: every wavefront launch then costs the latency of its
// slowest ray (profiles/r1g: 30-130 us per launch for <1% of the rays, 3 launches per bounce).  k_tail takes
// every path that is alive at bounce b0 to its end inside one launch: extend -> shade -> connect -> extend ...
// per lane, so the launch costs the slowest PATH, not the sum over bounces of the slowest ray.  Per path the
// order of operations (and of the float additions into accucolor) is the wavefront's, hence the reference's
// (pathtracer.h:276-290); randIdx advances per bounce exactly as the host loop does.
// Scratch: the path with queue index i keeps its current extension ray in ray buffer `cur`[i], writes the next
// one to the other buffer [i] and its shadow ray to shO/shD[i] (all dead wavefront storage by now).
#ifndef PT_TAIL_MINBLOCKS
#define PT_TAIL_MINBLOCKS 3
#endif
__device__ __forceinline__ float4 shfl4(const float4 v, const uint src) {
    return make_float4(__shfl_sync(0xffffffffu, v.x, src), __shfl_sync(0xffffffffu, v.y, src), __shfl_sync(0xffffffffu, v.z, src), __shfl_sync(0xffffffffu, v.w, src));
}
// Pair mode (sc.tailPair, chosen per launch when the paths fit the machine 16 to a warp): a path owns TWO lanes.  The even
// lane carries the extension rays and shades; the shadow ray of bounce b is handed (shuffles) to the odd lane, which traces
// it WHILE the even lane traces the extension ray of bounce b+1 -- the two are independent once shade(b) has run, exactly
// the pairing k_trace_pair gives the wavefront.  The even lane shades bounce b+1 only after its helper has finished
// connect(b), so the float additions into accucolor keep the wavefront's (the reference's) order: identical bits.  A lone
// sample on a small shard is latency bound and has the lanes to spare: the per-bounce latency becomes
// max(extend, connect) + shade instead of their sum.
// k_tail with the cooperative traversal (pt_trace.cuh traceLoopCoop): a PATH owns a group of 8 lanes -- in pair mode 16: the
// even group extends and shades, the odd group traces the shadow rays -- so a warp carries 4 (2) paths.  Used when the
// surviving paths fit the machine that way (multi-GPU shards, deep bounces); every lane of a group runs the same statements on
// the same values (shading included: it costs no extra issue slots), lane 0 of the group does the atomics and stores.
__device__ __forceinline__ void tailCoop(const ShadeParams& P, const Queues& q, const int inBuf, const int b0, const int maxBounces, uint* cursor, const uint n) {
    const DScene& sc = P.sc;
    const uint lane = laneId(), group = lane >> 3, sub = lane & 7u, gmask = 0xFFu << (group * 8u);
    const uint totalWarps = (gridDim.x * blockDim.x) >> 5;
    const bool pair = sc.tailPair != 0u && P.nee && n <= totalWarps * 2u * max(1u, sc.coopTail);
    const bool isHelper = pair && (group & 1u) != 0u;
    const uint ownerBits = pair ? 0x00010001u : 0x01010101u;                     // lane-0 bits of the groups that own paths
    const uint width = min(pair ? 2u : 4u, max(1u, (n + totalWarps - 1u) / totalWarps));
    const bool writer = sub == 0u;
    TravStack stack; stack.init();
    NoTally tally;
    LaneState L;
    L.ray = PT_NONE; L.ngroup = make_uint2(0u, 0u); L.tgroup = make_uint2(0u, 0u); L.curT = PT_NONE; L.sp = 0; L.spBase = 0; L.inBLAS = false;
    int bounce = b0, buf = inBuf;
    bool any = false, extPending = false, rayDone = false;
    bool exhausted = n == 0;
    const uint randIdx0 = P.randIdx;
    auto loadCur = [&](float4& o4, float4& d4) {
        if (any) { o4 = q.shO[L.ray]; d4 = q.shD[L.ray]; } else { o4 = q.rayO[buf][L.ray]; d4 = q.rayD[buf][L.ray]; }
    };
    auto addLight = [&]() {                                                     // unoccluded: kernels.h:806-809; one lane adds, the group sees it
        if (writer) {
            const uint slot = __float_as_uint(q.shD[L.ray].w);
            const float4 l = P.lights[slot];
            float4 a = P.accucolors[slot];
            a.x += l.x; a.y += l.y; a.z += l.z;
            P.accucolors[slot] = a;
        }
        __syncwarp(gmask);
    };
    while (true) {
        // ---- F: idle owner groups claim new paths ----
        const uint activeAll = __ballot_sync(0xffffffffu, L.ray != PT_NONE) & 0x01010101u;
        const uint activeMask = activeAll & ownerBits;
        const uint nActive = __popc(activeMask);
        if (!exhausted && nActive < width) {
            const uint want = width - nActive;
            uint base = 0;
            if (lane == 0) base = atomicAdd(cursor, want);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (base + want >= n) exhausted = true;
            const uint r = __popc(~activeAll & ownerBits & ((1u << (group * 8u)) - 1u));
            if (!isHelper && L.ray == PT_NONE && r < want && base + r < n) {
                L.ray = base + r; bounce = b0; buf = inBuf; any = false; extPending = false;
                float4 o4, d4; loadCur(o4, d4);
                rayDone = !beginRay(sc, L, o4, d4, false);
            }
        } else if (activeAll == 0u) break;
        // ---- C: a finished ray continues its path: shade, then the shadow ray, then the next extension ray ----
        bool helperBusy = false;
        if (pair) {
            if (isHelper && L.ray != PT_NONE && rayDone) {                      // the helper group's shadow ray is done
                if (L.hitId == PT_NONE) addLight();
                L.ray = PT_NONE; rayDone = false;
            }
            __syncwarp();                                                       // its accucolor update before the owner's next shade
            helperBusy = __shfl_xor_sync(0xffffffffu, (int)(L.ray != PT_NONE), 8) != 0;
        }
        bool handoff = false;
        uint hpath = PT_NONE;
        float4 hso = make_float4(0.f, 0.f, 0.f, 0.f), hsd = hso;
        while (!isHelper && L.ray != PT_NONE && rayDone && !helperBusy && !handoff) {
            rayDone = false;
            bool startExt = false;
            if (any) {
                if (L.hitId == PT_NONE) addLight();
                any = false;
                startExt = extPending;
            } else {
                float4 o4, d4; loadCur(o4, d4);
                const uint slot = __float_as_uint(d4.w);
                Hit h; h.type = L.hitType; h.id = L.hitId; h.inst = L.hitInst; h.t = L.tbest;
                ShadeParams Pb = P; Pb.bounce = bounce; Pb.randIdx = randIdx0 + (uint)(bounce - b0);
                ShadeOut so;
                shadeOne(Pb, f3(o4), f3(d4), slot, h, so, gmask, writer);
                __syncwarp(gmask);                                              // the writer's state stores before anybody of the group reads them again
                extPending = so.pushExt && bounce + 1 < maxBounces;
                if (so.pushExt && writer) atomicAdd(q.cnt + CNT_EXT(bounce + 1), 1u);
                if (extPending) {                                               // every lane stores the same values and later reads back its own store
                    q.rayO[buf ^ 1][L.ray] = make_float4(so.ext.o.x, so.ext.o.y, so.ext.o.z, so.ext.len);
                    q.rayD[buf ^ 1][L.ray] = make_float4(so.ext.d.x, so.ext.d.y, so.ext.d.z, __uint_as_float(slot));
                }
                if (so.pushShadow) {
                    if (writer) atomicAdd(q.cnt + CNT_SHADOW(bounce), 1u);
                    const float4 so4 = make_float4(so.shadow.o.x, so.shadow.o.y, so.shadow.o.z, so.shadow.len);
                    const float4 sd4 = make_float4(so.shadow.d.x, so.shadow.d.y, so.shadow.d.z, __uint_as_float(slot));
                    q.shO[L.ray] = so4; q.shD[L.ray] = sd4;
                    if (pair) { handoff = true; hpath = L.ray; hso = so4; hsd = sd4; startExt = extPending; }
                    else { any = true; rayDone = !beginRay(sc, L, so4, sd4, true); }
                } else startExt = extPending;
            }
            if (startExt) {
                extPending = false; bounce++; buf ^= 1;
                float4 o4, d4; loadCur(o4, d4);
                rayDone = !beginRay(sc, L, o4, d4, false);
            } else if (!any) L.ray = PT_NONE;                                    // path ended
        }
        if (pair) {                                                             // hand the new shadow rays to the helper groups
            const uint src = lane & ~8u;
            const bool take = __shfl_sync(0xffffffffu, (int)handoff, src) != 0;
            const uint path = __shfl_sync(0xffffffffu, hpath, src);
            const float4 o4 = shfl4(hso, src), d4 = shfl4(hsd, src);
            __syncwarp();                                                       // q.shO/shD[path] (re-read when an instance is left) written by the owner
            if (isHelper && take) { L.ray = path; any = true; rayDone = !beginRay(sc, L, o4, d4, true); }
        }
        // ---- T: top level ----
        if (L.ray != PT_NONE && !rayDone && !hasBlasWork(L)) rayDone = tlasTrip(sc, L, stack, loadCur, tally);
        // ---- B: bottom level, cooperative ----
        while (true) {
            const bool work = hasBlasWork(L);
            if (__ballot_sync(0xffffffffu, work) == 0u) break;
            if (work && coopBlasTrip(sc, L, stack, any, sub, gmask, tally)) rayDone = true;
        }
    }
}

__global__ void __launch_bounds__(128, PT_TAIL_MINBLOCKS) k_tail(ShadeParams P, Queues q, int inBuf, int b0, int maxBounces, uint* cursor) {
    pdlWait(); pdlLaunch();
    const DScene& sc = P.sc;
    const uint n = q.cnt[CNT_EXT(b0)];
    const uint lane = laneId();
    const uint totalWarps = (gridDim.x * blockDim.x) >> 5;
    if (sc.coopTail != 0u && n <= totalWarps * 4u * sc.coopTail) { tailCoop(P, q, inBuf, b0, maxBounces, cursor, n); return; }
    const bool pair = sc.tailPair != 0u && P.nee && n <= totalWarps * sc.tailPair;     // tailPair: paths per warp up to which a path gets two lanes (16 = no queueing)
    const bool isHelper = pair && (lane & 1u) != 0u;
    const uint ownerMask = pair ? 0x55555555u : 0xffffffffu;
    const uint width = min(pair ? 16u : 32u, max(max(1u, pair ? 1u : sc.minWidth), (n + totalWarps - 1u) / totalWarps));
    TravStack stack; stack.init();
    NoTally tally;
    LaneState L;
    L.ray = PT_NONE; L.ngroup = make_uint2(0u, 0u); L.tgroup = make_uint2(0u, 0u); L.curT = PT_NONE; L.sp = 0; L.spBase = 0; L.inBLAS = false;
    int bounce = b0, buf = inBuf;
    bool any = false, extPending = false, rayDone = false;
    bool exhausted = n == 0;
    const uint randIdx0 = P.randIdx;
    auto loadCur = [&](float4& o4, float4& d4) {
        if (any) { o4 = q.shO[L.ray]; d4 = q.shD[L.ray]; } else { o4 = q.rayO[buf][L.ray]; d4 = q.rayD[buf][L.ray]; }
    };
    auto addLight = [&]() {                                                     // unoccluded: kernels.h:806-809
        const uint slot = __float_as_uint(q.shD[L.ray].w);
        const float4 l = P.lights[slot];
        float4 a = P.accucolors[slot];
        a.x += l.x; a.y += l.y; a.z += l.z;
        P.accucolors[slot] = a;
    };
    while (true) {
        // ---- F: idle lanes (pair mode: idle even lanes) claim new paths ----
        const uint activeAll = __ballot_sync(0xffffffffu, L.ray != PT_NONE);
        const uint activeMask = activeAll & ownerMask;
        const uint nActive = __popc(activeMask);
        if (!exhausted && nActive < width) {
            const uint want = width - nActive;
            uint base = 0;
            if (lane == 0) base = atomicAdd(cursor, want);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (base + want >= n) exhausted = true;
            const uint r = __popc(~activeAll & ownerMask & ((1u << lane) - 1u));
            if (!isHelper && L.ray == PT_NONE && r < want && base + r < n) {
                L.ray = base + r; bounce = b0; buf = inBuf; any = false; extPending = false;
                float4 o4, d4; loadCur(o4, d4);
                rayDone = !beginRay(sc, L, o4, d4, false);
            }
        } else if (activeAll == 0u) break;
        // ---- C: a finished ray continues its path: shade, then the shadow ray, then the next extension ray ----
        bool helperBusy = false;
        if (pair) {
            if (isHelper && L.ray != PT_NONE && rayDone) {                      // the helper's shadow ray is done
                if (L.hitId == PT_NONE) addLight();
                L.ray = PT_NONE; rayDone = false;
            }
            __syncwarp();                                                       // its accucolor update before the owner's next shade
            helperBusy = __shfl_xor_sync(0xffffffffu, (int)(L.ray != PT_NONE), 1) != 0;
        }
        bool handoff = false;
        uint hpath = PT_NONE;
        float4 hso = make_float4(0.f, 0.f, 0.f, 0.f), hsd = hso;
        while (!isHelper && L.ray != PT_NONE && rayDone && !helperBusy && !handoff) {
            rayDone = false;
            bool startExt = false;
            if (any) {
                if (L.hitId == PT_NONE) addLight();
                any = false;
                startExt = extPending;
            } else {
                float4 o4, d4; loadCur(o4, d4);
                const uint slot = __float_as_uint(d4.w);
                Hit h; h.type = L.hitType; h.id = L.hitId; h.inst = L.hitInst; h.t = L.tbest;
                ShadeParams Pb = P; Pb.bounce = bounce; Pb.randIdx = randIdx0 + (uint)(bounce - b0);
                ShadeOut so;
                shadeOne(Pb, f3(o4), f3(d4), slot, h, so);
                extPending = so.pushExt && bounce + 1 < maxBounces;
                if (so.pushExt) atomicAdd(q.cnt + CNT_EXT(bounce + 1), 1u);
                if (extPending) {
                    q.rayO[buf ^ 1][L.ray] = make_float4(so.ext.o.x, so.ext.o.y, so.ext.o.z, so.ext.len);
                    q.rayD[buf ^ 1][L.ray] = make_float4(so.ext.d.x, so.ext.d.y, so.ext.d.z, __uint_as_float(slot));
                }
                if (so.pushShadow) {
                    atomicAdd(q.cnt + CNT_SHADOW(bounce), 1u);
                    const float4 so4 = make_float4(so.shadow.o.x, so.shadow.o.y, so.shadow.o.z, so.shadow.len);
                    const float4 sd4 = make_float4(so.shadow.d.x, so.shadow.d.y, so.shadow.d.z, __uint_as_float(slot));
                    q.shO[L.ray] = so4; q.shD[L.ray] = sd4;
                    if (pair) { handoff = true; hpath = L.ray; hso = so4; hsd = sd4; startExt = extPending; }
                    else { any = true; rayDone = !beginRay(sc, L, so4, sd4, true); }
                } else startExt = extPending;
            }
            if (startExt) {
                extPending = false; bounce++; buf ^= 1;
                float4 o4, d4; loadCur(o4, d4);
                rayDone = !beginRay(sc, L, o4, d4, false);
            } else if (!any) L.ray = PT_NONE;                                    // path ended
        }
        if (pair) {                                                             // hand the new shadow rays to the helpers
            const uint src = lane & ~1u;
            const bool take = __shfl_sync(0xffffffffu, (int)handoff, src) != 0;
            const uint path = __shfl_sync(0xffffffffu, hpath, src);
            const float4 o4 = shfl4(hso, src), d4 = shfl4(hsd, src);
            __syncwarp();                                                       // q.shO/shD[path] (re-read when an instance is left) written by the owner
            if (isHelper && take) { L.ray = path; any = true; rayDone = !beginRay(sc, L, o4, d4, true); }
        }
        // ---- T: top level ----
        if (L.ray != PT_NONE && !rayDone && !hasBlasWork(L)) rayDone = tlasTrip(sc, L, stack, loadCur, tally);
        // ---- B: bottom level ----
        const uint nB = __popc(__ballot_sync(0xffffffffu, hasBlasWork(L)));
        const uint minKeep = max(1u, (nB * PT_KEEP_NUM + PT_KEEP_DEN - 1u) / PT_KEEP_DEN);
        while (true) {
            const bool work = hasBlasWork(L);
            if (__popc(__ballot_sync(0xffffffffu, work)) < minKeep) break;
            if (work && blasTrip(sc, L, stack, any, tally)) rayDone = true;
        }
    }
}

// ---- accumulate + statistics -------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_accumulate(TileMap map, const float4* __restrict__ accucolors, const float4* __restrict__ albedoSample,
                                                    float4* image, float4* albedo, const uint* __restrict__ cnt, unsigned long long* stats, uint maxBounces, uint waveBounces,
                                                    volatile uint* feedback) {
    const uint slot = map.slotBase + blockIdx.x * blockDim.x + threadIdx.x;
    uint x, y;
    pdlWait(); pdlLaunch();
    if (slotToPixel(map, slot, x, y)) {
        const float4 c = accucolors[slot];
        const uint pix = x + y * map.W;
        const float4 p = image[pix];
        image[pix] = make_float4(fmaxf(0.0f, p.x) + c.x, fmaxf(0.0f, p.y) + c.y, fmaxf(0.0f, p.z) + c.z, p.w + 1.0f);
        if ((__float_as_uint(c.w) & 0xffu) != 0u) {                                // albedoSet: kernels.h:56-62, deferred from shade
            const float4 nv = albedoSample[slot];
            const float4 a = albedo[pix];
            albedo[pix] = make_float4(fmaxf(0.0f, a.x) + nv.x, fmaxf(0.0f, a.y) + nv.y, fmaxf(0.0f, a.z) + nv.z, a.w + 1.0f);
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {           // one thread per launch; launches of one sample's tile ranges run concurrently: atomics
        unsigned long long ext = 0, sh = 0;
        unsigned long long wext = 0, wsh = 0;
        for (uint b = 0; b < maxBounces; b++) {
            ext += cnt[CNT_EXT(b)]; sh += cnt[CNT_SHADOW(b)];
            if (b < waveBounces) { wext += cnt[CNT_EXT(b)]; wsh += cnt[CNT_SHADOW(b)]; }
        }
        if (feedback) {                              // per-bounce queue sizes for the host's next scheduling decision (mapped host memory)
            for (uint b = 0; b < maxBounces; b++) feedback[1 + b] = cnt[CNT_EXT(b)];
            __threadfence_system();
            feedback[0] = feedback[0] + 1u;
        }
        atomicAdd(stats + 0, (unsigned long long)cnt[CNT_EXT(0)]); atomicAdd(stats + 1, ext); atomicAdd(stats + 2, sh); atomicAdd(stats + 3, ext);
        if (map.slotBase == 0u) atomicAdd(stats + 4, 1ull);
        atomicAdd(stats + 5, wext); atomicAdd(stats + 6, wsh);
    }
}
// this rank's pixels in slot order (the tile partition): what a rank ships to the host / to rank 0
__global__ void __launch_bounds__(256) k_pack_tiles(TileMap map, const float4* __restrict__ image, float4* out) {
    const uint slot = blockIdx.x * blockDim.x + threadIdx.x;
    uint x, y;
    out[slot] = slotToPixel(map, slot, x, y) ? image[x + y * map.W] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
}
__global__ void k_clear_image(float4* a, float4* b, uint n) {
    const uint i = blockIdx.x * blockDim.x + threadIdx.x;
    pdlWait(); pdlLaunch();
    if (i < n) { a[i] = make_float4(0, 0, 0, 0); b[i] = make_float4(0, 0, 0, 0); }
}

// ---- radiance cache (path guiding) ----------------------------------------------------------------------------------
// Layout: rcRadiance = 3 x float4 per triangle {radianceCache[0..3]}{radianceCache[4..7]}{radianceTotal,-,-,-};
// rcAdditions = SoA planes [16][T] (additionCache[8] | additionCacheCount[8]) -- one contiguous buffer,
// which is what the multi-GPU all-reduce sums.
// ---- top-level BVH on the device: BuildTopLevelBVH / FindBestMatch / transformBox (scene.h:27-118) ------------------
// One block.  Same agglomerative clustering as the reference's host code, node for node: leaves sit at 2n-2-i, the walk
// A -> bestMatch(A) -> ... merges a pair when the relation is mutual, new nodes take the next lower index.  The best
// match is the first minimum in ascending node order (the reference iterates a std::map with a strict <), found here
// as the lexicographic minimum of (merged area, node index) over the block.
struct TlasBox { float lo[3], hi[3]; };
__device__ __forceinline__ TlasBox tlasLoadBox(const pt_tlas_node* n) {
    TlasBox b; b.lo[0] = n->vmin.x; b.lo[1] = n->vmin.y; b.lo[2] = n->vmin.z; b.hi[0] = n->vmax.x; b.hi[1] = n->vmax.y; b.hi[2] = n->vmax.z; return b;
}
__device__ __forceinline__ float tlasMergedArea(const TlasBox& a, const TlasBox& o) {
    float lo[3], hi[3];
    for (int k = 0; k < 3; k++) { lo[k] = fminf(a.lo[k], o.lo[k]); hi[k] = fmaxf(a.hi[k], o.hi[k]); }
    const float lx = hi[0] - lo[0], ly = hi[1] - lo[1], lz = hi[2] - lo[2];
    return 2 * lx * ly + 2 * lx * lz + 2 * ly * lz;                                          // Box::getSurfaceArea, types.h:154-161
}
__device__ uint tlasBestMatch(const pt_tlas_node* dest, const unsigned char* alive, uint lowest, uint total, uint node, float* sArea, uint* sIdx) {
    const TlasBox nb = tlasLoadBox(dest + node);
    float best = 3.402823466e+38f; uint bi = 0xffffffffu;
    for (uint i = lowest + threadIdx.x; i < total; i += blockDim.x) {
        if (!alive[i] || i == node) continue;
        const float a = tlasMergedArea(nb, tlasLoadBox(dest + i));
        if (a < best) { best = a; bi = i; }
    }
    for (int off = 16; off > 0; off >>= 1) {
        const float oa = __shfl_down_sync(0xffffffffu, best, off); const uint oi = __shfl_down_sync(0xffffffffu, bi, off);
        if (oi != 0xffffffffu && (bi == 0xffffffffu || oa < best || (oa == best && oi < bi))) { best = oa; bi = oi; }
    }
    __syncthreads();                                                                         // previous readers of sArea / sIdx are done
    if ((threadIdx.x & 31) == 0) { sArea[threadIdx.x >> 5] = best; sIdx[threadIdx.x >> 5] = bi; }
    __syncthreads();
    best = sArea[0]; bi = sIdx[0];
    for (uint w = 1; w < blockDim.x / 32; w++) {
        const float oa = sArea[w]; const uint oi = sIdx[w];
        if (oi != 0xffffffffu && (bi == 0xffffffffu || oa < best || (oa == best && oi < bi))) { best = oa; bi = oi; }
    }
    return bi == 0xffffffffu ? node : bi;
}
__global__ void __launch_bounds__(256) k_build_tlas(const pt_instance* inst, const DModel* models, const float4* modelBox, uint n, pt_tlas_node* dest,
                                                   unsigned char* alive, uint* depthScratch, uint* maxDepthOut) {
    __shared__ float sArea[8]; __shared__ uint sIdx[8];
    const uint total = 2 * n - 1;
    for (uint i = threadIdx.x; i < total; i += blockDim.x) alive[i] = 0;
    __syncthreads();
    for (uint i = threadIdx.x; i < n; i += blockDim.x) {                                     // leaves: transformBox of the model's root box, scene.h:27-52
        const float4 lo = modelBox[2 * inst[i].model_id], hi = modelBox[2 * inst[i].model_id + 1];
        const float blo[3] = {lo.x, lo.y, lo.z}, bhi[3] = {hi.x, hi.y, hi.z};
        const float d[3] = {bhi[0] - blo[0], bhi[1] - blo[1], bhi[2] - blo[2]};
        float pts[8][3];
        for (int a = 0; a < 3; a++) { pts[0][a] = blo[a]; pts[4][a] = bhi[a]; }
        for (int k = 0; k < 3; k++) for (int a = 0; a < 3; a++) {
            pts[1 + k][a] = blo[a] + (a == k ? d[a] : 0.0f);
            pts[5 + k][a] = bhi[a] - (a == k ? d[a] : 0.0f);
        }
        const pt_mat4x3 t = inst[i].transform;
        float rlo[3] = {9999999, 9999999, 9999999}, rhi[3] = {-9999999, -9999999, -9999999};  // Box::insideOut, types.h:170-176
        for (int p = 0; p < 8; p++) {
            const float wp[3] = {pts[p][0] * t.r1.x + pts[p][1] * t.r1.y + pts[p][2] * t.r1.z + 1.0f * t.r1.w,
                                 pts[p][0] * t.r2.x + pts[p][1] * t.r2.y + pts[p][2] * t.r2.z + 1.0f * t.r2.w,
                                 pts[p][0] * t.r3.x + pts[p][1] * t.r3.y + pts[p][2] * t.r3.z + 1.0f * t.r3.w};
            for (int a = 0; a < 3; a++) { rlo[a] = fminf(wp[a], rlo[a]); rhi[a] = fmaxf(wp[a], rhi[a]); }
        }
        pt_tlas_node nd;
        nd.vmin = pt_float3{rlo[0], rlo[1], rlo[2]}; nd.vmax = pt_float3{rhi[0], rhi[1], rhi[2]}; nd.children = 0; nd.leaf = i;
        dest[total - 1 - i] = nd; alive[total - 1 - i] = 1;
    }
    __syncthreads();
    uint nodeCount = n - 1, live = n;
    uint A = n - 1;                                                                          // list.begin(): the lowest live index
    uint B = live > 1 ? tlasBestMatch(dest, alive, nodeCount, total, A, sArea, sIdx) : A;
    while (live > 1) {
        const uint Cc = tlasBestMatch(dest, alive, nodeCount, total, B, sArea, sIdx);
        if (A == Cc) {
            const uint idx = --nodeCount;
            if (threadIdx.x == 0) {
                const TlasBox a = tlasLoadBox(dest + A), b = tlasLoadBox(dest + B);
                float lo[3] = {9999999, 9999999, 9999999}, hi[3] = {-9999999, -9999999, -9999999};                      // Box::merged, types.h:178-184
                for (int k = 0; k < 3; k++) { lo[k] = fminf(lo[k], a.lo[k]); hi[k] = fmaxf(hi[k], a.hi[k]); }
                for (int k = 0; k < 3; k++) { lo[k] = fminf(lo[k], b.lo[k]); hi[k] = fmaxf(hi[k], b.hi[k]); }
                pt_tlas_node nd = dest[idx];
                nd.vmin = pt_float3{lo[0], lo[1], lo[2]}; nd.vmax = pt_float3{hi[0], hi[1], hi[2]}; nd.children = A | (B << 16);
                dest[idx] = nd;
                alive[A] = 0; alive[B] = 0; alive[idx] = 1;
            }
            __syncthreads();
            live--;
            A = idx;
            B = live > 1 ? tlasBestMatch(dest, alive, nodeCount, total, A, sArea, sIdx) : A;
        } else { A = B; B = Cc; }
    }
    __syncthreads();
    if (threadIdx.x == 0) {                                                                  // depth in edges (children always sit above their parent)
        uint deepest = 0;
        depthScratch[0] = 0;
        for (uint i = 0; i < total; i++) {
            const uint c = dest[i].children;
            if (c == 0) { deepest = max(deepest, depthScratch[i]); continue; }
            depthScratch[c & 0xffffu] = depthScratch[i] + 1; depthScratch[(c >> 16) & 0xffffu] = depthScratch[i] + 1;
        }
        *maxDepthOut = deepest;
    }
}

__global__ void k_rc_init(float4* rc, float* add, uint T) {                          // kernels.h:848-861
    const uint i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= T) return;
    rc[(size_t)i * 3] = make_float4(0.1f, 0.1f, 0.1f, 0.1f);
    rc[(size_t)i * 3 + 1] = make_float4(0.1f, 0.1f, 0.1f, 0.1f);
    rc[(size_t)i * 3 + 2] = make_float4(8 * 0.1f, 0.0f, 0.0f, 0.0f);
    for (uint b = 0; b < 16; b++) add[(size_t)b * T + i] = 0.0f;
}
__global__ void __launch_bounds__(256) k_rc_update(TileMap map, const float4* __restrict__ accucolors,
                                                   const pt_sample_cache* __restrict__ sampleCache, float* add, uint T) {   // kernels.h:863-882
    const uint slot = map.slotBase + blockIdx.x * blockDim.x + threadIdx.x;
    uint x, y;
    pdlWait(); pdlLaunch();
    if (!slotToPixel(map, slot, x, y)) return;
    const float3 totalEnergy = f3(accucolors[slot]);
    for (uint bounce = 0; bounce < PT_MAX_CACHE_DEPTH; bounce++) {
        const pt_sample_cache c = sampleCache[(size_t)bounce * map.nSlots + slot];
        if (c.sample_type == 1) return;
        if (c.sample_type == 0) continue;
        const float energy = fminf(100.0f, luminance(totalEnergy / f3(c.cum_mask.x, c.cum_mask.y, c.cum_mask.z)));
        const uint bk = PT_IDX(c.cache_bucket_id, 8u, PT_CHK_CACHE), tr = PT_IDX(c.triangle_id, T, PT_CHK_CACHE);
        atomicAdd(add + (size_t)bk * T + tr, energy);          // RED.ADD.F32, no return value used
        atomicAdd(add + (size_t)(8 + bk) * T + tr, 1.0f);
    }
}
__global__ void __launch_bounds__(256) k_rc_propagate(float4* rc, float* add, uint T) {                                   // kernels.h:884-905
    const uint i = blockIdx.x * blockDim.x + threadIdx.x;
    pdlWait(); pdlLaunch();
    if (i >= T) return;
    float4 r0 = rc[(size_t)i * 3], r1 = rc[(size_t)i * 3 + 1], r2 = rc[(size_t)i * 3 + 2];
    float v[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
    float total = r2.x;
    const float alpha = 0.95f;
    bool dirty = false;
#pragma unroll
    for (uint t = 0; t < 8; t++) {
        const float additionCount = add[(size_t)(8 + t) * T + i];
        if (additionCount < PT_EPS) continue;
        const float oldValue = v[t];
        const float incomingEnergy = add[(size_t)t * T + i] / additionCount;
        const float newValue = clampf(alpha * oldValue + (1 - alpha) * incomingEnergy, 0.1f, 2.0f);
        total += newValue - oldValue;
        v[t] = newValue;
        add[(size_t)t * T + i] = 0.0f;
        add[(size_t)(8 + t) * T + i] = 0.0f;
        dirty = true;
    }
    if (dirty) {
        rc[(size_t)i * 3] = make_float4(v[0], v[1], v[2], v[3]);
        rc[(size_t)i * 3 + 1] = make_float4(v[4], v[5], v[6], v[7]);
        rc[(size_t)i * 3 + 2] = make_float4(total, 0.0f, 0.0f, 0.0f);
    }
}

// ---- multi-GPU: reduce + propagate + broadcast over NVLink peer memory in place of all-reduce + propagate -----------------
// Every rank maps every peer's radiance-cache block (cudaIpc, pt_peer_connect).  Triangles are split into `world`
// contiguous ranges; the OWNER of a range reads all ranks' additions for it straight out of their HBM (fixed rank order:
// deterministic sums), runs the reference's propagate step once (kernels.h:884-905) and stores the new radianceCache[8] /
// radianceTotal into every rank's copy.  Per rank and frame that moves 64 B/triangle in and <= 48 B/triangle out over
// NVLink instead of the ~2 x 64 B of a ring all-reduce, needs no intermediate buffer, and the cross-GPU ordering is three
// flag words per peer:  arrive[r] = "rank r's additions of this epoch are complete",  done[r] = "owner r has read every
// rank's additions and written every rank's cache".  Flags live at the head of each rank's block and are written by the
// peers with release stores at system scope; waiting is a bounded spin (a missing peer sets an error word, never hangs).
#define PT_PEER_MAX 16
#define PT_PEER_ARRIVE 0        /* uint[16] at the head of the block */
#define PT_PEER_DONE 64         /* uint[16] */
#define PT_PEER_COUNTER 128     /* blocks of k_peer_reduce_propagate that finished */
#define PT_PEER_ERROR 129       /* set when a wait timed out */
#define PT_PEER_MAGIC 130       /* 0x50544230 + rank, checked by pt_peer_connect through the mapped pointer */
struct PeerSet {
    unsigned char* block[PT_PEER_MAX];      // base of every rank's block (own rank: the local pointer)
    uint rank, world;
    size_t addOffset, rcOffset;             // byte offsets of rcAdditions / rcRadiance inside a block (same on every rank)
};
__device__ __forceinline__ uint* peerFlags(const PeerSet& ps, uint r) { return reinterpret_cast<uint*>(ps.block[r]); }
__device__ __forceinline__ void stReleaseSys(uint* p, uint v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint ldAcquireSys(const uint* p) { uint v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ float ldPeer(const float* p) { float v; asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void stPeer4(float4* p, float4 v) { asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory"); }
__device__ __forceinline__ unsigned long long globalTimerNs() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
// threads 0..world-1 of the block wait until flag word [base + t] of the LOCAL block has reached `epoch`.  Returns false
// (block-uniform) when a wait timed out or an error was already raised for this exchange: the caller then skips its
// part of the epoch.  The error word is raised on EVERY rank (system-scope exchange over NVLink), so no rank goes on to
// consume a half-finished exchange silently; pt_finish reports and clears it.
__device__ __forceinline__ bool peerWait(const PeerSet& ps, uint base, uint epoch, unsigned long long timeoutNs) {
    __shared__ int peerWaitFailed;
    if (threadIdx.x == 0) peerWaitFailed = 0;
    __syncthreads();
    if (threadIdx.x < ps.world) {
        uint* f = peerFlags(ps, ps.rank);
        const unsigned long long t0 = globalTimerNs();
        while ((int)(ldAcquireSys(f + base + threadIdx.x) - epoch) < 0) {
            if (ldAcquireSys(f + PT_PEER_ERROR) != 0u) { peerWaitFailed = 1; break; }
            if (globalTimerNs() - t0 > timeoutNs) {
                for (uint r = 0; r < ps.world; r++) atomicExch_system(peerFlags(ps, r) + PT_PEER_ERROR, 1u + base + threadIdx.x);
                peerWaitFailed = 1;
                break;
            }
            __nanosleep(64);
        }
        if (threadIdx.x == 0 && ldAcquireSys(f + PT_PEER_ERROR) != 0u) peerWaitFailed = 1;      // raised by another rank meanwhile
    }
    __syncthreads();
    return peerWaitFailed == 0;
}
__global__ void k_peer_arrive(PeerSet ps, uint epoch) {
    pdlWait(); pdlLaunch();
    __threadfence_system();                                  // this rank's k_rc_update (previous kernel on the stream) before the flag
    if (threadIdx.x < ps.world) stReleaseSys(peerFlags(ps, threadIdx.x) + PT_PEER_ARRIVE + ps.rank, epoch);
}
__global__ void __launch_bounds__(256) k_peer_reduce_propagate(PeerSet ps, uint epoch, uint T, uint lo, uint hi, unsigned long long timeoutNs) {
    pdlWait(); pdlLaunch();
    const bool ok = peerWait(ps, PT_PEER_ARRIVE, epoch, timeoutNs);     // false: a rank is missing -- nothing is summed, stored or cleared this epoch
    const uint i = lo + blockIdx.x * blockDim.x + threadIdx.x;
    if (ok && i < hi) {
        float cnt[8], en[8];
#pragma unroll
        for (uint t = 0; t < 8; t++) { cnt[t] = 0.0f; en[t] = 0.0f; }
        // Fixed rank order: the sums do not depend on arrival order.  NVLink round trips are the cost here, so the loads
        // are issued in batches that do not wait for each other: the 8 counts of rank r+1 go out before the energies of
        // rank r are requested (only where the count is non-zero: a zero count means a zero energy sum).
        float cNext[8];
        {
            const float* add0 = reinterpret_cast<const float*>(ps.block[0] + ps.addOffset);
#pragma unroll
            for (uint t = 0; t < 8; t++) cNext[t] = ldPeer(add0 + (size_t)(8 + t) * T + i);
        }
        for (uint r = 0; r < ps.world; r++) {
            const float* add = reinterpret_cast<const float*>(ps.block[r] + ps.addOffset);
            float c[8], e[8];
#pragma unroll
            for (uint t = 0; t < 8; t++) c[t] = cNext[t];
            if (r + 1 < ps.world) {
                const float* addN = reinterpret_cast<const float*>(ps.block[r + 1] + ps.addOffset);
#pragma unroll
                for (uint t = 0; t < 8; t++) cNext[t] = ldPeer(addN + (size_t)(8 + t) * T + i);
            }
#pragma unroll
            for (uint t = 0; t < 8; t++) { e[t] = 0.0f; if (c[t] != 0.0f) e[t] = ldPeer(add + (size_t)t * T + i); }
#pragma unroll
            for (uint t = 0; t < 8; t++) { cnt[t] += c[t]; en[t] += e[t]; }
        }
        float4* rc = reinterpret_cast<float4*>(ps.block[ps.rank] + ps.rcOffset);
        const float4 r0 = rc[(size_t)i * 3], r1 = rc[(size_t)i * 3 + 1], r2 = rc[(size_t)i * 3 + 2];
        float v[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
        float total = r2.x;
        const float alpha = 0.95f;
        bool dirty = false;
#pragma unroll
        for (uint t = 0; t < 8; t++) {                       // kernels.h:890-903, as k_rc_propagate
            if (cnt[t] < PT_EPS) continue;
            const float oldValue = v[t];
            const float incomingEnergy = en[t] / cnt[t];
            const float newValue = clampf(alpha * oldValue + (1 - alpha) * incomingEnergy, 0.1f, 2.0f);
            total += newValue - oldValue;
            v[t] = newValue;
            dirty = true;
        }
        if (dirty)
            for (uint r = 0; r < ps.world; r++) {
                float4* dst = reinterpret_cast<float4*>(ps.block[r] + ps.rcOffset) + (size_t)i * 3;
                stPeer4(dst, make_float4(v[0], v[1], v[2], v[3]));
                stPeer4(dst + 1, make_float4(v[4], v[5], v[6], v[7]));
                stPeer4(dst + 2, make_float4(total, 0.0f, 0.0f, 0.0f));
            }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        uint* f = peerFlags(ps, ps.rank);
        if (atomicAdd(f + PT_PEER_COUNTER, 1u) == gridDim.x - 1u) {             // last block of this owner: tell every rank
            f[PT_PEER_COUNTER] = 0u;
            __threadfence_system();
            for (uint r = 0; r < ps.world; r++) stReleaseSys(peerFlags(ps, r) + PT_PEER_DONE + ps.rank, epoch);
        }
    }
}
// Once every owner is done nobody reads this rank's additions any more: clear them for the next frame (k_rc_propagate
// clears the entries it consumed; untouched ones are zero already, so the state is the same).
__global__ void __launch_bounds__(256) k_peer_finish(PeerSet ps, uint epoch, uint T, unsigned long long timeoutNs) {
    pdlWait(); pdlLaunch();
    if (!peerWait(ps, PT_PEER_DONE, epoch, timeoutNs)) return;          // failed exchange: keep the additions, pt_finish reports the error
    float* add = reinterpret_cast<float*>(ps.block[ps.rank] + ps.addOffset);
    const size_t n = (size_t)16 * T;
    for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (size_t)gridDim.x * blockDim.x) add[k] = 0.0f;
}

// ---- multi-GPU: one framebuffer on rank 0 --------------------------------------------------------------------------------
// pt_gather_image_async: every rank writes its tiles of the accumulated image straight into a slot of the image ring in
// rank 0's HBM (NVLink stores through the mapped block), then raises its ARRIVE flag; rank 0 waits for all flags, ships the
// slot to the host with ONE copy and tells the ranks (CONSUMED) that the slot may be reused PT_GATHER_RING epochs later.
// Flags live in the header of the radiance-cache block every rank already maps (pt_peer_connect).
__global__ void __launch_bounds__(256) k_gather_push(PeerSet ps, TileMap map, const float4* __restrict__ image, float4* rootSlot, uint epoch, uint ring,
                                                     uint arriveWord, uint consumedWord, uint counterWord, unsigned long long timeoutNs) {
    __shared__ int failed;
    if (threadIdx.x == 0) {                                  // the slot's previous content (epoch - ring) must have left for the host
        failed = 0;
        uint* f = peerFlags(ps, ps.rank);
        const unsigned long long t0 = globalTimerNs();
        while (epoch > ring && (int)(ldAcquireSys(f + consumedWord) - (epoch - ring)) < 0) {
            if (globalTimerNs() - t0 > timeoutNs) { atomicExch_system(f + PT_PEER_ERROR, 1000u + consumedWord); failed = 1; break; }
            __nanosleep(64);
        }
    }
    __syncthreads();
    const uint slot = blockIdx.x * blockDim.x + threadIdx.x;
    uint x, y;
    if (!failed && slotToPixel(map, slot, x, y)) stPeer4(rootSlot + (size_t)y * map.W + x, image[(size_t)y * map.W + x]);
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        uint* f = peerFlags(ps, ps.rank);
        if (atomicAdd(f + counterWord, 1u) == gridDim.x - 1u) {
            f[counterWord] = 0u;
            __threadfence_system();
            stReleaseSys(peerFlags(ps, 0) + arriveWord + ps.rank, epoch);
        }
    }
}
__global__ void k_gather_wait(PeerSet ps, uint epoch, uint arriveWord, unsigned long long timeoutNs) {      // rank 0, one block
    if (threadIdx.x < ps.world) {
        uint* f = peerFlags(ps, ps.rank);
        const unsigned long long t0 = globalTimerNs();
        while ((int)(ldAcquireSys(f + arriveWord + threadIdx.x) - epoch) < 0) {
            if (globalTimerNs() - t0 > timeoutNs) { atomicExch_system(f + PT_PEER_ERROR, 2000u + threadIdx.x); break; }
            __nanosleep(64);
        }
    }
}
// N ranks, one framebuffer in HOST memory, every rank through its own PCIe link (pt_deliver_image_async).  `packed` holds this
// rank's pixels in slot order (k_pack_tiles: the image itself is free again right after that copy).  One block per local tile;
// a warp stores whole tile rows (32 pixels = 512 contiguous bytes) straight into the mapped, process-shared host image.  The
// epoch is published by k_publish_host, a separate launch behind this one in the stream: a kernel boundary makes the stores
// visible at system scope, so a consumer that sees the epoch sees the pixels, without a fence per block.
__global__ void __launch_bounds__(256) k_push_host(TileMap map, const float4* __restrict__ packed, float4* hostImage) {
    const uint lt = blockIdx.x, w = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    const uint gt = lt * map.world + map.rank;
    const uint ty = gt / map.tilesX, tx = gt - ty * map.tilesX;
    const uint x = tx * 32u + lane;
#pragma unroll
    for (uint k = 0; k < 4u; k++) {
        const uint py = w + 8u * k, y = ty * 32u + py;
        const uint j = (((py >> 2) * 4u + (lane >> 3)) << 5) | ((py & 3u) << 3) | (lane & 7u);      // slot of pixel (lane, py) inside the tile (slotToPixel)
        if (x < map.W && y < map.H) hostImage[(size_t)y * map.W + x] = packed[(size_t)lt * 1024u + j];
    }
}
// Band delivery (pt_deliver_image_async with connected peers): the frame is assembled in rank 0's HBM ring as for the gather, then
// every rank copies ONE contiguous band of it (rows [r H / N, (r+1) H / N)) to the shared host framebuffer with its own copy
// engine -- NVLink in, its own PCIe link out.  k_flags_wait: all ranks' flags `word` on rank `onRank` have reached the epoch
// (any rank may poll rank 0's header through the mapped block); k_band_done: this rank's band has landed -- tells rank 0 (which
// hands the HBM slot back once all bands have left) and publishes the epoch in the host header.
__global__ void k_flags_wait(PeerSet ps, uint onRank, uint word, uint epoch, unsigned long long timeoutNs) {
    if (threadIdx.x < ps.world) {
        uint* f = peerFlags(ps, onRank);
        const unsigned long long t0 = globalTimerNs();
        while ((int)(ldAcquireSys(f + word + threadIdx.x) - epoch) < 0) {
            if (globalTimerNs() - t0 > timeoutNs) { atomicExch_system(peerFlags(ps, ps.rank) + PT_PEER_ERROR, 3000u + word + threadIdx.x); break; }
            __nanosleep(64);
        }
    }
}
__global__ void k_band_done(PeerSet ps, uint word, uint epoch, volatile uint* hostArrived, uint hostEpoch) {
    __threadfence_system();
    stReleaseSys(peerFlags(ps, 0) + word + ps.rank, epoch);
    hostArrived[ps.rank] = hostEpoch;
}
__global__ void k_publish_host(volatile uint* arrived, uint rank, uint epoch) {
    __threadfence_system();
    arrived[rank] = epoch;
}
__global__ void k_gather_release(PeerSet ps, uint epoch, uint consumedWord) {                                 // rank 0, after the copy to the host
    if (threadIdx.x < ps.world) stReleaseSys(peerFlags(ps, threadIdx.x) + consumedWord, epoch);
}

}  // namespace ptb
