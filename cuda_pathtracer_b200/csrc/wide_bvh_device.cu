// wide_bvh_device.cu -- BVH2 -> 8-wide compressed BVH on the device ("output feeds the
// wide-BVH converter directly").  Same node format, same collapse / slot / quantisation rules and the same
// breadth-first numbering as the host converter in wide_bvh.cpp, so the two produce byte-identical arrays
// (tests/test_gpu_bvh_build.py compares them).  One thread builds one wide node; a level of wide nodes is one
// launch pair (count, then emit) with a prefix sum in between that hands out child-node and triangle-record
// indices in the host converter's queue order.  Compiled with -fmad=false like the rest of the library.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <algorithm>
#include <cub/device/device_scan.cuh>
#include "wide_bvh.h"

namespace ptb {
namespace {

typedef unsigned int uint;
typedef unsigned long long u64;
#define WB_THREADS 128

// A node of the temporary binary tree of wide_bvh.cpp, by reference: b == 0 -> BVH2 node a; b > 0 -> triangles of a BVH2 leaf that
// was cut down to <= 3 triangles per leaf: the range [a, a + (b & 0xFFFFFF)), or -- b >> 24 != 0 -- the triangles a + i of the set
// bits i of that 4-bit mask (one side of a 4-triangle leaf cut by wbBestCut4, wide_bvh.h).
struct TRef { uint a, b; };
struct TInfo {
    float lo[3], hi[3];
    TRef self, left, right;
    uint first, count, mask;      // mask != 0: count = its population; the triangles are first + i of its set bits
    bool internal;
};
__constant__ int c_cut4;          // 1: 4-triangle leaves are cut by wbBestCut4, 0: by index (PTB200_LEAF_INDEX_CUT, as the host converter)

__device__ __forceinline__ float stdmin(float a, float b) { return (b < a) ? b : a; }     // std::min / std::max operand rules
__device__ __forceinline__ float stdmax(float a, float b) { return (a < b) ? b : a; }

__device__ void rangeBounds(const pt_triangle_v* tris, uint first, uint count, float* lo, float* hi) {
    for (int a = 0; a < 3; a++) { lo[a] = 3.0e38f; hi[a] = -3.0e38f; }
    for (uint i = first; i < first + count; i++) {
        const float* p = &tris[i].v0.x;                                                     // v0, v1, v2 are contiguous
        for (int v = 0; v < 3; v++)
            for (int a = 0; a < 3; a++) { lo[a] = stdmin(lo[a], p[3 * v + a]); hi[a] = stdmax(hi[a], p[3 * v + a]); }
    }
}

__device__ void maskBounds(const pt_triangle_v* tris, uint first, uint mask, float* lo, float* hi) {
    for (int a = 0; a < 3; a++) { lo[a] = 3.0e38f; hi[a] = -3.0e38f; }
    for (uint i = 0; i < 4; i++) {
        if (!(mask >> i & 1u)) continue;
        const float* p = &tris[first + i].v0.x;
        for (int v = 0; v < 3; v++)
            for (int a = 0; a < 3; a++) { lo[a] = stdmin(lo[a], p[3 * v + a]); hi[a] = stdmax(hi[a], p[3 * v + a]); }
    }
}
__device__ uint bestCut4(const pt_triangle_v* tris, uint first) {
    float area[16];
    for (uint m = 1; m < 15; m++) {
        float lo[3], hi[3]; maskBounds(tris, first, m, lo, hi);
        const float x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2];
        area[m] = (x < 0 || y < 0 || z < 0) ? 0.0f : 2.0f * (x * y + x * z + y * z);
    }
    return wbBestCut4(area);
}

__device__ void loadRef(const float4* bvh2, const pt_triangle_v* tris, TRef r, TInfo& t) {
    t.self = r; t.mask = 0;
    if (r.b == 0) {
        const float4 lo = bvh2[2 * (size_t)r.a], hi = bvh2[2 * (size_t)r.a + 1];
        t.lo[0] = lo.x; t.lo[1] = lo.y; t.lo[2] = lo.z; t.hi[0] = hi.x; t.hi[1] = hi.y; t.hi[2] = hi.z;
        t.first = __float_as_uint(lo.w); t.count = __float_as_uint(hi.w);
        if (t.count == 0) { t.internal = true; t.left = TRef{t.first, 0}; t.right = TRef{t.first + 1, 0}; return; }
    } else if (r.b >> 24) {
        t.first = r.a; t.mask = r.b >> 24; t.count = (uint)__popc(t.mask);
        maskBounds(tris, r.a, t.mask, t.lo, t.hi);
    } else {
        t.first = r.a; t.count = r.b;
        rangeBounds(tris, r.a, r.b, t.lo, t.hi);
    }
    t.internal = t.count > 3;
    if (t.internal && t.count == 4 && c_cut4) {
        const uint lm = bestCut4(tris, t.first), rm = 15u ^ lm;
        t.left = TRef{t.first, (uint)__popc(lm) | (lm << 24)}; t.right = TRef{t.first, (uint)__popc(rm) | (rm << 24)};
    } else if (t.internal) { const uint lc = t.count / 2; t.left = TRef{t.first, lc}; t.right = TRef{t.first + lc, t.count - lc}; }
}

__device__ __forceinline__ float boxArea(const TInfo& t) {
    const float x = t.hi[0] - t.lo[0], y = t.hi[1] - t.lo[1], z = t.hi[2] - t.lo[2];
    return (x < 0 || y < 0 || z < 0) ? 0.0f : 2.0f * (x * y + x * z + y * z);
}

// ---- SAH-optimal collapse (wide_bvh.h): tables ------------------------------------------------------------------------------
__device__ __forceinline__ float rangeArea(const pt_triangle_v* tris, uint first, uint count) {
    float lo[3], hi[3]; rangeBounds(tris, first, count, lo, hi);
    const float x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2];
    return (x < 0 || y < 0 || z < 0) ? 0.0f : 2.0f * (x * y + x * z + y * z);
}
// table of the triangle range of a cut reference leaf (wide_bvh.cpp virtTable); `area` = surface area of the range's own box
template <int DEPTH>
__device__ void rangeTable(const pt_triangle_v* tris, uint first, uint count, float area, WideDP& out) {
    if (count <= 3) { wdpLeaf(out); return; }
    if constexpr (DEPTH == 0) { wdpOpaque(out, area); }
    else {
        const uint lc = count / 2;
        WideDP L, R;
        rangeTable<DEPTH - 1>(tris, first, lc, lc > 3 ? rangeArea(tris, first, lc) : 0.0f, L);
        rangeTable<DEPTH - 1>(tris, first + lc, count - lc, count - lc > 3 ? rangeArea(tris, first + lc, count - lc) : 0.0f, R);
        wdpCombine(out, L, R, area);
    }
}
struct DPView { const WideDP* table; };      // table == nullptr: greedy collapse
__device__ void tableOf(const DPView& dp, const pt_triangle_v* tris, const TInfo& t, WideDP& out) {
    if (t.self.b == 0) out = dp.table[t.self.a];
    else rangeTable<WB_DP_RANGE_DEPTH>(tris, t.first, t.count, boxArea(t), out);
}
// one level of the reference tree, children before parents (k_dp_level runs the levels bottom-up)
__global__ void __launch_bounds__(WB_THREADS) k_dp_expand(const float4* bvh2, uint* order, uint* lvl, uint level, uint maxNodes) {
    const uint begin = lvl[level], end = lvl[level + 1];
    for (uint i = begin + blockIdx.x * WB_THREADS + threadIdx.x; i < end; i += gridDim.x * WB_THREADS) {
        const uint n = order[i];
        const float4 lo = bvh2[2 * (size_t)n], hi = bvh2[2 * (size_t)n + 1];
        if (__float_as_uint(hi.w) != 0u) continue;                     // leaf
        const uint first = __float_as_uint(lo.w);
        const uint pos = atomicAdd(&lvl[WB_DP_MAX_LEVELS + 2], 2u);    // running end of the list (order inside a level is free)
        if (pos + 1 < maxNodes) { order[pos] = first; order[pos + 1] = first + 1; }
    }
}
__global__ void k_dp_close(uint* lvl, uint level, uint maxNodes) { lvl[level + 2] = min(lvl[WB_DP_MAX_LEVELS + 2], maxNodes); }
__global__ void __launch_bounds__(WB_THREADS) k_dp_level(const float4* bvh2, const pt_triangle_v* tris, const uint* order, const uint* lvl, uint level, WideDP* table) {
    const uint begin = lvl[level], end = lvl[level + 1];
    for (uint i = begin + blockIdx.x * WB_THREADS + threadIdx.x; i < end; i += gridDim.x * WB_THREADS) {
        const uint n = order[i];
        const float4 lo = bvh2[2 * (size_t)n], hi = bvh2[2 * (size_t)n + 1];
        const float x = hi.x - lo.x, y = hi.y - lo.y, z = hi.z - lo.z;
        const float area = (x < 0 || y < 0 || z < 0) ? 0.0f : 2.0f * (x * y + x * z + y * z);
        const uint first = __float_as_uint(lo.w), count = __float_as_uint(hi.w);
        WideDP d;
        if (count == 0) wdpCombine(d, table[first], table[first + 1], area);
        else rangeTable<WB_DP_RANGE_DEPTH>(tris, first, count, area, d);
        table[n] = d;
    }
}

// collapse + slot assignment of one wide node (wide_bvh.cpp "collapse" and "assign children to slots")
struct Collapsed { TInfo root; TInfo ch[8]; int childIn[8]; int nch; };

__device__ void collapse(const float4* bvh2, const pt_triangle_v* tris, TRef ref, Collapsed& c, const DPView dp) {
    loadRef(bvh2, tris, ref, c.root);
    c.nch = 0;
    if (!c.root.internal) c.ch[c.nch++] = c.root;                      // degenerate model: the root itself is a leaf
    else if (dp.table) {
        // optimal cover (wide_bvh.cpp Cover::run): depth-first, left before right, every subtree within its budget
        WideDP d; tableOf(dp, tris, c.root, d);
        const int k = d.k[7];
        TRef stackRef[8]; int stackBudget[8]; int sp = 0;
        stackRef[sp] = c.root.right; stackBudget[sp++] = 8 - k;
        stackRef[sp] = c.root.left; stackBudget[sp++] = k;
        while (sp > 0) {
            --sp;
            const int budget = stackBudget[sp];
            TInfo t; loadRef(bvh2, tris, stackRef[sp], t);
            int b = 1;
            if (t.internal) { tableOf(dp, tris, t, d); b = wdpEffective(d, budget); }
            if (b == 1) { c.ch[c.nch++] = t; continue; }
            const int kl = d.k[b - 1];
            stackRef[sp] = t.right; stackBudget[sp++] = b - kl;
            stackRef[sp] = t.left; stackBudget[sp++] = kl;
        }
    }
    else { loadRef(bvh2, tris, c.root.left, c.ch[0]); loadRef(bvh2, tris, c.root.right, c.ch[1]); c.nch = 2; }
    while (!dp.table && c.nch < 8) {
        int best = -1; float bestArea = -1.0f;
        for (int i = 0; i < c.nch; i++)
            if (c.ch[i].internal) { const float a = boxArea(c.ch[i]); if (a > bestArea) { bestArea = a; best = i; } }
        if (best < 0) break;
        const TRef l = c.ch[best].left, r = c.ch[best].right;
        loadRef(bvh2, tris, l, c.ch[best]); loadRef(bvh2, tris, r, c.ch[c.nch++]);
    }
    float ctr[3];
    for (int a = 0; a < 3; a++) ctr[a] = 0.5f * (c.root.lo[a] + c.root.hi[a]);
    float cost[8][8];
    for (int i = 0; i < c.nch; i++) {
        float d[3];
        for (int a = 0; a < 3; a++) d[a] = 0.5f * (c.ch[i].lo[a] + c.ch[i].hi[a]) - ctr[a];
        for (int s = 0; s < 8; s++) {
            const float sx = (s & 4) ? 1.0f : -1.0f, sy = (s & 2) ? 1.0f : -1.0f, sz = (s & 1) ? 1.0f : -1.0f;
            cost[i][s] = d[0] * sx + d[1] * sy + d[2] * sz;
        }
    }
    int slotOf[8];
    for (int i = 0; i < 8; i++) { slotOf[i] = -1; c.childIn[i] = -1; }
    for (int k = 0; k < c.nch; k++) {
        int bi = -1, bs = -1; float bc = -3.0e38f;
        for (int i = 0; i < c.nch; i++) if (slotOf[i] < 0)
            for (int s = 0; s < 8; s++) if (c.childIn[s] < 0 && cost[i][s] > bc) { bc = cost[i][s]; bi = i; bs = s; }
        slotOf[bi] = bs; c.childIn[bs] = bi;
    }
}

// pass 1 of a level: how many internal children / leaf triangles each wide node will emit
__global__ void __launch_bounds__(WB_THREADS) k_wide_count(const float4* bvh2, const pt_triangle_v* tris, const TRef* refs, uint levelBegin, uint levelSize, u64* counts, const DPView dp) {
    const uint i = blockIdx.x * WB_THREADS + threadIdx.x;
    if (i > levelSize) return;
    if (i == levelSize) { counts[i] = 0; return; }
    Collapsed c;
    collapse(bvh2, tris, refs[levelBegin + i], c, dp);
    uint nInternal = 0, nTris = 0;
    for (int s = 0; s < 8; s++) {
        if (c.childIn[s] < 0) continue;
        const TInfo& cn = c.ch[c.childIn[s]];
        if (cn.internal) nInternal++; else nTris += cn.count;
    }
    counts[i] = ((u64)nInternal << 32) | nTris;
}

// pass 2: write the node, its triangle records and the references of its internal children
__global__ void __launch_bounds__(WB_THREADS) k_wide_emit(const float4* bvh2, const pt_triangle_v* tris, TRef* refs, uint levelBegin, uint levelSize, const u64* scan,
                                                           uint nodeBase, uint triCursor, WideNode* nodes, WideTri* wtris, const DPView dp) {
    const uint i = blockIdx.x * WB_THREADS + threadIdx.x;
    if (i >= levelSize) return;
    Collapsed c;
    collapse(bvh2, tris, refs[levelBegin + i], c, dp);
    const TInfo& r = c.root;
    uint ebyte[3]; float scale[3];
    for (int a = 0; a < 3; a++) {                                                           // quantisation frame
        const float ext = r.hi[a] - r.lo[a];
        int e = -126;
        if (ext > 0.0f) { int ex; const float m = frexpf(ext / 255.0f, &ex); e = m == 0.5f ? ex - 1 : ex; }
        e = max(-126, min(127, e));
        while (e < 127 && ceilf(ext / __uint_as_float((uint)(e + 127) << 23)) > 255.0f) e++;
        ebyte[a] = (uint)(e + 127);
        scale[a] = __uint_as_float((uint)(e + 127) << 23);                                  // ldexp(1, e), e >= -126
    }
    uint w[20];
    for (int k = 0; k < 20; k++) w[k] = 0;
    unsigned char* bytes = (unsigned char*)w;
    const uint nextLevelBegin = levelBegin + levelSize;                                     // local index of the first node of the next level
    const uint childLocal = nextLevelBegin + (uint)(scan[i] >> 32);
    const uint triBase = triCursor + (uint)(scan[i] & 0xffffffffull);
    uint imask = 0, nInternal = 0, triOffset = 0;
    for (int s = 0; s < 8; s++) {
        if (c.childIn[s] < 0) continue;
        const TInfo& cn = c.ch[c.childIn[s]];
        unsigned char meta;
        if (cn.internal) {
            imask |= 1u << s; meta = (unsigned char)(0x20 | (24 + s));
            refs[childLocal + nInternal] = cn.self;
            nInternal++;
        } else {
            const unsigned char unary[4] = {0, 1, 3, 7};
            meta = (unsigned char)((unary[cn.count] << 5) | triOffset);
            uint k = 0;
            for (uint t = cn.first; t < cn.first + (cn.mask ? 4u : cn.count); t++) {
                if (cn.mask && !(cn.mask >> (t - cn.first) & 1u)) continue;
                const pt_triangle_v tv = tris[t];
                WideTri wt;
                wt.v0[0] = tv.v0.x; wt.v0[1] = tv.v0.y; wt.v0[2] = tv.v0.z; wt.id = t;
                wt.e1[0] = tv.v1.x - tv.v0.x; wt.e1[1] = tv.v1.y - tv.v0.y; wt.e1[2] = tv.v1.z - tv.v0.z; wt.pad1 = 0;   // kernels.h:171
                wt.e2[0] = tv.v2.x - tv.v0.x; wt.e2[1] = tv.v2.y - tv.v0.y; wt.e2[2] = tv.v2.z - tv.v0.z; wt.pad2 = 0;   // kernels.h:172
                wtris[triBase + triOffset + k++] = wt;
            }
            triOffset += cn.count;
        }
        bytes[24 + s] = meta;
        for (int a = 0; a < 3; a++) {
            float qlo = floorf((cn.lo[a] - r.lo[a]) / scale[a]);
            float qhi = ceilf((cn.hi[a] - r.lo[a]) / scale[a]);
            while (qlo > 0.0f && r.lo[a] + qlo * scale[a] > cn.lo[a]) qlo -= 1.0f;
            while (qhi < 255.0f && r.lo[a] + qhi * scale[a] < cn.hi[a]) qhi += 1.0f;
            qlo = stdmax(0.0f, stdmin(255.0f, qlo)); qhi = stdmax(0.0f, stdmin(255.0f, qhi));
            bytes[32 + a * 8 + s] = (unsigned char)qlo;
            bytes[56 + a * 8 + s] = (unsigned char)qhi;
        }
    }
    w[0] = __float_as_uint(r.lo[0]); w[1] = __float_as_uint(r.lo[1]); w[2] = __float_as_uint(r.lo[2]);
    w[3] = ebyte[0] | (ebyte[1] << 8) | (ebyte[2] << 16) | (imask << 24);
    w[4] = nodeBase + childLocal; w[5] = triBase;
    uint4* dst = (uint4*)(nodes + nodeBase + levelBegin + i);
    for (int k = 0; k < 5; k++) dst[k] = make_uint4(w[4 * k], w[4 * k + 1], w[4 * k + 2], w[4 * k + 3]);
}

}  // namespace

// Upper bound on the wide nodes of one model: every wide node is an internal node of the temporary binary tree.
size_t wideNodeBound(uint32_t n_bvh2, uint32_t n_tris) { return (size_t)(n_bvh2 / 2) + n_tris + 2; }

// Appends one model's wide BVH to device arrays.  d_bvh2 / d_tris are device pointers (d_tris = the scene's whole
// TriangleV array in reference order); nodes are written at nodes[nodeBase ...], triangle records at wtris[triBase ...].
// `refs` (wideNodeBound entries of 8 bytes), `counts` and `scan` (wideNodeBound + 1 u64 each) and `tmp` are scratch.
int buildWideBVHDevice(cudaStream_t st, const pt_bvh_node* d_bvh2, uint32_t n_bvh2, const pt_triangle_v* d_tris, WideNode* d_nodes, uint32_t nodeBase,
                       WideTri* d_wtris, uint32_t triBase, void* refsScratch, void* countsScratch, void* scanScratch, void* tmp, size_t tmpBytes,
                       uint32_t* nodesUsed, uint32_t* trisUsed, uint32_t* maxDepth, const char** err, void* dpScratch) {
    TRef* refs = (TRef*)refsScratch; u64* counts = (u64*)countsScratch; u64* scan = (u64*)scanScratch;
    const TRef root{0, 0};
    cudaError_t e;
#define WCK(call) do { e = (call); if (e != cudaSuccess) { *err = cudaGetErrorString(e); return PT_ERR_CUDA; } } while (0)
    WCK(cudaMemcpyAsync(refs, &root, sizeof root, cudaMemcpyHostToDevice, st));
    { const int cut4 = getenv("PTB200_LEAF_INDEX_CUT") ? 0 : 1; WCK(cudaMemcpyToSymbolAsync(c_cut4, &cut4, sizeof cut4, 0, cudaMemcpyHostToDevice, st)); }
    // ---- SAH-optimal collapse: the reference tree level by level (lists), then the tables from the deepest level up ----
    DPView dp{nullptr};
    if (dpScratch && n_bvh2) {
        // scratch: [table n_bvh2 x 36 B][order n_bvh2 x 4 B][lvl (WB_DP_MAX_LEVELS + 3) x 4 B: level starts, then the running end]
        WideDP* table = (WideDP*)dpScratch;
        uint* order = (uint*)((char*)dpScratch + (((size_t)n_bvh2 * sizeof(WideDP) + 255) & ~(size_t)255));
        uint* lvl = order + (((size_t)n_bvh2 + 63) & ~(size_t)63);
        uint head[WB_DP_MAX_LEVELS + 3];
        for (int i = 0; i < WB_DP_MAX_LEVELS + 3; i++) head[i] = 0;
        head[1] = 1; head[WB_DP_MAX_LEVELS + 2] = 1;                       // level 0 = [0, 1): the root, order[0] = 0
        WCK(cudaMemcpyAsync(lvl, head, sizeof head, cudaMemcpyHostToDevice, st));
        WCK(cudaMemsetAsync(order, 0, sizeof(uint), st));
        const uint grid = std::max(1u, std::min((n_bvh2 + WB_THREADS - 1) / WB_THREADS, 1184u));
        // levels in chunks of 16 launches pairs; after each chunk the level starts come back and the walk stops at the first empty level
        uint levels = 0; bool tooDeep = true;
        for (uint l0 = 0; l0 < WB_DP_MAX_LEVELS && tooDeep; l0 += 16) {
            for (uint l = l0; l < l0 + 16 && l < WB_DP_MAX_LEVELS; l++) {
                k_dp_expand<<<grid, WB_THREADS, 0, st>>>((const float4*)d_bvh2, order, lvl, l, n_bvh2);
                k_dp_close<<<1, 1, 0, st>>>(lvl, l, n_bvh2);
            }
            WCK(cudaMemcpyAsync(head, lvl, sizeof head, cudaMemcpyDeviceToHost, st));
            WCK(cudaStreamSynchronize(st));
            const uint upTo = std::min<uint>(l0 + 16, WB_DP_MAX_LEVELS);
            for (uint l = l0; l <= upTo; l++)
                if (head[l] == head[l + 1]) { levels = l; tooDeep = false; break; }      // level l is empty: the tree has l levels
        }
        if (!tooDeep) {                                                     // nodes on level WB_DP_MAX_LEVELS: too deep, greedy rule (as the host converter)
            for (int l = (int)levels - 1; l >= 0; l--)
                k_dp_level<<<grid, WB_THREADS, 0, st>>>((const float4*)d_bvh2, d_tris, order, lvl, (uint)l, table);
            dp.table = table;
        }
    }
    uint levelBegin = 0, levelSize = 1, triCursor = triBase, depth = 0;
    while (levelSize) {
        const uint grid1 = (levelSize + 1 + WB_THREADS - 1) / WB_THREADS, grid = (levelSize + WB_THREADS - 1) / WB_THREADS;
        k_wide_count<<<grid1, WB_THREADS, 0, st>>>((const float4*)d_bvh2, d_tris, refs, levelBegin, levelSize, counts, dp);
        size_t need = tmpBytes;
        WCK(cub::DeviceScan::ExclusiveSum(tmp, need, counts, scan, (int)(levelSize + 1), st));
        k_wide_emit<<<grid, WB_THREADS, 0, st>>>((const float4*)d_bvh2, d_tris, refs, levelBegin, levelSize, scan, nodeBase, triCursor, d_nodes, d_wtris, dp);
        u64 totals = 0;
        WCK(cudaMemcpyAsync(&totals, scan + levelSize, sizeof totals, cudaMemcpyDeviceToHost, st));
        WCK(cudaStreamSynchronize(st));
        levelBegin += levelSize; levelSize = (uint)(totals >> 32); triCursor += (uint)(totals & 0xffffffffull); depth++;
    }
    WCK(cudaGetLastError());
#undef WCK
    *nodesUsed = levelBegin; *trisUsed = triCursor - triBase; *maxDepth = depth;
    return PT_OK;
}

size_t wideDpScratchBytes(uint32_t n_bvh2) {
    return (((size_t)n_bvh2 * sizeof(WideDP) + 255) & ~(size_t)255) + ((((size_t)n_bvh2 + 63) & ~(size_t)63) + WB_DP_MAX_LEVELS + 3) * sizeof(uint32_t);
}

size_t wideScanTempBytes(size_t maxItems) {
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, (u64*)nullptr, (u64*)nullptr, (int)maxItems);
    return bytes;
}

}  // namespace ptb
