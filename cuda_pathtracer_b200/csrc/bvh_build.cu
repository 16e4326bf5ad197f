// bvh_build.cu -- binned-SAH BVH2 build on the device (sm_100a), SURVEY.md 8(f) row 2.
//
// Produces, for the same triangles, the SAME tree as the host builder in host_scene.cpp (which is pinned
// against the reference's createBVHBinned, src/bvhBuilder.h:35-274): the same split decisions, the same
// triangle order (and therefore the same triangle ids), the same boxes and the same node numbering.
// The reference builds depth-first with an explicit stack; this builder works level by level over all
// open nodes at once.  Why the result is still identical:
//   * every quantity a split decision reads (centroid bounds, bin counts, bin bounds) is a min/max/count
//     over the node's triangle SET, so the order of accumulation does not matter; they are accumulated
//     with integer atomics on order-preserving keys.  The one order-sensitive detail is the sign of a zero
//     coordinate (SSE min returns its second operand on +0/-0 ties); boxes may differ from the host
//     builder in the SIGN of a zero, never in value.
//   * the 16-bin sweep, the SAH costs and the child boxes are computed by one thread per node with the
//     host builder's expression order (this file is compiled with -fmad=false, IEEE div).
//   * the reference's two-pointer partition (bvhBuilder.h:197-220) swaps the k-th misplaced right-side
//     element counted from the left with the k-th misplaced left-side element counted from the right;
//     that pairing is reproduced with one prefix sum per level.
//   * node ids: the reference hands out two ids per split in pre-order; pre-order among the inner nodes is
//     the order of (first triangle ascending, triangle count descending), so a final radix sort of the
//     inner nodes by that key yields the reference's numbering.
// Memory traffic per level is a few passes over 4..48 bytes per triangle (HBM bound, coalesced except
// the gathers through the order array); there is no tensor-core shape here.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <algorithm>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_radix_sort.cuh>
#include "../../include/ptb200.h"

namespace {

typedef unsigned int uint;
typedef unsigned long long u64;
#define BB_K 16                 /* bins, bvhBuilder.h:121 */
#define BB_EPS 0.001f           /* constants.h:8 */
#define BB_THREADS 256
#define BB_SCRATCH 118          /* words per open node: 6 centroid-bound keys, 16 bin counts, 16 x 6 bin-bound keys */
enum { ST_ACTIVE = 0, ST_LEAF = 1, ST_INNER = 2 };

// order-preserving float <-> uint keys (so min/max can be integer atomics)
__host__ __device__ __forceinline__ uint fkey(float f) {
    uint u;
#ifdef __CUDA_ARCH__
    u = __float_as_uint(f);
#else
    memcpy(&u, &f, 4);
#endif
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float fkeyInv(uint k) { return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k); }

// SSE min/max semantics of the reference builder (_mm_min_ps(a,b) = a<b?a:b)
__device__ __forceinline__ float smin(float a, float b) { return a < b ? a : b; }
__device__ __forceinline__ float smax(float a, float b) { return a > b ? a : b; }

struct Bounds {                 // SSEBox, types.h:205-256 (its "empty" constants are as written there)
    float lo[3], hi[3];
    __device__ static Bounds empty() { Bounds b; b.lo[0] = 999999; b.lo[1] = 99999; b.lo[2] = 99999; b.hi[0] = -999999; b.hi[1] = -99999; b.hi[2] = -99999; return b; }
    __device__ void add(const Bounds& o) { for (int a = 0; a < 3; a++) { lo[a] = smin(lo[a], o.lo[a]); hi[a] = smax(hi[a], o.hi[a]); } }
    __device__ float area() const { float lx = hi[0] - lo[0], ly = hi[1] - lo[1], lz = hi[2] - lo[2]; return 2 * lx * ly + 2 * lx * lz + 2 * ly * lz; }
};

struct Nodes {                  // open-addressed by a temporary node id (allocation order), SoA
    uint* start; uint* count; uint* child; uint* state; uint* left;
    int* axis; float* bmin; float* binFac; int* split;
    float4* lo; float4* hi;
};

__device__ __forceinline__ void emptyScratch(uint* s) {
    const uint klo[3] = {fkey(999999.f), fkey(99999.f), fkey(99999.f)}, khi[3] = {fkey(-999999.f), fkey(-99999.f), fkey(-99999.f)};
    for (int a = 0; a < 3; a++) { s[a] = klo[a]; s[3 + a] = khi[a]; }
    for (int k = 0; k < BB_K; k++) {
        s[6 + k] = 0;
        for (int a = 0; a < 3; a++) { s[22 + k * 6 + a] = klo[a]; s[22 + k * 6 + 3 + a] = khi[a]; }
    }
}

__global__ void k_init_scratch(uint* scratch, uint slots) {
    const uint i = blockIdx.x * BB_THREADS + threadIdx.x;
    if (i < slots) emptyScratch(scratch + (size_t)i * BB_SCRATCH);
}

// per triangle: centroid, bounds (bvhBuilder.h:50-63), identity order, root bounds
__global__ void __launch_bounds__(BB_THREADS) k_prepare(uint n, const float4* tv, float4* cen, float4* tbLo, float4* tbHi, uint* order, uint* posNode, uint* rootKeys) {
    __shared__ uint red[BB_THREADS / 32][6];
    const uint i = blockIdx.x * BB_THREADS + threadIdx.x;
    uint k[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    if (i < n) {
        const float4 q0 = tv[3 * (size_t)i], q1 = tv[3 * (size_t)i + 1], q2 = tv[3 * (size_t)i + 2];
        const float a[3] = {q0.x, q0.y, q0.z}, b[3] = {q0.w, q1.x, q1.y}, c[3] = {q1.z, q1.w, q2.x};
        const float third[3] = {(float)0.333333, (float)0.3333333, (float)0.3333333};     // bvhBuilder.h:57 (sic: x differs)
        float ce[3], lo[3], hi[3];
        for (int l = 0; l < 3; l++) {
            ce[l] = third[l] * (a[l] + (b[l] + c[l]));
            lo[l] = smin(a[l], smin(b[l], c[l])); hi[l] = smax(a[l], smax(b[l], c[l]));
            k[l] = fkey(lo[l]); k[3 + l] = fkey(hi[l]);
        }
        cen[i] = make_float4(ce[0], ce[1], ce[2], 0.f);
        tbLo[i] = make_float4(lo[0], lo[1], lo[2], 0.f); tbHi[i] = make_float4(hi[0], hi[1], hi[2], 0.f);
        order[i] = i; posNode[i] = 0;
    }
    for (int j = 0; j < 3; j++) { k[j] = __reduce_min_sync(0xffffffffu, k[j]); k[3 + j] = __reduce_max_sync(0xffffffffu, k[3 + j]); }
    if ((threadIdx.x & 31) == 0) for (int j = 0; j < 6; j++) red[threadIdx.x >> 5][j] = k[j];
    __syncthreads();
    if (threadIdx.x < 6) {
        uint v = red[0][threadIdx.x];
        for (int w = 1; w < BB_THREADS / 32; w++) v = threadIdx.x < 3 ? min(v, red[w][threadIdx.x]) : max(v, red[w][threadIdx.x]);
        if (threadIdx.x < 3) atomicMin(rootKeys + threadIdx.x, v); else atomicMax(rootKeys + threadIdx.x, v);
    }
}

__global__ void k_root(uint n, Nodes N, const uint* rootKeys, uint* counter) {
    N.start[0] = 0; N.count[0] = n; N.child[0] = 0xffffffffu; N.left[0] = 0;
    N.state[0] = n <= 4 ? ST_LEAF : ST_ACTIVE;                                              // bvhBuilder.h:110
    N.lo[0] = make_float4(fkeyInv(rootKeys[0]), fkeyInv(rootKeys[1]), fkeyInv(rootKeys[2]), 0.f);
    N.hi[0] = make_float4(fkeyInv(rootKeys[3]), fkeyInv(rootKeys[4]), fkeyInv(rootKeys[5]), 0.f);
    *counter = 1;
}

// Level pass 1: positions whose node split in the previous level step down to their child; centroid bounds
// of every open node (bvhBuilder.h:113-119).  One atomic per block / warp when it covers a single node.
__global__ void __launch_bounds__(BB_THREADS) k_centroid_bounds(uint n, const uint* order, uint* posNode, const float4* cen, Nodes N, uint levelBegin, uint* scratch) {
    __shared__ uint red[BB_THREADS / 32][6];
    __shared__ uint node0;
    const uint i = blockIdx.x * BB_THREADS + threadIdx.x;
    uint node = 0xffffffffu; bool active = false;
    uint k[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    if (i < n) {
        node = posNode[i];
        if (N.state[node] == ST_INNER) { const uint c = N.child[node]; node = (i < N.start[node] + N.left[node]) ? c : c + 1; posNode[i] = node; }
        active = N.state[node] == ST_ACTIVE;
        if (active) { const float4 c = cen[order[i]]; k[0] = k[3] = fkey(c.x); k[1] = k[4] = fkey(c.y); k[2] = k[5] = fkey(c.z); }
    }
    if (threadIdx.x == 0) node0 = node;
    __syncthreads();
    const bool blockUniform = __syncthreads_and(node == node0) != 0;
    const bool warpUniform = __all_sync(0xffffffffu, node == __shfl_sync(0xffffffffu, node, 0));
    if (blockUniform || warpUniform) for (int j = 0; j < 3; j++) { k[j] = __reduce_min_sync(0xffffffffu, k[j]); k[3 + j] = __reduce_max_sync(0xffffffffu, k[3 + j]); }
    if (blockUniform) {
        if ((threadIdx.x & 31) == 0) for (int j = 0; j < 6; j++) red[threadIdx.x >> 5][j] = k[j];
        __syncthreads();
        if (active && threadIdx.x < 6) {
            uint v = red[0][threadIdx.x];
            for (int w = 1; w < BB_THREADS / 32; w++) v = threadIdx.x < 3 ? min(v, red[w][threadIdx.x]) : max(v, red[w][threadIdx.x]);
            uint* s = scratch + (size_t)(node - levelBegin) * BB_SCRATCH;
            if (threadIdx.x < 3) atomicMin(s + threadIdx.x, v); else atomicMax(s + threadIdx.x, v);
        }
    } else if (active) {
        uint* s = scratch + (size_t)(node - levelBegin) * BB_SCRATCH;
        if (warpUniform) {
            const uint l = threadIdx.x & 31;
            if (l < 3) atomicMin(s + l, k[l]); else if (l < 6) atomicMax(s + l, k[l]);
        } else {
            for (int j = 0; j < 3; j++) { atomicMin(s + j, k[j]); atomicMax(s + 3 + j, k[3 + j]); }
        }
    }
}

// Level pass 2, one thread per open node: split axis and bin scale, or a leaf when the centroids are too close
// (bvhBuilder.h:121-131)
__global__ void k_choose_axis(uint levelBegin, uint levelEnd, Nodes N, const uint* scratch) {
    const uint node = levelBegin + blockIdx.x * BB_THREADS + threadIdx.x;
    if (node >= levelEnd || N.state[node] != ST_ACTIVE) return;
    const uint* s = scratch + (size_t)(node - levelBegin) * BB_SCRATCH;
    float lo[3], hi[3];
    for (int a = 0; a < 3; a++) { lo[a] = fkeyInv(s[a]); hi[a] = fkeyInv(s[3 + a]); }
    const float d0 = hi[0] - lo[0], d1 = hi[1] - lo[1], d2 = hi[2] - lo[2];
    const int axis = (d0 > d1 && d0 > d2) ? 0 : ((d1 > d0 && d1 > d2) ? 1 : 2);
    const float bmin = lo[axis], bmax = hi[axis];
    if (bmax - bmin < (float)BB_K * BB_EPS) { N.state[node] = ST_LEAF; return; }
    N.axis[node] = axis; N.bmin[node] = bmin;
    N.binFac[node] = (float)BB_K * (1 - BB_EPS) / (bmax - bmin);
}

// Level pass 3: bin every triangle of an open node (bvhBuilder.h:133-152).  A block that lies inside one node
// accumulates its 16 bins in shared memory first.
__global__ void __launch_bounds__(BB_THREADS) k_bin(uint n, const uint* order, const uint* posNode, const float4* cen, const float4* tbLo, const float4* tbHi,
                                                    Nodes N, uint levelBegin, uint* scratch, unsigned char* binOf) {
    __shared__ uint sb[BB_K * 7];          // [0,16): counts, then 16 x 6 keys
    __shared__ uint node0;
    const uint i = blockIdx.x * BB_THREADS + threadIdx.x;
    uint node = 0xffffffffu; bool active = false; uint bin = 0; uint k[6];
    if (i < n) {
        node = posNode[i];
        active = N.state[node] == ST_ACTIVE;
        if (active) {
            const uint t = order[i];
            const float4 c = cen[t];
            const int axis = N.axis[node];
            const float cv = axis == 0 ? c.x : (axis == 1 ? c.y : c.z);
            bin = (uint)(int)(signed char)((cv - N.bmin[node]) * N.binFac[node]);
            bin = min(bin, (uint)BB_K - 1u);
            binOf[i] = (unsigned char)bin;
            const float4 lo = tbLo[t], hi = tbHi[t];
            k[0] = fkey(lo.x); k[1] = fkey(lo.y); k[2] = fkey(lo.z); k[3] = fkey(hi.x); k[4] = fkey(hi.y); k[5] = fkey(hi.z);
        }
    }
    if (threadIdx.x == 0) node0 = node;
    __syncthreads();
    const bool blockUniform = __syncthreads_and(node == node0) != 0;
    if (blockUniform) {
        if (!active) return;
        if (threadIdx.x < BB_K * 7) {
            const uint j = threadIdx.x;
            sb[j] = j < BB_K ? 0u : (((j - BB_K) % 6) < 3 ? 0xffffffffu : 0u);
        }
        __syncthreads();
        atomicAdd(sb + bin, 1u);
        for (int j = 0; j < 3; j++) { atomicMin(sb + BB_K + bin * 6 + j, k[j]); atomicMax(sb + BB_K + bin * 6 + 3 + j, k[3 + j]); }
        __syncthreads();
        if (threadIdx.x < BB_K * 7) {
            const uint j = threadIdx.x;
            uint* s = scratch + (size_t)(node - levelBegin) * BB_SCRATCH + 6;
            if (j < BB_K) { if (sb[j]) atomicAdd(s + j, sb[j]); }
            else if (sb[(j - BB_K) / 6]) { if (((j - BB_K) % 6) < 3) atomicMin(s + j, sb[j]); else atomicMax(s + j, sb[j]); }
        }
    } else if (active) {
        uint* s = scratch + (size_t)(node - levelBegin) * BB_SCRATCH + 6;
        atomicAdd(s + bin, 1u);
        for (int j = 0; j < 3; j++) { atomicMin(s + BB_K + bin * 6 + j, k[j]); atomicMax(s + BB_K + bin * 6 + 3 + j, k[3 + j]); }
    }
}

// Level pass 4, one thread per open node: the SAH sweep (bvhBuilder.h:154-195) with the host builder's expression
// order; a split allocates two adjacent children and hands them their boxes (bvhBuilder.h:222-262).
__global__ void k_split(uint levelBegin, uint levelEnd, Nodes N, const uint* scratch, uint* counter) {
    const uint node = levelBegin + blockIdx.x * BB_THREADS + threadIdx.x;
    if (node >= levelEnd || N.state[node] != ST_ACTIVE) return;
    const uint* s = scratch + (size_t)(node - levelBegin) * BB_SCRATCH + 6;
    const uint count = N.count[node];
    const float4 plo = N.lo[node], phi = N.hi[node];
    const float px = phi.x - plo.x, py = phi.y - plo.y, pz = phi.z - plo.z;                  // Box::getSurfaceArea, types.h:154-161
    const float invParent = 1.0f / (2 * px * py + 2 * px * pz + 2 * py * pz);
    float costL[BB_K], costR[BB_K];
    Bounds L = Bounds::empty(), R = Bounds::empty();
    uint nl = 0, nr = 0;
    auto binBounds = [&](int k) { Bounds b; for (int a = 0; a < 3; a++) { b.lo[a] = fkeyInv(s[BB_K + k * 6 + a]); b.hi[a] = fkeyInv(s[BB_K + k * 6 + 3 + a]); } return b; };
    for (int k = 0; k < BB_K; k++) {
        costL[k] = nl * L.area() * invParent;                                                // left exclusive
        L.add(binBounds(k)); nl += s[k];
        R.add(binBounds(BB_K - k - 1)); nr += s[BB_K - k - 1];                               // right inclusive
        costR[BB_K - k - 1] = nr * R.area() * invParent;
    }
    float best = (float)count; int split = -1;
    for (int k = 0; k < BB_K; k++) { const float c = costL[k] + costR[k] + BB_EPS; if (c < best) { best = c; split = k; } }
    if (split < 0) { N.state[node] = ST_LEAF; return; }
    L = Bounds::empty(); R = Bounds::empty(); nl = 0;
    for (int k = 0; k < split; k++) { L.add(binBounds(k)); nl += s[k]; }
    for (int k = BB_K - 1; k >= split; k--) R.add(binBounds(k));
    const uint c = atomicAdd(counter, 2u);
    const uint start = N.start[node];
    N.child[node] = c; N.split[node] = split; N.left[node] = nl; N.state[node] = ST_INNER;
    N.start[c] = start; N.count[c] = nl; N.child[c] = 0xffffffffu; N.left[c] = 0; N.state[c] = nl <= 4 ? ST_LEAF : ST_ACTIVE;
    N.lo[c] = make_float4(L.lo[0], L.lo[1], L.lo[2], 0.f); N.hi[c] = make_float4(L.hi[0], L.hi[1], L.hi[2], 0.f);
    N.start[c + 1] = start + nl; N.count[c + 1] = count - nl; N.child[c + 1] = 0xffffffffu; N.left[c + 1] = 0; N.state[c + 1] = (count - nl) <= 4 ? ST_LEAF : ST_ACTIVE;
    N.lo[c + 1] = make_float4(R.lo[0], R.lo[1], R.lo[2], 0.f); N.hi[c + 1] = make_float4(R.hi[0], R.hi[1], R.hi[2], 0.f);
}

// Level pass 5: the reference's two-pointer partition (bvhBuilder.h:197-220) as a pairing.  Inside a node that
// split, the left zone is its first `left` positions.  flags = (right-side element in the left zone) << 32 |
// (left-side element in the right zone); after an exclusive sum the k-th misplaced element from the left swaps with
// the k-th misplaced element from the right.
__global__ void __launch_bounds__(BB_THREADS) k_flags(uint n, const uint* posNode, Nodes N, const unsigned char* binOf, u64* flags) {
    const uint i = blockIdx.x * BB_THREADS + threadIdx.x;
    if (i > n) return;
    u64 f = 0;
    if (i < n) {
        const uint p = posNode[i];
        if (N.state[p] == ST_INNER) {
            const bool isLeft = (int)binOf[i] < N.split[p];
            const bool leftZone = i < N.start[p] + N.left[p];
            if (leftZone && !isLeft) f = 1ull << 32;
            if (!leftZone && isLeft) f = 1ull;
        }
    }
    flags[i] = f;
}
__global__ void __launch_bounds__(BB_THREADS) k_list_right(uint n, const u64* flags, const u64* scan, uint* listL) {
    const uint i = blockIdx.x * BB_THREADS + threadIdx.x;
    if (i < n && (flags[i] & 0xffffffffull)) listL[(uint)(scan[i] & 0xffffffffull)] = i;
}
__global__ void __launch_bounds__(BB_THREADS) k_swap(uint n, const uint* posNode, Nodes N, const u64* flags, const u64* scan, const uint* listL, uint* order) {
    const uint i = blockIdx.x * BB_THREADS + threadIdx.x;
    if (i >= n || !(flags[i] >> 32)) return;
    const uint p = posNode[i];
    const uint s = N.start[p], e = s + N.count[p];
    const uint k = (uint)(scan[i] >> 32) - (uint)(scan[s] >> 32);
    const uint m = (uint)(scan[e] >> 32) - (uint)(scan[s] >> 32);
    const uint j = listL[(uint)(scan[s] & 0xffffffffull) + m - 1u - k];
    const uint a = order[i], b = order[j];
    order[i] = b; order[j] = a;
}

// Numbering: inner nodes sorted by (first triangle ascending, count descending) are in the reference's pre-order.
__global__ void k_sort_keys(uint total, Nodes N, u64* keys, uint* vals) {
    const uint x = blockIdx.x * BB_THREADS + threadIdx.x;
    if (x >= total) return;
    keys[x] = N.state[x] == ST_INNER ? (((u64)N.start[x] << 32) | (u64)(0xffffffffu - N.count[x])) : ~0ull;
    vals[x] = x;
}
__global__ void k_new_ids(uint numInner, Nodes N, const uint* sorted, uint* newId) {
    const uint r = blockIdx.x * BB_THREADS + threadIdx.x;
    if (r == 0) newId[0] = 0;
    if (r >= numInner) return;
    const uint c = N.child[sorted[r]];
    newId[c] = 1 + 2 * r; newId[c + 1] = 2 + 2 * r;                                          // c1 = used++, c2 = used++ (bvhBuilder.h:222-223)
}
__global__ void k_emit(uint total, Nodes N, const uint* newId, uint triOffset, float4* out) {
    const uint x = blockIdx.x * BB_THREADS + threadIdx.x;
    if (x >= total) return;
    const bool inner = N.state[x] == ST_INNER;
    float4 lo = N.lo[x], hi = N.hi[x];
    lo.w = __uint_as_float(inner ? newId[N.child[x]] : N.start[x] + triOffset);
    hi.w = __uint_as_float(inner ? 0u : N.count[x]);
    out[2 * (size_t)newId[x]] = lo; out[2 * (size_t)newId[x] + 1] = hi;
}

thread_local char g_buildErr[512] = "";
#define BCK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { snprintf(g_buildErr, sizeof g_buildErr, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); rc = e_ == cudaErrorMemoryAllocation ? PT_ERR_OOM : PT_ERR_CUDA; goto done; } } while (0)

struct DevMem {                 // one device allocation carved into 256-byte aligned pieces; freed on every exit path
    struct Want { void** out; size_t bytes; };
    Want want[40]; int n = 0; void* base = nullptr;
    template <typename T> void add(T** out, size_t count) { want[n++] = Want{(void**)out, (std::max<size_t>(count, 1) * sizeof(T) + 255) & ~(size_t)255}; }
    cudaError_t commit() {
        size_t total = 0;
        for (int i = 0; i < n; i++) total += want[i].bytes;
        const cudaError_t e = cudaMalloc(&base, total);
        if (e != cudaSuccess) return e;
        size_t off = 0;
        for (int i = 0; i < n; i++) { *want[i].out = (char*)base + off; off += want[i].bytes; }
        return cudaSuccess;
    }
    ~DevMem() { if (base) cudaFree(base); }
};

}  // namespace

extern "C" {

PT_API const char* pt_build_bvh_last_error(void) { return g_buildErr; }

// createBVHBinned (bvhBuilder.h:35-274) on the device.  tris: n TriangleV records (host memory); nodes: room for
// max(1, 2n-1) records; order[i] = index of the input triangle that ends up at position i (the caller applies it to
// its TriangleV / TriangleD arrays, == permuteTriangles bvhBuilder.h:17-27).
PT_API int pt_build_bvh(int device, const pt_triangle_v* tris, uint32_t n, uint32_t tri_offset, pt_bvh_node* nodes, uint32_t* n_nodes,
                        uint32_t* order, float* device_ms, uint32_t* levels_out) {
    int rc = PT_OK;
    g_buildErr[0] = 0;
    if (!nodes || !n_nodes || (n && (!tris || !order))) { snprintf(g_buildErr, sizeof g_buildErr, "pt_build_bvh: null argument"); return PT_ERR_INVALID; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0 || device < 0 || device >= ndev) {
        snprintf(g_buildErr, sizeof g_buildErr, "pt_build_bvh: no CUDA device %d (there is no CPU fallback; the host builder is pts_add_model_*)", device);
        return PT_ERR_NO_DEVICE;
    }
    DevMem mem;
    cudaStream_t st = nullptr; cudaEvent_t e0 = nullptr, e1 = nullptr;
    const uint cap = n ? 2 * n - 1 : 1;
    const uint maxSlots = n / 5 * 2 + 4;       // open nodes of one level hold > 4 triangles each and are disjoint; their children are at most twice as many
    float4 *dTv, *cen, *tbLo, *tbHi, *out; uint *ord, *posNode, *rootKeys, *counter, *scratch, *listL, *vals, *valsSorted, *newId;
    unsigned char* binOf; u64 *flags, *scan, *keys, *keysSorted; void* tmp = nullptr; size_t tmpScan = 0, tmpSort = 0, tmpBytes;
    Nodes N;
    uint levelBegin = 0, levelEnd = 1, levels = 0, total = 1;
    const uint gridN = (n + BB_THREADS - 1) / BB_THREADS, gridN1 = (n + 1 + BB_THREADS - 1) / BB_THREADS;
    const uint hostRootKeys[6] = {fkey(999999.f), fkey(99999.f), fkey(99999.f), fkey(-999999.f), fkey(-99999.f), fkey(-99999.f)};
    BCK(cudaSetDevice(device));
    BCK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    BCK(cudaEventCreate(&e0)); BCK(cudaEventCreate(&e1));
    mem.add(&dTv, 3 * (size_t)n); mem.add(&cen, n); mem.add(&tbLo, n); mem.add(&tbHi, n); mem.add(&out, 2 * (size_t)cap);
    mem.add(&ord, n); mem.add(&posNode, n); mem.add(&rootKeys, 6); mem.add(&counter, 1); mem.add(&scratch, (size_t)maxSlots * BB_SCRATCH);
    mem.add(&listL, n); mem.add(&vals, cap); mem.add(&valsSorted, cap); mem.add(&newId, cap); mem.add(&binOf, n);
    mem.add(&flags, (size_t)n + 1); mem.add(&scan, (size_t)n + 1); mem.add(&keys, cap); mem.add(&keysSorted, cap);
    mem.add(&N.start, cap); mem.add(&N.count, cap); mem.add(&N.child, cap); mem.add(&N.state, cap); mem.add(&N.left, cap);
    mem.add(&N.axis, cap); mem.add(&N.bmin, cap); mem.add(&N.binFac, cap); mem.add(&N.split, cap); mem.add(&N.lo, cap); mem.add(&N.hi, cap);
    flags = scan = keys = keysSorted = nullptr; vals = valsSorted = nullptr;      // cub only sizes its temporary storage in these calls
    BCK(cub::DeviceScan::ExclusiveSum(nullptr, tmpScan, flags, scan, (int)(n + 1), st));
    BCK(cub::DeviceRadixSort::SortPairs(nullptr, tmpSort, keys, keysSorted, vals, valsSorted, (int)cap, 0, 64, st));
    tmpBytes = std::max(tmpScan, tmpSort);
    { unsigned char* t8 = nullptr; mem.add(&t8, tmpBytes); BCK(mem.commit()); tmp = t8; }
    if (n) BCK(cudaMemcpyAsync(dTv, tris, (size_t)n * sizeof(pt_triangle_v), cudaMemcpyHostToDevice, st));
    BCK(cudaMemcpyAsync(rootKeys, hostRootKeys, sizeof hostRootKeys, cudaMemcpyHostToDevice, st));
    BCK(cudaEventRecord(e0, st));
    if (n) k_prepare<<<gridN, BB_THREADS, 0, st>>>(n, dTv, cen, tbLo, tbHi, ord, posNode, rootKeys);
    k_root<<<1, 1, 0, st>>>(n, N, rootKeys, counter);
    while (levelEnd > levelBegin && n > 4) {
        const uint slots = levelEnd - levelBegin, gridS = (slots + BB_THREADS - 1) / BB_THREADS;
        if (slots > maxSlots) { snprintf(g_buildErr, sizeof g_buildErr, "pt_build_bvh: level of %u nodes exceeds the scratch bound %u", slots, maxSlots); rc = PT_ERR_INVALID; goto done; }
        k_init_scratch<<<gridS, BB_THREADS, 0, st>>>(scratch, slots);
        k_centroid_bounds<<<gridN, BB_THREADS, 0, st>>>(n, ord, posNode, cen, N, levelBegin, scratch);
        k_choose_axis<<<gridS, BB_THREADS, 0, st>>>(levelBegin, levelEnd, N, scratch);
        k_bin<<<gridN, BB_THREADS, 0, st>>>(n, ord, posNode, cen, tbLo, tbHi, N, levelBegin, scratch, binOf);
        k_split<<<gridS, BB_THREADS, 0, st>>>(levelBegin, levelEnd, N, scratch, counter);
        k_flags<<<gridN1, BB_THREADS, 0, st>>>(n, posNode, N, binOf, flags);
        BCK(cub::DeviceScan::ExclusiveSum(tmp, tmpScan, flags, scan, (int)(n + 1), st));
        k_list_right<<<gridN, BB_THREADS, 0, st>>>(n, flags, scan, listL);
        k_swap<<<gridN, BB_THREADS, 0, st>>>(n, posNode, N, flags, scan, listL, ord);
        BCK(cudaMemcpyAsync(&total, counter, sizeof(uint), cudaMemcpyDeviceToHost, st));      // how many nodes exist now = where the next level ends
        BCK(cudaStreamSynchronize(st));
        levelBegin = levelEnd; levelEnd = total; levels++;
    }
    {
        const uint gridT = (total + BB_THREADS - 1) / BB_THREADS, numInner = (total - 1) / 2;
        k_sort_keys<<<gridT, BB_THREADS, 0, st>>>(total, N, keys, vals);
        BCK(cub::DeviceRadixSort::SortPairs(tmp, tmpSort, keys, keysSorted, vals, valsSorted, (int)total, 0, 64, st));
        k_new_ids<<<(numInner + BB_THREADS) / BB_THREADS, BB_THREADS, 0, st>>>(numInner, N, valsSorted, newId);
        k_emit<<<gridT, BB_THREADS, 0, st>>>(total, N, newId, tri_offset, out);
    }
    BCK(cudaEventRecord(e1, st));
    BCK(cudaMemcpyAsync(nodes, out, (size_t)total * sizeof(pt_bvh_node), cudaMemcpyDeviceToHost, st));
    if (n) BCK(cudaMemcpyAsync(order, ord, (size_t)n * sizeof(uint), cudaMemcpyDeviceToHost, st));
    BCK(cudaStreamSynchronize(st));
    BCK(cudaGetLastError());
    *n_nodes = total;
    if (levels_out) *levels_out = levels;
    if (device_ms) BCK(cudaEventElapsedTime(device_ms, e0, e1));
done:
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    if (st) cudaStreamDestroy(st);
    return rc;
}

}  // extern "C"
