// pt_device.cuh -- device-side math + data views shared by the stage kernels (sm_100a only).
// This translation unit is compiled with -fmad=false: every expression written with the helpers
// below rounds exactly like the reference's HOST build (IEEE float, no contraction; reference
// src/cutil_math.h operators), which is what hit-id parity is defined against.  Where a
// fused multiply-add is wanted (wide-BVH box tests, which are not parity-relevant) it is written
// explicitly with __fmaf_rn.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/ptb200.h"

namespace ptb {

typedef unsigned int uint;
#define PT_EPS 0.001f                       /* constants.h:8 */
#define PT_PI 3.14159265358979323846f       /* constants.h:9 */
#define PT_MAX_CACHE_DEPTH 3                /* constants.h:14 */
#define PT_MAX_BOUNCES 64

__device__ __forceinline__ uint laneId() { uint l; asm("mov.u32 %0, %%laneid;" : "=r"(l)); return l; }

// Programmatic dependent launch (sm_90+): a kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may be
// scheduled while its predecessor on the stream still drains; pdlWait() blocks until the predecessor has completed and its
// writes are visible (a no-op for a normal launch), pdlLaunch() lets the successor's blocks take freed SM slots early.
// Every stage kernel calls pdlWait() before it touches anything an earlier kernel wrote.
__device__ __forceinline__ void pdlWait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdlLaunch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ---- float3 helpers with the reference's operator semantics (src/cutil_math.h) -------------------
__device__ __forceinline__ float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
__device__ __forceinline__ float3 f3(float s) { return make_float3(s, s, s); }
__device__ __forceinline__ float3 f3(const float4& a) { return make_float3(a.x, a.y, a.z); }
__device__ __forceinline__ float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ float3 operator*(float s, float3 a) { return f3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ float3 operator/(float3 a, float s) { float inv = 1.0f / s; return a * inv; }   // cutil_math.h:361-365
__device__ __forceinline__ float3 operator/(float3 a, float3 b) { return f3(a.x / b.x, a.y / b.y, a.z / b.z); }
__device__ __forceinline__ float3 rcp3(float3 a) { return f3(1.0f / a.x, 1.0f / a.y, 1.0f / a.z); }        // 1.0f / float3, cutil_math.h:366-369
__device__ __forceinline__ void operator+=(float3& a, float3 b) { a = a + b; }
__device__ __forceinline__ void operator*=(float3& a, float3 b) { a = a * b; }
__device__ __forceinline__ void operator*=(float3& a, float s) { a = a * s; }
__device__ __forceinline__ float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ float3 cross(float3 a, float3 b) { return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
__device__ __forceinline__ float length(float3 v) { return sqrtf(dot(v, v)); }
// The reference's normalize() is rsqrtf(dot)*v (cutil_math.h:438-441); its HOST rsqrtf is
// (float)(1.0/sqrt((double)x)) (crt/math_functions.hpp:3333).  Double div/sqrt are IEEE on the device,
// so this reproduces the host value bit for bit.
__device__ __forceinline__ float host_rsqrtf(float x) { return __double2float_rn(1.0 / sqrt((double)x)); }
__device__ __forceinline__ float3 normalize(float3 v) { return host_rsqrtf(dot(v, v)) * v; }
__device__ __forceinline__ float3 reflect(float3 i, float3 n) { return i - 2.0f * n * dot(n, i); }
__device__ __forceinline__ float fmaxcompf(float3 a) { return fmaxf(a.x, fmaxf(a.y, a.z)); }
__device__ __forceinline__ float clampf(float f, float a, float b) { return fmaxf(a, fminf(f, b)); }
__device__ __forceinline__ float luminance(float3 v) { return 0.299f * v.x + 0.587f * v.y + 0.114f * v.z; }

// ---- RNG: use_cuda.h:61-101 ----------------------------------------------------------------------
__device__ __forceinline__ uint wang_hash(uint seed) {
    seed = (seed ^ 61) ^ (seed >> 16);
    seed *= 9;
    seed = seed ^ (seed >> 4);
    seed *= 0x27d4eb2d;
    seed = seed ^ (seed >> 15);
    return seed;
}
__device__ __forceinline__ float rnd(uint& seed) {
    seed ^= (seed << 13); seed ^= (seed >> 17); seed ^= (seed << 5);
    return seed * 2.3283064365387e-10f;
}
__device__ __forceinline__ uint getSeed(uint x, uint y, uint randIdx, uint W) { return wang_hash(wang_hash(x + W * y) + randIdx); }

struct RandState { uint sampleIdx, randIdx, seed, blueFrames; float blueNoiseSample; };
__device__ __forceinline__ float rnd(RandState& rs) {          // device branch of kernels.h:20-29
    if (rs.sampleIdx < rs.blueFrames) { float k = (float)(rs.randIdx++); return fmodf(rs.blueNoiseSample + PT_PI * k, 1.0f); }
    return rnd(rs.seed);
}

// ---- pixel <-> slot mapping (tile partition) ---------------------------------------
// The image is cut into 32x32 tiles (row-major tile id t); rank r of `world` owns tiles with
// t % world == r.  A local slot is (local tile, 8x4 block, lane): one warp = one 8x4 pixel block.
struct TileMap { uint W, H, tilesX, tilesY, rank, world, nLocalTiles, nSlots, slotBase; };   // slotBase: first slot of a launch that covers a range of this rank's tiles (split frames)
__host__ __device__ __forceinline__ bool slotToPixel(const TileMap& m, uint slot, uint& x, uint& y) {
    uint lt = slot >> 10, j = slot & 1023u;
    uint gt = lt * m.world + m.rank;
    uint ty = gt / m.tilesX, tx = gt - ty * m.tilesX;
    uint blk = j >> 5, lane = j & 31u;
    x = tx * 32u + (blk & 3u) * 8u + (lane & 7u);
    y = ty * 32u + (blk >> 2) * 4u + (lane >> 3);
    return x < m.W && y < m.H;
}

// ---- bounds-checked build (make variant NAME=_chk DEFS=-DPT_BOUNDS_CHECK) ------------------------------------------------------
// compute-sanitizer is closed on the GPU pool, so out-of-range indices are looked for with checks of our own: every index the
// kernels form into the node / triangle / instance / material / queue / per-slot arrays and every traversal-stack push goes
// through PT_IDX, which in this build compares it with the array's extent (DScene::chk, set by the host per context), records the first failure in
// g_chkFail and substitutes index 0 so that the run goes on and the report can be read (pt_debug_check_report).  In the
// default build PT_IDX(i, ...) is i.
#if defined(PT_BOUNDS_CHECK)
struct CheckLimits { uint numNodes, numWtris, numTlas, numModels, numMaterials, queueCap, nSlots, stack; };
static __device__ unsigned int g_chkFail[4];        // [0] failures, [1] code of the first, [2] its index, [3] its limit
__device__ __forceinline__ uint ptCheckedIndex(uint i, uint limit, uint code) {
    if (i < limit) return i;
    if (atomicAdd(&g_chkFail[0], 1u) == 0u) { g_chkFail[1] = code; g_chkFail[2] = i; g_chkFail[3] = limit; }
    return 0u;
}
#define PT_IDX(i, limit, code) ptCheckedIndex((uint)(i), (uint)(limit), (uint)(code))
#define PT_LIMS(scene, field) ((scene).chk.field)
#else
#define PT_IDX(i, limit, code) (i)
#define PT_LIMS(scene, field) 0u
#endif
#define PT_LIM(field) PT_LIMS(sc, field)
enum { PT_CHK_STACK = 1, PT_CHK_NODE, PT_CHK_WTRI, PT_CHK_TLAS, PT_CHK_INSTANCE, PT_CHK_MODEL, PT_CHK_QUEUE, PT_CHK_SLOT, PT_CHK_TRIANGLE, PT_CHK_MATERIAL, PT_CHK_CACHE, PT_CHK_LIGHT };

// ---- device views ------------------------------------------------------------------------------------
struct DModel { uint rootNode; uint triangleStart, nrTriangles, pad; };
struct DScene {
    const float4* nodes;        // wide BVH, 5 x float4 per node
    const float4* wtris;        // wide-leaf-order triangle records, 3 x float4
    const pt_triangle_v* vertices;      // reference order (48 B)
    const pt_triangle_d* vertexData;    // reference order (80 B)
    const pt_material* materials;
    const pt_instance* instances;
    const pt_tlas_node* tlas;
    const DModel* models;
    const pt_sphere* spheres;
    const pt_plane* planes;
    const pt_triangle_light* lights;
    float4* lightTab;           // 6 x float4 per light, everything NEE needs that depends on the light only (k_prepare_lights)
    float4* rcRadiance;         // 3 x float4 per triangle: radianceCache[8], radianceTotal, pad[3]
    float* rcAdditions;         // SoA [16][num_triangles]: additionCache[8] then additionCacheCount[8]
    uint num_triangles, num_instances, num_spheres, num_planes, num_lights;
    uint minWidth;              // scheduling: least number of rays a warp claims from a short queue (1..32)
    uint tailPair;              // scheduling: k_tail may give a path two lanes (extension + shadow ray in parallel) when the queue is short
    float3 sortLo, sortScale;   // ray sorting (pt_kernels.cuh sortBin): cell = (origin - sortLo) * sortScale, clamped
    uint sortBits;              // cells per axis = 1 << sortBits
    uint coop;                  // scheduling: a trace launch whose queue holds <= coop x (4 rays per warp) runs the cooperative 8-lanes-per-ray traversal; 0: never
    uint coopTail;              // the same bound for k_tail (paths instead of rays)
    uint anyFarFirst;           // any-hit rays visit the children far end first (order only: same boolean)
    uint mixPair;               // scheduling: k_trace_pair treats its two queues as one work list (latency mode)
#if defined(PT_BOUNDS_CHECK)
    CheckLimits chk;            // extents for PT_IDX
#endif
};

struct DCamera { float3 eye, viewDir, lt, u, v; float d, focalLength, aperture; };

}  // namespace ptb
