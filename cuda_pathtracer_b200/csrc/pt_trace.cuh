// pt_trace.cuh -- two-level traversal: reference-exact top level (spheres, planes, TLAS, instance
// transform; reference src/kernels.h:277-387) over an 8-wide compressed bottom level (wide_bvh.h) with
// the reference's Moeller-Trumbore arithmetic (src/kernels.h:169-188) so that hit ids and t are
// bit-identical for every triangle that is tested.
#pragma once
#include "pt_device.cuh"
#if defined(PT_BYTE_HALF)
#include <cuda_fp16.h>
#endif

namespace ptb {

struct Hit { uint type, id, inst; float t; };

__device__ __forceinline__ uint extractByte(uint x, uint i) { return (x >> (i * 8u)) & 0xffu; }
__device__ __forceinline__ uint signExtendS8x4(uint x) {
    uint r; asm("prmt.b32 %0, %1, 0x0, 0x0000BA98;" : "=r"(r) : "r"(x)); return r;   // byte msb -> 0xFF / 0x00
}
__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }
// byte j of x as a float.  Default: I2FP.F32.U32 with a byte selector (one instruction, XU pipe).  The
// PRMT + FADD form (bits of 2^23 + b, minus 2^23; exact for 0..255) avoids the XU pipe at the price of
// one more instruction per value; measured slower on B200 (tune run r1e), kept as a build knob.
#if defined(PT_BYTE_PRMT)
__device__ __forceinline__ float byteToFloat(uint x, uint j) { return __uint_as_float(__byte_perm(x, 0x4B000000u, 0x7540u + j)) - 8388608.0f; }
#elif defined(PT_BYTE_HALF)
// Tuning variant: bytes j, j+1 of x become the half2 {1024 + b_j, 1024 + b_j+1} with one PRMT (0x64 in the exponent byte), each
// half widens to float on the FMA pipe (HADD2.F32); the caller folds the 1024 into the FMA's addend.
__device__ __forceinline__ float byteToFloat(uint x, uint j) {
    const uint h2 = __byte_perm(x, 0x64646464u, j < 2 ? 0x4140u : 0x4342u);
    const __half2 h = *reinterpret_cast<const __half2*>(&h2);
    return (j & 1u) ? __high2float(h) : __low2float(h);
}
#else
__device__ __forceinline__ float byteToFloat(uint x, uint j) { return (float)((x >> (j * 8u)) & 0xffu); }
#endif

// kernels.h:120-135
__device__ __forceinline__ bool raySphereIntersect(float3 o, float3 d, const pt_sphere& s, float& t) {
    float3 OC = o - f3(s.pos.x, s.pos.y, s.pos.z);
    float a = dot(d, d);
    if (fabsf(a) < 0.001) return false;
    float b = 2 * dot(d, OC);
    float c = dot(OC, OC) - s.radius * s.radius;
    float det = b * b - 4 * a * c;
    if (det < 0) return false;
    det = sqrtf(det);
    float tmin = (-b - det) / (2 * a);
    float tmax = (-b + det) / (2 * a);
    t = tmin;
    if (tmin < 0) t = tmax;
    return tmax > 0.0f;
}
// kernels.h:137-143
__device__ __forceinline__ bool rayPlaneIntersect(float3 o, float3 d, const pt_plane& p, float& t) {
    float3 n = f3(p.normal.x, p.normal.y, p.normal.z);
    float q = dot(normalize(d), n);
    if (fabsf(q) < PT_EPS) return false;
    t = -(dot(o, n) + p.d) / q;
    return t > 0.0f;
}
// host (SSE) branch of slabTest, kernels.h:155-165: _mm_min_ps(a,b) = a<b?a:b, _mm_max_ps(a,b) = a>b?a:b
__device__ __forceinline__ bool slabTestRef(float3 o, float3 inv, const float* vmin, const float* vmax, float& tmin) {
    float t0x = (vmin[0] - o.x) * inv.x, t0y = (vmin[1] - o.y) * inv.y, t0z = (vmin[2] - o.z) * inv.z;
    float t1x = (vmax[0] - o.x) * inv.x, t1y = (vmax[1] - o.y) * inv.y, t1z = (vmax[2] - o.z) * inv.z;
    float mnx = t0x < t1x ? t0x : t1x, mny = t0y < t1y ? t0y : t1y, mnz = t0z < t1z ? t0z : t1z;
    float mxx = t0x > t1x ? t0x : t1x, mxy = t0y > t1y ? t0y : t1y, mxz = t0z > t1z ? t0z : t1z;
    tmin = fmaxf(fmaxf(mnx, mny), mnz);
    float tmax = fminf(fminf(mxx, mxy), mxz);
    return tmax >= fmaxf(0.0f, tmin);
}
// mat4x3::mul, types.h:401-405
__device__ __forceinline__ float dot4(float3 v, float w, const pt_float4& r) { return v.x * r.x + v.y * r.y + v.z * r.z + w * r.w; }
__device__ __forceinline__ float3 matmul(const pt_mat4x3& m, float3 v, float w) { return f3(dot4(v, w, m.r1), dot4(v, w, m.r2), dot4(v, w, m.r3)); }

// Moeller-Trumbore with the reference's operation order (kernels.h:169-188); e1/e2 are v1-v0, v2-v0.
__device__ __forceinline__ bool triIntersect(float3 o, float3 d, float3 v0, float3 e1, float3 e2, float& t, float& u, float& v) {
    float3 pvec = cross(d, e2);
    float det = dot(e1, pvec);
    if (fabsf(det) < 0.0001f) return false;
    float invDet = 1.0f / det;
    float3 tvec = o - v0;
    u = dot(tvec, pvec) * invDet;
    if (u < 0.0f || u > 1.0f) return false;
    float3 qvec = cross(tvec, e1);
    v = dot(d, qvec) * invDet;
    if (v < 0.0f || u + v > 1.0f) return false;
    t = dot(e2, qvec) * invDet;
    return t > 0.0f;
}


#define PT_STACK 40            /* unified stack: TLAS entries below the node groups of the instance being traversed */
#ifndef PT_SMEM_STACK
#define PT_SMEM_STACK 0        /* tuning variant: the first PT_SMEM_STACK entries of every lane's stack live in shared memory */
#endif
#define PT_NONE 0xffffffffu
#ifndef PT_TRIS_PER_ITER
#define PT_TRIS_PER_ITER 2     /* triangle tests a lane performs per loop trip before node work resumes */
#endif
#ifndef PT_REFILL_DIV
#define PT_REFILL_DIV 1        /* fetch new rays once width/PT_REFILL_DIV lanes of the warp are idle (1: when all are) */
#endif
#ifndef PT_KEEP_NUM
#define PT_KEEP_NUM 1          /* the bottom-level loop runs while >= PT_KEEP_NUM/PT_KEEP_DEN of the lanes that */
#define PT_KEEP_DEN 32         /* entered it still have bottom-level work; the others wait for the next top-level phase */
#endif

// The traversal stack of one lane.  Default: a local-memory array (L1-resident).  With PT_SMEM_STACK = n the n entries
// nearest the bottom sit in shared memory, one column per thread (entry i of thread t at [i][t]: conflict-free 8-byte
// accesses), deeper entries in local memory.  All trace kernels run 128-thread blocks.
struct TravStack {
#if PT_SMEM_STACK > 0
    uint2* sh;
    uint2 loc[PT_STACK - PT_SMEM_STACK];
    __device__ __forceinline__ void init() {
        __shared__ uint2 shStack[PT_SMEM_STACK * 128];
        sh = shStack + threadIdx.x;
    }
    __device__ __forceinline__ void push(int& sp, const uint2 v) { if (sp < PT_SMEM_STACK) sh[sp * 128] = v; else loc[sp - PT_SMEM_STACK] = v; sp++; }
    __device__ __forceinline__ uint2 pop(int& sp) { --sp; return sp < PT_SMEM_STACK ? sh[sp * 128] : loc[sp - PT_SMEM_STACK]; }
    __device__ __forceinline__ uint2 peek(int sp) const { --sp; return sp < PT_SMEM_STACK ? sh[sp * 128] : loc[sp - PT_SMEM_STACK]; }
#else
    uint2 loc[PT_STACK];
    __device__ __forceinline__ void init() {}
    __device__ __forceinline__ void push(int& sp, const uint2 v) { loc[PT_IDX(sp, PT_STACK, PT_CHK_STACK)] = v; sp++; }
    __device__ __forceinline__ uint2 pop(int& sp) { return loc[--sp]; }
    __device__ __forceinline__ uint2 peek(int sp) const { return loc[sp - 1]; }
#endif
};

// Per-lane traversal state.  A lane is idle (ray == PT_NONE), in the top level (curT / TLAS entries
// on the stack), or inside an instance (ngroup / tgroup / node groups above spBase on the stack).
// Logical per-ray order == the reference's nested loops (kernels.h:277-387 outside, one bottom-level
// traversal per TLAS leaf inside): TLAS entries sit below the BLAS entries of the instance being
// traversed, so an instance is finished before the top level resumes.
struct LaneState {
    float3 o, d, idir;          // ray in the CURRENT space (world in the top level, model inside an instance)
    float tbest, tcull;
    uint ray;                   // queue index of the ray in flight, PT_NONE when the lane is idle
    uint hitId, hitInst, hitType;
    uint curT;                  // TLAS node to process next (PT_NONE: none)
    uint curInst;               // instance being traversed
    uint octinv, orderFlip;     // orderFlip: 7 reverses the child visiting order (far end first) for this ray, 0 keeps near-first
    uint2 ngroup, tgroup;
    int sp, spBase;
    bool inBLAS, found;
};

// Ray sources/sinks.  load() must be repeatable (the world-space ray is re-read when an instance is left).
struct QueueIO {                // the wavefront queues: SoA float4 streams
    const float4* rayO; const float4* rayD; float4* hits; const float4* lights; float4* accucolors;
    const uint* perm;           // ray sorting (k_sort_*): fetch position -> queue index, nullptr = queue order
#if defined(PT_BOUNDS_CHECK)
    struct { struct { uint queueCap, nSlots; } chk; } sc;      // what PT_LIM looks at inside this struct
#endif
    __device__ __forceinline__ void limits(const DScene& scene) {
#if defined(PT_BOUNDS_CHECK)
        sc.chk.queueCap = scene.chk.queueCap; sc.chk.nSlots = scene.chk.nSlots;
#else
        (void)scene;
#endif
    }
    static constexpr bool kStreams = true;      // rays are two contiguous float4 streams: a fetch window can be staged with bulk copies
    __device__ __forceinline__ uint index(uint i) const { return perm ? perm[i] : i; }
    __device__ __forceinline__ void load(uint i, float4& o4, float4& d4) const { i = PT_IDX(i, PT_LIM(queueCap), PT_CHK_QUEUE); o4 = rayO[i]; d4 = rayD[i]; }
    __device__ __forceinline__ void window(uint base, float4*& o, float4*& d) const { o = const_cast<float4*>(rayO + base); d = const_cast<float4*>(rayD + base); }
    template <bool ANY> __device__ __forceinline__ void store(uint i, const LaneState& L) const {
        if (ANY) {
            if (L.hitId == PT_NONE) {                                         // unoccluded: kernels.h:806-809
                const uint slot = PT_IDX(__float_as_uint(rayD[PT_IDX(i, PT_LIM(queueCap), PT_CHK_QUEUE)].w), PT_LIM(nSlots), PT_CHK_SLOT);
                const float4 l = lights[slot];
                float4 a = accucolors[slot];
                a.x += l.x; a.y += l.y; a.z += l.z;
                accucolors[slot] = a;
            }
        } else hits[PT_IDX(i, PT_LIM(queueCap), PT_CHK_QUEUE)] = make_float4(__uint_as_float(L.hitType), __uint_as_float(L.hitId), __uint_as_float(L.hitInst), L.tbest);
    }
};
struct DebugIO {                // reference-format rays in, reference-format hits out (pt_debug_trace)
    const pt_ray* rays; pt_hit* out;
    static constexpr bool kStreams = false;
    __device__ __forceinline__ uint index(uint i) const { return i; }
    __device__ __forceinline__ void load(uint i, float4& o4, float4& d4) const {
        o4 = *reinterpret_cast<const float4*>(&rays[i].origin); d4 = *reinterpret_cast<const float4*>(&rays[i].direction);
    }
    __device__ __forceinline__ void window(uint, float4*& o, float4*& d) const { o = nullptr; d = nullptr; }
    template <bool ANY> __device__ __forceinline__ void store(uint i, const LaneState& L) const {
        pt_hit p; p.primitive_type = L.hitType; p.primitive_id = L.hitId; p.instance_id = L.hitInst; p.t = L.tbest;
        if (L.hitId == PT_NONE) { p.primitive_type = 0; p.instance_id = 0; }
        else if (ANY) { p.primitive_type = 0; p.instance_id = 0; p.primitive_id = 0; }
        else if (L.hitType != 0) p.instance_id = 0;
        out[i] = p;
    }
};

__device__ __forceinline__ void setModelSpaceDir(LaneState& L) {
    const float ooeps = 8.271806125530277e-25f;   // 2^-80
    L.idir.x = 1.0f / (fabsf(L.d.x) > ooeps ? L.d.x : copysignf(ooeps, L.d.x));
    L.idir.y = 1.0f / (fabsf(L.d.y) > ooeps ? L.d.y : copysignf(ooeps, L.d.y));
    L.idir.z = 1.0f / (fabsf(L.d.z) > ooeps ? L.d.z : copysignf(ooeps, L.d.z));
    L.octinv = ((L.d.x < 0.0f ? 0u : 4u) | (L.d.y < 0.0f ? 0u : 2u) | (L.d.z < 0.0f ? 0u : 1u)) ^ L.orderFlip;
}

// Start of a ray: spheres, planes, TLAS root (kernels.h:281-318).  Returns false when the ray is already finished.
__device__ __forceinline__ bool beginRay(const DScene& sc, LaneState& L, const float4 o4, const float4 d4, const bool ANY) {
    L.o = f3(o4); L.d = f3(d4);
    L.tbest = o4.w; L.hitId = PT_NONE; L.hitType = 0; L.hitInst = 0;
    L.sp = 0; L.ngroup = make_uint2(0u, 0u); L.tgroup = make_uint2(0u, 0u); L.curT = PT_NONE; L.inBLAS = false; L.found = false;
    // An any-hit ray only asks WHETHER something lies on the segment, so its visiting order is free.  Shadow rays start on the light
    // (kernels.h:735-747): far-end-first starts the search at the shaded surface, where the occluders of a shadowed point tend to be.
    L.orderFlip = (ANY && sc.anyFarFirst) ? 7u : 0u;
    float t;
    for (uint i = 0; i < sc.num_spheres; i++)
        if (raySphereIntersect(L.o, L.d, sc.spheres[i], t) && t < L.tbest) {
            L.hitId = i; if (ANY) return false;
            L.hitType = 1; L.tbest = t;
        }
    for (uint i = 0; i < sc.num_planes; i++)
        if (rayPlaneIntersect(L.o, L.d, sc.planes[i], t) && t < L.tbest) {
            L.hitId = i; if (ANY) return false;
            L.hitType = 2; L.tbest = t;
        }
    if (sc.num_instances == 0) return false;
    L.idir = rcp3(L.d);
    const pt_tlas_node root = sc.tlas[0];
    float tnear;
    if (!(slabTestRef(L.o, L.idir, &root.vmin.x, &root.vmax.x, tnear) && tnear < L.tbest)) return false;
    L.curT = 0;
    return true;
}

// Work counters of the counting kernels (k_trace_count): node steps, triangle tests, top-level steps per lane, summed per
// warp and flushed with one atomic per counter.  NoTally compiles to nothing.
struct NoTally {
    __device__ __forceinline__ void node() {}
    __device__ __forceinline__ void tri() {}
    __device__ __forceinline__ void top() {}
    __device__ __forceinline__ void begin() {}
    __device__ __forceinline__ void end() {}
};
struct Tally {
    uint nNode = 0, nTri = 0, nTop = 0, cur = 0, longest = 0;      // longest: most node steps + triangle tests any ONE ray took (what a short launch costs)
    __device__ __forceinline__ void node() { nNode++; cur++; }
    __device__ __forceinline__ void tri() { nTri++; cur++; }
    __device__ __forceinline__ void top() { nTop++; }
    __device__ __forceinline__ void begin() { cur = 0; }
    __device__ __forceinline__ void end() { longest = max(longest, cur); }
    __device__ __forceinline__ void flush(unsigned long long* out, unsigned long long* outLongest) {
        const uint a = __reduce_add_sync(0xffffffffu, nNode), b = __reduce_add_sync(0xffffffffu, nTri), c = __reduce_add_sync(0xffffffffu, nTop);
        const uint m = __reduce_max_sync(0xffffffffu, longest);
        if (laneId() == 0) { atomicAdd(out, (unsigned long long)a); atomicAdd(out + 1, (unsigned long long)b); atomicAdd(out + 2, (unsigned long long)c); atomicMax(outLongest, (unsigned long long)m); }
    }
};

// One top-level step for the TLAS node L.curT (kernels.h:320-386): a leaf enters its instance
// (transformRay, kernels.h:81-86), an inner node tests both children in the reference's order.
__device__ __forceinline__ void tlasStep(const DScene& sc, LaneState& L, TravStack& stack) {
    const pt_tlas_node cur = sc.tlas[PT_IDX(L.curT, PT_LIM(numTlas), PT_CHK_TLAS)];
    L.curT = PT_NONE;
    if (cur.children == 0) {
        const pt_instance& inst = sc.instances[PT_IDX(cur.leaf, sc.num_instances, PT_CHK_INSTANCE)];
        const float3 nd = normalize(matmul(inst.invTransform, L.d, 0.0f));
        const float3 no = matmul(inst.invTransform, L.o, 1.0f);
        L.o = no; L.d = nd;
        setModelSpaceDir(L);
        L.tcull = L.tbest * 1.000001f;      // slack: box distances are not computed like triangle distances
        L.curInst = cur.leaf; L.inBLAS = true; L.found = false; L.spBase = L.sp;
        L.ngroup = make_uint2(sc.models[PT_IDX(inst.model_id, PT_LIM(numModels), PT_CHK_MODEL)].rootNode, 0x80000000u);
    } else {
        const uint near_id = cur.children & 0xffffu, far_id = (cur.children >> 16) & 0xffffu;
        const pt_tlas_node nn = sc.tlas[PT_IDX(near_id, PT_LIM(numTlas), PT_CHK_TLAS)];
        const pt_tlas_node ff = sc.tlas[PT_IDX(far_id, PT_LIM(numTlas), PT_CHK_TLAS)];
        float tnear, tfar;
        const bool bnear = slabTestRef(L.o, L.idir, &nn.vmin.x, &nn.vmax.x, tnear) && tnear < L.tbest;
        const bool bfar = slabTestRef(L.o, L.idir, &ff.vmin.x, &ff.vmax.x, tfar) && tfar < L.tbest;
        if (bnear && bfar) {
            const bool rev = tnear > tfar;
            stack.push(L.sp, make_uint2(rev ? near_id : far_id, 0u));
            L.curT = rev ? far_id : near_id;
        } else if (bnear) L.curT = near_id;
        else if (bfar) L.curT = far_id;
    }
}

// One bottom-level step: take the nearest pending child of the node group, intersect its 8 quantised
// child boxes (Ylitie et al. 2017), producing a new node group and a triangle group.
__device__ __forceinline__ void blasNodeStep(const DScene& sc, LaneState& L, TravStack& stack) {
    uint2 ngroup = L.ngroup;
    const uint bit = 31u - __clz(ngroup.y);
    ngroup.y &= ~(1u << bit);
    if (ngroup.y > 0x00FFFFFFu) { stack.push(L.sp, ngroup); }
    const uint slot = (bit - 24u) ^ L.octinv;
    const uint rel = __popc(ngroup.y & ~(0xFFFFFFFFu << slot) & 0xFFu);
    const float4* np = sc.nodes + (size_t)PT_IDX(ngroup.x + rel, PT_LIM(numNodes), PT_CHK_NODE) * 5;
    const float4 n0 = ldg4(np), n1 = ldg4(np + 1), n2 = ldg4(np + 2), n3 = ldg4(np + 3), n4 = ldg4(np + 4);
    const uint ew = __float_as_uint(n0.w);
    const float sx = __uint_as_float((ew & 0xffu) << 23), sy = __uint_as_float(((ew >> 8) & 0xffu) << 23), sz = __uint_as_float(((ew >> 16) & 0xffu) << 23);
    const float aix = sx * L.idir.x, aiy = sy * L.idir.y, aiz = sz * L.idir.z;
#if defined(PT_BYTE_HALF)
    const float aox = __fmaf_rn(-1024.0f, aix, (n0.x - L.o.x) * L.idir.x), aoy = __fmaf_rn(-1024.0f, aiy, (n0.y - L.o.y) * L.idir.y), aoz = __fmaf_rn(-1024.0f, aiz, (n0.z - L.o.z) * L.idir.z);
#else
    const float aox = (n0.x - L.o.x) * L.idir.x, aoy = (n0.y - L.o.y) * L.idir.y, aoz = (n0.z - L.o.z) * L.idir.z;
#endif
    const uint octinv4 = L.octinv * 0x01010101u;
    const bool negx = L.d.x < 0.0f, negy = L.d.y < 0.0f, negz = L.d.z < 0.0f;
    uint hitmask = 0;
#pragma unroll
    for (int half = 0; half < 2; half++) {
        const uint meta4 = __float_as_uint(half ? n1.w : n1.z);
        const uint is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        const uint inner_mask4 = signExtendS8x4(is_inner4 << 3);
        const uint bit_index4 = (meta4 ^ (octinv4 & inner_mask4)) & 0x1F1F1F1Fu;
        const uint child_bits4 = (meta4 >> 5) & 0x07070707u;
        const uint qlx = __float_as_uint(half ? n2.y : n2.x), qly = __float_as_uint(half ? n2.w : n2.z), qlz = __float_as_uint(half ? n3.y : n3.x);
        const uint qhx = __float_as_uint(half ? n3.w : n3.z), qhy = __float_as_uint(half ? n4.y : n4.x), qhz = __float_as_uint(half ? n4.w : n4.z);
        const uint nx = negx ? qhx : qlx, fx = negx ? qlx : qhx;
        const uint ny = negy ? qhy : qly, fy = negy ? qly : qhy;
        const uint nz = negz ? qhz : qlz, fz = negz ? qlz : qhz;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const float tnx = __fmaf_rn(byteToFloat(nx, j), aix, aox), tny = __fmaf_rn(byteToFloat(ny, j), aiy, aoy), tnz = __fmaf_rn(byteToFloat(nz, j), aiz, aoz);
            const float tfx = __fmaf_rn(byteToFloat(fx, j), aix, aox), tfy = __fmaf_rn(byteToFloat(fy, j), aiy, aoy), tfz = __fmaf_rn(byteToFloat(fz, j), aiz, aoz);
            const float cmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.0f));
            const float cmax = fminf(fminf(tfx, tfy), fminf(tfz, L.tcull));
            if (cmin <= cmax) hitmask |= extractByte(child_bits4, j) << extractByte(bit_index4, j);
        }
    }
    L.ngroup = make_uint2(__float_as_uint(n1.x), (hitmask & 0xFF000000u) | (ew >> 24));
    L.tgroup = make_uint2(__float_as_uint(n1.y), hitmask & 0x00FFFFFFu);
}

// Up to PT_TRIS_PER_ITER Moeller-Trumbore tests from the lane's pending triangle group.  Returns true when an
// any-hit ray found its occluder.
template <class TALLY>
__device__ __forceinline__ bool triStep(const DScene& sc, LaneState& L, const bool ANY, TALLY& tally) {
#pragma unroll 1
    for (int k = 0; k < PT_TRIS_PER_ITER && L.tgroup.y != 0u; k++) {
        tally.tri();
        const uint ti = 31u - __clz(L.tgroup.y);
        L.tgroup.y &= ~(1u << ti);
        const float4* tp = sc.wtris + (size_t)PT_IDX(L.tgroup.x + ti, PT_LIM(numWtris), PT_CHK_WTRI) * 3;
        const float4 a = ldg4(tp), b = ldg4(tp + 1), c = ldg4(tp + 2);
        float t, u, v;
        if (triIntersect(L.o, L.d, f3(a), f3(b), f3(c), t, u, v)) {
            const uint id = __float_as_uint(a.w);
            if (t < L.tbest || (!ANY && t == L.tbest && L.found && id < L.hitId)) {
                L.tbest = t; L.hitId = id; L.hitInst = L.curInst; L.hitType = 0; L.found = true;
                L.tcull = t * 1.000001f;
                if (ANY) return true;
            }
        }
    }
    return false;
}
// Bottom-level work of one lane for one loop trip: a node step or triangle tests, then a pop if both ran dry.
// Returns true when an any-hit ray is finished (occluded).
template <class TALLY>
__device__ __forceinline__ bool blasTrip(const DScene& sc, LaneState& L, TravStack& stack, const bool ANY, TALLY& tally) {
    if (L.tgroup.y == 0u) { blasNodeStep(sc, L, stack); tally.node(); }
    if (triStep(sc, L, ANY, tally)) { L.sp = 0; L.tgroup.y = 0u; L.ngroup.y = 0u; return true; }
    if (L.tgroup.y == 0u && L.ngroup.y <= 0x00FFFFFFu && L.sp > L.spBase) L.ngroup = stack.pop(L.sp);
    return false;
}
// Top-level work of a lane that is not inside an instance: TLAS steps until it enters one (returns false) or
// the ray is finished (returns true).  reload() re-reads the world-space ray after an instance was left.
template <class Reload, class TALLY>
__device__ __forceinline__ bool tlasTrip(const DScene& sc, LaneState& L, TravStack& stack, const Reload& reload, TALLY& tally) {
    if (L.inBLAS && L.sp > 0) {
        float4 o4, d4; reload(o4, d4);
        L.o = f3(o4); L.d = f3(d4); L.idir = rcp3(L.d);
    }
    L.inBLAS = false;
    while (true) {
        if (L.curT == PT_NONE) {
            if (L.sp == 0) return true;
            L.curT = stack.pop(L.sp).x;
        }
        tlasStep(sc, L, stack); tally.top();
        if (L.inBLAS) return false;
    }
}

// Warp-cooperative two-phase traversal with dynamic ray fetch.  All 32 lanes must call it.
//   F  idle lanes claim new queue entries (ONE atomic per warp) once enough of them are idle;
//   T  lanes that are in the top level run TLAS steps until they enter an instance or finish;
//   B  all lanes inside an instance run the bottom-level loop together; a lane that leaves its
//      instance waits, and the loop is left once fewer than PT_KEEP_NUM/PT_KEEP_DEN of its lanes remain,
// so every phase executes with most lanes participating (SIMT efficiency) on incoherent rays too.
// When the queue is shorter than the machine is wide, only ceil(n / warps) lanes per warp are used so
// the rays spread over all SMs instead of serialising inside a few warps.
// Closest hit keeps the smallest t (strict <, like kernels.h:228); among exactly equal t INSIDE one
// instance the lowest reference triangle id wins, which makes the result independent of visit order.
__device__ __forceinline__ bool hasBlasWork(const LaneState& L) { return L.ray != PT_NONE && (L.tgroup.y | (L.ngroup.y & 0xFF000000u)) != 0u; }

// ---- bottom-level phase with a warp-wide triangle queue (build knob PT_TRI_QUEUE) --------------------------------------------
// In blasTrip a lane tests its own triangles right after the node step that found them: on incoherent rays 4-5 of 32 lanes run
// the Moeller-Trumbore section, which then takes a third of the warp's issue slots.  Here the lanes only PUSH (owner lane,
// triangle) pairs into a per-warp queue in shared memory and go on with node steps; once 32 pairs are waiting the whole warp
// tests one pair per lane (the owner's model-space ray comes over with six shuffles) and merges the hits into the owner's slot
// with a 64-bit atomicMin on (t bits, triangle id).  That key order IS the selection rule of triStep: smallest t, lowest id among
// equal t inside an instance; the slot starts at (tbest, 0) when the ray has no hit in this instance yet, so a tie with a hit
// of an earlier instance loses exactly as the strict < does.  Same hits, different schedule.  The queue is drained before the
// phase ends, so no pair outlives the instance its ray was in.
#if defined(PT_TRI_QUEUE)
#ifndef PT_TQ_PUSH
#define PT_TQ_PUSH 3                                   /* pairs a lane pushes per trip; the rest of its triangle group waits */
#endif
#ifndef PT_TQ_MIN
#define PT_TQ_MIN 32                                   /* a test round starts once this many pairs are waiting */
#endif
#define PT_TQ_CAP (32 + 32 * PT_TQ_PUSH)
struct TriQueue {
    uint* q; unsigned long long* best; uint n;
    __device__ __forceinline__ void init() {
        __shared__ uint qs[4][PT_TQ_CAP];
        __shared__ unsigned long long bs[4][32];
        const uint w = (threadIdx.x >> 5) & 3u;
        q = qs[w]; best = bs[w]; n = 0;
    }
};
__device__ __forceinline__ unsigned long long hitKey(const LaneState& L) {
    return ((unsigned long long)__float_as_uint(L.tbest) << 32) | (L.found ? L.hitId : 0u);
}
// one test round: up to 32 queued pairs, one per lane.  All 32 lanes must call it.
template <class TALLY>
__device__ __forceinline__ void triQueueRound(const DScene& sc, const LaneState& L, TriQueue& Q, const uint lane, TALLY& tally) {
    __syncwarp();
    const uint take = min(Q.n, 32u);
    Q.n -= take;
    const bool valid = lane < take;
    const uint e = valid ? Q.q[Q.n + lane] : (lane << 27);
    const uint owner = e >> 27;
    const float ox = __shfl_sync(0xffffffffu, L.o.x, owner), oy = __shfl_sync(0xffffffffu, L.o.y, owner), oz = __shfl_sync(0xffffffffu, L.o.z, owner);
    const float dx = __shfl_sync(0xffffffffu, L.d.x, owner), dy = __shfl_sync(0xffffffffu, L.d.y, owner), dz = __shfl_sync(0xffffffffu, L.d.z, owner);
    if (valid) {
        tally.tri();
        const float4* tp = sc.wtris + (size_t)PT_IDX(e & 0x07FFFFFFu, PT_LIM(numWtris), PT_CHK_WTRI) * 3;
        const float4 a = ldg4(tp), b = ldg4(tp + 1), c = ldg4(tp + 2);
        float t, u, v;
        if (triIntersect(f3(ox, oy, oz), f3(dx, dy, dz), f3(a), f3(b), f3(c), t, u, v))
            atomicMin(Q.best + owner, ((unsigned long long)__float_as_uint(t) << 32) | __float_as_uint(a.w));
    }
    __syncwarp();
}
// The owner takes over what the rounds found for it.  Returns true when an any-hit ray is finished (occluded).
__device__ __forceinline__ bool triQueueRefresh(LaneState& L, const TriQueue& Q, const uint lane, const bool any) {
    if (L.ray == PT_NONE || !L.inBLAS) return false;
    const unsigned long long k = Q.best[lane];
    if (k == hitKey(L)) return false;
    L.tbest = __uint_as_float((uint)(k >> 32)); L.hitId = (uint)k; L.hitInst = L.curInst; L.hitType = 0; L.found = true;
    L.tcull = L.tbest * 1.000001f;
    if (any) { L.sp = 0; L.tgroup.y = 0u; L.ngroup.y = 0u; return true; }
    return false;
}
template <class Finish, class TALLY>
__device__ __forceinline__ void blasPhaseQueued(const DScene& sc, LaneState& L, TravStack& stack, TriQueue& Q, const uint lane, const bool any, const Finish& finish, TALLY& tally) {
    if (__ballot_sync(0xffffffffu, hasBlasWork(L)) == 0u) return;
    if (L.ray != PT_NONE && L.inBLAS) Q.best[lane] = hitKey(L);
    const uint below = (1u << lane) - 1u;
    while (true) {
        const bool work = hasBlasWork(L);
        if (__ballot_sync(0xffffffffu, work) == 0u) {
            if (Q.n == 0u) break;
            while (Q.n > 0u) triQueueRound(sc, L, Q, lane, tally);
            if (triQueueRefresh(L, Q, lane, any)) { tally.end(); finish(); }
            break;
        }
        if (work && L.tgroup.y == 0u) { blasNodeStep(sc, L, stack); tally.node(); }
        const uint c = work ? min((uint)__popc(L.tgroup.y), (uint)PT_TQ_PUSH) : 0u;
#pragma unroll
        for (uint k = 0; k < PT_TQ_PUSH; k++) {
            const uint m = __ballot_sync(0xffffffffu, c > k);
            if (c > k) {
                const uint ti = 31u - __clz(L.tgroup.y);
                L.tgroup.y &= ~(1u << ti);
                Q.q[Q.n + __popc(m & below)] = (lane << 27) | (L.tgroup.x + ti);
            }
            Q.n += __popc(m);
        }
        if (work && L.tgroup.y == 0u && L.ngroup.y <= 0x00FFFFFFu && L.sp > L.spBase) L.ngroup = stack.pop(L.sp);
        if (Q.n >= (uint)PT_TQ_MIN) {
            do triQueueRound(sc, L, Q, lane, tally); while (Q.n >= 32u);
            if (triQueueRefresh(L, Q, lane, any)) { tally.end(); finish(); }
        }
    }
}
#endif

template <bool ANY, class IO, class TALLY>
__device__ __forceinline__ void traceLoop(const DScene& sc, const uint n, uint* cursor, const IO io, TALLY& tally) {
    const uint lane = laneId();
    const uint totalWarps = (gridDim.x * blockDim.x) >> 5;
    const uint width = min(32u, max(max(1u, sc.minWidth), (n + totalWarps - 1u) / totalWarps));
    const uint refillAt = max(1u, width / PT_REFILL_DIV);
    TravStack stack; stack.init();
    LaneState L;
    L.ray = PT_NONE; L.ngroup = make_uint2(0u, 0u); L.tgroup = make_uint2(0u, 0u); L.curT = PT_NONE; L.sp = 0; L.spBase = 0; L.inBLAS = false;
    bool exhausted = n == 0;
#if defined(PT_TRI_QUEUE)
    TriQueue triQ; triQ.init();
#endif
#if defined(PT_QUEUE_BULK)
    // Build knob (A/B for the north star's "ray buffers staged through shared memory or TMA"): a warp's fetch window -- up to 32
    // consecutive queue entries, 16 B each in the two streams -- is brought into shared memory by two bulk copies
    // (cp.async.bulk, the 1-D TMA path) that complete on the warp's mbarrier, instead of one float4 load per lane and stream.
    __shared__ __align__(16) float4 stageO[4][32];
    __shared__ __align__(16) float4 stageD[4][32];
    __shared__ __align__(8) unsigned long long stageBar[4];
    const uint wib = (threadIdx.x >> 5) & 3u;
    const uint barAddr = (uint)__cvta_generic_to_shared(&stageBar[wib]);
    uint stagePhase = 0;
    if (IO::kStreams) {
        if (lane == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(barAddr) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncwarp();
    }
#endif
    while (true) {
        // ---- F: fetch ----
        const uint activeMask = __ballot_sync(0xffffffffu, L.ray != PT_NONE);
        const uint nActive = __popc(activeMask);
        if (!exhausted && nActive < width && (nActive == 0 || width - nActive >= refillAt)) {
            const uint want = width - nActive;
            uint base = 0;
            if (lane == 0) base = atomicAdd(cursor, want);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (base + want >= n) exhausted = true;
            const uint r = __popc(~activeMask & ((1u << lane) - 1u));          // rank among the idle lanes
#if defined(PT_QUEUE_BULK)
            const bool staged = IO::kStreams && io.index(0u) == 0u && io.index(1u) == 1u && base < n;      // queue order (no fetch permutation)
            if (staged) {
                const uint cnt = min(want, n - base);
                __syncwarp();                                                  // every lane is done with the previous window
                if (lane == 0) {
                    float4* srcO; float4* srcD; io.window(base, srcO, srcD);
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(barAddr), "r"(cnt * 32u) : "memory");
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 ::"r"((uint)__cvta_generic_to_shared(&stageO[wib][0])), "l"(srcO), "r"(cnt * 16u), "r"(barAddr) : "memory");
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 ::"r"((uint)__cvta_generic_to_shared(&stageD[wib][0])), "l"(srcD), "r"(cnt * 16u), "r"(barAddr) : "memory");
                }
                uint done = 0;
                do {
                    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(barAddr), "r"(stagePhase) : "memory");
                } while (!done);
                stagePhase ^= 1u;
            }
#endif
            if (L.ray == PT_NONE && r < want && base + r < n) {
                L.ray = io.index(base + r);
                tally.begin();
                float4 o4, d4;
#if defined(PT_QUEUE_BULK)
                if (staged) { o4 = stageO[wib][r]; d4 = stageD[wib][r]; } else
#endif
                io.load(L.ray, o4, d4);
                if (!beginRay(sc, L, o4, d4, ANY)) { io.template store<ANY>(L.ray, L); L.ray = PT_NONE; }
            }
        } else if (nActive == 0) break;
        // ---- T: top level ----
        if (L.ray != PT_NONE && !hasBlasWork(L)) {
            const uint ray = L.ray;
            if (tlasTrip(sc, L, stack, [&](float4& o4, float4& d4) { io.load(ray, o4, d4); }, tally)) { tally.end(); io.template store<ANY>(L.ray, L); L.ray = PT_NONE; }
        }
        // ---- B: bottom level ----
#if defined(PT_TRI_QUEUE)
        blasPhaseQueued(sc, L, stack, triQ, lane, ANY, [&]() { io.template store<ANY>(L.ray, L); L.ray = PT_NONE; }, tally);
        continue;
#endif
        const uint nB = __popc(__ballot_sync(0xffffffffu, hasBlasWork(L)));
        const uint minKeep = max(1u, (nB * PT_KEEP_NUM + PT_KEEP_DEN - 1u) / PT_KEEP_DEN);
        while (true) {
            const bool work = hasBlasWork(L);
            if (__popc(__ballot_sync(0xffffffffu, work)) < minKeep) break;
            if (work && blasTrip(sc, L, stack, ANY, tally)) { tally.end(); io.template store<ANY>(L.ray, L); L.ray = PT_NONE; }
        }
    }
}

// ---- cooperative traversal: 8 lanes per ray ------------------------------------------------------------------------------
// When a queue is much shorter than the machine is wide (deep bounces, small multi-GPU shards) most lanes of traceLoop have
// no ray, and the launch costs the latency of its slowest ray: ~50-100 dependent steps of "fetch a node, test 8 boxes, fetch
// triangles, test them".  Here a ray owns a GROUP of 8 lanes (4 rays per warp).  Ray state and stack are replicated in the 8
// lanes (identical instructions on identical values, so no broadcasts are needed); the lanes split the work of a step:
//   node step:  lane j tests child j of the 8-wide node (2 vector loads of the node header, 7 byte loads of its own
//               quantised box and meta byte), a group-wide OR (REDUX) assembles the hit mask;
//   triangle step: lane j tests the j-th pending triangle of the node's leaves (up to 8 per pass, up to 24 pending), a
//               group-wide min over (t, id) picks the winner -- the same arithmetic per triangle (triIntersect) and the same
//               selection rule as traceLoop (smallest t, lowest id among equal t inside an instance), hence the same hits.
// A step then costs one dependent load + ~1/6 of the instructions of the one-lane form.
__device__ __forceinline__ void coopNodeStep(const DScene& sc, LaneState& L, TravStack& stack, const uint sub, const uint gmask) {
    uint2 ngroup = L.ngroup;
    const uint bit = 31u - __clz(ngroup.y);
    ngroup.y &= ~(1u << bit);
    if (ngroup.y > 0x00FFFFFFu) { stack.push(L.sp, ngroup); }
    const uint slot = (bit - 24u) ^ L.octinv;
    const uint rel = __popc(ngroup.y & ~(0xFFFFFFFFu << slot) & 0xFFu);
    const float4* np = sc.nodes + (size_t)PT_IDX(ngroup.x + rel, PT_LIM(numNodes), PT_CHK_NODE) * 5;
    const float4 n0 = ldg4(np), n1 = ldg4(np + 1);
    const unsigned char* nb = reinterpret_cast<const unsigned char*>(np);
    const uint meta = __ldg(nb + 24 + sub);
    const uint qlx = __ldg(nb + 32 + sub), qly = __ldg(nb + 40 + sub), qlz = __ldg(nb + 48 + sub);
    const uint qhx = __ldg(nb + 56 + sub), qhy = __ldg(nb + 64 + sub), qhz = __ldg(nb + 72 + sub);
    const uint ew = __float_as_uint(n0.w);
    const float sx = __uint_as_float((ew & 0xffu) << 23), sy = __uint_as_float(((ew >> 8) & 0xffu) << 23), sz = __uint_as_float(((ew >> 16) & 0xffu) << 23);
    const float aix = sx * L.idir.x, aiy = sy * L.idir.y, aiz = sz * L.idir.z;
    const float aox = (n0.x - L.o.x) * L.idir.x, aoy = (n0.y - L.o.y) * L.idir.y, aoz = (n0.z - L.o.z) * L.idir.z;
    const bool negx = L.d.x < 0.0f, negy = L.d.y < 0.0f, negz = L.d.z < 0.0f;
    const float tnx = __fmaf_rn((float)(negx ? qhx : qlx), aix, aox), tny = __fmaf_rn((float)(negy ? qhy : qly), aiy, aoy), tnz = __fmaf_rn((float)(negz ? qhz : qlz), aiz, aoz);
    const float tfx = __fmaf_rn((float)(negx ? qlx : qhx), aix, aox), tfy = __fmaf_rn((float)(negy ? qly : qhy), aiy, aoy), tfz = __fmaf_rn((float)(negz ? qlz : qhz), aiz, aoz);
    const float cmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.0f));
    const float cmax = fminf(fminf(tfx, tfy), fminf(tfz, L.tcull));
    const bool inner = (meta & 0x18u) == 0x18u;                      // is_inner: meta & (meta << 1) & 0x10
    const uint bitIndex = (meta ^ (inner ? L.octinv : 0u)) & 0x1Fu;
    const uint mine = (cmin <= cmax) ? (((meta >> 5) & 7u) << bitIndex) : 0u;
    const uint hitmask = __reduce_or_sync(gmask, mine);
    L.ngroup = make_uint2(__float_as_uint(n1.x), (hitmask & 0xFF000000u) | (ew >> 24));
    L.tgroup = make_uint2(__float_as_uint(n1.y), hitmask & 0x00FFFFFFu);
}
// One pass over up to 8 pending triangles of the group's ray.  Returns true when an any-hit ray found its occluder.
template <class TALLY>
__device__ __forceinline__ bool coopTriStep(const DScene& sc, LaneState& L, const bool ANY, const uint sub, const uint gmask, TALLY& tally) {
    const uint pending = L.tgroup.y;
    const uint ti = __fns(pending, 0u, sub + 1u);                    // position of this lane's triangle (the (sub+1)-th set bit), ~0u: none
    const uint p8 = __fns(pending, 0u, 8u);
    L.tgroup.y = (p8 >= 31u) ? 0u : (pending & (0xFFFFFFFEu << p8));  // the 8 lowest set bits are taken by this pass
    unsigned long long key = ~0ull;
    if (ti != 0xffffffffu) {
        tally.tri();
        const float4* tp = sc.wtris + (size_t)PT_IDX(L.tgroup.x + ti, PT_LIM(numWtris), PT_CHK_WTRI) * 3;
        const float4 a = ldg4(tp), b = ldg4(tp + 1), c = ldg4(tp + 2);
        float t, u, v;
        if (triIntersect(L.o, L.d, f3(a), f3(b), f3(c), t, u, v)) key = ((unsigned long long)__float_as_uint(t) << 32) | __float_as_uint(a.w);   // t > 0: bit order == numeric order
    }
#pragma unroll
    for (uint off = 1; off < 8; off <<= 1) {
        const unsigned long long o = __shfl_xor_sync(gmask, key, off);
        key = o < key ? o : key;
    }
    if (key != ~0ull) {
        const float t = __uint_as_float((uint)(key >> 32));
        const uint id = (uint)key;
        if (t < L.tbest || (!ANY && t == L.tbest && L.found && id < L.hitId)) {
            L.tbest = t; L.hitId = id; L.hitInst = L.curInst; L.hitType = 0; L.found = true;
            L.tcull = t * 1.000001f;
            if (ANY) return true;
        }
    }
    return false;
}
// Build knob PT_COOP_PREFETCH: while the group tests the triangles a node step found, the node it will visit next (nearest
// pending child, else the top of the stack) is already on its way into L1 -- in the cooperative form the chain of dependent
// loads IS the launch time, and a step with triangles otherwise pays two of them (triangles, then the next node).
__device__ __forceinline__ void coopPrefetchNext(const DScene& sc, const LaneState& L, const TravStack& stack, const uint sub) {
    uint2 g = L.ngroup;
    if (g.y <= 0x00FFFFFFu) { if (L.sp <= L.spBase) return; g = stack.peek(L.sp); }
    const uint bit = 31u - __clz(g.y);
    const uint rest = g.y & ~(1u << bit);
    const uint slot = (bit - 24u) ^ L.octinv;
    const uint rel = __popc(rest & ~(0xFFFFFFFFu << slot) & 0xFFu);
    const float4* np = sc.nodes + (size_t)(g.x + rel) * 5;
    if (sub < 2u) asm volatile("prefetch.global.L1 [%0];" ::"l"(np + sub * 4u));
}
template <class TALLY>
__device__ __forceinline__ bool coopBlasTrip(const DScene& sc, LaneState& L, TravStack& stack, const bool ANY, const uint sub, const uint gmask, TALLY& tally) {
    if (L.tgroup.y == 0u) { coopNodeStep(sc, L, stack, sub, gmask); tally.node(); }
#if defined(PT_COOP_PREFETCH)
    if (L.tgroup.y != 0u) coopPrefetchNext(sc, L, stack, sub);
#endif
    if (L.tgroup.y != 0u && coopTriStep(sc, L, ANY, sub, gmask, tally)) { L.sp = 0; L.tgroup.y = 0u; L.ngroup.y = 0u; return true; }
    if (L.tgroup.y == 0u && L.ngroup.y <= 0x00FFFFFFu && L.sp > L.spBase) L.ngroup = stack.pop(L.sp);
    return false;
}

// The warp loop of the cooperative form: F idle GROUPS claim queue entries (one atomic per warp), T groups in the top level
// run TLAS steps (all 8 lanes in lockstep), B groups inside an instance run the cooperative bottom-level loop.
template <bool ANY, class IO, class TALLY>
__device__ __forceinline__ void traceLoopCoop(const DScene& sc, const uint n, uint* cursor, const IO io, TALLY& tally) {
    const uint lane = laneId(), group = lane >> 3, sub = lane & 7u, gmask = 0xFFu << (group * 8u);
    const uint totalWarps = (gridDim.x * blockDim.x) >> 5;
    const uint width = min(4u, max(1u, (n + totalWarps - 1u) / totalWarps));           // groups of a warp that take rays
    TravStack stack; stack.init();
    LaneState L;
    L.ray = PT_NONE; L.ngroup = make_uint2(0u, 0u); L.tgroup = make_uint2(0u, 0u); L.curT = PT_NONE; L.sp = 0; L.spBase = 0; L.inBLAS = false;
    bool exhausted = n == 0;
    while (true) {
        const uint activeMask = __ballot_sync(0xffffffffu, L.ray != PT_NONE) & 0x01010101u;      // one bit per group (its lane 0)
        const uint nActive = __popc(activeMask);
        if (!exhausted && nActive < width) {
            const uint want = width - nActive;
            uint base = 0;
            if (lane == 0) base = atomicAdd(cursor, want);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (base + want >= n) exhausted = true;
            const uint r = __popc(~activeMask & 0x01010101u & ((1u << (group * 8u)) - 1u));      // rank among the idle groups
            if (L.ray == PT_NONE && r < want && base + r < n) {
                L.ray = io.index(base + r);
                float4 o4, d4; io.load(L.ray, o4, d4);
                if (!beginRay(sc, L, o4, d4, ANY)) { if (sub == 0) io.template store<ANY>(L.ray, L); L.ray = PT_NONE; }
            }
        } else if (nActive == 0) break;
        if (L.ray != PT_NONE && !hasBlasWork(L)) {
            const uint ray = L.ray;
            if (tlasTrip(sc, L, stack, [&](float4& o4, float4& d4) { io.load(ray, o4, d4); }, tally)) { if (sub == 0) io.template store<ANY>(L.ray, L); L.ray = PT_NONE; }
        }
        while (true) {
            const bool work = hasBlasWork(L);
            if (__ballot_sync(0xffffffffu, work) == 0u) break;
            if (work && coopBlasTrip(sc, L, stack, ANY, sub, gmask, tally)) { if (sub == 0) io.template store<ANY>(L.ray, L); L.ray = PT_NONE; }
        }
    }
}

// ---- extension and shadow queue as ONE work list (latency mode of k_trace_pair) ------------------------------------------
// extend(b+1) and connect(b) are independent once shade(b) has run.  Draining one queue after the other costs a warp the
// longest ray of its extension batch AND of its shadow batches; when a sample runs alone on a small shard that is most of the
// launch.  Here both queues form one index space [0, nE + nS): entry i < nE is extension ray i, the others are shadow rays, so
// every warp takes a few batches of whatever comes next and the two kinds are in flight side by side.  Per lane (per group)
// `any` says which kind it carries; the per-ray work is traceLoop's (same functions), hence the same results.
template <class TALLY>
__device__ __forceinline__ void traceLoopMixed(const DScene& sc, const uint nE, const uint nS, uint* cursor, const QueueIO ioE, const QueueIO ioS, TALLY& tally) {
    const uint lane = laneId();
    const uint n = nE + nS;
    const uint totalWarps = (gridDim.x * blockDim.x) >> 5;
    const uint width = min(32u, max(max(1u, sc.minWidth), (n + totalWarps - 1u) / totalWarps));
    TravStack stack; stack.init();
    LaneState L;
    L.ray = PT_NONE; L.ngroup = make_uint2(0u, 0u); L.tgroup = make_uint2(0u, 0u); L.curT = PT_NONE; L.sp = 0; L.spBase = 0; L.inBLAS = false;
    bool any = false;
    bool exhausted = n == 0;
#if defined(PT_TRI_QUEUE)
    TriQueue triQ; triQ.init();
#endif
    auto finish = [&]() { if (any) ioS.template store<true>(L.ray, L); else ioE.template store<false>(L.ray, L); L.ray = PT_NONE; };
    while (true) {
        const uint activeMask = __ballot_sync(0xffffffffu, L.ray != PT_NONE);
        const uint nActive = __popc(activeMask);
        if (!exhausted && nActive == 0) {
            uint base = 0;
            if (lane == 0) base = atomicAdd(cursor, width);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (base + width >= n) exhausted = true;
            const uint i = base + lane;
            if (lane < width && i < n) {
                any = i >= nE;
                L.ray = any ? ioS.index(i - nE) : ioE.index(i);
                float4 o4, d4;
                if (any) ioS.load(L.ray, o4, d4); else ioE.load(L.ray, o4, d4);
                if (!beginRay(sc, L, o4, d4, any)) finish();
            }
        } else if (nActive == 0) break;
        if (L.ray != PT_NONE && !hasBlasWork(L)) {
            const uint ray = L.ray;
            if (tlasTrip(sc, L, stack, [&](float4& o4, float4& d4) { if (any) ioS.load(ray, o4, d4); else ioE.load(ray, o4, d4); }, tally)) finish();
        }
#if defined(PT_TRI_QUEUE)
        blasPhaseQueued(sc, L, stack, triQ, lane, any, finish, tally);
        continue;
#endif
        while (true) {
            const bool work = hasBlasWork(L);
            if (__ballot_sync(0xffffffffu, work) == 0u) break;
            if (work && blasTrip(sc, L, stack, any, tally)) finish();
        }
    }
}
template <class TALLY>
__device__ __forceinline__ void traceLoopCoopMixed(const DScene& sc, const uint nE, const uint nS, uint* cursor, const QueueIO ioE, const QueueIO ioS, TALLY& tally) {
    const uint lane = laneId(), group = lane >> 3, sub = lane & 7u, gmask = 0xFFu << (group * 8u);
    const uint n = nE + nS;
    const uint totalWarps = (gridDim.x * blockDim.x) >> 5;
    const uint width = min(4u, max(1u, (n + totalWarps - 1u) / totalWarps));
    TravStack stack; stack.init();
    LaneState L;
    L.ray = PT_NONE; L.ngroup = make_uint2(0u, 0u); L.tgroup = make_uint2(0u, 0u); L.curT = PT_NONE; L.sp = 0; L.spBase = 0; L.inBLAS = false;
    bool any = false;
    bool exhausted = n == 0;
    auto finish = [&]() { if (sub == 0u) { if (any) ioS.template store<true>(L.ray, L); else ioE.template store<false>(L.ray, L); } L.ray = PT_NONE; };
    while (true) {
        const uint activeMask = __ballot_sync(0xffffffffu, L.ray != PT_NONE) & 0x01010101u;
        const uint nActive = __popc(activeMask);
        if (!exhausted && nActive < width) {
            const uint want = width - nActive;
            uint base = 0;
            if (lane == 0) base = atomicAdd(cursor, want);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (base + want >= n) exhausted = true;
            const uint r = __popc(~activeMask & 0x01010101u & ((1u << (group * 8u)) - 1u));
            const uint i = base + r;
            if (L.ray == PT_NONE && r < want && i < n) {
                any = i >= nE;
                L.ray = any ? ioS.index(i - nE) : ioE.index(i);
                float4 o4, d4;
                if (any) ioS.load(L.ray, o4, d4); else ioE.load(L.ray, o4, d4);
                if (!beginRay(sc, L, o4, d4, any)) finish();
            }
        } else if (nActive == 0) break;
        if (L.ray != PT_NONE && !hasBlasWork(L)) {
            const uint ray = L.ray;
            if (tlasTrip(sc, L, stack, [&](float4& o4, float4& d4) { if (any) ioS.load(ray, o4, d4); else ioE.load(ray, o4, d4); }, tally)) finish();
        }
        while (true) {
            const bool work = hasBlasWork(L);
            if (__ballot_sync(0xffffffffu, work) == 0u) break;
            if (work && coopBlasTrip(sc, L, stack, any, sub, gmask, tally)) finish();
        }
    }
}

}  // namespace ptb
