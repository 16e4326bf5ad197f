// pt_trace.cuh -- two-level traversal: reference-exact top level (spheres, planes, TLAS, instance
// transform; reference src/kernels.h:277-387) over an 8-wide compressed bottom level (wide_bvh.h) with
// the reference's Moeller-Trumbore arithmetic (src/kernels.h:169-188) so that hit ids and t are
// bit-identical for every triangle that is tested.
#pragma once
#include "pt_device.cuh"

namespace ptb {

struct Hit { uint type, id, inst; float t; };

__device__ __forceinline__ uint extractByte(uint x, uint i) { return (x >> (i * 8u)) & 0xffu; }
__device__ __forceinline__ uint signExtendS8x4(uint x) {
    uint r; asm("prmt.b32 %0, %1, 0x0, 0x0000BA98;" : "=r"(r) : "r"(x)); return r;   // byte msb -> 0xFF / 0x00
}
__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }

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

#define PT_STACK 32

// Bottom level: 8-wide compressed BVH, one thread per ray, node-group stack (Ylitie et al. 2017).
// Closest hit keeps the candidate with the smallest t (strict <, like kernels.h:228); among exactly
// equal t the lowest reference triangle id wins, which makes the result independent of visit order.
template <bool ANY>
__device__ __forceinline__ bool traceBLAS(const DScene& sc, uint rootNode, float3 o, float3 d, float& tbest, uint& hitId) {
    const float ooeps = 8.271806125530277e-25f;   // 2^-80
    float3 idir;
    idir.x = 1.0f / (fabsf(d.x) > ooeps ? d.x : copysignf(ooeps, d.x));
    idir.y = 1.0f / (fabsf(d.y) > ooeps ? d.y : copysignf(ooeps, d.y));
    idir.z = 1.0f / (fabsf(d.z) > ooeps ? d.z : copysignf(ooeps, d.z));
    const uint octinv = (d.x < 0.0f ? 0u : 4u) | (d.y < 0.0f ? 0u : 2u) | (d.z < 0.0f ? 0u : 1u);
    const uint octinv4 = octinv * 0x01010101u;
    float tcull = tbest * 1.000001f;               // slack: box distances are not computed like triangle distances

    uint2 stack[PT_STACK];
    int sp = 0;
    uint2 ngroup = make_uint2(rootNode, 0x80000000u);
    bool found = false;

    while (true) {
        uint2 tgroup = make_uint2(0u, 0u);
        if (ngroup.y > 0x00FFFFFFu) {
            const uint bit = 31u - __clz(ngroup.y);
            ngroup.y &= ~(1u << bit);
            if (ngroup.y > 0x00FFFFFFu) { stack[sp++] = ngroup; }
            const uint slot = (bit - 24u) ^ octinv;
            const uint rel = __popc(ngroup.y & ~(0xFFFFFFFFu << slot) & 0xFFu);
            const float4* np = sc.nodes + (size_t)(ngroup.x + rel) * 5;
            const float4 n0 = ldg4(np), n1 = ldg4(np + 1), n2 = ldg4(np + 2), n3 = ldg4(np + 3), n4 = ldg4(np + 4);
            const uint ew = __float_as_uint(n0.w);
            const float sx = __uint_as_float((ew & 0xffu) << 23), sy = __uint_as_float(((ew >> 8) & 0xffu) << 23), sz = __uint_as_float(((ew >> 16) & 0xffu) << 23);
            const float aix = sx * idir.x, aiy = sy * idir.y, aiz = sz * idir.z;
            const float aox = (n0.x - o.x) * idir.x, aoy = (n0.y - o.y) * idir.y, aoz = (n0.z - o.z) * idir.z;
            ngroup.x = __float_as_uint(n1.x);
            tgroup.x = __float_as_uint(n1.y);
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
                const uint nx = d.x < 0.0f ? qhx : qlx, fx = d.x < 0.0f ? qlx : qhx;
                const uint ny = d.y < 0.0f ? qhy : qly, fy = d.y < 0.0f ? qly : qhy;
                const uint nz = d.z < 0.0f ? qhz : qlz, fz = d.z < 0.0f ? qlz : qhz;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const float tnx = __fmaf_rn((float)extractByte(nx, j), aix, aox), tny = __fmaf_rn((float)extractByte(ny, j), aiy, aoy), tnz = __fmaf_rn((float)extractByte(nz, j), aiz, aoz);
                    const float tfx = __fmaf_rn((float)extractByte(fx, j), aix, aox), tfy = __fmaf_rn((float)extractByte(fy, j), aiy, aoy), tfz = __fmaf_rn((float)extractByte(fz, j), aiz, aoz);
                    const float cmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.0f));
                    const float cmax = fminf(fminf(tfx, tfy), fminf(tfz, tcull));
                    if (cmin <= cmax) hitmask |= extractByte(child_bits4, j) << extractByte(bit_index4, j);
                }
            }
            ngroup.y = (hitmask & 0xFF000000u) | (ew >> 24);
            tgroup.y = hitmask & 0x00FFFFFFu;
        }
        while (tgroup.y != 0u) {
            const uint ti = 31u - __clz(tgroup.y);
            tgroup.y &= ~(1u << ti);
            const float4* tp = sc.wtris + (size_t)(tgroup.x + ti) * 3;
            const float4 a = ldg4(tp), b = ldg4(tp + 1), c = ldg4(tp + 2);
            float t, u, v;
            if (triIntersect(o, d, f3(a), f3(b), f3(c), t, u, v)) {
                const uint id = __float_as_uint(a.w);
                if (t < tbest || (!ANY && t == tbest && found && id < hitId)) {
                    if (ANY) return true;
                    tbest = t; hitId = id; found = true;
                    tcull = tbest * 1.000001f;
                }
            }
        }
        if (ngroup.y <= 0x00FFFFFFu) {
            if (sp == 0) break;
            ngroup = stack[--sp];
        }
    }
    return found;
}

// Top level, restating kernels.h:277-387 operation for operation (same visit order and culling as the
// reference: for unit-scale instances the result is order independent, for scaled instances the
// reference's mixed-space t comparison makes it order DEPENDENT, so the order is kept).
template <bool ANY>
__device__ __forceinline__ Hit traceTop(const DScene& sc, float3 o, float3 d, float length) {
    Hit hit; hit.type = 0; hit.inst = 0; hit.id = 0xffffffffu; hit.t = length;
    float t, tnear, tfar;
    for (uint i = 0; i < sc.num_spheres; i++)
        if (raySphereIntersect(o, d, sc.spheres[i], t) && t < hit.t) {
            if (ANY) { hit.id = i; return hit; }
            hit.id = i; hit.type = 1; hit.t = t;
        }
    for (uint i = 0; i < sc.num_planes; i++)
        if (rayPlaneIntersect(o, d, sc.planes[i], t) && t < hit.t) {
            if (ANY) { hit.id = i; return hit; }
            hit.id = i; hit.type = 2; hit.t = t;
        }
    if (sc.num_instances == 0) return hit;
    const float3 inv = rcp3(d);
    uint stack[PT_STACK];
    int sp = 0;
    bool needPop = false;
    pt_tlas_node cur = sc.tlas[0];
    if (!(slabTestRef(o, inv, &cur.vmin.x, &cur.vmax.x, tnear) && tnear < hit.t)) return hit;
    while (true) {
        if (cur.children == 0) {
            const pt_instance& inst = sc.instances[cur.leaf];
            const float3 nd = normalize(matmul(inst.invTransform, d, 0.0f));      // transformRay, kernels.h:81-86
            const float3 no = matmul(inst.invTransform, o, 1.0f);
            uint id = 0xffffffffu;
            if (traceBLAS<ANY>(sc, sc.models[inst.model_id].rootNode, no, nd, hit.t, id)) {
                if (ANY) { hit.id = 0; return hit; }
                hit.id = id; hit.type = 0; hit.inst = cur.leaf;
            }
            needPop = true;
        } else {
            const uint near_id = cur.children & 0xffffu, far_id = (cur.children >> 16) & 0xffffu;
            const pt_tlas_node nn = sc.tlas[near_id];
            const pt_tlas_node ff = sc.tlas[far_id];
            const bool bnear = slabTestRef(o, inv, &nn.vmin.x, &nn.vmax.x, tnear) && tnear < hit.t;
            const bool bfar = slabTestRef(o, inv, &ff.vmin.x, &ff.vmax.x, tfar) && tfar < hit.t;
            if (bnear && bfar) {
                const bool rev = tnear > tfar;
                if (sp < PT_STACK) stack[sp++] = rev ? near_id : far_id;
                cur = rev ? ff : nn;
            } else if (bnear) cur = nn;
            else if (bfar) cur = ff;
            else needPop = true;
        }
        if (needPop) {
            if (sp == 0) return hit;
            cur = sc.tlas[stack[--sp]];
            needPop = false;
        }
    }
}

}  // namespace ptb
