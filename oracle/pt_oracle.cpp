// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// build, load or call anything in this file.  The product (libptb200.so) never links it.
//
// pt_oracle.cpp: a plain-C++ CPU RESTATEMENT of the wavefront path tracer of
// HugoPeters1024/cuda_pathtracer (reference paths are relative to /root/reference/):
//   RNG / seeds            src/use_cuda.h:61-101, src/kernels.h:20-29
//   camera                 src/types.h:590-600,641-658,669-676
//   intersection           src/kernels.h:120-200
//   BVH2 / TLAS traversal  src/kernels.h:202-387
//   samplers, refraction   src/kernels.h:390-483
//   __global__ stages      src/kernels.h:485-905 (clear_state, primary, extend, shade, connect,
//                          add_to_screen, init/update/propagate radiance cache)
//   frame loop             src/pathtracer.h:224-302
// Arithmetic follows the reference's HOST build: IEEE float, no FMA contraction
// (-ffp-contract=off), normalize() = (float)(1.0/sqrt((double)dot)) * v because CUDA's host
// rsqrtf is defined that way (crt/math_functions.hpp:3333), slabTest = the SSE branch
// (kernels.h:155-165).  rand(RandState&) restates the DEVICE branch (blue noise while
// sampleIdx < blue_noise_frames, kernels.h:22-27) because that is what the product runs.
//
// PARITY PINNING: every function here that has a host-callable counterpart in the reference is
// checked bit-for-bit against oracle/_ref (the reference's own headers compiled as host code) by
// tests/test_oracle_vs_ref.py, and against the committed golden vectors in tests/golden/ that were
// generated from oracle/_ref.  The __global__ bodies (shade, connect, guiding) have no host
// counterpart in the reference and no reference test: their restatement is checked by review
// against the cited lines only.  glm (instance matrices) is un-vendored: unpinned there.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <xmmintrin.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "../include/ptb200.h"

#define ORA_API extern "C" __attribute__((visibility("default")))

namespace ora {

typedef uint32_t uint;
static const float EPS = 0.001f;                      // constants.h:8
static const float PI = 3.14159265358979323846f;      // constants.h:9
static const int MAX_CACHE_DEPTH = 3;                 // constants.h:14

struct f3 { float x, y, z; };
struct f2 { float x, y; };
static inline f3 mk3(float x, float y, float z) { f3 r = {x, y, z}; return r; }
static inline f3 mk3(float s) { return mk3(s, s, s); }
static inline f3 mk3(const pt_float3& a) { return mk3(a.x, a.y, a.z); }
static inline f3 mk3(const pt_float4& a) { return mk3(a.x, a.y, a.z); }
static inline f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline f3 operator-(f3 a) { return mk3(-a.x, -a.y, -a.z); }
static inline f3 operator*(f3 a, f3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline f3 operator*(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
static inline f3 operator*(float s, f3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
static inline f3 operator/(f3 a, float s) { float inv = 1.0f / s; return mk3(a.x * inv, a.y * inv, a.z * inv); }   // cutil_math.h:361-365: multiply by reciprocal
static inline f3 operator/(f3 a, f3 b) { return mk3(a.x / b.x, a.y / b.y, a.z / b.z); }
static inline f3 operator/(float s, f3 a) { return mk3(s / a.x, s / a.y, s / a.z); }
static inline f3& operator+=(f3& a, f3 b) { a = a + b; return a; }
static inline f3& operator*=(f3& a, f3 b) { a = a * b; return a; }
static inline f3& operator*=(f3& a, float s) { a = a * s; return a; }
static inline f3& operator/=(f3& a, float s) { a = a / s; return a; }   // cutil_math.h:385-389
static inline float dot(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }            // cutil_math.h dot: left-to-right
static inline f3 cross(f3 a, f3 b) { return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
static inline float length(f3 v) { return sqrtf(dot(v, v)); }
static inline float host_rsqrtf(float a) { return (float)(1.0 / sqrt((double)a)); }         // crt/math_functions.hpp:3333
static inline f3 normalize(f3 v) { return host_rsqrtf(dot(v, v)) * v; }                     // cutil_math.h:438-441
static inline f3 reflect(f3 i, f3 n) { return i - 2.0f * n * dot(n, i); }                   // cutil_math.h:450-453
static inline float fmaxcompf(f3 a) { return fmaxf(a.x, fmaxf(a.y, a.z)); }
static inline float clampf(float f, float a, float b) { return fmaxf(a, fminf(f, b)); }      // cutil_math.h:68-71
static inline float luminance(f3 v) { return 0.299f * v.x + 0.587f * v.y + 0.114f * v.z; }   // kernels.h:51-54

// ---- RNG: use_cuda.h:61-101 -------------------------------------------------------------------
static inline uint wang_hash(uint seed) {
    seed = (seed ^ 61) ^ (seed >> 16);
    seed *= 9;
    seed = seed ^ (seed >> 4);
    seed *= 0x27d4eb2d;
    seed = seed ^ (seed >> 15);
    return seed;
}
static inline uint rand_xorshift(uint seed) {
    seed ^= (seed << 13);
    seed ^= (seed >> 17);
    seed ^= (seed << 5);
    return seed;
}
static inline float rnd(uint& seed) {
    seed = rand_xorshift(seed);
    return seed * 2.3283064365387e-10f;    // uint -> float conversion (round to nearest), then float multiply
}
static inline uint getSeed(uint x, uint y, uint randIdx, uint W) { return wang_hash(wang_hash(x + W * y) + randIdx); }

// RandState (types.h:679-687) + the DEVICE branch of rand(RandState&) (kernels.h:20-29)
struct RandState {
    uint sampleIdx, randIdx;
    float blueNoiseSample;
    uint seed;
    uint blue_noise_frames;   // reference: 1
};
static inline float rnd(RandState& rs) {
    if (rs.sampleIdx < rs.blue_noise_frames) {
        float k = (float)(rs.randIdx++);
        return fmodf(rs.blueNoiseSample + PI * k, 1.0f);
    }
    return rnd(rs.seed);
}

// ---- Ray / Hit --------------------------------------------------------------------------------
struct Ray { f3 origin, direction; float length; uint pixeli; };
static inline Ray mkRay(f3 o, f3 d, uint pixeli) { Ray r; r.origin = o; r.direction = d; r.length = 9999999; r.pixeli = pixeli; return r; }   // types.h:267
static inline Ray unpack(const pt_ray& p) {
    Ray r; r.origin = mk3(p.origin); r.direction = mk3(p.direction); r.length = p.origin.w;
    memcpy(&r.pixeli, &p.direction.w, 4); return r;
}
static inline pt_ray pack(const Ray& r) {
    pt_ray p; p.origin.x = r.origin.x; p.origin.y = r.origin.y; p.origin.z = r.origin.z; p.origin.w = r.length;
    p.direction.x = r.direction.x; p.direction.y = r.direction.y; p.direction.z = r.direction.z; memcpy(&p.direction.w, &r.pixeli, 4);
    return p;
}
enum { TRIANGLE = 0, SPHERE = 1, PLANE = 2 };
struct Hit { uint primitive_type, primitive_id, instance_id; float t; bool intersected() const { return primitive_id != 0xffffffffu; } };

// ---- Camera: types.h:590-600 (recalculate), 641-658 (getRay), 669-676 (distort) ------------------
struct Camera { f3 eye, viewDir, lt, u, v; float d, focalLength, aperture; };
static Camera camFromPod(const pt_camera& c) {
    Camera r; r.eye = mk3(c.eye); r.viewDir = mk3(c.viewDir); r.lt = mk3(c.lt); r.u = mk3(c.u); r.v = mk3(c.v);
    r.d = c.d; r.focalLength = c.focalLength; r.aperture = c.aperture; return r;
}
static void camRecalculate(Camera& c, uint W, uint H) {
    f3 center = c.eye + c.d * c.viewDir;
    c.u = normalize(cross(mk3(0, 1, 0), c.viewDir));
    c.v = normalize(cross(c.viewDir, c.u));
    float ar = (float)W / (float)H;
    c.lt = center - c.u * ar - c.v;
    c.u = 2 * ar * c.u;      // (2*ar) evaluated first: int*float, then float*float3
    c.v = 2 * c.v;
}
static f3 camDistort(const Camera& c, f3 p) {
    f3 center = c.eye + c.d * c.viewDir;
    f3 fromCenter = p - center;
    float r = length(p - center);
    float rd = r + 0.2f * r * r * r;
    return center + fromCenter * (rd / fmaxf(0.0001f, r));
}
static Ray camGetRay(const Camera& c, uint x, uint y, uint& seed, uint W, uint H) {
    float xf = ((float)x + rnd(seed)) / W;
    float yf = ((float)y + rnd(seed)) / H;
    f3 origin = camDistort(c, c.lt + xf * c.u + yf * c.v);
    f3 direction = origin - c.eye;
    const float correction = length(direction);
    direction /= correction;
    const f3 focalPoint = origin + (c.focalLength - c.d) * direction;
    const float offsetR = sqrtf(rnd(seed));
    const float offsetA = rnd(seed) * 2.0f * PI;
    const f2 focalOffset = {offsetR * sinf(offsetA), offsetR * cosf(offsetA)};
    origin += c.aperture * (focalOffset.x * c.u + focalOffset.y * c.v);
    direction = normalize(focalPoint - origin);
    if (dot(direction, c.viewDir) < 0) direction = -direction;
    return mkRay(origin - correction * direction, direction, x + y * W);
}

// ---- Scene view (host pointers; mirrors SceneBuffers, types.h:689-703) ---------------------------
struct Tex { const float* data; uint w, h, ch; };   // ch = 4 (rgba, linear filter) or 1 (point filter)
struct Scene {
    pt_scene_desc d;
    std::vector<pt_triangle_light> lights;     // pathtracer.h:155-167
    std::vector<pt_radiance_cache> rc;         // kernels.h:848-861
    Tex sky; f3 skyConst; bool hasSky;
    std::vector<float> skyPdf, skyCdf; bool skySample = false;   // CDF of pathtracer.h:85-122; sampling is this repo's opt-in (pt_set_sky_sampling)
    Tex blue; bool hasBlue;
    std::vector<Tex> textures;                 // material.texture / normal_texture = index+1 into this table
};

// ---- Intersection: kernels.h:120-200 --------------------------------------------------------------
static bool raySphereIntersect(const Ray& ray, const pt_sphere& s, float& t) {
    f3 OC = ray.origin - mk3(s.pos);
    float a = dot(ray.direction, ray.direction);
    if (fabsf(a) < 0.001) return false;          // double literal: float promoted, same decision
    float b = 2 * dot(ray.direction, OC);
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
static bool rayPlaneIntersect(const Ray& ray, const pt_plane& p, float& t) {
    float q = dot(normalize(ray.direction), mk3(p.normal));
    if (fabsf(q) < EPS) return false;
    t = -(dot(ray.origin, mk3(p.normal)) + p.d) / q;
    return t > 0.0f;
}
// host branch (SSE) of slabTest, kernels.h:155-165.  _mm_min_ps/_mm_max_ps return the SECOND operand
// when either is NaN -- kept by using the intrinsics themselves.
static inline bool slabTest(f3 o, f3 inv, const float* vmin, const float* vmax, float& tmin) {
    __m128 bmin = _mm_setr_ps(vmin[0], vmin[1], vmin[2], 0.0f);
    __m128 bmax = _mm_setr_ps(vmax[0], vmax[1], vmax[2], 0.0f);
    __m128 so = _mm_setr_ps(o.x, o.y, o.z, 0.0f);
    __m128 si = _mm_setr_ps(inv.x, inv.y, inv.z, 0.0f);
    __m128 t0 = _mm_mul_ps(_mm_sub_ps(bmin, so), si);
    __m128 t1 = _mm_mul_ps(_mm_sub_ps(bmax, so), si);
    __m128 mn = _mm_min_ps(t0, t1);
    __m128 mx = _mm_max_ps(t0, t1);
    float a[4], b[4];
    _mm_storeu_ps(a, mn); _mm_storeu_ps(b, mx);
    tmin = fmaxf(fmaxf(a[0], a[1]), a[2]);
    float tmax = fminf(fminf(b[0], b[1]), b[2]);
    return tmax >= fmaxf(0.0f, tmin);
}
static inline bool rayTriangleIntersect(const Ray& ray, const pt_triangle_v& tri, float& t, float& u, float& v) {
    f3 v0 = mk3(tri.v0);
    f3 v0v1 = mk3(tri.v1) - v0;
    f3 v0v2 = mk3(tri.v2) - v0;
    f3 pvec = cross(ray.direction, v0v2);
    float det = dot(v0v1, pvec);
    if (fabsf(det) < 0.0001f) return false;
    float invDet = 1.0f / det;
    f3 tvec = ray.origin - v0;
    u = dot(tvec, pvec) * invDet;
    if (u < 0.0f || u > 1.0f) return false;
    f3 qvec = cross(tvec, v0v1);
    v = dot(ray.direction, qvec) * invDet;
    if (v < 0.0f || u + v > 1.0f) return false;
    t = dot(v0v2, qvec) * invDet;
    return t > 0.0f;
}
static inline bool boxtest(const float* vmin, const float* vmax, f3 o, f3 inv, float& tmin, const Hit& h) {
    return slabTest(o, inv, vmin, vmax, tmin) && tmin < h.t;
}

// mat4x3::mul, types.h:401-405: dot(float4(target, fourth), row) left-to-right
static inline float dot4(float x, float y, float z, float w, const pt_float4& r) { return x * r.x + y * r.y + z * r.z + w * r.w; }
static inline f3 matmul(const pt_mat4x3& m, f3 v, float w) {
    return mk3(dot4(v.x, v.y, v.z, w, m.r1), dot4(v.x, v.y, v.z, w, m.r2), dot4(v.x, v.y, v.z, w, m.r3));
}
static inline Ray transformRay(const Ray& ray, const pt_mat4x3& m) {     // kernels.h:81-86 (length reset to 9999999 by the ctor)
    f3 nd = normalize(matmul(m, ray.direction, 0.0f));
    f3 no = matmul(m, ray.origin, 1.0f);
    return mkRay(no, nd, ray.pixeli);
}

struct Counters { uint64_t inner, tris, tlas, maxStack; };   // maxStack: deepest BVH2 stack the ray needed (the reference's array holds 18 entries, kernels.h:207)   // algorithmic-bytes counters; not in the reference

// kernels.h:202-275
template <bool ANY>
static bool traverseBVHStack(const Scene& sc, const Ray& ray, Hit& hit, const pt_bvh_node* bvh, Counters* cnt) {
    float t, u, v;
    uint stack[64];      // reference: 18 entries, unchecked (kernels.h:207); larger here so a deeper tree cannot corrupt the oracle
    uint* sp = stack;
    float tnear, tfar;
    const f3 inv = 1.0f / ray.direction;
    pt_bvh_node cur = bvh[0];
    bool needPop = false, intersected = false;
    while (true) {
        uint tcount; memcpy(&tcount, &cur.vmax.w, 4);
        uint first; memcpy(&first, &cur.vmin.w, 4);
        if (tcount > 0) {
            const uint end = first + tcount;
            for (uint i = first; i < end; i++) {
                if (cnt) cnt->tris++;
                if (rayTriangleIntersect(ray, sc.d.vertices[i], t, u, v) && t < hit.t) {
                    if (ANY) return true;
                    hit.primitive_id = i; hit.t = t; intersected = true;
                }
            }
            needPop = true;
        } else {
            if (cnt) cnt->inner++;
            const uint near_id = first;
            const pt_bvh_node& nn = bvh[near_id];
            const pt_bvh_node& ff = bvh[near_id + 1];
            const bool bnear = boxtest(&nn.vmin.x, &nn.vmax.x, ray.origin, inv, tnear, hit);
            const bool bfar = boxtest(&ff.vmin.x, &ff.vmax.x, ray.origin, inv, tfar, hit);
            if (bnear & bfar) {
                const bool rev = tnear > tfar;
                cur = rev ? ff : nn;
                *sp++ = near_id + !rev;
                if (cnt && (uint64_t)(sp - stack) > cnt->maxStack) cnt->maxStack = (uint64_t)(sp - stack);
            } else if (bnear) cur = nn;
            else if (bfar) cur = ff;
            else needPop = true;
        }
        if (needPop) {
            if (stack != sp) cur = bvh[*--sp];
            else return intersected;
            needPop = false;
        }
    }
}

// kernels.h:277-387
template <bool ANY>
static Hit traverseTopLevel(const Scene& sc, const Ray& ray, Counters* cnt) {
    Hit hit; hit.primitive_type = TRIANGLE; hit.instance_id = 0;
    hit.primitive_id = 0xffffffffu;
    hit.t = ray.length;
    float t, tnear, tfar;
    for (uint i = 0; i < sc.d.num_spheres; i++)
        if (raySphereIntersect(ray, sc.d.spheres[i], t) && t < hit.t) {
            if (ANY) { hit.primitive_id = i; return hit; }
            hit.primitive_id = i; hit.primitive_type = SPHERE; hit.t = t;
        }
    for (uint i = 0; i < sc.d.num_planes; i++)
        if (rayPlaneIntersect(ray, sc.d.planes[i], t) && t < hit.t) {
            if (ANY) { hit.primitive_id = i; return hit; }
            hit.primitive_id = i; hit.primitive_type = PLANE; hit.t = t;
        }
    if (sc.d.num_instances == 0) return hit;
    const f3 inv = 1.0f / ray.direction;
    uint stack[64];
    uint* sp = stack;
    bool needPop = false;
    pt_tlas_node cur = sc.d.topBvh[0];
    if (cnt) cnt->tlas++;
    if (!boxtest(&cur.vmin.x, &cur.vmax.x, ray.origin, inv, tnear, hit)) return hit;
    while (true) {
        if (cur.children == 0) {
            const pt_instance& inst = sc.d.instances[cur.leaf];
            const Ray tr = transformRay(ray, inst.invTransform);
            if (traverseBVHStack<ANY>(sc, tr, hit, sc.d.models[inst.model_id].bvh, cnt)) {
                if (ANY) { hit.primitive_id = 0; return hit; }
                hit.primitive_type = TRIANGLE; hit.instance_id = cur.leaf;
            }
            needPop = true;
        } else {
            const uint near_id = cur.children & 0xffffu, far_id = (cur.children >> 16) & 0xffffu;
            const pt_tlas_node nn = sc.d.topBvh[near_id];
            const pt_tlas_node ff = sc.d.topBvh[far_id];
            if (cnt) cnt->tlas += 2;
            const bool bnear = boxtest(&nn.vmin.x, &nn.vmax.x, ray.origin, inv, tnear, hit);
            const bool bfar = boxtest(&ff.vmin.x, &ff.vmax.x, ray.origin, inv, tfar, hit);
            if (bnear && bfar) {
                bool rev = tnear > tfar;
                *sp++ = rev ? near_id : far_id;
                cur = rev ? ff : nn;
            } else if (bnear) cur = nn;
            else if (bfar) cur = ff;
            else needPop = true;
        }
        if (needPop) {
            if (stack != sp) cur = sc.d.topBvh[*--sp];
            else return hit;
            needPop = false;
        }
    }
}


// ---- Samplers / BSDF helpers: kernels.h:31-36, 390-483 ---------------------------------------------
static inline f2 normalToUv(f3 n) {
    f2 r; r.x = atan2f(n.x, n.z) / (2.0f * PI); r.y = -acosf(n.y) / PI; return r;
}
static f3 SampleHemisphereCosine(f3 normal, float r0, float r1) {
    float r = sqrtf(r0);
    float theta = 2 * PI * r1;
    float x = r * cosf(theta);
    float y = r * sinf(theta);
    f3 sample = mk3(x, y, sqrtf(1 - r0));
    const f3 w = normal;
    f3 u = normalize(cross((fabsf(w.x) > .1f ? mk3(0, 1, 0) : mk3(1, 0, 0)), w));
    f3 v = normalize(cross(w, u));
    return normalize(mk3(dot(sample, mk3(u.x, v.x, w.x)), dot(sample, mk3(u.y, v.y, w.y)), dot(sample, mk3(u.z, v.z, w.z))));
}
static f3 SampleHemisphereCached(f3 normal, RandState& rs, const pt_radiance_cache& rc, int& sampleBucket, float& invprob) {
    const float sample = rnd(rs) * rc.radianceTotal;
    float acc = EPS;
    sampleBucket = -1;
    do { acc += rc.radianceCache[++sampleBucket]; } while (acc < sample && sampleBucket < 7);   // `&& < 7`: guard the oracle against the reference's unchecked overrun (kernels.h:413-417)
    const float r0Min = (sampleBucket < 4) ? 0 : 0.5f;
    const float r0Max = (sampleBucket < 4) ? 0.5f : 1.0f;
    const uint r1i = sampleBucket % 4;
    const float r1Min = r1i * (1.0f / 4.0f);
    const float r1Max = (r1i + 1.0f) * (1.0f / 4.0f);
    const float r0R = rnd(rs);
    const float r1R = rnd(rs);
    const float r0 = r0Min * r0R + r0Max * (1.0f - r0R);
    const float r1 = r1Min * r1R + r1Max * (1.0f - r1R);
    invprob = rc.radianceTotal / (rc.radianceCache[sampleBucket] * 8.0f);
    return SampleHemisphereCosine(normal, r0, r1);
}
static Ray getReflectRay(const Ray& ray, f3 normal, f3 pos) {
    const f3 nd = reflect(ray.direction, normal);
    return mkRay(pos + EPS * nd, nd, ray.pixeli);
}
static Ray getRefractRay(const Ray& ray, f3 normal, f3 pos, float ior, bool inside, float& reflected) {
    const float n1 = inside ? ior : 1.0f;
    const float n2 = inside ? 1.0f : ior;
    const float eta = n1 / n2;
    const float costi = dot(normal, -ray.direction);
    const float k = 1 - (eta * eta) * (1 - costi * costi);
    if (k < 0) { reflected = 1; return mkRay(mk3(0), mk3(0), 0); }
    const f3 refractDir = normalize(eta * ray.direction + normal * (eta * costi - sqrtf(k)));
    const float sinti = sqrtf(fmaxf(0.0f, 1.0f - costi - costi));      // sic: kernels.h:476
    const float costt = sqrtf(1.0f - eta * eta * sinti * sinti);
    const float spol = (n1 * costi - n2 * costt) / (n1 * costi + n2 * costt);
    const float ppol = (n1 * costt - n2 * costi) / (n1 * costt + n2 * costi);
    reflected = 0.5f * (spol * spol + ppol * ppol);
    return mkRay(pos + EPS * refractDir, refractDir, ray.pixeli);
}

// ---- Texture fetches (CPU model of the CUDA texture unit the reference samples through) -----------
// Normalised coordinates, wrap addressing (use_cuda.h:147-151).  Linear filter: CUDA programming
// guide "Texture Fetching": xB = N*frac(x) - 0.5, i = floor(xB), alpha = frac(xB) held in 9-bit fixed
// point with 8 fractional bits.  Point filter (blue noise, use_cuda.h:187): T[floor(N*frac(x))].
static inline float fracf(float x) { return x - floorf(x); }
static inline void tex2D4(const Tex& t, float u, float v, float out[4]) {
    float xb = (float)t.w * fracf(u) - 0.5f, yb = (float)t.h * fracf(v) - 0.5f;
    float fi = floorf(xb), fj = floorf(yb);
    float a = floorf((xb - fi) * 256.0f + 0.5f) / 256.0f, b = floorf((yb - fj) * 256.0f + 0.5f) / 256.0f;
    int i0 = (int)fi, j0 = (int)fj;
    int i1 = i0 + 1, j1 = j0 + 1;
    int W = (int)t.w, H = (int)t.h;
    i0 = ((i0 % W) + W) % W; i1 = ((i1 % W) + W) % W; j0 = ((j0 % H) + H) % H; j1 = ((j1 % H) + H) % H;
    const float* p00 = t.data + 4 * ((size_t)j0 * W + i0); const float* p10 = t.data + 4 * ((size_t)j0 * W + i1);
    const float* p01 = t.data + 4 * ((size_t)j1 * W + i0); const float* p11 = t.data + 4 * ((size_t)j1 * W + i1);
    for (int c = 0; c < 4; c++)
        out[c] = (1 - a) * (1 - b) * p00[c] + a * (1 - b) * p10[c] + (1 - a) * b * p01[c] + a * b * p11[c];
}
static inline float tex2D1point(const Tex& t, float u, float v) {
    int i = (int)floorf((float)t.w * fracf(u)), j = (int)floorf((float)t.h * fracf(v));
    if (i >= (int)t.w) i = t.w - 1;
    if (j >= (int)t.h) j = t.h - 1;
    return t.data[(size_t)j * t.w + i];
}


// ---- The oracle's renderer state (what Pathtracer owns, pathtracer.h:46-71,192-209) -----------------
struct F4 { float x, y, z, w; };
struct TraceState { f3 mask, accucolor, light; bool fromSpecular, albedoSet; };   // types.h:461-468
struct Oracle {
    uint W, H, NP;
    pt_config cfg;
    Scene sc;
    bool HNEE, HCACHE, DNEE, DCACHE;          // globals.h:8-9,19-20 (device copies synced on shouldClear, pathtracer.h:242-243)
    uint sampleIdx, randIdx;
    std::vector<F4> masks, accucolors, lights;  // TraceStateSOA, types.h:470-499
    std::vector<pt_ray> rayQ, rayQNew, shadowQ; // globals.h:11-15
    std::vector<pt_hit> hits;
    std::vector<pt_sample_cache> sampleCache;   // NR_PIXELS * MAX_CACHE_DEPTH, pathtracer.h:209
    std::vector<F4> image, albedo;              // stand-ins for the two GL RGBA32F textures (main.cpp:214-260)
    Counters cnt, cntConn;      // extend / connect traversal counters (algorithmic bytes)
    pt_stats stats;
    int threads;
};

static inline bool wbit(float w) { unsigned char b; memcpy(&b, &w, 1); return b != 0; }   // reinterpret_cast<bool&>(float), types.h:488-489
static inline float bitw(bool b) { float w = 0; unsigned char c = b ? 1 : 0; memcpy(&w, &c, 1); return w; }   // types.h:495-496
static TraceState getState(const Oracle& o, uint i) {
    TraceState s; const F4 &m = o.masks[i], &a = o.accucolors[i], &l = o.lights[i];
    s.mask = mk3(m.x, m.y, m.z); s.accucolor = mk3(a.x, a.y, a.z); s.light = mk3(l.x, l.y, l.z);
    s.fromSpecular = wbit(m.w); s.albedoSet = wbit(a.w); return s;
}
static void setState(Oracle& o, uint i, const TraceState& s) {
    F4 m = {s.mask.x, s.mask.y, s.mask.z, bitw(s.fromSpecular)}; o.masks[i] = m;
    F4 a = {s.accucolor.x, s.accucolor.y, s.accucolor.z, bitw(s.albedoSet)}; o.accucolors[i] = a;
    F4 l = {s.light.x, s.light.y, s.light.z, 0}; o.lights[i] = l;
}
static void updateAlbedo(Oracle& o, f3 nv, uint x, uint y) {   // kernels.h:56-62
    F4& p = o.albedo[x + y * o.W];
    F4 r = {fmaxf(0.0f, p.x) + nv.x, fmaxf(0.0f, p.y) + nv.y, fmaxf(0.0f, p.z) + nv.z, p.w + 1.0f};
    p = r;
}
static inline f3 skyLookup(const Oracle& o, f3 dir) {           // kernels.h:527-528
    if (!o.sc.hasSky) return o.sc.skyConst;
    f2 uv = normalToUv(dir); float c[4]; tex2D4(o.sc.sky, uv.x, uv.y, c); return mk3(c[0], c[1], c[2]);
}

// kernels.h:65-79 completed (see binarySearchCdf in csrc/pt_shade.cuh): first texel whose cumulative value reaches target
static inline uint binarySearchCdf(const float* values, uint nrValues, float target) {
    uint L = 0, R = nrValues - 1;
    while (L + 1 < R) { const uint m = (L + R) / 2; if (values[m] >= target) R = m; else L = m; }
    return values[L] >= target ? L : R;
}
static inline f3 skyUvToDir(float u, float v, float& sinPolar) {     // inverse of normalToUv under wrap addressing
    const float theta = u * 2.0f * PI;
    const float a = PI * (1.0f - v);
    sinPolar = sinf(a);
    return mk3(sinf(theta) * sinPolar, cosf(a), cosf(theta) * sinPolar);
}

struct ShadeOut { bool pushExt, pushShadow; pt_ray ext, shadow; };

// kernel_shade, kernels.h:513-796, for queue entry i.  Queue pushes are returned instead of atomically appended.
static void shadeOne(Oracle& o, uint i, int bounce, uint randIdx, ShadeOut& out) {
    const Scene& sc = o.sc;
    out.pushExt = out.pushShadow = false;
    const Ray ray = unpack(o.rayQ[i]);
    TraceState state = getState(o, ray.pixeli);
    pt_sample_cache cache; memset(&cache, 0, sizeof(cache));
    cache.sample_type = 1;                                   // SAMPLE_TERMINATE
    const uint cache_id = ray.pixeli + (bounce * o.NP);
    const uint x = ray.pixeli % o.W, y = ray.pixeli / o.W;
    Hit hit; { const pt_hit& h = o.hits[i]; hit.primitive_type = h.primitive_type; hit.primitive_id = h.primitive_id; hit.instance_id = h.instance_id; hit.t = h.t; }
    if (!hit.intersected()) {                                // kernels.h:526-538
        f3 sk = skyLookup(o, ray.direction);
        if (!state.albedoSet) { updateAlbedo(o, state.mask * sk, x, y); state.albedoSet = true; }
        if (!(o.sc.skySample && o.sc.hasSky) || state.fromSpecular) state.accucolor += state.mask * sk;
        setState(o, ray.pixeli, state);
        if (bounce < MAX_CACHE_DEPTH) o.sampleCache[cache_id] = cache;
        return;
    }
    RandState rs; rs.sampleIdx = o.sampleIdx; rs.randIdx = randIdx; rs.blue_noise_frames = o.cfg.blue_noise_frames;
    rs.seed = getSeed(x, y, rs.randIdx, o.W);                // kernels.h:540
    rs.blueNoiseSample = 0.0f;
    if (sc.hasBlue) {                                        // kernels.h:541-542
        float bu = (float)x * (1.0f / (float)sc.blue.w), bv = (float)y * (1.0f / (float)sc.blue.h);
        rs.blueNoiseSample = tex2D1point(sc.blue, bu, bv);
    }
    const pt_instance* instance = sc.d.instances + (hit.primitive_type == TRIANGLE ? hit.instance_id : 0);
    const f3 intersectionPos = ray.origin + hit.t * ray.direction;
    uint material_id;                                        // getColliderMaterialID, kernels.h:88-99
    f3 originalNormal;                                       // getColliderNormal, kernels.h:101-118
    if (hit.primitive_type == TRIANGLE) {
        material_id = instance->material_id == 0xffffffffu ? sc.d.vertexData[hit.primitive_id].material : instance->material_id;
        originalNormal = mk3(sc.d.vertexData[hit.primitive_id].normal);
        originalNormal = normalize(matmul(instance->transform, originalNormal, 0.0f));     // kernels.h:553-556
    } else if (hit.primitive_type == SPHERE) {
        material_id = sc.d.spheres[hit.primitive_id].material;
        originalNormal = normalize(intersectionPos - mk3(sc.d.spheres[hit.primitive_id].pos));
    } else {
        material_id = sc.d.planes[hit.primitive_id].material;
        originalNormal = mk3(sc.d.planes[hit.primitive_id].normal);
    }
    pt_material material = sc.d.materials[material_id];
    bool inside = dot(ray.direction, originalNormal) > 0;
    f3 surfaceNormal = inside ? -originalNormal : originalNormal;
    f3 colliderNormal = surfaceNormal;
    const bool NEE = o.DNEE && o.sc.lights.size() > 0;      // device _NEE, globals.h:23
    const uint numLights = (uint)o.sc.lights.size();

    if (fmaxcompf(mk3(material.emission)) > EPS) {           // kernels.h:563-576
        if (!NEE || state.fromSpecular) {
            state.accucolor += state.mask * mk3(material.emission);
            if (!state.albedoSet) { updateAlbedo(o, state.mask * mk3(material.emission), x, y); state.albedoSet = true; }
            setState(o, ray.pixeli, state);
        }
        return;
    }
    f3 diffuse = mk3(material.diffuse_color);
    if (hit.primitive_type == PLANE) {                       // kernels.h:578-582
        uint px = (uint)(fabsf(intersectionPos.x / 4 + 1000));
        uint py = (uint)(fabsf(intersectionPos.z / 4 + 1000));
        diffuse = (px + py) % 2 == 0 ? mk3(1) : mk3(0.2);
    }
    if (material.hasTexture || material.hasNormalMap) {      // kernels.h:586-619
        const pt_triangle_d& td = sc.d.vertexData[hit.primitive_id];
        const pt_triangle_v& tv = sc.d.vertices[hit.primitive_id];
        float t, u = 0, v = 0;
        Ray tr = transformRay(ray, instance->invTransform);
        rayTriangleIntersect(tr, tv, t, u, v);
        f2 uv; uv.x = td.uv0.x * (1 - u - v) + td.uv1.x * u + td.uv2.x * v; uv.y = td.uv0.y * (1 - u - v) + td.uv1.y * u + td.uv2.y * v;
        if (material.hasTexture) {
            float c[4]; tex2D4(sc.textures[material.texture - 1], uv.x, uv.y, c);
            diffuse = diffuse * mk3(c[0], c[1], c[2]);
        }
        if (material.hasNormalMap) {
            float c[4]; tex2D4(sc.textures[material.normal_texture - 1], uv.x, uv.y, c);
            f3 tnT = (mk3(c[0], c[1], c[2]) * 2.0f) - mk3(1.0f);
            f3 tn = mk3(dot(tnT, mk3(td.tangent.x, td.bitangent.x, td.normal.x)), dot(tnT, mk3(td.tangent.y, td.bitangent.y, td.normal.y)),
                        dot(tnT, mk3(td.tangent.z, td.bitangent.z, td.normal.z)));
            tn = normalize(matmul(instance->transform, tn, 0.0f));
            colliderNormal = dot(tn, colliderNormal) < 0.0f ? -tn : tn;
        }
    }
    Ray secondary = mkRay(mk3(0), mk3(0), ray.pixeli);
    float random = rnd(rs);
    f3 BRDF = diffuse / PI;
    cache.sample_type = 0;                                   // SAMPLE_IGNORE
    float russianP = 1.0f;
    if (random < material.transmit) {                        // kernels.h:633-653
        state.fromSpecular = true;
        if (inside) {
            const f3 c = mk3(material.absorption);
            state.mask *= mk3(expf(-c.x * hit.t), expf(-c.y * hit.t), expf(-c.z * hit.t));
        }
        float reflected;
        secondary = getRefractRay(ray, colliderNormal, intersectionPos, material.refractive_index, inside, reflected);
        if (rnd(rs) < reflected) {
            state.mask *= diffuse;
            secondary = getReflectRay(ray, colliderNormal, intersectionPos);
        }
        // C++ leaves the evaluation order of the two rand() arguments unspecified.  kernel_shade only
        // exists as DEVICE code, and nvcc's device front end evaluates call arguments left to right
        // (checked in PTX: the first xorshift feeds param0); gcc on the host would go right to left.
        float ra = rnd(rs); float rb = rnd(rs);
        f3 noiseDir = SampleHemisphereCosine(secondary.direction, ra, rb);
        secondary.direction = secondary.direction * (1.0f - material.glossy) + material.glossy * noiseDir;
    } else if (random - material.transmit < material.reflect) {   // kernels.h:654-661
        state.fromSpecular = true;
        state.mask *= diffuse;
        secondary = getReflectRay(ray, colliderNormal, intersectionPos);
        float ra = rnd(rs); float rb = rnd(rs);
        f3 noiseDir = SampleHemisphereCosine(secondary.direction, ra, rb);
        secondary.direction = secondary.direction * (1.0f - material.glossy) + material.glossy * noiseDir;
    } else {                                                 // kernels.h:662-782
        if (!state.albedoSet) { updateAlbedo(o, state.mask * diffuse, x, y); state.albedoSet = true; }
        state.fromSpecular = false;
        // skydome importance sampling, opt-in: same statements as shadeOne in csrc/pt_shade.cuh
        bool lightNee = NEE, skyNee = o.sc.skySample && o.sc.hasSky;
        float pickScale = 1.0f;
        if (lightNee && skyNee) { pickScale = 2.0f; if (rnd(rs) < 0.5f) lightNee = false; else skyNee = false; }
        if (skyNee) {
            const uint SW = o.sc.sky.w, SH = o.sc.sky.h;
            const uint m = binarySearchCdf(o.sc.skyCdf.data(), SW * SH, rnd(rs));
            const uint tx = m % SW, ty = m / SW;
            const float sa = rnd(rs); const float sb = rnd(rs);
            const float su = ((float)tx + sa) / (float)SW, sv = ((float)ty + sb) / (float)SH;
            float sinPolar;
            const f3 sdir = skyUvToDir(su, sv, sinPolar);
            const float pdf = o.sc.skyPdf[m] * (float)(SW * SH) / (2.0f * PI * PI * fmaxf(sinPolar, 0.0001f));
            const float NL = dot(colliderNormal, sdir);
            if (NL > 0 && dot(sdir, surfaceNormal) > 0 && pdf > 0) {
                const f3 sk = skyLookup(o, sdir);
                state.light = state.mask * (BRDF * sk * (NL * pickScale / pdf));
                const float f = NL * NL * NL;
                Ray shadowRay = mkRay(intersectionPos + EPS * f * sdir + EPS * (1 - f) * colliderNormal, sdir, ray.pixeli);
                out.pushShadow = true; out.shadow = pack(shadowRay);
            }
        }
        if (lightNee) {                                      // kernels.h:672-753
            uint successIdx = 0; float valid = 0.0f;
            for (uint k = 0; k < 4; k++) {
                uint lightSource = uint(rnd(rs) * numLights) % numLights;
                const pt_triangle_light& light = sc.lights[lightSource];
                const pt_triangle_v& lv = sc.d.vertices[light.triangle_index];
                const pt_mat4x3& lt = sc.d.instances[light.instance_index].transform;
                f3 centroid = (mk3(lv.v0) + mk3(lv.v1) + mk3(lv.v2)) / 3.0f;
                centroid = matmul(lt, centroid, 1.0f);
                f3 lightNormal = mk3(sc.d.vertexData[light.triangle_index].normal);
                lightNormal = normalize(matmul(lt, lightNormal, 0.0f));
                const f3 fromLight = normalize(intersectionPos - centroid);
                if (dot(lightNormal, fromLight) > 0) { valid += 1.0f; successIdx = lightSource; }
            }
            if (valid > 0) {
                const pt_triangle_light& light = sc.lights[successIdx];
                const pt_triangle_v& lv = sc.d.vertices[light.triangle_index];
                const pt_triangle_d& ld = sc.d.vertexData[light.triangle_index];
                const pt_instance& li = sc.d.instances[light.instance_index];
                const pt_mat4x3& lt = li.transform;
                const f3 v0 = matmul(lt, mk3(lv.v0), 1.0f), v1 = matmul(lt, mk3(lv.v1), 1.0f), v2 = matmul(lt, mk3(lv.v2), 1.0f);
                const f3 v0v1 = v1 - v0, v0v2 = v2 - v0;
                const f3 cr = cross(v0v1, v0v2);
                const float crLength = length(cr);
                float u = rnd(rs), v = rnd(rs);
                if (u + v > 1.0f) { u = 1.0f - u; v = 1.0f - v; }
                const f3 samplePoint = v0 + u * v0v1 + v * v0v2;
                f3 shadowDir = intersectionPos - samplePoint;
                const float shadowLength = length(shadowDir);
                const float invShadowLength = 1.0f / shadowLength;
                shadowDir *= invShadowLength;
                const f3 lightNormal = cr / crLength;
                const float NL = dot(colliderNormal, -shadowDir);
                const float LNL = dot(lightNormal, shadowDir);
                if (NL > 0 && dot(-shadowDir, surfaceNormal) > 0 && LNL > 0) {
                    const uint mid = li.material_id == 0xffffffffu ? ld.material : li.material_id;
                    const f3 emission = mk3(sc.d.materials[mid].emission);
                    float A = 0.5f * crLength;
                    float SA = LNL * A * invShadowLength * invShadowLength;
                    // NL * SA * size * BRDF * emission * (valid/4): left-to-right, float*uint -> float
                    state.light = state.mask * (NL * SA * numLights * BRDF * emission * (valid / 4.0f));
                    if (pickScale != 1.0f) state.light *= pickScale;
                    const float f = LNL * LNL * LNL;
                    Ray shadowRay = mkRay(samplePoint + f * EPS * shadowDir + (1 - f) * EPS * lightNormal, shadowDir, ray.pixeli);
                    shadowRay.length = shadowLength - 2 * EPS;
                    out.pushShadow = true; out.shadow = pack(shadowRay);
                }
            }
        }
        f3 r;
        if (o.DCACHE && hit.primitive_type == TRIANGLE && dot(colliderNormal, originalNormal) > 0) {   // kernels.h:757-766
            const pt_radiance_cache rc = sc.rc[hit.primitive_id];
            float invprob; int bucket;
            r = SampleHemisphereCached(colliderNormal, rs, rc, bucket, invprob);
            cache.cache_bucket_id = bucket;
            state.mask *= invprob;
            cache.sample_type = 2;                           // SAMPLE_BUCKET
            cache.triangle_id = hit.primitive_id;
            cache.cum_mask.x = state.mask.x; cache.cum_mask.y = state.mask.y; cache.cum_mask.z = state.mask.z;
        } else {
            float ra = rnd(rs); float rb = rnd(rs);          // argument evaluation order, see above
            r = SampleHemisphereCosine(colliderNormal, ra, rb);
        }
        if (dot(r, surfaceNormal) < 0) state.mask = mk3(0.0f);
        float f = fmaxf(dot(colliderNormal, r), 0.0f);
        f = f * f * f;
        secondary = mkRay(intersectionPos + EPS * f * r + EPS * (1 - f) * colliderNormal, r, ray.pixeli);
        state.mask *= PI * BRDF;
        russianP = clampf(fmaxcompf(diffuse), 0.1f, 0.9f);
    }
    if (fmaxcompf(state.mask) > 0.0001f && rnd(rs) < russianP) {   // kernels.h:785-793
        state.mask = state.mask / russianP;
        out.pushExt = true; out.ext = pack(secondary);
    } else cache.sample_type = 1;
    setState(o, ray.pixeli, state);
    if (bounce < MAX_CACHE_DEPTH) o.sampleCache[cache_id] = cache;
}


// kernel_clear_state, kernels.h:485-491
static void clearState(Oracle& o) {
    for (uint i = 0; i < o.NP; i++) {
        F4 a = {0, 0, 0, 0}; o.accucolors[i] = a;                        // __int_as_float(false) == +0.0f
        F4 m = {1, 1, 1, 0}; uint one = 1; memcpy(&m.w, &one, 4); o.masks[i] = m;   // __int_as_float(1): fromSpecular = true
    }
}
// kernel_generate_primary_rays, kernels.h:493-501 (queue order = pixel order here; the GPU's is not defined)
static void generatePrimary(Oracle& o, const Camera& cam, uint randIdx, std::vector<pt_ray>& q) {
    q.resize(o.NP);
#pragma omp parallel for schedule(static)
    for (int y = 0; y < (int)o.H; y++)
        for (uint x = 0; x < o.W; x++) {
            uint seed = getSeed(x, y, randIdx, o.W);
            q[x + y * o.W] = pack(camGetRay(cam, x, y, seed, o.W, o.H));
        }
}
// kernel_extend, kernels.h:503-510
static void extend(Oracle& o) {
    const size_t n = o.rayQ.size();
    o.hits.resize(n);
    uint64_t inner = 0, tris = 0, tlas = 0;
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : inner, tris, tlas)
    for (size_t i = 0; i < n; i++) {
        Counters c = {0, 0, 0};
        Hit h = traverseTopLevel<false>(o.sc, unpack(o.rayQ[i]), &c);
        pt_hit p; p.primitive_type = h.primitive_type; p.primitive_id = h.primitive_id; p.instance_id = h.instance_id; p.t = h.t;
        o.hits[i] = p;
        inner += c.inner; tris += c.tris; tlas += c.tlas;
    }
    o.cnt.inner += inner; o.cnt.tris += tris; o.cnt.tlas += tlas;
    o.stats.extend_rays += n;
}
static void shade(Oracle& o, int bounce, uint randIdx) {
    const size_t n = o.rayQ.size();
    std::vector<ShadeOut> outs(n);
    // each pixel owns at most one entry per queue, so entries are independent (SURVEY.md "Determinism");
    // albedo/state writes go to that pixel only.
#pragma omp parallel for schedule(dynamic, 64)
    for (size_t i = 0; i < n; i++) shadeOne(o, (uint)i, bounce, randIdx, outs[i]);
    for (size_t i = 0; i < n; i++) {
        if (outs[i].pushShadow) o.shadowQ.push_back(outs[i].shadow);
        if (outs[i].pushExt) o.rayQNew.push_back(outs[i].ext);
    }
    o.stats.shade_calls += n;
}
// kernel_connect, kernels.h:799-810
static void connect(Oracle& o) {
    const size_t n = o.shadowQ.size();
    uint64_t inner = 0, tris = 0, tlas = 0;
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : inner, tris, tlas)
    for (size_t i = 0; i < n; i++) {
        Counters c = {0, 0, 0};
        const Ray sr = unpack(o.shadowQ[i]);
        Hit h = traverseTopLevel<true>(o.sc, sr, &c);
        inner += c.inner; tris += c.tris; tlas += c.tlas;
        if (h.intersected()) continue;
        const F4 l = o.lights[sr.pixeli];
        F4& a = o.accucolors[sr.pixeli];
        a.x += l.x; a.y += l.y; a.z += l.z; a.w += 0.0f;
    }
    o.cntConn.inner += inner; o.cntConn.tris += tris; o.cntConn.tlas += tlas;
    o.stats.shadow_rays += n;
}
// kernel_add_to_screen, kernels.h:812-824
static void addToScreen(Oracle& o) {
    for (uint i = 0; i < o.NP; i++) {
        const F4 c = o.accucolors[i];
        F4& p = o.image[i];
        F4 r = {fmaxf(0.0f, p.x) + c.x, fmaxf(0.0f, p.y) + c.y, fmaxf(0.0f, p.z) + c.z, p.w + 1.0f};
        p = r;
    }
}
// kernel_init_radiance_cache, kernels.h:848-861
static void initRadianceCache(Oracle& o) {
    o.sc.rc.resize(o.sc.d.num_triangles);
    for (auto& r : o.sc.rc) {
        memset(&r, 0, sizeof(r));
        for (int t = 0; t < 8; t++) { r.radianceCache[t] = 0.1f; r.additionCacheCount[t] = 0.0f; r.additionCache[t] = 0.0f; }
        r.radianceTotal = 8 * 0.1f;
    }
}
// kernel_update_buckets, kernels.h:863-882 (pixel order here; float atomics on the GPU are unordered)
static void updateBuckets(Oracle& o) {
    for (uint i = 0; i < o.NP; i++) {
        const f3 totalEnergy = mk3(o.accucolors[i].x, o.accucolors[i].y, o.accucolors[i].z);
        for (uint bounce = 0; bounce < (uint)MAX_CACHE_DEPTH; bounce++) {
            const pt_sample_cache cache = o.sampleCache[i + (o.NP * bounce)];
            if (cache.sample_type == 1) break;
            if (cache.sample_type == 0) continue;
            pt_radiance_cache& rc = o.sc.rc[cache.triangle_id];
            const float energy = fminf(100.0f, luminance(totalEnergy / mk3(cache.cum_mask)));
            rc.additionCache[cache.cache_bucket_id] += energy;
            rc.additionCacheCount[cache.cache_bucket_id] += 1.0f;
        }
    }
}
// kernel_propagate_buckets, kernels.h:884-905
static void propagateBuckets(Oracle& o) {
    for (auto& rcr : o.sc.rc) {
        pt_radiance_cache rc = rcr;
        const float alpha = 0.95f;
        for (uint t = 0; t < 8; t++) {
            float additionCount = rc.additionCacheCount[t];
            if (additionCount < EPS) continue;
            const float oldValue = rc.radianceCache[t];
            const float incomingEnergy = rc.additionCache[t] / additionCount;
            const float newValue = clampf(alpha * oldValue + (1 - alpha) * incomingEnergy, 0.1f, 2.0f);
            const float deltaValue = newValue - oldValue;
            rc.radianceCache[t] = newValue;
            rc.radianceTotal += deltaValue;
            rc.additionCache[t] = 0;
            rc.additionCacheCount[t] = 0;
        }
        rcr = rc;
    }
}

// Pathtracer::Render, pathtracer.h:224-302
static void render(Oracle& o, const pt_camera& pcam, bool shouldClear) {
    const Camera cam = camFromPod(pcam);
#ifdef _OPENMP
    if (o.threads > 0) omp_set_num_threads(o.threads);
#endif
    if (shouldClear) {
        o.DNEE = o.HNEE; o.DCACHE = o.HCACHE;                            // :242-243
        F4 z = {0, 0, 0, 0};
        std::fill(o.image.begin(), o.image.end(), z);                    // :248-249
        std::fill(o.albedo.begin(), o.albedo.end(), z);
        o.sampleIdx = 0; o.randIdx = 0;                                  // :250-251
    }
    const uint nsamples = shouldClear ? o.cfg.interactive_depth : 1;    // :255 (sic: interactive_depth)
    for (uint sample = 0; sample < nsamples; sample++) {
        clearState(o);                                                   // :258
        o.rayQ.clear(); o.shadowQ.clear(); o.rayQNew.clear();            // :265
        generatePrimary(o, cam, o.randIdx, o.rayQ);                      // :266
        o.stats.primary_rays += o.rayQ.size();
        o.randIdx++;                                                     // :267
        uint max_bounces;                                                // :270-274, host _NEE == HNEE
        if (o.HNEE) max_bounces = shouldClear ? o.cfg.interactive_depth : o.cfg.max_ray_depth;
        else max_bounces = shouldClear ? o.cfg.interactive_depth + 1 : o.cfg.max_ray_depth;
        for (uint bounce = 0; bounce < max_bounces; bounce++) {         // :276-290
            extend(o);
            shade(o, (int)bounce, o.randIdx);
            o.randIdx++;
            if (o.HNEE || o.sc.skySample) connect(o);
            o.rayQ.swap(o.rayQNew); o.rayQNew.clear(); o.shadowQ.clear();
        }
        if (!shouldClear && o.HCACHE && o.sampleIdx < o.cfg.guiding_frames) {   // :292-296
            updateBuckets(o);
            propagateBuckets(o);
        }
        addToScreen(o);                                                  // :299
        o.sampleIdx++;                                                   // :300
        o.stats.samples++;
    }
}

}  // namespace ora
using namespace ora;

// ---------------------------------------------------------------------------------------------------
// C entry points (loaded with ctypes by tests/, smoke() and bench.py's CPU legs only)
// ---------------------------------------------------------------------------------------------------
ORA_API void* ora_create(const pt_config* cfg) {
    Oracle* o = new Oracle();
    o->cfg = *cfg; o->W = cfg->width; o->H = cfg->height; o->NP = o->W * o->H;
    o->HNEE = o->HCACHE = true;            // pathtracer.h:213-214
    o->DNEE = o->DCACHE = false;           // __constant__ zero-initialised until the first shouldClear frame
    o->sampleIdx = o->randIdx = 0;
    F4 z = {0, 0, 0, 0};
    o->masks.assign(o->NP, z); o->accucolors.assign(o->NP, z); o->lights.assign(o->NP, z);
    o->image.assign(o->NP, z); o->albedo.assign(o->NP, z);
    pt_sample_cache sc0; memset(&sc0, 0, sizeof(sc0));
    o->sampleCache.assign((size_t)o->NP * MAX_CACHE_DEPTH, sc0);
    memset(&o->sc.d, 0, sizeof(o->sc.d));
    o->sc.hasSky = false; o->sc.skyConst = mk3(1, 1, 1); o->sc.hasBlue = false;
    memset(&o->cnt, 0, sizeof(o->cnt)); memset(&o->cntConn, 0, sizeof(o->cntConn)); memset(&o->stats, 0, sizeof(o->stats));
    o->threads = 0;
    return o;
}
ORA_API void ora_destroy(void* h) { delete (Oracle*)h; }
ORA_API void ora_set_threads(void* h, int n) { ((Oracle*)h)->threads = n; }
// Keeps the caller's host pointers (caller keeps them alive), extracts emissive triangles
// (pathtracer.h:155-167) and initialises the radiance cache (pathtracer.h:220).
ORA_API int ora_upload_scene(void* h, const pt_scene_desc* d) {
    Oracle& o = *(Oracle*)h;
    o.sc.d = *d;
    o.sc.lights.clear();
    for (uint i = 0; i < d->num_instances; i++) {
        const pt_instance& inst = d->instances[i];
        const pt_model& model = d->models[inst.model_id];
        for (uint t = model.triangleStart; t < model.triangleStart + model.nrTriangles; t++) {
            const uint mat_id = inst.material_id == 0xffffffffu ? d->vertexData[t].material : inst.material_id;
            if (fmaxcompf(mk3(d->materials[mat_id].emission)) < EPS) continue;
            pt_triangle_light l = {t, i}; o.sc.lights.push_back(l);
        }
    }
    initRadianceCache(o);
    return (int)o.sc.lights.size();
}
ORA_API void ora_update_instances(void* h, const pt_instance* inst, uint32_t n, const pt_tlas_node* tlas, uint32_t nn) {
    Oracle& o = *(Oracle*)h; o.sc.d.instances = inst; o.sc.d.num_instances = n; o.sc.d.topBvh = tlas; o.sc.d.num_tlas_nodes = nn;
}
ORA_API void ora_set_toggles(void* h, int nee, int guiding) { Oracle& o = *(Oracle*)h; o.HNEE = nee != 0; o.HCACHE = guiding != 0; }
ORA_API void ora_set_skydome(void* h, const float* rgba, uint32_t w, uint32_t hh) {
    Oracle& o = *(Oracle*)h; o.sc.sky.data = rgba; o.sc.sky.w = w; o.sc.sky.h = hh; o.sc.sky.ch = 4; o.sc.hasSky = rgba != nullptr;
}
ORA_API int ora_set_sky_sampling(void* h, int on) {               // builds the CDF exactly as pathtracer.h:85-122 does
    Oracle& o = *(Oracle*)h;
    if (on && !o.sc.hasSky) return -1;
    if (on) {
        const size_t n = (size_t)o.sc.sky.w * o.sc.sky.h;
        o.sc.skyPdf.resize(n); o.sc.skyCdf.resize(n);
        float totalEnergy = 0.0f;
        for (size_t i = 0; i < n; i++) {
            const float energy = fmaxf(o.sc.sky.data[4 * i], fmaxf(o.sc.sky.data[4 * i + 1], o.sc.sky.data[4 * i + 2]));
            o.sc.skyPdf[i] = energy; totalEnergy += energy; o.sc.skyCdf[i] = totalEnergy;
        }
        for (size_t i = 0; i < n; i++) { o.sc.skyPdf[i] /= totalEnergy; o.sc.skyCdf[i] /= totalEnergy; }
    }
    o.sc.skySample = on != 0;
    return 0;
}
ORA_API void ora_set_sky_constant(void* h, float r, float g, float b) { Oracle& o = *(Oracle*)h; o.sc.hasSky = false; o.sc.skyConst = mk3(r, g, b); }
ORA_API void ora_set_blue_noise(void* h, const float* lum, uint32_t w, uint32_t hh) {
    Oracle& o = *(Oracle*)h; o.sc.blue.data = lum; o.sc.blue.w = w; o.sc.blue.h = hh; o.sc.blue.ch = 1; o.sc.hasBlue = lum != nullptr;
}
ORA_API uint64_t ora_add_texture(void* h, const float* rgba, uint32_t w, uint32_t hh) {
    Oracle& o = *(Oracle*)h; Tex t; t.data = rgba; t.w = w; t.h = hh; t.ch = 4; o.sc.textures.push_back(t); return o.sc.textures.size();
}
ORA_API void ora_render(void* h, const pt_camera* cam, int should_clear) { render(*(Oracle*)h, *cam, should_clear != 0); }
ORA_API void ora_read_image(void* h, float* out, int which) {
    Oracle& o = *(Oracle*)h; memcpy(out, which == 0 ? (void*)o.image.data() : (void*)o.albedo.data(), (size_t)o.NP * 16);
}
ORA_API void ora_read_radiance_cache(void* h, pt_radiance_cache* out) { Oracle& o = *(Oracle*)h; memcpy(out, o.sc.rc.data(), o.sc.rc.size() * sizeof(pt_radiance_cache)); }
ORA_API void ora_get_stats(void* h, pt_stats* st, uint64_t* counters3) {
    Oracle& o = *(Oracle*)h; if (st) *st = o.stats;
    if (counters3) { counters3[0] = o.cnt.inner; counters3[1] = o.cnt.tris; counters3[2] = o.cnt.tlas;
                     counters3[3] = o.cntConn.inner; counters3[4] = o.cntConn.tris; counters3[5] = o.cntConn.tlas; }   // [0..2] extend, [3..5] connect
}
ORA_API void ora_reset_stats(void* h) { Oracle& o = *(Oracle*)h; memset(&o.cnt, 0, sizeof(o.cnt)); memset(&o.cntConn, 0, sizeof(o.cntConn)); memset(&o.stats, 0, sizeof(o.stats)); }
ORA_API void ora_get_rand_state(void* h, uint32_t* sampleIdx, uint32_t* randIdx) { Oracle& o = *(Oracle*)h; *sampleIdx = o.sampleIdx; *randIdx = o.randIdx; }

// Camera as the reference's host prepares it: ctor + update() with no key down (types.h:609-637):
// viewDir = normalize(viewDir); recalculate().   in9 = eye3, viewDir3, d, focalLength, aperture
ORA_API void ora_camera_make(const float* in9, uint32_t W, uint32_t H, pt_camera* out) {
    Camera c; c.eye = mk3(in9[0], in9[1], in9[2]); c.viewDir = normalize(mk3(in9[3], in9[4], in9[5]));
    c.d = in9[6]; c.focalLength = in9[7]; c.aperture = in9[8];
    camRecalculate(c, W, H);
    memset(out, 0, sizeof(*out));
    out->eye.x = c.eye.x; out->eye.y = c.eye.y; out->eye.z = c.eye.z;
    out->viewDir.x = c.viewDir.x; out->viewDir.y = c.viewDir.y; out->viewDir.z = c.viewDir.z;
    out->lt.x = c.lt.x; out->lt.y = c.lt.y; out->lt.z = c.lt.z;
    out->u.x = c.u.x; out->u.y = c.u.y; out->u.z = c.u.z;
    out->v.x = c.v.x; out->v.y = c.v.y; out->v.z = c.v.z;
    out->d = c.d; out->focalLength = c.focalLength; out->aperture = c.aperture;
}
ORA_API void ora_primary_rays(void* h, const pt_camera* cam, uint32_t randIdx, pt_ray* out) {
    Oracle& o = *(Oracle*)h; std::vector<pt_ray> q; generatePrimary(o, camFromPod(*cam), randIdx, q);
    memcpy(out, q.data(), q.size() * sizeof(pt_ray));
}
// traverseTopLevel over a ray buffer.  counters3 (optional) += {inner nodes popped, triangles tested, TLAS nodes fetched}.
ORA_API void ora_trace(void* h, const pt_ray* rays, uint32_t n, int any_hit, pt_hit* out, int threads, uint64_t* counters3) {
    Oracle& o = *(Oracle*)h;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
    uint64_t inner = 0, tris = 0, tlas = 0;
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : inner, tris, tlas)
    for (uint32_t i = 0; i < n; i++) {
        Counters c = {0, 0, 0};
        const Ray r = unpack(rays[i]);
        Hit hh = any_hit ? traverseTopLevel<true>(o.sc, r, &c) : traverseTopLevel<false>(o.sc, r, &c);
        if (!hh.intersected()) { hh.primitive_type = 0; hh.instance_id = 0; }
        else if (any_hit) { hh.primitive_type = 0; hh.instance_id = 0; hh.primitive_id = 0; }
        else if (hh.primitive_type != TRIANGLE) hh.instance_id = 0;
        pt_hit p; p.primitive_type = hh.primitive_type; p.primitive_id = hh.primitive_id; p.instance_id = hh.instance_id; p.t = hh.t;
        out[i] = p;
        inner += c.inner; tris += c.tris; tlas += c.tlas;
    }
    if (counters3) { counters3[0] += inner; counters3[1] += tris; counters3[2] += tlas; }
}

// Per ray: the deepest bottom-level stack the reference's traversal order needs (its own array has 18 unchecked entries).
ORA_API void ora_trace_stack_depth(void* h, const pt_ray* rays, uint32_t n, uint32_t* depth_out) {
    Oracle& o = *(Oracle*)h;
#pragma omp parallel for schedule(dynamic, 64)
    for (uint32_t i = 0; i < n; i++) {
        Counters c = {0, 0, 0, 0};
        traverseTopLevel<false>(o.sc, unpack(rays[i]), &c);
        depth_out[i] = (uint32_t)c.maxStack;
    }
}

// ---- single-function pins (same signatures as oracle/ref_harness.cu's ref_*) -------------------------
ORA_API unsigned ora_wang_hash(unsigned s) { return wang_hash(s); }
ORA_API unsigned ora_get_seed(unsigned x, unsigned y, unsigned randIdx, unsigned W) { return getSeed(x, y, randIdx, W); }
ORA_API float ora_rand(unsigned* seed) { return rnd(*seed); }
ORA_API void ora_normalize(const float* v, float* out) { f3 r = normalize(mk3(v[0], v[1], v[2])); out[0] = r.x; out[1] = r.y; out[2] = r.z; }
ORA_API void ora_sample_cosine(const float* n, float r0, float r1, float* out) {
    f3 r = SampleHemisphereCosine(mk3(n[0], n[1], n[2]), r0, r1); out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
ORA_API void ora_sample_cached(const float* n, unsigned* seed, const float* rc9, float* out_dir, int* bucket, float* invprob) {
    pt_radiance_cache rc; memset(&rc, 0, sizeof(rc));
    for (int i = 0; i < 8; i++) rc.radianceCache[i] = rc9[i];
    rc.radianceTotal = rc9[8];
    RandState rs; rs.seed = *seed; rs.sampleIdx = 1; rs.randIdx = 0; rs.blueNoiseSample = 0; rs.blue_noise_frames = 1;
    f3 r = SampleHemisphereCached(mk3(n[0], n[1], n[2]), rs, rc, *bucket, *invprob);
    *seed = rs.seed; out_dir[0] = r.x; out_dir[1] = r.y; out_dir[2] = r.z;
}
ORA_API void ora_refract(const float* ray6, const float* n, const float* pos, float ior, int inside, float* out7) {
    Ray ray = mkRay(mk3(ray6[0], ray6[1], ray6[2]), mk3(ray6[3], ray6[4], ray6[5]), 0);
    float reflected = 0;
    Ray r = getRefractRay(ray, mk3(n[0], n[1], n[2]), mk3(pos[0], pos[1], pos[2]), ior, inside != 0, reflected);
    out7[0] = r.origin.x; out7[1] = r.origin.y; out7[2] = r.origin.z; out7[3] = r.direction.x; out7[4] = r.direction.y; out7[5] = r.direction.z; out7[6] = reflected;
}
ORA_API void ora_reflect(const float* ray6, const float* n, const float* pos, float* out6) {
    Ray ray = mkRay(mk3(ray6[0], ray6[1], ray6[2]), mk3(ray6[3], ray6[4], ray6[5]), 0);
    Ray r = getReflectRay(ray, mk3(n[0], n[1], n[2]), mk3(pos[0], pos[1], pos[2]));
    out6[0] = r.origin.x; out6[1] = r.origin.y; out6[2] = r.origin.z; out6[3] = r.direction.x; out6[4] = r.direction.y; out6[5] = r.direction.z;
}
ORA_API void ora_normal_to_uv(const float* n, float* out2) { f2 r = normalToUv(mk3(n[0], n[1], n[2])); out2[0] = r.x; out2[1] = r.y; }
ORA_API int ora_tri_intersect(const float* ray6, const float* tri9, float* tuv) {
    Ray ray = mkRay(mk3(ray6[0], ray6[1], ray6[2]), mk3(ray6[3], ray6[4], ray6[5]), 0);
    pt_triangle_v t; memset(&t, 0, sizeof(t)); memcpy(&t.v0, tri9, 12); memcpy(&t.v1, tri9 + 3, 12); memcpy(&t.v2, tri9 + 6, 12);
    return rayTriangleIntersect(ray, t, tuv[0], tuv[1], tuv[2]) ? 1 : 0;
}
ORA_API int ora_sphere_intersect(const float* ray6, const float* pos, float radius, float* t) {
    Ray ray = mkRay(mk3(ray6[0], ray6[1], ray6[2]), mk3(ray6[3], ray6[4], ray6[5]), 0);
    pt_sphere s; s.pos.x = pos[0]; s.pos.y = pos[1]; s.pos.z = pos[2]; s.radius = radius; s.material = 0;
    return raySphereIntersect(ray, s, *t) ? 1 : 0;
}
ORA_API int ora_plane_intersect(const float* ray6, const float* n, float d, float* t) {
    Ray ray = mkRay(mk3(ray6[0], ray6[1], ray6[2]), mk3(ray6[3], ray6[4], ray6[5]), 0);
    pt_plane p; p.normal.x = n[0]; p.normal.y = n[1]; p.normal.z = n[2]; p.d = d; p.material = 0;
    return rayPlaneIntersect(ray, p, *t) ? 1 : 0;
}
ORA_API int ora_slab(const float* origin, const float* invdir, const float* box6, float* tmin) {
    return slabTest(mk3(origin[0], origin[1], origin[2]), mk3(invdir[0], invdir[1], invdir[2]), box6, box6 + 3, *tmin) ? 1 : 0;
}

// ---- display post-process: CPU restatement of the reference's GLSL (src/main.cpp:46-171), chained as main.cpp:316-376.
// lum / albedo: float4 per pixel (rgb sums, w = counts).  blur != 0: gauss_horz (:111-141) -> gauss_vert (:143-171) ->
// quad_fs_blurred (:76-109); blur == 0: quad_fs (:46-74).  The shaders' hard-coded 640/480 bounds are W/H.
// Parity UNPINNED: the GLSL cannot run on this box (no graphics stack); checked by review against the shader source.
ORA_API void ora_postprocess(const float* lum, const float* albedo, int W, int H, float nrSamples, int blur, float* out) {
    const size_t n = (size_t)W * H;
    std::vector<float> horz(n * 4), vert(n * 4);
    const float spread = std::max(0.7f, nrSamples / 200.0f);
    auto gauss = [&](int i) { return expf(-(float)(i * i) * 0.5f * spread) / sqrtf(2.0f * 3.1415926f); };
    const float* src = lum;
    if (blur) {
        for (int y = 0; y < H; y++) for (int x = 0; x < W; x++) {
            float sum = 0.0f, cs[3] = {0, 0, 0};
            for (int i = -3; i < 4; i++) {
                if (x + i < 0 || x + i >= W) continue;
                const float* l = lum + ((size_t)y * W + x + i) * 4; const float* a = albedo + ((size_t)y * W + x + i) * 4;
                const float g = gauss(i);
                for (int k = 0; k < 3; k++) cs[k] = cs[k] + ((l[k] / std::max(a[k], 0.001f)) * nrSamples) * g;
                sum += g;
            }
            float* d = &horz[((size_t)y * W + x) * 4];
            d[0] = cs[0] / sum; d[1] = cs[1] / sum; d[2] = cs[2] / sum; d[3] = nrSamples;
        }
        for (int y = 0; y < H; y++) for (int x = 0; x < W; x++) {
            float sum = 0.0f, cs[3] = {0, 0, 0};
            for (int i = -4; i < 4; i++) {
                if (y + i < 0 || y + i >= H) continue;
                const float* l = &horz[((size_t)(y + i) * W + x) * 4];
                const float g = gauss(i);
                for (int k = 0; k < 3; k++) cs[k] = cs[k] + l[k] * g;
                sum += g;
            }
            float* d = &vert[((size_t)y * W + x) * 4];
            d[0] = cs[0] / sum; d[1] = cs[1] / sum; d[2] = cs[2] / sum; d[3] = nrSamples;
        }
        src = vert.data();
    }
    for (int y = 0; y < H; y++) for (int x = 0; x < W; x++) {
        const float u = ((float)x + 0.5f) / (float)W - 0.5f, v = ((float)y + 0.5f) / (float)H - 0.5f;
        const float* l = src + ((size_t)y * W + x) * 4; const float* a = albedo + ((size_t)y * W + x) * 4;
        const float vig = 1.0f - (u * u + v * v);
        for (int k = 0; k < 3; k++) {
            float c = l[k] / l[3];
            if (blur) c = c * (a[k] / a[3]);
            out[((size_t)y * W + x) * 4 + k] = powf(c, 1.0f / 2.0f) * vig;
        }
        out[((size_t)y * W + x) * 4 + 3] = vig;
    }
}
