// TEST INFRASTRUCTURE -- NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load the library built from this file.
//
// oracle/_ref harness: compiles the UNMODIFIED reference headers where they lie under
// /root/reference/src (nothing is copied) as HOST code through nvcc, exactly as the reference
// compiles its single translation unit (reference CMakeLists.txt:28), and exposes the reference's
// own HYBRID functions through a C ABI so that the CPU restatement (oracle/pt_oracle.cpp), the
// product's host code and the CUDA kernels can be pinned against them:
//   Scene / createBVHBinned / BuildTopLevelBVH / ConvertToInstance   (scene.h, bvhBuilder.h)
//   Camera::getRay / getSeed / rand                                  (types.h:641-658, use_cuda.h:61-101)
//   traverseTopLevel<false|true>                                     (kernels.h:277-387)
//   SampleHemisphereCosine / SampleHemisphereCached / getRefractRay / getReflectRay / normalToUv
// Shims (oracle/shim, oracle/glfwstub.cpp) replace GL typedefs, glm and three GLFW symbols.
#include <map>
#include <fstream>
#include <sstream>
#include <cstring>
#include <cassert>
#include <string>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "types.h"
#include "bvhBuilder.h"
#include "scene.h"
#include "globals.h"
#include "kernels.h"

#define REF_API extern "C" __attribute__((visibility("default")))

struct RefScene {
    Scene scene;
    bool finalized = false;
};

REF_API int ref_width() { return WINDOW_WIDTH; }
REF_API int ref_height() { return WINDOW_HEIGHT; }

REF_API void ref_sizes(unsigned* out) {
    // struct sizes pinned by tests against include/ptb200.h
    out[0] = sizeof(Material);   out[1] = sizeof(TriangleV);   out[2] = sizeof(TriangleD);
    out[3] = sizeof(RadianceCache); out[4] = sizeof(BVHNode);  out[5] = sizeof(Model);
    out[6] = sizeof(Instance);   out[7] = sizeof(TopLevelBVH); out[8] = sizeof(Sphere);
    out[9] = sizeof(Plane);      out[10] = sizeof(RayPacked);  out[11] = sizeof(HitInfo);
    out[12] = sizeof(SampleCache); out[13] = sizeof(Camera);   out[14] = sizeof(RandState);
    out[15] = sizeof(SceneBuffers); out[16] = sizeof(TriangleLight); out[17] = sizeof(TraceState);
}

REF_API void* ref_scene_new() { return new RefScene(); }
REF_API void ref_scene_free(void* h) { delete (RefScene*)h; }

// m = diffuse[3], specular[3], emission[3], reflect, glossy, transmit, ior, absorption[3]   (16 floats)
REF_API unsigned ref_add_material(void* h, const float* m) {
    Material mat = Material::DIFFUSE(make_float3(m[0], m[1], m[2]));
    mat.specular_color = make_float3(m[3], m[4], m[5]);
    mat.emission = make_float3(m[6], m[7], m[8]);
    mat.reflect = m[9]; mat.glossy = m[10]; mat.transmit = m[11]; mat.refractive_index = m[12];
    mat.absorption = make_float3(m[13], m[14], m[15]);
    mat.texture = 0; mat.normal_texture = 0;
    return ((RefScene*)h)->scene.addMaterial(mat);
}

// Scene::addModel (scene.h:159-347) on an OBJ file.  useMtl with textures needs a CUDA device
// (loadTexture, use_cuda.h:108) and is therefore only usable for texture-free MTL files here.
REF_API unsigned ref_add_model_obj(void* h, const char* path, float scale, const float* rot, const float* off,
                                   unsigned material, int useMtl) {
    return ((RefScene*)h)->scene.addModel(std::string(path), scale, make_float3(rot[0], rot[1], rot[2]),
                                         make_float3(off[0], off[1], off[2]), material, useMtl != 0);
}

// The tail of Scene::addModel (scene.h:255-346) for geometry that does not come from an OBJ file
// (procedural proxies for the missing meshes): push TriangleV/TriangleD, run the reference's
// createBVHBinned on them, push the Model.  v = n*9 floats, nrm = n*3 floats (flat normal per
// triangle) or NULL => normalize(cross(e1,e2)) as scene.h:297-299; uv = n*6 floats or NULL.
REF_API unsigned ref_add_model_soup(void* h, const float* v, const float* nrm, const float* uv,
                                    const unsigned* mats, unsigned n, unsigned material) {
    Scene& scene = ((RefScene*)h)->scene;
    Model model;
    model.triangleStart = scene.allVertices.size();
    model.nrTriangles = 0;
    for (unsigned i = 0; i < n; i++) {
        float3 v0 = make_float3(v[i * 9 + 0], v[i * 9 + 1], v[i * 9 + 2]);
        float3 v1 = make_float3(v[i * 9 + 3], v[i * 9 + 4], v[i * 9 + 5]);
        float3 v2 = make_float3(v[i * 9 + 6], v[i * 9 + 7], v[i * 9 + 8]);
        float3 normal;
        if (nrm) normal = make_float3(nrm[i * 3], nrm[i * 3 + 1], nrm[i * 3 + 2]);
        else normal = normalize(cross(v1 - v0, v2 - v0));
        float2 uv0 = make_float2(0), uv1 = make_float2(0), uv2 = make_float2(0);
        if (uv) { uv0 = make_float2(uv[i*6], uv[i*6+1]); uv1 = make_float2(uv[i*6+2], uv[i*6+3]); uv2 = make_float2(uv[i*6+4], uv[i*6+5]); }
        scene.allVertices.push_back(TriangleV(v0, v1, v2));
        scene.allVertexData.push_back(TriangleD(normal, make_float3(0), make_float3(0), uv0, uv1, uv2, mats ? mats[i] : material));
        model.nrTriangles++;
    }
    model.bvh = createBVHBinned(scene.allVertices.data() + model.triangleStart, scene.allVertexData.data() + model.triangleStart,
                                model.nrTriangles, model.triangleStart, &model.nrBvhNodes);
    scene.models.push_back(model);
    return scene.models.size() - 1;
}

REF_API void ref_add_object(void* h, unsigned model, unsigned material, const float* pos, const float* rot, const float* scl) {
    GameObject o(model, material);
    o.position = make_float3(pos[0], pos[1], pos[2]);
    o.rotation = make_float3(rot[0], rot[1], rot[2]);
    o.scale = make_float3(scl[0], scl[1], scl[2]);
    ((RefScene*)h)->scene.addObject(o);
}
REF_API void ref_set_object(void* h, unsigned idx, const float* pos, const float* rot, const float* scl) {
    GameObject& o = ((RefScene*)h)->scene.objects[idx];
    o.position = make_float3(pos[0], pos[1], pos[2]);
    o.rotation = make_float3(rot[0], rot[1], rot[2]);
    o.scale = make_float3(scl[0], scl[1], scl[2]);
}
REF_API void ref_add_sphere(void* h, const float* pos, float radius, unsigned material) {
    ((RefScene*)h)->scene.addSphere(Sphere(make_float3(pos[0], pos[1], pos[2]), radius, material));
}
REF_API void ref_add_plane(void* h, const float* normal, float d, unsigned material) {
    ((RefScene*)h)->scene.addPlane(Plane(make_float3(normal[0], normal[1], normal[2]), d, material));
}
REF_API void ref_finalize(void* h) {
    RefScene* s = (RefScene*)h;
    s->scene.finalize();   // scene.h:354-365
    s->finalized = true;
}
// The data half of Scene::update (scene.h:396-400): instances from GameObjects, then the TLAS.
REF_API void ref_update(void* h) {
    Scene& scene = ((RefScene*)h)->scene;
    for (size_t i = 0; i < scene.objects.size(); i++) scene.instances[i] = ConvertToInstance(scene.objects[i]);
    BuildTopLevelBVH(scene.topLevelBVH.data(), scene.instances, scene.models.data(), scene.objects.size());
}

// counts: [0]=triangles [1]=models [2]=instances [3]=tlas nodes [4]=materials [5]=spheres [6]=planes
REF_API void ref_counts(void* h, unsigned* out) {
    Scene& s = ((RefScene*)h)->scene;
    out[0] = s.allVertices.size(); out[1] = s.models.size(); out[2] = s.objects.size();
    out[3] = s.topLevelBVH.size(); out[4] = s.materials.size(); out[5] = s.spheres.size(); out[6] = s.planes.size();
}
REF_API const void* ref_ptr(void* h, int what, unsigned idx) {
    Scene& s = ((RefScene*)h)->scene;
    switch (what) {
        case 0: return s.allVertices.data();
        case 1: return s.allVertexData.data();
        case 2: return s.models.data();
        case 3: return s.instances;
        case 4: return s.topLevelBVH.data();
        case 5: return s.materials.data();
        case 6: return s.spheres.data();
        case 7: return s.planes.data();
        case 8: return s.models[idx].bvh;
    }
    return nullptr;
}

static SceneBuffers hostBuffers(Scene& scene) {   // raytracer.h:35-44
    SceneBuffers b;
    memset(&b, 0, sizeof(b));
    b.vertices = scene.allVertices.data();
    b.vertexData = scene.allVertexData.data();
    b.num_triangles = scene.allVertices.size();
    b.instances = scene.instances;
    b.models = scene.models.data();
    b.materials = scene.materials.data();
    b.topBvh = scene.topLevelBVH.data();
    b.spheres = scene.spheres.data();
    b.num_spheres = scene.spheres.size();
    b.planes = scene.planes.data();
    b.num_planes = scene.planes.size();
    return b;
}

// traverseTopLevel<false/true> (kernels.h:277-387) over a RayPacked buffer, parallelised like the
// reference's CPU tracer (raytracer.h:66, `omp parallel for schedule(dynamic)`), thread count
// chosen by the caller (the reference hard-codes 8, raytracer.h:48).  out = HitInfo[n] (16 B each:
// primitive_type, primitive_id, instance_id, t).  For any_hit, primitive_id != ~0u <=> occluded.
REF_API void ref_trace(void* h, const RayPacked* rays, unsigned n, int any_hit, HitInfo* out, int threads) {
    Scene& scene = ((RefScene*)h)->scene;
    const SceneBuffers b = hostBuffers(scene);
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#endif
#pragma omp parallel for schedule(dynamic, 64)
    for (unsigned i = 0; i < n; i++) {
        const Ray ray = rays[i].getRay();
        HitInfo hi;
        hi.primitive_type = TRIANGLE; hi.instance_id = 0;
        if (any_hit) hi = traverseTopLevel<true>(b, ray);
        else hi = traverseTopLevel<false>(b, ray);
        if (!hi.intersected()) { hi.primitive_type = TRIANGLE; hi.instance_id = 0; }
        else if (any_hit) { hi.primitive_type = TRIANGLE; hi.instance_id = 0; hi.primitive_id = 0; }
        else if (hi.primitive_type != TRIANGLE) hi.instance_id = 0;
        out[i] = hi;
    }
}

// Camera as the reference's host builds it: ctor (types.h:609) then update(window) with no key
// pressed (types.h:612-637) == viewDir = normalize(viewDir); recalculate().
// in  = eye3, viewDir3, d, focalLength, aperture (9);   out = eye3, viewDir3, lt3, u3, v3, d, focal, aperture (18)
static Camera makeCamera(const float* in) {
    Camera cam(make_float3(in[0], in[1], in[2]), make_float3(in[3], in[4], in[5]), in[6], in[7], in[8]);
    cam.update(nullptr);
    return cam;
}
REF_API void ref_camera(const float* in, float* out) {
    Camera c = makeCamera(in);
    const float3 f[5] = {c.eye, c.viewDir, c.lt, c.u, c.v};
    for (int i = 0; i < 5; i++) { out[i*3] = f[i].x; out[i*3+1] = f[i].y; out[i*3+2] = f[i].z; }
    out[15] = c.d; out[16] = c.focalLength; out[17] = c.aperture;
}

// kernel_generate_primary_rays (kernels.h:493-501) for every pixel, written at index pixeli.
REF_API void ref_primary_rays(const float* cam_in, unsigned randIdx, RayPacked* out) {
    const Camera cam = makeCamera(cam_in);
#pragma omp parallel for schedule(static)
    for (int y = 0; y < WINDOW_HEIGHT; y++)
        for (int x = 0; x < WINDOW_WIDTH; x++) {
            uint seed = getSeed(x, y, randIdx);
            out[x + y * WINDOW_WIDTH] = RayPacked(cam.getRay(x, y, seed));
        }
}

// ---- single-function pins ------------------------------------------------------------------
REF_API unsigned ref_wang_hash(unsigned s) { return wang_hash(s); }
REF_API unsigned ref_get_seed(unsigned x, unsigned y, unsigned randIdx) { return getSeed(x, y, randIdx); }
REF_API float ref_rand(unsigned* seed) { return rand(*seed); }
REF_API void ref_normalize(const float* v, float* out) {
    float3 r = normalize(make_float3(v[0], v[1], v[2])); out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
REF_API void ref_sample_cosine(const float* n, float r0, float r1, float* out) {
    float3 r = SampleHemisphereCosine(make_float3(n[0], n[1], n[2]), r0, r1); out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
// rc = radianceCache[8] + radianceTotal (9 floats); host rand(RandState&) == xorshift (kernels.h:20-29)
REF_API void ref_sample_cached(const float* n, unsigned* seed, const float* rc9, float* out_dir, int* bucket, float* invprob) {
    RadianceCache rc; memset(&rc, 0, sizeof(rc));
    for (int i = 0; i < 8; i++) rc.radianceCache[i] = rc9[i];
    rc.radianceTotal = rc9[8];
    RandState rs; memset(&rs, 0, sizeof(rs)); rs.seed = *seed; rs.sampleIdx = 1;
    float3 r = SampleHemisphereCached(make_float3(n[0], n[1], n[2]), rs, rc, *bucket, *invprob);
    *seed = rs.seed; out_dir[0] = r.x; out_dir[1] = r.y; out_dir[2] = r.z;
}
// ray = origin3, dir3;  out = origin3, dir3, reflected
REF_API void ref_refract(const float* ray6, const float* n, const float* pos, float ior, int inside, float* out7) {
    Ray ray(make_float3(ray6[0], ray6[1], ray6[2]), make_float3(ray6[3], ray6[4], ray6[5]), 0u);
    Material m = Material::DIFFUSE(make_float3(1)); m.refractive_index = ior;
    float reflected = 0;
    bool in = inside != 0;
    Ray r = getRefractRay(ray, make_float3(n[0], n[1], n[2]), make_float3(pos[0], pos[1], pos[2]), m, in, reflected);
    out7[0] = r.origin.x; out7[1] = r.origin.y; out7[2] = r.origin.z;
    out7[3] = r.direction.x; out7[4] = r.direction.y; out7[5] = r.direction.z; out7[6] = reflected;
}
REF_API void ref_reflect(const float* ray6, const float* n, const float* pos, float* out6) {
    Ray ray(make_float3(ray6[0], ray6[1], ray6[2]), make_float3(ray6[3], ray6[4], ray6[5]), 0u);
    Ray r = getReflectRay(ray, make_float3(n[0], n[1], n[2]), make_float3(pos[0], pos[1], pos[2]));
    out6[0] = r.origin.x; out6[1] = r.origin.y; out6[2] = r.origin.z;
    out6[3] = r.direction.x; out6[4] = r.direction.y; out6[5] = r.direction.z;
}
REF_API void ref_normal_to_uv(const float* n, float* out2) {
    float2 r = normalToUv(make_float3(n[0], n[1], n[2])); out2[0] = r.x; out2[1] = r.y;
}
REF_API int ref_tri_intersect(const float* ray6, const float* tri9, float* tuv) {
    Ray ray(make_float3(ray6[0], ray6[1], ray6[2]), make_float3(ray6[3], ray6[4], ray6[5]), 0u);
    TriangleV t(make_float3(tri9[0], tri9[1], tri9[2]), make_float3(tri9[3], tri9[4], tri9[5]), make_float3(tri9[6], tri9[7], tri9[8]));
    return rayTriangleIntersect(ray, t, tuv[0], tuv[1], tuv[2]) ? 1 : 0;
}
REF_API int ref_sphere_intersect(const float* ray6, const float* pos, float radius, float* t) {
    Ray ray(make_float3(ray6[0], ray6[1], ray6[2]), make_float3(ray6[3], ray6[4], ray6[5]), 0u);
    Sphere s(make_float3(pos[0], pos[1], pos[2]), radius, 0);
    return raySphereIntersect(ray, s, *t) ? 1 : 0;
}
REF_API int ref_plane_intersect(const float* ray6, const float* n, float d, float* t) {
    Ray ray(make_float3(ray6[0], ray6[1], ray6[2]), make_float3(ray6[3], ray6[4], ray6[5]), 0u);
    Plane p(make_float3(n[0], n[1], n[2]), d, 0);
    return rayPlaneIntersect(ray, p, *t) ? 1 : 0;
}
REF_API int ref_slab(const float* origin, const float* invdir, const float* box6, float* tmin) {
    Box b(make_float3(box6[0], box6[1], box6[2]), make_float3(box6[3], box6[4], box6[5]));
    return slabTest(make_float3(origin[0], origin[1], origin[2]), make_float3(invdir[0], invdir[1], invdir[2]), b, *tmin) ? 1 : 0;
}
// createBVHBinned wall time (bvhBuilder.h:35-274) on a private copy of the soup; returns ms, node count in *nodes.
REF_API double ref_time_bvh_build(const float* v, unsigned n, unsigned* nodes) {
    std::vector<TriangleV> tv; std::vector<TriangleD> td;
    for (unsigned i = 0; i < n; i++) {
        tv.push_back(TriangleV(make_float3(v[i*9], v[i*9+1], v[i*9+2]), make_float3(v[i*9+3], v[i*9+4], v[i*9+5]), make_float3(v[i*9+6], v[i*9+7], v[i*9+8])));
        td.push_back(TriangleD());
    }
    double t0 = glfwGetTime();
    BVHNode* b = createBVHBinned(tv.data(), td.data(), n, 0, nodes);
    double t1 = glfwGetTime();
    free(b);
    return (t1 - t0) * 1000.0;
}
