// pt_shade.cuh -- per-bounce BSDF shading, NEE light sampling and path-guided direction sampling.
// Behaviour follows reference src/kernels.h:513-796 (kernel_shade) statement by statement; the code
// is organised differently (no early returns, queue pushes hoisted to one warp-aggregated append per
// warp) so that a whole warp compacts its survivors with a single atomic.
#pragma once
#include "pt_trace.cuh"

namespace ptb {

struct FrameTex {
    cudaTextureObject_t sky;        // float4, linear, wrap, normalised (use_cuda.h:202-263); 0 => constant sky
    cudaTextureObject_t blue;       // float, point, wrap, normalised (use_cuda.h:166-200); 0 => 0.0
    float3 skyConst;
    float2 invBlueSize;
    // skydome importance sampling (opt-in, pt_set_sky_sampling): the CDF Pathtracer::Init builds (pathtracer.h:85-122)
    const float* skyPdf;            // values[]: max(r,g,b) of texel i / total
    const float* skyCdf;            // cumValues[]: running sum / total
    uint skyW, skyH, skySample;
};

struct TraceState { float3 mask, accucolor, light; bool fromSpecular, albedoSet; };   // types.h:461-468

// Per-light terms of next event estimation (kernels.h:676-716) depend on the light only, not on the shading point:
// they are computed once per scene / instance update with the SAME expressions the reference evaluates per ray
// (bit-identical values), PT_LIGHT_STRIDE float4 per light:
//   [0] centroid (world)  [1] normalize(T * normal)  [2] v0  [3] v1-v0  [4] v2-v0  [5] cross/|cross|, |cross|  [6] emission
#define PT_LIGHT_STRIDE 8
__device__ __forceinline__ void prepareLight(const DScene& sc, uint i) {
    const pt_triangle_light light = sc.lights[i];
    const pt_triangle_v lv = sc.vertices[light.triangle_index];
    const pt_instance& li = sc.instances[light.instance_index];
    const pt_mat4x3& T = li.transform;
    float3 centroid = (f3(lv.v0.x, lv.v0.y, lv.v0.z) + f3(lv.v1.x, lv.v1.y, lv.v1.z) + f3(lv.v2.x, lv.v2.y, lv.v2.z)) / 3.0f;
    centroid = matmul(T, centroid, 1.0f);
    const pt_float3 ln = sc.vertexData[light.triangle_index].normal;
    const float3 lightNormal = normalize(matmul(T, f3(ln.x, ln.y, ln.z), 0.0f));
    const float3 v0 = matmul(T, f3(lv.v0.x, lv.v0.y, lv.v0.z), 1.0f);
    const float3 v1 = matmul(T, f3(lv.v1.x, lv.v1.y, lv.v1.z), 1.0f);
    const float3 v2 = matmul(T, f3(lv.v2.x, lv.v2.y, lv.v2.z), 1.0f);
    const float3 v0v1 = v1 - v0, v0v2 = v2 - v0;
    const float3 cr = cross(v0v1, v0v2);
    const float crLength = length(cr);
    const float3 n2 = cr / crLength;
    const uint mid = li.material_id == 0xffffffffu ? sc.vertexData[light.triangle_index].material : li.material_id;
    const pt_float3 em = sc.materials[mid].emission;
    float4* o = sc.lightTab + (size_t)i * PT_LIGHT_STRIDE;
    o[0] = make_float4(centroid.x, centroid.y, centroid.z, 0.0f);
    o[1] = make_float4(lightNormal.x, lightNormal.y, lightNormal.z, 0.0f);
    o[2] = make_float4(v0.x, v0.y, v0.z, 0.0f);
    o[3] = make_float4(v0v1.x, v0v1.y, v0v1.z, 0.0f);
    o[4] = make_float4(v0v2.x, v0v2.y, v0v2.z, 0.0f);
    o[5] = make_float4(n2.x, n2.y, n2.z, crLength);
    o[6] = make_float4(em.x, em.y, em.z, 0.0f);
    o[7] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
}

// kernels.h:31-36
__device__ __forceinline__ float2 normalToUv(float3 n) {
    return make_float2(atan2f(n.x, n.z) / (2.0f * PT_PI), -acosf(n.y) / PT_PI);
}
// kernels.h:390-406
__device__ __forceinline__ float3 sampleHemisphereCosine(float3 normal, float r0, float r1) {
    float r = sqrtf(r0);
    float theta = 2 * PT_PI * r1;
    float x = r * cosf(theta);
    float y = r * sinf(theta);
    float3 sample = f3(x, y, sqrtf(1 - r0));
    const float3 w = normal;
    float3 u = normalize(cross((fabsf(w.x) > .1f ? f3(0, 1, 0) : f3(1, 0, 0)), w));
    float3 v = normalize(cross(w, u));
    return normalize(f3(dot(sample, f3(u.x, v.x, w.x)), dot(sample, f3(u.y, v.y, w.y)), dot(sample, f3(u.z, v.z, w.z))));
}
// kernels.h:408-431; rc = radianceCache[8] (two float4) + radianceTotal
__device__ __forceinline__ float3 sampleHemisphereCached(float3 normal, RandState& rs, const float4 rc0, const float4 rc1, float total,
                                                         int& sampleBucket, float& invprob) {
    const float rcv[8] = {rc0.x, rc0.y, rc0.z, rc0.w, rc1.x, rc1.y, rc1.z, rc1.w};
    const float sample = rnd(rs) * total;
    float acc = PT_EPS;
    sampleBucket = -1;
    float picked = 0.0f;
#pragma unroll
    for (int b = 0; b < 8; b++) {             // do { acc += rc[++bucket]; } while (acc < sample);  bounded to the 8 buckets
        if (sampleBucket == b - 1 && (b == 0 || acc < sample)) { acc += rcv[b]; sampleBucket = b; picked = rcv[b]; }
    }
    const float r0Min = (sampleBucket < 4) ? 0 : 0.5f;
    const float r0Max = (sampleBucket < 4) ? 0.5f : 1.0f;
    const uint r1i = sampleBucket % 4;
    const float r1Min = r1i * (1.0f / 4.0f);
    const float r1Max = (r1i + 1.0f) * (1.0f / 4.0f);
    const float r0R = rnd(rs);
    const float r1R = rnd(rs);
    const float r0 = r0Min * r0R + r0Max * (1.0f - r0R);
    const float r1 = r1Min * r1R + r1Max * (1.0f - r1R);
    invprob = total / (picked * 8.0f);
    return sampleHemisphereCosine(normal, r0, r1);
}

// kernels.h:65-79 (binarySearch over CDF::cumValues).  The reference returns its last probe, which is the texel just
// BELOW the one containing `target` whenever the search ends by raising L (no kernel ever called it, so it went unnoticed);
// here the search is completed: the result is the first texel whose cumulative value reaches target.
__device__ __forceinline__ uint binarySearchCdf(const float* values, uint nrValues, float target) {
    uint L = 0, R = nrValues - 1;
    while (L + 1 < R) {
        const uint m = (L + R) / 2;
        if (values[m] >= target) R = m; else L = m;
    }
    return values[L] >= target ? L : R;
}
// inverse of normalToUv (kernels.h:31-36) under the wrap addressing of the skydome texture: u = atan2(n.x, n.z) / 2pi,
// v = -acos(n.y) / pi (mod 1).  (The reference's uvToNormal, kernels.h:39-49, mirrors n.y and is not that inverse.)
__device__ __forceinline__ float3 skyUvToDir(float u, float v, float& sinPolar) {
    const float theta = u * 2.0f * PT_PI;
    const float a = PT_PI * (1.0f - v);
    sinPolar = sinf(a);
    return f3(sinf(theta) * sinPolar, cosf(a), cosf(theta) * sinPolar);
}

struct ShadeRay { float3 o, d; float len; };

// kernels.h:452-456
__device__ __forceinline__ ShadeRay reflectRay(float3 dir, float3 normal, float3 pos) {
    ShadeRay r; r.d = reflect(dir, normal); r.o = pos + PT_EPS * r.d; r.len = 9999999; return r;
}
// kernels.h:458-483
__device__ __forceinline__ ShadeRay refractRay(float3 dir, float3 normal, float3 pos, float ior, bool inside, float& reflected) {
    ShadeRay r; r.len = 9999999;
    const float n1 = inside ? ior : 1.0f;
    const float n2 = inside ? 1.0f : ior;
    const float eta = n1 / n2;
    const float costi = dot(normal, -dir);
    const float k = 1 - (eta * eta) * (1 - costi * costi);
    if (k < 0) { reflected = 1; r.o = f3(0); r.d = f3(0); return r; }
    const float3 refractDir = normalize(eta * dir + normal * (eta * costi - sqrtf(k)));
    const float sinti = sqrtf(fmaxf(0.0f, 1.0f - costi - costi));
    const float costt = sqrtf(1.0f - eta * eta * sinti * sinti);
    const float spol = (n1 * costi - n2 * costt) / (n1 * costi + n2 * costt);
    const float ppol = (n1 * costt - n2 * costi) / (n1 * costt + n2 * costi);
    reflected = 0.5f * (spol * spol + ppol * ppol);
    r.o = pos + PT_EPS * refractDir; r.d = refractDir;
    return r;
}

// updateAlbedo (kernels.h:56-62) happens at most once per pixel per sample (albedoSet).  The value is parked per
// slot and k_accumulate applies the reference's read-modify-write (max(0,old)+v, w+1) in sample order, so two
// samples in flight (pt_api.cu) cannot race on the albedo image and the float sums keep the reference's order.
__device__ __forceinline__ void updateAlbedo(float4* albedoSample, uint slot, float3 nv) {
    albedoSample[slot] = make_float4(nv.x, nv.y, nv.z, 0.0f);
}

struct ShadeParams {
    DScene sc;
    TileMap map;
    FrameTex tex;
    float4 *masks, *accucolors, *lights;    // TraceStateSOA (types.h:470-499), indexed by local slot
    pt_sample_cache* sampleCache;           // [MAX_CACHE_DEPTH][nSlots]
    float4* albedoSample;                   // this sample's albedo contribution per local slot (applied by k_accumulate)
    uint sampleIdx, randIdx, blueFrames;
    int bounce;
    bool nee, cache;                        // device _NEE (globals.h:23) and DCACHE
    bool cacheWrite;                        // false: this sample's SampleCache entries are never read (overlapped non-training sample)
    uint* sortKeys[2];                      // ray sorting: bin per pushed extension [0] / shadow [1] ray, nullptr = off
    uint* sortHist[2];
};

struct ShadeOut { bool pushExt, pushShadow; ShadeRay ext, shadow; };

// groupMask != 0: the call is made redundantly by the lanes named in it (k_tail_coop: 8 lanes share a path and compute identical
// values); they are synchronised between the reads of the slot's state and its writes, and only `writer` lanes store.
__device__ __forceinline__ void shadeOne(const ShadeParams& P, float3 ro, float3 rd, uint slot, const Hit hitIn, ShadeOut& out, const uint groupMask = 0u, const bool writer = true) {
    const DScene& sc = P.sc;
    Hit hit = hitIn;
#if defined(PT_BOUNDS_CHECK)
    slot = PT_IDX(slot, PT_LIM(nSlots), PT_CHK_SLOT);
    if (hit.id != 0xffffffffu) {
        if (hit.type == 0) { hit.id = PT_IDX(hit.id, sc.num_triangles, PT_CHK_TRIANGLE); hit.inst = PT_IDX(hit.inst, sc.num_instances, PT_CHK_INSTANCE); }
        else if (hit.type == 1) hit.id = PT_IDX(hit.id, sc.num_spheres, PT_CHK_TRIANGLE);
        else hit.id = PT_IDX(hit.id, sc.num_planes, PT_CHK_TRIANGLE);
    }
#endif
    out.pushExt = out.pushShadow = false;
    uint x, y; slotToPixel(P.map, slot, x, y);
    TraceState state;
    if (P.bounce == 0) {
        // kernel_clear_state (kernels.h:485-491) folded in: the first shade of a sample starts from the cleared state instead of
        // reading it back (mask 1, fromSpecular, nothing accumulated; `light` is only ever read after the shade that sets it)
        state.mask = f3(1.0f); state.accucolor = f3(0.0f); state.light = f3(0.0f);
        state.fromSpecular = true; state.albedoSet = false;
    } else {
        const float4 m = P.masks[slot], a = P.accucolors[slot], l = P.lights[slot];
        state.mask = f3(m); state.accucolor = f3(a); state.light = f3(l);
        state.fromSpecular = (__float_as_uint(m.w) & 0xffu) != 0; state.albedoSet = (__float_as_uint(a.w) & 0xffu) != 0;
    }
    pt_sample_cache cache; cache.sample_type = 1; cache.triangle_id = 0; cache.cache_bucket_id = 0; cache.cum_mask.x = cache.cum_mask.y = cache.cum_mask.z = 0;
    bool writeState = true, writeCache = true;

    if (hit.id == 0xffffffffu) {                                              // kernels.h:526-538
        float3 sk = P.tex.skyConst;
        if (P.tex.sky) { float2 uv = normalToUv(rd); sk = f3(tex2D<float4>(P.tex.sky, uv.x, uv.y)); }
        if (!state.albedoSet) { updateAlbedo(P.albedoSample, slot, state.mask * sk); state.albedoSet = true; }
        // with skydome sampling on, a diffuse vertex has already collected the sky through its shadow ray (like emitters under NEE, kernels.h:567)
        if (!(P.tex.skySample && P.tex.sky) || state.fromSpecular) state.accucolor += state.mask * sk;
    } else {
        RandState rs; rs.sampleIdx = P.sampleIdx; rs.randIdx = P.randIdx; rs.blueFrames = P.blueFrames;
        rs.seed = getSeed(x, y, rs.randIdx, P.map.W);
        rs.blueNoiseSample = 0.0f;
        if (P.tex.blue) rs.blueNoiseSample = tex2D<float>(P.tex.blue, (float)x * P.tex.invBlueSize.x, (float)y * P.tex.invBlueSize.y);

        const pt_instance* instance = sc.instances + (hit.type == 0 ? hit.inst : 0);
        const float3 intersectionPos = ro + hit.t * rd;
        uint material_id; float3 originalNormal;
        if (hit.type == 0) {
            const uint im = instance->material_id;
            material_id = im == 0xffffffffu ? sc.vertexData[hit.id].material : im;
            const pt_float3 n = sc.vertexData[hit.id].normal;
            originalNormal = normalize(matmul(instance->transform, f3(n.x, n.y, n.z), 0.0f));
        } else if (hit.type == 1) {
            const pt_sphere s = sc.spheres[hit.id];
            material_id = s.material;
            originalNormal = normalize(intersectionPos - f3(s.pos.x, s.pos.y, s.pos.z));
        } else {
            const pt_plane p = sc.planes[hit.id];
            material_id = p.material; originalNormal = f3(p.normal.x, p.normal.y, p.normal.z);
        }
        const pt_material material = sc.materials[PT_IDX(material_id, PT_LIM(numMaterials), PT_CHK_MATERIAL)];
        const bool inside = dot(rd, originalNormal) > 0;
        const float3 surfaceNormal = inside ? -originalNormal : originalNormal;
        float3 colliderNormal = surfaceNormal;
        const float3 emission = f3(material.emission.x, material.emission.y, material.emission.z);

        if (fmaxcompf(emission) > PT_EPS) {                                   // kernels.h:563-576
            writeCache = false; writeState = false;
            if (!P.nee || state.fromSpecular) {
                state.accucolor += state.mask * emission;
                if (!state.albedoSet) { updateAlbedo(P.albedoSample, slot, state.mask * emission); state.albedoSet = true; }
                writeState = true;
            }
        } else {
            float3 diffuse = f3(material.diffuse_color.x, material.diffuse_color.y, material.diffuse_color.z);
            if (hit.type == 2) {                                              // kernels.h:578-582
                uint px = (uint)(fabsf(intersectionPos.x / 4 + 1000));
                uint py = (uint)(fabsf(intersectionPos.z / 4 + 1000));
                diffuse = (px + py) % 2 == 0 ? f3(1) : f3(0.2);
            }
            if (material.hasTexture || material.hasNormalMap) {               // kernels.h:586-619
                const pt_triangle_d td = sc.vertexData[hit.id];
                const pt_triangle_v tv = sc.vertices[hit.id];
                float t, u = 0, v = 0;
                const float3 nd = normalize(matmul(instance->invTransform, rd, 0.0f));
                const float3 no = matmul(instance->invTransform, ro, 1.0f);
                const float3 v0 = f3(tv.v0.x, tv.v0.y, tv.v0.z);
                triIntersect(no, nd, v0, f3(tv.v1.x, tv.v1.y, tv.v1.z) - v0, f3(tv.v2.x, tv.v2.y, tv.v2.z) - v0, t, u, v);
                const float uvx = td.uv0.x * (1 - u - v) + td.uv1.x * u + td.uv2.x * v;
                const float uvy = td.uv0.y * (1 - u - v) + td.uv1.y * u + td.uv2.y * v;
                if (material.hasTexture) diffuse = diffuse * f3(tex2D<float4>((cudaTextureObject_t)material.texture, uvx, uvy));
                if (material.hasNormalMap) {
                    const float3 tnT = (f3(tex2D<float4>((cudaTextureObject_t)material.normal_texture, uvx, uvy)) * 2.0f) - f3(1.0f);
                    float3 tn = f3(dot(tnT, f3(td.tangent.x, td.bitangent.x, td.normal.x)), dot(tnT, f3(td.tangent.y, td.bitangent.y, td.normal.y)),
                                   dot(tnT, f3(td.tangent.z, td.bitangent.z, td.normal.z)));
                    tn = normalize(matmul(instance->transform, tn, 0.0f));
                    colliderNormal = dot(tn, colliderNormal) < 0.0f ? -tn : tn;
                }
            }
            ShadeRay secondary; secondary.o = f3(0); secondary.d = f3(0); secondary.len = 9999999;
            const float random = rnd(rs);
            const float3 BRDF = diffuse / PT_PI;
            cache.sample_type = 0;
            float russianP = 1.0f;
            if (random < material.transmit) {                                 // kernels.h:633-653
                state.fromSpecular = true;
                if (inside) {
                    const float3 c = f3(material.absorption.x, material.absorption.y, material.absorption.z);
                    state.mask *= f3(expf(-c.x * hit.t), expf(-c.y * hit.t), expf(-c.z * hit.t));
                }
                float reflected;
                secondary = refractRay(rd, colliderNormal, intersectionPos, material.refractive_index, inside, reflected);
                if (rnd(rs) < reflected) { state.mask *= diffuse; secondary = reflectRay(rd, colliderNormal, intersectionPos); }
                const float ra = rnd(rs), rb = rnd(rs);                       // device call-argument order: left to right
                const float3 noiseDir = sampleHemisphereCosine(secondary.d, ra, rb);
                secondary.d = secondary.d * (1.0f - material.glossy) + material.glossy * noiseDir;
            } else if (random - material.transmit < material.reflect) {       // kernels.h:654-661
                state.fromSpecular = true;
                state.mask *= diffuse;
                secondary = reflectRay(rd, colliderNormal, intersectionPos);
                const float ra = rnd(rs), rb = rnd(rs);
                const float3 noiseDir = sampleHemisphereCosine(secondary.d, ra, rb);
                secondary.d = secondary.d * (1.0f - material.glossy) + material.glossy * noiseDir;
            } else {                                                          // kernels.h:662-782
                if (!state.albedoSet) { updateAlbedo(P.albedoSample, slot, state.mask * diffuse); state.albedoSet = true; }
                state.fromSpecular = false;
                // Skydome importance sampling (opt-in; the reference builds the CDF but no kernel samples it): a pixel owns one
                // shadow ray per bounce, so a vertex estimates EITHER the emissive triangles (below) OR the sky, chosen with
                // probability 1/2 when both exist, the estimate divided by that probability.
                bool lightNee = P.nee, skyNee = P.tex.skySample != 0u && P.tex.sky != 0;
                float pickScale = 1.0f;
                if (lightNee && skyNee) { pickScale = 2.0f; if (rnd(rs) < 0.5f) lightNee = false; else skyNee = false; }
                if (skyNee) {
                    const uint m = binarySearchCdf(P.tex.skyCdf, P.tex.skyW * P.tex.skyH, rnd(rs));
                    const uint tx = m % P.tex.skyW, ty = m / P.tex.skyW;
                    const float su = ((float)tx + rnd(rs)) / (float)P.tex.skyW, sv = ((float)ty + rnd(rs)) / (float)P.tex.skyH;
                    float sinPolar;
                    const float3 sdir = skyUvToDir(su, sv, sinPolar);
                    const float pdf = P.tex.skyPdf[m] * (float)(P.tex.skyW * P.tex.skyH) / (2.0f * PT_PI * PT_PI * fmaxf(sinPolar, 0.0001f));
                    const float NL = dot(colliderNormal, sdir);
                    if (NL > 0 && dot(sdir, surfaceNormal) > 0 && pdf > 0) {
                        const float2 uv = normalToUv(sdir);
                        const float3 sk = f3(tex2D<float4>(P.tex.sky, uv.x, uv.y));
                        state.light = state.mask * (BRDF * sk * (NL * pickScale / pdf));
                        const float f = NL * NL * NL;
                        out.pushShadow = true;
                        out.shadow.o = intersectionPos + PT_EPS * f * sdir + PT_EPS * (1 - f) * colliderNormal;
                        out.shadow.d = sdir;
                        out.shadow.len = 9999999;
                    }
                }
                if (lightNee) {                                               // kernels.h:672-753
                    uint successIdx = 0; float valid = 0.0f;
                    for (uint k = 0; k < 4; k++) {                            // kernels.h:676-690; per-light terms come from k_prepare_lights
                        const uint lightSource = uint(rnd(rs) * sc.num_lights) % sc.num_lights;
                        const float3 centroid = f3(sc.lightTab[lightSource * PT_LIGHT_STRIDE]);
                        const float3 lightNormal = f3(sc.lightTab[lightSource * PT_LIGHT_STRIDE + 1]);
                        const float3 fromLight = normalize(intersectionPos - centroid);
                        if (dot(lightNormal, fromLight) > 0) { valid += 1.0f; successIdx = lightSource; }
                    }
                    if (valid > 0) {
                        const float4* lt = sc.lightTab + successIdx * PT_LIGHT_STRIDE;
                        const float3 v0 = f3(lt[2]), v0v1 = f3(lt[3]), v0v2 = f3(lt[4]);
                        const float4 ln4 = lt[5];
                        const float3 lightNormal = f3(ln4);                   // cr / crLength
                        const float crLength = ln4.w;
                        float u = rnd(rs), v = rnd(rs);
                        if (u + v > 1.0f) { u = 1.0f - u; v = 1.0f - v; }
                        const float3 samplePoint = v0 + u * v0v1 + v * v0v2;
                        float3 shadowDir = intersectionPos - samplePoint;
                        const float shadowLength = length(shadowDir);
                        const float invShadowLength = 1.0f / shadowLength;
                        shadowDir *= invShadowLength;
                        const float NL = dot(colliderNormal, -shadowDir);
                        const float LNL = dot(lightNormal, shadowDir);
                        if (NL > 0 && dot(-shadowDir, surfaceNormal) > 0 && LNL > 0) {
                            const float4 em = lt[6];
                            const float A = 0.5f * crLength;
                            const float SA = LNL * A * invShadowLength * invShadowLength;
                            state.light = state.mask * (NL * SA * sc.num_lights * BRDF * f3(em.x, em.y, em.z) * (valid / 4.0f));
                            if (pickScale != 1.0f) state.light *= pickScale;
                            const float f = LNL * LNL * LNL;
                            out.pushShadow = true;
                            out.shadow.o = samplePoint + f * PT_EPS * shadowDir + (1 - f) * PT_EPS * lightNormal;
                            out.shadow.d = shadowDir;
                            out.shadow.len = shadowLength - 2 * PT_EPS;
                        }
                    }
                }
                float3 r;
                if (P.cache && hit.type == 0 && dot(colliderNormal, originalNormal) > 0) {   // kernels.h:757-766
                    const float4 rc0 = sc.rcRadiance[(size_t)hit.id * 3], rc1 = sc.rcRadiance[(size_t)hit.id * 3 + 1], rc2 = sc.rcRadiance[(size_t)hit.id * 3 + 2];
                    float invprob; int bucket;
                    r = sampleHemisphereCached(colliderNormal, rs, rc0, rc1, rc2.x, bucket, invprob);
                    cache.cache_bucket_id = bucket;
                    state.mask *= invprob;
                    cache.sample_type = 2;
                    cache.triangle_id = hit.id;
                    cache.cum_mask.x = state.mask.x; cache.cum_mask.y = state.mask.y; cache.cum_mask.z = state.mask.z;
                } else {
                    const float ra = rnd(rs), rb = rnd(rs);
                    r = sampleHemisphereCosine(colliderNormal, ra, rb);
                }
                if (dot(r, surfaceNormal) < 0) state.mask = f3(0.0f);
                float f = fmaxf(dot(colliderNormal, r), 0.0f);
                f = f * f * f;
                secondary.o = intersectionPos + PT_EPS * f * r + PT_EPS * (1 - f) * colliderNormal;
                secondary.d = r;
                state.mask *= PT_PI * BRDF;
                russianP = clampf(fmaxcompf(diffuse), 0.1f, 0.9f);
            }
            if (fmaxcompf(state.mask) > 0.0001f && rnd(rs) < russianP) {       // kernels.h:785-793
                state.mask = state.mask / russianP;
                out.pushExt = true; out.ext = secondary;
            } else cache.sample_type = 1;
        }
    }
    if (groupMask) __syncwarp(groupMask);
    if (writeState && writer) {                                                // setState, types.h:493-498
        P.masks[slot] = make_float4(state.mask.x, state.mask.y, state.mask.z, __uint_as_float(state.fromSpecular ? 1u : 0u));
        P.accucolors[slot] = make_float4(state.accucolor.x, state.accucolor.y, state.accucolor.z, __uint_as_float(state.albedoSet ? 1u : 0u));
        P.lights[slot] = make_float4(state.light.x, state.light.y, state.light.z, 0.0f);
    }
    if (writeCache && writer && P.cacheWrite && P.bounce < PT_MAX_CACHE_DEPTH) P.sampleCache[(size_t)P.bounce * P.map.nSlots + slot] = cache;
}

}  // namespace ptb
