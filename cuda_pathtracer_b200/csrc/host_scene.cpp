// host_scene.cpp -- host side of the scene surface the reference keeps in front of the kernels
// (reference src/scene.h, src/bvhBuilder.h, src/types.h Camera): the Scene container, the binned-SAH
// BVH2 builder whose output feeds the wide-BVH converter, the agglomerative top-level BVH, instance
// matrices, the camera frame and an OBJ/MTL reader.  Plain C++ (no CUDA): usable without a GPU.
// Everything that decides triangle ORDER or box values is arithmetic-identical to the reference so
// that triangle ids, BVH2 nodes and TLAS nodes can be compared bit for bit (tests/test_host_scene.py).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <chrono>
#include <map>
#include <string>
#include <vector>
#include "../../include/ptb200.h"
#include "../../include/ptb200_scene.h"

namespace {

typedef uint32_t uint;
const float kEps = 0.001f;    // constants.h:8

// SSE min/max semantics (_mm_min_ps(a,b) = a<b?a:b), which the reference's builder uses throughout.
inline float smin(float a, float b) { return a < b ? a : b; }
inline float smax(float a, float b) { return a > b ? a : b; }

struct V4 { float v[4]; };   // one SSE register; lane 3 is carried along like the reference does
inline V4 vmin4(const V4& a, const V4& b) { V4 r; for (int i = 0; i < 4; i++) r.v[i] = smin(a.v[i], b.v[i]); return r; }
inline V4 vmax4(const V4& a, const V4& b) { V4 r; for (int i = 0; i < 4; i++) r.v[i] = smax(a.v[i], b.v[i]); return r; }

struct Bounds {               // SSEBox, types.h:205-256
    V4 lo, hi;
    static Bounds empty() { Bounds b; b.lo = V4{{999999, 99999, 99999, 0}}; b.hi = V4{{-999999, -99999, -99999, 0}}; return b; }   // types.h:209-215 (sic)
    void add(const Bounds& o) { lo = vmin4(lo, o.lo); hi = vmax4(hi, o.hi); }
    void addPoint(const V4& p) { lo = vmin4(lo, p); hi = vmax4(hi, p); }
    float area() const {       // types.h:248-255
        float lx = hi.v[0] - lo.v[0], ly = hi.v[1] - lo.v[1], lz = hi.v[2] - lo.v[2];
        return 2 * lx * ly + 2 * lx * lz + 2 * ly * lz;
    }
};

inline void setNode(pt_bvh_node& n, const float* lo, const float* hi, uint a, uint b) {
    n.vmin.x = lo[0]; n.vmin.y = lo[1]; n.vmin.z = lo[2]; memcpy(&n.vmin.w, &a, 4);
    n.vmax.x = hi[0]; n.vmax.y = hi[1]; n.vmax.z = hi[2]; memcpy(&n.vmax.w, &b, 4);
}

// Binned SAH builder == createBVHBinned (bvhBuilder.h:35-274): 16 bins on the dominant centroid axis,
// leaf when count <= 4, when the centroid extent is < 16*EPS or when no split beats the leaf cost,
// children allocated adjacently, boxes assigned forward from the bin sweeps, triangles permuted at the end.
std::vector<pt_bvh_node> buildBinnedSAH(pt_triangle_v* tv, pt_triangle_d* td, uint n, uint triOffset) {
    const uint K = 16;
    std::vector<pt_bvh_node> nodes(n ? 2 * (size_t)n - 1 : 1);
    std::vector<uint> order(n);
    std::vector<Bounds> tb(n);
    std::vector<V4> cen(n);
    std::vector<signed char> binOf(n);
    Bounds root = Bounds::empty();
    for (uint i = 0; i < n; i++) {
        order[i] = i;
        const V4 a{{tv[i].v0.x, tv[i].v0.y, tv[i].v0.z, 0}}, b{{tv[i].v1.x, tv[i].v1.y, tv[i].v1.z, 0}}, c{{tv[i].v2.x, tv[i].v2.y, tv[i].v2.z, 0}};
        const float k[4] = {(float)0.333333, (float)0.3333333, (float)0.3333333, 0};     // bvhBuilder.h:57 (sic: x differs)
        for (int l = 0; l < 4; l++) cen[i].v[l] = k[l] * (a.v[l] + (b.v[l] + c.v[l]));
        tb[i].lo = vmin4(a, vmin4(b, c)); tb[i].hi = vmax4(a, vmax4(b, c));              // types.h:217-226
        root.add(tb[i]);
    }
    setNode(nodes[0], root.lo.v, root.hi.v, 0, 0);
    struct Job { uint node, start, count; };
    std::vector<Job> jobs; jobs.push_back({0, 0, n});
    uint used = 1;
    float costL[K], costR[K]; uint cnt[K]; Bounds bins[K], accL[K], accR[K];
    while (!jobs.empty()) {
        const Job j = jobs.back(); jobs.pop_back();
        const float plo[3] = {nodes[j.node].vmin.x, nodes[j.node].vmin.y, nodes[j.node].vmin.z};
        const float phi[3] = {nodes[j.node].vmax.x, nodes[j.node].vmax.y, nodes[j.node].vmax.z};
        auto makeLeaf = [&]() { setNode(nodes[j.node], plo, phi, j.start + triOffset, j.count); };
        if (j.count <= 4) { makeLeaf(); continue; }
        const float px = phi[0] - plo[0], py = phi[1] - plo[1], pz = phi[2] - plo[2];      // Box::getSurfaceArea, types.h:154-161
        const float invParent = 1.0f / (2 * px * py + 2 * px * pz + 2 * py * pz);
        Bounds cb = Bounds::empty();
        for (uint i = j.start; i < j.start + j.count; i++) cb.addPoint(cen[order[i]]);
        const float d0 = cb.hi.v[0] - cb.lo.v[0], d1 = cb.hi.v[1] - cb.lo.v[1], d2 = cb.hi.v[2] - cb.lo.v[2];
        const int axis = (d0 > d1 && d0 > d2) ? 0 : ((d1 > d0 && d1 > d2) ? 1 : 2);
        const float bmin = cb.lo.v[axis], bmax = cb.hi.v[axis];
        if (bmax - bmin < K * kEps) { makeLeaf(); continue; }
        for (uint k = 0; k < K; k++) { cnt[k] = 0; bins[k] = Bounds::empty(); }
        const float binFac = K * (1 - kEps) / (bmax - bmin);
        for (uint i = j.start; i < j.start + j.count; i++) {
            const uint t = order[i];
            const signed char b = (signed char)((cen[t].v[axis] - bmin) * binFac);
            binOf[t] = b; cnt[(int)b]++; bins[(int)b].add(tb[t]);
        }
        Bounds L = Bounds::empty(), R = Bounds::empty();
        uint nl = 0, nr = 0;
        for (uint k = 0; k < K; k++) {
            costL[k] = nl * L.area() * invParent; accL[k] = L;                            // left exclusive
            L.add(bins[k]); nl += cnt[k];
            R.add(bins[K - k - 1]); nr += cnt[K - k - 1];                                 // right inclusive
            costR[K - k - 1] = nr * R.area() * invParent; accR[K - k - 1] = R;
        }
        float best = (float)j.count; int split = -1;
        for (uint k = 0; k < K; k++) { const float c = costL[k] + costR[k] + kEps; if (c < best) { best = c; split = (int)k; } }
        if (split < 0) { makeLeaf(); continue; }
        // two-pointer partition, bvhBuilder.h:197-220 (the exact swap sequence decides triangle order)
        uint* pl = order.data() + j.start; uint* pr = order.data() + j.start + j.count - 1;
        int state = 0;
        while (pl != pr) {
            if (state == 0) { if (binOf[*pl] < split) { pl++; continue; } else state = 1; }
            if (state == 1) { if (binOf[*pr] >= split) { pr--; continue; } }
            std::swap(*pl, *pr); state = 0;
        }
        const uint leftCount = (uint)(pl - (order.data() + j.start));
        const uint c1 = used++, c2 = used++;
        setNode(nodes[c1], accL[split].lo.v, accL[split].hi.v, 0, 0);
        setNode(nodes[c2], accR[split].lo.v, accR[split].hi.v, 0, 0);
        jobs.push_back({c2, j.start + leftCount, j.count - leftCount});
        jobs.push_back({c1, j.start, leftCount});
        setNode(nodes[j.node], plo, phi, c1, 0);
    }
    std::vector<pt_triangle_v> tv2(tv, tv + n); std::vector<pt_triangle_d> td2(td, td + n);   // permuteTriangles, bvhBuilder.h:17-27
    for (uint i = 0; i < n; i++) { tv[i] = tv2[order[i]]; td[i] = td2[order[i]]; }
    nodes.resize(used);
    return nodes;
}

// ---- 4x4 matrices --------------------------------------------------------------------------------------
// Column-major m[c][r].  Two families are restated because the reference uses two libraries:
//  * linmath.h (vendored, src/linmath.h:282-391) for the bake-in transform of addModel (scene.h:176)
//  * glm (un-vendored, version unpinned) for instance matrices (scene.h:9-25): translate, rotate, scale,
//    inverse restated from glm's published scalar algorithms; parity is UNPINNED at this boundary.
struct M4 { float m[4][4]; };
M4 identity() { M4 r; for (int c = 0; c < 4; c++) for (int k = 0; k < 4; k++) r.m[c][k] = c == k ? 1.f : 0.f; return r; }
M4 lmMul(const M4& a, const M4& b) {       // mat4x4_mul, linmath.h:282-291 (accumulates from 0.f)
    M4 t;
    for (int c = 0; c < 4; c++) for (int r = 0; r < 4; r++) {
        t.m[c][r] = 0.f;
        for (int k = 0; k < 4; k++) t.m[c][r] += a.m[k][r] * b.m[c][k];
    }
    return t;
}
void lmMulVec(float r[4], const M4& M, const float v[4]) {     // mat4x4_mul_vec4, linmath.h:292-299
    for (int j = 0; j < 4; j++) { r[j] = 0.f; for (int i = 0; i < 4; i++) r[j] += M.m[i][j] * v[i]; }
}
M4 lmRot(int axis, float angle) {          // mat4x4_rotate_X/Y/Z on identity, linmath.h:357-391
    const float s = sinf(angle), c = cosf(angle);
    M4 R = identity();
    if (axis == 0) { R.m[1][1] = c; R.m[1][2] = s; R.m[2][1] = -s; R.m[2][2] = c; }
    else if (axis == 1) { R.m[0][0] = c; R.m[0][2] = s; R.m[2][0] = -s; R.m[2][2] = c; }
    else { R.m[0][0] = c; R.m[0][1] = s; R.m[1][0] = -s; R.m[1][1] = c; }
    return lmMul(identity(), R);
}
M4 bakeTransform(float scale, const float rot[3], const float off[3]) {   // scene.h:176 + vec.h Matrix4 helpers
    M4 T = identity(); T.m[3][0] = off[0]; T.m[3][1] = off[1]; T.m[3][2] = off[2];
    M4 S = identity(); for (int c = 0; c < 3; c++) for (int k = 0; k < 4; k++) S.m[c][k] = S.m[c][k] * scale;
    M4 R = lmMul(lmMul(lmRot(0, rot[0]), lmRot(1, rot[1])), lmRot(2, rot[2]));
    return lmMul(lmMul(T, S), R);
}

struct C4 { float v[4]; };
inline C4 col(const M4& m, int c) { C4 r; for (int k = 0; k < 4; k++) r.v[k] = m.m[c][k]; return r; }
inline C4 operator*(const C4& a, float s) { C4 r; for (int k = 0; k < 4; k++) r.v[k] = a.v[k] * s; return r; }
inline C4 operator*(const C4& a, const C4& b) { C4 r; for (int k = 0; k < 4; k++) r.v[k] = a.v[k] * b.v[k]; return r; }
inline C4 operator+(const C4& a, const C4& b) { C4 r; for (int k = 0; k < 4; k++) r.v[k] = a.v[k] + b.v[k]; return r; }
inline C4 operator-(const C4& a, const C4& b) { C4 r; for (int k = 0; k < 4; k++) r.v[k] = a.v[k] - b.v[k]; return r; }
inline void setCol(M4& m, int c, const C4& v) { for (int k = 0; k < 4; k++) m.m[c][k] = v.v[k]; }

M4 glmTranslate(const M4& m, const float v[3]) {
    M4 r = m; setCol(r, 3, col(m, 0) * v[0] + col(m, 1) * v[1] + col(m, 2) * v[2] + col(m, 3)); return r;
}
M4 glmRotate(const M4& m, float angle, const float axisIn[3]) {
    const float c = std::cos(angle), s = std::sin(angle);
    const float il = 1.0f / std::sqrt(axisIn[0] * axisIn[0] + axisIn[1] * axisIn[1] + axisIn[2] * axisIn[2]);
    const float axis[3] = {axisIn[0] * il, axisIn[1] * il, axisIn[2] * il};
    const float temp[3] = {(1.0f - c) * axis[0], (1.0f - c) * axis[1], (1.0f - c) * axis[2]};
    float R[3][3];
    R[0][0] = c + temp[0] * axis[0]; R[0][1] = temp[0] * axis[1] + s * axis[2]; R[0][2] = temp[0] * axis[2] - s * axis[1];
    R[1][0] = temp[1] * axis[0] - s * axis[2]; R[1][1] = c + temp[1] * axis[1]; R[1][2] = temp[1] * axis[2] + s * axis[0];
    R[2][0] = temp[2] * axis[0] + s * axis[1]; R[2][1] = temp[2] * axis[1] - s * axis[0]; R[2][2] = c + temp[2] * axis[2];
    M4 r;
    setCol(r, 0, col(m, 0) * R[0][0] + col(m, 1) * R[0][1] + col(m, 2) * R[0][2]);
    setCol(r, 1, col(m, 0) * R[1][0] + col(m, 1) * R[1][1] + col(m, 2) * R[1][2]);
    setCol(r, 2, col(m, 0) * R[2][0] + col(m, 1) * R[2][1] + col(m, 2) * R[2][2]);
    setCol(r, 3, col(m, 3));
    return r;
}
M4 glmScale(const M4& m, const float v[3]) {
    M4 r; setCol(r, 0, col(m, 0) * v[0]); setCol(r, 1, col(m, 1) * v[1]); setCol(r, 2, col(m, 2) * v[2]); setCol(r, 3, col(m, 3)); return r;
}
M4 glmInverse(const M4& mm) {
    const float (*m)[4] = mm.m;
    const float c00 = m[2][2] * m[3][3] - m[3][2] * m[2][3], c02 = m[1][2] * m[3][3] - m[3][2] * m[1][3], c03 = m[1][2] * m[2][3] - m[2][2] * m[1][3];
    const float c04 = m[2][1] * m[3][3] - m[3][1] * m[2][3], c06 = m[1][1] * m[3][3] - m[3][1] * m[1][3], c07 = m[1][1] * m[2][3] - m[2][1] * m[1][3];
    const float c08 = m[2][1] * m[3][2] - m[3][1] * m[2][2], c10 = m[1][1] * m[3][2] - m[3][1] * m[1][2], c11 = m[1][1] * m[2][2] - m[2][1] * m[1][2];
    const float c12 = m[2][0] * m[3][3] - m[3][0] * m[2][3], c14 = m[1][0] * m[3][3] - m[3][0] * m[1][3], c15 = m[1][0] * m[2][3] - m[2][0] * m[1][3];
    const float c16 = m[2][0] * m[3][2] - m[3][0] * m[2][2], c18 = m[1][0] * m[3][2] - m[3][0] * m[1][2], c19 = m[1][0] * m[2][2] - m[2][0] * m[1][2];
    const float c20 = m[2][0] * m[3][1] - m[3][0] * m[2][1], c22 = m[1][0] * m[3][1] - m[3][0] * m[1][1], c23 = m[1][0] * m[2][1] - m[2][0] * m[1][1];
    const C4 F0{{c00, c00, c02, c03}}, F1{{c04, c04, c06, c07}}, F2{{c08, c08, c10, c11}}, F3{{c12, c12, c14, c15}}, F4{{c16, c16, c18, c19}}, F5{{c20, c20, c22, c23}};
    const C4 V0{{m[1][0], m[0][0], m[0][0], m[0][0]}}, V1{{m[1][1], m[0][1], m[0][1], m[0][1]}}, V2{{m[1][2], m[0][2], m[0][2], m[0][2]}}, V3{{m[1][3], m[0][3], m[0][3], m[0][3]}};
    const C4 I0 = V1 * F0 - V2 * F1 + V3 * F2, I1 = V0 * F0 - V2 * F3 + V3 * F4, I2 = V0 * F1 - V1 * F3 + V3 * F5, I3 = V0 * F2 - V1 * F4 + V2 * F5;
    const C4 SA{{+1, -1, +1, -1}}, SB{{-1, +1, -1, +1}};
    M4 inv; setCol(inv, 0, I0 * SA); setCol(inv, 1, I1 * SB); setCol(inv, 2, I2 * SA); setCol(inv, 3, I3 * SB);
    const C4 row0{{inv.m[0][0], inv.m[1][0], inv.m[2][0], inv.m[3][0]}};
    const C4 d0 = col(mm, 0) * row0;
    const float det = (d0.v[0] + d0.v[1]) + (d0.v[2] + d0.v[3]);
    const float ood = 1.0f / det;
    M4 r; for (int c = 0; c < 4; c++) setCol(r, c, col(inv, c) * ood);
    return r;
}
pt_mat4x3 toMat4x3(const M4& m) {           // mat4x3::fromGLM, types.h:391-399
    pt_mat4x3 r;
    r.r1 = pt_float4{m.m[0][0], m.m[1][0], m.m[2][0], m.m[3][0]};
    r.r2 = pt_float4{m.m[0][1], m.m[1][1], m.m[2][1], m.m[3][1]};
    r.r3 = pt_float4{m.m[0][2], m.m[1][2], m.m[2][2], m.m[3][2]};
    return r;
}

struct GameObject { uint model_id, material_id; float pos[3], rot[3], scl[3]; };   // types.h:416-429

pt_instance convertToInstance(const GameObject& o) {   // ConvertToInstance, scene.h:9-25
    M4 t = identity();
    t = glmTranslate(t, o.pos);
    const float X[3] = {1, 0, 0}, Y[3] = {0, 1, 0}, Z[3] = {0, 0, 1};
    t = glmRotate(t, o.rot[0], X); t = glmRotate(t, o.rot[1], Y); t = glmRotate(t, o.rot[2], Z);
    t = glmScale(t, o.scl);
    pt_instance inst; memset(&inst, 0, sizeof(inst));
    inst.model_id = o.model_id; inst.transform = toMat4x3(t); inst.invTransform = toMat4x3(glmInverse(t)); inst.material_id = o.material_id;
    return inst;
}

// ---- top-level BVH: BuildTopLevelBVH / FindBestMatch / transformBox, scene.h:27-118 ------------------------------
struct Box { float lo[3], hi[3]; };
inline Box insideOut() { Box b; for (int a = 0; a < 3; a++) { b.lo[a] = 9999999; b.hi[a] = -9999999; } return b; }   // types.h:170-176
inline void consume(Box& b, const Box& o) { for (int a = 0; a < 3; a++) { b.lo[a] = fminf(b.lo[a], o.lo[a]); b.hi[a] = fmaxf(b.hi[a], o.hi[a]); } }
inline float boxArea(const Box& b) { float lx = b.hi[0] - b.lo[0], ly = b.hi[1] - b.lo[1], lz = b.hi[2] - b.lo[2]; return 2 * lx * ly + 2 * lx * lz + 2 * ly * lz; }
inline float dot4(const float p[3], float w, const pt_float4& r) { return p[0] * r.x + p[1] * r.y + p[2] * r.z + w * r.w; }
Box transformBox(const Box& box, const pt_mat4x3& t) {
    float pts[8][3];
    const float d[3] = {box.hi[0] - box.lo[0], box.hi[1] - box.lo[1], box.hi[2] - box.lo[2]};
    for (int a = 0; a < 3; a++) { pts[0][a] = box.lo[a]; pts[4][a] = box.hi[a]; }
    for (int k = 0; k < 3; k++) for (int a = 0; a < 3; a++) {
        pts[1 + k][a] = box.lo[a] + (a == k ? d[a] : 0.0f);
        pts[5 + k][a] = box.hi[a] - (a == k ? d[a] : 0.0f);
    }
    Box r = insideOut();
    for (int i = 0; i < 8; i++) {
        const float wp[3] = {dot4(pts[i], 1.0f, t.r1), dot4(pts[i], 1.0f, t.r2), dot4(pts[i], 1.0f, t.r3)};
        for (int a = 0; a < 3; a++) { r.lo[a] = fminf(wp[a], r.lo[a]); r.hi[a] = fmaxf(wp[a], r.hi[a]); }
    }
    return r;
}
struct Entry { uint index; Box box; };
Entry bestMatch(uint node, const std::map<uint, Entry>& list) {
    float minSurface = 3.402823466e+38f; bool found = false; Entry match{};
    const Box nb = list.at(node).box;
    for (const auto& kv : list) {
        if (kv.second.index == node) continue;
        Box merged = nb; consume(merged, kv.second.box);
        const float a = boxArea(merged);
        if (a < minSurface) { minSurface = a; match = kv.second; found = true; }
    }
    return found ? match : list.at(node);
}
void setTlas(pt_tlas_node& n, const Box& b, uint children, uint leaf) {
    n.vmin = pt_float3{b.lo[0], b.lo[1], b.lo[2]}; n.vmax = pt_float3{b.hi[0], b.hi[1], b.hi[2]}; n.children = children; n.leaf = leaf;
}
void buildTopLevel(pt_tlas_node* dest, const pt_instance* inst, const std::vector<std::vector<pt_bvh_node>>& bvhs, uint n) {
    if (n == 0) return;
    std::map<uint, Entry> list;
    uint nodeCount = 2 * n - 1;
    for (uint i = 0; i < n; i++) {
        nodeCount--;
        const pt_bvh_node& rootNode = bvhs[inst[i].model_id][0];
        Box mb; mb.lo[0] = rootNode.vmin.x; mb.lo[1] = rootNode.vmin.y; mb.lo[2] = rootNode.vmin.z; mb.hi[0] = rootNode.vmax.x; mb.hi[1] = rootNode.vmax.y; mb.hi[2] = rootNode.vmax.z;
        const Box b = transformBox(mb, inst[i].transform);
        setTlas(dest[nodeCount], b, 0, i);
        list[nodeCount] = Entry{nodeCount, b};
    }
    Entry A = list.begin()->second;
    Entry B = bestMatch(A.index, list);
    while (list.size() > 1) {
        const Entry Cc = bestMatch(B.index, list);
        if (A.index == Cc.index) {
            list.erase(A.index); list.erase(B.index);
            Box m = insideOut(); consume(m, A.box); consume(m, B.box);          // Box::merged, types.h:178-184
            const uint idx = --nodeCount;
            setTlas(dest[idx], m, A.index | (B.index << 16), dest[idx].leaf);
            A = Entry{idx, m};
            list[idx] = A;
            B = bestMatch(A.index, list);
        } else { A = B; B = Cc; }
    }
}

// ---- camera frame: Camera ctor + update() with no key pressed + recalculate(), types.h:590-637 ---------------------
inline float hostRsqrt(float a) { return (float)(1.0 / sqrt((double)a)); }      // CUDA host rsqrtf, crt/math_functions.hpp:3333
inline void norm3(float v[3]) { const float s = hostRsqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); for (int a = 0; a < 3; a++) v[a] = s * v[a]; }
inline void cross3(const float a[3], const float b[3], float r[3]) { r[0] = a[1] * b[2] - a[2] * b[1]; r[1] = a[2] * b[0] - a[0] * b[2]; r[2] = a[0] * b[1] - a[1] * b[0]; }

struct SceneImpl {
    std::vector<pt_triangle_v> verts;
    std::vector<pt_triangle_d> vdata;
    std::vector<pt_material> materials;
    std::vector<std::string> texNames, normalTexNames;      // per material (OBJ/MTL path), "" when none
    std::vector<pt_model> models;
    std::vector<std::vector<pt_bvh_node>> bvhs;
    std::vector<GameObject> objects;
    std::vector<pt_instance> instances;
    std::vector<pt_tlas_node> tlas;
    std::vector<pt_sphere> spheres;
    std::vector<pt_plane> planes;
    std::string err;
    double lastBuildMs = 0;
    int buildDevice = -1;       // >= 0: BVH2 built by pt_build_bvh on that GPU
    bool finalized = false;
};

double nowMs() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// tail of Scene::addModel (scene.h:255-346): the triangles are already appended at [start, start+n)
int finishModel(SceneImpl& s, uint start, uint n) {
    pt_model m; memset(&m, 0, sizeof(m));
    m.triangleStart = start; m.nrTriangles = n;
    const double t0 = nowMs();
    if (s.buildDevice >= 0) {
        std::vector<pt_bvh_node> nodes(n ? 2 * (size_t)n - 1 : 1);
        std::vector<uint> order(n);
        uint used = 0;
        const int r = pt_build_bvh(s.buildDevice, s.verts.data() + start, n, start, nodes.data(), &used, order.data(), nullptr, nullptr);
        if (r != PT_OK) { s.err = std::string("device BVH build failed: ") + pt_build_bvh_last_error(); s.verts.resize(start); s.vdata.resize(start); return r; }
        std::vector<pt_triangle_v> tv2(s.verts.begin() + start, s.verts.end()); std::vector<pt_triangle_d> td2(s.vdata.begin() + start, s.vdata.end());
        for (uint i = 0; i < n; i++) { s.verts[start + i] = tv2[order[i]]; s.vdata[start + i] = td2[order[i]]; }    // permuteTriangles, bvhBuilder.h:17-27
        nodes.resize(used);
        s.bvhs.push_back(std::move(nodes));
    } else {
        s.bvhs.push_back(buildBinnedSAH(s.verts.data() + start, s.vdata.data() + start, n, start));
    }
    s.lastBuildMs = nowMs() - t0;
    m.nrBvhNodes = (uint)s.bvhs.back().size();
    s.models.push_back(m);
    for (size_t i = 0; i < s.models.size(); i++) s.models[i].bvh = s.bvhs[i].data();
    return (int)s.models.size() - 1;
}

// ---- OBJ / MTL reader ------------------------------------------------------------------------------------
// Own reader for the subset of Wavefront OBJ/MTL that Scene::addModel consumes through tinyobjloader
// 2.0.0 (scene.h:159-347): v / vt / vn / f (all index forms, negative = relative), usemtl, mtllib;
// MTL: newmtl, Kd, Ks, d, Tr, Ns, Ni, map_Kd (with -o), norm.  Faces come out in file order; polygons
// with more than three vertices are ear-clipped in the projection plane chosen from the first
// non-degenerate corner, the way tinyobjloader's triangulation does, so that triangle order (and with it
// every triangle id) matches the reference for the assets it ships (teapot, cube, house).
struct ObjIndex { int v, vt, vn; };
struct MtlEntry { std::string name; float Kd[3] = {0, 0, 0}, Ks[3] = {0, 0, 0}; float d = 1, Ns = 1, Ni = 1; std::string mapKd, norm; float off[2] = {0, 0}; };

inline const char* skipWs(const char* p) { while (*p == ' ' || *p == '\t') p++; return p; }
inline float parseF(const char*& p) { p = skipWs(p); char* e; double d = strtod(p, &e); if (e == p) { return 0.0f; } p = e; return (float)d; }
inline bool fixIdx(int idx, int n, int& out) { if (idx > 0) { out = idx - 1; return true; } if (idx == 0) return false; out = n + idx; return true; }
bool parseIndex(const char*& p, int nv, int nvt, int nvn, ObjIndex& o) {
    o.v = o.vt = o.vn = -1;
    if (!fixIdx(atoi(p), nv, o.v)) return false;
    p += strcspn(p, "/ \t\r");
    if (*p != '/') return true;
    p++;
    if (*p == '/') { p++; if (!fixIdx(atoi(p), nvn, o.vn)) return false; p += strcspn(p, "/ \t\r"); return true; }
    if (!fixIdx(atoi(p), nvt, o.vt)) return false;
    p += strcspn(p, "/ \t\r");
    if (*p != '/') return true;
    p++;
    if (!fixIdx(atoi(p), nvn, o.vn)) return false;
    p += strcspn(p, "/ \t\r");
    return true;
}
inline std::string restOfLine(const char* p) { p = skipWs(p); std::string s(p); while (!s.empty() && (s.back() == '\r' || s.back() == '\n' || s.back() == ' ' || s.back() == '\t')) s.pop_back(); return s; }
inline std::string dirOf(const std::string& path) { size_t k = path.find_last_of("/\\"); return k == std::string::npos ? std::string() : path.substr(0, k + 1); }

bool loadMtl(const std::string& file, std::vector<MtlEntry>& out, std::map<std::string, int>& byName) {
    FILE* f = fopen(file.c_str(), "r");
    if (!f) return false;
    char line[4096];
    MtlEntry cur; bool have = false;
    auto flush = [&]() { if (have) { byName[cur.name] = (int)out.size(); out.push_back(cur); } };
    while (fgets(line, sizeof(line), f)) {
        const char* p = skipWs(line);
        if (*p == '#' || *p == '\n' || *p == '\r' || *p == 0) continue;
        if (!strncmp(p, "newmtl", 6) && (p[6] == ' ' || p[6] == '\t')) { flush(); cur = MtlEntry(); cur.name = restOfLine(p + 7); have = true; continue; }
        if (!strncmp(p, "Kd", 2) && (p[2] == ' ' || p[2] == '\t')) { p += 2; for (int i = 0; i < 3; i++) cur.Kd[i] = parseF(p); continue; }
        if (!strncmp(p, "Ks", 2) && (p[2] == ' ' || p[2] == '\t')) { p += 2; for (int i = 0; i < 3; i++) cur.Ks[i] = parseF(p); continue; }
        if (!strncmp(p, "Ni", 2) && (p[2] == ' ' || p[2] == '\t')) { p += 2; cur.Ni = parseF(p); continue; }
        if (!strncmp(p, "Ns", 2) && (p[2] == ' ' || p[2] == '\t')) { p += 2; cur.Ns = parseF(p); continue; }
        if (p[0] == 'd' && (p[1] == ' ' || p[1] == '\t')) { p += 1; cur.d = parseF(p); continue; }
        if (!strncmp(p, "Tr", 2) && (p[2] == ' ' || p[2] == '\t')) { p += 2; cur.d = 1.0f - parseF(p); continue; }
        const bool isKd = !strncmp(p, "map_Kd", 6) && (p[6] == ' ' || p[6] == '\t');
        const bool isNorm = !strncmp(p, "norm", 4) && (p[4] == ' ' || p[4] == '\t');
        if (isKd || isNorm) {
            p += isKd ? 6 : 4;
            std::string name;
            while (true) {                                   // texture options: only -o (origin offset) is consumed by the reference
                p = skipWs(p);
                if (!strncmp(p, "-o", 2) && (p[2] == ' ' || p[2] == '\t')) { p += 2; float o0 = parseF(p), o1 = parseF(p); parseF(p); if (isKd) { cur.off[0] = o0; cur.off[1] = o1; } continue; }
                name = restOfLine(p); break;
            }
            if (isKd) cur.mapKd = name; else cur.norm = name;
            continue;
        }
    }
    flush();
    fclose(f);
    return true;
}
// point-in-polygon (crossing number) used by the ear test, as in tinyobjloader's pnpoly
inline bool pnpoly(int nvert, const float* vx, const float* vy, float tx, float ty) {
    bool c = false;
    for (int i = 0, j = nvert - 1; i < nvert; j = i++)
        if (((vy[i] > ty) != (vy[j] > ty)) && (tx < (vx[j] - vx[i]) * (ty - vy[i]) / (vy[j] - vy[i]) + vx[i])) c = !c;
    return c;
}
// Polygon -> triangles (tiny_obj_loader.h:1394-1598).  pos = flat xyz array.
void triangulateFace(const std::vector<ObjIndex>& face, const std::vector<float>& pos, std::vector<ObjIndex>& out) {
    size_t n = face.size();
    if (n < 3) return;
    if (n == 3) { out.push_back(face[0]); out.push_back(face[1]); out.push_back(face[2]); return; }
    size_t axes[2] = {1, 2};
    for (size_t k = 0; k < n; ++k) {
        const size_t a = (size_t)face[k % n].v, b = (size_t)face[(k + 1) % n].v, c = (size_t)face[(k + 2) % n].v;
        if (3 * a + 2 >= pos.size() || 3 * b + 2 >= pos.size() || 3 * c + 2 >= pos.size()) continue;
        const float e0x = pos[b * 3] - pos[a * 3], e0y = pos[b * 3 + 1] - pos[a * 3 + 1], e0z = pos[b * 3 + 2] - pos[a * 3 + 2];
        const float e1x = pos[c * 3] - pos[b * 3], e1y = pos[c * 3 + 1] - pos[b * 3 + 1], e1z = pos[c * 3 + 2] - pos[b * 3 + 2];
        const float cx = std::fabs(e0y * e1z - e0z * e1y), cy = std::fabs(e0z * e1x - e0x * e1z), cz = std::fabs(e0x * e1y - e0y * e1x);
        const float eps = 1.1920929e-07f;
        if (cx > eps || cy > eps || cz > eps) {
            if (cx > cy && cx > cz) {} else { axes[0] = 0; if (cz > cx && cz > cy) axes[1] = 1; }
            break;
        }
    }
    float area = 0;
    for (size_t k = 0; k < n; ++k) {
        const size_t a = (size_t)face[k % n].v, b = (size_t)face[(k + 1) % n].v;
        if (a * 3 + axes[0] >= pos.size() || a * 3 + axes[1] >= pos.size() || b * 3 + axes[0] >= pos.size() || b * 3 + axes[1] >= pos.size()) continue;
        area += (pos[a * 3 + axes[0]] * pos[b * 3 + axes[1]] - pos[a * 3 + axes[1]] * pos[b * 3 + axes[0]]) * 0.5f;
    }
    std::vector<ObjIndex> rem = face;
    size_t guess = 0, remainingIterations = face.size(), prevCount = rem.size();
    ObjIndex ind[3]; float vx[3], vy[3];
    while (rem.size() > 3 && remainingIterations > 0) {
        n = rem.size();
        if (guess >= n) guess -= n;
        if (prevCount != n) { prevCount = n; remainingIterations = n; } else remainingIterations--;
        for (size_t k = 0; k < 3; k++) {
            ind[k] = rem[(guess + k) % n];
            const size_t vi = (size_t)ind[k].v;
            if (vi * 3 + axes[0] >= pos.size() || vi * 3 + axes[1] >= pos.size()) { vx[k] = 0; vy[k] = 0; }
            else { vx[k] = pos[vi * 3 + axes[0]]; vy[k] = pos[vi * 3 + axes[1]]; }
        }
        const float e0x = vx[1] - vx[0], e0y = vy[1] - vy[0], e1x = vx[2] - vx[1], e1y = vy[2] - vy[1];
        const float cr = e0x * e1y - e0y * e1x;
        if (cr * area < 0.0f) { guess += 1; continue; }
        bool overlap = false;
        for (size_t other = 3; other < n; ++other) {
            const size_t idx = (guess + other) % n;
            if (idx >= rem.size()) continue;
            const size_t ovi = (size_t)rem[idx].v;
            if (ovi * 3 + axes[0] >= pos.size() || ovi * 3 + axes[1] >= pos.size()) continue;
            if (pnpoly(3, vx, vy, pos[ovi * 3 + axes[0]], pos[ovi * 3 + axes[1]])) { overlap = true; break; }
        }
        if (overlap) { guess += 1; continue; }
        out.push_back(ind[0]); out.push_back(ind[1]); out.push_back(ind[2]);
        size_t removed = (guess + 1) % n;
        while (removed + 1 < n) { rem[removed] = rem[removed + 1]; removed += 1; }
        rem.pop_back();
    }
    if (rem.size() == 3) { out.push_back(rem[0]); out.push_back(rem[1]); out.push_back(rem[2]); }
}
inline float clamp01(float f) { return fmaxf(0.0f, fminf(f, 1.0f)); }

// Scene::addModel, scene.h:159-347
int loadObjModel(SceneImpl& s, const char* path, float scale, const float* rot, const float* off, uint material, bool useMtl) {
    FILE* f = fopen(path, "r");
    if (!f) { s.err = std::string("cannot open OBJ file ") + path; return PT_ERR_IO; }
    std::vector<float> pos, nrm, tex;
    struct Face { std::vector<ObjIndex> idx; int mtl; };
    std::vector<Face> faces;
    std::vector<MtlEntry> mtls; std::map<std::string, int> mtlByName;
    int curMtl = -1;
    std::vector<char> lineBuf(1 << 16);
    while (fgets(lineBuf.data(), (int)lineBuf.size(), f)) {
        const char* p = skipWs(lineBuf.data());
        if (p[0] == 'v' && (p[1] == ' ' || p[1] == '\t')) { p += 2; for (int i = 0; i < 3; i++) pos.push_back(parseF(p)); continue; }
        if (p[0] == 'v' && p[1] == 'n' && (p[2] == ' ' || p[2] == '\t')) { p += 3; for (int i = 0; i < 3; i++) nrm.push_back(parseF(p)); continue; }
        if (p[0] == 'v' && p[1] == 't' && (p[2] == ' ' || p[2] == '\t')) { p += 3; for (int i = 0; i < 2; i++) tex.push_back(parseF(p)); continue; }
        if (p[0] == 'f' && (p[1] == ' ' || p[1] == '\t')) {
            p += 2;
            Face fc; fc.mtl = curMtl;
            while (true) {
                p = skipWs(p);
                if (*p == 0 || *p == '\n' || *p == '\r') break;
                ObjIndex oi;
                if (!parseIndex(p, (int)pos.size() / 3, (int)tex.size() / 2, (int)nrm.size() / 3, oi)) { fclose(f); s.err = "bad face index in OBJ"; return PT_ERR_IO; }
                fc.idx.push_back(oi);
            }
            faces.push_back(fc);
            continue;
        }
        if (!strncmp(p, "usemtl", 6) && (p[6] == ' ' || p[6] == '\t')) {
            auto it = mtlByName.find(restOfLine(p + 7));
            curMtl = it == mtlByName.end() ? -1 : it->second;
            continue;
        }
        if (!strncmp(p, "mtllib", 6) && (p[6] == ' ' || p[6] == '\t')) { loadMtl(dirOf(path) + restOfLine(p + 7), mtls, mtlByName); continue; }
    }
    fclose(f);
    {   // every index must name an element the file defines (positive indices may point forward; relative ones may not
        // reach before the first element): pos / tex / nrm are indexed directly below
        const int nv = (int)pos.size() / 3, nvt = (int)tex.size() / 2, nvn = (int)nrm.size() / 3;
        for (const Face& fc : faces)
            for (const ObjIndex& oi : fc.idx)
                if (oi.v < 0 || oi.v >= nv || oi.vt < -1 || oi.vt >= nvt || oi.vn < -1 || oi.vn >= nvn) { s.err = std::string("face index out of range in OBJ file ") + path; return PT_ERR_IO; }
    }

    std::vector<uint> matIds(mtls.size(), 0);
    if (useMtl) {                                              // scene.h:180-248
        for (size_t m = 0; m < mtls.size(); m++) {
            const MtlEntry& e = mtls[m];
            pt_material mat; memset(&mat, 0, sizeof(mat));
            mat.diffuse_color = pt_float3{clamp01(e.Kd[0]), clamp01(e.Kd[1]), clamp01(e.Kd[2])};
            mat.specular_color = pt_float3{e.Ks[0], e.Ks[1], e.Ks[2]};
            mat.transmit = 1 - e.d;
            mat.reflect = (e.Ks[0] + e.Ks[1] + e.Ks[2]) / 3.0f;
            mat.glossy = e.Ns / 4000.0f;
            const float sum = mat.transmit + mat.reflect;
            if (sum > 1) { const float k = 1.0f / sum; mat.transmit *= k; mat.reflect *= k; }
            if (mat.transmit > kEps) mat.diffuse_color = pt_float3{1, 1, 1};
            mat.refractive_index = e.Ni;
            mat.hasTexture = !e.mapKd.empty(); mat.hasNormalMap = !e.norm.empty();
            s.materials.push_back(mat); s.texNames.push_back(e.mapKd); s.normalTexNames.push_back(e.norm);
            matIds[m] = (uint)s.materials.size() - 1;
        }
    }
    const M4 xf = bakeTransform(scale, rot, off);
    const bool hasUvs = !tex.empty();
    const uint start = (uint)s.verts.size();
    uint count = 0;
    std::vector<ObjIndex> tri;
    for (const Face& fc : faces) {
        tri.clear();
        triangulateFace(fc.idx, pos, tri);
        for (size_t i = 0; i + 2 < tri.size(); i += 3) {
            const ObjIndex it[3] = {tri[i], tri[i + 1], tri[i + 2]};
            float v[3][3]; pt_float2 uv[3];
            for (int k = 0; k < 3; k++) {
                const float in[4] = {pos[it[k].v * 3], pos[it[k].v * 3 + 1], pos[it[k].v * 3 + 2], 1};
                float o4[4]; lmMulVec(o4, xf, in);                 // scene.h:285-291
                v[k][0] = o4[0]; v[k][1] = o4[1]; v[k][2] = o4[2];
                uv[k] = hasUvs && it[k].vt >= 0 ? pt_float2{tex[it[k].vt * 2], tex[it[k].vt * 2 + 1]} : pt_float2{0, 0};
                if (useMtl && fc.mtl >= 0) { uv[k].x = uv[k].x + mtls[fc.mtl].off[0]; uv[k].y = uv[k].y + mtls[fc.mtl].off[1]; }
            }
            pt_triangle_v tv; memset(&tv, 0, sizeof(tv));
            tv.v0 = pt_float3{v[0][0], v[0][1], v[0][2]}; tv.v1 = pt_float3{v[1][0], v[1][1], v[1][2]}; tv.v2 = pt_float3{v[2][0], v[2][1], v[2][2]};
            pt_triangle_d td; memset(&td, 0, sizeof(td));
            const float e1[3] = {v[1][0] - v[0][0], v[1][1] - v[0][1], v[1][2] - v[0][2]}, e2[3] = {v[2][0] - v[0][0], v[2][1] - v[0][1], v[2][2] - v[0][2]};
            if (it[0].vn < 0 || it[1].vn < 0 || it[2].vn < 0) { float c[3]; cross3(e1, e2, c); norm3(c); td.normal = pt_float3{c[0], c[1], c[2]}; }
            else td.normal = pt_float3{nrm[it[0].vn * 3], nrm[it[0].vn * 3 + 1], nrm[it[0].vn * 3 + 2]};   // scene.h:302 (v0's normal, untransformed)
            const uint mat = useMtl ? (fc.mtl >= 0 ? matIds[fc.mtl] : 0u) : material;
            if (useMtl && mat >= s.materials.size()) { s.err = std::string("OBJ face without a material and no material defined: ") + path; return PT_ERR_IO; }
            if (useMtl && s.materials[mat].hasNormalMap) {          // scene.h:308-328
                const float d1x = uv[1].x - uv[0].x, d1y = uv[1].y - uv[0].y, d2x = uv[2].x - uv[0].x, d2y = uv[2].y - uv[0].y;
                const float fi = 1.0f / (d1x * d2y - d2x * d1y);
                float tg[3], bt[3];
                for (int a = 0; a < 3; a++) { tg[a] = fi * (d2y * e1[a] - d1y * e2[a]); bt[a] = fi * (d1x * e2[a] - d2x * e1[a]); }
                bool bad = false; for (int a = 0; a < 3; a++) bad = bad || tg[a] != tg[a] || bt[a] != bt[a];
                if (bad) {
                    const float w[3] = {td.normal.x, td.normal.y, td.normal.z};
                    const float ax[3] = {fabs(w[0]) > .1 ? 0.0f : 1.0f, fabs(w[0]) > .1 ? 1.0f : 0.0f, 0.0f};
                    cross3(ax, w, tg); norm3(tg); cross3(w, tg, bt); norm3(bt);
                }
                td.tangent = pt_float3{tg[0], tg[1], tg[2]}; td.bitangent = pt_float3{bt[0], bt[1], bt[2]};
            }
            td.uv0 = uv[0]; td.uv1 = uv[1]; td.uv2 = uv[2]; td.material = mat;
            s.verts.push_back(tv); s.vdata.push_back(td); count++;
        }
    }
    if (count == 0) { s.err = std::string("OBJ file has no faces: ") + path; return PT_ERR_IO; }
    return finishModel(s, start, count);
}

}  // namespace

struct pts_scene { SceneImpl s; };

extern "C" {

PT_API pts_scene* pts_scene_new(void) { return new pts_scene(); }
PT_API void pts_scene_free(pts_scene* h) { delete h; }
PT_API const char* pts_last_error(const pts_scene* h) { return h ? h->s.err.c_str() : ""; }
PT_API double pts_last_build_ms(const pts_scene* h) { return h ? h->s.lastBuildMs : 0.0; }
PT_API void pts_set_build_device(pts_scene* h, int device) { if (h) h->s.buildDevice = device; }

PT_API uint32_t pts_add_material(pts_scene* h, const pt_material* m) {      // Scene::addMaterial, scene.h:140-144
    h->s.materials.push_back(*m); h->s.texNames.push_back(""); h->s.normalTexNames.push_back("");
    return (uint32_t)h->s.materials.size() - 1;
}
PT_API int pts_set_material(pts_scene* h, uint32_t idx, const pt_material* m) {
    if (idx >= h->s.materials.size()) return PT_ERR_INVALID;
    h->s.materials[idx] = *m; return PT_OK;
}

PT_API int pts_add_model_soup(pts_scene* h, const float* v, const float* nrm, const float* uv, const uint32_t* mats, uint32_t n, uint32_t material) {
    SceneImpl& s = h->s;
    if (!v || n == 0) { s.err = "pts_add_model_soup: empty model"; return PT_ERR_INVALID; }
    const uint start = (uint)s.verts.size();
    for (uint32_t i = 0; i < n; i++) {
        pt_triangle_v tv; memset(&tv, 0, sizeof(tv));
        tv.v0 = pt_float3{v[i * 9], v[i * 9 + 1], v[i * 9 + 2]}; tv.v1 = pt_float3{v[i * 9 + 3], v[i * 9 + 4], v[i * 9 + 5]}; tv.v2 = pt_float3{v[i * 9 + 6], v[i * 9 + 7], v[i * 9 + 8]};
        pt_triangle_d td; memset(&td, 0, sizeof(td));
        if (nrm) td.normal = pt_float3{nrm[i * 3], nrm[i * 3 + 1], nrm[i * 3 + 2]};
        else {                                                  // scene.h:297-299
            const float e1[3] = {tv.v1.x - tv.v0.x, tv.v1.y - tv.v0.y, tv.v1.z - tv.v0.z}, e2[3] = {tv.v2.x - tv.v0.x, tv.v2.y - tv.v0.y, tv.v2.z - tv.v0.z};
            float c[3]; cross3(e1, e2, c); norm3(c); td.normal = pt_float3{c[0], c[1], c[2]};
        }
        if (uv) { td.uv0 = pt_float2{uv[i * 6], uv[i * 6 + 1]}; td.uv1 = pt_float2{uv[i * 6 + 2], uv[i * 6 + 3]}; td.uv2 = pt_float2{uv[i * 6 + 4], uv[i * 6 + 5]}; }
        td.material = mats ? mats[i] : material;
        s.verts.push_back(tv); s.vdata.push_back(td);
    }
    return finishModel(s, start, n);
}

PT_API int pts_add_model_obj(pts_scene* h, const char* path, float scale, const float* rot, const float* off, uint32_t material, int use_mtl) {
    return loadObjModel(h->s, path, scale, rot, off, material, use_mtl != 0);
}
PT_API const char* pts_material_texture_name(const pts_scene* h, uint32_t material, int normal_map) {
    if (material >= h->s.materials.size()) return "";
    return normal_map ? h->s.normalTexNames[material].c_str() : h->s.texNames[material].c_str();
}

PT_API uint32_t pts_add_object(pts_scene* h, uint32_t model, uint32_t material, const float* pos, const float* rot, const float* scl) {   // GameObject, types.h:416-429
    GameObject o; o.model_id = model; o.material_id = material;
    for (int a = 0; a < 3; a++) { o.pos[a] = pos ? pos[a] : 0.0f; o.rot[a] = rot ? rot[a] : 0.0f; o.scl[a] = scl ? scl[a] : 1.0f; }
    h->s.objects.push_back(o);
    return (uint32_t)h->s.objects.size() - 1;
}
PT_API int pts_set_object(pts_scene* h, uint32_t idx, const float* pos, const float* rot, const float* scl) {
    if (idx >= h->s.objects.size()) return PT_ERR_INVALID;
    GameObject& o = h->s.objects[idx];
    for (int a = 0; a < 3; a++) { if (pos) o.pos[a] = pos[a]; if (rot) o.rot[a] = rot[a]; if (scl) o.scl[a] = scl[a]; }
    return PT_OK;
}
PT_API void pts_add_sphere(pts_scene* h, const float* pos, float radius, uint32_t material) {
    pt_sphere s; s.pos = pt_float3{pos[0], pos[1], pos[2]}; s.radius = radius; s.material = material; h->s.spheres.push_back(s);
}
PT_API void pts_add_plane(pts_scene* h, const float* n, float d, uint32_t material) {
    pt_plane p; p.normal = pt_float3{n[0], n[1], n[2]}; p.d = d; p.material = material; h->s.planes.push_back(p);
}
// data half of Scene::update (scene.h:396-400)
PT_API int pts_update(pts_scene* h) {
    SceneImpl& s = h->s;
    if (!s.finalized) { s.err = "pts_update before pts_finalize"; return PT_ERR_INVALID; }
    for (size_t i = 0; i < s.objects.size(); i++) {
        if (s.objects[i].model_id >= s.models.size()) { s.err = "object refers to a missing model"; return PT_ERR_INVALID; }
        s.instances[i] = convertToInstance(s.objects[i]);
    }
    buildTopLevel(s.tlas.data(), s.instances.data(), s.bvhs, (uint)s.objects.size());
    return PT_OK;
}
// Scene::finalize (scene.h:354-365) followed by the first update (main.cpp:313 runs it after frame 0;
// a headless caller needs the TLAS before the first render)
PT_API int pts_finalize(pts_scene* h) {
    SceneImpl& s = h->s;
    if (s.verts.size() != s.vdata.size()) { s.err = "vertex/vertexData size mismatch"; return PT_ERR_INVALID; }
    s.tlas.assign(s.objects.empty() ? 0 : 2 * s.objects.size() - 1, pt_tlas_node{});
    s.instances.assign(s.objects.size(), pt_instance{});
    s.finalized = true;
    return pts_update(h);
}
PT_API int pts_get_desc(pts_scene* h, pt_scene_desc* d) {
    SceneImpl& s = h->s;
    memset(d, 0, sizeof(*d));
    for (size_t i = 0; i < s.models.size(); i++) s.models[i].bvh = s.bvhs[i].data();
    d->vertices = s.verts.data(); d->vertexData = s.vdata.data(); d->num_triangles = (uint32_t)s.verts.size();
    d->models = s.models.data(); d->num_models = (uint32_t)s.models.size();
    d->materials = s.materials.data(); d->num_materials = (uint32_t)s.materials.size();
    d->instances = s.instances.data(); d->num_instances = (uint32_t)s.instances.size();
    d->topBvh = s.tlas.data(); d->num_tlas_nodes = (uint32_t)s.tlas.size();
    d->spheres = s.spheres.data(); d->num_spheres = (uint32_t)s.spheres.size();
    d->planes = s.planes.data(); d->num_planes = (uint32_t)s.planes.size();
    return PT_OK;
}
// in9 = eye3, viewDir3, d, focalLength, aperture
PT_API void pts_camera_make(const float* in9, uint32_t W, uint32_t H, pt_camera* out) {
    float eye[3] = {in9[0], in9[1], in9[2]}, dir[3] = {in9[3], in9[4], in9[5]};
    norm3(dir);                                                       // types.h:634
    const float d = in9[6];
    float center[3], u[3], v[3], lt[3];
    for (int a = 0; a < 3; a++) center[a] = eye[a] + d * dir[a];      // types.h:591
    const float up[3] = {0, 1, 0};
    cross3(up, dir, u); norm3(u);                                     // :592
    cross3(dir, u, v); norm3(v);                                      // :593
    const float ar = (float)W / (float)H;                             // :595
    for (int a = 0; a < 3; a++) lt[a] = center[a] - u[a] * ar - v[a]; // :596
    const float k = 2 * ar;
    for (int a = 0; a < 3; a++) { u[a] = u[a] * k; v[a] = v[a] * 2; } // :598-599
    memset(out, 0, sizeof(*out));
    out->eye = pt_float3{eye[0], eye[1], eye[2]}; out->viewDir = pt_float3{dir[0], dir[1], dir[2]};
    out->lt = pt_float3{lt[0], lt[1], lt[2]}; out->u = pt_float3{u[0], u[1], u[2]}; out->v = pt_float3{v[0], v[1], v[2]};
    out->d = d; out->focalLength = in9[7]; out->aperture = in9[8];
}

}  // extern "C"
