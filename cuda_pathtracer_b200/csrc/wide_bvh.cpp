// wide_bvh.cpp -- BVH2 (reference createBVHBinned output) -> 8-wide compressed BVH.  See wide_bvh.h.
#include "wide_bvh.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <deque>
#include <cstdlib>
#include <utility>

namespace ptb {
namespace {

struct Box3 {
    float lo[3], hi[3];
    void reset() { for (int a = 0; a < 3; a++) { lo[a] = 3.0e38f; hi[a] = -3.0e38f; } }
    void grow(const float* p) { for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], p[a]); hi[a] = std::max(hi[a], p[a]); } }
    void grow(const Box3& b) { for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], b.lo[a]); hi[a] = std::max(hi[a], b.hi[a]); } }
    float area() const {
        float x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2];
        return (x < 0 || y < 0 || z < 0) ? 0.0f : 2.0f * (x * y + x * z + y * z);
    }
};

// Temporary binary tree: the reference BVH2 with every leaf cut down to <= 3 triangles (the wide node
// encodes a leaf's triangle count in 3 unary bits).  Triangle ranges stay in reference order.
struct TNode { Box3 box; int left, right; uint32_t first, count; bool virt; uint32_t mask; };   // mask != 0: the triangles first + i of the set bits i (a 4-triangle reference leaf cut by wbBestCut4); count = their number   // virt: made by cutting a reference leaf (not a node of the reference tree)

static Box3 triBounds(const pt_triangle_v* tris, uint32_t first, uint32_t count) {
    Box3 b; b.reset();
    for (uint32_t i = first; i < first + count; i++) { b.grow(&tris[i].v0.x); b.grow(&tris[i].v1.x); b.grow(&tris[i].v2.x); }
    return b;
}

static bool wbCut4Enabled() { return getenv("PTB200_LEAF_INDEX_CUT") == nullptr; }      // developer knob: round 1's cut by index (A/B)

static Box3 maskBounds(const pt_triangle_v* tris, uint32_t first, uint32_t mask) {
    Box3 b; b.reset();
    for (uint32_t i = 0; i < 4; i++) if (mask >> i & 1u) { b.grow(&tris[first + i].v0.x); b.grow(&tris[first + i].v1.x); b.grow(&tris[first + i].v2.x); }
    return b;
}

static int splitLeaf(std::vector<TNode>& t, const pt_triangle_v* tris, uint32_t first, uint32_t count, const Box3* known) {
    TNode n; n.left = n.right = -1; n.first = first; n.count = count; n.virt = true; n.mask = 0;
    n.box = known ? *known : triBounds(tris, first, count);
    int idx = (int)t.size(); t.push_back(n);
    if (count == 4 && wbCut4Enabled()) {
        // the common leaf of the reference builder (it stops at <= 4 triangles): of the seven ways to cut four triangles into two leaves
        // of <= 3, take the one with the least area x count (wide_bvh.h wbBestCut4; the device converter evaluates the same floats)
        float area[16];
        for (uint32_t m = 1; m < 15; m++) area[m] = maskBounds(tris, first, m).area();
        const uint32_t lm = wbBestCut4(area);
        for (int side = 0; side < 2; side++) {
            TNode c; c.left = c.right = -1; c.first = first; c.mask = side ? (15u ^ lm) : lm; c.count = (uint32_t)__builtin_popcount(c.mask); c.virt = true;
            c.box = maskBounds(tris, first, c.mask);
            (side ? t[idx].right : t[idx].left) = (int)t.size(); t.push_back(c);
        }
        t[idx].count = 0;
    } else if (count > 3) {
        uint32_t lc = count / 2;
        int l = splitLeaf(t, tris, first, lc, nullptr);
        int r = splitLeaf(t, tris, first + lc, count - lc, nullptr);
        t[idx].left = l; t[idx].right = r; t[idx].count = 0;
    }
    return idx;
}

static int convert(std::vector<TNode>& t, const pt_bvh_node* bvh2, const pt_triangle_v* tris, uint32_t root) {
    // explicit stack: reference trees can be deep (no depth bound in createBVHBinned)
    struct Item { uint32_t node; int parent; bool isRight; };
    std::vector<Item> stack; stack.push_back({root, -1, false});
    int rootIdx = -1;
    while (!stack.empty()) {
        Item it = stack.back(); stack.pop_back();
        const pt_bvh_node& b = bvh2[it.node];
        uint32_t first, count; memcpy(&first, &b.vmin.w, 4); memcpy(&count, &b.vmax.w, 4);
        Box3 box; box.lo[0] = b.vmin.x; box.lo[1] = b.vmin.y; box.lo[2] = b.vmin.z; box.hi[0] = b.vmax.x; box.hi[1] = b.vmax.y; box.hi[2] = b.vmax.z;
        int idx;
        if (count > 0) idx = splitLeaf(t, tris, first, count, &box);
        else {
            TNode n; n.box = box; n.left = n.right = -1; n.first = 0; n.count = 0; n.virt = false; n.mask = 0;
            idx = (int)t.size(); t.push_back(n);
            stack.push_back({first + 1, idx, true});
            stack.push_back({first, idx, false});
        }
        if (it.parent < 0) rootIdx = idx;
        else if (it.isRight) t[it.parent].right = idx;
        else t[it.parent].left = idx;
    }
    return rootIdx;
}

// table of a cut reference leaf, looked at `depthLeft` cuts deep (wide_bvh.h WB_DP_RANGE_DEPTH; the device converter's rangeTable)
static void virtTable(const std::vector<TNode>& t, int n, int depthLeft, WideDP& out) {
    if (t[n].left < 0) { wdpLeaf(out); return; }
    if (depthLeft == 0) { wdpOpaque(out, t[n].box.area()); return; }
    WideDP L, R;
    virtTable(t, t[n].left, depthLeft - 1, L); virtTable(t, t[n].right, depthLeft - 1, R);
    wdpCombine(out, L, R, t[n].box.area());
}

static inline uint32_t fbits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float bitsf(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

}  // namespace

uint32_t buildWideBVH(const pt_bvh_node* bvh2, uint32_t n_bvh2, const pt_triangle_v* triangles,
                      std::vector<WideNode>& nodes, std::vector<WideTri>& tris, WideBuildStats* stats, bool optimal) {
    (void)n_bvh2;
    std::vector<TNode> t;
    t.reserve(n_bvh2 * 2);
    const int troot = convert(t, bvh2, triangles, 0);

    const uint32_t rootIndex = (uint32_t)nodes.size();
    nodes.push_back(WideNode());
    struct Work { int tnode; uint32_t wide; uint32_t depth; };
    std::deque<Work> queue;
    queue.push_back({troot, rootIndex, 1});
    uint32_t nLeaves = 0, maxDepth = 0;
    double sah = 0.0;
    const float rootArea = std::max(t[troot].box.area(), 1e-30f);

    // ---- SAH-optimal collapse (wide_bvh.h): tables of the reference tree's nodes, children before parents ----
    std::vector<WideDP> D;
    bool dp = optimal;
    if (dp) {
        std::vector<int> order; order.reserve(t.size());
        std::vector<std::pair<int, uint32_t>> st; st.push_back({troot, 0});
        uint32_t deepest = 0;
        while (!st.empty()) {                                  // reference-tree nodes only: a cut leaf is one unit here
            const int n = st.back().first; const uint32_t lvl = st.back().second; st.pop_back();
            order.push_back(n); deepest = std::max(deepest, lvl);
            if (!t[n].virt && t[n].left >= 0) { st.push_back({t[n].left, lvl + 1}); st.push_back({t[n].right, lvl + 1}); }
        }
        if (deepest >= WB_DP_MAX_LEVELS) dp = false;           // the device converter walks a bounded number of levels: same rule here
        else {
            D.resize(t.size());
            for (size_t q = order.size(); q-- > 0;) {
                const int n = order[q];
                if (t[n].virt || t[n].left < 0) virtTable(t, n, WB_DP_RANGE_DEPTH, D[n]);
                else wdpCombine(D[n], D[t[n].left], D[t[n].right], t[n].box.area());
            }
        }
    }
    // children of the wide node rooted at a subtree: depth-first, left before right, each subtree within its budget
    struct Cover {
        const std::vector<TNode>& t; const std::vector<WideDP>& D; int* ch; int& nch;
        void table(int n, WideDP& out) const { if (t[n].virt) virtTable(t, n, WB_DP_RANGE_DEPTH, out); else out = D[n]; }
        void run(int n, int budget) {
            if (t[n].left < 0) { ch[nch++] = n; return; }
            WideDP d; table(n, d);
            const int b = wdpEffective(d, budget);
            if (b == 1) { ch[nch++] = n; return; }
            const int k = d.k[b - 1];
            run(t[n].left, k); run(t[n].right, b - k);
        }
    };

    while (!queue.empty()) {
        Work w = queue.front(); queue.pop_front();
        maxDepth = std::max(maxDepth, w.depth);
        const TNode& r = t[w.tnode];
        // ---- collapse: open the child with the largest surface area until 8 children ----
        int ch[8]; int nch = 0;
        if (r.left < 0) ch[nch++] = w.tnode;                 // degenerate model: the root itself is a leaf
        else if (dp) { Cover cov{t, D, ch, nch}; WideDP d; cov.table(w.tnode, d); const int k = d.k[7]; cov.run(r.left, k); cov.run(r.right, 8 - k); }
        else { ch[nch++] = r.left; ch[nch++] = r.right; }
        while (!dp && nch < 8) {
            int best = -1; float bestArea = -1.0f;
            for (int i = 0; i < nch; i++)
                if (t[ch[i]].left >= 0) { float a = t[ch[i]].box.area(); if (a > bestArea) { bestArea = a; best = i; } }
            if (best < 0) break;
            int c = ch[best];
            ch[best] = t[c].left; ch[nch++] = t[c].right;
        }
        // leaf triangles per node are limited to 24 (3 x 8) by construction.
        // ---- assign children to slots so that slot ^ octant orders them front to back ----
        float ctr[3]; for (int a = 0; a < 3; a++) ctr[a] = 0.5f * (r.box.lo[a] + r.box.hi[a]);
        float cost[8][8];
        for (int i = 0; i < nch; i++) {
            const Box3& b = t[ch[i]].box;
            float d[3]; for (int a = 0; a < 3; a++) d[a] = 0.5f * (b.lo[a] + b.hi[a]) - ctr[a];
            for (int s = 0; s < 8; s++) {
                float sx = (s & 4) ? 1.0f : -1.0f, sy = (s & 2) ? 1.0f : -1.0f, sz = (s & 1) ? 1.0f : -1.0f;
                cost[i][s] = d[0] * sx + d[1] * sy + d[2] * sz;
            }
        }
        int slotOf[8]; int childIn[8];
        for (int i = 0; i < 8; i++) { slotOf[i] = -1; childIn[i] = -1; }
        for (int k = 0; k < nch; k++) {
            int bi = -1, bs = -1; float bc = -3.0e38f;
            for (int i = 0; i < nch; i++) if (slotOf[i] < 0)
                for (int s = 0; s < 8; s++) if (childIn[s] < 0 && cost[i][s] > bc) { bc = cost[i][s]; bi = i; bs = s; }
            slotOf[bi] = bs; childIn[bs] = ch[bi];
        }
        // ---- quantisation frame ----
        uint32_t ebyte[3]; float scale[3];
        for (int a = 0; a < 3; a++) {
            float ext = r.box.hi[a] - r.box.lo[a];
            // e = ceil(log2(ext / 255)), from the float's own exponent (exact; the device converter does the same)
            int e = -126;
            if (ext > 0.0f) { int ex; const float m = std::frexp(ext / 255.0f, &ex); e = m == 0.5f ? ex - 1 : ex; }
            e = std::max(-126, std::min(127, e));
            // make sure ceil((hi-lo)/2^e) <= 255 despite rounding
            while (e < 127 && std::ceil(ext / std::ldexp(1.0f, e)) > 255.0f) e++;
            ebyte[a] = (uint32_t)(e + 127);
            scale[a] = std::ldexp(1.0f, e);
        }
        WideNode out; memset(&out, 0, sizeof(out));
        uint8_t* bytes = (uint8_t*)out.w;
        uint32_t imask = 0, nInternal = 0, triOffset = 0;
        const uint32_t triBase = (uint32_t)tris.size();
        uint32_t childBase = (uint32_t)nodes.size();
        for (int s = 0; s < 8; s++) {
            int c = childIn[s];
            if (c < 0) continue;                                 // meta = 0, empty box (lo=255? keep 0/0 with meta 0 => never hit)
            const TNode& cn = t[c];
            uint8_t meta;
            if (cn.left >= 0) {
                imask |= 1u << s; meta = (uint8_t)(0x20 | (24 + s));
                nodes.push_back(WideNode());
                queue.push_back({c, childBase + nInternal, w.depth + 1});
                nInternal++;
            } else {
                static const uint8_t unary[4] = {0, 1, 3, 7};
                meta = (uint8_t)((unary[cn.count] << 5) | triOffset);
                for (uint32_t i = cn.first; i < cn.first + (cn.mask ? 4u : cn.count); i++) {
                    if (cn.mask && !(cn.mask >> (i - cn.first) & 1u)) continue;
                    const pt_triangle_v& tv = triangles[i];
                    WideTri wt; memset(&wt, 0, sizeof(wt));
                    wt.v0[0] = tv.v0.x; wt.v0[1] = tv.v0.y; wt.v0[2] = tv.v0.z; wt.id = i;
                    wt.e1[0] = tv.v1.x - tv.v0.x; wt.e1[1] = tv.v1.y - tv.v0.y; wt.e1[2] = tv.v1.z - tv.v0.z;   // kernels.h:171
                    wt.e2[0] = tv.v2.x - tv.v0.x; wt.e2[1] = tv.v2.y - tv.v0.y; wt.e2[2] = tv.v2.z - tv.v0.z;   // kernels.h:172
                    tris.push_back(wt);
                }
                triOffset += cn.count;
                nLeaves++;
                sah += (double)cn.box.area() / rootArea * cn.count;
            }
            bytes[24 + s] = meta;                                // w1 bytes 8..15 = meta[0..7]
            for (int a = 0; a < 3; a++) {
                float qlo = std::floor((cn.box.lo[a] - r.box.lo[a]) / scale[a]);
                float qhi = std::ceil((cn.box.hi[a] - r.box.lo[a]) / scale[a]);
                // guard against rounding in the division: the decoded box must contain the child box
                while (qlo > 0.0f && r.box.lo[a] + qlo * scale[a] > cn.box.lo[a]) qlo -= 1.0f;
                while (qhi < 255.0f && r.box.lo[a] + qhi * scale[a] < cn.box.hi[a]) qhi += 1.0f;
                qlo = std::max(0.0f, std::min(255.0f, qlo)); qhi = std::max(0.0f, std::min(255.0f, qhi));
                bytes[32 + a * 8 + s] = (uint8_t)qlo;           // w2.. : qlo_x[8] qlo_y[8] qlo_z[8]
                bytes[56 + a * 8 + s] = (uint8_t)qhi;           //        qhi_x[8] qhi_y[8] qhi_z[8]
            }
            sah += (double)cn.box.area() / rootArea * (cn.left >= 0 ? 1.0 : 0.0);
        }
        out.w[0] = fbits(r.box.lo[0]); out.w[1] = fbits(r.box.lo[1]); out.w[2] = fbits(r.box.lo[2]);
        out.w[3] = ebyte[0] | (ebyte[1] << 8) | (ebyte[2] << 16) | (imask << 24);
        out.w[4] = childBase; out.w[5] = triBase;
        nodes[w.wide] = out;
    }
    if (stats) {
        stats->nodes = (uint32_t)nodes.size() - rootIndex; stats->leaves = nLeaves; stats->max_depth = maxDepth;
        stats->tris = 0; for (const TNode& n : t) if (n.left < 0) stats->tris += n.count;
        stats->sah_cost = sah;
    }
    return rootIndex;
}

}  // namespace ptb
