// wide_bvh.h -- 8-wide compressed BVH ("CWBVH"-style, Ylitie/Karras/Laine 2017) built from the
// reference's binned-SAH BVH2 (createBVHBinned output, reference src/bvhBuilder.h:35-274; node
// format src/types.h:347-375).  north_star: "a wide, compressed BVH node layout built from the
// reference's SAH builder output and fetched with coalesced, vectorised 16-byte loads".
//
// Node = 80 bytes = five 16-byte words (all fetched with LDG.128):
//   w0: origin p.xyz (float) | ex, ey, ez (biased float exponents of the quantisation grid), imask
//   w1: child_base (absolute node index of the first internal child) | tri_base (index of the first
//       triangle record of this node) | meta[0..3] | meta[4..7]
//   w2: qlo_x[8] qlo_y[8]      w3: qlo_z[8] qhi_x[8]      w4: qhi_y[8] qhi_z[8]     (uint8 each)
// meta[slot]: 0 = empty; internal child: 0x20 | (24 + slot); leaf: (unary triangle count << 5) | offset
// from tri_base (at most 3 triangles per leaf child, 24 per node).
//
// Triangle record = 48 bytes = three 16-byte words, laid out in wide-leaf order:
//   (v0.xyz, reference triangle id) (v0v1.xyz, 0) (v0v2.xyz, 0)
// The edges are the float differences v1-v0 and v2-v0 exactly as the reference's Moeller-Trumbore
// computes them per test (src/kernels.h:171-172), so the intersection arithmetic stays bit-identical.
#ifndef PTB200_WIDE_BVH_H
#define PTB200_WIDE_BVH_H
#include <cstdint>
#include <vector>
#include "../../include/ptb200.h"

namespace ptb {

struct WideNode { uint32_t w[20]; };   // 80 bytes
struct WideTri { float v0[3]; uint32_t id; float e1[3]; uint32_t pad1; float e2[3]; uint32_t pad2; };   // 48 bytes
static_assert(sizeof(WideNode) == 80, "node size");
static_assert(sizeof(WideTri) == 48, "tri size");

// ---- SAH-optimal collapse (Ylitie, Karras, Laine 2017, section 4.1) ---------------------------------------------------------
// Which nodes of the binary tree become the (up to 8) children of a wide node is decided by dynamic programming instead of
// "open the child with the largest area": c[i-1] = least cost of covering a subtree with at most i children of ONE wide node,
// where a child that is an inner node costs its surface area (the chance of a node step) plus the best 8-way cover of its own
// subtree, and leaf children (<= 3 triangles) cost the same whatever the cover (their term is dropped).  k[i-1] = how many of
// the i children the left subtree gets (0: budget i is no better than i - 1); k[7] = the split of the node's own 8 children.
// Host and device converter share these functions, so their tables -- and with them the trees -- are identical.  Measured
// (sibenik proxy, profiles/r2/ab_wide_dp.log): -4.8 % node steps per extension ray, -8.3 % per shadow ray, -2.4..2.9 % frame time.
#ifdef __CUDACC__
#define WB_HD __host__ __device__
#else
#define WB_HD
#endif
#define WB_DP_RANGE_DEPTH 3     /* leaves of the reference tree with more than 3 triangles are cut in halves; the table of such a
                                   range looks this many cuts deep (exact up to 24 triangles), deeper ranges count as one inner child */
#define WB_DP_MAX_LEVELS 96     /* binary trees deeper than this fall back to the greedy rule (both converters) */
struct WideDP { float c[7]; unsigned char k[8]; };
static_assert(sizeof(WideDP) == 36, "WideDP size");
WB_HD inline void wdpLeaf(WideDP& d) { for (int i = 0; i < 7; i++) { d.c[i] = 0.0f; d.k[i] = 0; } d.k[7] = 0; }
WB_HD inline void wdpOpaque(WideDP& d, float area) { for (int i = 0; i < 7; i++) { d.c[i] = area; d.k[i] = 0; } d.k[7] = 0; }
WB_HD inline float wdpSplit(const WideDP& L, const WideDP& R, int budget, unsigned char& kbest) {      // budget 2..8 over the two subtrees
    float best = 3.0e38f; kbest = 1;
    for (int k = 1; k < budget; k++) {
        if (k > 7 || budget - k > 7) continue;
        const float c = L.c[k - 1] + R.c[budget - k - 1];
        if (c < best) { best = c; kbest = (unsigned char)k; }
    }
    return best;
}
WB_HD inline void wdpCombine(WideDP& d, const WideDP& L, const WideDP& R, float area) {
    d.c[0] = area + wdpSplit(L, R, 8, d.k[7]); d.k[0] = 0;
    for (int i = 2; i <= 7; i++) {
        unsigned char kb; const float c = wdpSplit(L, R, i, kb);
        if (c < d.c[i - 2]) { d.c[i - 1] = c; d.k[i - 1] = kb; } else { d.c[i - 1] = d.c[i - 2]; d.k[i - 1] = 0; }
    }
}
// children a subtree may take out of `budget`: the largest budget that still pays (k != 0), 1 = the subtree stays one child
WB_HD inline int wdpEffective(const WideDP& d, int budget) { while (budget > 1 && d.k[budget - 1] == 0) budget--; return budget; }

// ---- cutting the reference builder's usual leaf -------------------------------------------------------------------------------
// createBVHBinned stops at <= 4 triangles, a wide leaf child holds <= 3: three of four reference leaves (sibenik proxy: 28,585 of 37,583)
// have to be cut in two.  Round 1 cut by index (2 + 2 in reference order); wbBestCut4 picks, of the seven cuts {a | bcd} x 4 and
// {ab | cd} x 3, the one with the least sum of area x triangle count.  area[m] = surface area of the box of the triangles in bit mask m
// (1..14).  Returns the mask of the left leaf; ties go to the first cut in this order.  Shared by both converters.
WB_HD inline uint32_t wbBestCut4(const float* area) {
    const uint32_t cuts[7] = {3u, 5u, 6u, 1u, 2u, 4u, 8u};          // left masks: the three pairings (the one holding triangle 0 on the left), then each single
    uint32_t best = 3u; float bestCost = 3.0e38f;
    for (int i = 0; i < 7; i++) {
        const uint32_t l = cuts[i], r = 15u ^ l;
        const float nl = i < 3 ? 2.0f : 1.0f, nr = i < 3 ? 2.0f : 3.0f;
        const float c = area[l] * nl + area[r] * nr;
        if (c < bestCost) { bestCost = c; best = l; }
    }
    return best;
}

struct WideBuildStats { uint32_t nodes, leaves, tris, max_depth; double sah_cost; };

// Appends the wide BVH of one model to `nodes` / `tris` (indices stored in nodes are absolute).
// Returns the absolute index of the model's root node.
uint32_t buildWideBVH(const pt_bvh_node* bvh2, uint32_t n_bvh2, const pt_triangle_v* triangles /* global array */,
                      std::vector<WideNode>& nodes, std::vector<WideTri>& tris, WideBuildStats* stats, bool optimal = true);


// The same conversion on the device (wide_bvh_device.cu); byte-identical output.  Pointers are device pointers.
size_t wideNodeBound(uint32_t n_bvh2, uint32_t n_tris);
size_t wideScanTempBytes(size_t maxItems);
#ifdef __CUDACC__
int buildWideBVHDevice(cudaStream_t st, const pt_bvh_node* d_bvh2, uint32_t n_bvh2, const pt_triangle_v* d_tris, WideNode* d_nodes, uint32_t nodeBase,
                       WideTri* d_wtris, uint32_t triBase, void* refsScratch, void* countsScratch, void* scanScratch, void* tmp, size_t tmpBytes,
                       uint32_t* nodesUsed, uint32_t* trisUsed, uint32_t* maxDepth, const char** err, void* dpScratch = nullptr);
// dpScratch: wideDpScratchBytes(n_bvh2) bytes for the optimal collapse, nullptr = greedy collapse (as buildWideBVH with optimal = false)
size_t wideDpScratchBytes(uint32_t n_bvh2);
#endif

}  // namespace ptb
#endif
