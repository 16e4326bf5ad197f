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

struct WideBuildStats { uint32_t nodes, leaves, tris, max_depth; double sah_cost; };

// Appends the wide BVH of one model to `nodes` / `tris` (indices stored in nodes are absolute).
// Returns the absolute index of the model's root node.
uint32_t buildWideBVH(const pt_bvh_node* bvh2, uint32_t n_bvh2, const pt_triangle_v* triangles /* global array */,
                      std::vector<WideNode>& nodes, std::vector<WideTri>& tris, WideBuildStats* stats);


// The same conversion on the device (wide_bvh_device.cu); byte-identical output.  Pointers are device pointers.
size_t wideNodeBound(uint32_t n_bvh2, uint32_t n_tris);
size_t wideScanTempBytes(size_t maxItems);
#ifdef __CUDACC__
int buildWideBVHDevice(cudaStream_t st, const pt_bvh_node* d_bvh2, uint32_t n_bvh2, const pt_triangle_v* d_tris, WideNode* d_nodes, uint32_t nodeBase,
                       WideTri* d_wtris, uint32_t triBase, void* refsScratch, void* countsScratch, void* scanScratch, void* tmp, size_t tmpBytes,
                       uint32_t* nodesUsed, uint32_t* trisUsed, uint32_t* maxDepth, const char** err);
#endif

}  // namespace ptb
#endif
