// bvh8.h -- compressed 8-wide BVH (80-byte nodes) used by the sm_100a traversal kernels.
//
// Replaces the reference's GPU acceleration structure: `Node2` (64 B binary node,
// src/artic/traversal/mapping_gpu.art:3-7) + `Tri1` (48 B, src/artic/shapes/trimesh.art:107-114) built by
// src/runtime/bvh/TriBVHAdapter.h:163-202 / SceneBVHAdapter.h:109-135 with madmann91/bvh.
// Here: own binned-SAH binary build -> greedy surface-area collapse to 8 children -> octant slot
// assignment -> 8-bit quantised child boxes in a per-node power-of-two grid.
// Triangle records stay 48 B and Tri1-compatible (v0, e1 = v0 - v1, e2 = v2 - v0, prim_id) so the
// Moeller-Trumbore arithmetic is the reference's (src/artic/traversal/intersection.art:71-101).
#pragma once
#include <cstdint>
#include <vector>

namespace igb {

// 80 bytes = 5 x 16 B, fetched as five coalescable 128-bit loads.
struct alignas(16) Node8 {
    float    p[3];       // grid origin (parent box min)
    uint8_t  e[3];       // per-axis grid scale exponent (biased, as IEEE f32 exponent field)
    uint8_t  imask;      // bit s set: slot s holds an inner node
    uint32_t child_base; // index of the first inner child (children are stored contiguously, in slot order)
    uint32_t prim_base;  // index of the first primitive record referenced by this node's leaves
    uint8_t  meta[8];    // inner: 0b001'11sss (sss = slot); leaf: unary count in bits 7..5, first-prim offset in 4..0; empty: 0
    uint8_t  qlo[3][8];  // quantised child box minima, [axis][slot]
    uint8_t  qhi[3][8];  // quantised child box maxima
};
static_assert(sizeof(Node8) == 80, "Node8 must be 80 bytes");

// 48 bytes, byte-compatible with the reference's Tri1.
struct alignas(16) TriRecord {
    float   v0[3];
    int32_t pad0;
    float   e1[3];
    int32_t pad1;
    float   e2[3];
    int32_t prim_id;
};
static_assert(sizeof(TriRecord) == 48, "TriRecord must be 48 bytes");

struct Bvh8 {
    std::vector<Node8>    nodes;        // nodes[0] is the root
    std::vector<uint32_t> prim_indices; // primitive records in leaf-contiguous order (indices into the input)
    uint32_t              max_depth = 0;
    // SAH cost of the wide tree, split in its two sums (areas relative to the root's): sum over wide nodes of A_n and
    // sum over leaf children of A_l * N_l.  cost = c_node * node_term + c_prim * prim_term.
    double                sah_node_term = 0.0, sah_prim_term = 0.0;
};

// boxes: n x {min xyz, max xyz}. max_leaf in 1..3 (the meta encoding holds at most 3 per leaf).
// An empty input yields a single node with no children.
// balanced: object-median splits (binary depth <= ceil(log2 n)) instead of binned SAH -- the fallback when the SAH tree of a
// degenerate mesh is deeper than the traversal stack allows.
void build_bvh8(const float* boxes, uint32_t n, uint32_t max_leaf, Bvh8& out, bool balanced = false);

// Dequantised box of child `slot` (for the conservative-bounds property test).
void node8_child_box(const Node8& node, int slot, float* out_min_max6);

} // namespace igb
