// TEST INFRASTRUCTURE ONLY (see oracle.h).
// Binary BVH + the reference's binary->N-ary collapse, restated.
//
// The reference builds its binary trees with madmann91/bvh tag v1 (cmake/GetDependencies.cmake:53-58;
// call sites src/runtime/bvh/TriBVHAdapter.h:186-198, SceneBVHAdapter.h:120-132).  That library is not
// vendored, so the binary tree here comes from a plain binned-SAH builder ("tree shape: parity
// unpinned").  Everything AFTER the binary tree follows the reference:
//   - node record {bounds[6] = lo_x,hi_x,lo_y,hi_y,lo_z,hi_z; primitive_count; first_child_or_primitive}
//     as consumed by src/runtime/bvh/NArityBvh.h:12-36,
//   - collapse_to_narity(): NArityBvh.h:65-130 (breadth-first pull-up of grand-children),
//   - the writers in oracle/layout.cpp: BvhNAdapter.h:16-80, TriBVHAdapter.h:50-116, SceneBVHAdapter.h:56-105.
#pragma once
#include <cstdint>
#include <vector>

namespace igo {

struct BinNode {
    float    bounds[6];          // lo_x, hi_x, lo_y, hi_y, lo_z, hi_z
    uint32_t primitive_count;    // 0 => inner node
    uint32_t first_child_or_primitive;
    bool is_leaf() const { return primitive_count != 0; }
};

struct BinBvh {
    std::vector<BinNode>  nodes; // children of an inner node are adjacent: first, first+1
    std::vector<uint32_t> primitive_indices;
};

struct NaryNode {
    float    bounds[6];
    int32_t  primitive_or_child_count; // < 0: leaf with -count primitives
    uint32_t first_child_or_primitive;
    bool     is_leaf() const { return primitive_or_child_count < 0; }
    uint32_t primitive_count() const { return (uint32_t)-primitive_or_child_count; }
};

struct NaryBvh {
    std::vector<NaryNode> nodes;
    std::vector<uint32_t> primitive_indices;
};

// boxes: n * 6 floats (min xyz, max xyz); centers: n * 3 floats.
void build_binary_sah(const float* boxes, const float* centers, uint32_t n, uint32_t max_leaf, BinBvh& out);

// NArityBvh.h:65-130, N in {4, 8}.
void collapse_to_narity(const BinBvh& in, int N, NaryBvh& out);

} // namespace igo
