// debug_api.cpp -- host-only test hook: builds a BVH8 and verifies its structural invariants.
#include "../../include/ignis_b200.h"
#include "bvh8.h"

#include <algorithm>
#include <cfloat>
#include <chrono>
#include <cstring>
#include <vector>

namespace {
struct Checker {
    const igb::Bvh8& bvh;
    const float*     boxes;
    std::vector<uint32_t> seen;
    uint64_t violations = 0, leaves = 0, inner = 0;

    // returns exact bounds of the subtree rooted at node `id`
    void visit(uint32_t id, float* lo, float* hi)
    {
        const igb::Node8& n = bvh.nodes[id];
        for (int k = 0; k < 3; ++k) { lo[k] = FLT_MAX; hi[k] = -FLT_MAX; }
        for (int s = 0; s < 8; ++s) {
            const uint8_t meta = n.meta[s];
            if (meta == 0) continue;
            float clo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, chi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
            if ((meta & 0x18) == 0x18 && (meta >> 5) == 1) {
                if ((meta & 7) != s || !((n.imask >> s) & 1)) violations++;
                const uint32_t rel = __builtin_popcount(n.imask & ((1u << s) - 1u));
                visit(n.child_base + rel, clo, chi);
                inner++;
            } else {
                const uint32_t cnt = __builtin_popcount(meta >> 5), off = meta & 0x1F;
                if (cnt == 0 || cnt > 3 || off + cnt > 24 || ((n.imask >> s) & 1)) violations++;
                for (uint32_t i = 0; i < cnt; ++i) {
                    const uint32_t p = bvh.prim_indices[n.prim_base + off + i];
                    seen[p]++;
                    for (int k = 0; k < 3; ++k) {
                        clo[k] = std::min(clo[k], boxes[6 * (size_t)p + k]);
                        chi[k] = std::max(chi[k], boxes[6 * (size_t)p + 3 + k]);
                    }
                }
                leaves++;
            }
            float q[6];
            igb::node8_child_box(n, s, q);
            for (int k = 0; k < 3; ++k) {
                if (!(q[k] <= clo[k]) || !(q[3 + k] >= chi[k])) violations++;
                lo[k] = std::min(lo[k], clo[k]);
                hi[k] = std::max(hi[k], chi[k]);
            }
        }
    }
};
} // namespace

extern "C" int igb200_debug_check_bvh8(const float* boxes, uint32_t n, uint32_t max_leaf, uint64_t* out5)
{
    igb::Bvh8 bvh;
    igb::build_bvh8(boxes, n, max_leaf, bvh);
    Checker c{ bvh, boxes, std::vector<uint32_t>(n, 0) };
    float   lo[3], hi[3];
    c.visit(0, lo, hi);
    uint64_t missing = 0;
    for (uint32_t s : c.seen) missing += (s != 1);
    if (out5) {
        out5[0] = bvh.nodes.size();
        out5[1] = bvh.max_depth;
        out5[2] = c.violations;
        out5[3] = missing;
        out5[4] = c.leaves;
    }
    return (c.violations == 0 && missing == 0) ? IGB200_OK : IGB200_E_INVALID;
}

extern "C" int igb200_debug_bvh8_quality(const float* boxes, uint32_t n, uint32_t max_leaf, double* out4)
{
    igb::Bvh8 bvh;
    const auto t0 = std::chrono::steady_clock::now();
    igb::build_bvh8(boxes, n, max_leaf, bvh);
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (out4) {
        out4[0] = bvh.sah_node_term;
        out4[1] = bvh.sah_prim_term;
        out4[2] = (double)bvh.nodes.size();
        out4[3] = ms;
    }
    return IGB200_OK;
}
