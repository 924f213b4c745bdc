// TEST INFRASTRUCTURE ONLY (see oracle.h, bvh_build.h).
#include "bvh_build.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <queue>

namespace igo {
namespace {

struct Box {
    float lo[3], hi[3];
    void reset()
    {
        for (int k = 0; k < 3; ++k) { lo[k] = FLT_MAX; hi[k] = -FLT_MAX; }
    }
    void extend(const float* b)
    {
        for (int k = 0; k < 3; ++k) {
            lo[k] = std::min(lo[k], b[k]);
            hi[k] = std::max(hi[k], b[3 + k]);
        }
    }
    void extend(const Box& b)
    {
        for (int k = 0; k < 3; ++k) {
            lo[k] = std::min(lo[k], b.lo[k]);
            hi[k] = std::max(hi[k], b.hi[k]);
        }
    }
    float half_area() const
    {
        const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        return dx * dy + dy * dz + dz * dx;
    }
};

void store_bounds(BinNode& n, const Box& b)
{
    for (int k = 0; k < 3; ++k) {
        n.bounds[2 * k + 0] = b.lo[k];
        n.bounds[2 * k + 1] = b.hi[k];
    }
}

struct Builder {
    static constexpr int NBins = 16;
    const float* boxes;
    const float* centers;
    uint32_t     max_leaf;
    BinBvh&      bvh;

    void build_node(uint32_t node_id, uint32_t begin, uint32_t end)
    {
        uint32_t* idx = bvh.primitive_indices.data();
        Box bb, cb;
        bb.reset();
        cb.reset();
        for (uint32_t i = begin; i < end; ++i) {
            bb.extend(boxes + 6 * idx[i]);
            const float* c = centers + 3 * idx[i];
            for (int k = 0; k < 3; ++k) {
                cb.lo[k] = std::min(cb.lo[k], c[k]);
                cb.hi[k] = std::max(cb.hi[k], c[k]);
            }
        }
        store_bounds(bvh.nodes[node_id], bb);
        const uint32_t count = end - begin;

        auto make_leaf = [&]() {
            bvh.nodes[node_id].primitive_count          = count;
            bvh.nodes[node_id].first_child_or_primitive = begin;
        };
        if (count <= 1) {
            make_leaf();
            return;
        }

        // Binned SAH over the three axes.
        float    best_cost = FLT_MAX;
        int      best_axis = -1, best_bin = -1;
        for (int axis = 0; axis < 3; ++axis) {
            const float ext = cb.hi[axis] - cb.lo[axis];
            if (!(ext > 0)) continue;
            Box      bin_box[NBins];
            uint32_t bin_cnt[NBins] = {};
            for (auto& b : bin_box) b.reset();
            const float scale = (float)NBins / ext;
            for (uint32_t i = begin; i < end; ++i) {
                int b = (int)((centers[3 * idx[i] + axis] - cb.lo[axis]) * scale);
                b     = std::min(std::max(b, 0), NBins - 1);
                bin_box[b].extend(boxes + 6 * idx[i]);
                bin_cnt[b]++;
            }
            float    right_area[NBins];
            uint32_t right_cnt[NBins];
            Box      acc;
            acc.reset();
            uint32_t c = 0;
            for (int b = NBins - 1; b > 0; --b) {
                acc.extend(bin_box[b]);
                c += bin_cnt[b];
                right_area[b] = acc.half_area();
                right_cnt[b]  = c;
            }
            acc.reset();
            c = 0;
            for (int b = 0; b < NBins - 1; ++b) {
                acc.extend(bin_box[b]);
                c += bin_cnt[b];
                if (c == 0 || right_cnt[b + 1] == 0) continue;
                const float cost = acc.half_area() * (float)c + right_area[b + 1] * (float)right_cnt[b + 1];
                if (cost < best_cost) {
                    best_cost = cost;
                    best_axis = axis;
                    best_bin  = b;
                }
            }
        }

        uint32_t mid;
        if (best_axis < 0) {
            // All centres coincide: leaf if allowed, otherwise split in the middle.
            if (count <= max_leaf) { make_leaf(); return; }
            mid = begin + count / 2;
        } else {
            // Leaf when not splitting is cheaper (traversal cost 1, intersection cost 1).
            const float leaf_cost = bb.half_area() * ((float)count - 1.0f);
            if (count <= max_leaf && best_cost >= leaf_cost) { make_leaf(); return; }
            const float ext   = cb.hi[best_axis] - cb.lo[best_axis];
            const float scale = (float)NBins / ext;
            const float lo    = cb.lo[best_axis];
            uint32_t*   m     = std::partition(idx + begin, idx + end, [&](uint32_t p) {
                int b = (int)((centers[3 * p + best_axis] - lo) * scale);
                b     = std::min(std::max(b, 0), NBins - 1);
                return b <= best_bin;
            });
            mid = (uint32_t)(m - idx);
            if (mid == begin || mid == end) mid = begin + count / 2;
        }

        const uint32_t first = (uint32_t)bvh.nodes.size();
        bvh.nodes.emplace_back();
        bvh.nodes.emplace_back();
        bvh.nodes[node_id].primitive_count          = 0;
        bvh.nodes[node_id].first_child_or_primitive = first;
        build_node(first + 0, begin, mid);
        build_node(first + 1, mid, end);
    }
};

NaryNode clone_node(const BinNode& o)
{
    NaryNode n;
    for (int i = 0; i < 6; ++i) n.bounds[i] = o.bounds[i];
    n.primitive_or_child_count = -(int32_t)o.primitive_count;
    n.first_child_or_primitive = o.first_child_or_primitive;
    return n;
}

int ilog2(uint32_t n)
{
    int i = -1;
    while (n) { ++i; n >>= 1; }
    return i;
}

// NArityBvh.h:65-116
void convert_node(const BinBvh& original, const BinNode& node, int N, NaryBvh& bvh, uint32_t cur_id)
{
    const size_t max_iter = (size_t)1 << (ilog2((uint32_t)N) - 1);
    if (node.is_leaf()) return;

    std::queue<BinNode> queue;
    queue.push(original.nodes[node.first_child_or_primitive + 0]);
    queue.push(original.nodes[node.first_child_or_primitive + 1]);

    std::vector<BinNode> children;
    for (size_t k = 0; k < max_iter && !queue.empty(); ++k) {
        const BinNode cur = queue.front();
        queue.pop();
        if (cur.is_leaf()) {
            children.push_back(cur);
        } else {
            queue.push(original.nodes[cur.first_child_or_primitive + 0]);
            queue.push(original.nodes[cur.first_child_or_primitive + 1]);
        }
    }
    while (!queue.empty()) {
        children.push_back(queue.front());
        queue.pop();
    }

    bvh.nodes[cur_id].primitive_or_child_count = (int32_t)children.size();
    bvh.nodes[cur_id].first_child_or_primitive = (uint32_t)bvh.nodes.size();
    for (auto& c : children) bvh.nodes.push_back(clone_node(c));

    for (size_t i = 0; i < children.size(); ++i)
        convert_node(original, children[i], N, bvh, bvh.nodes[cur_id].first_child_or_primitive + (uint32_t)i);
}

} // namespace

void build_binary_sah(const float* boxes, const float* centers, uint32_t n, uint32_t max_leaf, BinBvh& out)
{
    out.nodes.clear();
    out.primitive_indices.resize(n);
    for (uint32_t i = 0; i < n; ++i) out.primitive_indices[i] = i;
    out.nodes.reserve(2 * (size_t)n + 1);
    out.nodes.emplace_back();
    Builder b{ boxes, centers, max_leaf, out };
    b.build_node(0, 0, n);
}

void collapse_to_narity(const BinBvh& in, int N, NaryBvh& out)
{
    out.nodes.clear();
    out.primitive_indices = in.primitive_indices;
    out.nodes.reserve(in.nodes.size());
    out.nodes.push_back(clone_node(in.nodes[0]));
    convert_node(in, in.nodes[0], N, out, 0);
}

} // namespace igo
