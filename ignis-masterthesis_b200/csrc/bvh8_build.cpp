// bvh8_build.cpp -- host builder of the compressed BVH8 (see bvh8.h).
#include "bvh8.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <deque>

namespace igb {
namespace {

struct Aabb {
    float lo[3], hi[3];
    void  reset()
    {
        for (int k = 0; k < 3; ++k) { lo[k] = FLT_MAX; hi[k] = -FLT_MAX; }
    }
    void grow(const float* b)
    {
        for (int k = 0; k < 3; ++k) {
            lo[k] = std::min(lo[k], b[k]);
            hi[k] = std::max(hi[k], b[3 + k]);
        }
    }
    void grow(const Aabb& b)
    {
        for (int k = 0; k < 3; ++k) {
            lo[k] = std::min(lo[k], b.lo[k]);
            hi[k] = std::max(hi[k], b.hi[k]);
        }
    }
    float area() const
    {
        const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        return dx * dy + dy * dz + dz * dx;
    }
};

struct BNode {
    Aabb     box;
    uint32_t left = 0, right = 0; // inner: child ids
    uint32_t first = 0, count = 0; // leaf: range in the index array (count > 0)
};

struct BinaryBuilder {
    static constexpr int NB = 32;
    const float*          boxes;
    uint32_t              max_leaf;
    std::vector<uint32_t> idx;
    std::vector<float>    cent; // 3 per primitive
    std::vector<BNode>    nodes;

    struct Task { uint32_t node, begin, end; };

    void run(uint32_t n)
    {
        idx.resize(n);
        cent.resize(3 * (size_t)n);
        for (uint32_t i = 0; i < n; ++i) {
            idx[i] = i;
            for (int k = 0; k < 3; ++k) cent[3 * (size_t)i + k] = 0.5f * (boxes[6 * (size_t)i + k] + boxes[6 * (size_t)i + 3 + k]);
        }
        nodes.reserve(2 * (size_t)n);
        nodes.emplace_back();
        std::vector<Task> todo{ Task{ 0, 0, n } };
        while (!todo.empty()) {
            const Task t = todo.back();
            todo.pop_back();
            split(t, todo);
        }
    }

    void split(const Task& t, std::vector<Task>& todo)
    {
        Aabb bb, cb;
        bb.reset();
        cb.reset();
        for (uint32_t i = t.begin; i < t.end; ++i) {
            bb.grow(boxes + 6 * (size_t)idx[i]);
            const float* c = &cent[3 * (size_t)idx[i]];
            for (int k = 0; k < 3; ++k) {
                cb.lo[k] = std::min(cb.lo[k], c[k]);
                cb.hi[k] = std::max(cb.hi[k], c[k]);
            }
        }
        nodes[t.node].box    = bb;
        const uint32_t count = t.end - t.begin;
        auto           leaf  = [&]() { nodes[t.node].first = t.begin; nodes[t.node].count = count; };
        if (count == 1) { leaf(); return; }

        float best = FLT_MAX;
        int   best_axis = -1, best_bin = -1;
        for (int axis = 0; axis < 3; ++axis) {
            const float ext = cb.hi[axis] - cb.lo[axis];
            if (!(ext > 0)) continue;
            Aabb     bbox[NB];
            uint32_t bcnt[NB] = {};
            for (auto& b : bbox) b.reset();
            const float scale = (float)NB / ext;
            for (uint32_t i = t.begin; i < t.end; ++i) {
                int b = (int)((cent[3 * (size_t)idx[i] + axis] - cb.lo[axis]) * scale);
                b     = std::min(std::max(b, 0), NB - 1);
                bbox[b].grow(boxes + 6 * (size_t)idx[i]);
                bcnt[b]++;
            }
            float    rarea[NB];
            uint32_t rcnt[NB];
            Aabb     acc;
            acc.reset();
            uint32_t c = 0;
            for (int b = NB - 1; b > 0; --b) {
                acc.grow(bbox[b]);
                c += bcnt[b];
                rarea[b] = c ? acc.area() : 0.0f;
                rcnt[b]  = c;
            }
            acc.reset();
            c = 0;
            for (int b = 0; b < NB - 1; ++b) {
                acc.grow(bbox[b]);
                c += bcnt[b];
                if (c == 0 || rcnt[b + 1] == 0) continue;
                const float cost = acc.area() * (float)c + rarea[b + 1] * (float)rcnt[b + 1];
                if (cost < best) { best = cost; best_axis = axis; best_bin = b; }
            }
        }

        uint32_t mid;
        if (best_axis < 0) {
            if (count <= max_leaf) { leaf(); return; }
            mid = t.begin + count / 2;
        } else {
            // traversal cost of one extra (1/8th of a wide) node step ~ 0.3 triangle tests
            const float leaf_cost = bb.area() * ((float)count - 0.3f);
            if (count <= max_leaf && best >= leaf_cost) { leaf(); return; }
            const float lo    = cb.lo[best_axis];
            const float scale = (float)NB / (cb.hi[best_axis] - lo);
            auto        m     = std::partition(idx.begin() + t.begin, idx.begin() + t.end, [&](uint32_t p) {
                int b = (int)((cent[3 * (size_t)p + best_axis] - lo) * scale);
                b     = std::min(std::max(b, 0), NB - 1);
                return b <= best_bin;
            });
            mid = (uint32_t)(m - idx.begin());
            if (mid == t.begin || mid == t.end) mid = t.begin + count / 2;
        }
        const uint32_t l = (uint32_t)nodes.size();
        nodes.emplace_back();
        nodes.emplace_back();
        nodes[t.node].left  = l;
        nodes[t.node].right = l + 1;
        todo.push_back(Task{ l + 1, mid, t.end });
        todo.push_back(Task{ l, t.begin, mid });
    }
};

inline float exp_scale(uint8_t e)
{
    const uint32_t bits = (uint32_t)e << 23;
    float          f;
    std::memcpy(&f, &bits, 4);
    return f;
}

// Quantise the children of one wide node. child[s] < 0: empty slot.
void encode_node(Node8& out, const Aabb& box, const Aabb* cbox, const bool* used)
{
    for (int k = 0; k < 3; ++k) {
        out.p[k]        = box.lo[k];
        const float ext = box.hi[k] - box.lo[k];
        int         e   = 1; // smallest normal scale for flat boxes
        if (ext > 0) {
            int ex;
            std::frexp((double)ext / 255.0, &ex); // ext/255 = m * 2^ex, m in [0.5, 1) -> 2^ex >= ext/255
            e = std::min(std::max(ex + 127, 1), 254);
        }
        for (;;) {
            const float scale = exp_scale((uint8_t)e);
            bool        ok    = true;
            for (int s = 0; s < 8 && ok; ++s) {
                if (!used[s]) { out.qlo[k][s] = 255; out.qhi[k][s] = 0; continue; }
                double qlo = std::floor(((double)cbox[s].lo[k] - (double)box.lo[k]) / (double)scale);
                double qhi = std::ceil(((double)cbox[s].hi[k] - (double)box.lo[k]) / (double)scale);
                qlo        = std::max(qlo, 0.0);
                qhi        = std::max(qhi, qlo);
                // float decode as seen by a consumer must stay conservative
                while (qlo > 0 && box.lo[k] + (float)qlo * scale > cbox[s].lo[k]) qlo -= 1;
                while (qhi < 256 && box.lo[k] + (float)qhi * scale < cbox[s].hi[k]) qhi += 1;
                if (qhi > 255) { ok = false; break; }
                out.qlo[k][s] = (uint8_t)qlo;
                out.qhi[k][s] = (uint8_t)qhi;
            }
            if (ok) break;
            ++e;
        }
        out.e[k] = (uint8_t)e;
    }
}

} // namespace

void node8_child_box(const Node8& node, int slot, float* o)
{
    for (int k = 0; k < 3; ++k) {
        const float scale = exp_scale(node.e[k]);
        o[k]              = node.p[k] + (float)node.qlo[k][slot] * scale;
        o[3 + k]          = node.p[k] + (float)node.qhi[k][slot] * scale;
    }
}

void build_bvh8(const float* boxes, uint32_t n, uint32_t max_leaf, Bvh8& out)
{
    out.nodes.clear();
    out.prim_indices.clear();
    out.max_depth = 0;
    max_leaf      = std::min(std::max(max_leaf, 1u), 3u);
    if (n == 0) {
        Node8 root;
        std::memset(&root, 0, sizeof(root));
        for (int k = 0; k < 3; ++k) {
            root.e[k] = 1;
            for (int s = 0; s < 8; ++s) { root.qlo[k][s] = 255; root.qhi[k][s] = 0; }
        }
        out.nodes.push_back(root);
        return;
    }

    BinaryBuilder bb{ boxes, max_leaf, {}, {}, {} };
    bb.run(n);
    const std::vector<BNode>& bn = bb.nodes;

    struct Work { uint32_t node8, bnode, depth; };
    std::deque<Work> queue;
    out.nodes.emplace_back();
    out.prim_indices.reserve(n);

    // A leaf root becomes a wide node with one leaf child.
    queue.push_back(Work{ 0, 0, 1 });
    while (!queue.empty()) {
        const Work w = queue.front();
        queue.pop_front();
        out.max_depth = std::max(out.max_depth, w.depth);

        // Greedy collapse: open the inner child with the largest surface area until 8 children.
        uint32_t kids[8];
        int      nk = 0;
        if (bn[w.bnode].count > 0) {
            kids[nk++] = w.bnode;
        } else {
            kids[nk++] = bn[w.bnode].left;
            kids[nk++] = bn[w.bnode].right;
            while (nk < 8) {
                int   pick = -1;
                float pa   = -1.0f;
                for (int i = 0; i < nk; ++i) {
                    if (bn[kids[i]].count == 0) {
                        const float a = bn[kids[i]].box.area();
                        if (a > pa) { pa = a; pick = i; }
                    }
                }
                if (pick < 0) break;
                const uint32_t open = kids[pick];
                kids[pick]          = bn[open].left;
                kids[nk++]          = bn[open].right;
            }
        }

        // Octant slot assignment (greedy auction): slot bit 2/1/0 set <=> child lies towards +x/+y/+z.
        const Aabb& box = bn[w.bnode].box;
        float       cost[8][8];
        for (int c = 0; c < nk; ++c) {
            float d[3];
            for (int k = 0; k < 3; ++k)
                d[k] = 0.5f * (bn[kids[c]].box.lo[k] + bn[kids[c]].box.hi[k]) - 0.5f * (box.lo[k] + box.hi[k]);
            for (int s = 0; s < 8; ++s)
                cost[c][s] = ((s & 4) ? d[0] : -d[0]) + ((s & 2) ? d[1] : -d[1]) + ((s & 1) ? d[2] : -d[2]);
        }
        int  slot_child[8];
        bool child_done[8] = {};
        for (int s = 0; s < 8; ++s) slot_child[s] = -1;
        for (int it = 0; it < nk; ++it) {
            float best = -FLT_MAX;
            int   bc = -1, bs = -1;
            for (int c = 0; c < nk; ++c) {
                if (child_done[c]) continue;
                for (int s = 0; s < 8; ++s) {
                    if (slot_child[s] >= 0) continue;
                    if (cost[c][s] > best) { best = cost[c][s]; bc = c; bs = s; }
                }
            }
            slot_child[bs] = bc;
            child_done[bc] = true;
        }

        Node8 node;
        std::memset(&node, 0, sizeof(node));
        node.child_base = (uint32_t)out.nodes.size();
        node.prim_base  = (uint32_t)out.prim_indices.size();
        Aabb cbox[8];
        bool used[8];
        uint32_t prim_off = 0;
        for (int s = 0; s < 8; ++s) {
            used[s] = slot_child[s] >= 0;
            if (!used[s]) continue;
            const BNode& c = bn[kids[slot_child[s]]];
            cbox[s]        = c.box;
            if (c.count == 0) {
                node.imask |= (uint8_t)(1u << s);
                node.meta[s] = (uint8_t)(0x38 | s);
                out.nodes.emplace_back();
                queue.push_back(Work{ (uint32_t)out.nodes.size() - 1, kids[slot_child[s]], w.depth + 1 });
            } else {
                node.meta[s] = (uint8_t)((((1u << c.count) - 1u) << 5) | prim_off);
                for (uint32_t i = 0; i < c.count; ++i) out.prim_indices.push_back(bb.idx[c.first + i]);
                prim_off += c.count;
            }
        }
        encode_node(node, box, cbox, used);
        out.nodes[w.node8] = node;
    }
}

} // namespace igb
