// bvh8_build.cpp -- host builder of the compressed BVH8 (see bvh8.h).
#include "bvh8.h"

#include <algorithm>
#include <cfloat>
#include <chrono>
#include <cstdio>
#include <cmath>
#include <cstring>
#include <atomic>
#include <cstdlib>
#include <deque>
#include <thread>

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
    uint32_t left = 0, right = 0;  // inner: child ids (left == 0: leaf)
    uint32_t first = 0, count = 0; // range in the index array: the leaf's primitives, or everything below an inner node
    bool     leaf() const { return left == 0; }
};

struct BinaryBuilder {
    static constexpr int NB = 32;
    const float*          boxes;
    uint32_t              max_leaf;
    std::vector<uint32_t> idx;
    std::vector<float>    cent; // 3 per primitive
    std::vector<BNode>    nodes;
    bool                  median_only = false; // balanced object-median splits instead of binned SAH (depth fallback)

    struct Task { uint32_t node, begin, end; };

    // Top of the tree on the calling thread until there are enough independent subtrees, then one subtree per task on
    // `threads` workers; each worker appends to its own node vector, which is spliced into `nodes` afterwards (the index
    // ranges of different subtrees are disjoint, so `idx` is partitioned in place without locks).
    void run(uint32_t n, unsigned threads)
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
        const uint32_t    grain = 1u << 14;
        if (threads <= 1 || n < 4 * grain) {
            while (!todo.empty()) {
                const Task t = todo.back();
                todo.pop_back();
                split(t, todo, nodes);
            }
            return;
        }
        // phase 1: split the largest task until there are 8 subtrees per thread (or everything is small)
        std::vector<Task> big{ todo[0] }, ready;
        while (!big.empty() && big.size() + ready.size() < 8 * (size_t)threads) {
            size_t pick = 0;
            for (size_t i = 1; i < big.size(); ++i)
                if (big[i].end - big[i].begin > big[pick].end - big[pick].begin) pick = i;
            const Task t = big[pick];
            big.erase(big.begin() + pick);
            std::vector<Task> kids;
            split(t, kids, nodes, threads);
            for (const Task& k : kids) (k.end - k.begin > grain ? big : ready).push_back(k);
        }
        ready.insert(ready.end(), big.begin(), big.end());
        std::sort(ready.begin(), ready.end(), [](const Task& a, const Task& b) { return a.end - a.begin > b.end - b.begin; });
        // phase 2: subtrees in parallel, local node ids (0 = the subtree's root)
        std::vector<std::vector<BNode>> local(ready.size());
        std::atomic<size_t>             next{ 0 };
        auto worker = [&]() {
            for (;;) {
                const size_t i = next.fetch_add(1);
                if (i >= ready.size()) return;
                std::vector<BNode>& out = local[i];
                out.reserve(2 * (size_t)(ready[i].end - ready[i].begin));
                out.emplace_back();
                std::vector<Task> st{ Task{ 0, ready[i].begin, ready[i].end } };
                while (!st.empty()) {
                    const Task t = st.back();
                    st.pop_back();
                    split(t, st, out);
                }
            }
        };
        std::vector<std::thread> pool;
        for (unsigned t = 1; t < threads; ++t) pool.emplace_back(worker);
        worker();
        for (auto& t : pool) t.join();
        // phase 3: splice (local id 0 replaces the placeholder, local id k > 0 becomes base + k - 1)
        for (size_t i = 0; i < ready.size(); ++i) {
            const uint32_t base = (uint32_t)nodes.size();
            auto           fix  = [&](BNode b) {
                if (!b.leaf()) { b.left += base - 1; b.right += base - 1; }
                return b;
            };
            nodes[ready[i].node] = fix(local[i][0]);
            for (size_t k = 1; k < local[i].size(); ++k) nodes.push_back(fix(local[i][k]));
        }
    }

    // Runs fn(chunk, begin, end) over [begin, end) in `parts` contiguous chunks, on that many threads.
    template <typename F>
    static void chunks(uint32_t begin, uint32_t end, unsigned parts, F fn)
    {
        if (parts <= 1) { fn(0u, begin, end); return; }
        std::vector<std::thread> pool;
        const uint64_t           n = end - begin;
        for (unsigned c = 1; c < parts; ++c)
            pool.emplace_back(fn, c, begin + (uint32_t)(n * c / parts), begin + (uint32_t)(n * (c + 1) / parts));
        fn(0u, begin, begin + (uint32_t)(n / parts));
        for (auto& t : pool) t.join();
    }

    struct Bins {
        Aabb     box[3][NB];
        uint32_t cnt[3][NB];
        Bins()
        {
            for (auto& a : box)
                for (auto& b : a) b.reset();
            std::memset(cnt, 0, sizeof(cnt));
        }
    };

    // Splits task t; `out` is the node vector t.node indexes and new children are appended to.  Three passes over the
    // range (bounds, binning of all three axes at once, partition); each pass of a large range runs on `par` threads.
    void split(const Task& t, std::vector<Task>& todo, std::vector<BNode>& out, unsigned par = 1)
    {
        const uint32_t count = t.end - t.begin;
        if (count < (1u << 17)) par = 1;
        par = std::min(par, 32u);
        Aabb bbs[32], cbs[32];
        chunks(t.begin, t.end, par, [&](unsigned c, uint32_t b, uint32_t e) {
            Aabb bb, cb;
            bb.reset();
            cb.reset();
            for (uint32_t i = b; i < e; ++i) {
                bb.grow(boxes + 6 * (size_t)idx[i]);
                const float* ce = &cent[3 * (size_t)idx[i]];
                for (int k = 0; k < 3; ++k) {
                    cb.lo[k] = std::min(cb.lo[k], ce[k]);
                    cb.hi[k] = std::max(cb.hi[k], ce[k]);
                }
            }
            bbs[c] = bb;
            cbs[c] = cb;
        });
        Aabb bb = bbs[0], cb = cbs[0];
        for (unsigned c = 1; c < par; ++c) { bb.grow(bbs[c]); cb.grow(cbs[c]); }
        out[t.node].box   = bb;
        out[t.node].first = t.begin;
        out[t.node].count = count;
        auto leaf = [&]() { out[t.node].left = out[t.node].right = 0; };
        if (count == 1) { leaf(); return; }
        if (median_only) {
            // Balanced fallback (depth <= ceil(log2 n)): object median along the longest centroid axis.  Used when the SAH tree
            // of a mesh comes out deeper than the traversal stack allows (slivers, exponentially spaced geometry).
            int axis = 0;
            for (int k = 1; k < 3; ++k)
                if (cb.hi[k] - cb.lo[k] > cb.hi[axis] - cb.lo[axis]) axis = k;
            const uint32_t half = t.begin + count / 2;
            if (cb.hi[axis] > cb.lo[axis])
                std::nth_element(idx.begin() + t.begin, idx.begin() + half, idx.begin() + t.end,
                                 [&](uint32_t a, uint32_t b) { return cent[3 * (size_t)a + axis] < cent[3 * (size_t)b + axis]; });
            const uint32_t l = (uint32_t)out.size();
            out.emplace_back();
            out.emplace_back();
            out[t.node].left  = l;
            out[t.node].right = l + 1;
            todo.push_back(Task{ l + 1, half, t.end });
            todo.push_back(Task{ l, t.begin, half });
            return;
        }

        float scale[3];
        bool  any_axis = false;
        for (int k = 0; k < 3; ++k) {
            const float ext = cb.hi[k] - cb.lo[k];
            scale[k]        = ext > 0 ? (float)NB / ext : 0.0f;
            any_axis |= ext > 0;
        }
        float best = FLT_MAX;
        int   best_axis = -1, best_bin = -1;
        if (any_axis) {
            std::vector<Bins> part(par);
            chunks(t.begin, t.end, par, [&](unsigned c, uint32_t b, uint32_t e) {
                Bins& bins = part[c];
                for (uint32_t i = b; i < e; ++i) {
                    const float* pb = boxes + 6 * (size_t)idx[i];
                    const float* ce = &cent[3 * (size_t)idx[i]];
                    for (int k = 0; k < 3; ++k) {
                        if (scale[k] == 0.0f) continue;
                        const int bi = std::min(std::max((int)((ce[k] - cb.lo[k]) * scale[k]), 0), NB - 1);
                        bins.box[k][bi].grow(pb);
                        bins.cnt[k][bi]++;
                    }
                }
            });
            Bins& bins = part[0];
            for (unsigned c = 1; c < par; ++c)
                for (int k = 0; k < 3; ++k)
                    for (int b = 0; b < NB; ++b) {
                        bins.box[k][b].grow(part[c].box[k][b]);
                        bins.cnt[k][b] += part[c].cnt[k][b];
                    }
            for (int axis = 0; axis < 3; ++axis) {
                if (scale[axis] == 0.0f) continue;
                float    rarea[NB];
                uint32_t rcnt[NB];
                Aabb     acc;
                acc.reset();
                uint32_t c = 0;
                for (int b = NB - 1; b > 0; --b) {
                    acc.grow(bins.box[axis][b]);
                    c += bins.cnt[axis][b];
                    rarea[b] = c ? acc.area() : 0.0f;
                    rcnt[b]  = c;
                }
                acc.reset();
                c = 0;
                for (int b = 0; b < NB - 1; ++b) {
                    acc.grow(bins.box[axis][b]);
                    c += bins.cnt[axis][b];
                    if (c == 0 || rcnt[b + 1] == 0) continue;
                    const float cost = acc.area() * (float)c + rarea[b + 1] * (float)rcnt[b + 1];
                    if (cost < best) { best = cost; best_axis = axis; best_bin = b; }
                }
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
            const float lo = cb.lo[best_axis], sc = scale[best_axis];
            auto        goes_left = [&](uint32_t p) {
                const int b = std::min(std::max((int)((cent[3 * (size_t)p + best_axis] - lo) * sc), 0), NB - 1);
                return b <= best_bin;
            };
            if (par <= 1) {
                auto m = std::partition(idx.begin() + t.begin, idx.begin() + t.end, goes_left);
                mid    = (uint32_t)(m - idx.begin());
            } else { // count per chunk, prefix, scatter into a copy: left parts in chunk order, then right parts
                std::vector<uint32_t> src(idx.begin() + t.begin, idx.begin() + t.end), nl(par, 0), lo_at(par), hi_at(par);
                chunks(0, count, par, [&](unsigned c, uint32_t b, uint32_t e) {
                    uint32_t k = 0;
                    for (uint32_t i = b; i < e; ++i) k += goes_left(src[i]) ? 1u : 0u;
                    nl[c] = k;
                });
                uint32_t total_left = 0;
                for (unsigned c = 0; c < par; ++c) { lo_at[c] = total_left; total_left += nl[c]; }
                uint32_t right = total_left;
                for (unsigned c = 0; c < par; ++c) {
                    const uint32_t b = (uint32_t)((uint64_t)count * c / par), e = (uint32_t)((uint64_t)count * (c + 1) / par);
                    hi_at[c] = right;
                    right += (e - b) - nl[c];
                }
                chunks(0, count, par, [&](unsigned c, uint32_t b, uint32_t e) {
                    uint32_t l = t.begin + lo_at[c], r = t.begin + hi_at[c];
                    for (uint32_t i = b; i < e; ++i) (goes_left(src[i]) ? idx[l++] : idx[r++]) = src[i];
                });
                mid = t.begin + total_left;
            }
            if (mid == t.begin || mid == t.end) mid = t.begin + count / 2;
        }
        const uint32_t l = (uint32_t)out.size();
        out.emplace_back();
        out.emplace_back();
        out[t.node].left  = l;
        out[t.node].right = l + 1;
        todo.push_back(Task{ l + 1, mid, t.end });
        todo.push_back(Task{ l, t.begin, mid });
    }
};

// ---- Reinsertion optimisation of the binary tree (Bittner, Hapala, Havran, "Fast Insertion-Based Optimization of Bounding
// Volume Hierarchies", CGF 2013; the reference runs madmann91/bvh's parallel variant of it after its SBVH build,
// src/runtime/bvh/TriBVHAdapter.h:163-202).  One pass: every non-root node whose parent is not the root, in order of decreasing
// surface area, is detached (its sibling takes the parent's place) and re-attached where the induced increase of the tree's
// surface-area sum is smallest -- a best-first search bounded by the best position found so far.  The original position is
// among the candidates, so a reinsertion never makes the sum larger.  Serial; meant for per-mesh trees, not for million-node ones.
#ifndef IGB_BVH_REINSERT_DEFAULT
#define IGB_BVH_REINSERT_DEFAULT 0
#endif
static void reinsertion_pass(std::vector<BNode>& bn, std::vector<uint32_t>& idx, int passes)
{
    const uint32_t n = (uint32_t)bn.size();
    if (n < 8 || passes <= 0) return;
    std::vector<uint32_t> parent(n, 0);
    {
        std::vector<uint32_t> st{ 0 };
        while (!st.empty()) {
            const uint32_t i = st.back();
            st.pop_back();
            if (!bn[i].leaf()) {
                parent[bn[i].left] = parent[bn[i].right] = i;
                st.push_back(bn[i].left);
                st.push_back(bn[i].right);
            }
        }
    }
    auto merged_area = [](const Aabb& a, const Aabb& b) {
        Aabb m = a;
        m.grow(b);
        return m.area();
    };
    auto refit_up = [&](uint32_t i) { // boxes and primitive counts from i to the root
        for (;;) {
            BNode& b = bn[i];
            b.box    = bn[b.left].box;
            b.box.grow(bn[b.right].box);
            b.count = bn[b.left].count + bn[b.right].count;
            if (i == 0) break;
            i = parent[i];
        }
    };
    struct Cand { float induced; uint32_t node; };
    auto worse = [](const Cand& a, const Cand& b) { return a.induced > b.induced; };
    std::vector<Cand> heap;
    for (int pass = 0; pass < passes; ++pass) {
        std::vector<uint32_t> order;
        order.reserve(n);
        for (uint32_t i = 1; i < n; ++i)
            if (parent[i] != 0) order.push_back(i);
        std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return bn[a].box.area() > bn[b].box.area(); });
        for (uint32_t l : order) {
            const uint32_t p = parent[l];
            if (p == 0) continue; // moved under the root by an earlier reinsertion of this pass
            const uint32_t g = parent[p];
            const uint32_t s = bn[p].left == l ? bn[p].right : bn[p].left;
            // detach l: s takes p's place under g
            (bn[g].left == p ? bn[g].left : bn[g].right) = s;
            parent[s]                                     = g;
            refit_up(g);
            // best-first search for the cheapest position
            const Aabb  lb    = bn[l].box;
            const float la    = lb.area();
            float       best  = FLT_MAX;
            uint32_t    where = s;
            heap.clear();
            heap.push_back(Cand{ 0.0f, 0u });
            while (!heap.empty()) {
                std::pop_heap(heap.begin(), heap.end(), worse);
                const Cand c = heap.back();
                heap.pop_back();
                if (c.induced + la >= best) break; // every remaining candidate is at least this expensive
                const float direct = merged_area(bn[c.node].box, lb);
                const float total  = c.induced + direct;
                if (c.node != 0 && total < best) {
                    best  = total;
                    where = c.node;
                }
                const float child_induced = total - bn[c.node].box.area();
                if (!bn[c.node].leaf() && child_induced + la < best) {
                    heap.push_back(Cand{ child_induced, bn[c.node].left });
                    std::push_heap(heap.begin(), heap.end(), worse);
                    heap.push_back(Cand{ child_induced, bn[c.node].right });
                    std::push_heap(heap.begin(), heap.end(), worse);
                }
            }
            // attach: p becomes the parent of (where, l) in where's place
            const uint32_t q = parent[where];
            (bn[q].left == where ? bn[q].left : bn[q].right) = p;
            parent[p]     = q;
            bn[p].left    = where;
            bn[p].right   = l;
            parent[where] = p;
            parent[l]     = p;
            refit_up(p);
        }
    }
    // the primitive ranges of subtrees are no longer contiguous in `idx`: relinearise in depth-first order
    std::vector<uint32_t> fresh;
    fresh.reserve(idx.size());
    struct Frame { uint32_t node; bool done; };
    std::vector<Frame> st{ Frame{ 0, false } };
    while (!st.empty()) {
        Frame f = st.back();
        st.pop_back();
        BNode& b = bn[f.node];
        if (b.leaf()) {
            const uint32_t first = (uint32_t)fresh.size();
            for (uint32_t k = 0; k < b.count; ++k) fresh.push_back(idx[b.first + k]);
            b.first = first;
        } else if (!f.done) {
            st.push_back(Frame{ f.node, true });
            st.push_back(Frame{ b.right, false });
            st.push_back(Frame{ b.left, false });
        } else {
            b.first = bn[b.left].first;
            b.count = bn[b.left].count + bn[b.right].count;
        }
    }
    idx.swap(fresh);
}

// Build switches (environment, read per build so that one binary serves the A/Bs of DESIGN.md):
//   IGB200_BVH_COLLAPSE = dp (default) | greedy      IGB200_BVH_CNODE / IGB200_BVH_CPRIM = SAH constants of the DP
//   IGB200_BVH_THREADS  = worker threads of the binary build (default: hardware concurrency, at most 16)
//   IGB200_BVH_REINSERT = passes of the reinsertion optimisation over the binary tree (0 = off)
struct BuildOptions { int collapse; float c_node, c_prim; unsigned threads; int reinsert; };
BuildOptions build_options()
{
    BuildOptions o{ 1, 1.0f, 0.7f, std::min(16u, std::max(1u, std::thread::hardware_concurrency())), IGB_BVH_REINSERT_DEFAULT };
    if (const char* e = std::getenv("IGB200_BVH_COLLAPSE")) o.collapse = std::strcmp(e, "greedy") == 0 ? 0 : 1;
    if (const char* e = std::getenv("IGB200_BVH_CNODE")) o.c_node = (float)std::atof(e);
    if (const char* e = std::getenv("IGB200_BVH_CPRIM")) o.c_prim = (float)std::atof(e);
    if (const char* e = std::getenv("IGB200_BVH_THREADS")) o.threads = (unsigned)std::max(1, std::atoi(e));
    if (const char* e = std::getenv("IGB200_BVH_REINSERT")) o.reinsert = std::max(0, std::atoi(e));
    return o;
}

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

void build_bvh8(const float* boxes, uint32_t n, uint32_t max_leaf, Bvh8& out, bool balanced)
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

    const BuildOptions opt = build_options();
    // DP collapse decides itself which subtrees (<= max_leaf primitives) become leaves, so the binary tree goes down to
    // single primitives; the greedy collapse takes the binary builder's SAH-terminated leaves as they are.
    using clk = std::chrono::steady_clock;
    const bool verbose = std::getenv("IGB200_BVH_VERBOSE") != nullptr;
    auto       t_last  = clk::now();
    auto       lap     = [&](const char* what) {
        if (!verbose) return;
        const auto now = clk::now();
        std::fprintf(stderr, "[bvh8] %-10s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
        t_last = now;
    };
    BinaryBuilder bb{ boxes, opt.collapse == 1 ? 1u : max_leaf, {}, {}, {} };
    bb.median_only = balanced || std::getenv("IGB200_BVH_BALANCED") != nullptr;
    bb.run(n, opt.threads);
    lap("binary");
    std::vector<BNode>& bn = bb.nodes;
    if (opt.reinsert > 0 && !bb.median_only && n <= (1u << 18)) { // per-mesh trees; a serial pass over millions of nodes takes seconds
        reinsertion_pass(bn, bb.idx, opt.reinsert);
        lap("reinsert");
    }

    // ---- SAH-optimal collapse to 8-wide nodes by dynamic programming over the binary tree (the formulation of Ylitie,
    // Karras, Laine, "Efficient Incoherent Ray Traversal on GPUs Through Compressed Wide BVHs", HPG 2017, section 4.1; the
    // reference's own collapse, NArityBvh.h:65-116, pulls grand-children breadth-first by area like the greedy path below):
    //   cost[n][i-1], i = 1..7 = cheapest way to represent the subtree of n with at most i wide-node children ("roots")
    //   i = 1: min(leaf: A_n * N_n * c_prim if N_n <= max_leaf,  internal: distribute(n, 8) + A_n * c_node)
    //   i > 1: min(distribute(n, i), cost[n][i-2]);   distribute(n, j) = min_{0<k<j} cost[left][k-1] + cost[right][j-k-1]
    struct Dp { float cost[7]; uint8_t choice[7]; uint8_t split8; }; // choice: 0 leaf, 0xFF internal, 0xFE = same as i-1, else k (left gets k)
    std::vector<Dp> dp;
    if (opt.collapse == 1) {
        dp.resize(bn.size());
        // children always have larger ids than their parent within a subtree vector, but splicing breaks global monotonicity:
        // post-order by explicit stack
        std::vector<uint32_t> order;
        order.reserve(bn.size());
        {
            std::vector<uint32_t> st{ 0 };
            while (!st.empty()) {
                const uint32_t i = st.back();
                st.pop_back();
                order.push_back(i);
                if (!bn[i].leaf()) { st.push_back(bn[i].left); st.push_back(bn[i].right); }
            }
        }
        for (size_t oi = order.size(); oi-- > 0;) {
            const uint32_t i = order[oi];
            const BNode&   b = bn[i];
            Dp&            d = dp[i];
            const float    area = b.box.area();
            if (b.leaf()) {
                for (int k = 0; k < 7; ++k) { d.cost[k] = area * (float)b.count * opt.c_prim; d.choice[k] = 0; }
                d.split8 = 0;
                continue;
            }
            const Dp& L = dp[b.left];
            const Dp& R = dp[b.right];
            auto distribute = [&](int j, uint8_t& arg) {
                float best = FLT_MAX;
                for (int k = 1; k < j; ++k) {
                    if (k > 7 || j - k > 7) continue;
                    const float c = L.cost[k - 1] + R.cost[j - k - 1];
                    if (c < best) { best = c; arg = (uint8_t)k; }
                }
                return best;
            };
            uint8_t     k8 = 1;
            const float internal = distribute(8, k8) + area * opt.c_node;
            const float as_leaf  = b.count <= max_leaf ? area * (float)b.count * opt.c_prim : FLT_MAX;
            d.split8    = k8;
            d.cost[0]   = std::min(internal, as_leaf);
            d.choice[0] = as_leaf <= internal ? 0 : 0xFF;
            for (int j = 2; j <= 7; ++j) {
                uint8_t     k = 1;
                const float c = distribute(j, k);
                if (c < d.cost[j - 2]) { d.cost[j - 1] = c; d.choice[j - 1] = k; }
                else { d.cost[j - 1] = d.cost[j - 2]; d.choice[j - 1] = 0xFE; }
            }
        }
    }
    lap("dp");
    // Collects the wide-node children of binary node `b` under a budget of `i` roots, following the DP's decisions.
    struct Kid { uint32_t bnode; bool leaf; };
    auto collect = [&](auto&& self, uint32_t b, int i, Kid* kids, int& nk) -> void {
        const Dp& d = dp[b];
        uint8_t   c = d.choice[i - 1];
        while (c == 0xFE) { --i; c = d.choice[i - 1]; }
        if (i == 1 || c == 0 || c == 0xFF) { kids[nk++] = Kid{ b, c == 0 }; return; }
        self(self, bn[b].left, c, kids, nk);
        self(self, bn[b].right, i - c, kids, nk);
    };

    struct Work { uint32_t node8, bnode, depth; };
    std::deque<Work> queue;
    out.nodes.emplace_back();
    out.prim_indices.reserve(n);
    double sah_nodes = 0.0, sah_prims = 0.0;
    const double root_area = std::max((double)bn[0].box.area(), 1e-300);

    // A leaf root becomes a wide node with one leaf child.
    queue.push_back(Work{ 0, 0, 1 });
    while (!queue.empty()) {
        const Work w = queue.front();
        queue.pop_front();
        out.max_depth = std::max(out.max_depth, w.depth);
        sah_nodes += (double)bn[w.bnode].box.area() / root_area;

        Kid kid[8];
        int nk = 0;
        if (opt.collapse == 1) {
            const bool root_is_leaf = bn[w.bnode].leaf() || (w.bnode == 0 && dp[0].choice[0] == 0);
            if (root_is_leaf) {
                kid[nk++] = Kid{ w.bnode, true };
            } else {
                const uint8_t k = dp[w.bnode].split8;
                collect(collect, bn[w.bnode].left, k, kid, nk);
                collect(collect, bn[w.bnode].right, 8 - k, kid, nk);
            }
        } else {
            // Greedy collapse: open the inner child with the largest surface area until 8 children.
            uint32_t kids[8];
            if (bn[w.bnode].leaf()) {
                kids[nk++] = w.bnode;
            } else {
                kids[nk++] = bn[w.bnode].left;
                kids[nk++] = bn[w.bnode].right;
                while (nk < 8) {
                    int   pick = -1;
                    float pa   = -1.0f;
                    for (int i = 0; i < nk; ++i) {
                        if (!bn[kids[i]].leaf()) {
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
            for (int i = 0; i < nk; ++i) kid[i] = Kid{ kids[i], bn[kids[i]].leaf() };
        }
        uint32_t kids[8];
        for (int i = 0; i < nk; ++i) kids[i] = kid[i].bnode;

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
            if (!kid[slot_child[s]].leaf) {
                node.imask |= (uint8_t)(1u << s);
                node.meta[s] = (uint8_t)(0x38 | s);
                out.nodes.emplace_back();
                queue.push_back(Work{ (uint32_t)out.nodes.size() - 1, kids[slot_child[s]], w.depth + 1 });
            } else {
                node.meta[s] = (uint8_t)((((1u << c.count) - 1u) << 5) | prim_off);
                for (uint32_t i = 0; i < c.count; ++i) out.prim_indices.push_back(bb.idx[c.first + i]);
                prim_off += c.count;
                sah_prims += (double)c.box.area() / root_area * c.count;
            }
        }
        encode_node(node, box, cbox, used);
        out.nodes[w.node8] = node;
    }
    lap("collapse");
    out.sah_node_term = sah_nodes;
    out.sah_prim_term = sah_prims;
}

} // namespace igb
