// TEST INFRASTRUCTURE ONLY (see oracle.h).
// Table parsing + the reference's BVH layout writers, restated:
//   write_node : src/runtime/bvh/BvhNAdapter.h:16-80 (child ids shifted by +1, leaves ~index, unused = +-Inf / 0)
//   Tri4 leaves: src/runtime/bvh/TriBVHAdapter.h:50-116
//   EntityLeaf1: src/runtime/bvh/SceneBVHAdapter.h:56-105
//   mesh table : src/runtime/shape/TriMeshProvider.cpp:577-598 <-> src/artic/shapes/trimesh.art:75-91
#include "oracle.h"
#include "scene.h"

#include <cfloat>
#include <limits>

namespace igo {
namespace {

template <typename LeafWriter>
struct NodeWriter {
    std::vector<Node4>& nodes;
    const NaryBvh&      bvh;
    LeafWriter&         write_leaf;

    void write_node(const NaryNode& node, int parent, size_t child)
    {
        const size_t node_id = nodes.size();
        if (parent >= 0) nodes[parent].child[child] = (int)node_id + 1;
        nodes.emplace_back();
        std::memset(&nodes[node_id], 0, sizeof(Node4));

        for (size_t i = 0; i < (size_t)node.primitive_or_child_count; ++i) {
            const NaryNode& src = bvh.nodes[node.first_child_or_primitive + i];
            for (int k = 0; k < 6; ++k) nodes[node_id].bounds[k][i] = src.bounds[k];
            if (src.is_leaf())
                write_leaf(src, node_id, i);
            else
                write_node(src, (int)node_id, i);
        }
        const float inf = std::numeric_limits<float>::infinity();
        for (size_t i = (size_t)node.primitive_or_child_count; i < 4; ++i) {
            Node4& dst       = nodes[node_id];
            dst.bounds[0][i] = inf;
            dst.bounds[2][i] = inf;
            dst.bounds[4][i] = inf;
            dst.bounds[1][i] = -inf;
            dst.bounds[3][i] = -inf;
            dst.bounds[5][i] = -inf;
            dst.child[i]     = 0;
        }
    }

    void adapt()
    {
        if (bvh.nodes[0].is_leaf()) {
            NaryNode root                 = bvh.nodes[0];
            root.primitive_or_child_count = 1;
            root.first_child_or_primitive = 0;
            write_node(root, -1, 0);
        } else {
            write_node(bvh.nodes[0], -1, 0);
        }
    }
};

bool build_trimesh(const uint8_t* data, uint64_t avail, PrimBvh& out)
{
    if (avail < 48) return false;
    uint32_t hdr[4];
    std::memcpy(hdr, data, 16);
    const uint64_t nf = hdr[0], nv = hdr[1], nn = hdr[2];
    const uint64_t need = 4 * (12 + nv * 4 + nn * 4 + nf * 4);
    if (nf == 0 || avail < need) return false;
    const float*    verts = reinterpret_cast<const float*>(data) + 12;
    const uint32_t* idx   = reinterpret_cast<const uint32_t*>(data) + 12 + nv * 4 + nn * 4;

    out.flat.resize(nf);
    std::vector<float> boxes(nf * 6), centers(nf * 3);
    for (uint64_t i = 0; i < nf; ++i) {
        const uint32_t i0 = idx[4 * i + 0], i1 = idx[4 * i + 1], i2 = idx[4 * i + 2];
        if (i0 >= nv || i1 >= nv || i2 >= nv) return false;
        const V3 p0 = v3(verts[4 * i0], verts[4 * i0 + 1], verts[4 * i0 + 2]);
        const V3 p1 = v3(verts[4 * i1], verts[4 * i1 + 1], verts[4 * i1 + 2]);
        const V3 p2 = v3(verts[4 * i2], verts[4 * i2 + 1], verts[4 * i2 + 2]);
        // madmann91/bvh v1 Triangle: e1 = p0 - p1, e2 = p2 - p0, n = cross(e1, e2)
        // (convention stated in src/artic/traversal/intersection.art:10-17)
        FlatTri& t = out.flat[i];
        t.v0       = p0;
        t.e1       = sub(p0, p1);
        t.e2       = sub(p2, p0);
        t.n        = cross(t.e1, t.e2);
        t.prim_id  = (int32_t)i;
        const float px[3] = { p0.x, p1.x, p2.x }, py[3] = { p0.y, p1.y, p2.y }, pz[3] = { p0.z, p1.z, p2.z };
        boxes[6 * i + 0]   = std::fmin(px[0], std::fmin(px[1], px[2]));
        boxes[6 * i + 1]   = std::fmin(py[0], std::fmin(py[1], py[2]));
        boxes[6 * i + 2]   = std::fmin(pz[0], std::fmin(pz[1], pz[2]));
        boxes[6 * i + 3]   = std::fmax(px[0], std::fmax(px[1], px[2]));
        boxes[6 * i + 4]   = std::fmax(py[0], std::fmax(py[1], py[2]));
        boxes[6 * i + 5]   = std::fmax(pz[0], std::fmax(pz[1], pz[2]));
        centers[3 * i + 0] = (p0.x + p1.x + p2.x) * (1.0f / 3.0f);
        centers[3 * i + 1] = (p0.y + p1.y + p2.y) * (1.0f / 3.0f);
        centers[3 * i + 2] = (p0.z + p1.z + p2.z) * (1.0f / 3.0f);
    }

    BinBvh bin;
    build_binary_sah(boxes.data(), centers.data(), (uint32_t)nf, 8, bin);
    NaryBvh nary;
    collapse_to_narity(bin, 4, nary);

    auto write_leaf = [&](const NaryNode& node, size_t parent, size_t child) {
        out.nodes[parent].child[child] = ~(int)out.tris.size();
        const size_t ref_count         = node.primitive_count();
        for (size_t i = 0; i < ref_count; i += 4) {
            const size_t c = i + 4 <= ref_count ? 4 : ref_count - i;
            Tri4 tri;
            std::memset(&tri, 0, sizeof(Tri4));
            for (size_t j = 0; j < c; ++j) {
                const FlatTri& in = out.flat[nary.primitive_indices[node.first_child_or_primitive + i + j]];
                tri.v0[0][j] = in.v0.x; tri.v0[1][j] = in.v0.y; tri.v0[2][j] = in.v0.z;
                tri.e1[0][j] = in.e1.x; tri.e1[1][j] = in.e1.y; tri.e1[2][j] = in.e1.z;
                tri.e2[0][j] = in.e2.x; tri.e2[1][j] = in.e2.y; tri.e2[2][j] = in.e2.z;
                tri.n[0][j]  = in.n.x;  tri.n[1][j]  = in.n.y;  tri.n[2][j]  = in.n.z;
                tri.prim_id[j] = in.prim_id;
            }
            for (size_t j = c; j < 4; ++j) tri.prim_id[j] = (int32_t)0xFFFFFFFFu;
            out.tris.push_back(tri);
        }
        out.tris.back().prim_id[3] |= (int32_t)0x80000000u;
    };
    NodeWriter<decltype(write_leaf)> w{ out.nodes, nary, write_leaf };
    w.adapt();
    return true;
}

void build_scene_bvh(const std::vector<igb200_entity_leaf>& in, SceneBvh& out)
{
    const uint32_t n = (uint32_t)in.size();
    std::vector<float> boxes(n * 6), centers(n * 3);
    for (uint32_t i = 0; i < n; ++i) {
        for (int k = 0; k < 3; ++k) {
            boxes[6 * i + k]     = in[i].bbox_min[k];
            boxes[6 * i + 3 + k] = in[i].bbox_max[k];
            centers[3 * i + k]   = (in[i].bbox_min[k] + in[i].bbox_max[k]) * 0.5f;
        }
    }
    BinBvh bin;
    build_binary_sah(boxes.data(), centers.data(), n, 1, bin);
    NaryBvh nary;
    collapse_to_narity(bin, 4, nary);

    auto write_leaf = [&](const NaryNode& node, size_t parent, size_t child) {
        out.nodes[parent].child[child] = ~(int)out.leaves.size();
        for (size_t i = 0; i < node.primitive_count(); ++i)
            out.leaves.push_back(in[nary.primitive_indices[node.first_child_or_primitive + i]]);
        out.leaves.back().entity_id |= (int32_t)0x80000000u; // sentinel
    };
    NodeWriter<decltype(write_leaf)> w{ out.nodes, nary, write_leaf };
    w.adapt();
}

} // namespace
} // namespace igo

extern "C" igo_scene* igo_scene_create(const igb200_scene* sc)
{
    using namespace igo;
    if (!sc || (sc->shape_count && (!sc->shape_lookups || !sc->shape_data))) return nullptr;
    igo_scene* s = new igo_scene();
    s->shapes.resize(sc->shape_count);
    s->spheres.resize(sc->shape_count, Sphere{ v3(0, 0, 0), 0.0f });
    s->shape_type.resize(sc->shape_count);
    s->lookups.assign(sc->shape_lookups, sc->shape_lookups + sc->shape_count);
    s->shape_data.assign(sc->shape_data, sc->shape_data + sc->shape_bytes);
    if (sc->entities) s->entities.assign(sc->entities, sc->entities + 36 * sc->entity_count);
    std::memcpy(s->bbox, sc->scene_bbox, sizeof(s->bbox));
    for (uint64_t i = 0; i < sc->shape_count; ++i) {
        const igb200_lookup_entry& e = sc->shape_lookups[i];
        s->shape_type[i]             = e.type_id;
        if (e.offset > sc->shape_bytes) { delete s; return nullptr; }
        const uint8_t* data  = sc->shape_data + e.offset;
        const uint64_t avail = sc->shape_bytes - e.offset;
        if (e.type_id == IGB200_PRIM_TRIMESH) {
            if (!build_trimesh(data, avail, s->shapes[i])) { delete s; return nullptr; }
        } else if (e.type_id == IGB200_PRIM_SPHERE) {
            if (avail < 16) { delete s; return nullptr; }
            float f[4];
            std::memcpy(f, data, 16);
            s->spheres[i] = Sphere{ v3(f[0], f[1], f[2]), f[3] };
        } else {
            delete s;
            return nullptr;
        }
    }
    for (int k = 0; k < 2; ++k) {
        for (uint64_t i = 0; i < sc->leaf_count[k]; ++i) {
            igb200_entity_leaf l = sc->leaves[k][i];
            l.entity_id &= 0x7FFFFFFF;
            if (l.shape_id < 0 || (uint64_t)l.shape_id >= sc->shape_count || s->shape_type[l.shape_id] != (uint32_t)k) {
                delete s;
                return nullptr;
            }
            s->flat_leaves[k].push_back(l);
        }
        if (!s->flat_leaves[k].empty()) build_scene_bvh(s->flat_leaves[k], s->scene[k]);
    }
    return s;
}

extern "C" void igo_scene_destroy(igo_scene* s) { delete s; }

extern "C" void igo_scene_info(const igo_scene* s, uint64_t* node4_count, uint64_t* tri4_count, uint64_t* tri_count)
{
    uint64_t n = 0, t4 = 0, t = 0;
    for (auto& p : s->shapes) { n += p.nodes.size(); t4 += p.tris.size(); t += p.flat.size(); }
    for (int k = 0; k < 2; ++k) n += s->scene[k].nodes.size();
    if (node4_count) *node4_count = n;
    if (tri4_count) *tri4_count = t4;
    if (tri_count) *tri_count = t;
}
