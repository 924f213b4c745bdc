#include "SceneLoader.h"

#include "Json.h"
#include "PngReader.h"
#include "SceneMesh.h"
#include "SkyModel.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <functional>
#include <limits>
#include <map>
#include <set>
#include <sstream>

namespace IG {
using namespace loader;
using json::Value;

namespace {
constexpr double kPiD = 3.14159265358979323846;
constexpr float  kPiF = 3.14159265358979323846f;

[[noreturn]] void unsupported(const std::string& what) { throw std::runtime_error(what + " is outside the supported subset"); }

std::string dirName(const std::string& path)
{
    const size_t p = path.find_last_of('/');
    return p == std::string::npos ? std::string(".") : path.substr(0, p);
}
std::string joinPath(const std::string& base, const std::string& file)
{
    if (!file.empty() && file[0] == '/') return file;
    return base + "/" + file;
}
std::string lower(std::string s)
{
    for (char& c : s) c = (char)std::tolower((unsigned char)c);
    return s;
}
bool endsWith(const std::string& s, const char* suffix)
{
    const size_t n = std::strlen(suffix);
    return s.size() >= n && s.compare(s.size() - n, n, suffix) == 0;
}
std::string readText(const std::string& path)
{
    std::ifstream in(path);
    if (!in) throw std::runtime_error("cannot open " + path);
    std::ostringstream ss;
    ss << in.rdbuf();
    return ss.str();
}

// ------------------------------------------------------------------------------------------------ value helpers
std::vector<double> numbers(const Value& v)
{
    std::vector<double> out;
    if (v.isNumber()) {
        out.push_back(v.number);
    } else if (v.isArray()) {
        for (const Value& e : v.array) {
            if (e.isArray()) {
                for (const Value& f : e.array) out.push_back(f.number); // nested rows of a matrix
            } else {
                out.push_back(e.number);
            }
        }
    }
    return out;
}
Vec3d vec3(const Value* v, const Vec3d& def)
{
    if (!v) return def;
    const auto n = numbers(*v);
    if (n.size() >= 3) return Vec3d{ n[0], n[1], n[2] };
    if (n.size() == 1) return Vec3d{ n[0], n[0], n[0] };
    return def;
}
Vec3f vec3f(const Value* v, const Vec3d& def)
{
    const Vec3d d = vec3(v, def);
    return Vec3f{ (float)d[0], (float)d[1], (float)d[2] };
}

// A colour property: number, [r, g, b], or the constant PExpr "color(r, g, b[, a])" (loader/Transpiler.cpp); any other
// expression is refused.
Vec3f color(const Value* v, double def)
{
    if (!v || v->isNull()) return Vec3f{ (float)def, (float)def, (float)def };
    if (v->isString()) {
        const std::string& s = v->string;
        const size_t       a = s.find("color("), b = s.rfind(')');
        if (a == std::string::npos || b == std::string::npos || b < a) unsupported("expression '" + s + "'");
        std::vector<double> c;
        std::istringstream  ss(s.substr(a + 6, b - a - 6));
        std::string         tok;
        while (std::getline(ss, tok, ',')) {
            char*        end = nullptr;
            const double x   = std::strtod(tok.c_str(), &end);
            if (end == tok.c_str()) unsupported("expression '" + s + "'");
            c.push_back(x);
        }
        if (c.size() < 3) unsupported("expression '" + s + "'");
        return Vec3f{ (float)c[0], (float)c[1], (float)c[2] };
    }
    const auto n = numbers(*v);
    if (n.size() == 1) return Vec3f{ (float)n[0], (float)n[0], (float)n[0] };
    if (n.size() >= 3) return Vec3f{ (float)n[0], (float)n[1], (float)n[2] };
    return Vec3f{ (float)def, (float)def, (float)def };
}

// ------------------------------------------------------------------------------------------------ transforms
// Parser.cpp:124-243: row-major 3x3 / 3x4 / 4x4 arrays, or op lists multiplied left to right in member order.
Mat4d lookAt(const Vec3d& eye, const Vec3d& center, const Vec3d& up)
{
    auto nrm = [](const Vec3d& a) { return std::sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]); };
    auto crs = [](const Vec3d& a, const Vec3d& b) { return Vec3d{ a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0] }; };
    Vec3d f{ center[0] - eye[0], center[1] - eye[1], center[2] - eye[2] };
    const double fl = nrm(f);
    f               = fl > 0 ? Vec3d{ f[0] / fl, f[1] / fl, f[2] / fl } : Vec3d{ 0, 0, 1 };
    const double ul = nrm(up);
    Vec3d        u{ up[0] / ul, up[1] / ul, up[2] / ul };
    Vec3d        s  = crs(f, u);
    const double sl = nrm(s);
    s               = Vec3d{ s[0] / sl, s[1] / sl, s[2] / sl };
    u               = crs(s, f);
    Mat4d m         = identity4();
    for (int i = 0; i < 3; ++i) {
        m[i][0] = s[i];
        m[i][1] = u[i];
        m[i][2] = f[i];
        m[i][3] = eye[i];
    }
    return m;
}
Mat4d rotAxis(int axis, double deg)
{
    const double a = deg * (kPiD / 180.0);
    const double c = std::cos(a), s = std::sin(a);
    Mat4d        m = identity4();
    const int    ij[3][2] = { { 1, 2 }, { 2, 0 }, { 0, 1 } };
    const int    i = ij[axis][0], j = ij[axis][1];
    m[i][i] = c;
    m[i][j] = -s;
    m[j][i] = s;
    m[j][j] = c;
    return m;
}

Mat4d parseTransform(const Value* t)
{
    if (!t || t->isNull()) return identity4();
    std::vector<const Value*> ops;
    if (t->isObject()) {
        ops.push_back(t);
    } else if (t->isArray() && !t->array.empty() && t->array[0].isObject()) {
        for (const Value& e : t->array) ops.push_back(&e);
    }
    if (!ops.empty()) {
        Mat4d m = identity4();
        for (const Value* op : ops)
            for (const json::Member& kv : op->members) {
                const std::string& k = kv.first;
                const Value&       v = kv.second;
                Mat4d              o = identity4();
                if (k == "translate") {
                    const Vec3d d = vec3(&v, { 0, 0, 0 });
                    for (int i = 0; i < 3; ++i) o[i][3] = d[i];
                } else if (k == "scale") {
                    const Vec3d d = vec3(&v, { 1, 1, 1 });
                    for (int i = 0; i < 3; ++i) o[i][i] = d[i];
                } else if (k == "rotate") {
                    const Vec3d d = vec3(&v, { 0, 0, 0 });
                    o             = mul(mul(rotAxis(0, d[0]), rotAxis(1, d[1])), rotAxis(2, d[2]));
                } else if (k == "qrotate") { // Eigen::Quaternionf(w, x, y, z), Parser.cpp:96-105,192-193
                    const auto q = numbers(v);
                    if (q.size() != 4) throw std::runtime_error("qrotate needs four numbers");
                    const double qw = q[0], qx = q[1], qy = q[2], qz = q[3];
                    const double r[3][3] = { { 1 - 2 * (qy * qy + qz * qz), 2 * (qx * qy - qz * qw), 2 * (qx * qz + qy * qw) },
                                             { 2 * (qx * qy + qz * qw), 1 - 2 * (qx * qx + qz * qz), 2 * (qy * qz - qx * qw) },
                                             { 2 * (qx * qz - qy * qw), 2 * (qy * qz + qx * qw), 1 - 2 * (qx * qx + qy * qy) } };
                    for (int i = 0; i < 3; ++i)
                        for (int j = 0; j < 3; ++j) o[i][j] = r[i][j];
                } else if (k == "lookat") {
                    const Vec3d origin = vec3(v.find("origin"), { 0, 0, 0 });
                    Vec3d       target;
                    if (v.has("direction")) {
                        const Vec3d d = vec3(v.find("direction"), { 0, 0, 1 });
                        target        = Vec3d{ d[0] + origin[0], d[1] + origin[1], d[2] + origin[2] };
                    } else {
                        target = vec3(v.find("target"), { 0, 1, 0 });
                    }
                    o = lookAt(origin, target, vec3(v.find("up"), { 0, 0, 1 }));
                } else if (k == "matrix") {
                    o = parseTransform(&v);
                } else {
                    throw std::runtime_error("unsupported transform op " + k);
                }
                m = mul(m, o);
            }
        return m;
    }
    const auto a = numbers(*t);
    Mat4d      m = identity4();
    if (a.size() == 16) {
        for (int i = 0; i < 4; ++i)
            for (int j = 0; j < 4; ++j) m[i][j] = a[(size_t)(4 * i + j)];
    } else if (a.size() == 12) {
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 4; ++j) m[i][j] = a[(size_t)(4 * i + j)];
    } else if (a.size() == 9) {
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) m[i][j] = a[(size_t)(3 * i + j)];
    } else {
        throw std::runtime_error("transform must have 9, 12 or 16 numbers");
    }
    return m;
}

// transform.linear().transpose().inverse() as a 3x3 (EnvironmentLight.cpp:44), float32, row major
std::array<float, 9> envTransform(const Value* t)
{
    const Mat4d m  = parseTransform(t);
    Mat4d       tr = identity4();
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) tr[i][j] = m[j][i];
    const Mat4d          inv = inverse(tr);
    std::array<float, 9> out{};
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) out[(size_t)(3 * i + j)] = (float)inv[i][j];
    return out;
}

// ------------------------------------------------------------------------------------------------ externals
// Parser.cpp:446: externals are read first "such that the current file can replace parts of it later on"; objects live
// in name-keyed maps (Scene.h:43-48), so a later definition of a name replaces the earlier one in place.
const char* const kListSections[] = { "shapes", "bsdfs", "textures", "entities", "lights", "media" };

Value mergeExternals(Value scene, const std::string& base)
{
    const Value* ext = scene.find("externals");
    if (!ext || !ext->isArray()) return scene;
    const std::vector<Value> externals = ext->array;
    for (const Value& e : externals) {
        const std::string path = joinPath(base, e.str("filename", ""));
        Value             sub  = mergeExternals(json::parse(readText(path)), dirName(path));
        for (const json::Member& kv : sub.members) {
            if (kv.second.isArray()) {
                std::vector<Value> merged;
                std::map<std::string, size_t> slot;
                std::vector<Value> all = kv.second.array;
                if (const Value* own = scene.find(kv.first); own && own->isArray())
                    for (const Value& o : own->array) all.push_back(o);
                for (const Value& obj : all) {
                    const Value* name = obj.isObject() ? obj.find("name") : nullptr;
                    if (name && name->isString() && slot.count(name->string)) {
                        merged[slot[name->string]] = obj;
                    } else {
                        if (name && name->isString()) slot[name->string] = merged.size();
                        merged.push_back(obj);
                    }
                }
                bool found = false;
                for (json::Member& m : scene.members)
                    if (m.first == kv.first) {
                        m.second.type  = Value::Array;
                        m.second.array = merged;
                        found          = true;
                        break;
                    }
                if (!found) {
                    Value arr;
                    arr.type  = Value::Array;
                    arr.array = merged;
                    scene.members.emplace_back(kv.first, arr);
                }
            } else if (!scene.has(kv.first)) {
                scene.members.push_back(kv);
            }
        }
    }
    (void)kListSections;
    return scene;
}

// ------------------------------------------------------------------------------------------------ scene tables
// DynTable "shapes" (TriMeshProvider.cpp:574-602, SphereProvider.cpp:33-45, DynTable.h:15-24), FixTable "entities"
// (LoaderEntity.cpp:150-176), EntityLeaf1 (SceneBVHAdapter.h:73-105).
struct TableBuilder {
    std::vector<uint8_t>                 data;
    std::vector<LookupEntry>             lookups;
    std::vector<std::pair<Vec3f, Vec3f>> shape_bbox;
    std::vector<float>                   entities;
    std::vector<igb200_entity_leaf>      leaves[2];
    std::map<int, int>                   materials;

    template <typename T>
    void append(const T* p, size_t n)
    {
        const uint8_t* b = reinterpret_cast<const uint8_t*>(p);
        data.insert(data.end(), b, b + n * sizeof(T));
    }
    void addLookup(uint32_t type_id)
    {
        if (!data.empty()) data.resize(data.size() + (16 - data.size() % 16), 0); // pads even when already aligned, like the reference
        lookups.push_back(LookupEntry{ type_id, 0u, (uint64)data.size() });
    }
    int addTrimesh(const Mesh& m)
    {
        if (m.v.empty() || m.f.empty()) throw std::runtime_error("mesh needs vertices and faces");
        Vec3f lo = m.v[0], hi = m.v[0];
        for (const Vec3f& p : m.v)
            for (int k = 0; k < 3; ++k) {
                lo[k] = std::min(lo[k], p[k]);
                hi[k] = std::max(hi[k], p[k]);
            }
        for (int k = 0; k < 3; ++k) { // bbox.inflate(1e-5f), TriMeshProvider.cpp:546-547
            lo[k] -= 1e-5f;
            hi[k] += 1e-5f;
        }
        addLookup(IGB200_PRIM_TRIMESH);
        const uint32_t hdr[4]  = { (uint32_t)m.f.size(), (uint32_t)m.v.size(), (uint32_t)m.n.size(), (uint32_t)m.t.size() };
        const float    bbox[8] = { lo[0], lo[1], lo[2], 0, hi[0], hi[1], hi[2], 0 };
        append(hdr, 4);
        append(bbox, 8);
        for (const Vec3f& p : m.v) {
            const float v4[4] = { p[0], p[1], p[2], 0 };
            append(v4, 4);
        }
        for (const Vec3f& p : m.n) {
            const float v4[4] = { p[0], p[1], p[2], 0 };
            append(v4, 4);
        }
        for (const auto& f : m.f) {
            const uint32_t f4[4] = { f[0], f[1], f[2], 0 };
            append(f4, 4);
        }
        for (const auto& t : m.t) append(t.data(), 2);
        shape_bbox.emplace_back(lo, hi);
        return (int)lookups.size() - 1;
    }
    int addSphere(const Vec3f& c, float r)
    {
        if (!(r > 0)) throw std::runtime_error("sphere radius must be positive"); // SphereProvider.cpp:23-26
        addLookup(IGB200_PRIM_SPHERE);
        const float rec[4] = { c[0], c[1], c[2], r };
        append(rec, 4);
        shape_bbox.emplace_back(Vec3f{ c[0] - r - 1e-5f, c[1] - r - 1e-5f, c[2] - r - 1e-5f }, Vec3f{ c[0] + r + 1e-5f, c[1] + r + 1e-5f, c[2] + r + 1e-5f });
        return (int)lookups.size() - 1;
    }
    int addEntity(int shape_id, Mat4d m, int material, uint32_t flags)
    {
        m[3] = { 0, 0, 0, 1 }; // Transformf::makeAffine
        const Mat4d inv = inverse(m);
        Mat4d       lin = identity4();
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) lin[i][j] = m[i][j];
        const Mat4d linv = inverse(lin);
        const int   mat_id = materials.emplace(material, (int)materials.size()).first->second;
        const int   ent_id = (int)(entities.size() / 36);
        float       row[36] = {};
        float       to_local[12], to_global[12];
        for (int c = 0; c < 4; ++c)
            for (int r = 0; r < 3; ++r) { // column major (Serializer::write(m, true))
                to_local[3 * c + r]  = (float)inv[r][c];
                to_global[3 * c + r] = (float)m[r][c];
            }
        std::memcpy(row, to_local, sizeof(to_local));
        std::memcpy(row + 12, to_global, sizeof(to_global));
        for (int c = 0; c < 3; ++c)
            for (int r = 0; r < 3; ++r) row[24 + 3 * c + r] = (float)linv[c][r]; // normal matrix = (M^-1)^T, column major
        const uint32_t ids[3] = { (uint32_t)shape_id, (uint32_t)mat_id, 0u };
        std::memcpy(row + 33, ids, sizeof(ids));
        entities.insert(entities.end(), row, row + 36);

        const Vec3f lo = shape_bbox[(size_t)shape_id].first, hi = shape_bbox[(size_t)shape_id].second;
        igb200_entity_leaf leaf{};
        for (int i = 0; i < 8; ++i) { // entity box = shape box transformed corner-wise (math/BoundingBox.h:84-95)
            const float c[3] = { (i & 1) ? hi[0] : lo[0], (i & 2) ? hi[1] : lo[1], (i & 4) ? hi[2] : lo[2] };
            for (int r = 0; r < 3; ++r) {
                const float w = c[0] * to_global[r] + c[1] * to_global[3 + r] + c[2] * to_global[6 + r] + to_global[9 + r];
                leaf.bbox_min[r] = i == 0 ? w : std::min(leaf.bbox_min[r], w);
                leaf.bbox_max[r] = i == 0 ? w : std::max(leaf.bbox_max[r], w);
            }
        }
        leaf.entity_id = ent_id;
        leaf.shape_id  = shape_id;
        std::memcpy(leaf.local, to_local, sizeof(to_local));
        leaf.flags  = flags;
        leaf.mat_id = mat_id;
        leaves[lookups[(size_t)shape_id].TypeID].push_back(leaf);
        return ent_id;
    }
    void finish(SceneFile& out)
    {
        SceneDatabase& db = out.database;
        float          bb[6] = { 0, 0, 0, 0, 0, 0 };
        bool           first = true;
        std::vector<int32> per_mat(materials.size(), 0);
        for (int k = 0; k < 2; ++k)
            for (const auto& l : leaves[k]) {
                for (int r = 0; r < 3; ++r) {
                    bb[r]     = first ? l.bbox_min[r] : std::min(bb[r], l.bbox_min[r]);
                    bb[3 + r] = first ? l.bbox_max[r] : std::max(bb[3 + r], l.bbox_max[r]);
                }
                first = false;
                per_mat[(size_t)l.mat_id]++;
            }
        db.DynTables["shapes"].assign(lookups, data);
        std::vector<uint8> ent(entities.size() * 4);
        if (!ent.empty()) std::memcpy(ent.data(), entities.data(), ent.size());
        db.FixTables["entities"].assign(entities.size() / 36, ent);
        const char* names[2] = { "trimesh", "sphere" };
        for (int k = 0; k < 2; ++k) {
            std::vector<uint8> raw(leaves[k].size() * sizeof(igb200_entity_leaf));
            if (!raw.empty()) std::memcpy(raw.data(), leaves[k].data(), raw.size());
            db.SceneBVHs[names[k]].Leaves = raw;
        }
        for (int r = 0; r < 3; ++r) {
            db.SceneBBox.min(r) = bb[r];
            db.SceneBBox.max(r) = bb[3 + r];
        }
        db.MaterialCount        = materials.size();
        out.entity_per_material = per_mat;
    }
};

// ------------------------------------------------------------------------------------------------ shading tables
// What the reference bakes into generated shader source (ShaderUtils.cpp:95-122, LoaderLight.cpp:106-250,
// PathTechnique.cpp:36-79, PerspectiveCamera.cpp:27-67) travels as POD tables (include/ignis_b200.h).
struct ShadingBuilder {
    ShadingTables& sh;
    explicit ShadingBuilder(ShadingTables& s) : sh(s)
    {
        sh.technique            = igb200_technique{};
        sh.technique.type       = IGB200_TECH_PATH;
        sh.technique.max_depth  = 64;
        sh.technique.min_depth  = 2;
        sh.technique.enable_nee = 1;
    }
    static void set3(float* dst, const Vec3f& v)
    {
        for (int k = 0; k < 3; ++k) dst[k] = v[k];
    }
    size_t blobAppend(const void* p, size_t n)
    {
        while (sh.blob.size() % 16) sh.blob.push_back(0);
        const size_t   off = sh.blob.size();
        const uint8_t* b   = static_cast<const uint8_t*>(p);
        sh.blob.insert(sh.blob.end(), b, b + n);
        return off;
    }
    int bsdf(uint32_t type, const Vec3f& c = { 0, 0, 0 }, const Vec3f& c2 = { 0, 0, 0 }, std::initializer_list<float> p = {})
    {
        igb200_bsdf b{};
        b.type = type;
        set3(b.color, c);
        set3(b.color2, c2);
        int i = 0;
        for (float x : p) b.p[i++] = x;
        sh.bsdfs.push_back(b);
        return (int)sh.bsdfs.size() - 1;
    }
    static void texTransform(igb200_texture& t, const Mat4d* m)
    {
        double a[3][3] = { { 1, 0, 0 }, { 0, 1, 0 }, { 0, 0, 1 } };
        if (m) { // LoaderUtils::inlineTransformAs2d: the 2-D affine part
            for (int i = 0; i < 2; ++i) {
                for (int j = 0; j < 2; ++j) a[i][j] = (*m)[i][j];
                a[i][2] = (*m)[i][3];
            }
        }
        for (int c = 0; c < 3; ++c)
            for (int r = 0; r < 3; ++r) t.transform[3 * c + r] = (float)a[r][c];
    }
    int checkerboard(float sx, float sy, const Vec3f& c0, const Vec3f& c1, const Mat4d* m)
    {
        igb200_texture t{};
        t.type = IGB200_TEX_CHECKERBOARD;
        set3(t.color0, c0);
        set3(t.color1, c1);
        t.scale[0] = sx;
        t.scale[1] = sy;
        texTransform(t, m);
        sh.textures.push_back(t);
        return (int)sh.textures.size() - 1;
    }
    int image(const void* px, uint32_t w, uint32_t h, uint32_t format, uint32_t filter, uint32_t wrap_u, uint32_t wrap_v, const Mat4d* m, bool opaque)
    {
        igb200_texture t{};
        t.type   = IGB200_TEX_IMAGE;
        t.filter = filter;
        t.wrap_u = wrap_u;
        t.wrap_v = wrap_v;
        t.opaque = opaque ? 1u : 0u;
        t.format = format;
        t.width  = w;
        t.height = h;
        t.offset = blobAppend(px, (size_t)w * h * (format == IGB200_IMAGE_RGBA8 ? 4 : 16));
        texTransform(t, m);
        sh.textures.push_back(t);
        return (int)sh.textures.size() - 1;
    }
    uint32_t cdf2d(const std::vector<float>& marginal, const std::vector<float>& conditional, uint32_t w, uint32_t h)
    {
        std::vector<uint8_t> raw(16 + 4 * (marginal.size() + conditional.size()));
        const uint32_t       head[4] = { w, h, 0, 0 };
        std::memcpy(raw.data(), head, 16);
        std::memcpy(raw.data() + 16, marginal.data(), 4 * marginal.size());
        std::memcpy(raw.data() + 16 + 4 * marginal.size(), conditional.data(), 4 * conditional.size());
        return (uint32_t)(blobAppend(raw.data(), raw.size()) / 4);
    }
    int light(std::vector<igb200_light>& table, uint32_t type, const Vec3f& c, const std::vector<float>& v, int entity = -1)
    {
        igb200_light l{};
        l.type   = type;
        l.entity = entity;
        set3(l.color, c);
        for (size_t i = 0; i < v.size() && i < 16; ++i) l.v[i] = v[i];
        table.push_back(l);
        return (int)table.size() - 1;
    }
};

// Image::loadAsPacked (src/runtime/Image.cpp:565-655): packed RGBA8, bottom row first, sRGB bytes mapped to linear
// (byte_color_to_linear, Image.cpp:31-42) unless `linear`.
std::vector<uint32_t> loadPackedImage(const std::string& path, bool linear, uint32_t& w, uint32_t& h, bool& opaque)
{
    PngImage    img;
    std::string err;
    if (!endsWith(lower(path), ".png")) unsupported("image file '" + path + "' (8-bit PNG only)");
    if (!loadPng8(path, img, err)) throw std::runtime_error(err);
    uint8_t lut[256];
    for (int c = 0; c < 256; ++c) {
        if (linear) {
            lut[c] = (uint8_t)c;
            continue;
        }
        const float x   = (float)c / 255.0f;
        const float lin = x <= 0.04045f ? x / 12.92f : std::pow((x + 0.055f) / 1.055f, 2.4f);
        lut[c]          = (uint8_t)std::min(255.0f, std::floor(lin * 255.0f));
    }
    w      = img.width;
    h      = img.height;
    opaque = img.opaque;
    std::vector<uint32_t> px((size_t)w * h);
    for (uint32_t y = 0; y < h; ++y)
        for (uint32_t x = 0; x < w; ++x) {
            const uint8_t* p = img.rgba.data() + 4 * ((size_t)(h - 1 - y) * w + x); // stbi_set_flip_vertically_on_load(1)
            px[(size_t)y * w + x] = (uint32_t)lut[p[0]] | ((uint32_t)lut[p[1]] << 8) | ((uint32_t)lut[p[2]] << 16) | ((uint32_t)p[3] << 24);
        }
    return px;
}

// CDF::computeForImage (src/runtime/CDF.cpp:42-133): image (h, w, 4) float32 -> marginal[h], conditional[h * w]
void computeCdf2d(const float* img, uint32_t w, uint32_t h, bool premultiply_sin, bool compensate, std::vector<float>& marginal, std::vector<float>& cond)
{
    float defect[3] = { 0, 0, 0 };
    if (compensate) {
        for (int c = 0; c < 3; ++c) {
            float s = 0.0f;
            for (size_t i = 0; i < (size_t)w * h; ++i) s += std::max(img[4 * i + c], 0.0f) / (float)w;
            defect[c] = s / (float)h;
        }
    }
    cond.assign((size_t)w * h, 0.0f);
    marginal.assign(h, 0.0f);
    std::vector<float> sums(h);
    for (uint32_t y = 0; y < h; ++y) {
        float acc = 0.0f;
        for (uint32_t x = 0; x < w; ++x) {
            const float* p   = img + 4 * ((size_t)y * w + x);
            const float  lum = (std::max(p[0] - defect[0], 0.0f) + std::max(p[1] - defect[1], 0.0f) + std::max(p[2] - defect[2], 0.0f)) / 3.0f;
            acc += lum;
            cond[(size_t)y * w + x] = acc;
        }
        sums[y]     = acc;
        marginal[y] = premultiply_sin ? acc * std::sin(kPiF * ((float)y + 0.5f) / (float)h) : acc;
    }
    for (uint32_t y = 0; y < h; ++y) {
        float* row = cond.data() + (size_t)y * w;
        if (sums[y] > 1e-5f) {
            const float inv = 1.0f / sums[y];
            for (uint32_t x = 0; x < w; ++x) row[x] *= inv;
        } else {
            const float step = 1.0f / (float)(w - 1);
            for (uint32_t x = 0; x < w; ++x) row[x] = (float)x * step;
        }
        row[w - 1] = 1.0f;
    }
    float acc = 0.0f;
    for (uint32_t y = 0; y < h; ++y) {
        acc += marginal[y];
        marginal[y] = acc;
    }
    if (acc > 1e-5f) {
        const float inv = 1.0f / acc;
        for (float& m : marginal) m *= inv;
    } else {
        const float step = 1.0f / (float)(h - 1);
        for (uint32_t y = 0; y < h; ++y) marginal[y] = (float)y * step;
    }
    marginal[h - 1] = 1.0f;
}

// bake_texture2d (src/artic/entrypoints/bake.art) for packed RGBA8 images with nearest / bilinear lookup and identity
// transform: the texture at uv = (x / (w - 1), y / (h - 1)); the loader bakes environment radiance at >= 1024 x 512 before
// building its CDF (EnvironmentLight.cpp:37-41).
std::vector<float> bakeImageTexture(const uint32_t* px, uint32_t iw, uint32_t ih, bool opaque, uint32_t filter, uint32_t wrap_u, uint32_t wrap_v,
                                    uint32_t width, uint32_t height)
{
    auto texel = [&](long x, long y, int c) {
        const uint32_t p = px[(size_t)y * iw + (size_t)x];
        if (c == 3 && opaque) return 1.0f;
        return (float)((p >> (8 * c)) & 255u) / 255.0f;
    };
    auto border = [](uint32_t mode, long x, long n) {
        if (mode == IGB200_WRAP_CLAMP) return std::min(std::max(x, 0l), n - 1);
        if (mode == IGB200_WRAP_MIRROR) {
            const long t = x < 0 ? -1 - x : x;
            const long i = t / n, k = t % n;
            return i % 2 == 0 ? n - 1 - k : k;
        }
        long m = x % n;
        return m < 0 ? m + n : m;
    };
    if (filter > IGB200_FILTER_BILINEAR) unsupported("baking bicubic textures");
    std::vector<float> out((size_t)width * height * 4);
    for (uint32_t y = 0; y < height; ++y)
        for (uint32_t x = 0; x < width; ++x) {
            const float u = (float)x / (float)(width - 1), v = (float)y / (float)(height - 1);
            float*      o = out.data() + 4 * ((size_t)y * width + x);
            if (filter == IGB200_FILTER_NEAREST) {
                const long ix = border(wrap_u, (long)std::floor(u * (float)iw), (long)iw), iy = border(wrap_v, (long)std::floor(v * (float)ih), (long)ih);
                for (int c = 0; c < 4; ++c) o[c] = texel(ix, iy, c);
            } else {
                const float fu = u * (float)iw - 0.5f, fv = v * (float)ih - 0.5f;
                const long  x0 = (long)std::floor(fu), y0 = (long)std::floor(fv);
                const float fx = fu - std::floor(fu), fy = fv - std::floor(fv);
                const long  xa = border(wrap_u, x0, (long)iw), xb = border(wrap_u, x0 + 1, (long)iw);
                const long  ya = border(wrap_v, y0, (long)ih), yb = border(wrap_v, y0 + 1, (long)ih);
                for (int c = 0; c < 4; ++c) {
                    const float a = (1 - fx) * texel(xa, ya, c) + fx * texel(xb, ya, c);
                    const float b = (1 - fx) * texel(xa, yb, c) + fx * texel(xb, yb, c);
                    o[c]          = (1 - fy) * a + fy * b;
                }
            }
        }
    return out;
}

// ------------------------------------------------------------------------------------------------ sun position, skies
// computeSunEA (src/runtime/skysun/SunLocation.cpp:17-103) with the defaults of SunLocation.h:8-36
} // namespace
void computeSunEA(int year, int month, int day, int hour, int minute, float seconds, float longitude, float latitude, float timezone, float& el, float& az)
{
    const double dec_hours = hour + (double)timezone + (minute + (double)seconds / 60.0) / 60.0;
    const int    aux1      = (month - 14) / 12;
    const int    aux2      = (1461 * (year + 4800 + aux1)) / 4 + (367 * (month - 2 - 12 * aux1)) / 12 - (3 * ((year + 4900 + aux1) / 100)) / 4 + day - 32075;
    const double elapsed   = ((double)aux2 - 0.5 + dec_hours / 24.0) - 2451545.0;
    const double omega     = 2.1429 - 0.0010394594 * elapsed;
    const double mean_long = 4.8950630 + 0.017202791698 * elapsed;
    const double anomaly   = 6.2400600 + 0.0172019699 * elapsed;
    const double ecl_long  = mean_long + 0.03341607 * std::sin(anomaly) + 0.00034894 * std::sin(2 * anomaly) - 0.0001134 - 0.0000203 * std::sin(omega);
    const double ecl_obl   = 0.4090928 - 6.2140e-9 * elapsed + 0.0000396 * std::cos(omega);
    const double s_el      = std::sin(ecl_long);
    double       ra        = std::atan2(std::cos(ecl_obl) * s_el, std::cos(ecl_long));
    const double PI        = (double)kPiF;
    if (ra < 0) ra += 2 * PI;
    const double decl    = std::asin(std::sin(ecl_obl) * s_el);
    const double gmst    = 6.6974243242 + 0.0657098283 * elapsed + dec_hours;
    const double deg2rad = (double)(float)(PI / 180);
    const double lmst    = deg2rad * (double)(float)(gmst * 15 - (double)longitude);
    const double lat     = deg2rad * (double)latitude;
    const double ha      = lmst - ra;
    double       zenith  = std::acos(std::cos(lat) * std::cos(ha) * std::cos(decl) + std::sin(decl) * std::sin(lat));
    double       azim    = std::atan2(-std::sin(ha), std::tan(decl) * std::cos(lat) - std::sin(lat) * std::cos(ha));
    if (azim < 0) azim += 2 * PI;
    zenith += (6371.01 / 149597890) * std::sin(zenith);
    el = (float)(PI / 2) - (float)zenith;
    az = (float)std::fmod((double)((float)azim + (float)PI), 2 * PI);
}
namespace {

// LoaderUtils::getEA (loader/LoaderUtils.cpp:140-151): direction, sun_direction, elevation / azimuth, or time and place
void lightEA(const Value& l, float& el, float& az)
{
    if (l.has("direction") || l.has("sun_direction")) { // ElevationAzimuth::fromDirectionYUp
        Vec3d        d  = vec3(l.has("direction") ? l.find("direction") : l.find("sun_direction"), { 0, 0, 1 });
        const double n  = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        d               = Vec3d{ d[0] / n, d[1] / n, d[2] / n };
        const double ph = std::atan2(d[0], -d[2]);
        el              = (float)(kPiD / 2 - std::acos(d[1]));
        az              = (float)(ph < 0 ? ph + 2 * kPiD : ph);
        return;
    }
    if (l.has("elevation") || l.has("azimuth")) {
        el = (float)l.num("elevation", 0);
        az = (float)l.num("azimuth", 0);
        return;
    }
    computeSunEA((int)l.num("year", 2020), (int)l.num("month", 5), (int)l.num("day", 6), (int)l.num("hour", 12), (int)l.num("minute", 0),
                 (float)l.num("seconds", 0.0), (float)l.num("longitude", -6.9965744), (float)l.num("latitude", 49.235422), (float)l.num("timezone", -2.0), el, az);
}
Vec3f sunDirectionYUp(float el, float az) // ElevationAzimuth::toDirectionYUp
{
    return Vec3f{ (float)(std::cos((double)el) * std::sin((double)az)), (float)std::sin((double)el), (float)(-std::cos((double)el) * std::cos((double)az)) };
}

// ------------------------------------------------------------------------------------------------ light hierarchy
// LightHierarchy::setup (src/runtime/light/LightHierarchy.cpp:46-127) over PointBvh::store
// (src/runtime/container/PointBvh.inl:23-95).
struct HierEntry {
    Vec3f pos, dir;
    bool  has_dir;
    float flux;
};
void buildLightHierarchy(const std::vector<HierEntry>& entries, std::vector<uint32_t>& codes, std::vector<float>& nodes)
{
    struct Inner {
        int   index;
        Vec3f lo, hi;
        int   axis;
    };
    struct Leaf {
        Vec3f p, d;
        float f;
        int   id;
    };
    std::vector<Inner> inner;
    std::vector<Leaf>  leaves;
    for (size_t k = 0; k < entries.size(); ++k) {
        const HierEntry& e = entries[k];
        const Vec3f      p = e.pos;
        leaves.push_back(Leaf{ p, e.has_dir ? e.dir : Vec3f{ 0, 0, 1 }, e.has_dir ? e.flux : -e.flux, (int)k });
        if (inner.empty()) {
            inner.push_back(Inner{ 0, p, p, -1 });
            continue;
        }
        int i = 0;
        for (;;) { // descend, growing the boxes on the way
            Inner& n = inner[(size_t)i];
            for (int c = 0; c < 3; ++c) {
                n.lo[c] = std::min(n.lo[c], p[c]);
                n.hi[c] = std::max(n.hi[c], p[c]);
            }
            if (n.axis < 0) break;
            const float mid = ((n.hi[n.axis] + n.lo[n.axis]) / 2.0f);
            i               = p[n.axis] < mid ? n.index : n.index + 1;
        }
        const int   old_leaf = inner[(size_t)i].index;
        const Vec3f lo = inner[(size_t)i].lo, hi = inner[(size_t)i].hi;
        const Vec3f diam{ hi[0] - lo[0], hi[1] - lo[1], hi[2] - lo[2] };
        int         axis = 0;
        for (int c = 1; c < 3; ++c)
            if (diam[c] > diam[axis]) axis = c;
        const float mid      = diam[axis] / 2.0f; // (sic) half the extent, not a position
        const int   left_idx = (int)inner.size();
        inner[(size_t)i].index = left_idx;
        inner[(size_t)i].axis  = axis;
        const float off = (hi[axis] - lo[axis]) * 0.5f; // BoundingBox::computeSplit
        Vec3f       lhi = hi, rlo = lo;
        lhi[axis] -= off;
        rlo[axis] += off;
        const int new_leaf = (int)leaves.size() - 1;
        const int li = p[axis] < mid ? new_leaf : old_leaf, ri = p[axis] < mid ? old_leaf : new_leaf;
        inner.push_back(Inner{ li, lo, lhi, -1 });
        inner.push_back(Inner{ ri, rlo, hi, -1 });
    }
    nodes.assign(inner.size() * 8, 0.0f);
    std::vector<int32_t> ids(inner.size(), 0);
    codes.assign(entries.size(), 0u);
    struct Ret {
        Vec3f d;
        float f;
    };
    std::function<Ret(int, uint32_t, int)> populate = [&](int i, uint32_t code, int depth) -> Ret { // populateInnerNodes
        const Inner n   = inner[(size_t)i];
        float*      row = nodes.data() + 8 * (size_t)i;
        if (n.axis < 0) {
            const Leaf& l        = leaves[(size_t)n.index];
            codes[(size_t)l.id]  = code;
            row[0] = l.p[0], row[1] = l.p[1], row[2] = l.p[2], row[3] = l.f;
            row[4] = l.d[0], row[5] = l.d[1], row[6] = l.d[2];
            ids[(size_t)i] = l.id;
            return Ret{ l.d, l.f };
        }
        const Ret L = populate(n.index, code, depth + 1);
        const Ret R = populate(n.index + 1, code | (1u << depth), depth + 1);
        for (int c = 0; c < 3; ++c) row[c] = (n.hi[c] + n.lo[c]) / 2.0f;
        ids[(size_t)i] = -(n.index + 1);
        Vec3f d{ 0, 0, 1 };
        float f;
        if (L.f < 0 && R.f < 0) {
            f = L.f + R.f;
        } else if (L.f < 0) {
            f = -(-L.f + R.f);
        } else if (R.f < 0) {
            f = -(L.f - R.f);
        } else {
            const Vec3f s{ L.d[0] + R.d[0], L.d[1] + R.d[1], L.d[2] + R.d[2] };
            const float len = std::sqrt(s[0] * s[0] + s[1] * s[1] + s[2] * s[2]);
            d               = Vec3f{ s[0] / len, s[1] / len, s[2] / len };
            f               = L.f + R.f;
        }
        row[3] = f;
        row[4] = d[0], row[5] = d[1], row[6] = d[2];
        return Ret{ d, f };
    };
    if (!inner.empty()) populate(0, 0u, 0);
    for (size_t i = 0; i < inner.size(); ++i) std::memcpy(&nodes[8 * i + 7], &ids[i], 4);
}

// ------------------------------------------------------------------------------------------------ shapes
struct BuiltShape {
    bool  is_sphere = false;
    Mesh  mesh;
    Vec3f center{ 0, 0, 0 };
    float radius = 1.0f;
};

BuiltShape buildShape(const Value& e, const std::string& base)
{
    const std::string ty = e.str("type", "");
    BuiltShape        out;
    Mesh&             mesh = out.mesh;
    if (ty == "rectangle") {
        if (e.has("p0")) {
            mesh = makeRectangle(vec3(e.find("p0"), { -1, -1, 0 }), vec3(e.find("p1"), { 1, -1, 0 }), vec3(e.find("p2"), { 1, 1, 0 }), vec3(e.find("p3"), { -1, 1, 0 }));
        } else {
            const double w = e.num("width", 2.0), h = e.num("height", 2.0);
            mesh           = makePlane(vec3(e.find("origin"), { -w / 2, -h / 2, 0 }), { w, 0, 0 }, { 0, h, 0 });
        }
    } else if (ty == "triangle") {
        mesh = makeTriangle(vec3(e.find("p0"), { 0, 0, 0 }), vec3(e.find("p1"), { 1, 0, 0 }), vec3(e.find("p2"), { 0, 1, 0 }));
    } else if (ty == "cube" || ty == "box") {
        const double w = e.num("width", 2.0), h = e.num("height", 2.0), d = e.num("depth", 2.0);
        mesh           = makeBox(vec3(e.find("origin"), { -w / 2, -h / 2, -d / 2 }), { w, 0, 0 }, { 0, h, 0 }, { 0, 0, d });
    } else if (ty == "obj" || ty == "ply" || ty == "external") {
        const std::string path = joinPath(base, e.str("filename", ""));
        mesh                   = endsWith(lower(path), ".obj") ? loadObj(path) : loadPly(path);
    } else if (ty == "icosphere") { // setup_mesh_ico_sphere, TriMeshProvider.cpp:52-58
        mesh = makeIcoSphere(vec3(e.find("center"), { 0, 0, 0 }), e.num("radius", 1.0), (int)e.num("subdivisions", 4));
    } else if (ty == "uvsphere") { // :60-67
        mesh = makeUVSphere(vec3f(e.find("center"), { 0, 0, 0 }), (float)e.num("radius", 1.0), (int)e.num("stacks", 32), (int)e.num("slices", 16));
    } else if (ty == "cylinder") { // :69-86
        const bool   has_r = e.has("radius");
        const double r0    = has_r ? e.num("radius", 1.0) : e.num("bottom_radius", 1.0);
        const double r1    = has_r ? r0 : e.num("top_radius", r0);
        mesh = makeCylinder(vec3f(e.find("p0"), { 0, 0, 0 }), (float)r0, vec3f(e.find("p1"), { 0, 0, 1 }), (float)r1, (int)e.num("sections", 32), e.flag("filled", true));
    } else if (ty == "cone") { // :88-96
        mesh = makeCone(vec3f(e.find("p0"), { 0, 0, 0 }), (float)e.num("radius", 1.0), vec3f(e.find("p1"), { 0, 0, 1 }), (int)e.num("sections", 32), e.flag("filled", true));
    } else if (ty == "disk") { // :98-105
        mesh = makeDisk(vec3f(e.find("origin"), { 0, 0, 0 }), vec3f(e.find("normal"), { 0, 0, 1 }), (float)e.num("radius", 1.0), (int)e.num("sections", 32));
    } else if (ty == "sphere") {
        out.is_sphere = true;
        out.center    = vec3f(e.find("center"), { 0, 0, 0 });
        out.radius    = (float)e.num("radius", 1.0);
        return out;
    } else {
        unsupported("shape type '" + ty + "'");
    }
    if (e.flag("flip_normals", false)) mesh.flipNormals();
    if (e.flag("face_normals", false)) // TriMeshProvider.cpp:529-536
        mesh.faceNormalsAsVertexNormals();
    else if (e.flag("smooth_normals", false))
        mesh.computeVertexNormals();
    if (e.has("transform")) {
        const Mat4d m = parseTransform(e.find("transform"));
        if (!isIdentity(m)) mesh.transform(m);
    }
    const int subdiv = (int)e.num("subdivision", 0);
    if (subdiv > 0) { // handleModification, TriMeshProvider.cpp:435-473 (refinement / displacement: not built)
        for (int i = 0; i < subdiv; ++i) mesh.subdivide();
        if (e.flag("smooth_normals", false))
            mesh.computeVertexNormals();
        else
            mesh.faceNormalsAsVertexNormals();
    }
    return out;
}

float degToRadF(double deg) { return (float)(deg * (kPiD / 180.0)); }

// ------------------------------------------------------------------------------------------------ CIE / Perez skies
struct SkyFunction {
    uint32_t kind = IGB200_SKY_CIE;
    bool     variant = false, has_ground = true;
    Vec3f    scale{ 1, 1, 1 }, zenith{ 1, 1, 1 }, ground{ 1, 1, 1 }, sun{ 0, 1, 0 };
    float    zenith_brightness = 1.0f, ground_brightness = 0.2f, c2 = 1.0f, extra = 0.0f;
};

// CIELight::serialize (runtime/light/CIELight.cpp:41-110): the arguments of make_cie_sky_light / make_cie_sunny_light
SkyFunction cieSky(const Value& l, const std::string& ty)
{
    float el, az;
    lightEA(l, el, az);
    SkyFunction q;
    q.sun               = sunDirectionYUp(el, az);
    q.scale             = color(l.find("scale"), 1.0);
    q.zenith            = color(l.find("zenith"), 1.0);
    q.ground            = color(l.find("ground"), 1.0);
    q.ground_brightness = (float)l.num("ground_brightness", 0.2);
    q.has_ground        = l.flag("has_ground", true);
    if (ty == "cieuniform" || ty == "ciecloudy") {
        q.kind    = IGB200_SKY_CIE;
        q.variant = ty == "ciecloudy";
        return q;
    }
    const float PI = kPiF;
    el             = std::min(el, 87.0f * PI / 180.0f); // "Sun too close to zenith"
    const bool  clear     = ty == "cieclear";
    const float turbidity = (float)l.num("turbidity", 2.45);
    float       zb        = (1.376f * turbidity - 1.81f) * (float)std::tan((double)el) + 0.38f;
    if (!clear) zb = (zb + 8.6f * q.sun[1] + 0.123f) / 2.0f;
    zb                  = std::max(0.0f, zb * 1000.0f / 203.0f);
    const float half_pi = PI / 2.0f;
    float       factor;
    if (clear)
        factor = 0.274f * (0.91f + 10.0f * (float)std::exp(-3 * (double)(half_pi - el)) + 0.45f * q.sun[1] * q.sun[1]);
    else
        factor = (2.739f + 0.9891f * (float)std::sin(0.3119 + 2.6 * (double)el)) * (float)std::exp(-(double)(half_pi - el) * (0.4441 + 1.48 * (double)el));
    const double arr_clear[5] = { 2.766521, 0.547665, -0.369832, 0.009237, 0.059229 }, arr_inter[5] = { 3.5556, -2.7152, -1.3081, 1.0660, 0.60227 };
    const double* arr = clear ? arr_clear : arr_inter;
    const double  x   = ((double)el - (double)PI / 4) / ((double)PI / 4); // skylight_normalization_factor
    double        f   = arr[4];
    for (int i = 3; i >= 0; --i) f = f * x + arr[i];
    const float norm_factor = (float)f / PI / factor;
    const float solar       = 1.5e9f / 208.0f * (1.147f - 0.147f / std::max(q.sun[1], 0.16f));
    const float additive    = 6e-5f / PI * solar * q.sun[1] * (clear ? 1.0f : 0.15f);
    q.kind                  = IGB200_SKY_CIE_SUNNY;
    q.variant               = clear;
    q.zenith_brightness     = zb / factor;
    q.c2                    = zb * norm_factor + additive;
    return q;
}

// Perez all-weather sky: runtime/skysun/PerezModel.cpp (coefficients of Perez, Seals, Michalsky 1993)
const double kPerezRanges[9] = { 1.000, 1.065, 1.230, 1.500, 1.950, 2.800, 4.500, 6.200, std::numeric_limits<double>::infinity() };
const double kPerezA[32] = { 1.3525, -0.2576, -0.2690, -1.4366, -1.2219, -0.7730, 1.4148, 1.1016, -1.1000, -0.2515, 0.8952, 0.0156, -0.5484, -0.6654, -0.2672, 0.7117,
                             -0.6000, -0.3566, -2.5000, 2.3250, -1.0156, -0.3670, 1.0078, 1.4051, -1.0000, 0.0211, 0.5025, -0.5119, -1.0500, 0.0289, 0.4260, 0.3590 };
const double kPerezB[32] = { -0.7670, 0.0007, 1.2734, -0.1233, -0.2054, 0.0367, -3.9128, 0.9156, 0.2782, -0.1812, -4.5000, 1.1766, 0.7234, -0.6219, -5.6812, 2.6297,
                             0.2937, 0.0496, -5.6812, 1.8415, 0.2875, -0.5328, -3.8500, 3.3750, -0.3000, 0.1922, 0.7023, -1.6317, -0.3250, 0.1156, 0.7781, 0.0025 };
const double kPerezC[32] = { 2.8000, 0.6004, 1.2375, 1.0000, 6.9750, 0.1774, 6.4477, -0.1239, 24.7219, -13.0812, -37.7000, 34.8438, 33.3389, -18.3000, -62.2500, 52.0781,
                             21.0000, -4.7656, -21.5906, 7.2492, 14.0000, -0.9999, -7.1406, 7.5469, 19.0000, -5.0000, 1.2438, -1.9094, 31.0625, -14.5000, -46.1148, 55.3750 };
const double kPerezD[32] = { 1.8734, 0.6297, 0.9738, 0.2809, -1.5798, -0.5081, -1.7812, 0.1080, -5.0000, 1.5218, 3.9229, -2.6204, -3.5000, 0.0016, 1.1477, 0.1062,
                             -3.5000, -0.1554, 1.4062, 0.3988, -3.4000, -0.1078, -1.0750, 1.5702, -4.0000, 0.0250, 0.3844, 0.2656, -7.2312, 0.4050, 13.3500, 0.6234 };
const double kPerezE[32] = { 0.0356, -0.1246, -0.5718, 0.9938, 0.2624, 0.0672, -0.2190, -0.4285, -0.0156, 0.1597, 0.4199, -0.5562, 0.4659, -0.3296, -0.0876, -0.0329,
                             0.0032, 0.0766, -0.0656, -0.1294, -0.0672, 0.4016, 0.3017, -0.4844, 1.0468, -0.3788, -2.4517, 1.4656, 1.5000, -0.6426, 1.8564, 0.5636 };
constexpr double kSolarE = 1367.0, kSolarL = 127500.0; // skysun/Illuminance.h

double perezAirMass(double z) // computeAirMass, PerezModel.cpp:184-188
{
    const double zd = std::min(z * (180.0 / kPiD), 90.0);
    return 1 / (std::cos(zd * (kPiD / 180.0)) + 0.15 * std::exp(std::log(93.885 - zd) * -1.253));
}
double perezEccentricity(int day) // computeEccentricity :190-194
{
    const double a = 2 * kPiD * std::min(std::max(day / 365.0, 0.0), 1.0);
    return 1.00011 + 0.034221 * std::cos(a) + 0.00128 * std::sin(a) + 0.000719 * std::cos(2 * a) + 0.000077 * std::sin(2 * a);
}
void perezFromSky(double brightness, double clearness, double z, double out[5]) // PerezModel::fromSky :146-182
{
    if (clearness > 1.065 && clearness < 2.8 && brightness < 0.2) brightness = 0.2;
    brightness = std::min(std::max(brightness, 0.01), 0.6);
    int bin    = 8;
    for (int k = 0; k < 8; ++k)
        if (kPerezRanges[k] <= clearness && clearness < kPerezRanges[k + 1]) {
            bin = k;
            break;
        }
    if (bin >= 8) throw std::runtime_error("perez: sky clearness outside [1, inf)"); // the reference reads past its tables
    auto lin = [&](const double* x) { return x[0] + x[1] * z + brightness * (x[2] + x[3] * z); };
    out[0]   = lin(kPerezA + 4 * bin);
    out[1]   = lin(kPerezB + 4 * bin);
    out[2]   = bin > 0 ? lin(kPerezC + 4 * bin) : std::exp(std::pow(brightness * (kPerezC[0] + kPerezC[1] * z), kPerezC[2])) - kPerezC[3];
    out[3]   = bin > 0 ? lin(kPerezD + 4 * bin) : -std::exp(brightness * (kPerezD[0] + kPerezD[1] * z)) + kPerezD[2] + brightness * kPerezD[3];
    out[4]   = lin(kPerezE + 4 * bin);
}
double perezEval(const double p[5], double cos_sun, double cos_theta) // PerezModel::eval :66-72
{
    return (1 + p[0] * std::exp(p[1] / std::max(0.01, cos_theta))) * (1 + p[2] * std::exp(p[3] * std::acos(cos_sun)) + p[4] * cos_sun * cos_sun);
}
double perezIntegrate(const double p[5], double z) // PerezModel::integrate :74-118: Radiance's 145-patch base
{
    const int    rows[8][2] = { { 84, 30 }, { 72, 30 }, { 60, 24 }, { 48, 24 }, { 36, 18 }, { 24, 12 }, { 12, 6 }, { 0, 1 } };
    const double cs = std::cos(z), ss = std::sin(z);
    double       total = 0.0;
    for (const auto& r : rows) {
        const double ct = std::cos(r[0] * (kPiD / 180.0)), st = std::sin(r[0] * (kPiD / 180.0));
        for (int k = 0; k < r[1]; ++k) {
            const double cp = std::cos((360.0 / r[1] * k) * (kPiD / 180.0));
            total += perezEval(p, std::min(1.0, cs * ct + ss * st * cp), ct) * ct;
        }
    }
    return 2 * kPiD * total / 145;
}
int dayOfYear(int year, int month, int day) // tm_yday
{
    static const int cum[12] = { 0, 31, 59, 90, 120, 151, 181, 212, 243, 273, 304, 334 };
    const bool       leap    = (year % 4 == 0 && year % 100 != 0) || year % 400 == 0;
    return cum[std::min(std::max(month, 1), 12) - 1] + day - 1 + ((leap && month > 2) ? 1 : 0);
}
// PerezLight::serialize / getModel (runtime/light/PerezLight.cpp:25-110)
SkyFunction perezSky(const Value& l)
{
    float el, az;
    lightEA(l, el, az);
    SkyFunction q;
    q.kind         = IGB200_SKY_PEREZ;
    q.sun          = sunDirectionYUp(el, az);
    const double z = std::acos(std::min(1.0, std::max(-1.0, (double)q.sun[1])));
    const int    day = dayOfYear((int)l.num("year", 2020), (int)l.num("month", 5), (int)l.num("day", 6));
    double       model[5], diff;
    if (l.has("clearness") || l.has("brightness")) {
        const double br = l.num("brightness", 0.2);
        perezFromSky(br, l.num("clearness", 1.0), z, model);
        diff = br * kSolarE * perezEccentricity(day) / perezAirMass(z);
    } else if (l.has("direct_irradiance") || l.has("diffuse_irradiance") || l.has("direct_illuminance") || l.has("diffuse_illuminance")) {
        double dif, dirr;
        if (l.has("direct_irradiance") || l.has("diffuse_irradiance")) {
            dif  = l.num("diffuse_irradiance", 1.0);
            dirr = l.num("direct_irradiance", 1.0);
        } else {
            dif  = l.num("diffuse_illuminance", 1.0) * kSolarE / kSolarL;
            dirr = l.num("direct_illuminance", 1.0) * kSolarE / kSolarL;
        }
        const double a3 = 1.041 * std::pow(z, 3);
        perezFromSky(dif * perezAirMass(z) / (kSolarE * perezEccentricity(day)), ((dif + dirr) / dif + a3) / (1 + a3), z, model);
        diff = dif;
    } else {
        const char* names[5] = { "a", "b", "c", "d", "e" };
        for (int i = 0; i < 5; ++i) model[i] = l.num(names[i], 1.0);
        diff = 1.0;
    }
    const double diffnorm = diff / perezIntegrate(model, z);
    Vec3f        lum;
    if (l.has("luminance")) {
        lum = color(l.find("luminance"), 1.0);
        for (float& c : lum) c *= (float)diffnorm;
    } else {
        const double sin_el = std::min(1.0, std::max(-1.0, -(double)q.sun[1]));
        lum                 = color(l.find("zenith"), 1.0);
        for (float& c : lum) c *= (float)(perezEval(model, sin_el, 1.0) * diffnorm); // calc_perez(sin_elevation, 1, ...)
    }
    // packing of Shading.add_perez_sky: scale = (e, 0, 0), zenith = lum, brightness slots a, b, c, d in the last word
    q.scale             = Vec3f{ (float)model[4], 0, 0 };
    q.zenith            = lum;
    q.ground            = color(l.find("ground"), 1.0);
    q.zenith_brightness = (float)model[0];
    q.ground_brightness = (float)model[1];
    q.c2                = (float)model[2];
    q.extra             = (float)model[3];
    q.has_ground        = l.flag("has_ground", true);
    return q;
}

int addSkyFunction(ShadingBuilder& sb, const SkyFunction& q, const std::array<float, 9>& m)
{
    float params[16] = {};
    for (int k = 0; k < 3; ++k) {
        params[k]      = q.scale[k];
        params[3 + k]  = q.zenith[k];
        params[6 + k]  = q.ground[k];
        params[12 + k] = q.sun[k];
    }
    params[9]  = q.zenith_brightness;
    params[10] = q.ground_brightness;
    params[11] = q.c2;
    params[15] = q.extra;
    std::vector<float> v(9);
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) v[(size_t)(3 * i + j)] = m[(size_t)(3 * j + i)]; // m.T.reshape(-1)
    const int i = sb.light(sb.sh.infinite, IGB200_LIGHT_ENV_FUNCTION, { 1, 1, 1 }, v);
    igb200_light& L = sb.sh.infinite[(size_t)i];
    L.tex   = q.kind;
    L.cdf   = (uint32_t)(sb.blobAppend(params, sizeof(params)) / 4);
    L.flags = (q.variant ? IGB200_SKYFLAG_VARIANT : 0u) | (q.has_ground ? IGB200_SKYFLAG_HAS_GROUND : 0u);
    return i;
}

} // namespace

// ------------------------------------------------------------------------------------------------ the loader
static void loadDocumentImpl(const Value& scene, const std::string& base, SceneFile& out, size_t width_override, size_t height_override)
{
    auto list = [&](const char* key) -> const std::vector<Value>& {
        static const std::vector<Value> empty;
        const Value* v = scene.find(key);
        return v && v->isArray() ? v->array : empty;
    };

    size_t width = 1000, height = 1000;
    if (const Value* film = scene.find("film"))
        if (const Value* size = film->find("size")) {
            const auto n = numbers(*size);
            if (n.size() >= 2) {
                width  = (size_t)n[0];
                height = (size_t)n[1];
            }
        }
    if (width_override) width = width_override;
    if (height_override) height = height_override;

    TableBuilder tb;
    out.shading = ShadingTables{};
    ShadingBuilder sb(out.shading);

    std::map<std::string, int>        shapes;
    std::map<std::string, BuiltShape> built;
    for (const Value& e : list("shapes")) {
        const std::string name = e.str("name", "");
        BuiltShape        b    = buildShape(e, base);
        shapes[name]           = b.is_sphere ? tb.addSphere(b.center, b.radius) : tb.addTrimesh(b.mesh);
        built[name]            = std::move(b);
    }

    // ---- textures (pattern/CheckerBoardPattern.cpp:14-35, pattern/ImagePattern.cpp)
    std::map<std::string, int>          tex_ids;
    std::map<std::string, const Value*> tex_elems;
    for (const Value& e : list("textures")) {
        const std::string ty = e.str("type", ""), name = e.str("name", "");
        Mat4d             m;
        const Mat4d*      mp = nullptr;
        if (e.has("transform")) {
            m  = parseTransform(e.find("transform"));
            mp = &m;
        }
        if (ty == "checkerboard") {
            tex_ids[name] = sb.checkerboard((float)e.num("scale_x", 2.0), (float)e.num("scale_y", 2.0), color(e.find("color0"), 0.0), color(e.find("color1"), 1.0), mp);
        } else if (ty == "image" || ty == "bitmap") {
            const std::string file = joinPath(base, e.str("filename", ""));
            if (endsWith(lower(file), ".exr") || endsWith(lower(file), ".hdr")) unsupported("float image textures");
            uint32_t w, h;
            bool     opaque;
            const auto px = loadPackedImage(file, e.flag("linear", false), w, h, opaque);
            auto wrapOf = [](const std::string& s) { return s == "mirror" ? IGB200_WRAP_MIRROR : (s == "clamp" ? IGB200_WRAP_CLAMP : IGB200_WRAP_REPEAT); };
            uint32_t wu, wv;
            if (e.has("wrap_mode_u")) {
                wu = wrapOf(e.str("wrap_mode_u", "repeat"));
                wv = wrapOf(e.str("wrap_mode_v", "repeat"));
            } else {
                wu = wv = wrapOf(e.str("wrap_mode", "repeat"));
            }
            const std::string ft     = e.str("filter_type", "bicubic");
            const uint32_t    filter = ft == "bilinear" ? IGB200_FILTER_BILINEAR : (ft == "nearest" ? IGB200_FILTER_NEAREST : IGB200_FILTER_BICUBIC);
            tex_ids[name]            = sb.image(px.data(), w, h, IGB200_IMAGE_RGBA8, filter, wu, wv, mp, opaque);
        } else {
            unsupported("texture type '" + ty + "'");
        }
        tex_elems[name] = &e;
    }

    // ---- BSDFs (runtime/bsdf/*.cpp); bump maps refer to other BSDFs by name, in any order
    std::map<std::string, int> bsdf_ids;
    std::vector<const Value*>  pending;
    for (const Value& e : list("bsdfs")) pending.push_back(&e);
    size_t stall = 0;
    while (!pending.empty()) {
        const Value*      e  = pending.front();
        const std::string ty = e->str("type", ""), name = e->str("name", "");
        pending.erase(pending.begin());
        if (ty == "bumpmap" && !bsdf_ids.count(e->str("bsdf", ""))) {
            bool later = false;
            for (const Value* o : pending) later = later || o->str("name", "") == e->str("bsdf", "");
            if (!later || ++stall > 4 * (pending.size() + 1)) throw std::runtime_error("bumpmap '" + name + "' refers to an unknown bsdf");
            pending.push_back(e);
            continue;
        }
        if (ty == "bumpmap") { // runtime/bsdf/MapBSDF.cpp
            if (!tex_ids.count(e->str("map", ""))) throw std::runtime_error("bumpmap '" + name + "' refers to an unknown texture");
            const int i              = sb.bsdf(IGB200_BSDF_BUMPMAP, { 0, 0, 0 }, { 0, 0, 0 }, { (float)e->num("strength", 1.0) });
            out.shading.bsdfs[(size_t)i].q[0].u = (uint32_t)tex_ids[e->str("map", "")] + 1u;
            out.shading.bsdfs[(size_t)i].q[1].u = (uint32_t)bsdf_ids[e->str("bsdf", "")] + 1u;
            bsdf_ids[name]           = i;
        } else if (ty == "diffuse") {
            const Value* refl = e->find("reflectance");
            if (refl && refl->isString() && tex_ids.count(refl->string)) {
                const int i = sb.bsdf(IGB200_BSDF_DIFFUSE);
                out.shading.bsdfs[(size_t)i].q[0].u = (uint32_t)tex_ids[refl->string] + 1u;
                bsdf_ids[name] = i;
            } else {
                bsdf_ids[name] = sb.bsdf(IGB200_BSDF_DIFFUSE, color(refl, 0.8));
            }
        } else if (ty == "dielectric" || ty == "glass") { // DielectricBSDF.cpp:14-35: n1 = ext_ior, n2 = int_ior (bk7)
            bsdf_ids[name] = sb.bsdf(e->flag("thin", false) ? IGB200_BSDF_THIN_DIELECTRIC : IGB200_BSDF_DIELECTRIC, color(e->find("specular_reflectance"), 1.0),
                                     color(e->find("specular_transmittance"), 1.0), { (float)e->num("ext_ior", 1.0), (float)e->num("int_ior", 1.5046) });
        } else if (ty == "mirror") {
            bsdf_ids[name] = sb.bsdf(IGB200_BSDF_MIRROR, color(e->find("specular_reflectance"), 1.0));
        } else if (ty == "conductor") { // ConductorBSDF.cpp:13-36, BSDF.cpp:53-99
            if (lower(e->str("material", "none")) != "none") unsupported("named conductor materials");
            const Vec3f eta = color(e->find("eta"), 0.0), k = color(e->find("k"), 1.0), ks = color(e->find("specular_reflectance"), 1.0);
            const bool  old = e->has("alpha") || e->has("alpha_u") || e->has("alpha_v");
            const std::string par = old ? "alpha" : "roughness";
            auto conductor = [&](float au, float av) {
                const int i = sb.bsdf(IGB200_BSDF_CONDUCTOR, ks, eta, { k[0], k[1], k[2] });
                out.shading.bsdfs[(size_t)i].q[0].f = au;
                out.shading.bsdfs[(size_t)i].q[1].f = av;
                return i;
            };
            auto explicitAlpha = [](float roughness, double anisotropic, float& au, float& av) { // microfacet::compute_explicit
                const float aspect = anisotropic == 0 ? 1.0f : std::sqrt(1.0f - (float)std::min(std::max(anisotropic, 0.0), 1.0) * 0.99f);
                au                 = roughness / aspect;
                av                 = roughness * aspect;
            };
            if (!e->has(par) && !e->has(par + "_u") && !e->has(par + "_v")) {
                bool mirror = true; // delta distribution: mirror when eta ~ 0 and k ~ 1 (bsdf/conductor.art:112-122)
                for (int c = 0; c < 3; ++c) mirror = mirror && std::fabs(eta[c]) <= 1e-4f && std::fabs(k[c] - 1.0f) <= 1e-4f;
                if (mirror) {
                    bsdf_ids[name] = sb.bsdf(IGB200_BSDF_MIRROR, ks);
                } else {
                    float au, av;
                    explicitAlpha(0.0f, 0.0, au, av);
                    bsdf_ids[name] = conductor(au, av);
                }
            } else if (e->has(par + "_u") || e->has(par + "_v")) {
                bsdf_ids[name] = conductor((float)e->num(par + "_u", 0.1), (float)e->num(par + "_v", 0.1));
            } else {
                float au, av;
                explicitAlpha((float)e->num(par, 0.1), e->num("anisotropic", 0.0), au, av);
                bsdf_ids[name] = conductor(au, av);
            }
        } else {
            unsupported("bsdf type '" + ty + "'");
        }
    }

    // ---- entities: material groups in first-appearance order; area-light entities get a material of their own
    // (LoaderEntity.cpp:42-103)
    const auto&           lights = list("lights");
    std::set<std::string> area_entities;
    for (const Value& l : lights)
        if (l.str("type", "") == "area") area_entities.insert(l.str("entity", ""));
    struct Group {
        bool                      area;
        std::string               key;
        std::vector<const Value*> members;
    };
    std::vector<Group> groups;
    for (const Value& e : list("entities")) {
        const std::string name = e.str("name", "");
        const bool        area = area_entities.count(name) > 0;
        const std::string key  = area ? name : e.str("bsdf", "");
        bool              done = false;
        if (!area)
            for (Group& g : groups)
                if (!g.area && g.key == key) {
                    g.members.push_back(&e);
                    done = true;
                    break;
                }
        if (!done) groups.push_back(Group{ area, key, { &e } });
    }
    std::map<std::string, int>          entity_ids;
    std::map<std::string, Mat4d>        entity_xform;
    std::map<std::string, const Value*> entity_elems;
    for (size_t mat_id = 0; mat_id < groups.size(); ++mat_id)
        for (const Value* e : groups[mat_id].members) {
            const std::string name = e->str("name", "");
            if (!shapes.count(e->str("shape", ""))) throw std::runtime_error("entity '" + name + "' refers to an unknown shape");
            const Mat4d    m     = parseTransform(e->find("transform"));
            const uint32_t flags = (e->flag("camera_visible", true) ? 1u : 0u) | (e->flag("light_visible", true) ? 2u : 0u) | (e->flag("bounce_visible", true) ? 4u : 0u)
                                   | (e->flag("shadow_visible", true) ? 8u : 0u);
            entity_ids[name]   = tb.addEntity(shapes[e->str("shape", "")], m, (int)mat_id, flags);
            entity_xform[name] = m;
            entity_elems[name] = e;
        }
    tb.finish(out);

    // ---- lights (LoaderLight.cpp registry, runtime/light/*.cpp): infinite and finite tables in scene order
    std::map<std::string, int> area_light_of_entity;
    std::vector<HierEntry>     hier; // Light::position / direction / computeFlux per finite light
    const float                SR = (float)(4 * kPiD);
    auto mean3 = [](const Vec3f& c) { return ((double)c[0] + (double)c[1] + (double)c[2]) / 3.0; };
    auto xformPoint = [](const Mat4d& m, const Vec3d& p) {
        return Vec3d{ m[0][0] * p[0] + m[0][1] * p[1] + m[0][2] * p[2] + m[0][3], m[1][0] * p[0] + m[1][1] * p[1] + m[1][2] * p[2] + m[1][3],
                      m[2][0] * p[0] + m[2][1] * p[1] + m[2][2] * p[2] + m[2][3] };
    };
    auto xformVec = [](const Mat4d& m, const Vec3d& p) {
        return Vec3d{ m[0][0] * p[0] + m[0][1] * p[1] + m[0][2] * p[2], m[1][0] * p[0] + m[1][1] * p[1] + m[1][2] * p[2], m[2][0] * p[0] + m[2][1] * p[1] + m[2][2] * p[2] };
    };
    auto normalizedDirF = [](const Vec3d& d) {
        const double n = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        return Vec3f{ (float)(d[0] / n), (float)(d[1] / n), (float)(d[2] / n) };
    };
    for (const Value& l : lights) {
        std::string ty = l.str("type", "");
        if (ty.rfind("cie_", 0) == 0) ty = "cie" + ty.substr(4); // both spellings are registered (LoaderLight.cpp:90-97)
        if (ty == "point") { // PointLight.cpp:16-49
            Vec3f c;
            if (l.has("power")) {
                c = color(l.find("power"), 4 * kPiD);
                for (float& x : c) x = x / SR;
            } else {
                c = color(l.find("intensity"), 1.0);
            }
            const Vec3f pos = vec3f(l.find("position"), { 0, 0, 0 });
            sb.light(out.shading.finite, IGB200_LIGHT_POINT, c, { pos[0], pos[1], pos[2] });
            const Vec3f cs{ c[0] * SR, c[1] * SR, c[2] * SR };
            hier.push_back(HierEntry{ pos, { 0, 0, 1 }, false, (float)mean3(cs) });
        } else if (ty == "spot") {
            const double cutoff = l.num("cutoff", 30.0), falloff = l.num("falloff", 20.0);
            Vec3f        c = color(l.find("intensity"), 1.0);
            const double f = 2 * kPiD * (1 - 0.5 * std::cos(falloff * (kPiD / 180.0)) - 0.5 * std::cos(cutoff * (kPiD / 180.0)));
            if (l.has("power")) { // spot_from_power, light/spot.art:1-6
                c = color(l.find("power"), 1.0);
                for (float& x : c) x = x / (float)f;
            }
            const Vec3f pos = vec3f(l.find("position"), { 0, 0, 0 });
            const Vec3f d   = normalizedDirF(vec3(l.find("direction"), { 0, 0, 1 }));
            sb.light(out.shading.finite, IGB200_LIGHT_SPOT, c, { pos[0], pos[1], pos[2], d[0], d[1], d[2], degToRadF(cutoff), degToRadF(falloff) });
            const Vec3d dd = vec3(l.find("direction"), { 0, 0, 1 });
            const double dn = std::sqrt(dd[0] * dd[0] + dd[1] * dd[1] + dd[2] * dd[2]);
            hier.push_back(HierEntry{ pos, Vec3f{ (float)(dd[0] / dn), (float)(dd[1] / dn), (float)(dd[2] / dn) }, true, (float)(mean3(c) * f) });
        } else if (ty == "sun") { // SunLight.cpp:8-46: direction through the EA round trip, angle 11.4 degrees unless `radius`
            float el, az;
            lightEA(l, el, az);
            const Vec3d d{ std::cos((double)el) * std::sin((double)az), std::sin((double)el), -std::cos((double)el) * std::cos((double)az) };
            const Vec3f df = normalizedDirF(d);
            float       cos_angle;
            if (l.has("radius")) {
                const double r = l.num("radius", 0.0);
                cos_angle      = (float)(1 / std::sqrt(r * r + 1));
            } else {
                cos_angle = (float)std::cos((double)((float)l.num("angle", 11.4) / 2.0f) * (kPiD / 180.0));
            }
            sb.light(out.shading.infinite, IGB200_LIGHT_SUN, color(l.find("irradiance"), 1.0), { df[0], df[1], df[2], cos_angle });
        } else if (ty == "directional" || ty == "direction" || ty == "distant") {
            const Vec3f d = normalizedDirF(vec3(l.find("direction"), { 0, 0, 1 }));
            sb.light(out.shading.infinite, IGB200_LIGHT_DIRECTIONAL, color(l.find("irradiance"), 1.0), { d[0], d[1], d[2] });
        } else if (ty == "env" || ty == "envmap" || ty == "uniform" || ty == "constant") {
            const Value* rad = l.find("radiance");
            if (rad && rad->isString() && tex_ids.count(rad->string)) { // EnvironmentLight.cpp:27-76
                const int             tid = tex_ids[rad->string];
                const igb200_texture  t   = out.shading.textures[(size_t)tid];
                uint32_t              cdf = IGB200_NO_CDF;
                if (l.flag("cdf", true) && t.type == IGB200_TEX_IMAGE && t.format == IGB200_IMAGE_RGBA8 && !tex_elems[rad->string]->has("transform")) {
                    std::vector<uint32_t> px((size_t)t.width * t.height);
                    std::memcpy(px.data(), out.shading.blob.data() + t.offset, px.size() * 4);
                    const uint32_t bw = std::max(1024u, t.width), bh = std::max(512u, t.height);
                    const auto     baked = bakeImageTexture(px.data(), t.width, t.height, t.opaque != 0, t.filter, t.wrap_u, t.wrap_v, bw, bh);
                    std::vector<float> marginal, cond;
                    computeCdf2d(baked.data(), bw, bh, true, l.flag("compensate", true), marginal, cond);
                    cdf = sb.cdf2d(marginal, cond, bw, bh);
                }
                const auto         m = envTransform(l.find("transform"));
                std::vector<float> v(9);
                for (int i = 0; i < 3; ++i)
                    for (int j = 0; j < 3; ++j) v[(size_t)(3 * i + j)] = m[(size_t)(3 * j + i)];
                const int i = sb.light(out.shading.infinite, IGB200_LIGHT_ENV_TEXTURED, color(l.find("scale"), 1.0), v);
                out.shading.infinite[(size_t)i].tex = (uint32_t)tid;
                out.shading.infinite[(size_t)i].cdf = cdf;
            } else {
                const Vec3f r = color(rad, 1.0), s = color(l.find("scale"), 1.0);
                sb.light(out.shading.infinite, IGB200_LIGHT_ENV, { r[0] * s[0], r[1] * s[1], r[2] * s[2] }, { 1, 0, 0, 0, 1, 0, 0, 0, 1 });
            }
        } else if (ty == "sky") { // runtime/light/SkyLight.cpp:33-77: Hosek-Wilkie bake -> image texture + 2-D CDF
            float el, az;
            lightEA(l, el, az);
            const Vec3f ground = color(l.find("ground"), 0.8);
            SkyModel    model(ground.data(), el, az, (float)l.num("turbidity", 3.0));
            std::vector<float> img(4 * RES_AZ * RES_EL);
            model.toDeviceImage(img.data());
            const int tex = sb.image(img.data(), (uint32_t)RES_AZ, (uint32_t)RES_EL, IGB200_IMAGE_RGBA32F, IGB200_FILTER_BILINEAR, IGB200_WRAP_REPEAT, IGB200_WRAP_REPEAT,
                                     nullptr, false);
            std::vector<float> marginal, cond;
            computeCdf2d(img.data(), (uint32_t)RES_AZ, (uint32_t)RES_EL, true, false, marginal, cond);
            const uint32_t     cdf = sb.cdf2d(marginal, cond, (uint32_t)RES_AZ, (uint32_t)RES_EL);
            const auto         m   = envTransform(l.find("transform"));
            std::vector<float> v(9);
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) v[(size_t)(3 * i + j)] = m[(size_t)(3 * j + i)];
            const int i = sb.light(out.shading.infinite, IGB200_LIGHT_ENV_TEXTURED, color(l.find("scale"), 1.0), v);
            out.shading.infinite[(size_t)i].tex = (uint32_t)tex;
            out.shading.infinite[(size_t)i].cdf = cdf;
        } else if (ty == "cieuniform" || ty == "ciecloudy" || ty == "cieclear" || ty == "cieintermediate") {
            addSkyFunction(sb, cieSky(l, ty), envTransform(l.find("transform")));
        } else if (ty == "perez") {
            addSkyFunction(sb, perezSky(l), envTransform(l.find("transform")));
        } else if (ty == "area") { // AreaLight.cpp:26-92,143-160
            const std::string name = l.str("entity", "");
            if (!entity_elems.count(name)) throw std::runtime_error("area light refers to an unknown entity '" + name + "'");
            const Value*      ent   = entity_elems[name];
            const BuiltShape& shape = built[ent->str("shape", "")];
            const Mat4d&      m     = entity_xform[name];
            const Vec3f       rad   = color(l.find("radiance"), 1.0);
            const bool        optimize = l.flag("optimize", true);
            if (shape.is_sphere && optimize) { // analytic sphere sampler (:54-55,73-82)
                area_light_of_entity[name] = sb.light(out.shading.finite, IGB200_LIGHT_AREA_SPHERE, rad, { shape.center[0], shape.center[1], shape.center[2], shape.radius },
                                                      entity_ids[name]);
                double ext[3];
                for (int k = 0; k < 3; ++k) {
                    Vec3d axis{ 0, 0, 0 };
                    axis[k]       = shape.radius;
                    const Vec3d a = xformVec(m, axis);
                    ext[k]        = std::sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
                }
                const double P = 1.6075; // approximate_ellipsoid_area, AreaLight.cpp:26-35
                const double area = 4 * kPiD * std::pow((std::pow(ext[0] * ext[1], P) + std::pow(ext[0] * ext[2], P) + std::pow(ext[1] * ext[2], P)) / 3, 1 / P);
                const Vec3d  c    = xformPoint(m, Vec3d{ shape.center[0], shape.center[1], shape.center[2] });
                hier.push_back(HierEntry{ Vec3f{ (float)c[0], (float)c[1], (float)c[2] }, { 0, 0, 1 }, false, (float)(mean3(rad) * area * kPiD) });
                continue;
            }
            if (shape.is_sphere) unsupported("sphere area lights without the analytic sampler");
            Vec3f po, px, py;
            if (optimize && meshAsPlane(shape.mesh, po, px, py)) { // plane emitter in WORLD space (:56-70)
                const Vec3d o = xformPoint(m, Vec3d{ po[0], po[1], po[2] }), x = xformVec(m, Vec3d{ px[0], px[1], px[2] }), y = xformVec(m, Vec3d{ py[0], py[1], py[2] });
                const Vec3f of{ (float)o[0], (float)o[1], (float)o[2] }, xf{ (float)x[0], (float)x[1], (float)x[2] }, yf{ (float)y[0], (float)y[1], (float)y[2] };
                Vec3f       n{ xf[1] * yf[2] - xf[2] * yf[1], xf[2] * yf[0] - xf[0] * yf[2], xf[0] * yf[1] - xf[1] * yf[0] };
                const float area = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
                n                = Vec3f{ n[0] / area, n[1] / area, n[2] / area };
                area_light_of_entity[name] = sb.light(out.shading.finite, IGB200_LIGHT_AREA_PLANE, rad,
                                                      { of[0], of[1], of[2], xf[0], xf[1], xf[2], yf[0], yf[1], yf[2], n[0], n[1], n[2], area }, entity_ids[name]);
                const Vec3d  nd{ x[1] * y[2] - x[2] * y[1], x[2] * y[0] - x[0] * y[2], x[0] * y[1] - x[1] * y[0] };
                const double ad = std::sqrt(nd[0] * nd[0] + nd[1] * nd[1] + nd[2] * nd[2]);
                const Vec3d  c{ o[0] + x[0] * 0.5 + y[0] * 0.5, o[1] + x[1] * 0.5 + y[1] * 0.5, o[2] + x[2] * 0.5 + y[2] * 0.5 };
                hier.push_back(HierEntry{ Vec3f{ (float)c[0], (float)c[1], (float)c[2] }, Vec3f{ (float)(nd[0] / ad), (float)(nd[1] / ad), (float)(nd[2] / ad) }, true,
                                          (float)(mean3(rad) * ad * kPiD) });
            } else { // shape emitter (:84-92; world-space area for the selection weight)
                area_light_of_entity[name] = sb.light(out.shading.finite, IGB200_LIGHT_AREA_SHAPE, rad, {}, entity_ids[name]);
                double area = 0.0;
                Vec3f  lo = shape.mesh.v[0], hi = shape.mesh.v[0];
                for (const Vec3f& p : shape.mesh.v)
                    for (int k = 0; k < 3; ++k) {
                        lo[k] = std::min(lo[k], p[k]);
                        hi[k] = std::max(hi[k], p[k]);
                    }
                for (const auto& f : shape.mesh.f) {
                    Vec3d w[3];
                    for (int k = 0; k < 3; ++k) w[k] = xformPoint(m, Vec3d{ shape.mesh.v[f[(size_t)k]][0], shape.mesh.v[f[(size_t)k]][1], shape.mesh.v[f[(size_t)k]][2] });
                    const Vec3d a{ w[1][0] - w[0][0], w[1][1] - w[0][1], w[1][2] - w[0][2] }, b{ w[2][0] - w[0][0], w[2][1] - w[0][1], w[2][2] - w[0][2] };
                    const Vec3d c{ a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0] };
                    area += std::sqrt(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]);
                }
                area /= 2;
                const Vec3d c0{ ((double)(float)(lo[0] + hi[0])) / 2, ((double)(float)(lo[1] + hi[1])) / 2, ((double)(float)(lo[2] + hi[2])) / 2 };
                const Vec3d c = xformPoint(m, c0);
                hier.push_back(HierEntry{ Vec3f{ (float)c[0], (float)c[1], (float)c[2] }, { 0, 0, 1 }, false, (float)(mean3(rad) * area * kPiD) });
            }
        } else {
            unsupported("light type '" + ty + "'");
        }
    }

    for (const Group& g : groups) {
        igb200_material mat{};
        if (g.area) {
            mat.bsdf  = bsdf_ids.at(entity_elems[g.key]->str("bsdf", ""));
            mat.light = area_light_of_entity.at(g.key);
        } else {
            if (!bsdf_ids.count(g.key)) throw std::runtime_error("entity refers to an unknown bsdf '" + g.key + "'");
            mat.bsdf  = bsdf_ids[g.key];
            mat.light = -1;
        }
        out.shading.materials.push_back(mat);
    }

    // ---- camera (PerspectiveCamera.cpp:27-78, FishLensCamera.cpp:59-66)
    {
        static const Value empty_obj = [] { Value v; v.type = Value::Object; return v; }();
        const Value*   camp = scene.find("camera");
        const Value&   cam  = camp ? *camp : empty_obj;
        const Mat4d    m    = parseTransform(cam.find("transform"));
        igb200_camera& c    = out.shading.camera;
        for (int k = 0; k < 3; ++k) {
            c.eye[k] = (float)m[k][3];
            c.dir[k] = (float)m[k][2];
            c.up[k]  = (float)m[k][1];
        }
        double near = cam.num("near_clip", 0.0), far = cam.num("far_clip", (double)std::numeric_limits<float>::max());
        if (far < near) std::swap(near, far);
        c.tmin               = (float)near;
        c.tmax               = (float)far;
        const std::string ct = cam.str("type", "perspective");
        const float aspect   = (float)width / (float)height;
        if (ct == "fishlens" || ct == "fisheye") { // mode is always Circular
            c.scale[0] = aspect < 1 ? 1.0f : aspect;
            c.scale[1] = aspect > 1 ? 1.0f : aspect;
            c.type     = IGB200_CAMERA_FISHLENS;
        } else {
            const bool   vertical = cam.has("vfov");
            const double fov      = cam.has("vfov") ? cam.num("vfov", 60.0) : (cam.has("hfov") ? cam.num("hfov", 60.0) : cam.num("fov", 60.0));
            const float  t        = (float)std::tan((double)((float)fov * (float)(kPiD / 180.0)) / 2);
            if (vertical) {
                c.scale[1] = t;
                c.scale[0] = t * aspect;
            } else {
                c.scale[0] = t;
                c.scale[1] = t / aspect;
            }
            c.type = IGB200_CAMERA_PERSPECTIVE;
        }
    }

    // ---- technique (PathTechnique.cpp:8-27, DebugTechnique, InfoBufferTechnique.cpp:6-12) and light selector
    // (LoaderLight::generateLightSelector, loader/LoaderLight.cpp:423-478)
    {
        static const Value def_tech = json::parse("{\"type\": \"path\"}");
        const Value*      tp   = scene.find("technique");
        const Value&      tech = tp ? *tp : def_tech;
        const std::string ty   = tech.str("type", "path");
        igb200_technique& t    = out.shading.technique;
        t                      = igb200_technique{};
        if (ty == "path") {
            t.type       = IGB200_TECH_PATH;
            t.max_depth  = (int32_t)tech.num("max_depth", 64);
            t.min_depth  = (int32_t)tech.num("min_depth", 2);
            t.clamp      = (float)tech.num("clamp", 0.0);
            t.enable_nee = tech.flag("nee", true) ? 1 : 0;
            t.flags      = tech.flag("aov_mis", false) ? IGB200_TECHFLAG_AOV_MIS : 0u;
            const std::string sel  = tech.str("light_selector", "uniform");
            const bool        many = out.shading.infinite.size() + out.shading.finite.size() > 1 && !out.shading.finite.empty();
            if (sel == "hierarchy" && many) {
                std::vector<uint32_t> codes;
                std::vector<float>    nodes;
                buildLightHierarchy(hier, codes, nodes);
                while (codes.size() % 4) codes.push_back(0u);
                std::vector<uint8_t> raw(4 * codes.size() + 4 * nodes.size());
                std::memcpy(raw.data(), codes.data(), 4 * codes.size());
                std::memcpy(raw.data() + 4 * codes.size(), nodes.data(), 4 * nodes.size());
                out.shading.hierarchy = (uint32_t)(sb.blobAppend(raw.data(), raw.size()) / 4);
                out.shading.selector  = IGB200_SELECTOR_HIERARCHY;
            } else if (sel == "simple" && many) { // CDF::computeForArray (runtime/CDF.cpp:11-40) over the estimated flux
                std::vector<float> cdf(hier.size());
                float              acc = 0.0f;
                for (size_t i = 0; i < hier.size(); ++i) {
                    acc += hier[i].flux;
                    cdf[i] = acc;
                }
                if (acc > 1e-5f) {
                    const float inv = 1.0f / acc;
                    for (float& x : cdf) x *= inv;
                } else {
                    for (size_t i = 0; i < cdf.size(); ++i) cdf[i] = (float)i * (1.0f / (float)cdf.size());
                }
                cdf.back()            = 1.0f;
                out.shading.hierarchy = (uint32_t)(sb.blobAppend(cdf.data(), 4 * cdf.size()) / 4);
                out.shading.selector  = IGB200_SELECTOR_CDF;
            } else if (sel != "uniform" && sel != "hierarchy" && sel != "simple" && !sel.empty()) {
                unsupported("light selector '" + sel + "'");
            }
        } else if (ty == "debug") {
            t.type       = IGB200_TECH_DEBUG;
            t.max_depth  = 1;
            t.min_depth  = 1;
            t.debug_mode = (int32_t)tech.num("mode", 0);
        } else if (ty == "infobuffer") {
            t.type      = IGB200_TECH_INFOBUFFER;
            t.max_depth = (int32_t)tech.num("max_depth", 64);
            t.min_depth = 1;
            t.flags     = tech.flag("handle_specular", false) ? IGB200_TECHFLAG_IB_HANDLE_SPECULAR : 0u;
        } else {
            unsupported("technique '" + ty + "'");
        }
    }
    out.width  = width;
    out.height = height;
}

bool loadSceneDocument(const json::Value& scene, const std::string& base_dir, SceneFile& out, std::string& error, size_t width, size_t height)
{
    try {
        loadDocumentImpl(scene, base_dir, out, width, height);
        return true;
    } catch (const std::exception& e) {
        error = e.what();
        return false;
    }
}

bool loadSceneJson(const std::string& path, SceneFile& out, std::string& error, size_t width, size_t height)
{
    try {
        const std::string base = dirName(path);
        loadDocumentImpl(mergeExternals(json::parse(readText(path)), base), base, out, width, height);
        return true;
    } catch (const std::exception& e) {
        error = e.what();
        return false;
    }
}

bool loadAnyScene(const std::string& path, SceneFile& out, std::string& error, size_t width, size_t height)
{
    if (endsWith(lower(path), ".json")) return loadSceneJson(path, out, error, width, height);
    return loadSceneFile(path, out, error);
}

bool saveSceneFile(const std::string& path, const SceneFile& scene, std::string& error)
{
    struct Section {
        std::string          name;
        std::vector<uint8_t> raw;
    };
    std::vector<Section> secs;
    auto add = [&](const char* name, const void* p, size_t n) {
        Section s;
        s.name = name;
        s.raw.assign(static_cast<const uint8_t*>(p), static_cast<const uint8_t*>(p) + n);
        secs.push_back(std::move(s));
    };
    const SceneDatabase& db  = scene.database;
    const DynTable&      shp = db.DynTables.at("shapes");
    add("lookups", shp.lookups().data(), shp.lookups().size() * sizeof(LookupEntry));
    add("shape_data", shp.data().data(), shp.data().size());
    add("entities", db.FixTables.at("entities").data().data(), db.FixTables.at("entities").data().size());
    add("leaves0", db.SceneBVHs.at("trimesh").Leaves.data(), db.SceneBVHs.at("trimesh").Leaves.size());
    add("leaves1", db.SceneBVHs.at("sphere").Leaves.data(), db.SceneBVHs.at("sphere").Leaves.size());
    const float bbox[6] = { db.SceneBBox.min(0), db.SceneBBox.min(1), db.SceneBBox.min(2), db.SceneBBox.max(0), db.SceneBBox.max(1), db.SceneBBox.max(2) };
    add("bbox", bbox, sizeof(bbox));
    const uint64_t mc = db.MaterialCount;
    add("material_count", &mc, 8);
    add("entity_per_material", scene.entity_per_material.data(), scene.entity_per_material.size() * 4);
    const ShadingTables& sh = scene.shading;
    add("camera", &sh.camera, sizeof(sh.camera));
    add("technique", &sh.technique, sizeof(sh.technique));
    add("bsdfs", sh.bsdfs.data(), sh.bsdfs.size() * sizeof(igb200_bsdf));
    add("materials", sh.materials.data(), sh.materials.size() * sizeof(igb200_material));
    add("infinite", sh.infinite.data(), sh.infinite.size() * sizeof(igb200_light));
    add("finite", sh.finite.data(), sh.finite.size() * sizeof(igb200_light));
    add("textures", sh.textures.data(), sh.textures.size() * sizeof(igb200_texture));
    add("blob", sh.blob.data(), sh.blob.size());
    const uint32_t sel[2] = { sh.selector, sh.hierarchy };
    add("selector", sel, 8);
    const uint32_t film[2] = { (uint32_t)scene.width, (uint32_t)scene.height };
    add("film", film, 8);

    auto   align  = [](size_t x) { return (x + 15) / 16 * 16; };
    size_t offset = align(16 + 40 * secs.size());
    std::vector<uint8_t> head(offset, 0);
    std::memcpy(head.data(), "IGB200DB", 8);
    const uint32_t hv[2] = { 1u, (uint32_t)secs.size() };
    std::memcpy(head.data() + 8, hv, 8);
    size_t at = offset;
    for (size_t i = 0; i < secs.size(); ++i) {
        uint8_t* e = head.data() + 16 + 40 * i;
        std::memcpy(e, secs[i].name.c_str(), std::min<size_t>(23, secs[i].name.size()));
        const uint64_t v[2] = { secs[i].raw.size(), at };
        std::memcpy(e + 24, v, 16);
        at += align(secs[i].raw.size());
    }
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) {
        error = "cannot write " + path;
        return false;
    }
    bool ok = std::fwrite(head.data(), 1, head.size(), f) == head.size();
    for (const Section& s : secs) {
        std::vector<uint8_t> padded(align(s.raw.size()), 0);
        if (!s.raw.empty()) std::memcpy(padded.data(), s.raw.data(), s.raw.size());
        ok = ok && (padded.empty() || std::fwrite(padded.data(), 1, padded.size(), f) == padded.size());
    }
    std::fclose(f);
    if (!ok) error = "short write to " + path;
    return ok;
}

} // namespace IG

extern "C" {
/// computeSunEA of the loader (tests pin it against the Python restatement and the reference's sky goldens).
void ig_b200_sun_ea(int year, int month, int day, int hour, int minute, float seconds, float longitude, float latitude, float timezone, float* elevation,
                    float* azimuth)
{
    IG::computeSunEA(year, month, day, hour, minute, seconds, longitude, latitude, timezone, *elevation, *azimuth);
}
}
