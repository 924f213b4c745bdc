#include "SceneMesh.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <tuple>

namespace IG {
namespace loader {

// ------------------------------------------------------------------------------------------------ small linear algebra
Mat4d identity4()
{
    Mat4d m{};
    for (int i = 0; i < 4; ++i) m[i][i] = 1.0;
    return m;
}

Mat4d mul(const Mat4d& a, const Mat4d& b)
{
    Mat4d r{};
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            double s = 0.0;
            for (int k = 0; k < 4; ++k) s += a[i][k] * b[k][j];
            r[i][j] = s;
        }
    return r;
}

Mat4d inverse(const Mat4d& m)
{
    double a[4][8];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            a[i][j]     = m[i][j];
            a[i][j + 4] = i == j ? 1.0 : 0.0;
        }
    for (int c = 0; c < 4; ++c) {
        int p = c;
        for (int r = c + 1; r < 4; ++r)
            if (std::fabs(a[r][c]) > std::fabs(a[p][c])) p = r;
        if (a[p][c] == 0.0) throw std::runtime_error("singular transform");
        if (p != c)
            for (int j = 0; j < 8; ++j) std::swap(a[p][j], a[c][j]);
        const double inv = 1.0 / a[c][c];
        for (int j = 0; j < 8; ++j) a[c][j] *= inv;
        for (int r = 0; r < 4; ++r) {
            if (r == c) continue;
            const double f = a[r][c];
            if (f == 0.0) continue;
            for (int j = 0; j < 8; ++j) a[r][j] -= f * a[c][j];
        }
    }
    Mat4d out{};
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) out[i][j] = a[i][j + 4];
    return out;
}

bool isIdentity(const Mat4d& m, double tol)
{
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j)
            if (std::fabs(m[i][j] - (i == j ? 1.0 : 0.0)) > tol + 1e-5 * (i == j ? 1.0 : 0.0)) return false;
    return true;
}

namespace {
inline Vec3d toD(const Vec3f& a) { return Vec3d{ a[0], a[1], a[2] }; }
inline Vec3f toF(const Vec3d& a) { return Vec3f{ (float)a[0], (float)a[1], (float)a[2] }; }
inline Vec3d sub(const Vec3d& a, const Vec3d& b) { return Vec3d{ a[0] - b[0], a[1] - b[1], a[2] - b[2] }; }
inline Vec3d add(const Vec3d& a, const Vec3d& b) { return Vec3d{ a[0] + b[0], a[1] + b[1], a[2] + b[2] }; }
inline Vec3d neg(const Vec3d& a) { return Vec3d{ -a[0], -a[1], -a[2] }; }
inline Vec3d cross(const Vec3d& a, const Vec3d& b)
{
    return Vec3d{ a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0] };
}
inline double norm(const Vec3d& a) { return std::sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]); }
inline Vec3d  scaled(const Vec3d& a, double s) { return Vec3d{ a[0] * s, a[1] * s, a[2] * s }; }

inline Vec3f subf(const Vec3f& a, const Vec3f& b) { return Vec3f{ a[0] - b[0], a[1] - b[1], a[2] - b[2] }; }
inline Vec3f addf(const Vec3f& a, const Vec3f& b) { return Vec3f{ a[0] + b[0], a[1] + b[1], a[2] + b[2] }; }
inline Vec3f crossf(const Vec3f& a, const Vec3f& b)
{
    return Vec3f{ a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0] };
}
inline float dotf(const Vec3f& a, const Vec3f& b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
inline Vec3f normalizedf(const Vec3f& a)
{
    const float l = std::sqrt(dotf(a, a));
    return Vec3f{ a[0] / l, a[1] / l, a[2] / l };
}
constexpr float kPi = 3.14159265358979323846f;

// Tangent::frame == frame_duff (src/runtime/math/Tangent.h:50-72)
void frame(const Vec3f& n, Vec3f& nx, Vec3f& ny)
{
    const float sign = std::copysign(1.0f, n[2]);
    const float a    = -1.0f / (sign + n[2]);
    const float b    = n[0] * n[1] * a;
    nx               = Vec3f{ 1.0f + sign * n[0] * n[0] * a, sign * b, -sign * n[0] };
    ny               = Vec3f{ b, sign + n[1] * n[1] * a, -n[1] };
}

// addGrid(..., 1, 1) == addPlane (TriMesh.cpp:700-745)
void addGrid(Mesh& m, const Vec3d& o, const Vec3d& x, const Vec3d& y)
{
    const uint32_t off = (uint32_t)m.v.size();
    Vec3d          nrm = cross(x, y);
    nrm                = scaled(nrm, 1.0 / norm(nrm));
    const Vec3d vs[4]  = { o, add(o, x), add(o, y), add(add(o, x), y) };
    const float ts[4][2] = { { 0, 0 }, { 1, 0 }, { 0, 1 }, { 1, 1 } };
    for (int i = 0; i < 4; ++i) {
        m.v.push_back(toF(vs[i]));
        m.n.push_back(toF(nrm));
        m.t.push_back({ ts[i][0], ts[i][1] });
    }
    m.f.push_back({ off + 0, off + 1, off + 3 });
    m.f.push_back({ off + 0, off + 3, off + 2 });
}

// addDisk (TriMesh.cpp:747-780)
void addDisk(Mesh& m, const Vec3f& origin, const Vec3f& n, const Vec3f& nx, const Vec3f& ny, float radius, int sections, bool fill_cap,
             bool flip = false)
{
    const float    step = 1.0f / (float)sections;
    const uint32_t off  = (uint32_t)m.v.size();
    if (fill_cap) {
        m.v.push_back(origin);
        m.n.push_back(n);
        m.t.push_back({ 0.0f, 0.0f });
    }
    for (int i = 0; i < sections; ++i) {
        const float ang = 2.0f * kPi * step * (float)i;
        const float x = std::cos(ang), y = std::sin(ang);
        Vec3f       p;
        for (int k = 0; k < 3; ++k) p[k] = radius * nx[k] * x + radius * ny[k] * y + origin[k];
        m.v.push_back(p);
        m.n.push_back(n);
        m.t.push_back({ 0.5f * (x + 1.0f), 0.5f * (y + 1.0f) });
    }
    if (!fill_cap) return;
    for (int i = 0; i < sections; ++i) {
        const uint32_t c = (uint32_t)i + 1, nc = (uint32_t)(i + 1 < sections ? i + 1 : 0) + 1;
        if (flip)
            m.f.push_back({ off, nc + off, c + off });
        else
            m.f.push_back({ off, c + off, nc + off });
    }
}
} // namespace

// ------------------------------------------------------------------------------------------------ Mesh
void Mesh::computeVertexNormals()
{
    std::vector<Vec3d> acc(v.size(), Vec3d{ 0, 0, 0 });
    std::vector<Vec3d> fn(f.size());
    for (size_t i = 0; i < f.size(); ++i) {
        const Vec3d a = toD(v[f[i][0]]), b = toD(v[f[i][1]]), c = toD(v[f[i][2]]);
        fn[i]         = cross(sub(b, a), sub(c, a));
    }
    for (int k = 0; k < 3; ++k)
        for (size_t i = 0; i < f.size(); ++i) acc[f[i][k]] = add(acc[f[i][k]], fn[i]);
    n.resize(v.size());
    for (size_t i = 0; i < v.size(); ++i) {
        const double l = std::max(norm(acc[i]), 1e-30);
        n[i]           = Vec3f{ (float)(acc[i][0] / l), (float)(acc[i][1] / l), (float)(acc[i][2] / l) };
    }
}

void Mesh::flipNormals() // TriMesh.cpp:34-43
{
    for (auto& tri : f) std::swap(tri[1], tri[2]);
    for (auto& x : n) x = Vec3f{ -x[0], -x[1], -x[2] };
}

void Mesh::transform(const Mat4d& m)
{
    for (auto& p : v) {
        const Vec3d q = toD(p);
        Vec3d       r;
        for (int i = 0; i < 3; ++i) r[i] = q[0] * m[i][0] + q[1] * m[i][1] + q[2] * m[i][2] + m[i][3];
        p = toF(r);
    }
    Mat4d lin = identity4();
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) lin[i][j] = m[i][j];
    const Mat4d inv = inverse(lin);
    for (auto& x : n) {
        const Vec3d q = toD(x);
        Vec3d       r;
        for (int j = 0; j < 3; ++j) r[j] = q[0] * inv[0][j] + q[1] * inv[1][j] + q[2] * inv[2][j]; // n * M^-1 == (M^-T) n
        const double l = std::max(norm(r), 1e-30);
        x              = Vec3f{ (float)(r[0] / l), (float)(r[1] / l), (float)(r[2] / l) };
    }
}

void Mesh::faceNormalsAsVertexNormals() // TriMesh.cpp:152-197
{
    Mesh out;
    out.v.reserve(3 * f.size());
    for (size_t i = 0; i < f.size(); ++i) {
        const Vec3f a = v[f[i][0]], b = v[f[i][1]], c = v[f[i][2]];
        Vec3f       fn = crossf(subf(b, a), subf(c, a));
        const float l  = std::sqrt(fn[0] * fn[0] + fn[1] * fn[1] + fn[2] * fn[2]);
        fn             = Vec3f{ fn[0] / l, fn[1] / l, fn[2] / l };
        for (int k = 0; k < 3; ++k) {
            out.v.push_back(v[f[i][k]]);
            out.n.push_back(fn);
            out.t.push_back(t[f[i][k]]);
        }
        const uint32_t o = (uint32_t)(3 * i);
        out.f.push_back({ o, o + 1, o + 2 });
    }
    *this = std::move(out);
}

void Mesh::subdivide() // TriMesh::subdivide without mask (TriMesh.cpp:351-401), edge ids in first-seen order (:300-321)
{
    std::map<std::pair<uint32_t, uint32_t>, uint32_t> edges;
    std::vector<std::pair<uint32_t, uint32_t>>        order;
    auto key = [](uint32_t a, uint32_t b) { return std::make_pair(std::min(a, b), std::max(a, b)); };
    for (const auto& tri : f)
        for (int j = 0; j < 3; ++j) {
            const auto k = key(tri[j], tri[(j + 1) % 3]);
            if (edges.emplace(k, (uint32_t)order.size()).second) order.push_back(k);
        }
    const uint32_t nv = (uint32_t)v.size();
    for (const auto& e : order) {
        const uint32_t a = e.first, b = e.second;
        v.push_back(Vec3f{ (v[a][0] + v[b][0]) / 2.0f, (v[a][1] + v[b][1]) / 2.0f, (v[a][2] + v[b][2]) / 2.0f });
        const Vec3f s = addf(n[a], n[b]);
        const float l = std::sqrt(s[0] * s[0] + s[1] * s[1] + s[2] * s[2]);
        n.push_back(Vec3f{ s[0] / l, s[1] / l, s[2] / l });
        t.push_back({ (t[a][0] + t[b][0]) / 2.0f, (t[a][1] + t[b][1]) / 2.0f });
    }
    std::vector<std::array<uint32_t, 3>> nf;
    nf.reserve(4 * f.size());
    for (const auto& tri : f) {
        const uint32_t v0 = tri[0], v1 = tri[1], v2 = tri[2];
        const uint32_t e01 = nv + edges[key(v0, v1)], e12 = nv + edges[key(v1, v2)], e20 = nv + edges[key(v2, v0)];
        nf.push_back({ v0, e01, e20 });
        nf.push_back({ v1, e12, e01 });
        nf.push_back({ v2, e20, e12 });
        nf.push_back({ e01, e12, e20 });
    }
    f = std::move(nf);
}

void Mesh::finish(bool has_normals, bool has_texcoords)
{
    if (!has_texcoords) t.assign(v.size(), { 0.0f, 0.0f });
    if (!has_normals) computeVertexNormals();
}

// ------------------------------------------------------------------------------------------------ procedural shapes
Mesh makePlane(const Vec3d& origin, const Vec3d& x_axis, const Vec3d& y_axis)
{
    Mesh m;
    addGrid(m, origin, x_axis, y_axis);
    return m;
}

Mesh makeBox(const Vec3d& o, const Vec3d& x, const Vec3d& y, const Vec3d& z) // MakeBox, TriMesh.cpp:967-1003
{
    const Vec3d hhh = add(add(add(o, x), y), z);
    Mesh        m;
    addGrid(m, o, y, x);
    addGrid(m, o, x, z);
    addGrid(m, o, z, y);
    addGrid(m, hhh, neg(x), neg(y));
    addGrid(m, hhh, neg(z), neg(x));
    addGrid(m, hhh, neg(y), neg(z));
    return m;
}

static void addTriangle(Mesh& m, const Vec3d& o, const Vec3d& a, const Vec3d& b)
{
    const uint32_t off = (uint32_t)m.v.size();
    Vec3d          n   = cross(a, b);
    n                  = scaled(n, 1.0 / norm(n));
    const Vec3d vs[3]  = { o, add(o, a), add(o, b) };
    const float ts[3][2] = { { 0, 0 }, { 1, 0 }, { 0, 1 } };
    for (int i = 0; i < 3; ++i) {
        m.v.push_back(toF(vs[i]));
        m.n.push_back(toF(n));
        m.t.push_back({ ts[i][0], ts[i][1] });
    }
    m.f.push_back({ off, off + 1, off + 2 });
}

Mesh makeTriangle(const Vec3d& p0, const Vec3d& p1, const Vec3d& p2)
{
    Mesh m;
    addTriangle(m, p0, sub(p1, p0), sub(p2, p0));
    return m;
}

Mesh makeRectangle(const Vec3d& p0, const Vec3d& p1, const Vec3d& p2, const Vec3d& p3) // MakeRectangle: two addTriangle
{
    Mesh m;
    addTriangle(m, p0, sub(p1, p0), sub(p3, p0));
    addTriangle(m, p1, sub(p2, p1), sub(p3, p1));
    return m;
}

Mesh makeDisk(const Vec3f& center, const Vec3f& normal, float radius, int sections) // MakeDisk :955-965
{
    sections = std::max(3, sections);
    Vec3f nx, ny;
    frame(normal, nx, ny);
    Mesh m;
    addDisk(m, center, normal, nx, ny, radius, sections, true);
    return m;
}

Mesh makeCone(const Vec3f& base_center, float base_radius, const Vec3f& tip, int sections, bool fill_cap) // MakeCone :1005-1031
{
    sections      = std::max(3, sections);
    const Vec3f h = normalizedf(subf(base_center, tip));
    Vec3f       nx, ny;
    frame(h, nx, ny);
    Mesh m;
    addDisk(m, base_center, h, nx, ny, base_radius, sections, fill_cap);
    m.v.push_back(tip);
    m.n.push_back(h);
    m.t.push_back({ 0.0f, 0.0f });
    const uint32_t start = fill_cap ? 1 : 0, tp = (uint32_t)m.v.size() - 1;
    for (int i = 0; i < sections; ++i) {
        const uint32_t c = (uint32_t)i + start, nc = (uint32_t)(i + 1 < sections ? i + 1 : 0) + start;
        m.f.push_back({ c, tp, nc });
    }
    m.computeVertexNormals();
    return m;
}

Mesh makeCylinder(const Vec3f& base_center, float base_radius, const Vec3f& top_center, float top_radius, int sections, bool fill_cap)
{ // MakeCylinder :1033-1057
    sections      = std::max(3, sections);
    const Vec3f h = normalizedf(subf(base_center, top_center));
    Vec3f       nx, ny;
    frame(h, nx, ny);
    Mesh m;
    addDisk(m, base_center, h, nx, ny, base_radius, sections, fill_cap);
    const uint32_t off = (uint32_t)m.v.size();
    addDisk(m, top_center, h, nx, ny, top_radius, sections, fill_cap, true);
    const uint32_t start = fill_cap ? 1 : 0;
    for (int i = 0; i < sections; ++i) {
        const uint32_t c = (uint32_t)i + start, nc = (uint32_t)(i + 1 < sections ? i + 1 : 0) + start;
        m.f.push_back({ c, c + off, nc });
        m.f.push_back({ c + off, nc + off, nc });
    }
    m.computeVertexNormals();
    return m;
}

Mesh makeUVSphere(const Vec3f& center, float radius, int stacks, int slices) // MakeUVSphere :782-835 (keeps its degenerate faces)
{
    const float drho = 3.141592f / (float)stacks, dtheta = 2.0f * 3.141592f / (float)slices;
    Mesh        m;
    for (int i = 0; i <= stacks; ++i) {
        const float rho  = (float)i * drho;
        const float srho = std::sin(rho), crho = std::cos(rho);
        for (int j = 0; j < slices; ++j) {
            const float theta  = (float)j * dtheta;
            const float stheta = -std::sin(theta), ctheta = std::cos(theta);
            const Vec3f n{ stheta * srho, ctheta * srho, crho };
            m.v.push_back(Vec3f{ n[0] * radius + center[0], n[1] * radius + center[1], n[2] * radius + center[2] });
            m.n.push_back(n);
            m.t.push_back({ 0.5f * theta / kPi, rho / kPi });
        }
    }
    for (int i = 0; i <= stacks; ++i) {
        const uint32_t cur = (uint32_t)(i * slices), nxt = (uint32_t)(((i + 1) % (stacks + 1)) * slices);
        for (int j = 0; j < slices; ++j) {
            const uint32_t nj  = (uint32_t)((j + 1) % slices);
            const uint32_t id0 = cur + (uint32_t)j, id1 = cur + nj, id2 = nxt + (uint32_t)j, id3 = nxt + nj;
            m.f.push_back({ id2, id3, id1 });
            m.f.push_back({ id2, id1, id0 });
        }
    }
    return m;
}

Mesh makeIcoSphere(const Vec3d& center, double radius, int subdivisions) // MakeIcoSphere :838-953
{
    const float        golden = 1.618033989f;
    std::vector<Vec3f> v;
    const int          signs[2] = { -1, 1 };
    for (int d = 0; d < 3; ++d)
        for (int s1 : signs)
            for (int s2 : signs) {
                Vec3f vec{ 0, 0, 0 };
                vec[(d + 1) % 3] = golden * (float)s1;
                vec[(d + 2) % 3] = (float)s2;
                v.push_back(normalizedf(vec));
            }
    auto idx = [](int d, int s1, int s2) { return (uint32_t)(d * 4 + (s1 + 1) + ((s2 + 1) >> 1)); };
    std::vector<std::array<uint32_t, 3>> f;
    for (int s1 : signs)
        for (int s2 : signs)
            for (int s3 : signs) {
                const bool     rev = s1 * s2 * s3 == -1;
                const uint32_t i1 = idx(0, s1, s2), i2 = idx(1, s2, s3), i3 = idx(2, s3, s1);
                if (rev)
                    f.push_back({ i1, i3, i2 });
                else
                    f.push_back({ i1, i2, i3 });
            }
    for (int d = 0; d < 3; ++d)
        for (int s1 : signs)
            for (int s2 : signs) {
                const bool     rev = s1 * s2 == 1;
                const uint32_t i2 = idx(d, s1, -1), i1 = idx(d, s1, 1), i3 = idx((d + 2) % 3, s2, s1);
                if (rev)
                    f.push_back({ i1, i3, i2 });
                else
                    f.push_back({ i1, i2, i3 });
            }
    for (int s = 0; s < subdivisions; ++s) {
        std::map<std::pair<uint32_t, uint32_t>, uint32_t> edge;
        for (const auto& tri : f)
            for (int j = 0; j < 3; ++j) {
                const uint32_t i1 = tri[j], i2 = tri[(j + 1) % 3];
                if (i1 >= i2) continue;
                edge[{ i1, i2 }] = (uint32_t)v.size();
                v.push_back(normalizedf(addf(v[i1], v[i2])));
            }
        std::vector<std::array<uint32_t, 3>> nf;
        nf.reserve(4 * f.size());
        for (const auto& tri : f) {
            uint32_t ec[3];
            for (int j = 0; j < 3; ++j) ec[j] = edge[{ std::min(tri[j], tri[(j + 1) % 3]), std::max(tri[j], tri[(j + 1) % 3]) }];
            nf.push_back({ ec[0], ec[1], ec[2] });
            for (int j = 0; j < 3; ++j) nf.push_back({ tri[j], ec[j % 3], ec[(j + 2) % 3] });
        }
        f = std::move(nf);
    }
    Mesh m;
    m.v = v;
    m.f = f;
    for (const auto& p : v) {
        const Vec3f n     = normalizedf(p);
        const float theta = std::acos(n[2]);
        float       phi   = std::atan2(-n[0], n[1]);
        if (phi < 0.0f) phi += 2.0f * kPi;
        m.n.push_back(n);
        m.t.push_back({ phi / (2.0f * kPi), theta / kPi });
    }
    if (center[0] != 0 || center[1] != 0 || center[2] != 0 || radius != 1.0) {
        Mat4d xf = identity4();
        for (int i = 0; i < 3; ++i) {
            xf[i][i] = radius;
            xf[i][3] = center[i];
        }
        m.transform(xf);
    }
    return m;
}

// ------------------------------------------------------------------------------------------------ OBJ / PLY
namespace {
std::string readAll(const std::string& path, bool binary)
{
    std::ifstream in(path, binary ? std::ios::binary : std::ios::in);
    if (!in) throw std::runtime_error("cannot open " + path);
    std::ostringstream ss;
    ss << in.rdbuf();
    return ss.str();
}
std::vector<std::string> splitWs(const std::string& line)
{
    std::vector<std::string> out;
    std::istringstream       ss(line);
    std::string              tok;
    while (ss >> tok) out.push_back(tok);
    return out;
}
} // namespace

Mesh loadObj(const std::string& path)
{
    std::vector<Vec3d>                 pos, nrm;
    std::vector<std::array<double, 2>> tex;
    using Key = std::tuple<int, int, int>; // (v, vn, vt)
    std::vector<std::array<Key, 3>> faces;
    std::istringstream              in(readAll(path, false));
    std::string                     line;
    while (std::getline(in, line)) {
        const auto p = splitWs(line);
        if (p.empty()) continue;
        auto num = [&](size_t i) { return i < p.size() ? std::strtod(p[i].c_str(), nullptr) : 0.0; };
        if (p[0] == "v") {
            pos.push_back({ num(1), num(2), num(3) });
        } else if (p[0] == "vn") {
            nrm.push_back({ num(1), num(2), num(3) });
        } else if (p[0] == "vt") {
            tex.push_back({ num(1), num(2) });
        } else if (p[0] == "f") {
            std::vector<Key> idx;
            for (size_t k = 1; k < p.size(); ++k) {
                std::string q[3];
                size_t      part = 0;
                for (char c : p[k]) {
                    if (c == '/') {
                        if (++part > 2) break;
                    } else {
                        q[part] += c;
                    }
                }
                auto fix = [](const std::string& s, size_t n) {
                    if (s.empty()) return -1;
                    const int i = std::atoi(s.c_str());
                    return i > 0 ? i - 1 : (int)n + i;
                };
                idx.emplace_back(fix(q[0], pos.size()), fix(q[2], nrm.size()), fix(q[1], tex.size()));
            }
            for (size_t k = 1; k + 1 < idx.size(); ++k) faces.push_back({ idx[0], idx[k], idx[k + 1] });
        }
    }
    bool has_n = false, has_t = false;
    for (const auto& fc : faces)
        for (const Key& k : fc) {
            has_n = has_n || std::get<1>(k) >= 0;
            has_t = has_t || std::get<2>(k) >= 0;
        }
    std::map<Key, uint32_t> remap;
    Mesh                    m;
    for (const auto& fc : faces) {
        std::array<uint32_t, 3> o{};
        for (int c = 0; c < 3; ++c) {
            const Key& k  = fc[c];
            auto       it = remap.find(k);
            if (it == remap.end()) {
                it = remap.emplace(k, (uint32_t)m.v.size()).first;
                if (std::get<0>(k) < 0 || (size_t)std::get<0>(k) >= pos.size()) throw std::runtime_error("OBJ vertex index out of range in " + path);
                m.v.push_back(toF(pos[(size_t)std::get<0>(k)]));
                if (has_n) m.n.push_back(std::get<1>(k) >= 0 && (size_t)std::get<1>(k) < nrm.size() ? toF(nrm[(size_t)std::get<1>(k)]) : Vec3f{ 0, 0, 1 });
                if (has_t) {
                    if (std::get<2>(k) >= 0 && (size_t)std::get<2>(k) < tex.size())
                        m.t.push_back({ (float)tex[(size_t)std::get<2>(k)][0], (float)tex[(size_t)std::get<2>(k)][1] });
                    else
                        m.t.push_back({ 0.0f, 0.0f });
                }
            }
            o[c] = it->second;
        }
        m.f.push_back(o);
    }
    m.finish(has_n, has_t);
    return m;
}

namespace {
struct PlyProp {
    std::string type, name, count_type, index_type; // list properties carry count / index types
    bool        list = false;
};
struct PlyElem {
    std::string          name;
    size_t               count = 0;
    std::vector<PlyProp> props;
};
size_t plySize(const std::string& t)
{
    if (t == "char" || t == "uchar" || t == "int8" || t == "uint8") return 1;
    if (t == "short" || t == "ushort" || t == "int16" || t == "uint16") return 2;
    if (t == "int" || t == "uint" || t == "float" || t == "int32" || t == "uint32" || t == "float32") return 4;
    if (t == "double" || t == "float64") return 8;
    throw std::runtime_error("PLY: unknown property type " + t);
}
double plyRead(const uint8_t* p, const std::string& t, bool big)
{
    uint8_t      b[8];
    const size_t n = plySize(t);
    for (size_t i = 0; i < n; ++i) b[i] = big ? p[n - 1 - i] : p[i];
    if (t == "char" || t == "int8") return (double)*reinterpret_cast<int8_t*>(b);
    if (t == "uchar" || t == "uint8") return (double)b[0];
    if (t == "short" || t == "int16") { int16_t x; std::memcpy(&x, b, 2); return x; }
    if (t == "ushort" || t == "uint16") { uint16_t x; std::memcpy(&x, b, 2); return x; }
    if (t == "int" || t == "int32") { int32_t x; std::memcpy(&x, b, 4); return x; }
    if (t == "uint" || t == "uint32") { uint32_t x; std::memcpy(&x, b, 4); return x; }
    if (t == "float" || t == "float32") { float x; std::memcpy(&x, b, 4); return x; }
    double x;
    std::memcpy(&x, b, 8);
    return x;
}
} // namespace

Mesh loadPly(const std::string& path)
{
    const std::string data = readAll(path, true);
    const size_t      hend = data.find("end_header\n");
    if (hend == std::string::npos) throw std::runtime_error("PLY: no end_header in " + path);
    const size_t          body = hend + 11;
    std::istringstream    hs(data.substr(0, body));
    std::string           line, fmt;
    std::vector<PlyElem>  elems;
    while (std::getline(hs, line)) {
        const auto p = splitWs(line);
        if (p.empty()) continue;
        if (p[0] == "format" && p.size() > 1) {
            fmt = p[1];
        } else if (p[0] == "element" && p.size() > 2) {
            PlyElem e;
            e.name  = p[1];
            e.count = (size_t)std::strtoull(p[2].c_str(), nullptr, 10);
            elems.push_back(e);
        } else if (p[0] == "property" && !elems.empty()) {
            PlyProp pr;
            if (p.size() >= 5 && p[1] == "list") {
                pr.list       = true;
                pr.count_type = p[2];
                pr.index_type = p[3];
                pr.name       = p[4];
            } else if (p.size() >= 3) {
                pr.type = p[1];
                pr.name = p[2];
            }
            elems.back().props.push_back(pr);
        }
    }
    std::map<std::string, std::vector<double>> cols;
    std::vector<std::array<uint32_t, 3>>       faces;
    size_t                                     nverts = 0;
    if (fmt == "ascii") {
        std::istringstream bs(data.substr(body));
        for (const PlyElem& e : elems) {
            if (e.name == "vertex") {
                nverts = e.count;
                for (const auto& pr : e.props) cols[pr.name].resize(e.count);
                for (size_t i = 0; i < e.count; ++i)
                    for (const auto& pr : e.props) {
                        std::string tok;
                        bs >> tok;
                        cols[pr.name][i] = std::strtod(tok.c_str(), nullptr);
                    }
            } else if (e.name == "face") {
                for (size_t i = 0; i < e.count; ++i) {
                    long n = 0;
                    bs >> n;
                    std::vector<uint32_t> idx((size_t)std::max(0l, n));
                    for (auto& x : idx) {
                        long v;
                        bs >> v;
                        x = (uint32_t)v;
                    }
                    for (size_t k = 1; k + 1 < idx.size(); ++k) faces.push_back({ idx[0], idx[k], idx[k + 1] });
                }
            } else {
                for (size_t i = 0; i < e.count; ++i)
                    for (const auto& pr : e.props) {
                        std::string tok;
                        bs >> tok;
                        if (pr.list) {
                            const long n = std::strtol(tok.c_str(), nullptr, 10);
                            for (long k = 0; k < n; ++k) bs >> tok;
                        }
                    }
            }
        }
    } else if (fmt == "binary_little_endian" || fmt == "binary_big_endian") {
        const bool     big = fmt == "binary_big_endian";
        const uint8_t* p   = reinterpret_cast<const uint8_t*>(data.data()) + body;
        const uint8_t* end = reinterpret_cast<const uint8_t*>(data.data()) + data.size();
        auto need = [&](size_t n) { if ((size_t)(end - p) < n) throw std::runtime_error("PLY: truncated body in " + path); };
        for (const PlyElem& e : elems) {
            if (e.name == "vertex") {
                nverts = e.count;
                for (const auto& pr : e.props) cols[pr.name].resize(e.count);
                for (size_t i = 0; i < e.count; ++i)
                    for (const auto& pr : e.props) {
                        if (pr.list) throw std::runtime_error("PLY: list property on vertices");
                        const size_t s = plySize(pr.type);
                        need(s);
                        cols[pr.name][i] = plyRead(p, pr.type, big);
                        p += s;
                    }
            } else {
                for (size_t i = 0; i < e.count; ++i)
                    for (const auto& pr : e.props) {
                        if (!pr.list) {
                            const size_t s = plySize(pr.type);
                            need(s);
                            p += s;
                            continue;
                        }
                        const size_t cs = plySize(pr.count_type), is = plySize(pr.index_type);
                        need(cs);
                        const size_t n = (size_t)plyRead(p, pr.count_type, big);
                        p += cs;
                        need(n * is);
                        if (e.name == "face" && &pr == &e.props[0]) {
                            std::vector<uint32_t> idx(n);
                            for (size_t k = 0; k < n; ++k) idx[k] = (uint32_t)plyRead(p + k * is, pr.index_type, big);
                            for (size_t k = 1; k + 1 < n; ++k) faces.push_back({ idx[0], idx[k], idx[k + 1] });
                        }
                        p += n * is;
                    }
            }
        }
    } else {
        throw std::runtime_error("PLY: unsupported format '" + fmt + "' in " + path);
    }
    if (!cols.count("x") || !cols.count("y") || !cols.count("z")) throw std::runtime_error("PLY: no vertex positions in " + path);
    Mesh m;
    m.v.resize(nverts);
    for (size_t i = 0; i < nverts; ++i) m.v[i] = Vec3f{ (float)cols["x"][i], (float)cols["y"][i], (float)cols["z"][i] };
    const bool has_n = cols.count("nx") && cols.count("ny") && cols.count("nz");
    if (has_n) {
        m.n.resize(nverts);
        for (size_t i = 0; i < nverts; ++i) m.n[i] = Vec3f{ (float)cols["nx"][i], (float)cols["ny"][i], (float)cols["nz"][i] };
    }
    const char* tu = cols.count("u") ? "u" : (cols.count("s") ? "s" : nullptr);
    const char* tv = tu ? (std::string(tu) == "u" ? "v" : "t") : nullptr;
    const bool  has_t = tu && cols.count(tv);
    if (has_t) {
        m.t.resize(nverts);
        for (size_t i = 0; i < nverts; ++i) m.t[i] = { (float)cols[tu][i], (float)cols[tv][i] };
    }
    for (const auto& tri : faces)
        for (uint32_t i : tri)
            if (i >= nverts) throw std::runtime_error("PLY: face index out of range in " + path);
    m.f = std::move(faces);
    m.finish(has_n, has_t);
    return m;
}

// ------------------------------------------------------------------------------------------------ plane detection
bool meshAsPlane(const Mesh& m, Vec3f& origin, Vec3f& x_axis, Vec3f& y_axis)
{
    if (m.f.size() != 2 || m.v.size() != 4) return false;
    auto fn = [&](const std::array<uint32_t, 3>& t) { return crossf(subf(m.v[t[1]], m.v[t[0]]), subf(m.v[t[2]], m.v[t[0]])); };
    const Vec3f n0 = normalizedf(fn(m.f[0])), n1 = normalizedf(fn(m.f[1]));
    for (int k = 0; k < 3; ++k)
        if (!(std::fabs((double)n0[k] - (double)n1[k]) <= 1e-5 + 1e-5 * std::fabs((double)n1[k]))) return false;
    origin = m.v[0];
    auto angle = [&](int start) {
        const Vec3f a = subf(m.v[(size_t)((start + 0) % 3 + 1)], origin), b = subf(m.v[(size_t)((start + 1) % 3 + 1)], origin);
        const float d = dotf(normalizedf(a), normalizedf(b));
        return std::fabs(std::acos((double)std::min(1.0f, std::max(-1.0f, d))));
    };
    const double a12 = angle(0), a23 = angle(1), a31 = angle(2);
    int          sel;
    if (a12 >= a23 && a12 >= a31)
        sel = 0;
    else if (a23 >= a31 && a23 >= a12)
        sel = 1;
    else
        sel = 2;
    x_axis = subf(m.v[(size_t)((sel + 0) % 3 + 1)], origin);
    y_axis = subf(m.v[(size_t)((sel + 1) % 3 + 1)], origin);
    if (dotf(n0, crossf(x_axis, y_axis)) < 0.0f) std::swap(x_axis, y_axis);
    return true;
}

} // namespace loader
} // namespace IG
