// SceneMesh.h -- triangle meshes of the scene loader: procedural shapes, OBJ / PLY readers, normals, subdivision.
//
// Vertex and face ORDER follows the reference (primitive ids are face order): src/runtime/mesh/TriMesh.cpp:700-1057
// (MakePlane / MakeBox / addDisk / MakeUVSphere / MakeIcoSphere / MakeDisk / MakeCone / MakeCylinder), :34-43 (flipNormals),
// :152-197 (setupFaceNormalsAsVertexNormals), :300-401 (subdivide), src/runtime/mesh/ObjFile.cpp:113-196 (fan
// triangulation, first-appearance numbering of (v, vn, vt) triples), src/runtime/mesh/PlyFile.cpp (ascii and binary).
// Arithmetic is float32 where the reference computes in Eigen::Vector3f, double where it only matters to rounding.
#pragma once

#include <array>
#include <cstdint>
#include <string>
#include <vector>

namespace IG {
namespace loader {

using Vec3d = std::array<double, 3>;
using Vec3f = std::array<float, 3>;
using Mat4d = std::array<std::array<double, 4>, 4>;

Mat4d identity4();
Mat4d mul(const Mat4d& a, const Mat4d& b);
Mat4d inverse(const Mat4d& m); // affine or general 4x4, Gauss-Jordan with partial pivoting
bool  isIdentity(const Mat4d& m, double tol = 1e-8);

struct Mesh {
    std::vector<Vec3f>                   v, n;
    std::vector<std::array<float, 2>>    t;
    std::vector<std::array<uint32_t, 3>> f;

    void computeVertexNormals();
    void flipNormals();
    void transform(const Mat4d& m);
    void faceNormalsAsVertexNormals();
    void subdivide();
    /// fills missing normals (area-weighted vertex normals) and texture coordinates (zero)
    void finish(bool has_normals, bool has_texcoords);
};

Mesh makePlane(const Vec3d& origin, const Vec3d& x_axis, const Vec3d& y_axis);
Mesh makeBox(const Vec3d& origin, const Vec3d& x, const Vec3d& y, const Vec3d& z);
Mesh makeTriangle(const Vec3d& p0, const Vec3d& p1, const Vec3d& p2);
Mesh makeRectangle(const Vec3d& p0, const Vec3d& p1, const Vec3d& p2, const Vec3d& p3);
Mesh makeDisk(const Vec3f& center, const Vec3f& normal, float radius, int sections);
Mesh makeCone(const Vec3f& base_center, float base_radius, const Vec3f& tip, int sections, bool fill_cap);
Mesh makeCylinder(const Vec3f& base_center, float base_radius, const Vec3f& top_center, float top_radius, int sections, bool fill_cap);
Mesh makeUVSphere(const Vec3f& center, float radius, int stacks, int slices);
Mesh makeIcoSphere(const Vec3d& center, double radius, int subdivisions);

/// throw std::runtime_error on unreadable or malformed files
Mesh loadObj(const std::string& path);
Mesh loadPly(const std::string& path);

/// TriMesh::getAsPlane (src/runtime/mesh/TriMesh.cpp:520-628) for the 4-vertex / 2-face case
bool meshAsPlane(const Mesh& m, Vec3f& origin, Vec3f& x_axis, Vec3f& y_axis);

} // namespace loader
} // namespace IG
