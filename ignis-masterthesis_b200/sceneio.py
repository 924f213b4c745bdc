"""Minimal scene ingest for the reference's JSON scene format -- harness level, not the product.

Covers what the BASELINE configs and the evaluation scenes used as golden images need: perspective camera,
path / debug technique, diffuse / dielectric / mirror / conductor BSDFs, rectangle / triangle / cube / sphere /
obj / ply shapes, entities with transforms, point / spot / directional / env (constant) / area lights.
It follows the reference loader where ordering matters for parity:
  * transforms: src/runtime/loader/Parser.cpp:124-243 (row-major 3x4 / 4x4 arrays; op lists multiply left to right)
  * procedural meshes: src/runtime/mesh/TriMesh.cpp:700-745,967-1003; flip_normals :34-43
  * OBJ: fan triangulation + first-appearance vertex numbering (src/runtime/mesh/ObjFile.cpp:113-180)
  * entity / material numbering: src/runtime/loader/LoaderEntity.cpp:42-103 (material groups, area-light
    entities get a material of their own)
  * lights: infinite / finite tables in scene order (src/runtime/loader/LoaderLight.cpp)
"""
import json
import math
import os
import struct

import numpy as np

from .shading import Shading, as_plane
from .tables import SceneBuilder


# ---------------------------------------------------------------------------------------- transforms
def _look_at(eye, center, up):
    eye, center, up = [np.asarray(a, np.float64) for a in (eye, center, up)]
    f = center - eye
    f = f / np.linalg.norm(f) if np.linalg.norm(f) > 0 else np.array([0, 0, 1.0])
    u = up / np.linalg.norm(up)
    s = np.cross(f, u)
    s = s / np.linalg.norm(s)
    u = np.cross(s, f)
    m = np.eye(4)
    m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = s, u, f, eye
    return m


def _rot(axis, deg):
    a = math.radians(deg)
    c, s = math.cos(a), math.sin(a)
    m = np.eye(4)
    i, j = [(1, 2), (2, 0), (0, 1)][axis]
    m[i, i], m[i, j], m[j, i], m[j, j] = c, -s, s, c
    return m


class _Pairs(dict):
    """A JSON object that remembers repeated keys (rapidjson keeps every member; a transform may hold two "qrotate" ops)."""

    def __init__(self, pairs):
        super().__init__(pairs)
        self.pairs = list(pairs)


def _strip_json_comments(text):
    """rapidjson's kParseCommentsFlag | kParseTrailingCommasFlag (Parser.cpp:546): // and /* */ outside strings, `,]` `,}`."""
    out, i, n, in_str = [], 0, len(text), False
    while i < n:
        c = text[i]
        if in_str:
            out.append(c)
            if c == "\\" and i + 1 < n:
                out.append(text[i + 1])
                i += 1
            elif c == '"':
                in_str = False
        elif c == '"':
            in_str = True
            out.append(c)
        elif text.startswith("//", i):
            while i < n and text[i] != "\n":
                i += 1
            continue
        elif text.startswith("/*", i):
            i = text.find("*/", i + 2)
            i = n if i < 0 else i + 2
            continue
        else:
            out.append(c)
        i += 1
    import re
    return re.sub(r",(\s*[\]}])", r"\1", "".join(out))


def load_json(path_or_text, is_text=False):
    """Scene JSON the way the reference's parser reads it: comments, trailing commas, repeated keys in transform objects."""
    if not is_text:
        with open(path_or_text) as fh:
            path_or_text = fh.read()
    def hook(pairs):
        keys = [k for k, _ in pairs]
        return _Pairs(pairs) if len(set(keys)) != len(keys) else dict(pairs)
    return json.loads(_strip_json_comments(path_or_text), object_pairs_hook=hook)


def parse_transform(t):
    if t is None:
        return np.eye(4)
    if isinstance(t, (dict, _Pairs)):  # a single object of ops, e.g. {"scale": 0.5} (Parser.cpp: ops applied in member order)
        t = [t]
    if isinstance(t, list) and t and isinstance(t[0], (dict, _Pairs)):
        m = np.eye(4)
        for op in t:
            for k, v in (op.pairs if isinstance(op, _Pairs) else op.items()):
                if k == "translate":
                    o = np.eye(4); o[:3, 3] = v
                elif k == "scale":
                    o = np.eye(4)
                    o[0, 0], o[1, 1], o[2, 2] = (v, v, v) if np.isscalar(v) else v
                elif k == "rotate":
                    o = _rot(0, v[0]) @ _rot(1, v[1]) @ _rot(2, v[2])
                elif k == "qrotate":  # Eigen::Quaternionf(w, x, y, z), Parser.cpp:96-105,192-193
                    qw, qx, qy, qz = [float(c) for c in v]
                    o = np.eye(4)
                    o[:3, :3] = [[1 - 2 * (qy * qy + qz * qz), 2 * (qx * qy - qz * qw), 2 * (qx * qz + qy * qw)],
                                 [2 * (qx * qy + qz * qw), 1 - 2 * (qx * qx + qz * qz), 2 * (qy * qz - qx * qw)],
                                 [2 * (qx * qz - qy * qw), 2 * (qy * qz + qx * qw), 1 - 2 * (qx * qx + qy * qy)]]
                elif k == "lookat":
                    origin = v.get("origin", [0, 0, 0])
                    target = np.asarray(v["direction"], float) + np.asarray(origin, float) if "direction" in v \
                        else v.get("target", [0, 1, 0])
                    o = _look_at(origin, target, v.get("up", [0, 0, 1]))
                elif k == "matrix":
                    o = parse_transform(v)
                else:
                    raise ValueError(f"unsupported transform op {k}")
                m = m @ o
        return m
    a = np.asarray(t, np.float64).reshape(-1)
    m = np.eye(4)
    if a.size == 16:
        m = a.reshape(4, 4)
    elif a.size == 12:
        m[:3, :] = a.reshape(3, 4)
    elif a.size == 9:
        m[:3, :3] = a.reshape(3, 3)
    else:
        raise ValueError("transform must have 9, 12 or 16 numbers")
    return m


# ---------------------------------------------------------------------------------------- meshes
class Mesh:
    def __init__(self, v, f, n=None, t=None):
        self.v = np.asarray(v, np.float32).reshape(-1, 3)
        self.f = np.asarray(f, np.uint32).reshape(-1, 3)
        self.n = None if n is None else np.asarray(n, np.float32).reshape(-1, 3)
        self.t = np.zeros((self.v.shape[0], 2), np.float32) if t is None else np.asarray(t, np.float32).reshape(-1, 2)
        if self.n is None:
            self.compute_vertex_normals()

    def compute_vertex_normals(self):
        n = np.zeros_like(self.v, dtype=np.float64)
        v = self.v.astype(np.float64)
        fn = np.cross(v[self.f[:, 1]] - v[self.f[:, 0]], v[self.f[:, 2]] - v[self.f[:, 0]])
        for k in range(3):
            np.add.at(n, self.f[:, k], fn)
        ln = np.linalg.norm(n, axis=1, keepdims=True)
        self.n = (n / np.maximum(ln, 1e-30)).astype(np.float32)

    def flip_normals(self):
        self.f = self.f[:, [0, 2, 1]].copy()
        self.n = (-self.n).astype(np.float32)

    def transform(self, m):
        m = np.asarray(m, np.float64)
        self.v = (self.v.astype(np.float64) @ m[:3, :3].T + m[:3, 3]).astype(np.float32)
        nm = np.linalg.inv(m[:3, :3]).T
        n = self.n.astype(np.float64) @ nm.T
        self.n = (n / np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-30)).astype(np.float32)


def _grid(origin, x_axis, y_axis):  # addGrid(..., 1, 1) == addPlane
    o, x, y = [np.asarray(a, np.float64) for a in (origin, x_axis, y_axis)]
    v = [o, o + x, o + y, o + x + y]
    nrm = np.cross(x, y)
    nrm = nrm / np.linalg.norm(nrm)
    return v, [nrm] * 4, [(0, 0), (1, 0), (0, 1), (1, 1)], [(0, 1, 3), (0, 3, 2)]


def make_plane(origin, x_axis, y_axis):
    v, n, t, f = _grid(origin, x_axis, y_axis)
    return Mesh(v, f, n, t)


def make_box(origin, x, y, z):
    o, x, y, z = [np.asarray(a, np.float64) for a in (origin, x, y, z)]
    hhh = o + x + y + z
    vs, ns, ts, fs = [], [], [], []
    for org, a, b in ((o, y, x), (o, x, z), (o, z, y), (hhh, -x, -y), (hhh, -z, -x), (hhh, -y, -z)):
        v, n, t, f = _grid(org, a, b)
        off = len(vs)
        vs += v; ns += n; ts += t
        fs += [(i + off, j + off, k + off) for i, j, k in f]
    return Mesh(vs, fs, ns, ts)


# Procedural primitives: vertex and face ORDER follows src/runtime/mesh/TriMesh.cpp:747-1057 (primitive ids are the
# face order), arithmetic in float32 like the reference's Eigen Vector3f.
_F = np.float32
_PI = _F(3.14159265358979323846)


def _nrm(v):
    v = np.asarray(v, _F)
    return (v / _F(np.sqrt(np.dot(v, v), dtype=_F))).astype(_F)


def _frame(n):  # Tangent::frame == frame_duff, src/runtime/math/Tangent.h:50-72
    n = np.asarray(n, _F)
    sign = _F(math.copysign(1.0, float(n[2])))
    a = _F(-1.0) / (sign + n[2])
    b = n[0] * n[1] * a
    nx = np.array([_F(1) + sign * n[0] * n[0] * a, sign * b, -sign * n[0]], _F)
    ny = np.array([b, sign + n[1] * n[1] * a, -n[1]], _F)
    return nx, ny


class _Soup:
    def __init__(self):
        self.v, self.n, self.t, self.f = [], [], [], []

    def mesh(self, normals=True):
        return Mesh(self.v, self.f, self.n if normals else None, self.t)


def _add_disk(m, origin, n, nx, ny, radius, sections, fill_cap, flip=False):  # addDisk, TriMesh.cpp:747-780
    origin, radius = np.asarray(origin, _F), _F(radius)
    step = _F(1.0) / _F(sections)
    off = len(m.v)
    if fill_cap:
        m.v.append(origin); m.n.append(n); m.t.append((0, 0))
    for i in range(sections):
        ang = _F(2) * _PI * step * _F(i)
        x, y = _F(np.cos(ang, dtype=_F)), _F(np.sin(ang, dtype=_F))
        m.v.append(radius * nx * x + radius * ny * y + origin)
        m.n.append(n)
        m.t.append((_F(0.5) * (x + _F(1)), _F(0.5) * (y + _F(1))))
    if not fill_cap:
        return
    for i in range(sections):
        c, nc = i + 1, (i + 1 if i + 1 < sections else 0) + 1
        m.f.append((off, nc + off, c + off) if flip else (off, c + off, nc + off))


def make_disk(center, normal, radius, sections):  # MakeDisk :955-965
    sections = max(3, int(sections))
    n = np.asarray(normal, _F)
    nx, ny = _frame(n)
    m = _Soup()
    _add_disk(m, center, n, nx, ny, radius, sections, True)
    return m.mesh()


def make_cone(base_center, base_radius, tip, sections, fill_cap):  # MakeCone :1005-1031
    sections = max(3, int(sections))
    base_center, tip = np.asarray(base_center, _F), np.asarray(tip, _F)
    h = _nrm(base_center - tip)
    nx, ny = _frame(h)
    m = _Soup()
    _add_disk(m, base_center, h, nx, ny, base_radius, sections, fill_cap)
    m.v.append(tip); m.n.append(h); m.t.append((0, 0))
    start, tp = (1 if fill_cap else 0), len(m.v) - 1
    for i in range(sections):
        c, nc = i + start, (i + 1 if i + 1 < sections else 0) + start
        m.f.append((c, tp, nc))
    return m.mesh(normals=False)  # computeVertexNormals


def make_cylinder(base_center, base_radius, top_center, top_radius, sections, fill_cap):  # MakeCylinder :1033-1057
    sections = max(3, int(sections))
    base_center, top_center = np.asarray(base_center, _F), np.asarray(top_center, _F)
    h = _nrm(base_center - top_center)
    nx, ny = _frame(h)
    m = _Soup()
    _add_disk(m, base_center, h, nx, ny, base_radius, sections, fill_cap)
    off = len(m.v)
    _add_disk(m, top_center, h, nx, ny, top_radius, sections, fill_cap, True)
    start = 1 if fill_cap else 0
    for i in range(sections):
        c, nc = i + start, (i + 1 if i + 1 < sections else 0) + start
        m.f.append((c, c + off, nc))
        m.f.append((c + off, nc + off, nc))
    return m.mesh(normals=False)


def make_uv_sphere(center, radius, stacks, slices):  # MakeUVSphere :782-835 (keeps its degenerate pole/wrap faces)
    center, radius = np.asarray(center, _F), _F(radius)
    drho, dtheta = _F(3.141592) / _F(stacks), _F(2) * _F(3.141592) / _F(slices)
    m = _Soup()
    for i in range(stacks + 1):
        rho = _F(i) * drho
        srho, crho = _F(np.sin(rho, dtype=_F)), _F(np.cos(rho, dtype=_F))
        for j in range(slices):
            theta = _F(j) * dtheta
            stheta, ctheta = -_F(np.sin(theta, dtype=_F)), _F(np.cos(theta, dtype=_F))
            n = np.array([stheta * srho, ctheta * srho, crho], _F)
            m.v.append(n * radius + center); m.n.append(n)
            m.t.append((_F(0.5) * theta / _PI, rho / _PI))
    for i in range(stacks + 1):
        cur, nxt = i * slices, ((i + 1) % (stacks + 1)) * slices
        for j in range(slices):
            nj = (j + 1) % slices
            id0, id1, id2, id3 = cur + j, cur + nj, nxt + j, nxt + nj
            m.f.append((id2, id3, id1))
            m.f.append((id2, id1, id0))
    return m.mesh()


def make_ico_sphere(center, radius, subdivisions):  # MakeIcoSphere :838-953
    golden = _F(1.618033989)
    v = []
    for d in range(3):
        for s1 in (-1, 1):
            for s2 in (-1, 1):
                vec = np.zeros(3, _F)
                vec[(d + 1) % 3] = golden * _F(s1)
                vec[(d + 2) % 3] = _F(s2)
                v.append(_nrm(vec))
    idx = lambda d, s1, s2: d * 4 + (s1 + 1) + ((s2 + 1) >> 1)
    f = []
    for s1 in (-1, 1):
        for s2 in (-1, 1):
            for s3 in (-1, 1):
                rev = s1 * s2 * s3 == -1
                i1, i2, i3 = idx(0, s1, s2), idx(1, s2, s3), idx(2, s3, s1)
                f.append((i1, i3, i2) if rev else (i1, i2, i3))
    for d in range(3):
        for s1 in (-1, 1):
            for s2 in (-1, 1):
                rev = s1 * s2 == 1
                i2, i1, i3 = idx(d, s1, -1), idx(d, s1, 1), idx((d + 2) % 3, s2, s1)
                f.append((i1, i3, i2) if rev else (i1, i2, i3))
    for _ in range(int(subdivisions)):
        edge = {}
        for tri in f:
            for j in range(3):
                i1, i2 = tri[j], tri[(j + 1) % 3]
                if i1 >= i2:
                    continue
                edge[(i1, i2)] = len(v)
                v.append(_nrm(v[i1] + v[i2]))
        nf = []
        for tri in f:
            ec = [edge[(min(tri[j], tri[(j + 1) % 3]), max(tri[j], tri[(j + 1) % 3]))] for j in range(3)]
            nf.append((ec[0], ec[1], ec[2]))
            for j in range(3):
                nf.append((tri[j], ec[j % 3], ec[(j + 2) % 3]))
        f = nf
    v = np.asarray(v, _F)
    n = np.stack([_nrm(x) for x in v])
    theta = np.arccos(n[:, 2], dtype=_F)
    phi = np.arctan2(-n[:, 0], n[:, 1], dtype=_F)
    phi = np.where(phi < 0, phi + _F(2) * _PI, phi).astype(_F)
    t = np.stack([phi / (_F(2) * _PI), theta / _PI], axis=1)
    mesh = Mesh(v, f, n, t)
    center, radius = np.asarray(center, np.float64), float(radius)
    if np.any(center != 0) or radius != 1:
        xf = np.eye(4)
        xf[:3, :3] *= radius
        xf[:3, 3] = center
        mesh.transform(xf)
    return mesh


def subdivide(mesh):  # TriMesh::subdivide without mask, TriMesh.cpp:351-401 (edge ids in first-seen order :300-321)
    edges = {}
    for tri in mesh.f:
        for a, b in ((tri[0], tri[1]), (tri[1], tri[2]), (tri[2], tri[0])):
            edges.setdefault((min(a, b), max(a, b)), len(edges))
    nv = mesh.v.shape[0]
    ea = np.array([k[0] for k in edges], np.int64)
    eb = np.array([k[1] for k in edges], np.int64)
    v = np.concatenate([mesh.v, ((mesh.v[ea] + mesh.v[eb]) / _F(2)).astype(_F)])
    nsum = mesh.n[ea] + mesh.n[eb]
    n = np.concatenate([mesh.n, (nsum / np.sqrt((nsum * nsum).sum(1, keepdims=True), dtype=_F)).astype(_F)])
    t = np.concatenate([mesh.t, ((mesh.t[ea] + mesh.t[eb]) / _F(2)).astype(_F)])
    f = []
    for v0, v1, v2 in mesh.f:
        e01 = nv + edges[(min(v0, v1), max(v0, v1))]
        e12 = nv + edges[(min(v1, v2), max(v1, v2))]
        e20 = nv + edges[(min(v2, v0), max(v2, v0))]
        f += [(v0, e01, e20), (v1, e12, e01), (v2, e20, e12), (e01, e12, e20)]
    return Mesh(v, f, n, t)


def face_normals_as_vertex_normals(mesh):  # setupFaceNormalsAsVertexNormals, TriMesh.cpp:152-197
    idx = mesh.f.reshape(-1)
    v = mesh.v[idx]
    t = mesh.t[idx]
    e1, e2 = v[1::3] - v[0::3], v[2::3] - v[0::3]
    fn = np.cross(e1, e2).astype(_F)
    fn = (fn / np.sqrt((fn * fn).sum(1, keepdims=True), dtype=_F)).astype(_F)
    return Mesh(v, np.arange(v.shape[0], dtype=np.uint32).reshape(-1, 3), np.repeat(fn, 3, axis=0), t)


def load_obj(path):
    """OBJ subset: v / vn / vt / f, fan triangulation, vertices numbered by first appearance of the
    (v, vn, vt) triple (ObjFile.cpp:113-173); missing normals -> computeVertexNormals (ObjFile.cpp:193-196)."""
    pos, nrm, tex, faces = [], [], [], []
    with open(path) as fh:
        for line in fh:
            p = line.split()
            if not p:
                continue
            if p[0] == "v":
                pos.append([float(x) for x in p[1:4]])
            elif p[0] == "vn":
                nrm.append([float(x) for x in p[1:4]])
            elif p[0] == "vt":
                tex.append([float(x) for x in p[1:3]])
            elif p[0] == "f":
                idx = []
                for tok in p[1:]:
                    q = (tok.split("/") + ["", ""])[:3]
                    fix = lambda s, n: -1 if s == "" else (int(s) - 1 if int(s) > 0 else n + int(s))
                    idx.append((fix(q[0], len(pos)), fix(q[2], len(nrm)), fix(q[1], len(tex))))
                for k in range(1, len(idx) - 1):
                    faces.append((idx[0], idx[k], idx[k + 1]))
    has_n = any(i[1] >= 0 for f in faces for i in f)
    has_t = any(i[2] >= 0 for f in faces for i in f)
    remap, verts, norms, texs, out = {}, [], [], [], []
    for tri in faces:
        o = []
        for key in tri:
            if key not in remap:
                remap[key] = len(verts)
                verts.append(pos[key[0]])
                if has_n:
                    norms.append(nrm[key[1]] if key[1] >= 0 else [0.0, 0.0, 1.0])
                if has_t:
                    texs.append(tex[key[2]] if key[2] >= 0 else [0.0, 0.0])
            o.append(remap[key])
        out.append(o)
    return Mesh(verts, out, norms if has_n else None, texs if has_t else None)


def load_ply(path):
    with open(path, "rb") as fh:
        data = fh.read()
    end = data.index(b"end_header\n") + len(b"end_header\n")
    header = data[:end].decode("ascii", "replace").splitlines()
    fmt, elems, cur = None, [], None
    for line in header:
        p = line.split()
        if not p:
            continue
        if p[0] == "format":
            fmt = p[1]
        elif p[0] == "element":
            cur = {"name": p[1], "count": int(p[2]), "props": []}
            elems.append(cur)
        elif p[0] == "property":
            cur["props"].append(p[1:])
    tmap = {"float": "f", "float32": "f", "double": "d", "uchar": "B", "uint8": "B", "char": "b", "int": "i", "int32": "i",
            "uint": "I", "uint32": "I", "short": "h", "ushort": "H"}
    body = data[end:]
    verts, faces = None, []
    if fmt == "ascii":
        toks = body.split()
        pos = 0
        for e in elems:
            if e["name"] == "vertex":
                k = len(e["props"])
                arr = np.array(toks[pos:pos + k * e["count"]], dtype=np.float64).reshape(e["count"], k)
                pos += k * e["count"]
                verts = {p[-1]: arr[:, i] for i, p in enumerate(e["props"])}
            elif e["name"] == "face":
                for _ in range(e["count"]):
                    n = int(toks[pos]); idx = [int(x) for x in toks[pos + 1:pos + 1 + n]]; pos += 1 + n
                    for k in range(1, n - 1):
                        faces.append((idx[0], idx[k], idx[k + 1]))
    else:
        en = "<" if fmt == "binary_little_endian" else ">"
        off = 0
        for e in elems:
            if e["name"] == "vertex":
                dt = np.dtype([(p[-1], en + tmap[p[0]].replace("f", "f4").replace("d", "f8").replace("B", "u1")
                                .replace("i", "i4").replace("I", "u4")) for p in e["props"]])
                arr = np.frombuffer(body, dt, e["count"], off)
                off += dt.itemsize * e["count"]
                verts = {name: arr[name].astype(np.float64) for name in dt.names}
            elif e["name"] == "face":
                lp = e["props"][0]  # list <count type> <index type> name
                ct, it = tmap[lp[1]], tmap[lp[2]]
                for _ in range(e["count"]):
                    n = struct.unpack_from(en + ct, body, off)[0]; off += struct.calcsize(ct)
                    idx = struct.unpack_from(en + it * n, body, off); off += struct.calcsize(it) * n
                    for k in range(1, n - 1):
                        faces.append((idx[0], idx[k], idx[k + 1]))
    v = np.stack([verts["x"], verts["y"], verts["z"]], axis=1)
    n = np.stack([verts["nx"], verts["ny"], verts["nz"]], axis=1) if "nx" in verts else None
    tn = "u" if "u" in verts else ("s" if "s" in verts else None)
    t = np.stack([verts[tn], verts["v" if tn == "u" else "t"]], axis=1) if tn else None
    return Mesh(v, faces, n, t)


# ---------------------------------------------------------------------------------------- scene
def _merge_externals(scene, base):
    for ext in scene.get("externals", []):
        path = os.path.join(base, ext["filename"])
        sub = load_json(path)
        sub = _merge_externals(sub, os.path.dirname(path))
        for k, v in sub.items():
            if isinstance(v, list):
                # externals are handled first "such that the current file can replace parts of it later on" (Parser.cpp:446);
                # objects live in name-keyed maps (Scene.h:43-48), so a later definition of a name replaces the earlier one
                merged, slot = [], {}
                for obj in v + scene.get(k, []):
                    name = obj.get("name") if isinstance(obj, dict) else None
                    if name is not None and name in slot:
                        merged[slot[name]] = obj
                    else:
                        if name is not None:
                            slot[name] = len(merged)
                        merged.append(obj)
                scene[k] = merged
            elif k not in scene:
                scene[k] = v
    return scene


def _color(v, default):
    if v is None:
        v = default
    if isinstance(v, str):  # constant PExpr "color(r, g, b[, a])" (loader/Transpiler.cpp); anything else is not folded
        import re
        m = re.fullmatch(r"\s*color\(\s*([^,()]+),\s*([^,()]+),\s*([^,()]+)(?:,\s*[^,()]+)?\)\s*", v)
        if not m:
            raise ValueError(f"expression '{v}' is outside the supported subset")
        v = [float(m.group(i)) for i in (1, 2, 3)]
    a = np.asarray(v, np.float32).reshape(-1)
    return np.repeat(a, 3) if a.size == 1 else a[:3]


def build_shape(elem, base):
    ty = elem["type"]
    if ty == "rectangle":
        if "p0" in elem:
            p0, p1, p2, p3 = [np.asarray(elem.get(k, d), np.float64) for k, d in
                              (("p0", (-1, -1, 0)), ("p1", (1, -1, 0)), ("p2", (1, 1, 0)), ("p3", (-1, 1, 0)))]
            vs, ns, ts, fs = [], [], [], []
            for o, a, b in ((p0, p1 - p0, p3 - p0), (p1, p2 - p1, p3 - p1)):  # MakeRectangle: two addTriangle
                n = np.cross(a, b); n = n / np.linalg.norm(n)
                off = len(vs)
                vs += [o, o + a, o + b]; ns += [n] * 3; ts += [(0, 0), (1, 0), (0, 1)]
                fs.append((off, off + 1, off + 2))
            mesh = Mesh(vs, fs, ns, ts)
        else:
            w, h = elem.get("width", 2.0), elem.get("height", 2.0)
            mesh = make_plane(elem.get("origin", (-w / 2, -h / 2, 0)), (w, 0, 0), (0, h, 0))
    elif ty == "triangle":
        p0, p1, p2 = [np.asarray(elem.get(k, d), np.float64) for k, d in
                      (("p0", (0, 0, 0)), ("p1", (1, 0, 0)), ("p2", (0, 1, 0)))]
        n = np.cross(p1 - p0, p2 - p0); n = n / np.linalg.norm(n)
        mesh = Mesh([p0, p1, p2], [(0, 1, 2)], [n] * 3, [(0, 0), (1, 0), (0, 1)])
    elif ty in ("cube", "box"):
        w, h, d = elem.get("width", 2.0), elem.get("height", 2.0), elem.get("depth", 2.0)
        mesh = make_box(elem.get("origin", (-w / 2, -h / 2, -d / 2)), (w, 0, 0), (0, h, 0), (0, 0, d))
    elif ty in ("obj", "ply", "external"):
        path = os.path.join(base, elem["filename"])
        mesh = load_obj(path) if path.lower().endswith(".obj") else load_ply(path)
    elif ty == "icosphere":  # setup_mesh_ico_sphere, TriMeshProvider.cpp:52-58
        mesh = make_ico_sphere(elem.get("center", (0, 0, 0)), elem.get("radius", 1.0), elem.get("subdivisions", 4))
    elif ty == "uvsphere":  # :60-67
        mesh = make_uv_sphere(elem.get("center", (0, 0, 0)), elem.get("radius", 1.0), elem.get("stacks", 32), elem.get("slices", 16))
    elif ty == "cylinder":  # :69-86
        r0 = elem.get("radius", 1.0) if "radius" in elem else elem.get("bottom_radius", 1.0)
        r1 = r0 if "radius" in elem else elem.get("top_radius", r0)
        mesh = make_cylinder(elem.get("p0", (0, 0, 0)), r0, elem.get("p1", (0, 0, 1)), r1, elem.get("sections", 32),
                             elem.get("filled", True))
    elif ty == "cone":  # :88-96
        mesh = make_cone(elem.get("p0", (0, 0, 0)), elem.get("radius", 1.0), elem.get("p1", (0, 0, 1)), elem.get("sections", 32),
                         elem.get("filled", True))
    elif ty == "disk":  # :98-105
        mesh = make_disk(elem.get("origin", (0, 0, 0)), elem.get("normal", (0, 0, 1)), elem.get("radius", 1.0),
                         elem.get("sections", 32))
    elif ty == "sphere":
        return ("sphere", np.asarray(elem.get("center", (0, 0, 0)), np.float32), float(elem.get("radius", 1.0)))
    else:
        raise ValueError(f"shape type '{ty}' is outside the supported subset")
    if elem.get("flip_normals", False):
        mesh.flip_normals()
    if elem.get("face_normals", False):  # TriMeshProvider.cpp:529-536
        mesh = face_normals_as_vertex_normals(mesh)
    elif elem.get("smooth_normals", False):
        mesh.compute_vertex_normals()
    if "transform" in elem:
        m = parse_transform(elem["transform"])
        if not np.allclose(m, np.eye(4)):
            mesh.transform(m)
    if elem.get("subdivision", 0) > 0:  # handleModification, TriMeshProvider.cpp:435-473 (refinement/displacement: no)
        for _ in range(int(elem["subdivision"])):
            mesh = subdivide(mesh)
        if elem.get("smooth_normals", False):
            mesh.compute_vertex_normals()
        else:
            mesh = face_normals_as_vertex_normals(mesh)
    return ("trimesh", mesh)


# ---------------------------------------------------------------------------------------- images / textures
def _srgb_byte_to_linear(c):  # byte_color_to_linear, src/runtime/Image.cpp:31-42
    x = c.astype(np.float32) / np.float32(255.0)
    lin = np.where(x <= np.float32(0.04045), x / np.float32(12.92),
                   np.power((x + np.float32(0.055)) / np.float32(1.055), np.float32(2.4), dtype=np.float32))
    return np.minimum(255, np.floor(lin * np.float32(255))).astype(np.uint8)


def load_packed_image(path, linear=False):
    """Image::loadAsPacked (src/runtime/Image.cpp:565-655): (h, w) uint32 RGBA8, bottom row first, sRGB bytes mapped
    to linear unless `linear`; returns (pixels, opaque)."""
    import cv2
    img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    if img is None:
        raise FileNotFoundError(path)
    if img.dtype != np.uint8:
        raise ValueError("packed images are 8 bit")
    if img.ndim == 2:
        img = np.repeat(img[:, :, None], 3, axis=2)
    opaque = img.shape[2] == 3
    rgb = img[:, :, 2::-1] if img.shape[2] >= 3 else img
    alpha = img[:, :, 3] if img.shape[2] == 4 else np.full(img.shape[:2], 255, np.uint8)
    if not linear:
        rgb = _srgb_byte_to_linear(rgb)
    px = (rgb[:, :, 0].astype(np.uint32) | (rgb[:, :, 1].astype(np.uint32) << 8) | (rgb[:, :, 2].astype(np.uint32) << 16)
          | (alpha.astype(np.uint32) << 24))
    return np.ascontiguousarray(px[::-1]), opaque  # stbi_set_flip_vertically_on_load(1)


def _transform2d(t):  # LoaderUtils::inlineTransformAs2d: the 2-D affine part of a transform
    if t is None:
        return None
    m = parse_transform(t)
    out = np.eye(3)
    out[:2, :2] = m[:2, :2]
    out[:2, 2] = m[:2, 3]
    return out


_FILTERS = {"bilinear": 1, "nearest": 0}  # anything else is bicubic (pattern/ImagePattern.cpp:19-31)
_WRAPS = {"mirror": 1, "clamp": 2}


def build_texture(sh, e, base):
    ty = e["type"]
    if ty == "checkerboard":  # pattern/CheckerBoardPattern.cpp:14-35
        return sh.add_checkerboard((e.get("scale_x", 2.0), e.get("scale_y", 2.0)), _color(e.get("color0"), 0.0),
                                   _color(e.get("color1"), 1.0), _transform2d(e.get("transform")))
    if ty in ("image", "bitmap"):  # pattern/ImagePattern.cpp
        path = os.path.join(base, e["filename"])
        if path.lower().endswith((".exr", ".hdr")):
            raise ValueError("float image textures are outside the supported subset")
        px, opaque = load_packed_image(path, e.get("linear", False))
        if "wrap_mode_u" in e:
            wrap = (_WRAPS.get(e.get("wrap_mode_u", "repeat"), 0), _WRAPS.get(e.get("wrap_mode_v", "repeat"), 0))
        else:
            wrap = (_WRAPS.get(e.get("wrap_mode", "repeat"), 0),) * 2
        return sh.add_image_texture(px, _FILTERS.get(e.get("filter_type", "bicubic"), 2), wrap, _transform2d(e.get("transform")), opaque)
    raise ValueError(f"texture type '{ty}' is outside the supported subset")


def compute_cdf2d(image, premultiply_sin=True, compensate=False):
    """CDF::computeForImage (src/runtime/CDF.cpp:42-133): image (h, w, >=3) float32 -> (marginal[h], conditional[h, w])."""
    F = np.float32
    img = np.asarray(image, F)
    h, w = img.shape[:2]
    defect = np.zeros(3, F)
    if compensate:
        defect = (np.maximum(img[:, :, :3], 0).reshape(-1, 3) / F(w)).sum(axis=0, dtype=F) / F(h)
    lum = (np.maximum(img[:, :, 0] - defect[0], 0) + np.maximum(img[:, :, 1] - defect[1], 0) + np.maximum(img[:, :, 2] - defect[2], 0)) / F(3)
    cond = np.cumsum(lum.astype(F), axis=1, dtype=F)
    sums = cond[:, -1].copy()
    marginal = sums * np.sin(F(math.pi) * (np.arange(h, dtype=F) + F(0.5)) / F(h), dtype=F) if premultiply_sin else sums.copy()
    for y in range(h):
        if sums[y] > 1e-5:
            cond[y] *= F(1.0) / sums[y]
        else:
            cond[y] = np.arange(w, dtype=F) * (F(1.0) / F(w - 1))
        cond[y, -1] = 1
    marginal = np.cumsum(marginal.astype(F), dtype=F)
    if marginal[-1] > 1e-5:
        marginal *= F(1.0) / marginal[-1]
    else:
        marginal = np.arange(h, dtype=F) * (F(1.0) / F(h - 1))
    marginal[-1] = 1
    return marginal.astype(F), cond.astype(F)


# ---------------------------------------------------------------------------------------- sky (loader side, N3)
def compute_sun_ea(year=2020, month=5, day=6, hour=12, minute=0, seconds=0.0, longitude=-6.9965744, latitude=49.235422, timezone=-2.0):
    """computeSunEA (src/runtime/skysun/SunLocation.cpp:17-103) with the defaults of SunLocation.h:8-36.
    Returns (elevation, azimuth west of south) in radians, float32 like ElevationAzimuth."""
    F = np.float32
    # double arithmetic like the reference (`Hour + Timezone + (Minute + Seconds / 60.0) / 60.0` promotes to double); numpy float32
    # SCALARS would keep every following product in float32 and round the Julian day count to 0.25 days
    dec_hours = float(hour) + float(F(timezone)) + (float(minute) + float(F(seconds)) / 60.0) / 60.0
    aux1 = int((month - 14) / 12)
    aux2 = int((1461 * (year + 4800 + aux1)) / 4) + int((367 * (month - 2 - 12 * aux1)) / 12) - int((3 * int((year + 4900 + aux1) / 100)) / 4) + day - 32075
    elapsed = (float(aux2) - 0.5 + dec_hours / 24.0) - 2451545.0
    omega = 2.1429 - 0.0010394594 * elapsed
    mean_long = 4.8950630 + 0.017202791698 * elapsed
    anomaly = 6.2400600 + 0.0172019699 * elapsed
    ecl_long = mean_long + 0.03341607 * math.sin(anomaly) + 0.00034894 * math.sin(2 * anomaly) - 0.0001134 - 0.0000203 * math.sin(omega)
    ecl_obl = 0.4090928 - 6.2140e-9 * elapsed + 0.0000396 * math.cos(omega)
    s_el = math.sin(ecl_long)
    ra = math.atan2(math.cos(ecl_obl) * s_el, math.cos(ecl_long))
    PI = float(F(3.14159265358979323846))
    if ra < 0:
        ra += 2 * PI
    decl = math.asin(math.sin(ecl_obl) * s_el)
    gmst = 6.6974243242 + 0.0657098283 * elapsed + dec_hours
    deg2rad = float(F(PI / 180))
    lmst = deg2rad * float(F(gmst * 15 - float(F(longitude))))
    lat = deg2rad * float(F(latitude))
    ha = lmst - ra
    zenith = math.acos(math.cos(lat) * math.cos(ha) * math.cos(decl) + math.sin(decl) * math.sin(lat))
    az = math.atan2(-math.sin(ha), math.tan(decl) * math.cos(lat) - math.sin(lat) * math.cos(ha))
    if az < 0:
        az += 2 * PI
    zenith += (6371.01 / 149597890) * math.sin(zenith)
    return F(F(PI / 2) - F(zenith)), F(math.fmod(float(F(az) + F(PI)), 2 * PI))


def ea_from_direction_yup(d):  # ElevationAzimuth::fromDirectionYUp
    d = np.asarray(d, np.float64)
    d = d / np.linalg.norm(d)
    phi = math.atan2(d[0], -d[2])
    return np.float32(math.pi / 2 - math.acos(d[1])), np.float32(phi + 2 * math.pi if phi < 0 else phi)


def light_ea(light):
    """LoaderUtils::getEA (loader/LoaderUtils.cpp:140-151): direction, sun_direction, elevation / azimuth, or time and place."""
    F = np.float32
    if "direction" in light or "sun_direction" in light:
        return ea_from_direction_yup(light.get("direction", light.get("sun_direction")))
    if "elevation" in light or "azimuth" in light:
        return F(light.get("elevation", 0)), F(light.get("azimuth", 0))
    return compute_sun_ea(*[light.get(k, d) for k, d in (("year", 2020), ("month", 5), ("day", 6), ("hour", 12), ("minute", 0),
                                                         ("seconds", 0.0), ("longitude", -6.9965744), ("latitude", 49.235422),
                                                         ("timezone", -2.0))])


def cie_sky_parameters(light, kind):
    """CIELight::serialize (runtime/light/CIELight.cpp:41-110): the arguments of make_cie_sky_light / make_cie_sunny_light."""
    F = np.float32
    el, az = light_ea(light)
    sun = np.array([math.cos(el) * math.sin(az), math.sin(el), -math.cos(el) * math.cos(az)], np.float32)  # toDirectionYUp
    out = dict(scale=_color(light.get("scale"), 1.0), zenith=_color(light.get("zenith"), 1.0), ground=_color(light.get("ground"), 1.0),
               ground_brightness=F(light.get("ground_brightness", 0.2)), has_ground=bool(light.get("has_ground", True)), sun_dir=sun)
    if kind in ("cieuniform", "ciecloudy"):
        return out
    PI = F(3.14159265358979323846)
    el = F(min(float(el), float(F(87) * PI / F(180))))  # "Sun too close to zenith"
    clear = kind == "cieclear"
    turbidity = F(light.get("turbidity", 2.45))
    zb = (F(1.376) * turbidity - F(1.81)) * F(math.tan(el)) + F(0.38)
    if not clear:
        zb = (zb + F(8.6) * sun[1] + F(0.123)) / F(2)
    zb = max(F(0), F(zb * F(1000) / F(203)))
    half_pi = PI / F(2)
    if clear:
        factor = F(0.274) * (F(0.91) + F(10) * F(math.exp(-3 * float(half_pi - el))) + F(0.45) * sun[1] * sun[1])
    else:
        factor = (F(2.739) + F(0.9891) * F(math.sin(0.3119 + 2.6 * float(el)))) * F(math.exp(-float(half_pi - el) * (0.4441 + 1.48 * float(el))))
    arr = (2.766521, 0.547665, -0.369832, 0.009237, 0.059229) if clear else (3.5556, -2.7152, -1.3081, 1.0660, 0.60227)
    x = (float(el) - float(PI) / 4) / (float(PI) / 4)  # skylight_normalization_factor
    f = arr[4]
    for i in range(3, -1, -1):
        f = f * x + arr[i]
    norm_factor = F(f) / PI / factor
    solar = F(1.5e9) / F(208) * (F(1.147) - F(0.147) / max(sun[1], F(0.16)))
    additive = F(6e-5) / PI * solar * sun[1] * (F(1) if clear else F(0.15))
    out.update(zenith_brightness=F(zb / factor), c2=F(zb * norm_factor + additive), clear=clear)
    return out


# ---- Perez all-weather sky: runtime/skysun/PerezModel.cpp (coefficients of Perez, Seals, Michalsky 1993) ----
_PEREZ_RANGES = (1.000, 1.065, 1.230, 1.500, 1.950, 2.800, 4.500, 6.200, float("inf"))
_PEREZ = {
    "a": (1.3525, -0.2576, -0.2690, -1.4366, -1.2219, -0.7730, 1.4148, 1.1016, -1.1000, -0.2515, 0.8952, 0.0156, -0.5484, -0.6654, -0.2672, 0.7117,
          -0.6000, -0.3566, -2.5000, 2.3250, -1.0156, -0.3670, 1.0078, 1.4051, -1.0000, 0.0211, 0.5025, -0.5119, -1.0500, 0.0289, 0.4260, 0.3590),
    "b": (-0.7670, 0.0007, 1.2734, -0.1233, -0.2054, 0.0367, -3.9128, 0.9156, 0.2782, -0.1812, -4.5000, 1.1766, 0.7234, -0.6219, -5.6812, 2.6297,
          0.2937, 0.0496, -5.6812, 1.8415, 0.2875, -0.5328, -3.8500, 3.3750, -0.3000, 0.1922, 0.7023, -1.6317, -0.3250, 0.1156, 0.7781, 0.0025),
    "c": (2.8000, 0.6004, 1.2375, 1.0000, 6.9750, 0.1774, 6.4477, -0.1239, 24.7219, -13.0812, -37.7000, 34.8438, 33.3389, -18.3000, -62.2500, 52.0781,
          21.0000, -4.7656, -21.5906, 7.2492, 14.0000, -0.9999, -7.1406, 7.5469, 19.0000, -5.0000, 1.2438, -1.9094, 31.0625, -14.5000, -46.1148, 55.3750),
    "d": (1.8734, 0.6297, 0.9738, 0.2809, -1.5798, -0.5081, -1.7812, 0.1080, -5.0000, 1.5218, 3.9229, -2.6204, -3.5000, 0.0016, 1.1477, 0.1062,
          -3.5000, -0.1554, 1.4062, 0.3988, -3.4000, -0.1078, -1.0750, 1.5702, -4.0000, 0.0250, 0.3844, 0.2656, -7.2312, 0.4050, 13.3500, 0.6234),
    "e": (0.0356, -0.1246, -0.5718, 0.9938, 0.2624, 0.0672, -0.2190, -0.4285, -0.0156, 0.1597, 0.4199, -0.5562, 0.4659, -0.3296, -0.0876, -0.0329,
          0.0032, 0.0766, -0.0656, -0.1294, -0.0672, 0.4016, 0.3017, -0.4844, 1.0468, -0.3788, -2.4517, 1.4656, 1.5000, -0.6426, 1.8564, 0.5636),
}
_SOLAR_E, _SOLAR_L = 1367.0, 127500.0  # skysun/Illuminance.h


def _perez_air_mass(z):  # computeAirMass, PerezModel.cpp:184-188
    zd = min(math.degrees(z), 90.0)
    return 1 / (math.cos(math.radians(zd)) + 0.15 * math.exp(math.log(93.885 - zd) * -1.253))


def _perez_eccentricity(day):  # computeEccentricity :190-194
    a = 2 * math.pi * min(max(day / 365.0, 0.0), 1.0)
    return 1.00011 + 0.034221 * math.cos(a) + 0.00128 * math.sin(a) + 0.000719 * math.cos(2 * a) + 0.000077 * math.sin(2 * a)


def perez_from_sky(brightness, clearness, z):  # PerezModel::fromSky :146-182
    if 1.065 < clearness < 2.8 and brightness < 0.2:
        brightness = 0.2
    brightness = min(max(brightness, 0.01), 0.6)
    bin_ = next((k for k in range(8) if _PEREZ_RANGES[k] <= clearness < _PEREZ_RANGES[k + 1]), 8)
    if bin_ >= 8:  # clearness below 1: the reference reads past its tables; refuse instead
        raise ValueError("perez: sky clearness outside [1, inf)")
    lin = lambda x: x[0] + x[1] * z + brightness * (x[2] + x[3] * z)
    row = lambda k: _PEREZ[k][4 * bin_:4 * bin_ + 4]
    sc, sd = _PEREZ["c"], _PEREZ["d"]
    c = lin(row("c")) if bin_ > 0 else math.exp(math.pow(brightness * (sc[0] + sc[1] * z), sc[2])) - sc[3]
    d = lin(row("d")) if bin_ > 0 else -math.exp(brightness * (sd[0] + sd[1] * z)) + sd[2] + brightness * sd[3]
    return lin(row("a")), lin(row("b")), c, d, lin(row("e"))


def perez_eval(p, cos_sun, cos_theta):  # PerezModel::eval :66-72
    a, b, c, d, e = p
    return (1 + a * math.exp(b / max(0.01, cos_theta))) * (1 + c * math.exp(d * math.acos(cos_sun)) + e * cos_sun * cos_sun)


def perez_integrate(p, z):  # PerezModel::integrate :74-118: Radiance's 145-patch base
    rows = ((84, 30), (72, 30), (60, 24), (48, 24), (36, 18), (24, 12), (12, 6), (0, 1))
    cs, ss, total = math.cos(z), math.sin(z), 0.0
    for theta, n in rows:
        ct, st = math.cos(math.radians(theta)), math.sin(math.radians(theta))
        for k in range(n):
            cp = math.cos(math.radians(360.0 / n * k))
            total += perez_eval(p, min(1.0, cs * ct + ss * st * cp), ct) * ct
    return 2 * math.pi * total / 145


def perez_sky_parameters(light):
    """PerezLight::serialize / getModel (runtime/light/PerezLight.cpp:25-110): sun direction, (a..e), luminance colour, ground."""
    import datetime
    el, az = light_ea(light)
    sun = np.array([math.cos(el) * math.sin(az), math.sin(el), -math.cos(el) * math.cos(az)], np.float32)
    z = math.acos(min(1.0, max(-1.0, float(sun[1]))))
    day = datetime.date(int(light.get("year", 2020)), int(light.get("month", 5)), int(light.get("day", 6))).timetuple().tm_yday - 1  # tm_yday
    if "clearness" in light or "brightness" in light:
        br = light.get("brightness", 0.2)
        model = perez_from_sky(br, light.get("clearness", 1.0), z)
        diff = br * _SOLAR_E * _perez_eccentricity(day) / _perez_air_mass(z)
    elif "direct_irradiance" in light or "diffuse_irradiance" in light or "direct_illuminance" in light or "diffuse_illuminance" in light:
        if "direct_irradiance" in light or "diffuse_irradiance" in light:
            dif, dirr = light.get("diffuse_irradiance", 1.0), light.get("direct_irradiance", 1.0)
        else:
            dif, dirr = (light.get(k, 1.0) * _SOLAR_E / _SOLAR_L for k in ("diffuse_illuminance", "direct_illuminance"))
        a3 = 1.041 * z ** 3
        model = perez_from_sky(dif * _perez_air_mass(z) / (_SOLAR_E * _perez_eccentricity(day)), ((dif + dirr) / dif + a3) / (1 + a3), z)
        diff = dif
    else:
        model, diff = tuple(light.get(k, 1.0) for k in "abcde"), 1.0
    diffnorm = diff / perez_integrate(model, z)
    if "luminance" in light:
        lum = _color(light.get("luminance"), 1.0) * np.float32(diffnorm)
    else:
        sin_el = min(1.0, max(-1.0, -float(sun[1])))
        lum = _color(light.get("zenith"), 1.0) * np.float32(perez_eval(model, sin_el, 1.0) * diffnorm)  # calc_perez(sin_elevation, 1, ...)
    return sun, model, lum, _color(light.get("ground"), 1.0), bool(light.get("has_ground", True))


def bake_sky(light, width=512, height=256):
    """SkyLight.cpp:33-50 + SkyModel::SkyModel (src/runtime/skysun/SkyModel.cpp:9-55) at RES_AZ x RES_EL = 512 x 256: the
    product's own Hosek-Wilkie bake (host/SkyModel.cpp in libig_runtime_b200.so), followed by the EXR round trip of the
    reference (float32 on disk, flipped vertically when loaded, Image.cpp:559-560, alpha 1).
    Returns the (height, width, 4) float32 image as the device sees it."""
    import ctypes as C
    from . import _lib
    host = _lib.load_host()
    el, az = light_ea(light)
    ground = np.ascontiguousarray(_color(light.get("ground"), 0.8), np.float32)
    H, Wd = height, width
    img = np.empty((H, Wd, 4), np.float32)
    rc = host.ig_b200_bake_sky(ground.ctypes.data_as(C.POINTER(C.c_float)), float(el), float(az), float(light.get("turbidity", 3.0)),
                               Wd, H, img.ctypes.data_as(C.POINTER(C.c_float)))
    if rc != 0:
        raise RuntimeError("ig_b200_bake_sky failed")
    return img


def bake_image_texture(px, opaque, filt, wrap, width, height):
    """bake_texture2d (src/artic/entrypoints/bake.art): the texture at uv = (x / (w - 1), y / (h - 1)); the loader
    bakes environment radiance at >= 1024 x 512 before building its CDF (EnvironmentLight.cpp:37-41,
    ShadingTree.h:78-85).  Nearest and bilinear lookups of packed RGBA8 images with identity transform only."""
    F = np.float32
    ih, iw = px.shape
    rgba = np.stack([(px >> s) & 255 for s in (0, 8, 16, 24)], axis=-1).astype(F) / F(255)
    if opaque:
        rgba[:, :, 3] = 1

    def border(mode, x, n):
        if mode == 2:
            return np.clip(x, 0, n - 1)
        if mode == 1:
            t = np.where(x < 0, -1 - x, x)
            i, k = t // n, t % n
            return np.where(i % 2 == 0, n - 1 - k, k)
        return np.mod(x, n)

    u = (np.arange(width, dtype=F) / F(width - 1))[None, :].repeat(height, 0)
    v = (np.arange(height, dtype=F) / F(height - 1))[:, None].repeat(width, 1)
    if filt == 0:
        ix = border(wrap[0], np.floor(u * F(iw)).astype(np.int64), iw)
        iy = border(wrap[1], np.floor(v * F(ih)).astype(np.int64), ih)
        return rgba[iy, ix]
    if filt == 1:
        fu, fv = u * F(iw) - F(0.5), v * F(ih) - F(0.5)
        x0, y0 = np.floor(fu).astype(np.int64), np.floor(fv).astype(np.int64)
        fx, fy = (fu - np.floor(fu))[..., None], (fv - np.floor(fv))[..., None]
        g = lambda xx, yy: rgba[border(wrap[1], yy, ih), border(wrap[0], xx, iw)]
        a = (1 - fx) * g(x0, y0) + fx * g(x0 + 1, y0)
        b = (1 - fx) * g(x0, y0 + 1) + fx * g(x0 + 1, y0 + 1)
        return ((1 - fy) * a + fy * b).astype(F)
    raise ValueError("baking bicubic textures is outside the supported subset")


_CONDUCTORS = {"none": ((0.0, 0.0, 0.0), (1.0, 1.0, 1.0))}  # runtime/bsdf/BSDF.cpp:12-42 (named metals: not restated)


# ---------------------------------------------------------------------------------------- light hierarchy
def build_light_hierarchy(entries):
    """LightHierarchy::setup (src/runtime/light/LightHierarchy.cpp:46-127) over PointBvh::store
    (src/runtime/container/PointBvh.inl:23-95).  entries: (position, direction or None, flux) per finite light, in id
    order.  Returns (codes u32[n], nodes float32[m, 8])."""
    F = np.float32
    inner, leaves = [], []  # inner: dict(index, lo, hi, mid, axis)

    def descend(i, p):
        while True:
            n = inner[i]
            n["lo"], n["hi"] = np.minimum(n["lo"], p), np.maximum(n["hi"], p)
            if n["axis"] < 0:
                return i
            mid = ((n["hi"] + n["lo"]) / F(2))[n["axis"]]
            i = n["index"] if p[n["axis"]] < mid else n["index"] + 1

    for k, (pos, direction, flux) in enumerate(entries):
        p = np.asarray(pos, F)
        has_dir = direction is not None
        leaves.append((p, np.asarray(direction if has_dir else (0, 0, 1), F), F(flux) if has_dir else -F(flux), k))
        if not inner:
            inner.append(dict(index=0, lo=p.copy(), hi=p.copy(), mid=F(0), axis=-1))
            continue
        ni = descend(0, p)
        node = inner[ni]
        old_leaf, lo, hi = node["index"], node["lo"].copy(), node["hi"].copy()
        diam = hi - lo
        axis = int(np.argmax(diam))
        mid = diam[axis] / F(2)  # (sic) half the extent, not a position
        left_idx = len(inner)
        node.update(index=left_idx, axis=axis, mid=mid)
        off = (hi[axis] - lo[axis]) * F(0.5)  # BoundingBox::computeSplit
        lhi, rlo = hi.copy(), lo.copy()
        lhi[axis] -= off
        rlo[axis] += off
        new_leaf = len(leaves) - 1
        li, ri = (new_leaf, old_leaf) if p[axis] < mid else (old_leaf, new_leaf)
        inner.append(dict(index=li, lo=lo.copy(), hi=lhi, mid=F(0), axis=-1))
        inner.append(dict(index=ri, lo=rlo, hi=hi.copy(), mid=F(0), axis=-1))

    nodes = np.zeros((len(inner), 8), F)
    ids = np.zeros(len(inner), np.int32)
    codes = np.zeros(len(entries), np.uint32)

    def populate(i, code, depth):  # populateInnerNodes
        n = inner[i]
        if n["axis"] < 0:
            p, d, f, lid = leaves[n["index"]]
            codes[lid] = code
            nodes[i, 0:3], nodes[i, 3], nodes[i, 4:7] = p, f, d
            ids[i] = lid
            return d, f
        ld, lf = populate(n["index"], code, depth + 1)
        rd, rf = populate(n["index"] + 1, code | (1 << depth), depth + 1)
        nodes[i, 0:3] = (n["hi"] + n["lo"]) / F(2)
        ids[i] = -(n["index"] + 1)
        if lf < 0 and rf < 0:
            d, f = np.array([0, 0, 1], F), lf + rf
        elif lf < 0:
            d, f = np.array([0, 0, 1], F), -(-lf + rf)
        elif rf < 0:
            d, f = np.array([0, 0, 1], F), -(lf - rf)
        else:
            d = ld + rd
            d, f = (d / np.sqrt(np.dot(d, d), dtype=F)).astype(F), lf + rf
        nodes[i, 3], nodes[i, 4:7] = f, d
        return d, f

    populate(0, 0, 0)
    nodes[:, 7] = ids.view(F)
    return codes, nodes


def load_scene(scene, base=".", width=None, height=None):
    """scene: dict or path. Returns (SceneTables, Shading, width, height, info)."""
    if isinstance(scene, str):
        base = os.path.dirname(os.path.abspath(scene))
        scene = load_json(scene)
    scene = _merge_externals(dict(scene), base)
    film = scene.get("film", {}).get("size", [1000, 1000])
    width, height = width or int(film[0]), height or int(film[1])

    b = SceneBuilder()
    sh = Shading()

    shapes, shape_mesh, shape_sphere = {}, {}, {}
    for e in scene.get("shapes", []):
        built = build_shape(e, base)
        if built[0] == "sphere":
            shapes[e["name"]] = b.add_sphere(built[1], built[2])
            shape_sphere[e["name"]] = (built[1], built[2])
        else:
            m = built[1]
            shapes[e["name"]] = b.add_trimesh(m.v, m.f, m.n, m.t)
            shape_mesh[e["name"]] = m

    tex_ids = {e["name"]: build_texture(sh, e, base) for e in scene.get("textures", [])}

    bsdf_ids = {}
    pending = list(scene.get("bsdfs", []))
    while pending:  # bump maps refer to other BSDFs by name, in any order
        e = pending.pop(0)
        ty = e["type"]
        if ty == "bumpmap" and e.get("bsdf") not in bsdf_ids:
            if not any(o["name"] == e.get("bsdf") for o in pending):
                raise ValueError(f"bumpmap '{e['name']}' refers to an unknown bsdf")
            pending.append(e)
            continue
        if ty == "bumpmap":  # runtime/bsdf/MapBSDF.cpp
            bsdf_ids[e["name"]] = sh.add_bumpmap(bsdf_ids[e["bsdf"]], tex_ids[e["map"]], e.get("strength", 1.0))
        elif ty == "diffuse" and isinstance(e.get("reflectance"), str):
            bsdf_ids[e["name"]] = sh.add_diffuse((0, 0, 0), texture=tex_ids[e["reflectance"]])
        elif ty == "diffuse":
            bsdf_ids[e["name"]] = sh.add_diffuse(_color(e.get("reflectance"), 0.8))
        elif ty in ("dielectric", "glass"):
            bsdf_ids[e["name"]] = sh.add_dielectric(e.get("int_ior", 1.5046), e.get("ext_ior", 1.0),
                                                    _color(e.get("specular_reflectance"), 1.0),
                                                    _color(e.get("specular_transmittance"), 1.0), bool(e.get("thin", False)))
        elif ty == "mirror":
            bsdf_ids[e["name"]] = sh.add_mirror(_color(e.get("specular_reflectance"), 1.0))
        elif ty == "conductor":  # runtime/bsdf/ConductorBSDF.cpp:13-36, BSDF.cpp:53-99
            eta0, k0 = _CONDUCTORS[str(e.get("material", "none")).lower()]
            eta, k = _color(e.get("eta"), eta0), _color(e.get("k"), k0)
            ks = _color(e.get("specular_reflectance"), 1.0)
            old = any(n in e for n in ("alpha", "alpha_u", "alpha_v"))
            par = "alpha" if old else "roughness"
            if not any(n in e for n in (par, par + "_u", par + "_v")):
                # delta distribution: mirror when eta ~ 0 and k ~ 1 (make_conductor_bsdf, bsdf/conductor.art:112-122)
                if np.all(np.abs(eta) <= 1e-4) and np.all(np.abs(k - 1) <= 1e-4):
                    bsdf_ids[e["name"]] = sh.add_mirror(ks)
                else:
                    bsdf_ids[e["name"]] = sh.add_conductor(eta, k, ks)
            elif (par + "_u") in e or (par + "_v") in e:
                bsdf_ids[e["name"]] = sh.add_conductor(eta, k, ks, alpha=(e.get(par + "_u", 0.1), e.get(par + "_v", 0.1)))
            else:
                bsdf_ids[e["name"]] = sh.add_conductor(eta, k, ks, roughness=e.get(par, 0.1), anisotropic=e.get("anisotropic", 0.0))
        else:
            raise ValueError(f"bsdf type '{ty}' is outside the supported subset")

    lights = scene.get("lights", [])
    area_entities = {l["entity"] for l in lights if l["type"] == "area"}

    # material groups in first-appearance order; area-light entities get their own (LoaderEntity.cpp:87-103)
    groups, keys = [], []
    for e in scene.get("entities", []):
        key = ("area", e["name"]) if e["name"] in area_entities else ("bsdf", e["bsdf"])
        if key in keys and key[0] == "bsdf":
            groups[keys.index(key)].append(e)
        else:
            keys.append(key)
            groups.append([e])
    entity_ids, entity_xform, mat_of_entity = {}, {}, {}
    for mat_id, group in enumerate(groups):
        for e in group:
            m = parse_transform(e.get("transform"))
            flags = (1 if e.get("camera_visible", True) else 0) | (2 if e.get("light_visible", True) else 0) | \
                    (4 if e.get("bounce_visible", True) else 0) | (8 if e.get("shadow_visible", True) else 0)
            entity_ids[e["name"]] = b.add_entity(shapes[e["shape"]], m, material=mat_id, flags=flags)
            entity_xform[e["name"]] = m
            mat_of_entity[e["name"]] = mat_id
    tables = b.finish()

    # lights: finite ids are needed by the emissive materials
    area_light_of_entity = {}
    hier = []  # (position, direction or None, flux) per finite light: Light::position/direction/computeFlux
    SR = np.float32(4 * math.pi)
    for l in lights:
        ty = l["type"]
        if ty.startswith("cie_"):  # both spellings are registered (LoaderLight.cpp:90-97)
            ty = "cie" + ty[4:]
        if ty == "point":
            c = _color(l.get("power"), 4 * math.pi) / SR if "power" in l else _color(l.get("intensity"), 1.0)
            sh.add_point_light(l.get("position", (0, 0, 0)), c)
            hier.append((l.get("position", (0, 0, 0)), None, float(np.mean(c * SR))))  # PointLight.cpp:17-31
        elif ty == "spot":
            cutoff, falloff = l.get("cutoff", 30.0), l.get("falloff", 20.0)
            c = _color(l.get("intensity"), 1.0)
            if "power" in l:  # spot_from_power, light/spot.art:1-6
                f = 2 * math.pi * (1 - 0.5 * math.cos(math.radians(falloff)) - 0.5 * math.cos(math.radians(cutoff)))
                c = _color(l.get("power"), 1.0) / np.float32(f)
            sh.add_spot_light(l.get("position", (0, 0, 0)), l.get("direction", (0, 0, 1)), cutoff, falloff, c)
            f = 2 * math.pi * (1 - 0.5 * math.cos(math.radians(falloff)) - 0.5 * math.cos(math.radians(cutoff)))
            d = np.asarray(l.get("direction", (0, 0, 1)), np.float64)
            hier.append((l.get("position", (0, 0, 0)), d / np.linalg.norm(d), float(np.mean(c) * f)))
        elif ty == "sun":  # SunLight.cpp:8-46: direction through LoaderUtils::getDirection (EA round trip), angle 11.4 degrees unless `radius`
            el, az = light_ea(l)
            d = (math.cos(el) * math.sin(az), math.sin(el), -math.cos(el) * math.cos(az))
            sh.add_sun_light(d, _color(l.get("irradiance"), 1.0), l.get("angle", 11.4), l.get("radius"))
        elif ty in ("directional", "direction", "distant"):
            sh.add_directional_light(l.get("direction", (0, 0, 1)), _color(l.get("irradiance"), 1.0))
        elif ty in ("env", "envmap", "uniform", "constant"):
            if isinstance(l.get("radiance"), str) and l["radiance"] in tex_ids:  # EnvironmentLight.cpp:27-76
                tid = tex_ids[l["radiance"]]
                t = sh.textures[tid]
                te = next(e for e in scene["textures"] if e["name"] == l["radiance"])
                cdf = None
                if l.get("cdf", True) and t.type == 1 and t.format == 0 and "transform" not in te:
                    px = np.frombuffer(bytes(sh.blob[t.offset:t.offset + 4 * t.width * t.height]), np.uint32).reshape(t.height, t.width)
                    baked = bake_image_texture(px, bool(t.opaque), t.filter, (t.wrap_u, t.wrap_v), max(1024, t.width), max(512, t.height))
                    cdf = sh.add_cdf2d(*compute_cdf2d(baked, True, l.get("compensate", True)))
                m3 = parse_transform(l.get("transform"))[:3, :3]
                sh.add_env_textured(tid, _color(l.get("scale"), 1.0), cdf, np.linalg.inv(m3.T))
            elif isinstance(l.get("radiance"), str):
                c = _color(l["radiance"], 1.0)  # constant expression
                sh.add_env_light(c, _color(l.get("scale"), 1.0))
            else:
                sh.add_env_light(_color(l.get("radiance"), 1.0), _color(l.get("scale"), 1.0))
        elif ty == "sky":  # runtime/light/SkyLight.cpp:52-77
            img = bake_sky(l)
            tex = sh.add_image_texture(img, 1, (0, 0))  # make_repeat_border(), make_bilinear_filter()
            marginal, cond = compute_cdf2d(img, True, False)
            m3 = parse_transform(l.get("transform"))[:3, :3]
            sh.add_env_textured(tex, _color(l.get("scale"), 1.0), sh.add_cdf2d(marginal, cond), np.linalg.inv(m3.T))
        elif ty in ("cieuniform", "ciecloudy", "cieclear", "cieintermediate"):  # LoaderLight.cpp registry -> CIELight
            q = cie_sky_parameters(l, ty)
            m3 = parse_transform(l.get("transform"))[:3, :3]
            sunny = ty in ("cieclear", "cieintermediate")
            sh.add_sky_function(1 if sunny else 0, q.get("clear", False) if sunny else ty == "ciecloudy", q["has_ground"], q["scale"], q["zenith"],
                                q["ground"], q.get("zenith_brightness", 1.0), q["ground_brightness"], q.get("c2", 1.0), q["sun_dir"],
                                np.linalg.inv(m3.T))
        elif ty == "perez":
            sun, model, lum, ground, has_ground = perez_sky_parameters(l)
            m3 = parse_transform(l.get("transform"))[:3, :3]
            sh.add_perez_sky(sun, lum, ground, model, has_ground, np.linalg.inv(m3.T))
        elif ty == "area":
            name = l["entity"]
            ent = next(e for e in scene["entities"] if e["name"] == name)
            mesh = shape_mesh.get(ent["shape"])
            plane = None
            if mesh is not None and l.get("optimize", True):
                pl = as_plane(mesh.v, mesh.f)
                if pl is not None:
                    m = entity_xform[name]
                    o = (m[:3, :3] @ pl[0].astype(np.float64) + m[:3, 3])
                    plane = (o, m[:3, :3] @ pl[1].astype(np.float64), m[:3, :3] @ pl[2].astype(np.float64))
            m = entity_xform[name]
            sphere = shape_sphere.get(ent["shape"])
            if sphere is not None and l.get("optimize", True):  # AreaLight.cpp:54-55,73-82: analytic sphere sampler
                area_light_of_entity[name] = sh.add_sphere_area_light(entity_ids[name], _color(l.get("radiance"), 1.0), sphere[0], sphere[1])
                w, h_, d = (np.linalg.norm(m[:3, :3] @ (np.eye(3)[k] * sphere[1])) for k in range(3))
                P = 1.6075  # approximate_ellipsoid_area, AreaLight.cpp:26-35
                area = 4 * math.pi * (((w * h_) ** P + (w * d) ** P + (h_ * d) ** P) / 3) ** (1 / P)
                hier.append((m[:3, :3] @ np.asarray(sphere[0], np.float64) + m[:3, 3], None, float(np.mean(_color(l.get("radiance"), 1.0))) * area * math.pi))
                continue
            area_light_of_entity[name] = sh.add_area_light(entity_ids[name], _color(l.get("radiance"), 1.0), plane)
            if plane is not None:  # AreaLight.cpp:58-71
                n = np.cross(plane[1], plane[2])
                area = float(np.linalg.norm(n))
                hier.append((plane[0] + plane[1] * 0.5 + plane[2] * 0.5, n / area, float(np.mean(_color(l.get("radiance"), 1.0))) * area * math.pi))
            elif mesh is not None:  # :84-92 (world-space area instead of approximate_area_scale: selection weights only)
                wv = mesh.v.astype(np.float64) @ m[:3, :3].T + m[:3, 3]
                area = float(np.linalg.norm(np.cross(wv[mesh.f[:, 1]] - wv[mesh.f[:, 0]], wv[mesh.f[:, 2]] - wv[mesh.f[:, 0]]), axis=1).sum() / 2)
                c0 = (mesh.v.min(axis=0) + mesh.v.max(axis=0)).astype(np.float64) / 2
                hier.append((m[:3, :3] @ c0 + m[:3, 3], None, float(np.mean(_color(l.get("radiance"), 1.0))) * area * math.pi))
            else:
                raise ValueError("sphere area lights are outside the supported subset")
        else:
            raise ValueError(f"light type '{ty}' is outside the supported subset")

    for key in keys:
        if key[0] == "area":
            ent = next(e for e in scene["entities"] if e["name"] == key[1])
            sh.add_material(bsdf_ids[ent["bsdf"]], area_light_of_entity[key[1]])
        else:
            sh.add_material(bsdf_ids[key[1]])

    cam = scene.get("camera", {})
    vertical = "vfov" in cam
    fov = cam.get("vfov", cam.get("hfov", cam.get("fov", 60.0)))
    if cam.get("type", "perspective") in ("fishlens", "fisheye"):  # Camera.cpp registry; FishLensCamera.cpp: mode is always Circular
        sh.set_fishlens_from_transform(parse_transform(cam.get("transform")), width, height,
                                       cam.get("near_clip", 0.0), cam.get("far_clip", float(np.finfo(np.float32).max)))
    else:
        sh.set_perspective_from_transform(parse_transform(cam.get("transform")), fov, width, height,
                                          cam.get("near_clip", 0.0), cam.get("far_clip", float(np.finfo(np.float32).max)), vertical)
    tech = scene.get("technique", {"type": "path"})
    if tech.get("type", "path") == "path":
        sh.set_path(tech.get("max_depth", 64), tech.get("min_depth", 2), tech.get("clamp", 0.0), tech.get("nee", True),
                    aov_mis=bool(tech.get("aov_mis", False)))  # PathTechnique.cpp:16
        # LoaderLight::generateLightSelector (loader/LoaderLight.cpp:423-452)
        selector = tech.get("light_selector", "uniform")
        if selector == "hierarchy" and len(sh.infinite) + len(sh.finite) > 1 and sh.finite:
            codes, nodes = build_light_hierarchy(hier)
            sh.set_hierarchy(codes, nodes)
        elif selector == "simple" and len(sh.infinite) + len(sh.finite) > 1 and sh.finite:
            sh.set_light_cdf([h[2] for h in hier])  # LoaderLight::generateLightSelectionCDF: computeFlux per finite light
        elif selector not in ("uniform", "hierarchy", "simple", ""):
            raise ValueError(f"light selector '{selector}' is outside the supported subset")
    elif tech["type"] == "debug":
        sh.set_debug(tech.get("mode", 0))
    elif tech["type"] == "infobuffer":  # InfoBufferTechnique.cpp:9-12
        sh.set_infobuffer(tech.get("max_depth", 64), bool(tech.get("handle_specular", False)))
    else:
        raise ValueError(f"technique '{tech['type']}' is outside the supported subset")
    return tables, sh, width, height, {"entities": entity_ids, "materials": keys}
