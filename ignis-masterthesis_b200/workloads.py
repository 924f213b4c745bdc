"""Deterministic synthetic inputs of the BASELINE.json configs (definitions: SURVEY.md 8(d)).

C1: the reference's default `triangle` shape (src/runtime/shape/TriMeshProvider.cpp:19-25: p0=(0,0,0),
    p1=(1,0,0), p2=(0,1,0)) under an identity entity + N rays, splitmix64 seed 0x1615.
C3: 64 instances (4x4x4 grid, random rotations, splitmix64 seed 0x3D) of one 16 128-triangle UV sphere
    (64 stacks x 128 slices -- the same tessellation counts as scenes/meshes/SphereTri.ply, generated here
    because /root/reference does not exist on the GPU box) = 1 032 192 triangles; N uniform-direction rays
    from inside the scene box scaled x1.25, splitmix64 seed 0xB200, generation order (incoherent).
"""
import numpy as np

from .tables import RAY_DTYPE, SceneBuilder

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(seed, start, count):
    """Values start..start+count-1 of the splitmix64 stream with the given seed (vectorised)."""
    with np.errstate(over="ignore"):
        idx = np.arange(start + 1, start + count + 1, dtype=np.uint64)
        z = np.uint64(seed) + idx * np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return z


def uniform01(seed, start, count):
    """float32 in [0,1) with 24 random bits."""
    return ((splitmix64(seed, start, count) >> np.uint64(40)).astype(np.float32)) * np.float32(2.0 ** -24)


def uv_sphere(stacks=64, slices=128, radius=1.0):
    """Closed UV sphere: 2 poles + (stacks-1) rings of `slices` vertices; 2*slices*(stacks-1) triangles."""
    th = (np.arange(1, stacks, dtype=np.float64) / stacks) * np.pi
    ph = (np.arange(slices, dtype=np.float64) / slices) * 2 * np.pi
    ring = np.stack([np.outer(np.sin(th), np.cos(ph)), np.outer(np.sin(th), np.sin(ph)),
                     np.outer(np.cos(th), np.ones_like(ph))], axis=-1).reshape(-1, 3)
    verts = np.concatenate([[[0, 0, 1.0]], ring, [[0, 0, -1.0]]]) * radius
    faces = []
    top, bottom = 0, verts.shape[0] - 1
    rid = lambda r, s: 1 + r * slices + (s % slices)
    for s in range(slices):
        faces.append((top, rid(0, s), rid(0, s + 1)))
    for r in range(stacks - 2):
        for s in range(slices):
            a, b, c, d = rid(r, s), rid(r, s + 1), rid(r + 1, s), rid(r + 1, s + 1)
            faces.append((a, c, d))
            faces.append((a, d, b))
    for s in range(slices):
        faces.append((bottom, rid(stacks - 2, s + 1), rid(stacks - 2, s)))
    v = verts.astype(np.float32)
    n = v / np.maximum(np.linalg.norm(v, axis=1, keepdims=True), 1e-30)
    return v, np.asarray(faces, dtype=np.uint32), n.astype(np.float32)


def quat_to_mat(q):
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]], dtype=np.float64)


# ---- C1 -------------------------------------------------------------------------------------------
def c1_scene():
    b = SceneBuilder()
    shape = b.add_trimesh([[0, 0, 0], [1, 0, 0], [0, 1, 0]], [[0, 1, 2]])
    b.add_entity(shape)
    return b.finish()


def c1_rays(n=1_000_000, seed=0x1615):
    u = uniform01(seed, 0, 4 * n).reshape(n, 4)
    org = np.stack([u[:, 0] * 2 - 0.5, u[:, 1] * 2 - 0.5, np.ones(n, np.float32)], axis=1).astype(np.float32)
    tgt = np.stack([u[:, 2] * 2 - 0.5, u[:, 3] * 2 - 0.5, np.zeros(n, np.float32)], axis=1).astype(np.float32)
    d = tgt - org
    d = d / np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    rays = np.zeros(n, RAY_DTYPE)
    rays["org"] = org
    rays["dir"] = d.astype(np.float32)
    rays["tmin"] = 0.0
    rays["tmax"] = np.finfo(np.float32).max
    return rays


# ---- C3 -------------------------------------------------------------------------------------------
def c3_scene(grid=4, stacks=64, slices=128, seed=0x3D, flat=False):
    """Instanced (default) or flattened (one world-space mesh) 1M-triangle scene."""
    v, f, nrm = uv_sphere(stacks, slices)
    diag = float(np.linalg.norm(v.max(axis=0) - v.min(axis=0)))
    spacing = 2.5 * diag / np.sqrt(3.0)
    u = uniform01(seed, 0, 3 * grid ** 3).reshape(-1, 3).astype(np.float64)
    b = SceneBuilder()
    xforms = []
    k = 0
    for iz in range(grid):
        for iy in range(grid):
            for ix in range(grid):
                u1, u2, u3 = u[k]
                q = (np.sqrt(1 - u1) * np.sin(2 * np.pi * u2), np.sqrt(1 - u1) * np.cos(2 * np.pi * u2),
                     np.sqrt(u1) * np.sin(2 * np.pi * u3), np.sqrt(u1) * np.cos(2 * np.pi * u3))
                m = np.eye(4)
                m[:3, :3] = quat_to_mat(q)
                m[:3, 3] = (np.array([ix, iy, iz]) - (grid - 1) / 2.0) * spacing
                xforms.append(m)
                k += 1
    if flat:
        vs, fs = [], []
        for i, m in enumerate(xforms):
            vs.append((v.astype(np.float64) @ m[:3, :3].T + m[:3, 3]).astype(np.float32))
            fs.append(f + np.uint32(i * v.shape[0]))
        shape = b.add_trimesh(np.concatenate(vs), np.concatenate(fs))
        b.add_entity(shape)
    else:
        shape = b.add_trimesh(v, f, nrm)
        for m in xforms:
            b.add_entity(shape, m)
    return b.finish()


def c3_rays(tables, n=16_777_216, seed=0xB200, chunk=1 << 20):
    lo, hi = tables.bbox[:3].astype(np.float64), tables.bbox[3:].astype(np.float64)
    c, h = (lo + hi) / 2, (hi - lo) / 2 * 1.25
    rays = np.zeros(n, RAY_DTYPE)
    for start in range(0, n, chunk):
        m = min(chunk, n - start)
        u = uniform01(seed, 6 * start, 6 * m).reshape(m, 6)
        org = (c + (u[:, :3].astype(np.float64) * 2 - 1) * h).astype(np.float32)
        z = (1 - 2 * u[:, 3].astype(np.float64))
        r = np.sqrt(np.maximum(0.0, 1 - z * z))
        phi = 2 * np.pi * u[:, 4].astype(np.float64)
        d = np.stack([r * np.cos(phi), r * np.sin(phi), z], axis=1).astype(np.float32)
        sl = rays[start:start + m]
        sl["org"] = org
        sl["dir"] = d
    rays["tmin"] = 1e-4
    rays["tmax"] = np.finfo(np.float32).max
    return rays


def c3_dense_rays(tables, n=16_777_216, seed=0xDE5E, grid=4, chunk=1 << 20):
    """The high-hit-rate companion of c3_rays on the same scene: origins uniform in the scene box
    scaled x1.25 as before, but every ray is aimed at a point inside a randomly chosen sphere (centre + 0.9 r * uniform
    ball), so it hits that sphere or an occluder in front of it: hit rate ~1, long paths through the grid, rays still
    in generation order (incoherent: consecutive rays share neither origin nor target)."""
    lo, hi = tables.bbox[:3].astype(np.float64), tables.bbox[3:].astype(np.float64)
    c, h = (lo + hi) / 2, (hi - lo) / 2 * 1.25
    diag = 2.0 * np.sqrt(3.0)                       # bbox diagonal of the unit sphere (uv_sphere, radius 1)
    spacing = 2.5 * diag / np.sqrt(3.0)
    rays = np.zeros(n, RAY_DTYPE)
    for start in range(0, n, chunk):
        m = min(chunk, n - start)
        u = uniform01(seed, 8 * start, 8 * m).reshape(m, 8).astype(np.float64)
        org = c + (u[:, :3] * 2 - 1) * h
        cell = np.minimum((u[:, 3:6] * grid).astype(np.int64), grid - 1)
        centre = (cell - (grid - 1) / 2.0) * spacing
        # uniform point in the ball of radius 0.9 from three of the remaining numbers (direction from 6, 7; radius from the
        # fractional part of u3 * grid, which is independent of the cell index)
        z = 1 - 2 * u[:, 6]
        r = np.sqrt(np.maximum(0.0, 1 - z * z))
        phi = 2 * np.pi * u[:, 7]
        rad = 0.9 * np.cbrt(u[:, 3] * grid - cell[:, 0])
        tgt = centre + rad[:, None] * np.stack([r * np.cos(phi), r * np.sin(phi), z], axis=1)
        d = tgt - org
        d /= np.maximum(np.linalg.norm(d, axis=1, keepdims=True), 1e-30)
        sl = rays[start:start + m]
        sl["org"] = org.astype(np.float32)
        sl["dir"] = d.astype(np.float32)
    rays["tmin"] = 1e-4
    rays["tmax"] = np.finfo(np.float32).max
    return rays
