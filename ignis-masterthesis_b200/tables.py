"""Scene tables in the reference's `SceneDatabase` layout (the input contract of the backend).

Mirrors what the reference's loader produces, byte for byte:
  * DynTable "shapes"   -- src/runtime/shape/TriMeshProvider.cpp:574-602 (trimesh),
                           src/runtime/shape/SphereProvider.cpp:33-45 (sphere);
                           DynTable alignment/lookup rules src/runtime/table/DynTable.h:15-24
  * FixTable "entities" -- src/runtime/loader/LoaderEntity.cpp:150-176 (36 floats per entity)
  * EntityLeaf1 records -- src/runtime/bvh/SceneBVHAdapter.h:73-105, entity flags LoaderEntity.cpp:122-131,
                           entity bounding box = shape box transformed corner-wise (math/BoundingBox.h:84-95)
Entities are numbered in material-group order like the reference (LoaderEntity.cpp:99-178): callers add
entities grouped by material if they want the reference's `entity_per_material` invariant.
"""
import ctypes as C

import numpy as np

from . import _lib

DEFAULT_ALIGNMENT = 16  # IG::DefaultAlignment
PRIM_TRIMESH = 0
PRIM_SPHERE = 1
VISIBLE_ALL = 0xF

LEAF_DTYPE = np.dtype([
    ("bbox_min", "<f4", 3), ("entity_id", "<i4"),
    ("bbox_max", "<f4", 3), ("shape_id", "<i4"),
    ("local", "<f4", 12),
    ("flags", "<u4"), ("mat_id", "<i4"), ("user1", "<i4"), ("user2", "<i4"),
])
LOOKUP_DTYPE = np.dtype([("type_id", "<u4"), ("flags", "<u4"), ("offset", "<u8")])
RAY_DTYPE = np.dtype([("org", "<f4", 3), ("dir", "<f4", 3), ("tmin", "<f4"), ("tmax", "<f4")])
assert LEAF_DTYPE.itemsize == 96 and LOOKUP_DTYPE.itemsize == 16 and RAY_DTYPE.itemsize == 32


class SceneTables:
    """Immutable bundle of the table blobs + a ctypes `igb200_scene` view over them."""

    def __init__(self, lookups, shape_data, entities, leaves, bbox, material_count, entity_per_material):
        self.lookups = np.ascontiguousarray(lookups, dtype=LOOKUP_DTYPE)
        self.shape_data = np.ascontiguousarray(shape_data, dtype=np.uint8)
        self.entities = np.ascontiguousarray(entities, dtype=np.float32)
        self.leaves = [np.ascontiguousarray(l, dtype=LEAF_DTYPE) for l in leaves]
        self.bbox = np.asarray(bbox, dtype=np.float32)
        self.material_count = int(material_count)
        self.entity_per_material = list(entity_per_material)

    @property
    def entity_count(self):
        return self.entities.size // 36

    def c_scene(self):
        sc = _lib.Scene()
        sc.shape_lookups = self.lookups.ctypes.data if self.lookups.size else None
        sc.shape_count = self.lookups.size
        sc.shape_data = self.shape_data.ctypes.data if self.shape_data.size else None
        sc.shape_bytes = self.shape_data.size
        sc.entities = self.entities.ctypes.data if self.entities.size else None
        sc.entity_count = self.entity_count
        for k in range(2):
            sc.leaves[k] = self.leaves[k].ctypes.data if self.leaves[k].size else None
            sc.leaf_count[k] = self.leaves[k].size
        for k in range(6):
            sc.scene_bbox[k] = float(self.bbox[k])
        sc.material_count = self.material_count
        return sc


class SceneBuilder:
    def __init__(self):
        self._data = bytearray()
        self._lookups = []
        self._shape_bbox = []
        self._entities = []
        self._leaves = [[], []]
        self._materials = {}

    # -- DynTable::addLookup (table/DynTable.h:15-24) ------------------------------------------
    def _add_lookup(self, type_id):
        if len(self._data) != 0:
            # the reference pads by `alignment - size % alignment` even when already aligned
            self._data.extend(b"\0" * (DEFAULT_ALIGNMENT - len(self._data) % DEFAULT_ALIGNMENT))
        self._lookups.append((type_id, 0, len(self._data)))

    def add_trimesh(self, vertices, faces, normals=None, texcoords=None):
        """vertices (nv,3) f32, faces (nf,3) int. Returns the shape id."""
        v = np.ascontiguousarray(vertices, dtype=np.float32).reshape(-1, 3)
        f = np.ascontiguousarray(faces, dtype=np.uint32).reshape(-1, 3)
        if v.shape[0] == 0 or f.shape[0] == 0:
            raise ValueError("mesh needs vertices and faces")  # reference logs an error and skips the shape
        if normals is None:
            # TriMesh::computeFaceNormals fallback is per-vertex; geometry-only users do not read them
            normals = np.zeros_like(v)
            normals[:, 2] = 1.0
        n = np.ascontiguousarray(normals, dtype=np.float32).reshape(-1, 3)
        t = np.zeros((v.shape[0], 2), np.float32) if texcoords is None else \
            np.ascontiguousarray(texcoords, dtype=np.float32).reshape(-1, 2)
        lo = v.min(axis=0) - np.float32(1e-5)  # bbox.inflate(1e-5f), TriMeshProvider.cpp:546-547
        hi = v.max(axis=0) + np.float32(1e-5)
        self._add_lookup(PRIM_TRIMESH)
        hdr = np.array([f.shape[0], v.shape[0], n.shape[0], t.shape[0]], dtype=np.uint32)
        bbox = np.array([lo[0], lo[1], lo[2], 0, hi[0], hi[1], hi[2], 0], dtype=np.float32)
        v4 = np.zeros((v.shape[0], 4), np.float32); v4[:, :3] = v
        n4 = np.zeros((n.shape[0], 4), np.float32); n4[:, :3] = n
        f4 = np.zeros((f.shape[0], 4), np.uint32); f4[:, :3] = f
        for part in (hdr, bbox, v4, n4, f4, t):
            self._data.extend(part.tobytes())
        self._shape_bbox.append((lo.astype(np.float32), hi.astype(np.float32)))
        return len(self._lookups) - 1

    def add_sphere(self, center=(0, 0, 0), radius=1.0):
        if radius <= 0:
            raise ValueError("sphere radius must be positive")  # SphereProvider.cpp:23-26
        c = np.asarray(center, dtype=np.float32)
        r = np.float32(radius)
        self._add_lookup(PRIM_SPHERE)
        self._data.extend(np.array([c[0], c[1], c[2], r], dtype=np.float32).tobytes())
        self._shape_bbox.append((c - r - np.float32(1e-5), c + r + np.float32(1e-5)))
        return len(self._lookups) - 1

    def add_entity(self, shape_id, transform=None, material=0, flags=VISIBLE_ALL):
        """transform: 4x4 (or 3x4) object->world affine matrix. Returns the entity id."""
        m = np.eye(4, dtype=np.float64)
        if transform is not None:
            tr = np.asarray(transform, dtype=np.float64)
            m[:tr.shape[0], :] = tr
        m[3] = (0, 0, 0, 1)  # Transformf::makeAffine
        inv = np.linalg.inv(m)
        to_local = inv[:3, :].astype(np.float32)
        to_global = m[:3, :].astype(np.float32)
        normal = np.linalg.inv(m[:3, :3]).T.astype(np.float32)
        mat_id = self._materials.setdefault(material, len(self._materials))
        ent_id = len(self._entities)
        row = np.zeros(36, np.float32)
        row[0:12] = to_local.T.reshape(-1)    # column major (Serializer::write(m, true))
        row[12:24] = to_global.T.reshape(-1)
        row[24:33] = normal.T.reshape(-1)
        row[33:36] = np.array([shape_id, mat_id, 0], dtype=np.uint32).view(np.float32)
        self._entities.append(row)

        lo, hi = self._shape_bbox[shape_id]
        corners = np.array([[(hi if (i >> k) & 1 else lo)[k] for k in range(3)] for i in range(8)], dtype=np.float32)
        g = to_global.astype(np.float32)
        world = corners @ g[:, :3].T + g[:, 3]
        leaf = np.zeros((), dtype=LEAF_DTYPE)
        leaf["bbox_min"] = world.min(axis=0)
        leaf["bbox_max"] = world.max(axis=0)
        leaf["entity_id"] = ent_id
        leaf["shape_id"] = shape_id
        leaf["local"] = to_local.T.reshape(-1)
        leaf["flags"] = flags
        leaf["mat_id"] = mat_id
        self._leaves[self._lookups[shape_id][0]].append(leaf)
        return ent_id

    def finish(self):
        leaves = [np.array(l, dtype=LEAF_DTYPE) if l else np.zeros(0, LEAF_DTYPE) for l in self._leaves]
        all_l = np.concatenate(leaves) if (leaves[0].size + leaves[1].size) else np.zeros(0, LEAF_DTYPE)
        if all_l.size:
            bbox = np.concatenate([all_l["bbox_min"].min(axis=0), all_l["bbox_max"].max(axis=0)])
        else:
            bbox = np.zeros(6, np.float32)
        ents = np.concatenate(self._entities) if self._entities else np.zeros(0, np.float32)
        per_mat = [0] * len(self._materials)
        for l in all_l:
            per_mat[int(l["mat_id"])] += 1
        return SceneTables(np.array(self._lookups, dtype=LOOKUP_DTYPE), np.frombuffer(bytes(self._data), np.uint8),
                           ents, leaves, bbox, len(self._materials), per_mat)


def make_rays(org, direction, tmin=0.0, tmax=np.finfo(np.float32).max):
    """Pack arrays into StreamRay records (src/artic/traversal/ray.art:2-7)."""
    org = np.asarray(org, np.float32).reshape(-1, 3)
    rays = np.zeros(org.shape[0], RAY_DTYPE)
    rays["org"] = org
    rays["dir"] = np.asarray(direction, np.float32).reshape(-1, 3)
    rays["tmin"] = tmin
    rays["tmax"] = tmax
    return rays
