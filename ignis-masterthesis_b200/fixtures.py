"""Golden fixtures: scene tables + shading tables (+ an optional reference image) in one .npz file.

The reference tree does not exist on the GPU box, so scenes taken from it (scenes/evaluation/*.json,
scenes/diamond_scene.json) travel as the exact table blobs `sceneio.load_scene` produced here, together with
the reference renderer's image where the reference ships one (scenes/evaluation/references/*.exr).
Written by scripts/make_golden.py; read by tests/ and bench.py.
"""
import ctypes as C

import numpy as np

from . import shading as S
from .tables import LEAF_DTYPE, LOOKUP_DTYPE, SceneTables


def _bytes(obj):
    return np.frombuffer(bytes(obj), np.uint8).copy()


def _table(cls, items):
    return np.concatenate([_bytes(x) for x in items]) if items else np.zeros(0, np.uint8)


def save_fixture(path, tables, sh, width, height, ref=None, meta=""):
    np.savez_compressed(
        path, lookups=tables.lookups.view(np.uint8), shape_data=tables.shape_data, entities=tables.entities,
        leaves0=tables.leaves[0].view(np.uint8), leaves1=tables.leaves[1].view(np.uint8), bbox=tables.bbox,
        material_count=np.int64(tables.material_count), entity_per_material=np.asarray(tables.entity_per_material, np.int64),
        camera=_bytes(sh.camera), technique=_bytes(sh.technique), bsdfs=_table(S.Bsdf, sh.bsdfs),
        materials=_table(S.Material, sh.materials), infinite=_table(S.Light, sh.infinite), finite=_table(S.Light, sh.finite),
        textures=_table(S.Texture, sh.textures), blob=np.frombuffer(bytes(sh.blob), np.uint8).copy(),
        selector=np.asarray([sh.selector, sh.hierarchy], np.int64),
        size=np.asarray([width, height], np.int64), ref=np.zeros(0, np.float32) if ref is None else np.asarray(ref, np.float32),
        meta=np.asarray(meta))


def save_igdb(path, tables, sh, width, height):
    """The same tables as one `.igdb` file for the C++ frontends (layout: host/SceneFile.h)."""
    sections = [
        ("lookups", tables.lookups.view(np.uint8)), ("shape_data", tables.shape_data), ("entities", tables.entities.view(np.uint8)),
        ("leaves0", tables.leaves[0].view(np.uint8)), ("leaves1", tables.leaves[1].view(np.uint8)),
        ("bbox", np.asarray(tables.bbox, "<f4").view(np.uint8)), ("material_count", np.asarray([tables.material_count], "<u8").view(np.uint8)),
        ("entity_per_material", np.asarray(tables.entity_per_material, "<i4").view(np.uint8)),
        ("camera", _bytes(sh.camera)), ("technique", _bytes(sh.technique)), ("bsdfs", _table(S.Bsdf, sh.bsdfs)),
        ("materials", _table(S.Material, sh.materials)), ("infinite", _table(S.Light, sh.infinite)),
        ("finite", _table(S.Light, sh.finite)), ("textures", _table(S.Texture, sh.textures)),
        ("blob", np.frombuffer(bytes(sh.blob), np.uint8)), ("selector", np.asarray([sh.selector, sh.hierarchy], "<u4").view(np.uint8)),
        ("film", np.asarray([width, height], "<u4").view(np.uint8)),
    ]
    align = lambda x: (x + 15) // 16 * 16
    offset = align(16 + 40 * len(sections))
    toc, payload = bytearray(), bytearray()
    for name, data in sections:
        raw = np.ascontiguousarray(data).reshape(-1).tobytes()
        toc += name.encode().ljust(24, b"\0") + np.asarray([len(raw), offset + len(payload)], "<u8").tobytes()
        payload += raw + b"\0" * (align(len(raw)) - len(raw))
    head = b"IGB200DB" + np.asarray([1, len(sections)], "<u4").tobytes() + bytes(toc)
    with open(path, "wb") as f:
        f.write(head.ljust(offset, b"\0") + bytes(payload))


def _untable(cls, raw):
    n = raw.size // C.sizeof(cls)
    return [cls.from_buffer_copy(raw[i * C.sizeof(cls):(i + 1) * C.sizeof(cls)].tobytes()) for i in range(n)]


def load_fixture(path):
    """Returns (SceneTables, Shading, width, height, ref image or None)."""
    z = np.load(path)
    tables = SceneTables(z["lookups"].view(LOOKUP_DTYPE), z["shape_data"], z["entities"],
                         [z["leaves0"].view(LEAF_DTYPE), z["leaves1"].view(LEAF_DTYPE)], z["bbox"],
                         int(z["material_count"]), list(z["entity_per_material"]))
    sh = S.Shading()
    cam = z["camera"].tobytes()  # fixtures written before ABI v5 hold the 52-byte camera without `type` (= perspective)
    sh.camera = S.Camera.from_buffer_copy(cam.ljust(C.sizeof(S.Camera), b"\0"))
    sh.technique = S.Technique.from_buffer_copy(z["technique"].tobytes())
    sh.bsdfs = _untable(S.Bsdf, z["bsdfs"])
    sh.materials = _untable(S.Material, z["materials"])
    sh.infinite = _untable(S.Light, z["infinite"])
    sh.finite = _untable(S.Light, z["finite"])
    if "textures" in z.files:  # ABI v3 additions
        sh.textures = _untable(S.Texture, z["textures"])
        sh.blob = bytearray(z["blob"].tobytes())
        sh.selector, sh.hierarchy = [int(x) for x in z["selector"]]
    w, h = [int(x) for x in z["size"]]
    ref = z["ref"]
    return tables, sh, w, h, (ref if ref.size else None)
