"""C++ host side above the C ABI (ignis-masterthesis_b200/host: IG::Device / IG::Runtime mirror, igtrace, igcli):
file formats and error behaviour that need no device."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def _run(*cmd, **kw):
    return subprocess.run(list(cmd), capture_output=True, text=True, timeout=120, **kw)


@pytest.mark.parametrize("name", ["cbox-d6", "primitives", "many-point-lights"])
def test_igdb_round_trip_python_writer_cpp_reader(ig, host_bin, tmp_path, name):
    fx = ig.fixtures
    tables, sh, w, h, _ = fx.load_fixture(os.path.join(GOLDEN, name + ".npz"))
    path = str(tmp_path / (name + ".igdb"))
    fx.save_igdb(path, tables, sh, w, h)
    r = _run(os.path.join(host_bin, "igdbtool"), "inspect", path)
    assert r.returncode == 0, r.stderr
    seen = dict(line.split(" ", 1) for line in r.stdout.strip().splitlines())
    assert int(seen["shapes"]) == tables.lookups.size and int(seen["shape_bytes"]) == tables.shape_data.size
    assert int(seen["entities"]) == tables.entity_count
    assert int(seen["trimesh_leaves"]) == tables.leaves[0].size and int(seen["sphere_leaves"]) == tables.leaves[1].size
    assert int(seen["materials"]) == tables.material_count == len(sh.materials)
    assert int(seen["bsdfs"]) == len(sh.bsdfs) and int(seen["textures"]) == len(sh.textures)
    assert int(seen["infinite_lights"]) == len(sh.infinite) and int(seen["finite_lights"]) == len(sh.finite)
    assert int(seen["blob_bytes"]) == len(sh.blob) and int(seen["selector"]) == sh.selector
    assert int(seen["max_depth"]) == sh.technique.max_depth and (int(seen["width"]), int(seen["height"])) == (w, h)
    np.testing.assert_allclose([float(x) for x in seen["bbox"].split()], tables.bbox, rtol=1e-5)


def test_igdb_reader_rejects_garbage_and_truncation(ig, host_bin, tmp_path):
    fx = ig.fixtures
    tables, sh, w, h, _ = fx.load_fixture(os.path.join(GOLDEN, "cbox-d1.npz"))
    good = str(tmp_path / "good.igdb")
    fx.save_igdb(good, tables, sh, w, h)
    raw = open(good, "rb").read()
    for label, data in (("garbage", b"not an igdb file at all"), ("truncated", raw[: len(raw) // 2]), ("empty", b"")):
        bad = str(tmp_path / (label + ".igdb"))
        open(bad, "wb").write(data)
        r = _run(os.path.join(host_bin, "igdbtool"), "inspect", bad)
        assert r.returncode != 0 and "[ERROR]" in r.stderr, label
    assert _run(os.path.join(host_bin, "igdbtool"), "inspect", str(tmp_path / "missing.igdb")).returncode != 0
    # a table of contents that points outside the file (a multi-GB `bytes` field) is refused before any allocation
    import struct
    evil = bytearray(raw)
    evil[16 + 24:16 + 32] = struct.pack("<Q", 1 << 40)
    bad = str(tmp_path / "evil.igdb")
    open(bad, "wb").write(bytes(evil))
    r = _run(os.path.join(host_bin, "igdbtool"), "inspect", bad)
    assert r.returncode != 0 and "outside the file" in r.stderr
    # a file without its shading sections is an error, not a scene with empty tables
    n_sec = struct.unpack("<I", raw[12:16])[0]
    cut = bytearray(raw)
    for i in range(n_sec):
        e = 16 + 40 * i
        if cut[e:e + 24].rstrip(b"\0") == b"bsdfs":
            cut[e:e + 5] = b"xsdfs"
    nob = str(tmp_path / "nobsdfs.igdb")
    open(nob, "wb").write(bytes(cut))
    r = _run(os.path.join(host_bin, "igdbtool"), "inspect", nob)
    assert r.returncode != 0 and "bsdfs" in r.stderr


def test_exr_and_pfm_writers(host_bin, tmp_path):
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    import cv2
    w, h = 37, 11
    y, x = np.mgrid[0:h, 0:w].astype(np.float32)
    expect = np.stack([x, y, x * y + 0.5], axis=2)
    for ext in ("exr", "pfm"):
        path = str(tmp_path / ("img." + ext))
        assert _run(os.path.join(host_bin, "igdbtool"), "testimage", path, str(w), str(h)).returncode == 0
        img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
        assert img is not None and img.dtype == np.float32 and img.shape == (h, w, 3), ext
        assert np.array_equal(img[:, :, ::-1], expect), ext  # OpenCV hands back BGR; row 0 = top row


def test_frontends_fail_loudly_without_a_device(ig, host_bin, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    fx = ig.fixtures
    tables, sh, w, h, _ = fx.load_fixture(os.path.join(GOLDEN, "cbox-d1.npz"))
    scene = str(tmp_path / "s.igdb")
    fx.save_igdb(scene, tables, sh, w, h)
    rays = tmp_path / "rays.txt"
    rays.write_text("0 0 0 0 0 1\n")
    for cmd in ([os.path.join(host_bin, "igtrace"), scene, "-i", str(rays)],
                [os.path.join(host_bin, "igcli"), scene, "--spp", "1", "-o", str(tmp_path / "o.exr")]):
        r = _run(*cmd)
        assert r.returncode != 0 and "no CPU fallback" in r.stderr  # IG_LOG(L_FATAL) + abort, Device.cpp:271-274
    assert not (tmp_path / "o.exr").exists()


def test_frontend_argument_errors(host_bin, tmp_path):
    igtrace, igcli = os.path.join(host_bin, "igtrace"), os.path.join(host_bin, "igcli")
    assert _run(igtrace, "--help").returncode == 0
    r = _run(igtrace)
    assert r.returncode != 0 and "No input file given" in r.stderr
    empty = tmp_path / "empty.txt"
    empty.write_text("\n")
    r = _run(igtrace, "whatever.igdb", "-i", str(empty))
    assert r.returncode != 0 and "No rays given" in r.stderr  # trace/main.cpp:105-109
    r = _run(igcli, "whatever.igdb", "-o", "x.exr")
    assert r.returncode != 0 and "No valid spp count" in r.stderr  # cli/main.cpp:74-77
    r = _run(igcli, "whatever.igdb", "--spp", "4")
    assert r.returncode != 0 and "No output file given" in r.stderr
