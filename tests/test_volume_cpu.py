"""On-disk formats in front of the heterogeneous-medium path, no GPU: the OpenVDB reader
(host/vdb.cpp: Blosc + LZ4, active-mask compression, half floats), inflate, the `volume` /
`heterog` plugins' flatten hook."""
import json
import os
import re
import struct
import subprocess
import zlib

import ctypes as C
import numpy as np
import pytest

from conftest import golden

HERE = os.path.dirname(os.path.abspath(__file__))
VDB = "scenes/vol/fluid_data_0050.vdb"      # the reference's scenes/project/volume/fluid_data_0050.vdb, verbatim


def file_metadata(path, grid):
    """The writer's own statistics, parsed independently of the reader under test: OpenVDB stores
    file_bbox_min/max and file_voxel_count in each grid's metadata."""
    d = open(path, "rb").read()
    at = d.index(b"\x04\x00\x00\x00name\x06\x00\x00\x00string" + struct.pack("<I", len(grid)) + grid.encode())
    seg = d[:at]
    def find_val(key, typ, fmt):
        i = seg.rindex(struct.pack("<I", len(key)) + key + struct.pack("<I", len(typ)) + typ)
        i += 8 + len(key) + len(typ)
        n = struct.unpack_from("<I", seg, i)[0]
        assert n == struct.calcsize(fmt)
        return struct.unpack_from(fmt, seg, i + 4)
    return {"bbox_min": find_val(b"file_bbox_min", b"vec3i", "<3i"), "bbox_max": find_val(b"file_bbox_max", b"vec3i", "<3i"),
            "voxels": find_val(b"file_voxel_count", b"int64", "<q")[0]}


def test_vdb_grids_listed(nb, built_lib):
    assert nb.vdb_list(golden(VDB)) == [("velocity", "Tree_vec3s_5_4_3_HalfFloat"), ("density", "Tree_float_5_4_3_HalfFloat"),
                                       ("shadow", "Tree_float_5_4_3_HalfFloat"), ("temperature", "Tree_float_5_4_3_HalfFloat")]


@pytest.mark.parametrize("grid", ["density", "temperature", "shadow"])
def test_vdb_reader_agrees_with_the_files_own_statistics(nb, built_lib, grid):
    g = nb.vdb_read(golden(VDB), grid)
    meta = file_metadata(golden(VDB), grid)
    assert tuple(g["bbox_min"]) == meta["bbox_min"] and tuple(g["bbox_max"]) == meta["bbox_max"]
    assert g["active_voxels"] == meta["voxels"] == g["values"].size
    np.testing.assert_allclose(g["index_to_world"], np.hstack([np.eye(3) * 0.1875, np.zeros((3, 1))]))
    v = g["values"]
    assert np.isfinite(v).all() and v.min() == g["min"] >= 0.0 and v.max() == g["max"] <= 1.0
    # half-float payload: every value is exactly representable in fp16
    assert np.array_equal(v.astype(np.float16).astype(np.float32), v)
    # a smoke plume, not noise: neighbouring voxels are close
    assert np.abs(np.diff(v, axis=2)).mean() < 0.1 * (v.max() - v.min())


def test_vdb_density_regression(nb, built_lib):
    g = nb.vdb_read(golden(VDB), "density")
    assert np.array_equal(g["values"], np.load(golden("vdb_density_21x21x32.npy")))
    with pytest.raises(nb.NbError):
        nb.vdb_read(golden(VDB), "velocity")          # vec3 grid
    with pytest.raises(nb.NbError):
        nb.vdb_read(golden(VDB), "nope")


@pytest.mark.parametrize("level", [0, 1, 6, 9])
def test_inflate_round_trips(nb, built_lib, level):
    h = nb.load_host_library()
    rng = np.random.default_rng(level)
    for payload in (b"", b"a", bytes(rng.integers(0, 256, 70000, dtype=np.uint8)), b"abcabcabc" * 5000,
                    np.repeat(rng.integers(0, 4, 3000, dtype=np.uint8), 7).tobytes()):
        z = zlib.compress(payload, level)
        out = np.zeros(max(len(payload), 1), np.uint8)
        n = h.nbh_inflate(z, len(z), out.ctypes.data_as(C.c_void_p), out.size)
        assert n == len(payload), h.nbh_last_error()
        assert out[:n].tobytes() == payload
    bad = bytearray(zlib.compress(b"hello world" * 100)); bad[10] ^= 0xff
    assert h.nbh_inflate(bytes(bad[:20]), 20, out.ctypes.data_as(C.c_void_p), out.size) == -1     # truncated: error, no crash


def test_heterogeneous_medium_record(nb, built_lib):
    """`heterog` + `volume` (heterogmedium.cpp:49-83, NvdbVolume.cpp:15-18) -> nb_medium + density array."""
    s = nb.load_xml(golden("scenes/vol/cbox_heterog.xml"))            # the reference's scene, unchanged, .vdb input
    m = s["media"][0]
    assert m["type"] == nb.MEDIUM_HETEROGENEOUS and m["phase"] == nb.PHASE_ISOTROPIC
    assert np.allclose(m["mu_t"], 2.0) and m["density_scale"] == 1.0             # sigma_a = 2, sigma_s default 0
    assert list(m["grid_dims"]) == [25, 25, 36] and m["first_voxel"] == 0        # 21x21x32 active box + 2 voxels of border
    dens = s["density"].reshape(25, 25, 36)
    assert np.array_equal(dens[2:-2, 2:-2, 2:-2], np.load(golden("vdb_density_21x21x32.npy")))
    border = dens.copy(); border[2:-2, 2:-2, 2:-2] = 0
    assert not border.any()
    assert m["max_density"] == dens.max()
    w2g = m["world_to_grid"].reshape(3, 4)
    np.testing.assert_allclose(w2g, np.hstack([np.eye(3) / 0.1875, np.full((3, 1), 2.0)]), rtol=1e-6)
    assert s["shapes"][0]["medium"] == 0 and s["shapes"][0]["bsdf"] == -1
    # .nbv input with a permuting affine map (grid z up the scene's y)
    f = nb.load_xml(golden("scenes/vol/cbox_heterog_fluid_small.xml"))
    w2g = f["media"][0]["world_to_grid"].reshape(3, 4)
    p = w2g @ np.array([0.102, 0.03, -0.028, 1.0])
    np.testing.assert_allclose(p, [2, 2, 2], atol=1e-5)                          # the grid's index origin lands on the border offset
    p = w2g @ np.array([0.102 + 21 / 32, 0.03 + 32 / 32, -0.028 + 21 / 32, 1.0])
    np.testing.assert_allclose(p, [23, 23, 34], atol=1e-4)
    assert np.allclose(f["media"][0]["mu_t"], 5.0) and f["media"][0]["density_scale"] == 2.0


# ---- .nvdb: the NanoVDB cache file the reference reads and writes (NvdbVolume.vdb.cpp:9-61, :120-133) ----------------------
NVDB = "scenes/vol/fluid_data_0050.nvdb"    # written by the reference's own NanoVDB 28.1.2 (tests/golden/make_nvdb.py)
NVDB_TOOL = os.path.join(os.path.dirname(HERE), "oracle", "_ref", "nvdb_tool")


def test_nvdb_reader_on_a_file_written_by_nanovdb(nb, built_lib):
    g, v = nb.vdb_read(golden(NVDB), "density"), nb.vdb_read(golden(VDB), "density")
    assert np.array_equal(g["values"], np.load(golden("vdb_density_21x21x32.npy")))
    assert np.array_equal(g["bbox_min"], v["bbox_min"]) and np.array_equal(g["bbox_max"], v["bbox_max"])
    assert np.array_equal(g["index_to_world"], v["index_to_world"])
    assert g["active_voxels"] == 21 * 21 * 32 and g["background"] == 0.0 and g["max"] == v["max"] and g["min"] == v["min"]
    assert np.array_equal(nb.vdb_read(golden(NVDB), "")["values"], g["values"])     # readGrid(file, 0): the first grid
    with pytest.raises(nb.NbError, match="nope"):
        nb.vdb_read(golden(NVDB), "nope")


def test_nvdb_reader_rejects_damaged_files(nb, built_lib, tmp_path):
    data = open(golden(NVDB), "rb").read()
    for name, bad, what in (("short.nvdb", data[:100000], "truncated"), ("magic.nvdb", b"NanoVDB1" + data[8:], "magic"),
                            ("version.nvdb", data[:8] + b"\x1d\x00" + data[10:], "version"), ("empty.nvdb", b"", "no float grid")):
        p = tmp_path / name
        p.write_bytes(bad)
        with pytest.raises(nb.NbError, match=what):
            nb.vdb_read(str(p), "density")


def test_nvdb_writer_makes_the_file_nanovdb_makes(nb, built_lib, tmp_path):
    """host/nvdb.cpp's writer against nanovdb::io::writeGrid on the same grid: the same bytes except the checksum (not
    computed: ~0, which NanoVDB reads as "no checksum") and the average / standard deviation of the upper node and the
    root, where this writer stores the exact moments of the active voxels (NanoVDB 28.1.2's own differ in the 2nd digit
    from the mean of the values it stores; nothing on the path reads them)."""
    out = str(tmp_path / "mine.nvdb")
    nb.to_nvdb(golden(VDB), out, "density")
    mine, ref = open(out, "rb").read(), open(golden(NVDB), "rb").read()
    assert len(mine) == len(ref) == 235000
    grid = 16 + 160 + len("density") + 1
    root, upper = grid + 672 + 64, grid + 672 + 160
    free = set(range(grid + 8, grid + 16)) | set(range(root + 48, root + 56)) | set(range(upper + 8224 + 8, upper + 8224 + 16))
    diff = [i for i in np.nonzero(np.frombuffer(mine, np.uint8) != np.frombuffer(ref, np.uint8))[0] if i not in free]
    assert diff == []
    v = np.load(golden("vdb_density_21x21x32.npy")).astype(np.float64)
    avg, dev = np.frombuffer(mine[root + 48:root + 56], "<f4")
    assert abs(avg - v.mean()) < 1e-6 and abs(dev - v.std()) < 1e-6
    back = nb.vdb_read(out, "density")                                               # and through this side's reader
    assert np.array_equal(back["values"], v.astype(np.float32)) and back["active_voxels"] == v.size


@pytest.mark.parametrize("case", ["offset", "negative", "two_uppers", "single"])
def test_nvdb_writer_round_trips(nb, built_lib, tmp_path, case):
    """Boxes that start off a leaf boundary, below zero, across two 4096^3 root tiles, one voxel: written, read back
    here, and — where oracle/_ref/nvdb_tool exists (the build container) — read back by the reference's NanoVDB."""
    rng = np.random.default_rng(7)
    org, dims = {"offset": ((3, 9, 130), (13, 5, 20)), "negative": ((-11, -3, -140), (17, 9, 21)), "two_uppers": ((4090, 0, -5), (12, 3, 9)),
                 "single": ((5, 6, 7), (1, 1, 1))}[case]
    vals = (rng.random(dims, dtype=np.float32) + 0.25).astype("<f4")
    M = np.array([[0.5, 0, 0, 1.0], [0, 0, 0.25, -2.0], [0, -0.125, 0, 3.0]])
    nbv = tmp_path / "g.nbv"
    nbv.write_bytes(b"NBV1" + np.array(dims, "<i4").tobytes() + np.array(org, "<i4").tobytes() + M.astype("<f8").tobytes() +
                    np.float32(0.125).tobytes() + vals.tobytes())
    out = str(tmp_path / "g.nvdb")
    nb.to_nvdb(str(nbv), out, "smoke")
    g = nb.vdb_read(out, "smoke")
    assert g["active_voxels"] == vals.size and g["background"] == np.float32(0.125)
    assert np.array_equal(g["values"], vals) and list(g["bbox_min"]) == list(org)
    assert np.array_equal(g["index_to_world"], M) and g["min"] == vals.min() and g["max"] == vals.max()
    if not os.path.exists(NVDB_TOOL):
        return
    dump = tmp_path / "dump.bin"
    r = subprocess.run([NVDB_TOOL, "dump", out, str(dump), "smoke"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    info = json.loads(r.stdout)
    assert info["name"] == "smoke" and info["active"] == vals.size
    raw = dump.read_bytes()
    mn, dd = np.frombuffer(raw, "<i4", 3, 0), np.frombuffer(raw, "<i4", 3, 12)
    assert list(mn) == [o - 2 for o in org] and list(dd) == [d + 4 for d in dims]
    np.testing.assert_allclose(np.frombuffer(raw, "<f8", 12, 24).reshape(3, 4), M, atol=1e-12)
    bg, lo, hi = np.frombuffer(raw, "<f4", 3, 120)
    assert bg == np.float32(0.125) and lo == vals.min() and hi == vals.max()
    seen = np.frombuffer(raw, "<f4", int(np.prod(dd)), 140).reshape(dd)
    want = np.full(dd, np.float32(0.125)); want[2:-2, 2:-2, 2:-2] = vals
    assert np.array_equal(seen, want)


def test_scene_naming_a_nvdb_file(nb, built_lib):
    """<string name="filename" value="….nvdb"/> (NvdbVolume.vdb.cpp:36-37) gives the medium the .vdb gives."""
    a, b = nb.load_xml(golden("scenes/vol/cbox_heterog_nvdb.xml")), nb.load_xml(golden("scenes/vol/cbox_heterog.xml"))
    assert np.array_equal(a["density"], b["density"])
    for key in ("grid_dims", "world_to_grid", "mu_t", "max_density", "density_scale", "first_voxel"):
        assert np.array_equal(np.asarray(a["media"][0][key]), np.asarray(b["media"][0][key])), key


# ---- sparse grids: leaves kept as they are on the host, 10^3 bricks behind a table on the device (include/norib200.h) --------
def expand_bricks(density, m):
    """The virtual dense array a NB_MEDIUM_HETEROGENEOUS_BRICKS record describes, and the stencil property of every brick."""
    nx, ny, nz = (int(v) for v in m["grid_dims"])
    nb = [(v + 7) // 8 for v in (nx, ny, nz)]
    first = int(m["first_voxel"])
    entries = nb[0] * nb[1] * nb[2]
    table = density[first:first + entries].view("<u4").reshape(nb)
    bricks = density[first + entries:].reshape(-1, 10, 10, 10)[: int(table.max()) + 1]
    full = np.zeros([8 * b + 2 for b in nb], np.float32)              # array coordinates -1 .. 8 nb
    for (i, j, k), b in np.ndenumerate(table):
        full[8 * i + 1:8 * i + 9, 8 * j + 1:8 * j + 9, 8 * k + 1:8 * k + 9] = bricks[b][1:9, 1:9, 1:9]
    for (i, j, k), b in np.ndenumerate(table):                        # aprons repeat the neighbours' voxels
        assert np.array_equal(full[8 * i:8 * i + 10, 8 * j:8 * j + 10, 8 * k:8 * k + 10], bricks[b]), (i, j, k)
    return full[1:nx + 1, 1:ny + 1, 1:nz + 1], table, bricks


def test_brick_layout_of_the_fluid_equals_its_dense_array(nb, built_lib, monkeypatch):
    dense = nb.load_xml(golden("scenes/vol/cbox_heterog.xml"))
    monkeypatch.setenv("NB_VOLUME_LAYOUT", "bricks")
    s = nb.load_xml(golden("scenes/vol/cbox_heterog.xml"))
    m, d = s["media"][0], dense["media"][0]
    assert m["type"] == nb.MEDIUM_HETEROGENEOUS_BRICKS and d["type"] == nb.MEDIUM_HETEROGENEOUS
    for key in ("grid_dims", "world_to_grid", "mu_t", "max_density", "density_scale"):
        assert np.array_equal(np.asarray(m[key]), np.asarray(d[key])), key
    full, table, bricks = expand_bricks(s["density"], m)
    assert np.array_equal(full, dense["density"].reshape(25, 25, 36))
    assert not bricks[0].any() and table.shape == (4, 4, 5)


def test_sparse_nvdb_stays_sparse_on_the_way_to_the_device(nb, built_lib, monkeypatch):
    """sparse_smoke.nvdb (tests/golden/make_nvdb.py): 2 puffs in a 96 x 48 x 120 box. The host keeps the file's leaves, the
    device record is the brick layout, and both describe the voxels a dense read of the same file gives."""
    g = nb.vdb_read(golden("scenes/vol/sparse_smoke.nvdb"), "density")            # dense read: box of ACTIVE voxels
    assert g["active_voxels"] == np.count_nonzero(g["values"]) and 0 < g["active_voxels"] < g["values"].size // 8
    s = nb.load_xml(golden("scenes/vol/cbox_sparse_small.xml"))
    m = s["media"][0]
    assert m["type"] == nb.MEDIUM_HETEROGENEOUS_BRICKS and m["max_density"] == g["max"]
    assert list(m["grid_dims"]) == [d + 4 for d in g["values"].shape]
    full, table, bricks = expand_bricks(s["density"], m)
    assert np.array_equal(full[2:-2, 2:-2, 2:-2], g["values"]) and not (full[:2].any() or full[-2:].any() or full[:, :2].any() or full[:, :, -2:].any())
    used = len(bricks) - 1
    assert used == len(np.unique(table)) - 1 and used * 1000 + table.size < full.size // 2     # the point of the layout
    monkeypatch.setenv("NB_VOLUME_LAYOUT", "dense")                                # the same file, forced dense: same voxels
    d = nb.load_xml(golden("scenes/vol/cbox_sparse_small.xml"))
    assert d["media"][0]["type"] == nb.MEDIUM_HETEROGENEOUS
    assert np.array_equal(d["density"].reshape(full.shape), full)
    assert np.array_equal(d["media"][0]["world_to_grid"], m["world_to_grid"])


def test_sparse_nvdb_round_trip_keeps_leaves_and_masks(nb, built_lib, tmp_path, monkeypatch):
    """read sparse -> write -> the same bytes (leaf order, masks, statistics), whatever form the reader chose"""
    src = golden("scenes/vol/sparse_smoke.nvdb")
    want = open(src, "rb").read()
    nb.to_nvdb(src, str(tmp_path / "dense_form.nvdb"), "density", sparse=True)    # via the dense box, re-sparsified
    assert open(tmp_path / "dense_form.nvdb", "rb").read() == want
    if os.path.exists(NVDB_TOOL):
        r = subprocess.run([NVDB_TOOL, "dump", src, str(tmp_path / "dump.bin")], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        g = nb.vdb_read(src, "density")
        assert json.loads(r.stdout)["active"] == g["active_voxels"]
        raw = (tmp_path / "dump.bin").read_bytes()
        dd = np.frombuffer(raw, "<i4", 3, 12)
        seen = np.frombuffer(raw, "<f4", int(np.prod(dd)), 140).reshape(dd)
        assert np.array_equal(seen[2:-2, 2:-2, 2:-2], g["values"])


def test_a_1024_cubed_box_of_sparse_smoke_is_never_filled(nb, built_lib, tmp_path):
    """Three 8^3 blocks at the corners and the middle of a 1024^3 box: 4.3 GB dense; here the leaves stay leaves on the host,
    the device record is a 2 M-entry table plus the few bricks those blocks touch, and NanoVDB reads the file."""
    import shutil
    rng = np.random.default_rng(3)
    org = np.array([[0, 0, 0], [504, 512, 520], [1016, 1016, 1016]], "<i4")
    vals = (rng.random((3, 8, 8, 8), dtype=np.float32) + 0.5).astype("<f4")
    M = np.array([[0.7 / 1024, 0, 0, 0.08], [0, 1.04 / 1024, 0, 0.02], [0, 0, 0.7 / 1024, -0.05]])
    for name in ("meshes",):
        shutil.copytree(golden("scenes/vol/meshes"), tmp_path / name)
    nb.nvdb_write_leaves(str(tmp_path / "big.nvdb"), org, vals, M)
    (tmp_path / "scene.xml").write_text(open(golden("scenes/vol/cbox_sparse_small.xml")).read().replace("sparse_smoke.nvdb", "big.nvdb"))
    s = nb.load_xml(str(tmp_path / "scene.xml"))
    m = s["media"][0]
    assert m["type"] == nb.MEDIUM_HETEROGENEOUS_BRICKS and list(m["grid_dims"]) == [1028] * 3
    entries = 129 ** 3
    first = int(m["first_voxel"])
    table = s["density"][first:first + entries].view("<u4").reshape(129, 129, 129)
    bricks = s["density"][first + entries:].reshape(-1, 10, 10, 10)
    # a block at array offset 2 (the border) straddles 8 bricks when its apron counts; the three are far apart
    assert 3 * 8 <= len(bricks) - 1 <= 3 * 27 and s["density"].size < 3 * 10 ** 6
    for o, v in zip(org, vals):                                                    # voxel (1, 2, 3) of each block, through the table
        a = o + 2 + np.array([1, 2, 3])
        b = table[tuple(a >> 3)]
        assert b != 0 and bricks[b][tuple((a & 7) + 1)] == v[1, 2, 3]
    assert m["max_density"] == vals.max()
    if os.path.exists(NVDB_TOOL):
        r = subprocess.run([NVDB_TOOL, "info", str(tmp_path / "big.nvdb")], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        info = json.loads(r.stdout)
        assert info["active"] == 3 * 512 and info["min"] == [0, 0, 0] and info["max"] == [1023, 1023, 1023]
        assert info["probe"] == pytest.approx([float(vals[0][1, 2, 3]), float(vals[1][1, 2, 3]), float(vals[2][1, 2, 3]), 0.0])


def test_nvdb_scene_takes_grid_zero_whatever_its_name(nb, built_lib, tmp_path):
    """NvdbVolume::loadNanoVdb reads grid 0 of the file by INDEX (NvdbVolume.vdb.cpp:65): a cache whose first grid is not
    called "density" still loads."""
    import shutil
    shutil.copytree(golden("scenes/vol/meshes"), tmp_path / "meshes")
    nb.to_nvdb(golden(VDB), str(tmp_path / "other.nvdb"), "density")
    g = nb.vdb_read(str(tmp_path / "other.nvdb"), "density")
    org = (np.array(g["bbox_min"]) & ~7)
    # the same voxels under another grid name, through the leaf writer
    v = g["values"]; pad = [(int(b - o), int((-(b - o + s)) % 8)) for b, o, s in zip(g["bbox_min"], org, v.shape)]
    vp = np.pad(v, pad)
    leaves, origins = [], []
    for i in range(0, vp.shape[0], 8):
        for j in range(0, vp.shape[1], 8):
            for k in range(0, vp.shape[2], 8):
                leaves.append(vp[i:i + 8, j:j + 8, k:k + 8]); origins.append(org + [i, j, k])
    nb.nvdb_write_leaves(str(tmp_path / "smoke.nvdb"), np.array(origins), np.array(leaves), g["index_to_world"], grid="smoke")
    xml = open(golden("scenes/vol/cbox_heterog_nvdb.xml")).read().replace("fluid_data_0050.nvdb", "smoke.nvdb")
    (tmp_path / "s.xml").write_text(xml)
    a, b = nb.load_xml(str(tmp_path / "s.xml")), nb.load_xml(golden("scenes/vol/cbox_heterog_nvdb.xml"))
    assert a["media"][0]["max_density"] == b["media"][0]["max_density"]
    # the leaf writer marks zero voxels inactive, so the box is the box of the non-zero voxels: compare where both have voxels
    assert np.count_nonzero(a["density"]) == np.count_nonzero(b["density"])


def test_vdb_leaves_kept_as_they_are(nb, built_lib, tmp_path, monkeypatch):
    """The .vdb reader's sparse form: the file's own leaf nodes and active masks (what the reference's openToNanoVDB keeps,
    NvdbVolume.vdb.cpp:118-119) go to the .nvdb cache and to the device's bricks without a dense box in between; the voxels
    are the dense reader's."""
    dense = nb.vdb_read(golden(VDB), "density")
    nb.to_nvdb(golden(VDB), str(tmp_path / "leaves.nvdb"), "density", sparse="file")
    back = nb.vdb_read(str(tmp_path / "leaves.nvdb"), "density")
    for key in ("bbox_min", "bbox_max", "index_to_world"):
        assert np.array_equal(back[key], dense[key]), key
    assert np.array_equal(back["values"], dense["values"]) and back["active_voxels"] == dense["active_voxels"] == 14112
    assert back["max"] == dense["max"] and back["min"] == dense["min"]
    if os.path.exists(NVDB_TOOL):                                          # and NanoVDB agrees about the file
        r = subprocess.run([NVDB_TOOL, "info", str(tmp_path / "leaves.nvdb")], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        info = json.loads(r.stdout)
        assert info["active"] == 14112 and info["min"] == [0, 0, 0] and info["max"] == [20, 20, 31] and info["leaves"] == 36
    monkeypatch.setenv("NB_VOLUME_SPARSE", "1")                            # the scene, through leaves -> bricks
    s = nb.load_xml(golden("scenes/vol/cbox_heterog.xml"))
    monkeypatch.delenv("NB_VOLUME_SPARSE")
    d = nb.load_xml(golden("scenes/vol/cbox_heterog.xml"))
    assert s["media"][0]["type"] == nb.MEDIUM_HETEROGENEOUS_BRICKS and d["media"][0]["type"] == nb.MEDIUM_HETEROGENEOUS
    full, _, _ = expand_bricks(s["density"], s["media"][0])
    assert np.array_equal(full, d["density"].reshape(25, 25, 36))


def test_temperature_scale_is_accepted_and_has_no_effect(nb, built_lib, tmp_path):
    """heterog's temperatureScale feeds a local the reference never returns (heterogmedium.cpp:152, :165-167): a scene that sets
    it loads, and flattens to the scene without it."""
    import shutil
    shutil.copytree(golden("scenes/vol/meshes"), tmp_path / "meshes")
    shutil.copy(golden("scenes/vol/fluid.nbv"), tmp_path / "fluid.nbv")
    xml = open(golden("scenes/vol/cbox_heterog_fluid_small.xml")).read()
    assert '<float name="densityScale"' in xml
    (tmp_path / "s.xml").write_text(xml.replace('<float name="densityScale"', '<float name="temperatureScale" value="3.5"/>\n\t\t\t<float name="densityScale"', 1))
    a, b = nb.load_xml(str(tmp_path / "s.xml")), nb.load_xml(golden("scenes/vol/cbox_heterog_fluid_small.xml"))
    assert a["media"].tobytes() == b["media"].tobytes() and np.array_equal(a["density"], b["density"])
