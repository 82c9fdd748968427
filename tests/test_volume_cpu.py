"""On-disk formats in front of the heterogeneous-medium path, no GPU: the OpenVDB reader
(host/vdb.cpp: Blosc + LZ4, active-mask compression, half floats), inflate, the `volume` /
`heterog` plugins' flatten hook."""
import re
import struct
import zlib

import ctypes as C
import numpy as np
import pytest

from conftest import golden

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
