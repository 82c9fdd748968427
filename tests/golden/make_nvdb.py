"""tests/golden/scenes/vol/fluid_data_0050.nvdb — the NanoVDB cache of fluid_data_0050.vdb's density grid as the reference's own
NanoVDB 28.1.2 writes it (oracle/_ref/nvdb_tool = nanovdb::GridBuilder + nanovdb::io::writeGrid, ext/nanovdb): the fixture the
.nvdb READER is pinned on, and the bytes the WRITER is compared with. The voxels and the index-to-world map come from the
.vdb through host/vdb.cpp (itself pinned on the file's own statistics and the density regression array).
Run in the build container: needs /root/reference for `make -C oracle ref`."""
import importlib, json, os, subprocess, sys, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
nb = importlib.import_module("optix-renderer_b200")

subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "ref"], check=True, stdout=subprocess.DEVNULL)
vol = os.path.join(ROOT, "tests", "golden", "scenes", "vol")
h = nb.load_host_library()
with tempfile.TemporaryDirectory() as tmp:
    nbv = os.path.join(tmp, "density.nbv")
    assert h.nbh_vdb_to_nbv(os.fsencode(os.path.join(vol, "fluid_data_0050.vdb")), b"density", os.fsencode(nbv)) == 0, h.nbh_last_error()
    out = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "nvdb_tool"), "write", nbv, os.path.join(vol, "fluid_data_0050.nvdb"), "density"],
                         check=True, capture_output=True, text=True).stdout
print(json.loads(out))

# nvdb_probe_rays.bin / nvdb_probe_out.bin: `nori_ref medium` on cbox_heterog_nvdb.xml — the .nvdb through the reference's own
# nanovdb::io::readGrid and SampleFromVoxels<.., 2> (oracle/ref_heterog.cpp): density at points, collision probability, mean
# free path, mean transmittance on rays through the grid's world box (index box x voxel size 0.1875, sigma_t = 2)
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
rng = np.random.default_rng(23)
n = 192
lo, hi = np.zeros(3), np.array([21, 21, 32]) * 0.1875
o = rng.uniform(lo - 0.3, hi + 0.3, (n, 3))
d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
rays = np.concatenate([o, d, np.full((n, 1), 1e-4), rng.uniform(0.3, 4.0, (n, 1))], axis=1).astype("<f4")
rays.tofile(os.path.join(HERE, "nvdb_probe_rays.bin"))
r = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "nori_ref"), "medium", os.path.join(vol, "cbox_heterog_nvdb.xml"),
                    os.path.join(HERE, "nvdb_probe_rays.bin"), os.path.join(HERE, "nvdb_probe_out.bin"), "samples=16384"],
                   capture_output=True, text=True)
assert r.returncode == 0, r.stderr[-2000:]
print(r.stdout.strip().splitlines()[-1])
out = np.fromfile(os.path.join(HERE, "nvdb_probe_out.bin"), "<f4").reshape(-1, 4)
print("nonzero densities", int((out[:, 0] != 0).sum()), "collision probability range", out[:, 1].min(), out[:, 1].max())

# scenes/vol/sparse_smoke.nvdb + cbox_sparse_small.xml: two separate puffs in a 96 x 48 x 120 box, stored SPARSE (only the
# 8^3 leaves that hold smoke) by host/nvdb.cpp's writer; sparse_probe_rays.bin / _out.bin = `nori_ref medium` on that scene, the
# file read by the reference's own NanoVDB. The device gets the grid in the brick layout (NB_MEDIUM_HETEROGENEOUS_BRICKS).
dims, org = (96, 48, 120), (-24, 5, 100)
x, y, z = np.meshgrid(*(np.arange(d, dtype=np.float64) for d in dims), indexing="ij")
def puff(c, r, a):
    d2 = ((x - c[0]) / r[0]) ** 2 + ((y - c[1]) / r[1]) ** 2 + ((z - c[2]) / r[2]) ** 2
    return np.where(d2 < 1.0, a * (1.0 - d2) ** 2 * (1.0 + 0.3 * np.sin(0.9 * x) * np.cos(0.7 * z)), 0.0)
vals = (puff((20, 22, 25), (14, 12, 18), 0.9) + puff((75, 30, 95), (16, 13, 20), 0.6)).astype("<f4")
s = 0.7 / 120.0                                     # world: x = 0.08 + k s, y = 0.02 + i s', z = -0.05 + j s'' — the box of the fluid cube
M = np.array([[0, 0, 0.7 / 120, 0.08 - 100 * 0.7 / 120], [1.04 / 96, 0, 0, 0.02 + 24 * 1.04 / 96], [0, 0.7 / 48, 0, -0.05 - 5 * 0.7 / 48]])
with tempfile.TemporaryDirectory() as tmp:
    nbv = os.path.join(tmp, "smoke.nbv")
    with open(nbv, "wb") as f:
        f.write(b"NBV1" + np.array(dims, "<i4").tobytes() + np.array(org, "<i4").tobytes() + M.astype("<f8").tobytes() + np.float32(0).tobytes() + vals.tobytes())
    nb.to_nvdb(nbv, os.path.join(vol, "sparse_smoke.nvdb"), "density", sparse=True)
xml = open(os.path.join(vol, "cbox_heterog_fluid_small.xml")).read()
assert "fluid.nbv" in xml
open(os.path.join(vol, "cbox_sparse_small.xml"), "w").write(xml.replace("fluid.nbv", "sparse_smoke.nvdb"))
n = 256
lo, hi = np.array([0.08, 0.02, -0.05]), np.array([0.78, 1.06, 0.65])
o = rng.uniform(lo - 0.03, hi + 0.03, (n, 3))
centres = np.array([M @ np.array([org[0] + 20, org[1] + 22, org[2] + 25, 1.0]), M @ np.array([org[0] + 75, org[1] + 30, org[2] + 95, 1.0])])
o[: 2 * n // 3] = centres[rng.integers(0, 2, 2 * n // 3)] + rng.normal(scale=0.06, size=(2 * n // 3, 3))      # most of them in and around the puffs
d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
rays = np.concatenate([o, d, np.full((n, 1), 1e-4), rng.uniform(0.05, 0.9, (n, 1))], axis=1).astype("<f4")
rays.tofile(os.path.join(HERE, "sparse_probe_rays.bin"))
r = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "nori_ref"), "medium", os.path.join(vol, "cbox_sparse_small.xml"),
                    os.path.join(HERE, "sparse_probe_rays.bin"), os.path.join(HERE, "sparse_probe_out.bin"), "samples=16384"],
                   capture_output=True, text=True)
assert r.returncode == 0, r.stderr[-2000:]
out = np.fromfile(os.path.join(HERE, "sparse_probe_out.bin"), "<f4").reshape(-1, 4)
print("sparse: file bytes", os.path.getsize(os.path.join(vol, "sparse_smoke.nvdb")), "nonzero densities", int((out[:, 0] != 0).sum()),
      "collision probability range", out[:, 1].min(), out[:, 1].max())
