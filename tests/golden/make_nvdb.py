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
