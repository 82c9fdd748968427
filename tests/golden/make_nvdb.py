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
