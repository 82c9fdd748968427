#!/usr/bin/env python
"""Heterogeneous-medium fixtures (run in the build container; the GPU box only reads the results).

The reference's only heterogeneous scene, scenes/project/volume/cbox_heterog.xml, puts its grid
(world box [0,3.94]x[0,3.94]x[0,6], density zero below z-index 4) around a cube of [-0.25,0.25]^3:
the cube only ever sees empty voxels. So the fixtures are our own scenes on the reference's data:

  scenes/vol/fluid.nbv                 the "density" grid of fluid_data_0050.vdb, read by
                                       libnorib200_host's .vdb reader, re-mapped (voxel 1/32, grid z
                                       up the scene's y) so that it stands inside the Cornell box
  scenes/vol/const.nbv                 16^3 voxels of density 1 around the sphere of cbox_homog.xml
  scenes/vol/cbox_heterog_const_small.xml   cbox_homog_small.xml with its homogeneous medium replaced by
                                       `heterog` on const.nbv: same expected image as the REFERENCE's own
                                       homogeneous render (vol_homog_ref_128spp.npy) — the one pin
                                       the reference offers for this row
  scenes/vol/cbox_heterog_fluid_small.xml   the fluid in a cube, path_vol_mats; ..._mis_small.xml path_vol_mis
  vol_heterog_fluid_ref_128spp.npy, vol_heterog_fluid_mis_ref_128spp.npy
                                       renders by oracle/_ref/nori_ref with oracle/ref_heterog.cpp
                                       (the corrected restatement on the reference's NanoVDB; PARITY UNPINNED)
  vol_fluid_probe_rays.bin / _out.bin  `nori_ref medium`: density at points, free-path and
                                       transmittance means on fixed rays through the fluid
  vdb_density_21x21x32.npy             the decoded grid itself (for the CPU test of the .vdb reader)
"""
import ctypes as C
import os
import re
import shutil
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("NORI_REFERENCE", "/root/reference")
NORI_REF = os.path.join(ROOT, "oracle", "_ref", "nori_ref")
VOL = os.path.join(HERE, "scenes", "vol")
VDB = os.path.join(REF, "scenes/project/volume/fluid_data_0050.vdb")


def write_nbv(path, values, bbox_min, index_to_world, background=0.0):
    values = np.ascontiguousarray(values, dtype="<f4")
    with open(path, "wb") as f:
        f.write(b"NBV1")
        f.write(np.array(values.shape, dtype="<i4").tobytes())
        f.write(np.array(bbox_min, dtype="<i4").tobytes())
        f.write(np.array(index_to_world, dtype="<f8").reshape(12).tobytes())
        f.write(np.array([background], dtype="<f4").tobytes())
        f.write(values.tobytes())


def read_vdb(path, grid):
    lib = C.CDLL(os.path.join(ROOT, "optix-renderer_b200", "libnorib200_host.so"))
    lib.nbh_last_error.restype = C.c_char_p
    lib.nbh_vdb_read.restype = C.c_longlong
    bbox = (C.c_int32 * 6)(); xf = (C.c_double * 12)(); st = (C.c_float * 3)(); act = C.c_uint64()
    n = lib.nbh_vdb_read(path.encode(), grid.encode(), bbox, xf, st, C.byref(act), None, 0)
    if n < 0:
        raise RuntimeError(lib.nbh_last_error().decode())
    v = np.zeros(n, np.float32)
    lib.nbh_vdb_read(path.encode(), grid.encode(), bbox, xf, st, C.byref(act), v.ctypes.data_as(C.c_void_p), n)
    dims = [bbox[3 + k] - bbox[k] + 1 for k in range(3)]
    return v.reshape(dims), list(bbox), np.array(list(xf)).reshape(3, 4), list(st), act.value


HETEROG = """		<medium type="heterog">
			<color name="sigma_a" value="%s"/>
			<color name="sigma_s" value="%s"/>
			<float name="densityScale" value="%g"/>
			<volume type="volume">
				<string name="filename" value="%s"/>
			</volume>
		</medium>"""


def main():
    os.makedirs(os.path.join(VOL, "meshes"), exist_ok=True)
    shutil.copyfile(os.path.join(REF, "scenes/project/meshes/cube.obj"), os.path.join(VOL, "meshes", "cube.obj"))
    dens, bbox, xf, stats, active = read_vdb(VDB, "density")
    assert dens.shape == (21, 21, 32) and active == 21 * 21 * 32
    np.save(os.path.join(HERE, "vdb_density_21x21x32.npy"), dens)
    s = 1.0 / 32.0
    # index (i, j, k) -> world (x, y, z) = (i s + 0.102, k s + 0.03, j s - 0.028): the plume rises along y
    fluid_map = [[s, 0, 0, 0.102], [0, 0, s, 0.03], [0, s, 0, -0.028]]
    write_nbv(os.path.join(VOL, "fluid.nbv"), dens, bbox[:3], fluid_map)
    # constant grid around the sphere of cbox_homog.xml: centre (0.4458, 0.3321, 0.3767), r 0.3263
    c = np.array([0.4458, 0.3321, 0.3767]); v = 0.05
    o = c - 8 * v
    write_nbv(os.path.join(VOL, "const.nbv"), np.ones((16, 16, 16), np.float32), [0, 0, 0],
              [[v, 0, 0, o[0]], [0, v, 0, o[1]], [0, 0, v, o[2]]])

    homog = open(os.path.join(VOL, "cbox_homog_small.xml")).read()
    a = homog.index('<medium type="homog">', homog.index("<shape")); b = homog.index("</medium>", a) + len("</medium>")
    const_xml = homog[:a] + HETEROG.strip() % ("0.5, 0.5, 0.5", "0, 0, 0", 1.0, "const.nbv") + homog[b:]
    open(os.path.join(VOL, "cbox_heterog_const_small.xml"), "w").write(const_xml)

    # the fluid in a cube: replace the medium sphere by cube.obj around the grid's world box
    sa = homog.rindex("<shape", 0, a); sb = homog.index("</shape>", a) + len("</shape>")
    cube = """<shape type="obj">
		<string name="filename" value="meshes/cube.obj"/>
		<transform name="toWorld">
			<scale value="0.7,1.04,0.7"/>
			<translate value="0.43,0.54,0.3"/>
		</transform>
%s
	</shape>""" % (HETEROG % ("4, 4, 4", "1, 1, 1", 2.0, "fluid.nbv"))
    fluid_mis = homog[:sa] + cube + homog[sb:]
    assert 'type="path_vol_mis"' in fluid_mis
    open(os.path.join(VOL, "cbox_heterog_fluid_mis_small.xml"), "w").write(fluid_mis)
    open(os.path.join(VOL, "cbox_heterog_fluid_small.xml"), "w").write(fluid_mis.replace("path_vol_mis", "path_vol_mats"))
    # full-size variants (800 x 600, as scenes/project/volume/cbox_heterog.xml) for the throughput table
    for name in ("cbox_heterog_fluid", "cbox_heterog_fluid_mis"):
        t = open(os.path.join(VOL, name + "_small.xml")).read()
        t = re.sub(r'(name="width"\s+value=")\d+', r"\g<1>800", t); t = re.sub(r'(name="height"\s+value=")\d+', r"\g<1>600", t)
        open(os.path.join(VOL, name + ".xml"), "w").write(t)

    def ref(*args):
        r = subprocess.run([NORI_REF] + list(args), stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        if r.returncode != 0:
            raise RuntimeError(r.stderr[-2000:])
        print(r.stdout.strip().splitlines()[-1])

    ref("render", os.path.join(VOL, "cbox_heterog_fluid_small.xml"), os.path.join(HERE, "vol_heterog_fluid_ref_128spp.npy"), "spp=128")
    ref("render", os.path.join(VOL, "cbox_heterog_fluid_mis_small.xml"), os.path.join(HERE, "vol_heterog_fluid_mis_ref_128spp.npy"), "spp=128")
    # probes: rays starting inside the fluid's world box
    rng = np.random.default_rng(7)
    n = 256
    lo = np.array([0.102, 0.03, -0.028]); hi = lo + np.array([21, 32, 21]) * s
    o = rng.uniform(lo - 0.05, hi + 0.05, (n, 3))
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.concatenate([o, d, np.full((n, 1), 1e-4), rng.uniform(0.05, 0.8, (n, 1))], axis=1).astype("<f4")
    rays.tofile(os.path.join(HERE, "vol_fluid_probe_rays.bin"))
    ref("medium", os.path.join(VOL, "cbox_heterog_fluid_small.xml"), os.path.join(HERE, "vol_fluid_probe_rays.bin"),
        os.path.join(HERE, "vol_fluid_probe_out.bin"), "samples=16384")


if __name__ == "__main__":
    main()
