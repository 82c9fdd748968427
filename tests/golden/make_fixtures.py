#!/usr/bin/env python
"""Regenerates tests/golden/ from the reference (run in the build container, where
/root/reference and oracle/_ref/nori_ref exist; the GPU box only ever reads the results).

  scenes/<name>/...      the reference's scene XML + OBJ inputs, copied verbatim (data, not code)
                         plus *_small.xml variants (same scene, smaller image) for image goldens
  <name>.nbs             flattened scene dumped from the reference's own objects after
                         Scene::update() (`nori_ref flat`), with the BSDF records added here from
                         the XML (they are private to the reference's .cpp files)
  <name>_rays.bin        fixed ray set made by `nori_ref genrays` (n x 8 float32)
  <name>_hits.bin        `nori_ref hits` result reduced to nb_hit records (n x 24 bytes)
  <name>_ref_<spp>.npy   reference CPU render of the *_small scene
  rng_kat.txt            pcg32 known-answer streams (`nori_ref rng`)
  scenes/ttests/<pa>/    the reference's own Student-t known-answer scenes (scenes/pa3/tests,
                         scenes/pa4/tests: XML with the expected radiance per scene + OBJ inputs)
"""
import importlib
import json
import os
import re
import shutil
import subprocess
import sys
import xml.etree.ElementTree as ET

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("NORI_REFERENCE", "/root/reference")
NORI_REF = os.path.join(ROOT, "oracle", "_ref", "nori_ref")
sys.path.insert(0, ROOT)
nb = importlib.import_module("optix-renderer_b200")

SCENES = {
    # name: (xml relative to REF/scenes, small (w, h), golden spp, n primary rays)
    "cbox": ("pa4/cbox/cbox_path_mis.xml", (200, 150), 256, 6000),
    "cbox_mats": ("pa4/cbox/cbox_path_mats.xml", (200, 150), 256, 0),
    "table": ("pa4/table/table_path_mis.xml", (200, 150), 64, 6000),
    # depth-1 integrators on the reference's pa3 scenes (microfacet plates, 4 mesh area lights)
    "veach": ("pa3/veach_mi/veach_mis.xml", (192, 128), 64, 0),
    "veach_ems": ("pa3/veach_mi/veach_ems.xml", (192, 128), 64, 0),
    "veach_mats": ("pa3/veach_mi/veach_mats.xml", (192, 128), 64, 0),
    "odyssey": ("pa3/odyssey/odyssey_ems.xml", (192, 108), 64, 0),
    "odyssey_mis": ("pa3/odyssey/odyssey_mis.xml", (192, 108), 64, 0),
    # homogeneous participating media (project/volume): iso / Henyey-Greenstein phase, medium-only
    # shapes, an ambient medium, both volume integrators
    "vol_val_mats": ("project/volume/cbox_val_mats.xml", (150, 150), 128, 0),
    "vol_val_mis": ("project/volume/cbox_val_mis.xml", (150, 150), 128, 0),
    "vol_homog_aniso": ("project/volume/cbox_homog_aniso.xml", (150, 150), 128, 0),
    "vol_homog_ambient": ("project/volume/cbox_homog_ambient.xml", (150, 150), 128, 0),
    "vol_homog": ("project/volume/cbox_homog.xml", (150, 150), 128, 0),
}
# image-only fixtures: same geometry as their first variant, or (volumes) media the oracle's flat
# dump does not carry — those scenes reach the GPU through the C++ host loader
NO_FLAT = {"veach_ems", "veach_mats", "odyssey_mis", "vol_val_mats", "vol_val_mis", "vol_homog_aniso",
           "vol_homog_ambient", "vol_homog"}


def run(*args):
    r = subprocess.run([NORI_REF] + list(args), stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    if r.returncode != 0:
        raise RuntimeError("nori_ref %s failed:\n%s" % (" ".join(args), r.stderr[-2000:]))
    return json.loads(r.stdout.strip().splitlines()[-1]) if r.stdout.strip().startswith("{") else r.stdout


def floats(s):
    return [float(x) for x in re.split(r"[\s,]+", s.strip()) if x]


def bsdf_record(node):
    """nb_bsdf from a <bsdf> element (defaults as in the reference constructors)."""
    rec = np.zeros((), dtype=nb.bsdf_dtype)
    props = {c.get("name"): c.get("value") for c in node} if node is not None else {}
    kind = node.get("type") if node is not None else "diffuse"
    if kind == "diffuse":        # src/bsdf/diffuse.cpp:31-50
        rec["type"] = nb.BSDF_DIFFUSE
        rec["albedo"] = floats(props["albedo"]) if "albedo" in props else [0.5] * 3
    elif kind == "mirror":
        rec["type"] = nb.BSDF_MIRROR
    elif kind == "dielectric":   # src/bsdf/dielectric.cpp:28-35
        rec["type"] = nb.BSDF_DIELECTRIC
        rec["int_ior"] = float(props.get("intIOR", 1.5046)); rec["ext_ior"] = float(props.get("extIOR", 1.000277))
    elif kind == "microfacet":   # src/bsdf/microfacet.cpp:33-53
        rec["type"] = nb.BSDF_MICROFACET
        rec["alpha"] = float(props.get("alpha", 0.1))
        rec["int_ior"] = float(props.get("intIOR", 1.5046)); rec["ext_ior"] = float(props.get("extIOR", 1.000277))
        kd = np.array(floats(props["kd"]) if "kd" in props else [0.5] * 3, dtype=np.float32)
        rec["albedo"] = kd
        rec["ks"] = np.float32(1) - kd.max()
    else:
        raise ValueError("bsdf %s" % kind)
    return rec


def copy_scene(xml_rel, name):
    src = os.path.join(REF, "scenes", xml_rel)
    dst_dir = os.path.join(HERE, "scenes", name)
    os.makedirs(dst_dir, exist_ok=True)
    tree = ET.parse(src)
    text = open(src).read()
    for s in tree.getroot().iter("string"):
        if s.get("name") == "filename":
            rel = s.get("value")
            to = rel
            if rel.startswith(".."):         # inputs outside the scene's folder are gathered under meshes/
                to = "meshes/" + os.path.basename(rel)
                text = text.replace('"%s"' % rel, '"%s"' % to)
            os.makedirs(os.path.dirname(os.path.join(dst_dir, to)), exist_ok=True)
            shutil.copyfile(os.path.join(os.path.dirname(src), rel), os.path.join(dst_dir, to))
    dst = os.path.join(dst_dir, os.path.basename(src))
    open(dst, "w").write(text)
    return dst


def small_variant(xml, size):
    text = open(xml).read()
    text = re.sub(r'(name="width"\s+value=")\d+', r"\g<1>%d" % size[0], text)
    text = re.sub(r'(name="height"\s+value=")\d+', r"\g<1>%d" % size[1], text)
    out = xml[:-4] + "_small.xml"
    open(out, "w").write(text)
    return out


TTESTS = {"pa3": ["test-mesh.xml", "test-mesh-furnace.xml"], "pa4": ["test-direct.xml", "test-furnace.xml"]}


def copy_ttests():
    for pa, files in TTESTS.items():
        src_dir = os.path.join(REF, "scenes", pa, "tests")
        dst_dir = os.path.join(HERE, "scenes", "ttests", pa)
        os.makedirs(dst_dir, exist_ok=True)
        for f in files:
            shutil.copyfile(os.path.join(src_dir, f), os.path.join(dst_dir, f))
            for s in ET.parse(os.path.join(src_dir, f)).getroot().iter("string"):
                if s.get("name") == "filename":
                    shutil.copyfile(os.path.join(src_dir, s.get("value")), os.path.join(dst_dir, s.get("value")))


def main():
    os.chmod(HERE, 0o755)
    copy_ttests()
    only = set(sys.argv[1:])
    for name, (xml_rel, small, spp, n_primary) in SCENES.items():
        if only and name not in only:
            continue
        xml = copy_scene(xml_rel, name.split("_")[0] if name != "cbox_mats" else "cbox")
        sx = small_variant(xml, small)
        out = os.path.join(HERE, "%s_ref_%dspp.npy" % (name, spp))
        print(run("render", sx, out, "spp=%d" % spp))
        if name in NO_FLAT:
            continue
        nbs = os.path.join(HERE, name + ".nbs")
        print(run("flat", xml, nbs))
        sec = nb.read_nbs(nbs)
        shapes = sec["shapes"]
        bsdfs = []
        xshapes = [s for s in ET.parse(xml).getroot() if s.tag in ("shape", "mesh")]
        assert len(xshapes) == len(shapes)
        for i, xs in enumerate(xshapes):
            shapes[i]["bsdf"] = len(bsdfs)
            bsdfs.append(bsdf_record(xs.find("bsdf")))
        sec["bsdfs"] = np.array(bsdfs, dtype=nb.bsdf_dtype)
        integ = ET.parse(xml).getroot().find("integrator").get("type")
        sec["render"]["integrator"] = nb.INTEGRATORS[integ]
        nb.write_nbs(nbs, sec)
        if n_primary:
            rays = os.path.join(HERE, name + "_rays.bin"); hits_full = "/tmp/%s_hits_full.bin" % name
            print(run("genrays", xml, rays, str(n_primary), "2024"))
            print(run("hits", xml, rays, hits_full))
            full = np.fromfile(hits_full, dtype=np.uint8).reshape(-1, 128)
            h = np.zeros(len(full), dtype=nb.hit_dtype)
            h["shape"] = full[:, 0:4].copy().view("<i4")[:, 0]
            h["prim"] = full[:, 4:8].copy().view("<i4")[:, 0]
            tuv = full[:, 8:20].copy().view("<f4")
            h["t"], h["u"], h["v"] = tuv[:, 0], tuv[:, 1], tuv[:, 2]
            h["any_hit"] = full[:, 28:32].copy().view("<u4")[:, 0]
            # BVH::findShape (bvh.h:105-109) rewrites the index in place: prim is already shape-local
            hit = h["shape"] >= 0
            h.tofile(os.path.join(HERE, name + "_hits.bin"))
            # per-ray traversal work, for the roofline's bytes/ray
            nodes = full[:, 20:24].copy().view("<u4")[:, 0]; prims = full[:, 24:28].copy().view("<u4")[:, 0]
            json.dump({"rays": int(len(full)), "mean_nodes_visited": float(nodes.mean()),
                       "mean_prims_tested": float(prims.mean()), "hit_fraction": float(hit.mean())},
                      open(os.path.join(HERE, name + "_raystats.json"), "w"))
    open(os.path.join(HERE, "rng_kat.txt"), "w").write(run("rng", "16"))


if __name__ == "__main__":
    main()
