#!/usr/bin/env python
"""Round-2 fixtures: the parity pins VERDICT r01 asked for. Run in the build container (needs /root/reference and
oracle/_ref/nori_ref); the GPU box only reads what this writes.   python tests/golden/make_round2.py [only=<part>]

  pins       hi_<scene>_R<8N>.npy            the reference CPU renderer at 8N spp (SURVEY 8d's "R")
             hi_<scene>_cpu<N>.json          relRMSE / bias of an INDEPENDENT reference render at N spp against R
                                             (passes [8N, 9N) of the same run, `nori_ref render split=`): the figure the GPU
                                             has to stay within 1.10x of
  exr        exr_<name>.npz                  the reference's own EXR goldens (scenes/pa3/**/ref, scenes/pa4/table/ref),
                                             box-downsampled 4x4 to fp16 + the full-resolution channel means
             ../../profiles/r02_oracle_vs_exr.json   the oracle itself against those goldens (pins the oracle)
  probes     bsdf_{mirror,dielectric,microfacet_veach,microfacet_table}_{in,out}.bin   BSDF::eval/pdf/sample records
  scenes     scenes/round2/*.xml + r2_<name>_ref_<spp>.npy   thin-lens camera, Schlick phase (mats + mis), `direct` integrator
                                             images; filter_<name>.nbs: the reference's own filter tables (nori_ref flat)
             scenes/ttests/pa1/              scenes/pa1/test-direct.xml + disk.obj (the reference's `direct` known answers)
             scenes/round2/microfacet_ttest.xml   the BSDF of scenes/pa3/tests/ttest-microfacet.xml on a sphere
  pcg32      pcg32_demo_round1.txt           "Round 1" of ext/pcg32/pcg32-demo.out
"""
import json
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
R2 = os.path.join(HERE, "scenes", "round2")
ONLY = dict(a.split("=", 1) for a in sys.argv[1:]).get("only")


def run(*args):
    r = subprocess.run([NORI_REF] + list(args), stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    if r.returncode != 0:
        raise RuntimeError(" ".join(args) + "\n" + r.stderr[-2000:])
    return r.stdout.strip().splitlines()[-1]


def rel_rmse(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.sqrt(np.mean((a - b) ** 2 / (b ** 2 + 1e-2))))


def retarget(src, dst, **subs):
    """copy of a scene XML with regex substitutions (relative mesh paths keep working: same directory)"""
    t = open(src).read()
    for pat, rep in subs.values():
        t, n = re.subn(pat, rep, t, flags=re.S)
        assert n >= 1, (src, pat)
    open(dst, "w").write(t)


# ---- high-spp pins ------------------------------------------------------------------------------------
PINS = [  # name, xml (under scenes/), N
    ("cbox_path_mis", "cbox/cbox_path_mis_small.xml", 512),
    ("cbox_path_mats", "cbox/cbox_path_mats_small.xml", 512),
    ("veach_direct_mis", "veach/veach_mis_small.xml", 256),
    ("table_path_mis", "table/table_path_mis_small.xml", 128),
    ("val_path_vol_mis", "vol/cbox_val_mis_small.xml", 256),
    ("homog_aniso_path_vol_mats", "vol/cbox_homog_aniso_small.xml", 256),
]


def make_pins():
    for name, xml, n in PINS:
        src = os.path.join(HERE, "scenes", xml)
        r_path = os.path.join(HERE, "hi_%s_R%d.npy" % (name, 8 * n))
        tmp = "/tmp/hi_%s_cpu.npy" % name
        K = 5                                      # independent CPU_N renders: relRMSE of a heavy-tailed image is itself noisy
        print(name, run("render", src, r_path, "spp=%d" % ((8 + K) * n), "split=%d" % (8 * n), "seg=%d" % n, "out2=" + tmp))
        R = np.load(r_path)
        cs = [np.load("%s.%d" % (tmp, k)) for k in range(K)]
        errs = [rel_rmse(c, R) for c in cs]
        biases = [((c.mean(axis=(0, 1)) - R.mean(axis=(0, 1))) / R.mean(axis=(0, 1))).tolist() for c in cs]
        # standard error of R's channel means, from the per-pixel spread of the CPU_N around R (Var(R pixel) ~ Var(CPU_N pixel) / 8)
        d = np.mean([(c.astype(np.float64) - R) ** 2 for c in cs], axis=0)
        se_R = (np.sqrt(d.sum(axis=(0, 1)) / (1.0 + 1.0 / 8.0) / 8.0) / (R.shape[0] * R.shape[1]) / R.mean(axis=(0, 1))).tolist()
        j = {"scene": xml, "n_spp": n, "r_spp": 8 * n, "cpu_relrmse_vs_R": float(np.median(errs)), "cpu_relrmse_each": errs,
             "cpu_bias_vs_R": biases, "rel_se_of_R_mean": se_R,
             "how": "nori_ref render spp=%d split=%d seg=%d: R = passes [0,%d), CPU_N^k = passes [%d + k*%d, ...) of the same reference run, "
                    "k < %d; cpu_relrmse_vs_R is their median" % ((8 + K) * n, 8 * n, n, 8 * n, 8 * n, n, K)}
        json.dump(j, open(os.path.join(HERE, "hi_%s_cpu%d.json" % (name, n)), "w"), indent=1)
        print("  ", errs, se_R)


# ---- the reference's EXR goldens ---------------------------------------------------------------------------
EXRS = [  # fixture name, reference EXR, scene xml in the reference tree, spp of the golden
    ("table_path_mis", "scenes/pa4/table/ref/table_path_mis_512spp.exr", "scenes/pa4/table/table_path_mis.xml"),
    ("table_path_mats", "scenes/pa4/table/ref/table_path_mats_512spp.exr", "scenes/pa4/table/table_path_mats.xml"),
    ("veach_mis", "scenes/pa3/veach_mi/ref/veach_mis_128spp.exr", "scenes/pa3/veach_mi/veach_mis.xml"),
    ("veach_ems", "scenes/pa3/veach_mi/ref/veach_ems_256spp.exr", "scenes/pa3/veach_mi/veach_ems.xml"),
    ("veach_mats", "scenes/pa3/veach_mi/ref/veach_mats_256spp.exr", "scenes/pa3/veach_mi/veach_mats.xml"),
    ("odyssey_mis", "scenes/pa3/odyssey/ref/odyssey_mis_32spp.exr", "scenes/pa3/odyssey/odyssey_mis.xml"),
    ("odyssey_ems", "scenes/pa3/odyssey/ref/odyssey_ems_64spp.exr", "scenes/pa3/odyssey/odyssey_ems.xml"),
    ("odyssey_mats", "scenes/pa3/odyssey/ref/odyssey_mats_64spp.exr", "scenes/pa3/odyssey/odyssey_mats.xml"),
]


def pool4(img):
    h, w = img.shape[0] // 4 * 4, img.shape[1] // 4 * 4
    return img[:h, :w].reshape(h // 4, 4, w // 4, 4, 3).mean(axis=(1, 3))


def read_exr(path):
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    import cv2
    img = cv2.imread(path, cv2.IMREAD_UNCHANGED)
    assert img is not None, path
    return np.ascontiguousarray(img[..., 2::-1].astype(np.float32))     # BGR(A) -> RGB


def make_exr(check_oracle=True):
    import glob
    todo = list(EXRS)
    for f in sorted(glob.glob(os.path.join(REF, "scenes/pa3/sphere/ref/*.exr"))):
        base = os.path.basename(f)[:-4]
        xml = os.path.join("scenes/pa3/sphere", base + ".xml")
        if os.path.exists(os.path.join(REF, xml)):
            todo.append(("sphere_" + base, os.path.relpath(f, REF), xml))
    report = []
    for name, exr, xml in todo:
        img = read_exr(os.path.join(REF, exr))
        spp = int(re.search(r'name="sampleCount"\s+value="(\d+)"', open(os.path.join(REF, xml)).read()).group(1))
        np.savez_compressed(os.path.join(HERE, "exr_%s.npz" % name), pooled=pool4(img).astype(np.float16),
                            mean=img.reshape(-1, 3).mean(axis=0).astype(np.float64), shape=np.array(img.shape[:2]), spp=np.array(spp),
                            source=np.array(exr), scene=np.array(xml))
        row = {"golden": exr, "scene": xml, "spp": spp, "shape": list(img.shape[:2]), "golden_mean": img.reshape(-1, 3).mean(axis=0).tolist()}
        if check_oracle:
            out = "/tmp/exr_check.npy"
            j = json.loads(run("render", os.path.join(REF, xml), out))
            o = np.load(out)
            assert o.shape == img.shape, (o.shape, img.shape)
            gm = img.reshape(-1, 3).mean(axis=0)
            row.update({"oracle_mean": o.reshape(-1, 3).mean(axis=0).tolist(),
                        "oracle_bias": ((o.reshape(-1, 3).mean(axis=0) - gm) / gm).tolist(),
                        "oracle_relrmse_pooled4": rel_rmse(pool4(o), pool4(img)), "oracle_relrmse_fullres": rel_rmse(o, img),
                        "oracle_seconds": j["seconds"], "oracle_threads": j["threads"]})
            p = os.path.join(HERE, "exr_%s.npz" % name)              # the oracle's own figures travel with the fixture
            d = dict(np.load(p))
            d["oracle_mean"] = o.reshape(-1, 3).mean(axis=0).astype(np.float64); d["oracle_relrmse_pooled4"] = np.array(row["oracle_relrmse_pooled4"])
            np.savez_compressed(p, **d)
        print(json.dumps(row), flush=True)
        report.append(row)
    if check_oracle:
        json.dump(report, open(os.path.join(ROOT, "profiles", "r02_oracle_vs_exr.json"), "w"), indent=1)


# ---- BSDF probes for mirror / dielectric / microfacet ---------------------------------------------------------
def make_probes():
    rng = np.random.default_rng(20260211)
    n = 6000

    def sphere(k):
        v = rng.normal(size=(k, 3)).astype(np.float32)
        return v / np.linalg.norm(v, axis=1, keepdims=True)
    for name, xml, shape in [("mirror", "cbox/cbox_path_mis.xml", 3), ("dielectric", "cbox/cbox_path_mis.xml", 4),
                             ("microfacet_veach", "veach/veach_mis.xml", 4), ("microfacet_table", "table/table_path_mis.xml", 2)]:
        wi, wo = sphere(n), sphere(n)
        if name.startswith("microfacet"):        # most records in the lobe: both directions above the surface
            wi[: n * 3 // 4, 2] = np.abs(wi[: n * 3 // 4, 2]); wo[: n * 3 // 4, 2] = np.abs(wo[: n * 3 // 4, 2])
            # and a batch at grazing and at near-specular configurations
            k = n // 8
            wo[:k] = wi[:k] * np.array([-1, -1, 1], np.float32) + 0.02 * rng.normal(size=(k, 3)).astype(np.float32)
            wo[:k] /= np.linalg.norm(wo[:k], axis=1, keepdims=True)
        rec = np.concatenate([wi, wo, rng.random((n, 2), dtype=np.float32), rng.random((n, 2), dtype=np.float32)], axis=1).astype("<f4")
        fin, fout = os.path.join(HERE, "bsdf_%s_in.bin" % name), os.path.join(HERE, "bsdf_%s_out.bin" % name)
        rec.tofile(fin)
        print(name, run("bsdf", os.path.join(HERE, "scenes", xml), fin, fout, "index=%d" % shape))


# ---- small scenes for the untested branches -------------------------------------------------------------------
def make_scenes():
    os.makedirs(R2, exist_ok=True)
    cb = os.path.join(HERE, "scenes", "cbox")
    if not os.path.isdir(os.path.join(R2, "meshes")):
        shutil.copytree(os.path.join(cb, "meshes"), os.path.join(R2, "meshes"))
    # thin lens (perspective.cpp:112-127): focus on the spheres, walls out of focus
    retarget(os.path.join(cb, "cbox_path_mis_small.xml"), os.path.join(R2, "cbox_dof.xml"),
             a=(r'(<float name="fov" value="27.7856"/>)', r'\1\n\t\t<float name="lensRadius" value="0.06"/>\n\t\t<float name="focalDistance" value="4.9"/>'))
    # `direct` integrator (direct.cpp): point lights only contribute — the cbox with two point lights added
    retarget(os.path.join(cb, "cbox_path_mis_small.xml"), os.path.join(R2, "cbox_direct.xml"),
             a=(r'<integrator type="path_mis"/>', '<integrator type="direct"/>\n\t<emitter type="point"><point name="position" value="0.3,1.2,0.5"/>'
                '<color name="power" value="20,16,12"/></emitter>\n\t<emitter type="point"><point name="position" value="-0.5,0.4,1.0"/>'
                '<color name="power" value="4,8,12"/></emitter>'))
    vol = os.path.join(HERE, "scenes", "vol")
    if not os.path.isdir(os.path.join(R2, "vol_meshes")):
        pass
    # Schlick phase (schlickphase.cpp, registered as "schlick"): a scattering medium so that the phase function matters
    for integ in ("path_vol_mats", "path_vol_mis"):
        src = os.path.join(vol, "cbox_homog_aniso_small.xml")
        dst = os.path.join(vol, "cbox_homog_schlick_%s_small.xml" % integ.split("_")[-1])
        retarget(src, dst,
                 a=(r'<integrator type="path_vol_mats"/>', '<integrator type="%s"/>' % integ),
                 b=(r'<color name="sigma_a" value="2, 2, 2"/>\s*<phase type="anisophase">\s*<float name="g" value="-0.75"/>\s*</phase>',
                    '<color name="sigma_a" value="0.5, 0.5, 0.5"/>\n\t\t\t<color name="sigma_s" value="4, 4, 4"/>\n\t\t\t<phase type="schlick">\n\t\t\t\t<float name="g" value="0.7"/>\n\t\t\t</phase>'))
    for name, path, spp in [("dof", os.path.join(R2, "cbox_dof.xml"), 256), ("direct", os.path.join(R2, "cbox_direct.xml"), 64),
                            ("schlick_mats", os.path.join(vol, "cbox_homog_schlick_mats_small.xml"), 256),
                            ("schlick_mis", os.path.join(vol, "cbox_homog_schlick_mis_small.xml"), 256)]:
        print(name, run("render", path, os.path.join(HERE, "r2_%s_ref_%dspp.npy" % (name, spp)), "spp=%d" % spp))
    # reconstruction filters (rfilter.cpp:87-163): the reference's own tables through `nori_ref flat`
    for f, extra in [("mitchell", ""), ("tent", ""), ("box", ""), ("gaussian", '<float name="radius" value="3"/><float name="stddev" value="0.8"/>'),
                     ("mitchell_bc", '<float name="radius" value="2.5"/><float name="B" value="0.2"/><float name="C" value="0.4"/>')]:
        xml = os.path.join(R2, "filter_%s.xml" % f)
        retarget(os.path.join(cb, "cbox_path_mis_small.xml"), xml,
                 a=(r'(<integer name="width" value="200"/>)', r'\1\n\t\t<rfilter type="%s">%s</rfilter>' % (f.split("_")[0], extra)))
        print(f, run("flat", xml, os.path.join(HERE, "filter_%s.nbs" % f)))
    # scene files of the reference that its EXR goldens belong to and that were not fixtures yet (verbatim; the sphere
    # scenes name ../../pa1/plane.obj, which lives next to them here)
    shutil.copyfile(os.path.join(REF, "scenes/pa4/table/table_path_mats.xml"), os.path.join(HERE, "scenes", "table", "table_path_mats.xml"))
    for f in ("sphere_mats.xml", "sphere2_mats.xml", "sphere2_mesh_ems.xml"):
        t = open(os.path.join(REF, "scenes/pa3/sphere", f)).read().replace("../../pa1/plane.obj", "plane.obj")
        open(os.path.join(HERE, "scenes", "sphere", f), "w").write(t)
    # the reference's `direct` known answers
    d = os.path.join(HERE, "scenes", "ttests", "pa1")
    os.makedirs(d, exist_ok=True)
    shutil.copyfile(os.path.join(REF, "scenes/pa1/test-direct.xml"), os.path.join(d, "test-direct.xml"))
    shutil.copyfile(os.path.join(REF, "scenes/pa1/disk.obj"), os.path.join(d, "disk.obj"))
    # the BSDF of scenes/pa3/tests/ttest-microfacet.xml on a shape, so that nb_bsdf_probe can reach it
    t = open(os.path.join(REF, "scenes/pa3/tests/ttest-microfacet.xml")).read()
    bsdf = re.search(r"<bsdf.*?</bsdf>", t, flags=re.S).group(0)
    open(os.path.join(R2, "microfacet_ttest.xml"), "w").write(
        '<scene>\n\t<integrator type="direct_mis"/>\n\t<sampler type="independent"><integer name="sampleCount" value="4"/></sampler>\n'
        '\t<camera type="perspective"><transform name="toWorld"><lookat target="0,0,0" origin="0,0,4" up="0,1,0"/></transform>'
        '<integer name="width" value="16"/><integer name="height" value="16"/></camera>\n'
        '\t<shape type="sphere"><point name="center" value="0,0,0"/><float name="radius" value="1"/>\n\t' + bsdf + '\n\t</shape>\n'
        '\t<emitter type="point"><point name="position" value="2,3,4"/><color name="power" value="100,100,100"/></emitter>\n</scene>\n')
    json.dump({"angles": [float(x) for x in re.search(r'name="angles"\s+value="([^"]+)"', t).group(1).split(",")],
               "references": [float(x) for x in re.search(r'name="references"\s+value="([^"]+)"', t).group(1).split(",")],
               "source": "scenes/pa3/tests/ttest-microfacet.xml"}, open(os.path.join(HERE, "microfacet_ttest.json"), "w"), indent=1)


def make_pcg32():
    t = open(os.path.join(REF, "ext/pcg32/pcg32-demo.out")).read()
    m = re.search(r"Round 1:\s*\n\s*32bit:\s*(.*)", t)
    open(os.path.join(HERE, "pcg32_demo_round1.txt"), "w").write(m.group(1).strip() + "\n")


if __name__ == "__main__":
    parts = {"pcg32": make_pcg32, "probes": make_probes, "scenes": make_scenes, "pins": make_pins, "exr": make_exr}
    for k, f in parts.items():
        if ONLY is None or ONLY == k:
            print("==", k, flush=True)
            f()
