#!/usr/bin/env python
"""C3 of SURVEY 8(d): a synthetic 1 310 720-triangle mesh in the Cornell box.

Icosphere subdivided `levels` times (20 * 4^levels triangles; 8 -> 1 310 720), vertices displaced
radially by 1 + 0.15 * fbm(v) (value-noise fBm hashed from the lattice, seed 1234), scaled to
radius 0.42 and placed at the centre of the reference's Cornell box, replacing the two spheres;
diffuse albedo 0.6; light, walls, camera and integrator as scenes/pa4/cbox/cbox_path_mis.xml.
Written as OBJ + XML so that the reference loader (src/shapes/obj.cpp:77) and this repo's host
loader read the same bytes.

    python tools/make_c3_scene.py out_dir [levels=8] [width=800 height=600 spp=1024]
"""
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
CBOX = os.path.join(HERE, "..", "tests", "golden", "scenes", "cbox")


def icosphere(levels):
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], dtype=np.float64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2],
                  [10, 7, 6], [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5],
                  [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]], dtype=np.int64)
    for _ in range(levels):
        e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])
        e.sort(axis=1)
        key = e[:, 0] * (len(v) + 1) + e[:, 1]
        uniq, inv = np.unique(key, return_inverse=True)
        a, b = uniq // (len(v) + 1), uniq % (len(v) + 1)
        mid = v[a] + v[b]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        m = len(v) + inv.reshape(3, -1).T          # midpoint ids of edges 01, 12, 20 per face
        v = np.concatenate([v, mid])
        f = np.concatenate([np.stack([f[:, 0], m[:, 0], m[:, 2]], 1), np.stack([f[:, 1], m[:, 1], m[:, 0]], 1),
                            np.stack([f[:, 2], m[:, 2], m[:, 1]], 1), np.stack([m[:, 0], m[:, 1], m[:, 2]], 1)])
    return v, f


def hash3(i, seed):
    h = (i[..., 0].astype(np.uint64) * np.uint64(73856093)) ^ (i[..., 1].astype(np.uint64) * np.uint64(19349663)) \
        ^ (i[..., 2].astype(np.uint64) * np.uint64(83492791)) ^ np.uint64(seed)
    h = (h ^ (h >> np.uint64(13))) * np.uint64(0x9E3779B97F4A7C15)
    h ^= h >> np.uint64(29)
    return (h & np.uint64(0xFFFFFF)).astype(np.float64) / float(1 << 24)


def value_noise(p, seed):
    i = np.floor(p).astype(np.int64)
    f = p - i
    w = f * f * (3 - 2 * f)
    out = 0.0
    for dx in (0, 1):
        for dy in (0, 1):
            for dz in (0, 1):
                c = hash3(i + np.array([dx, dy, dz]), seed)
                out = out + c * np.where(dx, w[:, 0], 1 - w[:, 0]) * np.where(dy, w[:, 1], 1 - w[:, 1]) * np.where(dz, w[:, 2], 1 - w[:, 2])
    return out


def fbm(p, seed=1234, octaves=5):
    amp, freq, total, norm = 1.0, 3.0, 0.0, 0.0
    for o in range(octaves):
        total = total + amp * value_noise(p * freq + 17.0 * o, seed + o)
        norm += amp
        amp *= 0.5
        freq *= 2.0
    return total / norm


def main():
    out = sys.argv[1]
    opt = dict(a.split("=", 1) for a in sys.argv[2:])
    levels = int(opt.get("levels", 8))
    os.makedirs(os.path.join(out, "meshes"), exist_ok=True)
    v, f = icosphere(levels)
    r = 1.0 + 0.15 * (2.0 * fbm(v) - 1.0)
    v = v * r[:, None] * 0.42 + np.array([0.0, 0.5, 0.0])
    with open(os.path.join(out, "meshes", "blob.obj"), "w") as fo:
        fo.write("# C3 synthetic mesh: %d vertices, %d triangles\n" % (len(v), len(f)))
        np.savetxt(fo, v, fmt="v %.6f %.6f %.6f")
        np.savetxt(fo, f + 1, fmt="f %d %d %d")
    for m in ("walls.obj", "leftwall.obj", "rightwall.obj", "light.obj"):
        shutil.copyfile(os.path.join(CBOX, "meshes", m), os.path.join(out, "meshes", m))
    xml = open(os.path.join(CBOX, "cbox_path_mis.xml")).read()
    a = xml.index('<shape type="sphere">')
    b = xml.rindex('</shape>', 0, xml.index('meshes/light.obj')) + len('</shape>')
    blob = ('<shape type="obj">\n\t\t<string name="filename" value="meshes/blob.obj"/>\n\t\t<bsdf type="diffuse">\n'
            '\t\t\t<color name="albedo" value="0.6 0.6 0.6"/>\n\t\t</bsdf>\n\t</shape>')
    xml = xml[:a] + blob + xml[b:]
    import re
    xml = re.sub(r'(name="width"\s+value=")\d+', r"\g<1>%s" % opt.get("width", "800"), xml)
    xml = re.sub(r'(name="height"\s+value=")\d+', r"\g<1>%s" % opt.get("height", "600"), xml)
    xml = re.sub(r'(name="sampleCount"\s+value=")\d+', r"\g<1>%s" % opt.get("spp", "1024"), xml)
    open(os.path.join(out, "c3_path_mis.xml"), "w").write(xml)
    print({"triangles": int(len(f)), "vertices": int(len(v)), "xml": os.path.join(out, "c3_path_mis.xml")})


if __name__ == "__main__":
    main()
