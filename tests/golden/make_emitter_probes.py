#!/usr/bin/env python
"""Emitter::sample known answers from the reference's own emitter plugins (`nori_ref emitter`):
for fixed (reference point, 2D sample) records the sampled direction, radiance/pdf, pdf, shadow-ray
length and eval — area light on a mesh (cbox), on an analytic sphere (pa3 sphere_ems), on a sphere
mesh, a point light, the PNG environment map (luminance CDF + spherical lookup) and the constant one.
Run in the build container; writes tests/golden/emit_<name>_{in,out}.bin."""
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
NORI_REF = os.path.join(os.path.dirname(os.path.dirname(HERE)), "oracle", "_ref", "nori_ref")
CASES = {  # name: (scene, emitter index, box the reference points are drawn from)
    "cbox_area": ("scenes/cbox/cbox_path_mis.xml", 0, ([-0.9, 0.05, -0.9], [0.9, 1.5, 0.9])),
    "sphere_area": ("scenes/sphere/sphere_ems.xml", 0, ([-3, -3, -3], [3, 3, 3])),
    "sphere2_area": ("scenes/sphere/sphere2_ems.xml", 0, ([-3, -3, -3], [3, 3, 3])),
    "sphere_mesh_area": ("scenes/sphere/sphere_mesh_ems.xml", 0, ([-3, -3, -3], [3, 3, 3])),
    "point": ("scenes/sphere/point_ems.xml", 0, ([-3, -3, -1], [3, 3, 1])),
    "envmap_png": ("scenes/envmap/envmap_sphere_small.xml", 0, ([-25, -25, -25], [25, 25, 25])),
    "envmap_const": ("scenes/envmap/textures_small.xml", 0, ([-25, -25, 0], [25, 25, 25])),
}


def main():
    rng = np.random.default_rng(2024)
    for name, (scene, index, (lo, hi)) in CASES.items():
        n = 4096
        rec = np.concatenate([rng.uniform(lo, hi, (n, 3)), rng.uniform(0, 1, (n, 2))], axis=1).astype("<f4")
        fin, fout = os.path.join(HERE, "emit_%s_in.bin" % name), os.path.join(HERE, "emit_%s_out.bin" % name)
        rec.tofile(fin)
        r = subprocess.run([NORI_REF, "emitter", os.path.join(HERE, scene), fin, fout, "index=%d" % index],
                           stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        if r.returncode != 0:
            raise RuntimeError(r.stderr[-2000:])
        print(name, r.stdout.strip().splitlines()[-1])


if __name__ == "__main__":
    main()
