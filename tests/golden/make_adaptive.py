"""adaptive_cbox_ref.npy / adaptive_cbox_passes.npy: scenes/cbox/cbox_adaptive_small.xml (cbox_path_mis_small.xml with
<sampler type="adaptive"> sampleCount 8, initialUniform 2) through the reference's own AdaptiveSampler and block loop
(src/samplers/adaptive.cpp, render.cpp:319-335 as oracle/ref_main.cpp runs it): the image, and how many passes every 4 x 4
block rendered over the 8 rounds (in BlockGenerator order — only its distribution is compared). adaptive_cbox_ref2.npy: the
same with the per-block generators seeded differently is not available (adaptive.cpp:41-43 seeds from the block offset), so
the noise floor of the comparison comes from two device seeds. Run in the build container (needs oracle/_ref/nori_ref)."""
import json, os, subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
r = subprocess.run([os.path.join(ROOT, "oracle", "_ref", "nori_ref"), "render", os.path.join(HERE, "scenes", "cbox", "cbox_adaptive_small.xml"),
                    os.path.join(HERE, "adaptive_cbox_ref.npy"), "passes=" + os.path.join(HERE, "adaptive_cbox_passes.npy")],
                   capture_output=True, text=True)
assert r.returncode == 0, r.stderr[-2000:]
p = np.load(os.path.join(HERE, "adaptive_cbox_passes.npy"))
np.save(os.path.join(HERE, "adaptive_cbox_passes.npy"), p.astype("<u4"))
print(json.loads(r.stdout.strip().splitlines()[-1])["seconds"], "s;", "blocks", p.size, "capped at least once", int((p >= 1000).sum()),
      "median passes per round", float(np.median(p)) / 8)
