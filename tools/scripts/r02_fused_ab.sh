# fused closest + shadow scan (NB_FUSED_SCAN) against the shipped kernel: throughput and per-seed image identity, cbox path_mis
# variants first: printf 'fused -DNB_FUSED_SCAN\nfused7 -DNB_FUSED_SCAN -DNB_TILE_MIN_BLOCKS=6\n' > /tmp/v.txt && bash tools/scripts/build_variants.sh /tmp/v.txt
mkdir -p gpurun_out
for lib in optix-renderer_b200/libnorib200.so optix-renderer_b200/variants/libnorib200_fused.so optix-renderer_b200/variants/libnorib200_fused7.so; do
  echo "== $lib"
  for x in cbox/cbox_path_mis odyssey/odyssey_mis; do
    NORIB200_LIB=$lib python - $x <<'PY'
import sys, os, numpy as np
sys.path.insert(0, os.getcwd())
import norib200 as nb
x = sys.argv[1]
s = nb.load_xml("tests/golden/scenes/%s.xml" % x)
ctx = nb.Context(0); ctx.load(s)
ctx.render_begin(nb.make_render_params(s, seed=1))
best = None
for _ in range(4):
    ctx.film_clear(); ctx.render_subframe(0, 256); c = ctx.counters()
    if best is None or c["device_ms"] < best["device_ms"]: best = c
img = ctx.film_download(normalize=True)
tag = os.path.basename(os.environ["NORIB200_LIB"]) + "_" + x.replace("/", "_")
np.save("gpurun_out/fused_%s.npy" % tag, img.astype(np.float32))
print(x, "M samples/s", round(best["samples"] / best["device_ms"] / 1e3, 1), "rays/sample", round((best["rays_closest"] + best["rays_any"]) / best["samples"], 4), "mean", float(img.mean()))
PY
  done
done
python - <<'PY'
import numpy as np, glob
for x in ("cbox_cbox_path_mis", "odyssey_odyssey_mis"):
    a = np.load("gpurun_out/fused_libnorib200.so_%s.npy" % x); b = np.load("gpurun_out/fused_libnorib200_fused.so_%s.npy" % x)
    print(x, "max rel diff shipped vs fused", float(np.abs(a - b).max() / np.abs(a).max()))
PY
rm -f gpurun_out/fused_*.npy
