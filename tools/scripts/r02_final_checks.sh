# end-of-round checks on one GPU: adaptive throughput, the CLI on a million-triangle scene with the device-built tree
mkdir -p gpurun_out
python tools/adaptive_rate.py > gpurun_out/r02_adaptive_rate.jsonl 2> gpurun_out/adaptive_rate.err; cat gpurun_out/r02_adaptive_rate.jsonl; tail -2 gpurun_out/adaptive_rate.err
python tools/make_c3_scene.py /tmp/c3f spp=16 > /dev/null 2>&1
for b in sah host; do optix-renderer_b200/nori_b200 /tmp/c3f/c3_path_mis.xml /tmp/c3f/out_$b.npy spp=16 bvh=$b 2>&1 | tail -1 | cut -c1-400; done
python - <<'PY'
import numpy as np
a, b = np.load('/tmp/c3f/out_sah.npy'), np.load('/tmp/c3f/out_host.npy')
print("c3 cli sah vs host: mean", a.mean(), b.mean(), "max abs diff", float(np.abs(a - b).max()))
PY
