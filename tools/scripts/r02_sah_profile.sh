# where the 36 ms of the device SAH build of C3 go: ncu launch list of the builder's kernels (cold-cache, serialised: shares, not absolutes)
mkdir -p gpurun_out
python tools/make_c3_scene.py /tmp/c3s spp=16 > /dev/null 2>&1
python tools/sah_build_once.py /tmp/c3s/c3_path_mis.xml
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"nb_k_sah|nb_k_build|DeviceRadixSort|DeviceScan" --csv --log-file gpurun_out/r02_launches_sah_build.csv python tools/sah_build_once.py /tmp/c3s/c3_path_mis.xml > gpurun_out/ncu_sah.log 2>&1
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/r02_launches_sah_build.csv")) if len(r) > 10 and r[0].isdigit()]
t = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    name = r[4].split("(")[0].split("<")[0][-60:]; v = float(r[-1].replace(",", ""))
    unit = r[-2]
    ms = v / 1e6 if unit in ("nsecond", "ns") else v / 1e3 if unit in ("usecond", "us") else v
    t[name][0] += 1; t[name][1] += ms
tot = sum(v[1] for v in t.values())
with open("gpurun_out/r02_sah_build_breakdown.txt", "w") as f:
    for k, v in sorted(t.items(), key=lambda kv: -kv[1][1]):
        line = "%-62s launches %4d  %8.3f ms  %5.1f %%" % (k, v[0], v[1], 100 * v[1] / tot)
        print(line); f.write(line + "\n")
    f.write("total under ncu %.3f ms (serialised, cold cache)\n" % tot); print("total", tot)
PY
