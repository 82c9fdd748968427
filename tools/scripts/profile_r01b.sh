set -x
NCU="ncu --set full --clock-control none --import-source on"
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01_launches_bench.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
$NCU -k regex:nb_k_tile -c 1 -f -o gpurun_out/r01_bench_tile python bench.py --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_b.log 2>&1
python tools/render_xml.py tests/golden/scenes/vol/cbox_heterog_fluid_mis.xml spp=16 > gpurun_out/plain_het.log 2>&1 && \
$NCU -k regex:nb_k_tile -c 1 -f -o gpurun_out/r01_tile_heterog_mis python tools/render_xml.py tests/golden/scenes/vol/cbox_heterog_fluid_mis.xml spp=16 > gpurun_out/ncu_h.log 2>&1
$NCU -k regex:nb_k_tile -c 1 -f -o gpurun_out/r01_tile_heterog_mats python tools/render_xml.py tests/golden/scenes/vol/cbox_heterog_fluid.xml spp=16 > gpurun_out/ncu_h2.log 2>&1
python tools/make_c5_scene.py /tmp/c5p width=1920 height=1080 spp=4 > /dev/null
python tools/render_xml.py /tmp/c5p/c5_path_mis.xml spp=4 device_bvh=1 > gpurun_out/plain_c5.log 2>&1 && \
$NCU -k regex:nb_k_tile_walk -c 1 -f -o gpurun_out/r01_walk_c5 python tools/render_xml.py /tmp/c5p/c5_path_mis.xml spp=4 device_bvh=1 > gpurun_out/ncu_c5.log 2>&1
$NCU -k regex:"nb_k_build|DeviceRadixSort" -c 24 -f -o gpurun_out/r01_bvh_build python tools/render_xml.py /tmp/c5p/c5_path_mis.xml spp=1 device_bvh=1 > gpurun_out/ncu_bvh.log 2>&1
$NCU -k regex:"finalize" -c 1 -f -o gpurun_out/r01_film_finalize python tools/render_xml.py /tmp/c5p/c5_path_mis.xml spp=1 device_bvh=1 > gpurun_out/ncu_fin.log 2>&1
python tools/render_xml.py tests/golden/scenes/envmap/envmap_sphere.xml spp=64 > gpurun_out/plain_env.log 2>&1 && \
$NCU -k regex:nb_k_tile -c 1 -f -o gpurun_out/r01_tile_envmap python tools/render_xml.py tests/golden/scenes/envmap/envmap_sphere.xml spp=64 > gpurun_out/ncu_env.log 2>&1
ls -la gpurun_out/*.ncu-rep
