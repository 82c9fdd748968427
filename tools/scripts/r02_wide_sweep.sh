# one GPU call: the full gpu suite, the A/B of the traversal modes, the tuning variants of the wide walk kernel, ncu of it
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --durations=4 > gpurun_out/t_all.log 2>&1; echo suite rc=$?
grep -v "^$" gpurun_out/t_all.log | tail -14 | cut -c1-300
python tools/make_c3_scene.py /tmp/c3p > /dev/null 2>&1
for lib in optix-renderer_b200/variants/*.so; do
  echo "== $lib"
  NORIB200_LIB=$lib python tools/render_xml.py tests/golden/scenes/table/table_path_mis.xml spp=64 repeat=3 normalize=0 | tail -1 | cut -c1-150
  NORIB200_LIB=$lib python tools/render_xml.py /tmp/c3p/c3_path_mis.xml spp=16 repeat=3 normalize=0 | tail -1 | cut -c1-150
done
NCU="timeout 300 ncu --set full --clock-control none --import-source on"
$NCU -k regex:nb_k_tile_walk -c 1 -f -o gpurun_out/r02_walk_wide_table python tools/render_xml.py tests/golden/scenes/table/table_path_mis.xml spp=16 > gpurun_out/ncu_t.log 2>&1
$NCU -k regex:nb_k_tile_walk -c 1 -f -o gpurun_out/r02_walk_wide_c3 python tools/render_xml.py /tmp/c3p/c3_path_mis.xml spp=4 > gpurun_out/ncu_c3.log 2>&1
ls -la gpurun_out/*.ncu-rep
