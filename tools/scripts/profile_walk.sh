set -x
NCU="ncu --set full --clock-control none --import-source on"
python tools/render_xml.py tests/golden/scenes/table/table_path_mis.xml spp=16 > gpurun_out/plain_table.log 2>&1 || exit 1
$NCU -k regex:nb_k_tile_walk -c 1 -f -o gpurun_out/s15_walk_table python tools/render_xml.py tests/golden/scenes/table/table_path_mis.xml spp=16 > gpurun_out/ncu_t.log 2>&1
ls -la gpurun_out/*.ncu-rep
