# one-GPU measurement pass of the round: config table (GPU beside the reference CPU renderer), bench line,
# ncu launch list of the bench command, ncu full captures of the dominant kernels
set -x
mkdir -p gpurun_out
python tools/make_c3_scene.py c3cache > /dev/null 2>&1
python tools/make_c5_scene.py c5cache width=1920 height=1080 spp=16 > /dev/null 2>&1
timeout 420 python tools/config_table.py > gpurun_out/r01_config_table.jsonl 2> gpurun_out/config_table.err
timeout 300 python bench.py > gpurun_out/r01_bench.json 2> gpurun_out/bench.err || exit 1
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r01_bench_reference.json 2>> gpurun_out/bench.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01_launches_bench.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
NCU="timeout 300 ncu --set full --clock-control none --import-source on"
$NCU -k regex:nb_k_tile -c 1 -f -o gpurun_out/r01_tile_cbox python bench.py --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_b.log 2>&1
$NCU -k regex:nb_k_tile_walk -c 1 -f -o gpurun_out/r01_walk_table python tools/render_xml.py tests/golden/scenes/table/table_path_mis.xml spp=16 > gpurun_out/ncu_t.log 2>&1
$NCU -k regex:nb_k_tile_walk -c 1 -f -o gpurun_out/r01_walk_c3 python tools/render_xml.py c3cache/c3_path_mis.xml spp=4 > gpurun_out/ncu_c3.log 2>&1
ls -la gpurun_out/*.ncu-rep
