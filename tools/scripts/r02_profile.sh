# round-2 evidence pass on ONE GPU: bench line, launch list, ncu --set full of the shipped kernels (+ one headline capture
# without the L2-flush buffer), everything the summaries under profiles/ are made from
set -x
mkdir -p gpurun_out
python tools/make_c3_scene.py /tmp/c3p > /dev/null 2>&1
python bench.py > gpurun_out/r02_bench.json 2> gpurun_out/bench.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_reference.json 2>> gpurun_out/bench.err
python bench.py --impl reference --steps 1 --warmup 0 --ref-spp 512 > gpurun_out/r02_bench_reference_512spp.json 2>> gpurun_out/bench.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_bench.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-configs > gpurun_out/ncu_launch.log 2>&1
NCU="timeout 400 ncu --set full --clock-control none --import-source on"
$NCU -k regex:nb_k_tile -c 1 -f -o gpurun_out/r02_tile_cbox python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-configs > gpurun_out/ncu_b.log 2>&1
$NCU -k regex:nb_k_tile -c 1 -f -o gpurun_out/r02_tile_cbox_noflush python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-configs --no-flush > gpurun_out/ncu_b2.log 2>&1
$NCU -k regex:nb_k_tile_walk -c 1 -f -o gpurun_out/r02_walk_table python tools/render_xml.py tests/golden/scenes/table/table_path_mis.xml spp=16 > gpurun_out/ncu_t.log 2>&1
$NCU -k regex:nb_k_tile_walk -c 1 -f -o gpurun_out/r02_walk_c3 python tools/render_xml.py /tmp/c3p/c3_path_mis.xml spp=4 > gpurun_out/ncu_c3.log 2>&1
$NCU -k regex:nb_k_tile -c 1 -f -o gpurun_out/r02_tile_heterog_mis python tools/render_xml.py tests/golden/scenes/vol/cbox_heterog_fluid_mis.xml spp=16 > gpurun_out/ncu_h.log 2>&1
$NCU -k regex:nb_k_tile -c 1 -f -o gpurun_out/r02_tile_veach python tools/render_xml.py tests/golden/scenes/veach/veach_mis.xml spp=64 > gpurun_out/ncu_v.log 2>&1
python tools/config_table.py > gpurun_out/r02_config_table.jsonl 2> gpurun_out/config_table.err
# summaries are made on the box (gpurun brings back at most 64 MiB): metrics CSV + source attribution per capture, then the reports go
for r in tile_cbox tile_cbox_noflush walk_table walk_c3 tile_heterog_mis tile_veach; do
  python tools/ncu_summary.py gpurun_out/r02_$r.ncu-rep gpurun_out/r02_ncu_$r.csv
done
python tools/ncu_meta.py gpurun_out/r02_ncu_tile_cbox.csv $((800*600*512)) "bench.py --steps 1 --warmup 0: cbox path_mis 800x600x512spp, first launch"
python tools/ncu_meta.py gpurun_out/r02_ncu_tile_cbox_noflush.csv $((800*600*512)) "same without the 256 MB L2-flush buffer"
python tools/ncu_meta.py gpurun_out/r02_ncu_walk_table.csv $((800*600*16)) "render_xml table_path_mis spp=16"
python tools/ncu_meta.py gpurun_out/r02_ncu_walk_c3.csv $((800*600*4)) "render_xml c3_path_mis spp=4"
python tools/ncu_meta.py gpurun_out/r02_ncu_tile_heterog_mis.csv $((800*600*16)) "render_xml cbox_heterog_fluid_mis spp=16"
python tools/ncu_meta.py gpurun_out/r02_ncu_tile_veach.csv $((768*512*64)) "render_xml veach_mis spp=64"
python tools/ncu_attrib.py gpurun_out/r02_tile_cbox.ncu-rep optix-renderer_b200/libnorib200.so nb_mega nb_k_tileILi0ELb1ELb0 top=30 > gpurun_out/r02_attrib_tile_cbox.txt 2>&1
python tools/ncu_attrib.py gpurun_out/r02_walk_table.ncu-rep optix-renderer_b200/libnorib200.so nb_mega tile_walkILb1ELb0ELb1 top=30 > gpurun_out/r02_attrib_walk_table.txt 2>&1
python tools/ncu_attrib.py gpurun_out/r02_walk_c3.ncu-rep optix-renderer_b200/libnorib200.so nb_mega tile_walkILb1ELb0ELb1 top=30 > gpurun_out/r02_attrib_walk_c3.txt 2>&1
python tools/ncu_attrib.py gpurun_out/r02_tile_heterog_mis.ncu-rep optix-renderer_b200/libnorib200.so nb_mega nb_k_tileILi9ELb1ELb1 top=30 > gpurun_out/r02_attrib_tile_heterog_mis.txt 2>&1
rm -f gpurun_out/*.ncu-rep
ls -la gpurun_out | head -40
