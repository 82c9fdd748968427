# C4 rows of the config table and ncu captures of the volume kernels after the code-size work
set -x
timeout 600 python tools/config_table.py only=vol > gpurun_out/config_vol.jsonl 2> gpurun_out/config_vol.err
NCU="ncu --set full --clock-control none --import-source on"
timeout 300 $NCU -k regex:nb_k_tile -c 1 -f -o gpurun_out/r01_tile_heterog_mis_after python tools/render_xml.py tests/golden/scenes/vol/cbox_heterog_fluid_mis.xml spp=16 > gpurun_out/ncu_h2a.log 2>&1
timeout 300 $NCU -k regex:nb_k_tile -c 1 -f -o gpurun_out/r01_tile_homog_mis_after python tools/render_xml.py tests/golden/scenes/vol/cbox_val_mis.xml spp=64 > gpurun_out/ncu_h3a.log 2>&1
