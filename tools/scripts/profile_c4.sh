# ncu captures of the volume kernels (C4) and the smoke entry
set -x
NCU="ncu --set full --clock-control none --import-source on"
$NCU -k regex:nb_k_tile -c 1 -f -o gpurun_out/r01_tile_heterog_mats python tools/render_xml.py tests/golden/scenes/vol/cbox_heterog_fluid.xml spp=32 > gpurun_out/ncu_h1.log 2>&1
$NCU -k regex:nb_k_tile -c 1 -f -o gpurun_out/r01_tile_heterog_mis python tools/render_xml.py tests/golden/scenes/vol/cbox_heterog_fluid_mis.xml spp=16 > gpurun_out/ncu_h2.log 2>&1
$NCU -k regex:nb_k_tile -c 1 -f -o gpurun_out/r01_tile_homog_mis python tools/render_xml.py tests/golden/scenes/vol/cbox_val_mis.xml spp=64 > gpurun_out/ncu_h3.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()"
