# sparse-volume layout: parity tests, then the cost of the layout branch on the dense fluid (C4) and the bricks' own rate
mkdir -p gpurun_out
python -m pytest tests/test_volume_gpu.py -m gpu -q -x > gpurun_out/t_vol.log 2>&1; echo vol rc=$?
grep -v "^$" gpurun_out/t_vol.log | tail -8 | cut -c1-300
for layout in dense bricks; do for x in cbox_heterog_fluid_mis cbox_heterog_fluid; do
  echo "== $layout $x"; NB_VOLUME_LAYOUT=$layout python tools/render_xml.py tests/golden/scenes/vol/$x.xml spp=64 repeat=3 normalize=0 2>&1 | tail -1 | cut -c1-120; done; done
