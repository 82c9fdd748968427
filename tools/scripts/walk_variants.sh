# times every optix-renderer_b200/variants/*.so (tuning builds of the tile kernels) on cbox and the tree scenes
for lib in optix-renderer_b200/variants/*.so; do
  echo "== $lib"
  NORIB200_LIB=$lib python tools/render_xml.py tests/golden/scenes/cbox/cbox_path_mis.xml spp=512 repeat=3 normalize=0 | tail -1 | cut -c1-120
  NORIB200_LIB=$lib python tools/render_xml.py tests/golden/scenes/table/table_path_mis.xml spp=64 repeat=3 normalize=0 strategy=tile | tail -1 | cut -c1-120
  NORIB200_LIB=$lib python tools/render_xml.py tests/golden/scenes/veach/veach_mis.xml spp=128 repeat=3 normalize=0 | tail -1 | cut -c1-120
done
