# times every optix-renderer_b200/variants/*.so (tuning builds of the tile kernels) on the tree scenes
python tools/make_c3_scene.py /tmp/c3p > /dev/null 2>&1
for lib in optix-renderer_b200/variants/*.so; do
  echo "== $lib"
  NORIB200_LIB=$lib python tools/render_xml.py tests/golden/scenes/table/table_path_mis.xml spp=64 repeat=3 normalize=0 | tail -1 | cut -c1-120
  NORIB200_LIB=$lib python tools/render_xml.py /tmp/c3p/c3_path_mis.xml spp=16 repeat=3 normalize=0 | tail -1 | cut -c1-120
done
