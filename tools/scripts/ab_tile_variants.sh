# tuning builds (tools_build_variants.sh) against the default library: CTAs per SM of the EXT kernels
run() { echo "== $1 $2 $(timeout 90 python tools/render_xml.py tests/golden/scenes/$2.xml spp=64 size=800x600 repeat=3 normalize=0 2>&1 | tail -1 | grep -o "'samples_per_s': [0-9.]*")"; }
for v in cur e10 e12; do
  if [ $v = cur ]; then unset NORIB200_LIB; else export NORIB200_LIB=$PWD/optix-renderer_b200/variants/libnorib200_$v.so; fi
  for s in plugins/disney_spot_dir_path_mis plugins/disney_path_mats_area plugins/normalmap_direct_mis plugins/disney_spot_dir_direct_mis envmap/envmap_sphere plugins/preview plugins/volumelight_mis; do run $v $s; done
done
