# tuning builds (tools/scripts/build_variants.sh) against the default library: CTAs per SM of the flat path_mis kernel
run() { echo "== $1 $2 $(timeout 120 python tools/render_xml.py $2 spp=$3 size=800x600 repeat=4 normalize=0 2>&1 | tail -1 | grep -o "'samples_per_s': [0-9.]*")"; }
for rep in 1 2; do
for v in cur p9 p10 p12; do
  if [ $v = cur ]; then unset NORIB200_LIB; else export NORIB200_LIB=$PWD/optix-renderer_b200/variants/libnorib200_$v.so; fi
  run $v tests/golden/scenes/cbox/cbox_path_mis.xml 512
done
done
