# tuning builds (tools/scripts/build_variants.sh) against the default library: walk-kernel phase thresholds at 7 CTAs per SM
python tools/make_c3_scene.py /tmp/c3p > /dev/null 2>&1
run() { echo "== $1 $2 $(timeout 120 python tools/render_xml.py $2 spp=$3 repeat=4 normalize=0 2>&1 | tail -1 | grep -o "'samples_per_s': [0-9.]*")"; }
for rep in 1 2; do
for v in cur r12 l10 l12r12; do
  if [ $v = cur ]; then unset NORIB200_LIB; else export NORIB200_LIB=$PWD/optix-renderer_b200/variants/libnorib200_$v.so; fi
  run $v tests/golden/scenes/table/table_path_mis.xml 64
  run $v /tmp/c3p/c3_path_mis.xml 16
done
done
