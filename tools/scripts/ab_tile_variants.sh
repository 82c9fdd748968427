# tuning builds (tools/scripts/build_variants.sh) against the default library: filter table in shared memory
run() { echo "== $1 $2 $(timeout 120 python tools/render_xml.py tests/golden/scenes/$2.xml spp=$3 size=800x600 repeat=4 normalize=0 2>&1 | tail -1 | grep -o "'samples_per_s': [0-9.]*")"; }
for rep in 1 2; do
for v in cur smt; do
  if [ $v = cur ]; then unset NORIB200_LIB; else export NORIB200_LIB=$PWD/optix-renderer_b200/variants/libnorib200_$v.so; fi
  run $v cbox/cbox_path_mis 512; run $v cbox/cbox_path_mats 512; run $v vol/cbox_val_mis 128
done
done
