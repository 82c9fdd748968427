# A/B of libnorib200 variants on the volume scenes (C4): NORIB200_LIB picks the library
for v in base v1 cur; do
  if [ $v = cur ]; then unset NORIB200_LIB; else export NORIB200_LIB=$PWD/optix-renderer_b200/variants/libnorib200_$v.so; fi
  for s in vol/cbox_heterog_fluid_mis vol/cbox_heterog_fluid vol/cbox_val_mis plugins/volumelight_mis; do
    echo "== $v $s"
    python tools/render_xml.py tests/golden/scenes/$s.xml spp=64 size=800x600 repeat=3 normalize=1 2>&1 | tail -3
  done
done
unset NORIB200_LIB
python -m pytest tests -m gpu -x -q -k "vol or medium or plugin" 2>&1 | tail -5
