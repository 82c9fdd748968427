mkdir -p gpurun_out
for lib in "" optix-renderer_b200/variants/libnorib200_smem.so ""; do
  echo "== lib=$lib"
  NORIB200_LIB=$lib python bench.py --no-configs --no-cpu-baseline --steps 3 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']/1e6,1), 'Msamples/s e2e', round(d['e2e']['value']/1e6,1), d['ms_per_step'])"
  for x in tests/golden/scenes/cbox/cbox_path_mats.xml tests/golden/scenes/vol/cbox_val_mis.xml tests/golden/scenes/table/table_path_mis.xml; do
    NORIB200_LIB=$lib python tools/render_xml.py $x spp=64 repeat=3 normalize=0 2>&1 | tail -1 | cut -c1-70
  done
done
