for lib in "" optix-renderer_b200/variants/libnorib200_nocoop.so "" optix-renderer_b200/variants/libnorib200_nocoop.so; do
  echo "== lib=$lib"
  NORIB200_LIB=$lib python bench.py --no-configs --no-cpu-baseline --steps 3 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']/1e6,1), 'Msamples/s', d['ms_per_step'])"
  NORIB200_LIB=$lib python bench.py --no-configs --no-cpu-baseline --steps 3 --warmup 3 --spp 64 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('64spp', round(d['value']/1e6,1), 'Msamples/s', d['ms_per_step'])"
done
