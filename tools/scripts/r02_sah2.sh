mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_bvh_build_gpu.py tests/test_render_gpu.py tests/test_volume_gpu.py -m gpu -q -x > gpurun_out/t_sah2.log 2>&1; echo rc=$?
grep -v "^$" gpurun_out/t_sah2.log | tail -8 | cut -c1-300
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_sah.json 2> gpurun_out/bench_sah.err; tail -c 3000 gpurun_out/bench_sah.json
