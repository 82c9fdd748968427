mkdir -p gpurun_out
python -m pytest tests/test_adaptive_gpu.py tests/test_volume_gpu.py -m gpu -q -x -s > gpurun_out/t_adp.log 2>&1; echo adaptive rc=$?
grep -v "^$" gpurun_out/t_adp.log | tail -14 | cut -c1-400
python bench.py --steps 5 --warmup 3 --no-configs --no-cpu-baseline 2>/dev/null | tail -1 | cut -c1-400
