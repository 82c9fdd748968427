mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_bvh_build_gpu.py -m gpu -q -x -s > gpurun_out/t_sah.log 2>&1; echo sah rc=$?
grep -v "^$" gpurun_out/t_sah.log | tail -12 | cut -c1-400
timeout 900 python tools/ab_builders.py c3=1 c5=1 > gpurun_out/ab_builders.jsonl 2> gpurun_out/ab_builders.err; cut -c1-330 gpurun_out/ab_builders.jsonl; tail -3 gpurun_out/ab_builders.err
