mkdir -p gpurun_out
python -m pytest tests/test_wide_gpu.py tests/test_render_gpu.py tests/test_trace_gpu.py tests/test_c5_gpu.py -m gpu -q -x > gpurun_out/t_quick.log 2>&1; echo quick rc=$?
grep -v "^$" gpurun_out/t_quick.log | tail -6 | cut -c1-300
python tools/ab_wide.py c5=1 > gpurun_out/ab_wide.jsonl 2> gpurun_out/ab_wide.err; cut -c1-210 gpurun_out/ab_wide.jsonl; tail -3 gpurun_out/ab_wide.err
for w in 0 1; do python tools/render_xml.py tests/golden/scenes/veach/veach_mis.xml spp=128 repeat=3 wide=$w | tail -2 | head -1 | cut -c1-120; done
