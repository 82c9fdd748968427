# full GPU test suite, smoke and the bench line
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
echo "val_mis $(timeout 100 python tools/render_xml.py tests/golden/scenes/vol/cbox_val_mis.xml spp=128 size=800x600 repeat=3 normalize=0 | tail -1 | cut -c1-120)"
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py 2> gpurun_out/bench_final.err | tail -1 > gpurun_out/bench_final.json
cut -c1-240 gpurun_out/bench_final.json
