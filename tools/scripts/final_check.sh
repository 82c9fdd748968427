# full GPU test suite, the 8(f) / C4 rows of the config table, smoke and the bench line
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for k in heterog disney "normal maps" "volume lights" envmap; do timeout 300 python tools/config_table.py "only=$k" 2>> gpurun_out/config_ext.err; done > gpurun_out/config_ext.jsonl
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py 2> gpurun_out/bench_final.err | tail -1 > gpurun_out/bench_final.json
cut -c1-300 gpurun_out/bench_final.json
