#!/bin/bash
timeout 1500 python tools/config_table.py > gpurun_out/config_table.jsonl 2> gpurun_out/config_table.err
timeout 400 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
timeout 300 python bench.py --impl reference > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err
