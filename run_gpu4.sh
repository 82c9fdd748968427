#!/bin/bash
mkdir -p gpurun_out
python tools/render_once.py scene=cbox spp=128 strategy=tile repeat=3 > gpurun_out/strat.txt 2>&1
python tools/render_once.py scene=table spp=64 strategy=tile repeat=2 >> gpurun_out/strat.txt 2>&1
timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -5 >> gpurun_out/strat.txt
