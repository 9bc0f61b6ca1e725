#!/bin/bash
mkdir -p gpurun_out
timeout 800 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python tools/render_bench.py cfg2 --cpu | tee gpurun_out/render_cfg2.json
python tools/render_bench.py cfg5 100 --cpu | tee gpurun_out/render_cfg5.json
