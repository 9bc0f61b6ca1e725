#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_render.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -5
python tools/render_bench.py cfg2 --cpu | tee gpurun_out/render_cfg2.json
python tools/render_bench.py cfg5 100 --cpu | tee gpurun_out/render_cfg5.json
