#!/bin/bash
# 1-GPU visit: sampler build variants A/B (+ their parity tests), the extended microbenchmark, plugin host tests
set -u
mkdir -p gpurun_out
bash tools/gpu_r2_ab.sh sample "sample"
timeout 300 tools/microbench > gpurun_out/microbench2.log 2>&1; echo "micro rc=$?"
tail -30 gpurun_out/microbench2.log
timeout 600 python -m pytest tests/test_plugins.py tests/test_z_plugin_ad.py -m gpu -x -q 2>&1 | tail -3
