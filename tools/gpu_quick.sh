#!/bin/bash
# Parity tests + per-kernel timings of the current build.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python tools/quick_perf.py 26 2>&1 | grep -v "^{" | tee gpurun_out/quick_perf.log
