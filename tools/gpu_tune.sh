#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
timeout 300 python tools/quick_perf.py 26 2>&1 | grep -E "^(eval|pdf|bins|sample|backward)" | tee gpurun_out/quick4.log
