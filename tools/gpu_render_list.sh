#!/bin/bash
# Per-kernel launch list (ncu gpu__time_duration) of the render harness: cfg2 and cfg5 (20 tables to keep setup short).
set -u
mkdir -p gpurun_out
python tools/render_bench.py cfg2 | tee gpurun_out/render_cfg2.json
python tools/render_bench.py cfg5 100 | tee gpurun_out/render_cfg5.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 1500 -c 700 --csv --log-file gpurun_out/render_cfg2_launches.csv \
  python tools/render_bench.py cfg2 > gpurun_out/ncu_r2.log 2>&1; echo "ncu cfg2 rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 700 --csv --log-file gpurun_out/render_cfg5_launches.csv \
  python tools/render_bench.py cfg5 20 > gpurun_out/ncu_r5.log 2>&1; echo "ncu cfg5 rc=$?"
