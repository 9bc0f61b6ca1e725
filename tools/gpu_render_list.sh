#!/bin/bash
# Per-kernel launch list (ncu gpu__time_duration) of one steady-state render of cfg5 (100 tables) and cfg2.
set -u
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"rt_|stream_multi" -s 1856 -c 232 --csv --log-file gpurun_out/render_cfg5_launches.csv \
  python tools/render_bench.py cfg5 100 > gpurun_out/ncu_r5.log 2>&1; echo "ncu cfg5 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"rt_|stream_multi" -s 232 -c 116 --csv --log-file gpurun_out/render_cfg2_launches.csv \
  python tools/render_bench.py cfg2 > gpurun_out/ncu_r2.log 2>&1; echo "ncu cfg2 rc=$?"
