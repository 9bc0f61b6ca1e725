#!/bin/bash
# ncu --set full capture of the harness kernels (trace / scatter / shade) of one mid-path bounce of config 5.
set -u
mkdir -p gpurun_out
NT=${1:-100}
timeout 600 python tools/render_bench.py cfg5 $NT > gpurun_out/render_plain.json 2> gpurun_out/render_plain.err || { echo "plain run failed"; exit 1; }
cut -c1-200 gpurun_out/render_plain.json
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:rt_trace|rt_scatter|rt_shade' -s 39 -c 3 \
  -o gpurun_out/render_full -f python tools/render_bench.py cfg5 $NT > gpurun_out/ncu_render.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/ncu_render.log
