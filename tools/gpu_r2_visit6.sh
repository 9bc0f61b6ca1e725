#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python tools/coherent_bwd.py 64 > gpurun_out/coherent_bwd.log 2>&1; echo "coherent rc=$?"; grep -v "^{" gpurun_out/coherent_bwd.log
bash tools/gpu_render_list.sh
python tools/launch_breakdown.py gpurun_out/render_cfg5_launches.csv gpurun_out/render_cfg2_launches.csv
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:rt_primary|rt_trace|rt_shade|rt_scatter' -s 50 -c 5 \
  -o gpurun_out/render_full -f python tools/render_bench.py cfg5 100 > gpurun_out/ncu_render.log 2>&1; echo "ncu rc=$?"
