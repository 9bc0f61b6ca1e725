#!/bin/bash
# 1-GPU visit: renderer with the two-half trace kernels (hit queue) -- parity, throughput by wavefront size, launch list
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_render.py -m gpu -x -q > gpurun_out/pytest_render.log 2>&1; echo "render tests rc=$?"; tail -3 gpurun_out/pytest_render.log
for wf in 0 16777216 33554432 67108864; do
  for c in cfg2 "cfg5 100"; do
    echo "wavefront=$wf $c: $(MERL_WAVEFRONT=$wf timeout 300 python tools/render_bench.py $c 2>/dev/null | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["ms"], d["spp_mpix_per_s"], d["launches_per_rank"])')"
  done
done | tee gpurun_out/render_wavefront.log
bash tools/gpu_render_list.sh > /dev/null 2>&1
python tools/launch_breakdown.py gpurun_out/render_cfg5_launches.csv gpurun_out/render_cfg2_launches.csv
