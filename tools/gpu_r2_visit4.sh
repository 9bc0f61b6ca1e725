#!/bin/bash
# 1-GPU visit: full parity suite on the new build (CTA-aggregated backward, one Philox call per hit, shuffle scan,
# persistent primary kernel), sampler variants round 2, coherent-batch backward, render configs.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
bash tools/gpu_r2_ab.sh sample "sample"
timeout 300 python tools/coherent_bwd.py 64 > gpurun_out/coherent_bwd.log 2>&1; echo "coherent rc=$?"; grep -v "^{" gpurun_out/coherent_bwd.log
timeout 300 python tools/render_bench.py cfg2 > gpurun_out/render_cfg2.json 2> gpurun_out/render_cfg2.err; echo "cfg2 rc=$?"; cut -c1-300 gpurun_out/render_cfg2.json
timeout 600 python tools/render_bench.py cfg5 100 > gpurun_out/render_cfg5.json 2> gpurun_out/render_cfg5.err; echo "cfg5 rc=$?"; cut -c1-300 gpurun_out/render_cfg5.json
