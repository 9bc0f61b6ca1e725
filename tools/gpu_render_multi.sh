#!/bin/bash
# N-GPU render visit: both render configs under torchrun (films summed with one allreduce).
set -u
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tools/render_bench.py cfg2 > gpurun_out/render_cfg2_n$N.json 2> gpurun_out/render_cfg2_n$N.err; echo "cfg2 rc=$?"
timeout 900 $TR tools/render_bench.py cfg5 100 > gpurun_out/render_cfg5_n$N.json 2> gpurun_out/render_cfg5_n$N.err; echo "cfg5 rc=$?"
if [ "$N" = "2" ]; then timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -3; fi
cut -c1-300 gpurun_out/render_cfg2_n$N.json; cut -c1-300 gpurun_out/render_cfg5_n$N.json
