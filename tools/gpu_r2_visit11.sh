#!/bin/bash
# N-GPU tuning sweep of the switch kernel (tools/nvls_sweep.py)
set -u
N=${1:-8}
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 tools/nvls_sweep.py 2>gpurun_out/nvls_sweep_n$N.err | tee gpurun_out/nvls_sweep_n$N.log
tail -3 gpurun_out/nvls_sweep_n$N.err | cut -c1-300
