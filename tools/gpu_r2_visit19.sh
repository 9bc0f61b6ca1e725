#!/bin/bash
# 8-GPU diagnostic of the config-5 render leg: pass size default (adaptive) / explicit 32 Mi / 8 Mi, film in a peer buffer
set -u
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
P='import json,sys; d=json.loads(sys.stdin.read()); print(d["ms"], d["spp_mpix_per_s"], d["launches_per_rank"])'
for wf in 0 33554432 8388608; do
  echo "$N GPUs peer film wavefront=$wf cfg5: $(RD_FILM_PEER=1 MERL_WAVEFRONT=$wf timeout 300 $TR tools/render_bench.py cfg5 100 2>gpurun_out/rb.err | grep '^{' | tail -1 | python -c "$P"); $(grep '^# free' gpurun_out/rb.err)"
done | tee gpurun_out/render_n${N}_diag.log
