#!/bin/bash
# 2-GPU diagnostic of the render legs: film in a peer exchange buffer vs plain tensor, default pass size vs 8 Mi
set -u
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29516"
P='import json,sys; d=json.loads(sys.stdin.read()); print(d["ms"], d["launches_per_rank"])'
for peer in 1 0; do for wf in 0 8388608; do
  echo "2 GPUs peer_film=$peer wavefront=$wf cfg5: $(RD_FILM_PEER=$peer MERL_WAVEFRONT=$wf timeout 300 $TR tools/render_bench.py cfg5 100 2>gpurun_out/rb.err | grep '^{' | tail -1 | python -c "$P"); $(grep '^# free' gpurun_out/rb.err)"
done; done | tee gpurun_out/render_n2_diag.log
for wf in 0 8388608; do
  echo "1 GPU, share of 2, wavefront=$wf cfg5: $(RD_EMULATE_WORLD=2 MERL_WAVEFRONT=$wf timeout 300 python tools/render_bench.py cfg5 100 2>/dev/null | tail -1 | python -c "$P")"
done | tee -a gpurun_out/render_n2_diag.log
