#!/bin/bash
# Why is one rank's share of configs 2 / 5 on 8 GPUs slower than before?  Rank share emulated on one GPU:
# final library vs the library before the two-half trace kernels, default pass size vs 8 Mi-path passes.
set -u
mkdir -p gpurun_out
P=$PWD/mitsuba_customization_b200/variants/libmerl_presplit.so
for c in cfg2 "cfg5 100"; do
  for wf in 0 8388608; do
    echo "final    wavefront=$wf $c: $(RD_EMULATE_WORLD=8 MERL_WAVEFRONT=$wf timeout 300 python tools/render_bench.py $c 2>/dev/null | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["ms"], d["launches_per_rank"])')"
    echo "presplit wavefront=$wf $c: $(MERL_B200_LIB=$P RD_EMULATE_WORLD=8 MERL_WAVEFRONT=$wf timeout 300 python tools/render_bench.py $c 2>/dev/null | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["ms"], d["launches_per_rank"])')"
  done
done | tee gpurun_out/rank_share_ab.log
