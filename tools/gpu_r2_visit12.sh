#!/bin/bash
# 1-GPU visit on the final build: full GPU suite, bench line + reference arm, launch list of the bench, full ncu
# captures of eval_pdf / sample / backward, the time split of one rank's share of config 2 on 8 GPUs, quick_perf
set -u
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpu.txt 2>&1; nproc >> gpurun_out/gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 300 python tools/quick_perf.py 26 > gpurun_out/quick_perf.log 2>&1; echo "quick rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
  --log-file gpurun_out/launches.csv python bench.py --steps 5 --warmup 3 --preheat 0 --no-cpu-baseline --no-render \
  > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
bash tools/gpu_ncu.sh eval_pdf sample backward
# one rank's share of config 2 on 8 GPUs (8 of 64 spp): wall time per frame, then the kernels' own time
RD_EMULATE_WORLD=8 timeout 300 python tools/render_bench.py cfg2 2>/dev/null | tail -1 | cut -c1-400 | tee gpurun_out/cfg2_rank_share.json
RD_EMULATE_WORLD=8 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"rt_|stream_multi" -s 150 -c 50 --csv \
  --log-file gpurun_out/cfg2_rank_share_launches.csv python tools/render_bench.py cfg2 > /dev/null 2>&1; echo "ncu share rc=$?"
python tools/launch_breakdown.py gpurun_out/cfg2_rank_share_launches.csv
for c in cfg2 "cfg5 100"; do timeout 300 python tools/render_bench.py $c 2>/dev/null | tail -1 | cut -c1-330; done
cut -c1-2500 gpurun_out/bench.json
tail -12 gpurun_out/quick_perf.log
