#!/bin/bash
# Round-2 first GPU visit (one GPU): full parity suite incl. the full-size render tests, the bench line with the
# new legs, the reference arm, the microbenchmark grid sweep, the sustained-rate probe, an ncu launch list of the
# bench and a full capture of the sample kernel (the one furthest below its floor).
set -u
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpu.txt 2>&1; nproc >> gpurun_out/gpu.txt; nvidia-smi -q -d POWER | grep -i -E "limit|draw" | head -8 >> gpurun_out/gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 300 python tools/quick_perf.py 26 > gpurun_out/quick_perf.log 2>&1; echo "quick rc=$?"
if [ -x tools/microbench ]; then timeout 300 tools/microbench > gpurun_out/microbench.log 2>&1; echo "micro rc=$?"; fi
timeout 300 python tools/sustained_probe.py eval_pdf 1000 50 > gpurun_out/sustained_eval_pdf.log 2>&1; echo "sustained rc=$?"
timeout 200 python tools/sustained_probe.py copy 1000 50 > gpurun_out/sustained_copy.log 2>&1; echo "sustained copy rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
  --log-file gpurun_out/launches.csv python bench.py --steps 5 --warmup 3 --preheat 0 --no-cpu-baseline --no-render \
  > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
bash tools/gpu_ncu.sh sample
cut -c1-3000 gpurun_out/bench.json
tail -12 gpurun_out/quick_perf.log
tail -16 gpurun_out/microbench.log
