#!/bin/bash
# One GPU-box visit: parity tests, per-kernel timings, the bench line, the ncu launch list and one
# full capture of the eval_pdf kernel.  Everything lands in gpurun_out/.
set -u
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python tools/quick_perf.py 26 > gpurun_out/quick_perf.log 2>&1; echo "quick rc=$?"
if [ -x tools/microbench ]; then timeout 120 tools/microbench > gpurun_out/microbench.log 2>&1; fi
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file gpurun_out/launches.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-render \
  > gpurun_out/ncu_launches.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:stream_kernel -s 3 -c 1 \
  -o gpurun_out/eval_pdf_full -f python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras \
  > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -3 gpurun_out/pytest_gpu.log
cat gpurun_out/quick_perf.log | tail -15
cat gpurun_out/bench.json
