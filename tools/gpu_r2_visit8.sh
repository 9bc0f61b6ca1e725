#!/bin/bash
# 2-GPU visit: full GPU suite (incl. the nvls exchange), exchange microbenchmark (main build vs per-thread-fence build),
# render with the per-lane sphere masks, coherent backward incl. the flat-surface case, sampler with the guide default
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -s -k "two_gpu" > gpurun_out/pytest_multi.log 2>&1; echo "multi rc=$?"; tail -4 gpurun_out/pytest_multi.log | cut -c1-400
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512"
timeout 300 $TR tools/exchange_bench.py > gpurun_out/exchange_n2.log 2> gpurun_out/exchange_n2.err; echo "exchange rc=$?"
grep -v "^{" gpurun_out/exchange_n2.log | cut -c1-250
MERL_B200_LIB=$PWD/mitsuba_customization_b200/variants/libmerl_xfence.so timeout 300 $TR tools/exchange_bench.py > gpurun_out/exchange_n2_fence.log 2>&1; echo "exchange(fence) rc=$?"
grep -v "^{" gpurun_out/exchange_n2_fence.log | cut -c1-250
timeout 1500 python -m pytest tests -m gpu -x -q -k "not two_gpu" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for c in cfg2 "cfg5 100"; do timeout 300 python tools/render_bench.py $c 2>/dev/null | tail -1 | cut -c1-300; done
timeout 300 python tools/coherent_bwd.py 64 2>&1 | grep -v "^{" | tee gpurun_out/coherent_bwd.log
for r in 1 2; do timeout 200 python tools/one_kernel.py sample 26 20 2>&1 | tail -1; done
