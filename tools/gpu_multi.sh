#!/bin/bash
# N-GPU visit: the 2-GPU correctness tests (full output kept), then the bench line at N ranks (eval+pdf weak
# scaling, the gradient pass with its exchange step, both render configs, config 5 sharded by table).
set -u
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
if [ "$N" = "2" ]; then
  timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -v -s > gpurun_out/pytest_gpu_multi.log 2>&1; echo "pytest multi rc=$?" | tee -a gpurun_out/pytest_gpu_multi.log
  tail -5 gpurun_out/pytest_gpu_multi.log
fi
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,COLL,TUNING timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
grep -E "NVLS|Channel|Connected|algo|Algo|proto|NCCL version|comm 0x.* rank 0 " gpurun_out/bench_n$N.err | head -60 > gpurun_out/nccl_info_n$N.txt
timeout 300 $TR bench.py --impl reference --gpus $N --steps 5 --warmup 1 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; echo "ref rc=$?"
cut -c1-1500 gpurun_out/bench_n$N.json
tail -5 gpurun_out/bench_n$N.err
