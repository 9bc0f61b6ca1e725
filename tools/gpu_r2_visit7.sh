#!/bin/bash
# N-GPU visit: exchange microbenchmark by size and transport, then the bench line (gradient scaling, render legs)
set -u
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR tools/exchange_bench.py > gpurun_out/exchange_n$N.log 2> gpurun_out/exchange_n$N.err; echo "exchange rc=$?"
grep -v "^{" gpurun_out/exchange_n$N.log
timeout 900 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_n$N.err | cut -c1-300
python - <<PY
import json
b=json.loads([l for l in open("gpurun_out/bench_n$N.json") if l.startswith('{"metric"')][-1])
print(b["value"], b["roofline"]["frac"], b["e2e"]["value"], b["e2e"].get("host_link_ceiling"))
print(json.dumps(b["extras"].get("grad_scaling"), indent=1)[:3500])
print(json.dumps(b["extras"].get("render"), indent=1)[:2500])
PY
