#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out
timeout -s KILL 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 tools/mgpu_trace.py > gpurun_out/trace_n$N.log 2>&1; echo "trace N=$N rc=$?"; grep -A16 "^rank" gpurun_out/trace_n$N.log | head -60
timeout -s KILL 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n$N.log 2>&1; echo "bench N=$N rc=$?"; grep '"metric"' gpurun_out/bench_n$N.log | cut -c1-900; grep -v -i warn gpurun_out/bench_n$N.log | grep -i -E "error|Traceback" | head -5
