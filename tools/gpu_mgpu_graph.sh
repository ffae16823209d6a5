#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
QI_GRAPH=1 timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_parity.py > gpurun_out/mgpu_parity_graph_$N.log 2>&1; echo "parity graph N=$N rc=$?"; grep -E "MGPU_PARITY" gpurun_out/mgpu_parity_graph_$N.log | head -1; grep -v -i warn gpurun_out/mgpu_parity_graph_$N.log | grep -i -E "error" | head -5
timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 6 --warmup 3 > gpurun_out/bench_graph_$N.log 2>&1; echo "bench N=$N rc=$?"; grep '"metric"' gpurun_out/bench_graph_$N.log | cut -c1-300;  grep -v -i warn gpurun_out/bench_graph_$N.log | grep -i -E "error" | head -5
