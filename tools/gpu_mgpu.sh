#!/bin/bash
mkdir -p gpurun_out
N=${1:-2}
timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_parity.py > gpurun_out/mgpu_parity_$N.log 2>&1; echo "parity N=$N rc=$?"; grep -E "MGPU_PARITY|Error|error" gpurun_out/mgpu_parity_$N.log | head; tail -5 gpurun_out/mgpu_parity_$N.log
timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 6 --warmup 3 > gpurun_out/bench_$N.log 2>&1; echo "bench N=$N rc=$?"; tail -2 gpurun_out/bench_$N.log | cut -c1-1200
