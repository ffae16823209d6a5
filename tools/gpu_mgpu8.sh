#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_parity.py > gpurun_out/mgpu_parity_8.log 2>&1; echo "parity N=8 rc=$?"; grep -E "MGPU_PARITY" gpurun_out/mgpu_parity_8.log | head -2
for N in 8 4; do
timeout -s KILL 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps 6 --warmup 3 > gpurun_out/bench_$N.log 2>&1; echo "bench N=$N rc=$?"; grep '"metric"' gpurun_out/bench_$N.log | cut -c1-400
done
