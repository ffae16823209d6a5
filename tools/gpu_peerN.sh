#!/bin/bash
# N-GPU: parity (peer exchange, graph) then bench with peer (and optionally nccl)
N=${1:-4}
mkdir -p gpurun_out
QI_GRAPH=1 QI_SP_EXCHANGE=peer timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 tools/mgpu_parity.py > gpurun_out/par_peer_$N.log 2>&1
echo "parity peer N=$N rc=$?"; grep -E "MGPU_PARITY|timed out" gpurun_out/par_peer_$N.log | head -2; grep -v -i warn gpurun_out/par_peer_$N.log | grep -i -E "error|Traceback" | head -3
for ex in ${2:-peer}; do
QI_SP_EXCHANGE=$ex timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${ex}_$N.log 2>&1; echo "bench $ex N=$N rc=$?"; grep '"metric"' gpurun_out/bench_${ex}_$N.log | cut -c1-160; grep -v -i warn gpurun_out/bench_${ex}_$N.log | grep -i -E "error|Traceback" | head -3
done
