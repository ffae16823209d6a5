#!/bin/bash
# 2-GPU: op tests (incl. emulated scatter), pure-SP2 parity over both exchanges, ragged+batch, then bench in both modes
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_ops_gpu.py -q -x > gpurun_out/t_ops.log 2>&1; echo "ops rc=$?"; tail -3 gpurun_out/t_ops.log
run_par() { # name, env...
  name=$1; shift
  env "$@" timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 tools/mgpu_parity.py > gpurun_out/par_$name.log 2>&1
  echo "parity $name rc=$?"; grep -E "MGPU_PARITY|timed out" gpurun_out/par_$name.log | head -2; grep -v -i warn gpurun_out/par_$name.log | grep -i -E "error|Traceback" | head -3
}
run_par sp2_peer_eager QI_GRAPH=0 QI_CFG_PARALLEL=0 QI_SP_EXCHANGE=peer
run_par sp2_peer_graph QI_GRAPH=1 QI_CFG_PARALLEL=0 QI_SP_EXCHANGE=peer
run_par sp2_nccl_graph QI_GRAPH=1 QI_CFG_PARALLEL=0 QI_SP_EXCHANGE=nccl
run_par sp2_peer_ragged QI_GRAPH=1 QI_CFG_PARALLEL=0 QI_SP_EXCHANGE=peer QI_PAR_H=272 QI_PAR_W=240 QI_PAR_TXT=33 QI_PAR_BATCH=2
for ex in peer nccl; do
QI_CFG_PARALLEL=0 QI_SP_EXCHANGE=$ex timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 2 --steps 6 --warmup 3 --no-cpu-baseline > gpurun_out/bench_sp2_$ex.log 2>&1; echo "bench sp2 $ex rc=$?"; grep '"metric"' gpurun_out/bench_sp2_$ex.log | cut -c1-160; grep -v -i warn gpurun_out/bench_sp2_$ex.log | grep -i -E "error|Traceback" | head -3
done
