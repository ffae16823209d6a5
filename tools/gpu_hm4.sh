#!/bin/bash
# head-major exchange buffers on 4 GPUs: two parity cases (cfg2 x ulysses2 graph; pure ulysses4 ragged batch 2) + per-op trace
mkdir -p gpurun_out
run() { timeout -s KILL 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port $1 $2; }
QI_GRAPH=1 QI_SP_EXCHANGE=peer run 29581 tools/mgpu_parity.py > gpurun_out/hm_par1.log 2>&1; echo "parity cfg2xsp2 rc=$?"; grep MGPU_PARITY gpurun_out/hm_par1.log
QI_GRAPH=1 QI_CFG_PARALLEL=0 QI_SP_EXCHANGE=peer QI_PAR_H=272 QI_PAR_W=240 QI_PAR_TXT=33 QI_PAR_BATCH=2 run 29582 tools/mgpu_parity.py > gpurun_out/hm_par2.log 2>&1; echo "parity sp4 ragged rc=$?"; grep MGPU_PARITY gpurun_out/hm_par2.log
QI_CFG_PARALLEL=0 QI_TRACE_NO_CFG=1 QI_TRACE_LAYERS=12 run 29583 tools/mgpu_trace.py > gpurun_out/trace_hm.log 2>&1; echo "trace rc=$?"
grep -E "traced eager step|gemm_|attention_scatter|peer_barrier|adaln" gpurun_out/trace_hm.log
