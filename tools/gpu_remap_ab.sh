#!/bin/bash
# A/B of the destination-ordered scatter tiles on N GPUs (pure Ulysses, one CFG branch = the per-rank work of the
# 2N-GPU layout): per-op trace with the plain and the remapped column-tile order.  usage: gpu_remap_ab.sh N
N=${1:-4}
mkdir -p gpurun_out
timeout -s KILL 300 python -m pytest tests/test_ops_gpu.py -q -x -m gpu -k "scatter" > gpurun_out/t_scatter.log 2>&1; echo "scatter tests rc=$?"; tail -3 gpurun_out/t_scatter.log
for mode in 0 1 0 1; do
  QI_SCATTER_REMAP=$mode QI_CFG_PARALLEL=0 QI_TRACE_NO_CFG=1 QI_TRACE_LAYERS=12 timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2955$mode tools/mgpu_trace.py > gpurun_out/trace_remap${mode}.log 2>&1; echo "remap=$mode rc=$?"
  grep -E "traced eager step|gemm_qkv|attention_scatter|peer_barrier|gemm_out" gpurun_out/trace_remap${mode}.log
done
