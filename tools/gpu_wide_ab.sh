#!/bin/bash
# A/B of 128-byte vs 64-byte row segments in the QKV epilogue's Ulysses scatter, N GPUs, pure Ulysses with one CFG
# branch (= one rank's work in the 2N-GPU layout).  usage: gpu_wide_ab.sh N
N=${1:-4}
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_ops_gpu.py tests/test_model_gpu.py -q -x -m gpu -k "not multi_gpu" > gpurun_out/t_wide.log 2>&1; echo "ops+model tests rc=$?"; tail -3 gpurun_out/t_wide.log
for lib in narrow wide narrow wide; do
  if [ $lib = narrow ]; then export QI_LIB_PATH=$PWD/tools/ab/libqi_b200_narrow.so; else unset QI_LIB_PATH; fi
  QI_CFG_PARALLEL=0 QI_TRACE_NO_CFG=1 QI_TRACE_LAYERS=12 timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561 tools/mgpu_trace.py > gpurun_out/trace_$lib.log 2>&1; echo "$lib rc=$?"
  grep -E "traced eager step|gemm_qkv|attention_scatter|peer_barrier" gpurun_out/trace_$lib.log
done
