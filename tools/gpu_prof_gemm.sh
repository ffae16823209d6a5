#!/bin/bash
mkdir -p gpurun_out
for st in 3 4 5 6 7; do
  echo "== pair stages $st"; QI_GEMM_PAIR_STAGES=$st timeout -s KILL 200 python tools/gemm_sweep.py 1 1 pair256 2>&1 | grep -E "pair=1 bn=256"
done
timeout -s KILL 300 python tools/prof_kernels.py gemm > gpurun_out/plain_prof.log 2>&1 && \
timeout -s KILL 900 ncu --set full --clock-control none --import-source on -k regex:"gemm_pair|gemm_bf16" -c 6 -o gpurun_out/prof_pair python tools/prof_kernels.py gemm > gpurun_out/ncu_prof.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_prof.log
