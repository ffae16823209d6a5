#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests -q -x -m gpu > gpurun_out/t_gpu.log 2>&1; echo "gpu tests rc=$?"; tail -4 gpurun_out/t_gpu.log
QI_TRACE_LAYERS=8 timeout -s KILL 300 python tools/mgpu_trace.py > gpurun_out/trace1.log 2>&1; echo "trace rc=$?"; tail -16 gpurun_out/trace1.log
for tag in tuned untuned; do
  case $tag in tuned) export QI_GEMM_TUNE=1;; *) export QI_GEMM_TUNE=0;; esac
  timeout -s KILL 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_$tag.log 2>&1
  echo "$tag rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_$tag.log) $(grep -o '"clocks": {[^}]*}' gpurun_out/bench_$tag.log)"
done
