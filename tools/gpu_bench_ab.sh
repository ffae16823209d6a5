#!/bin/bash
# same-box A/B of two builds of the library on the 1-GPU bench
mkdir -p gpurun_out
for lib in narrow wide narrow wide; do
  if [ $lib = narrow ]; then export QI_LIB_PATH=$PWD/tools/ab/libqi_b200_narrow.so; else unset QI_LIB_PATH; fi
  timeout -s KILL 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_$lib.log 2>&1
  echo "$lib rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_$lib.log) $(grep -o '"gemm_qkv_qknorm_rope": {"us": [0-9.]*' gpurun_out/bench_$lib.log) $(grep -o '"gemm_mlp_down_gated_residual": {"us": [0-9.]*' gpurun_out/bench_$lib.log) $(grep -o '"sm_mhz": [0-9.]*' gpurun_out/bench_$lib.log)"
done
