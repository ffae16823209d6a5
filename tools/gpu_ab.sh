#!/bin/bash
# A/B on ONE box (power-capped clocks differ between boxes): fused vs separate QK-norm/RoPE
mkdir -p gpurun_out
for tag in fused unfused fused2 unfused2; do
  case $tag in fused*) export QI_FUSE_QKNORM=1;; *) export QI_FUSE_QKNORM=0;; esac
  timeout -s KILL 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_$tag.log 2>&1
  echo "$tag rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_$tag.log) $(grep -o '"clocks": {[^}]*}' gpurun_out/bench_$tag.log)"
done
