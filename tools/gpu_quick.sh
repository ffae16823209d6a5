#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests -q -x -m gpu > gpurun_out/t_gpu.log 2>&1; echo "gpu tests rc=$?"; tail -4 gpurun_out/t_gpu.log
timeout -s KILL 300 python tools/rank_shapes.py 4 1 > gpurun_out/rank_p4.log 2>&1; echo "rc=$?"; cat gpurun_out/rank_p4.log
timeout -s KILL 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench1.log 2>&1; echo "bench rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench1.log) $(grep -o '"clocks": {[^}]*}' gpurun_out/bench1.log)"; grep -o '"kernels": {.*' gpurun_out/bench1.log | cut -c1-1200
