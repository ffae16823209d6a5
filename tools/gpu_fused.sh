#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests -q -x -m gpu > gpurun_out/t_gpu.log 2>&1; echo "gpu tests rc=$?"; tail -6 gpurun_out/t_gpu.log
timeout -s KILL 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout -s KILL 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench1.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench1.log | cut -c1-3000
