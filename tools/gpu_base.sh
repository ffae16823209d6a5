#!/bin/bash
# 1-GPU: GPU tests, smoke, bench (PDL on / off)
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests -q -m gpu -x > gpurun_out/t_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -4 gpurun_out/t_gpu.log
timeout -s KILL 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout -s KILL 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench1.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench1.log | cut -c1-1200
QI_PDL=0 timeout -s KILL 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench1_nopdl.log 2>&1; echo "bench nopdl rc=$?"; tail -1 gpurun_out/bench1_nopdl.log | cut -c1-300
