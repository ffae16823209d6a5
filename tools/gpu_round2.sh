#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_model_gpu.py -q > gpurun_out/t_model.log 2>&1; echo "model rc=$?"
tail -15 gpurun_out/t_model.log
timeout -s KILL 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout -s KILL 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench1.log 2>&1; echo "bench rc=$?"; tail -5 gpurun_out/bench1.log
timeout -s KILL 600 python bench.py --steps 5 --warmup 3 --no-graph --no-cpu-baseline > gpurun_out/bench1_nograph.log 2>&1; echo "bench nograph rc=$?"; tail -2 gpurun_out/bench1_nograph.log
