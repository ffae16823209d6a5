#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_ops_gpu.py -q -x > gpurun_out/t_ops.log 2>&1; echo "ops tests rc=$?"; tail -3 gpurun_out/t_ops.log
timeout -s KILL 120 python tools/attn_sweep.py > gpurun_out/attn_sweep.log 2>&1; echo "sweep rc=$?"; cat gpurun_out/attn_sweep.log
timeout -s KILL 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench1.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench1.log | cut -c1-3000
