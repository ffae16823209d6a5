#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_ops_gpu.py -q > gpurun_out/t_ops.log 2>&1; echo "ops rc=$?"; tail -5 gpurun_out/t_ops.log
timeout -s KILL 300 python tools/time_kernels.py > gpurun_out/time_kernels.log 2>&1; echo "time rc=$?"; cat gpurun_out/time_kernels.log
