#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_ops_gpu.py -q -x -k "gemm" > gpurun_out/t_gemm.log 2>&1; echo "gemm tests rc=$?"; tail -3 gpurun_out/t_gemm.log
timeout -s KILL 400 python tools/gemm_sweep.py 4 1 > gpurun_out/gemm_sweep4.log 2>&1; echo "sweep4 rc=$?"; cat gpurun_out/gemm_sweep4.log
timeout -s KILL 400 python tools/gemm_sweep.py 2 1 > gpurun_out/gemm_sweep2.log 2>&1; echo "sweep2 rc=$?"; cat gpurun_out/gemm_sweep2.log
