#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_model_gpu.py -q -x -m gpu -k "multi_gpu_parity" -rs > gpurun_out/t_mgpu2.log 2>&1; echo "rc=$?"; tail -8 gpurun_out/t_mgpu2.log
