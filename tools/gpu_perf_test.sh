#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_perf_gpu.py "tests/test_model_gpu.py::test_decoded_image_psnr" -q -x -s -m gpu > gpurun_out/t_perf.log 2>&1; echo "rc=$?"; grep -E "LIBRARY_BASELINE|PSNR|passed|failed|Error|assert" gpurun_out/t_perf.log | head -20
