#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 180 python -m pytest tests/test_ops_gpu.py -q -x -k "cta_pair" > gpurun_out/t_pair.log 2>&1; rc=$?; echo "pair test rc=$rc"; tail -15 gpurun_out/t_pair.log
if [ $rc -ne 0 ]; then export QI_GEMM_PAIR=0; echo "PAIR DISABLED for the rest"; fi
timeout -s KILL 600 python -m pytest tests -q -m gpu -x --deselect tests/test_ops_gpu.py::test_gemm_cta_pair > gpurun_out/t_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -4 gpurun_out/t_gpu.log
timeout -s KILL 400 python tools/gemm_sweep.py 1 > gpurun_out/gemm_sweep1.log 2>&1; echo "sweep rc=$?"; cat gpurun_out/gemm_sweep1.log | tail -30
timeout -s KILL 400 python tools/gemm_sweep.py 4 > gpurun_out/gemm_sweep4.log 2>&1; echo "sweep4 rc=$?"; cat gpurun_out/gemm_sweep4.log | tail -30
timeout -s KILL 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench1.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench1.log | cut -c1-400
