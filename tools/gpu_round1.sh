#!/bin/bash
# First GPU contact: each group in its own process under a hard timeout so a deadlocked kernel
# cannot hang the box.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/smi.txt 2>&1
timeout -s KILL 240 python -m pytest tests/test_ops_gpu.py -q -k "not gemm and not attention" > gpurun_out/t_elem.log 2>&1; echo "elem rc=$?"
timeout -s KILL 120 python tools/diag.py gemm > gpurun_out/diag_gemm.log 2>&1; echo "diag gemm rc=$?"
timeout -s KILL 300 python -m pytest tests/test_ops_gpu.py -q -k "gemm" > gpurun_out/t_gemm.log 2>&1; echo "gemm rc=$?"
timeout -s KILL 120 python tools/diag.py attn > gpurun_out/diag_attn.log 2>&1; echo "diag attn rc=$?"
timeout -s KILL 300 python -m pytest tests/test_ops_gpu.py -q -k "attention" > gpurun_out/t_attn.log 2>&1; echo "attn rc=$?"
tail -5 gpurun_out/t_elem.log gpurun_out/diag_gemm.log gpurun_out/t_gemm.log gpurun_out/diag_attn.log gpurun_out/t_attn.log
