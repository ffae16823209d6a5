#!/bin/bash
# GPU check of the rows built after a-e: full GPU suite, smoke, VAE decode timing.
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests -q -x -m gpu -rs > gpurun_out/t_gpu.log 2>&1; echo "gpu tests rc=$?"; tail -25 gpurun_out/t_gpu.log
timeout -s KILL 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout -s KILL 300 python tools/vae_time.py > gpurun_out/vae_time.log 2>&1; echo "vae rc=$?"; tail -4 gpurun_out/vae_time.log
