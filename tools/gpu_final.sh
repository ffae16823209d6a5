#!/bin/bash
# final 1-GPU check of the tree: whole GPU suite, smoke, split-launch tuning at one rank's cfg2 x ulysses4 shapes
mkdir -p gpurun_out
timeout -s KILL 400 python -m pytest tests -q -x -m gpu -rs > gpurun_out/t_final.log 2>&1; echo "gpu tests rc=$?"; tail -6 gpurun_out/t_final.log
timeout -s KILL 120 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout -s KILL 120 python tools/split_tune.py 4 1 > gpurun_out/split_tune_p4.log 2>&1; echo "split rc=$?"; grep -E "SPLIT_TUNE|Error" gpurun_out/split_tune_p4.log
