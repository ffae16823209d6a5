#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 200 python tools/rank_shapes.py 4 1 > gpurun_out/rank_p4.log 2>&1; echo "rc=$?"; cat gpurun_out/rank_p4.log
timeout -s KILL 200 python tools/rank_shapes.py 8 2 > gpurun_out/rank_p8b2.log 2>&1; echo "rc=$?"; cat gpurun_out/rank_p8b2.log
timeout -s KILL 200 python tools/rank_shapes.py 2 1 > gpurun_out/rank_p2.log 2>&1; echo "rc=$?"; cat gpurun_out/rank_p2.log
