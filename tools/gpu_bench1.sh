#!/bin/bash
mkdir -p gpurun_out
SECONDS=0
timeout -s KILL 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_n1.log 2>&1; echo "bench rc=$? in ${SECONDS}s"; grep '"metric"' gpurun_out/bench_n1.log | tail -1
