#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests -q -m gpu > gpurun_out/t_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -8 gpurun_out/t_gpu.log
timeout -s KILL 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout -s KILL 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_graph.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_graph.log
# launch list of one eager step (skip schedule prep + 3 warm-up steps = 4 + 3*1209 launches)
timeout -s KILL 600 python bench.py --steps 1 --warmup 3 --no-graph --no-cpu-baseline > gpurun_out/plain_step.log 2>&1 &&
timeout -s KILL 1500 ncu --metrics gpu__time_duration.sum --clock-control none -s 3631 -c 1209 --csv --log-file gpurun_out/launches_step.csv python bench.py --steps 1 --warmup 3 --no-graph --no-cpu-baseline > gpurun_out/ncu_step.log 2>&1; echo "ncu list rc=$?"
timeout -s KILL 300 python tools/prof_kernels.py all > gpurun_out/plain_prof.log 2>&1 &&
timeout -s KILL 900 ncu --set full --clock-control none --import-source on -k regex:"gemm_bf16|attention_fwd|norm_rows|qknorm|cfg_euler" -o gpurun_out/prof_r1 python tools/prof_kernels.py all > gpurun_out/ncu_prof.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out | head -30
