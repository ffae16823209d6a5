#!/bin/bash
# 1-GPU verification + evidence: GPU tests, smoke, bench, ncu launch list of the timed region, ncu full of the hot kernels
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests -q -m gpu -x > gpurun_out/t_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -4 gpurun_out/t_gpu.log
timeout -s KILL 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout -s KILL 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench1.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench1.log | cut -c1-3000
timeout -s KILL 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "bench ref rc=$?"; tail -1 gpurun_out/bench_ref.log | cut -c1-600
timeout -s KILL 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_step.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_step.log 2>&1; echo "ncu list rc=$?"; wc -l gpurun_out/launches_step.csv
timeout -s KILL 300 python tools/prof_kernels.py all > gpurun_out/plain_prof.log 2>&1 &&
timeout -s KILL 900 ncu --set full --clock-control none --import-source on -k regex:"gemm_pair|gemm_bf16|attention_fwd|norm_rows|qknorm|cfg_euler" -c 30 -o gpurun_out/prof_r1c python tools/prof_kernels.py all > gpurun_out/ncu_prof.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out | head -30
