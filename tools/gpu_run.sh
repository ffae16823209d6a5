#!/bin/bash
# The GPU-box command lines behind every measured number, one target per kind (run as
#   gpurun --timeout T [--gpus N] -- 'bash tools/gpu_run.sh <target> [args]'
# from the repo root; every target writes into gpurun_out/ under a hard `timeout`).
#   tests            pytest -m gpu (whole suite)
#   bench [args]     bench.py with its default cfg-2 workload, JSON line -> gpurun_out/bench.json
#   benchN N [args]  torchrun bench.py on N GPUs
#   depth            60-block parity vs the fp32 oracle (tests/test_depth_gpu.py, prints the per-block growth)
#   kernels [batch]  per-kernel table alone (tools/kernels_json.py)
#   ncu_mlp_up       ncu --set full of the dominant launch -> profiles-ready CSV (bench.py's roofline.traffic)
#   ncu_kernels W    ncu --set full of prof_kernels.py W (gemm|qkv|attn|elem|all) -> summary CSV
#   ncu_step         launch list (gpu__time_duration) of the timed region of a 3-step bench
#   attn_pipe        attention half-tile pipeline: tests, sweep vs the default, wait counters, step bench
#   sass             per-kernel SASS mnemonic counts of the in-tree library (no GPU needed, works here too)
#   sanitizer        compute-sanitizer racecheck + synccheck on the tiny config (opt-in marker)
set -u
mkdir -p gpurun_out
T=${1:-tests}; shift || true
NCU_FULL="ncu --set full --clock-control none --import-source on"
case "$T" in
  tests)
    timeout -s KILL 1500 python -m pytest tests -x -q -m gpu "$@" > gpurun_out/tests.log 2>&1; echo "tests rc=$?"; tail -15 gpurun_out/tests.log;;
  bench)
    timeout -s KILL 1200 python bench.py "$@" > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -5 gpurun_out/bench.err
    python tools/show_bench.py gpurun_out/bench.json;;
  benchN)
    N=$1; shift
    timeout -s KILL 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus "$N" "$@" > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"; tail -5 gpurun_out/bench_n$N.err
    python tools/show_bench.py gpurun_out/bench_n$N.json;;
  depth)
    timeout -s KILL 900 python -m pytest tests/test_depth_gpu.py -x -q -s > gpurun_out/depth.log 2>&1; echo "depth rc=$?"
    grep -E "DEPTH|passed|failed|Error" gpurun_out/depth.log | cut -c1-1600;;
  kernels)
    timeout -s KILL 600 python tools/kernels_json.py "${1:-2}" > gpurun_out/kernels.json 2> gpurun_out/kernels.err; echo "kernels rc=$?"; tail -3 gpurun_out/kernels.err
    python tools/show_bench.py gpurun_out/kernels.json;;
  ncu_mlp_up)
    timeout -s KILL 600 python tools/prof_kernels.py mlp_up 2 > /dev/null 2>&1 || { echo "plain run failed"; exit 1; }
    timeout -s KILL 900 $NCU_FULL -k regex:"gemm_pair|gemm_bf16" -c 3 -f -o gpurun_out/ncu_mlp_up python tools/prof_kernels.py mlp_up 2 > gpurun_out/ncu_mlp_up.log 2>&1; echo "ncu rc=$?"
    ncu -i gpurun_out/ncu_mlp_up.ncu-rep --page raw --csv --print-units base > gpurun_out/ncu_mlp_up_raw.csv 2>/dev/null
    python tools/ncu_summary.py gpurun_out/ncu_mlp_up_raw.csv gpurun_out/ncu_mlp_up_batch2.csv \
      "ncu --set full --clock-control none, tools/prof_kernels.py mlp_up 2 (MLP-up + GELU pair GEMM, cfg-2 shapes, batch 2, img+txt in one launch), mean of 3 launches"
    cat gpurun_out/ncu_mlp_up_batch2.csv;;
  ncu_kernels)
    W=${1:-all}
    timeout -s KILL 600 python tools/prof_kernels.py "$W" 2 > /dev/null 2>&1 || { echo "plain run failed"; exit 1; }
    timeout -s KILL 1500 $NCU_FULL -k regex:"gemm_pair|gemm_bf16|attention_fwd|norm_rows|qknorm|cfg_euler" -c 24 -f -o gpurun_out/ncu_$W python tools/prof_kernels.py "$W" 2 > gpurun_out/ncu_$W.log 2>&1; echo "ncu rc=$?"
    ncu -i gpurun_out/ncu_$W.ncu-rep --page raw --csv --print-units base > gpurun_out/ncu_${W}_raw.csv 2>/dev/null
    python tools/ncu_summary.py gpurun_out/ncu_${W}_raw.csv gpurun_out/ncu_${W}_summary.csv "ncu --set full --clock-control none, tools/prof_kernels.py $W 2 (cfg-2 shapes, batch 2)"
    [ -n "$KEEP_REP" ] || rm -f gpurun_out/ncu_$W.ncu-rep     # gpurun_out/ comes back only below 64 MiB
    cat gpurun_out/ncu_${W}_summary.csv;;
  ncu_attn)
    timeout -s KILL 300 python tools/prof_attn.py > /dev/null 2>&1 || { echo "plain run failed"; exit 1; }
    timeout -s KILL 900 $NCU_FULL -c 12 -f -o gpurun_out/ncu_attn python tools/prof_attn.py > gpurun_out/ncu_attn.log 2>&1; echo "ncu rc=$?"
    ncu -i gpurun_out/ncu_attn.ncu-rep --page raw --csv --print-units base > gpurun_out/ncu_attn_raw.csv 2>/dev/null
    ncu -i gpurun_out/ncu_attn.ncu-rep --page details --csv --print-units base > gpurun_out/ncu_attn_details.csv 2>/dev/null
    python tools/ncu_summary.py gpurun_out/ncu_attn_raw.csv gpurun_out/ncu_attn_summary.csv "ncu --set full --clock-control none, tools/prof_attn.py (batch 2, 4608 tokens, 24x128 heads): ours then cuDNN SDPA"
    cat gpurun_out/ncu_attn_summary.csv;;
  ncu_step)
    timeout -s KILL 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-kernels --no-parity --e2e-steps 4 > gpurun_out/step_plain.json 2>&1 || { echo "plain run failed"; exit 1; }
    timeout -s KILL 1500 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 1400 --csv \
      --log-file gpurun_out/launch_list_step.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-kernels --no-parity --e2e-steps 4 > gpurun_out/ncu_step.log 2>&1; echo "ncu rc=$?"
    python tools/launch_share.py gpurun_out/launch_list_step.csv;;
  attn_pipe)
    # the Q-in-TMEM half-tile pipeline (variants 40+ / 56+): parity tests, interleaved sweep beside the current default, wait
    # counters, then a step bench with the fastest variant that passed.  These variants live in the experiments library
    # (build it first: make -C tools/experiments); the sweep, the counters and the bench below load it via QI_LIB_PATH.
    export QI_LIB_PATH=${QI_LIB_PATH:-$PWD/tools/experiments/libqi_b200_exp.so}
    timeout -s KILL 420 python -m pytest tests/test_ops_gpu.py -x -q -k "attention" > gpurun_out/attn_tests.log 2>&1; RC=$?; echo "attention tests rc=$RC"; tail -4 gpurun_out/attn_tests.log
    [ $RC -eq 0 ] || exit 1
    QI_SWEEP_VARIANTS=${QI_SWEEP_VARIANTS:-11,43,59,40,56,42,58,44,60} timeout -s KILL 300 python tools/attn_sweep.py > gpurun_out/attn_pipe_sweep.txt 2>&1; echo "sweep rc=$?"; cat gpurun_out/attn_pipe_sweep.txt
    timeout -s KILL 120 python tools/attn_waits.py 11 43 59 > gpurun_out/attn_pipe_waits.txt 2>&1; echo "waits rc=$?"; cat gpurun_out/attn_pipe_waits.txt
    BEST=$(python - <<'PY'
import re
best, bv = None, 0
for line in open("gpurun_out/attn_pipe_sweep.txt"):
    m = re.match(r"B=2 S=4608 H=24 variant=\s*(\d+).*median\s+([\d.]+) us", line)
    if m and (best is None or float(m.group(2)) < best):
        best, bv = float(m.group(2)), int(m.group(1))
print(bv)
PY
)
    echo "fastest variant at the step shape: $BEST"
    QI_ATTN_VARIANT=$BEST timeout -s KILL 600 python bench.py --no-cpu-baseline "$@" > gpurun_out/bench_attn_pipe.json 2> gpurun_out/bench_attn_pipe.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_attn_pipe.err
    python tools/show_bench.py gpurun_out/bench_attn_pipe.json;;
  sass)
    python tools/sass_summary.py > gpurun_out/sass_summary.txt; cat gpurun_out/sass_summary.txt;;
  sanitizer)
    QI_SANITIZER=1 timeout -s KILL 1500 python -m pytest tests/test_sanitizer_gpu.py -x -q -s > gpurun_out/sanitizer.log 2>&1; echo "sanitizer rc=$?"; tail -20 gpurun_out/sanitizer.log;;
  *)
    echo "unknown target $T"; exit 2;;
esac
