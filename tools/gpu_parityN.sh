#!/bin/bash
# N-GPU parity vs the oracle (tiny config): default layout (graph), pure SP (graph), pure SP ragged + batch 2
N=${1:-2}
mkdir -p gpurun_out
run_par() { name=$1; shift
  env "$@" timeout -s KILL 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 tools/mgpu_parity.py > gpurun_out/par_${N}_$name.log 2>&1
  echo "parity N=$N $name rc=$?"; grep -E "MGPU_PARITY|timed out" gpurun_out/par_${N}_$name.log | head -2; grep -v -i warn gpurun_out/par_${N}_$name.log | grep -i -E "error|Traceback" | head -3
}
run_par default_graph QI_GRAPH=1
run_par puresp_graph QI_GRAPH=1 QI_CFG_PARALLEL=0
run_par puresp_ragged_b2 QI_GRAPH=1 QI_CFG_PARALLEL=0 QI_PAR_H=272 QI_PAR_W=240 QI_PAR_TXT=33 QI_PAR_BATCH=2
run_par puresp_unfused QI_GRAPH=0 QI_CFG_PARALLEL=0 QI_FUSE_QKNORM=0
