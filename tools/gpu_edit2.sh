#!/bin/bash
# edit mode on 2 GPUs: pure Ulysses-2 (sharded reference tokens, ragged text) and CFG-parallel
mkdir -p gpurun_out
run() { timeout -s KILL 70 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 tools/mgpu_parity.py; }
QI_GRAPH=1 QI_CFG_PARALLEL=0 QI_SP_EXCHANGE=peer QI_PAR_EDIT=1 QI_PAR_TXT=33 run 29591 > gpurun_out/edit_sp2.log 2>&1; echo "sp2 rc=$?"; grep -E "MGPU_PARITY|Error|error" gpurun_out/edit_sp2.log | tail -3
QI_GRAPH=1 QI_PAR_EDIT=1 run 29592 > gpurun_out/edit_cfg2.log 2>&1; echo "cfg2 rc=$?"; grep -E "MGPU_PARITY|Error|error" gpurun_out/edit_cfg2.log | tail -3
