#!/bin/bash
# round 2, call 31 (2 GPUs): evp + transport_remap as two processes with peer-memory halos, bitwise against the oracle
mkdir -p gpurun_out
O=gpurun_out/c31
: > $O.rc
for g in tp1 tp025; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
      tests/run_multigpu_parity.py --grid $g --calls 1 --transport 1 > $O.$g.log 2>&1; echo "rc=$? $g" | tee -a $O.rc
  grep "MULTIGPU" $O.$g.log
done
