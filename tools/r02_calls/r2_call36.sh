#!/bin/bash
# round 2, call 36 (N GPUs): transport_remap timing on tp01 as N processes (slowest rank, CUDA events)
N=${1:-2}
mkdir -p gpurun_out
O=gpurun_out/c36_n$N
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 \
    tests/run_multigpu_parity.py --grid tp01 --scenario full --calls 1 --transport 2 --no-evp-check 1 --dt-scale 1 > $O.log 2>&1; echo "rc=$? n$N" | tee $O.rc
grep "MULTIGPU" $O.log
