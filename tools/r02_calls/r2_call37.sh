#!/bin/bash
# round 2, call 37 (2 GPUs): one ghost-update operation for all field planes: device tests, two-process parity, tp01 timing
mkdir -p gpurun_out
O=gpurun_out/c37
: > $O.rc
timeout 900 python -m pytest tests/test_gpu_zt_transport.py -m gpu -q -x > $O.transport.log 2>&1; echo "rc=$? transport" | tee -a $O.rc
tail -n 2 $O.transport.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 \
    tests/run_multigpu_parity.py --grid tp1 --calls 1 --transport 1 > $O.tp1.log 2>&1; echo "rc=$? tp1 parity" | tee -a $O.rc
grep "MULTIGPU" $O.tp1.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29522 \
    tests/run_multigpu_parity.py --grid tp01 --scenario full --calls 1 --transport 2 --no-evp-check 1 --dt-scale 1 > $O.tp01.log 2>&1; echo "rc=$? tp01 timing" | tee -a $O.rc
grep "MULTIGPU" $O.tp01.log
