#!/bin/bash
# round 2, call 40: evp + transport_remap step after step on the device-resident state (6 steps small tripole at 1 / 2 ranks,
# 12 steps on the 1-degree tripole with ndte = 120), every step bitwise against the oracle
mkdir -p gpurun_out
O=gpurun_out/c40
timeout 900 python -m pytest tests/test_gpu_zt_transport.py -m gpu -q -x -k "dynamics" --durations=3 > $O.steps.log 2>&1; echo "rc=$? steps" | tee $O.rc
tail -n 8 $O.steps.log
