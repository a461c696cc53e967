#!/bin/bash
# Round 2, call 50 (1 GPU, whatever seconds are left): the library's own rule now picks one row per item on gx3
mkdir -p gpurun_out
timeout 12 python tools/sweep_rpc_small.py gx3 0,2 > gpurun_out/c50.sweep.log 2>&1
echo "rc=$? sweep" | tee gpurun_out/c50.rc
cat gpurun_out/c50.sweep.log
