#!/bin/bash
# Round 2, call 48 (1 GPU): rows per work item of the loop kernel on the three L2-resident grids
mkdir -p gpurun_out
timeout 50 python tools/sweep_rpc_small.py > gpurun_out/c48.sweep.log 2>&1
echo "rc=$? sweep" | tee gpurun_out/c48.rc
cat gpurun_out/c48.sweep.log
