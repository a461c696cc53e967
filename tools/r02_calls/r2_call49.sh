#!/bin/bash
# Round 2, call 49 (1 GPU, the round's last GPU seconds): one row per work item (EVP_B200_RPC=1) on the L2-resident grids:
# bitwise parity on gx3 (real grid) and the device time per subcycle beside the rule's 2 rows
mkdir -p gpurun_out
O=gpurun_out/c49
EVP_B200_RPC=1 timeout 25 python -m pytest "tests/test_gpu_parity.py::test_one_call_bitwise_strict" -m gpu -q -x -k "gx3" > "$O.parity_rpc1.log" 2>&1
echo "rc=$? parity_rpc1" | tee "$O.rc"
timeout 20 python tools/sweep_rpc_small.py gx3,gx1,tp1 0,1 > "$O.sweep.log" 2>&1
echo "rc=$? sweep" | tee -a "$O.rc"
tail -n 2 "$O.parity_rpc1.log"; cat "$O.sweep.log"
