#!/bin/bash
# round 2, call 25: horizontal_remap on the device: parity tests, timing on tp01
mkdir -p gpurun_out
O=gpurun_out/c25
: > $O.rc
timeout 900 python -m pytest tests/test_gpu_zt_transport.py -m gpu -q -x > $O.transport.log 2>&1; echo "rc=$? transport" | tee -a $O.rc
tail -n 3 $O.transport.log
timeout 900 python tools/profile_remap.py tp01 3 > $O.remap.log 2>&1; echo "rc=$? remap" | tee -a $O.rc
tail -n 6 $O.remap.log
