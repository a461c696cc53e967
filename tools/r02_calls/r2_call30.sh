#!/bin/bash
# round 2, call 30: transport_remap on the device (tests); flux-kernel variants (edge in x / z, points in shared / local)
mkdir -p gpurun_out
O=gpurun_out/c30
: > $O.rc
timeout 900 python -m pytest tests/test_gpu_zt_transport.py -m gpu -q -x > $O.transport.log 2>&1; echo "rc=$? transport" | tee -a $O.rc
tail -n 2 $O.transport.log
timeout 900 python tools/profile_remap.py tp01 2 > $O.x1s0.log 2>&1; echo "rc=$? x1s0 (product)" | tee -a $O.rc
tail -n 2 $O.x1s0.log | head -n 1
for v in x1s1 x0s1 x0s0; do
  EVP_B200_LIB=cice5_b200/lib/exp_$v.so timeout 900 python tools/profile_remap.py tp01 2 > $O.$v.log 2>&1; echo "rc=$? $v" | tee -a $O.rc
  tail -n 2 $O.$v.log | head -n 1
done
