#!/bin/bash
# round 2, call 33: flux kernel with the tracers type by type; 8 / 6 / 5 blocks per SM
mkdir -p gpurun_out
O=gpurun_out/c33
: > $O.rc
timeout 900 python -m pytest tests/test_gpu_zt_transport.py -m gpu -q -x -k "remap" > $O.transport.log 2>&1; echo "rc=$? transport" | tee -a $O.rc
tail -n 2 $O.transport.log
timeout 900 python tools/profile_remap.py tp01 2 > $O.fb8.log 2>&1; echo "rc=$? fb8 (product)" | tee -a $O.rc
tail -n 2 $O.fb8.log | head -n 1
for v in fb6 fb5; do
  EVP_B200_LIB=cice5_b200/lib/exp_$v.so timeout 900 python tools/profile_remap.py tp01 2 > $O.$v.log 2>&1; echo "rc=$? $v" | tee -a $O.rc
  tail -n 2 $O.$v.log | head -n 1
done
