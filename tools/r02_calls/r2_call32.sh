#!/bin/bash
# round 2, call 32: flux kernel occupancy variants
mkdir -p gpurun_out
O=gpurun_out/c32
: > $O.rc
timeout 900 python tools/profile_remap.py tp01 2 > $O.fb1.log 2>&1; echo "rc=$? fb1 (product)" | tee -a $O.rc
tail -n 2 $O.fb1.log | head -n 1
for v in fb6 fb8 fb10; do
  EVP_B200_LIB=cice5_b200/lib/exp_$v.so timeout 900 python tools/profile_remap.py tp01 2 > $O.$v.log 2>&1; echo "rc=$? $v" | tee -a $O.rc
  tail -n 2 $O.$v.log | head -n 1
done
