#!/bin/bash
# round 2, call 23: direct reconstruct_kernel with L2 prefetch of the tracer planes ahead
mkdir -p gpurun_out
O=gpurun_out/c23
: > $O.rc
EVP_B200_KERNEL=sweep timeout 600 python tools/profile_reconstruct.py tp01 3 > $O.a0.log 2>&1; echo "rc=$? a0" | tee -a $O.rc
tail -n 1 $O.a0.log
for v in a2 a4 a11 a4h4 a4b4p a4b6; do
  EVP_B200_KERNEL=sweep EVP_B200_LIB=cice5_b200/lib/exp_$v.so timeout 600 python tools/profile_reconstruct.py tp01 3 > $O.$v.log 2>&1; echo "rc=$? $v" | tee -a $O.rc
  tail -n 1 $O.$v.log
done
