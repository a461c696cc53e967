#!/bin/bash
# round 2, call 19: reconstruction kernels after the zero-numerator shortcut: TMA (product), variants, direct
mkdir -p gpurun_out
O=gpurun_out/c19
: > $O.rc
timeout 600 python -m pytest tests/test_gpu_zt_transport.py -m gpu -q -x > $O.transport.log 2>&1; echo "rc=$? transport" | tee -a $O.rc
tail -n 2 $O.transport.log
timeout 600 python tools/profile_reconstruct.py tp01 3 > $O.product.log 2>&1; echo "rc=$? product" | tee -a $O.rc
tail -n 1 $O.product.log
for v in rc_b2 rc_t8b4 rc_t16b1; do
  EVP_B200_LIB=cice5_b200/lib/exp_$v.so timeout 600 python tools/profile_reconstruct.py tp01 3 > $O.$v.log 2>&1; echo "rc=$? $v" | tee -a $O.rc
  tail -n 1 $O.$v.log
done
EVP_B200_KERNEL=sweep timeout 600 python tools/profile_reconstruct.py tp01 3 > $O.direct.log 2>&1; echo "rc=$? direct" | tee -a $O.rc
tail -n 1 $O.direct.log
