#!/bin/bash
# round 2, call 21: direct reconstruct_kernel variants (blocks per SM x prefetch of the next tracer) and the TMA kernel
mkdir -p gpurun_out
O=gpurun_out/c21
: > $O.rc
timeout 600 python -m pytest tests/test_gpu_zt_transport.py -m gpu -q -x > $O.transport.log 2>&1; echo "rc=$? transport" | tee -a $O.rc
tail -n 2 $O.transport.log
timeout 600 python tools/profile_reconstruct.py tp01 3 > $O.tma.log 2>&1; echo "rc=$? tma" | tee -a $O.rc
tail -n 1 $O.tma.log
for v in d8 d8p d6 d6p d5p d4p; do
  EVP_B200_KERNEL=sweep EVP_B200_LIB=cice5_b200/lib/exp_$v.so timeout 600 python tools/profile_reconstruct.py tp01 3 > $O.$v.log 2>&1; echo "rc=$? $v" | tee -a $O.rc
  tail -n 1 $O.$v.log
done
