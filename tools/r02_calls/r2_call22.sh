#!/bin/bash
# round 2, call 22: direct reconstruct_kernel with 2-D tiles per block (rows share neighbour rows in L1)
mkdir -p gpurun_out
O=gpurun_out/c22
: > $O.rc
EVP_B200_KERNEL=sweep timeout 600 python -m pytest tests/test_gpu_zt_transport.py -m gpu -q -x -k construct > $O.transport.log 2>&1; echo "rc=$? transport(direct)" | tee -a $O.rc
tail -n 2 $O.transport.log
EVP_B200_KERNEL=sweep timeout 600 python tools/profile_reconstruct.py tp01 3 > $O.h1.log 2>&1; echo "rc=$? h1" | tee -a $O.rc
tail -n 1 $O.h1.log
for v in h2 h4 h4b4p h4b6 h8; do
  EVP_B200_KERNEL=sweep EVP_B200_LIB=cice5_b200/lib/exp_$v.so timeout 600 python tools/profile_reconstruct.py tp01 3 > $O.$v.log 2>&1; echo "rc=$? $v" | tee -a $O.rc
  tail -n 1 $O.$v.log
done
