#!/bin/bash
# round 2, call 16: reconstruct_kernel variants (launch bounds) on tp01, and the transport device tests
mkdir -p gpurun_out
O=gpurun_out/c16
: > $O.rc
for b in 4 5 6 8; do
  EVP_RECON_MINB=$b timeout 600 python tools/profile_reconstruct.py tp01 3 > $O.minb$b.log 2>&1; echo "rc=$? minb$b" | tee -a $O.rc
  tail -n 1 $O.minb$b.log
done
timeout 600 python -m pytest tests/test_gpu_zt_transport.py -m gpu -q -x > $O.transport.log 2>&1; echo "rc=$? transport" | tee -a $O.rc
tail -n 2 $O.transport.log
