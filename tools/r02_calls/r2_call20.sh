#!/bin/bash
# round 2, call 20: ncu --set full of the direct reconstruct_kernel after the division shortcut
mkdir -p gpurun_out
O=gpurun_out/c20
EVP_B200_KERNEL=sweep timeout 600 python tools/profile_reconstruct.py tp01 2 > $O.plain.log 2>&1; echo "rc=$? plain" | tee $O.rc
EVP_B200_KERNEL=sweep timeout 900 ncu --set full --clock-control none --import-source on -k regex:reconstruct_kernel -c 1 -o $O.reconstruct_direct -f \
    python tools/profile_reconstruct.py tp01 1 > $O.ncu.log 2>&1; echo "rc=$? ncu" | tee -a $O.rc
cat $O.plain.log; tail -n 3 $O.ncu.log
