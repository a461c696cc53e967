#!/bin/bash
# round 2, call 15: ncu --set full of reconstruct_kernel on tp01 (after the same command ran plainly)
mkdir -p gpurun_out
O=gpurun_out/c15
timeout 600 python tools/profile_reconstruct.py tp01 3 > $O.plain.log 2>&1; echo "rc=$? plain" | tee $O.rc
timeout 900 ncu --set full --clock-control none --import-source on -k regex:reconstruct_kernel -c 1 -o $O.reconstruct -f \
    python tools/profile_reconstruct.py tp01 1 > $O.ncu.log 2>&1; echo "rc=$? ncu" | tee -a $O.rc
cat $O.plain.log; tail -n 5 $O.ncu.log
