#!/bin/bash
# round 2, call 28: ncu --set full of edge_flux_kernel on tp01
mkdir -p gpurun_out
O=gpurun_out/c28
timeout 900 python tools/profile_remap.py tp01 1 > $O.plain.log 2>&1; echo "rc=$? plain" | tee $O.rc
timeout 900 ncu --set full --clock-control none --import-source on -k regex:edge_flux_kernel -c 1 -o $O.edge_flux -f \
    python tools/profile_remap.py tp01 1 > $O.ncu.log 2>&1; echo "rc=$? ncu" | tee -a $O.rc
tail -n 3 $O.plain.log; tail -n 3 $O.ncu.log
