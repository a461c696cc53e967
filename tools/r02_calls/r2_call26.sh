#!/bin/bash
# round 2, call 26: horizontal_remap timing on tp01 after the flux-kernel restructure; per-kernel times (ncu launch list)
mkdir -p gpurun_out
O=gpurun_out/c26
: > $O.rc
timeout 900 python -m pytest tests/test_gpu_zt_transport.py -m gpu -q -x -k remap > $O.transport.log 2>&1; echo "rc=$? transport" | tee -a $O.rc
tail -n 2 $O.transport.log
timeout 900 python tools/profile_remap.py tp01 3 > $O.remap.log 2>&1; echo "rc=$? remap" | tee -a $O.rc
tail -n 5 $O.remap.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"edge_flux|update_fields|reconstruct|departure|halo_.*multi" --csv --log-file $O.launches.csv \
    python tools/profile_remap.py tp01 1 > $O.ncu.log 2>&1; echo "rc=$? ncu" | tee -a $O.rc
grep -v "^==" $O.launches.csv | cut -d, -f5,15- | tail -n 14
