#!/bin/bash
# round 2, call 29: flux kernel with the points in shared memory and the two edge orientations side by side
mkdir -p gpurun_out
O=gpurun_out/c29
: > $O.rc
timeout 900 python -m pytest tests/test_gpu_zt_transport.py -m gpu -q -x -k remap > $O.transport.log 2>&1; echo "rc=$? transport" | tee -a $O.rc
tail -n 2 $O.transport.log
timeout 900 python tools/profile_remap.py tp01 2 > $O.remap.log 2>&1; echo "rc=$? remap" | tee -a $O.rc
tail -n 3 $O.remap.log
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"edge_flux|update_fields" --csv --log-file $O.launches.csv \
    python tools/profile_remap.py tp01 1 > $O.ncu.log 2>&1; echo "rc=$? ncu" | tee -a $O.rc
grep -v "^==" $O.launches.csv | cut -d, -f5,13- | tail -n 6
